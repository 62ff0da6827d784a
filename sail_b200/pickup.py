"""Picking and drag interaction without a DOM (mirror of /root/reference/src/core/pickup.js:1-123).

Same classes and methods as the reference -- ``Ray.generate / testBoundBox / intersectBoundBox`` and
``Pickup.pick / movingBegin / moving / movingEnd`` -- with two platform differences:

* ``pick`` asks the **device** which object the pixel's ray hits first (``sail_intersect_rays``: the very scan the
  trace kernel runs, so what is picked is what is drawn) instead of running the JS per-shape ``intersect`` loop
  (pickup.js:53-62, geometry.js:110-591, which uses other epsilons than the shader).  Objects whose
  ``boundbox()`` is ``False`` stay unpickable exactly as in the reference (geometry.js:51-53: the Cornell box).
* pixel coordinates are scaled by the renderer's size, not the hard-coded 512 of pickup.js:50-51.

The drag arithmetic (movement plane from the bound-box face that was hit, pickup.js:65-121) is host-side double
precision as in the reference.  A drag ends in ``scene.select.temporaryTranslate(..)`` with ``scene.moving`` set, which
makes the next ``Renderer.render`` restart the accumulation and re-upload the objects table
(renderer.js:57-60 -> ``sail_upload_objects``).
"""
import numpy as np

MINVALUE = 1e-4  # rollup.config.js:82-85
MAXVALUE = 1e5


class Ray:
    """pickup.js:3-38."""

    def __init__(self, origin, dir):
        self.origin = np.asarray(origin, dtype=np.float64).reshape(3)
        self.dir = np.asarray(dir, dtype=np.float64).reshape(3)

    @staticmethod
    def generate(eye, inmp, x, y):
        """pickup.js:9-12: un-project the NDC point (x, y, 0) and subtract the eye (direction not normalised)."""
        p = np.asarray(inmp, dtype=np.float64) @ np.array([x, y, 0.0, 1.0])
        return Ray(eye, p[:3] / p[3] - np.asarray(eye, dtype=np.float64).reshape(3))

    def _slabs(self, boundbox):
        with np.errstate(divide='ignore', invalid='ignore'):
            tMin = (np.asarray(boundbox['min'], dtype=np.float64) - self.origin) / self.dir
            tMax = (np.asarray(boundbox['max'], dtype=np.float64) - self.origin) / self.dir
        return np.minimum(tMin, tMax).max(), np.maximum(tMin, tMax).min()

    def testBoundBox(self, boundbox):
        """pickup.js:14-25."""
        tNear, tFar = self._slabs(boundbox)
        if tNear < -MINVALUE and tFar < -MINVALUE:
            return False
        return bool(tNear < tFar)

    def intersectBoundBox(self, boundbox):
        """pickup.js:27-38."""
        tNear, tFar = self._slabs(boundbox)
        if tNear > MINVALUE and tNear < tFar:
            return float(tNear)
        if tNear < tFar:
            return float(tFar)
        return MAXVALUE


class Pickup:
    """pickup.js:41-123.  ``renderer`` supplies the device context and the canvas size."""

    def __init__(self, scene, renderer):
        self.scene = scene
        self.renderer = renderer
        self.movementNormal = None
        self.movementDistance = 0.0
        self.originalHit = None

    def _ray(self, x, y):
        W, H = self.renderer.ctx.width, self.renderer.ctx.height
        inmp = np.linalg.inv(np.asarray(self.scene.mat, dtype=np.float64))
        return Ray.generate(self.scene.eye, inmp, (x / W) * 2 - 1, 1 - (y / H) * 2)

    def pick(self, x, y):
        """pickup.js:46-63.  ``(x, y)`` in canvas pixels, y down.  Sets ``scene.select``; True if something was hit."""
        ray = self._ray(x, y)
        self.scene.select = None
        ids, t, _, _ = self.renderer.ctx.intersect_rays(np.concatenate([ray.origin, ray.dir])[None, :])
        i = int(ids[0])
        if i < 0 or i >= len(self.scene.objects):
            return False
        obj = self.scene.objects[i]
        boundbox = obj.boundbox()
        if not boundbox or not ray.testBoundBox(boundbox):
            return False
        self.scene.select = obj
        return float(t[0]) < MAXVALUE

    def movingBegin(self, x, y):
        """pickup.js:65-90."""
        boundbox = self.scene.select.boundbox()
        ray = self._ray(x, y)
        t = ray.intersectBoundBox(boundbox)
        if t < MAXVALUE:
            hit = ray.origin + ray.dir * t
            mn, mx = boundbox['min'], boundbox['max']
            if abs(hit[0] - mn[0]) < MINVALUE: n = [-1, 0, 0]
            elif abs(hit[0] - mx[0]) < MINVALUE: n = [1, 0, 0]
            elif abs(hit[1] - mn[1]) < MINVALUE: n = [0, -1, 0]
            elif abs(hit[1] - mx[1]) < MINVALUE: n = [0, 1, 0]
            elif abs(hit[2] - mn[2]) < MINVALUE: n = [0, 0, -1]
            else: n = [0, 0, 1]
            self.movementNormal = np.array(n, dtype=np.float64)
            self.movementDistance = float(self.movementNormal @ hit)
            self.originalHit = hit
            self.scene.moving = True
            return True
        return False

    def _drag(self, x, y):
        ray = self._ray(x, y)
        t = (self.movementDistance - self.movementNormal @ ray.origin) / (self.movementNormal @ ray.dir)
        hit = ray.origin + ray.dir * t
        self.scene.select.temporaryTranslate(hit - self.originalHit)
        return hit

    def moving(self, x, y):
        """pickup.js:92-104."""
        self.originalHit = self._drag(x, y)

    def movingEnd(self, x, y):
        """pickup.js:106-121."""
        self._drag(x, y)
        self.scene.moving = False

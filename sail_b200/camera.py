"""Camera (mirror of /root/reference/src/scene/camera.js:6-62).  Matrices are float64 row-major
numpy arrays, as the reference's Sylvester-style ``Matrix`` holds doubles."""
import math

import numpy as np


def _unit(v):
    return v / math.sqrt(float(np.dot(v, v)))


class Camera:
    def __init__(self, eye, center, up=(0, 1, 0)):
        self.eye = np.asarray(eye, dtype=np.float64)
        self.center = np.asarray(center, dtype=np.float64)
        self.up = np.asarray(up, dtype=np.float64)
        self.makePerspective()
        self.makeLookAt()

    def makePerspective(self, fovy=55, aspect=1, znear=1, zfar=100):
        """camera.js:16-35."""
        top = znear * math.tan(fovy * math.pi / 360.0)
        bottom = -top
        left = bottom * aspect
        right = top * aspect
        X = 2 * znear / (right - left)
        Y = 2 * znear / (top - bottom)
        A = (right + left) / (right - left)
        B = (top + bottom) / (top - bottom)
        C = -(zfar + znear) / (zfar - znear)
        D = -2 * zfar * znear / (zfar - znear)
        self.projection = np.array([[X, 0, A, 0], [0, Y, B, 0], [0, 0, C, D], [0, 0, -1, 0]], dtype=np.float64)

    def makeLookAt(self):
        """camera.js:37-58.  Note the X axis is negated (mirror image)."""
        z = _unit(self.eye - self.center)
        x = _unit(np.cross(self.up, z))
        y = _unit(np.cross(z, x))
        x = x * -1
        m = np.array([[x[0], x[1], x[2], 0], [y[0], y[1], y[2], 0], [z[0], z[1], z[2], 0], [0, 0, 0, 1]], dtype=np.float64)
        t = np.array([[1, 0, 0, -self.eye[0]], [0, 1, 0, -self.eye[1]], [0, 0, 1, -self.eye[2]], [0, 0, 0, 1]], dtype=np.float64)
        self.modelview = m @ t

    def update(self):
        self.makeLookAt()

// Pickup without a DOM (src/core/pickup.js:41-123): pick() asks the device which object the pixel's ray hits first
// (native.intersectRays: the scan the trace kernel runs, so what is picked is what is drawn) instead of looping over the JS
// per-shape intersect(); objects whose boundbox() is false stay unpickable (geometry.js:51-53).  The drag arithmetic
// (movingBegin / moving / movingEnd, pickup.js:65-121) is inherited unchanged -- it reads the GLOBAL `scene` (pickup.js:66,
// SURVEY Q13), so the constructor publishes the scene there as the reference's demo page does.  Python twin: sail_b200/pickup.py.
import native from './native';
import {Pickup as ReferencePickup, Ray} from '../core/pickup';

class Pickup extends ReferencePickup {
    constructor(scene, renderer) {
        super(scene);
        this.renderer = renderer;
        globalThis.scene = scene;
    }

    // x, y: canvas pixels (y down), scaled by the renderer's size instead of the 512 of pickup.js:50-51
    pick(x, y) {
        const W = this.renderer.ctx.width, H = this.renderer.ctx.height;
        const ray = Ray.generate(this.scene.eye, this.scene.mat.inverse(), (x / W) * 2 - 1, 1 - (y / H) * 2);
        const rays = new Float32Array([...ray.origin.elements, ...ray.dir.elements]);
        const id = new Int32Array(1), t = new Float32Array(1);
        native.intersectRays(this.renderer.ctx.handle, rays, id, t);
        const hit = id[0] >= 0 ? this.scene.objects[id[0]] : null;
        this.scene.select = hit && hit.boundbox() ? hit : null;
        return this.scene.select !== null;
    }
}

export {Pickup, Ray};

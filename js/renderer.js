// Renderer on libsail_b200.so: the reference's src/core/renderer.js:8-73 with the canvas replaced by a size and
// the display pass returning pixels (Uint8ClampedArray, W*H*4, bottom row first) instead of drawing them.
import native from '../../napi/build/Release/sail_napi.node';
import {Tracer} from './tracer';

class Renderer {
    constructor(canvas, device = 0) {
        this.ctx = native.create(canvas.width, canvas.height, device);
        this.tracer = new Tracer(this.ctx);
    }

    updateObjects(scene) { this.tracer.updateObjects(scene); }

    update(scene) { this.tracer.update(scene); }

    render(scene) {
        if (scene.moving) {
            scene.sampleCount = 0;
            this.updateObjects(scene);
        }
        this.tracer.render(scene.mat, scene.eye, scene.sampleCount++);
        return native.present(this.ctx.handle, this.ctx.width, this.ctx.height);
    }

    destroy() { native.destroy(this.ctx.handle); }
}

export {Renderer};

// Renderer on libsail_b200.so: the reference's src/core/renderer.js:8-73 with the canvas replaced by a size and
// the display pass returning pixels (Uint8ClampedArray, W*H*4, bottom row first) instead of drawing them.
import native from './native';
import {Tracer} from './tracer';

class Renderer {
    // device: a CUDA ordinal, or {devices: [0, 1, ...], shard: 1 (sample index mod N) | 2 (row bands), bandRows, combine} for one
    // renderer over several GPUs (include/sail_b200.h sail_create_info.ndev; the split, the collective and the combined image
    // live behind the C ABI)
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

    // render() with the display double-buffered (sail_present_begin / _end): queues this frame's display pass and its copy to
    // the host, returns the PREVIOUS frame's pixels (null on the first call) -- the copy overlaps the next frame's tracing,
    // like the canvas reaching the screen one requestAnimationFrame later (ui/ui.js:67-70).  flush() returns the last frame.
    renderPipelined(scene) {
        if (scene.moving) {
            scene.sampleCount = 0;
            this.updateObjects(scene);
        }
        this.tracer.render(scene.mat, scene.eye, scene.sampleCount++);
        const prev = this.flush();
        if (!this.frames) {
            const bytes = this.ctx.width * this.ctx.height * 4;
            this.frames = [new Uint8ClampedArray(bytes), new Uint8ClampedArray(bytes)];
            this.slot = 1;
        }
        this.slot = 1 - this.slot;
        native.presentBegin(this.ctx.handle, this.frames[this.slot]);
        this.inflight = this.frames[this.slot];
        return prev;
    }

    flush() {
        if (!this.inflight) return null;
        native.presentEnd(this.ctx.handle);
        const done = this.inflight;
        this.inflight = null;
        return done;
    }

    destroy() { native.destroy(this.ctx.handle); }
}

export {Renderer};

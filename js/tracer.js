// Tracer on libsail_b200.so: same fields and methods as the reference's src/core/tracer.js:8-101, the WebGL
// texture/uniform plumbing replaced by calls into the N-API addon (napi/sail_napi.cc).
import native from './native';
import {Matrix, Vector} from '../utils/matrix';
import {lowerConfig} from './config';

const OBJECTS_LENGTH = 18, TEXPARAMS_LENGTH = 16;

class Tracer {
    constructor(ctx) {
        this.ctx = ctx;                 // {handle, width, height} from native.create()
        this.timeStart = new Date();
    }

    updateObjects(scene) {
        let objects = [];
        for (let object of scene.objects) objects.push(...object.gen());
        native.uploadObjects(this.ctx.handle, new Float32Array(objects));
    }

    update(scene) {
        native.configure(this.ctx.handle, lowerConfig(scene.tracerConfig(), scene.rendererConfig()));
        let objects = [], texparams = [], lights = [];
        for (let object of scene.objects) {
            objects.push(...object.gen(texparams.length / TEXPARAMS_LENGTH));
            texparams.push(...object.genTexparams());
        }
        for (let light of scene.lights) lights.push(...light.gen());
        native.uploadScene(this.ctx.handle, new Float32Array(objects), new Float32Array(texparams), new Float32Array(lights));
        this.n = objects.length / OBJECTS_LENGTH;
    }

    render(modelviewProjection, eye, sampleCount) {
        const W = this.ctx.width, H = this.ctx.height;
        let jitter = new Vector([(Math.random() * 2 - 1) / W, (Math.random() * 2 - 1) / H, 0]);
        let matrix = Matrix.Translation(jitter).multiply(modelviewProjection).inverse();
        native.render(this.ctx.handle, new Float32Array(matrix.flatten()),
                      new Float32Array([eye.elements[0], eye.elements[1], eye.elements[2]]),
                      sampleCount / (sampleCount + 1), (new Date() - this.timeStart) * 0.001);
    }
}

export {Tracer};

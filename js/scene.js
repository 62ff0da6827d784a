// Scene of the reference (src/scene/scene.js:11-119) with the one change a host without a global `scene` needs:
// Scene.update() resets ITS OWN sampleCount (the reference writes to the global variable `scene`, scene.js:65-68; SURVEY Q13).
import {Scene as ReferenceScene} from '../scene/scene';

class Scene extends ReferenceScene {
    update() {
        this.camera.update();
        this.sampleCount = 0;
    }
}

export {Scene};

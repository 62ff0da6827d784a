// The Sail namespace for Node.js: index.js:15-49 of the reference without `window` -- the same thirty names, with Renderer
// (and the Tracer behind it) on libsail_b200.so, Scene / Color carrying the two SURVEY Q13 fixes, and Control, whose handlers
// are DOM events (src/core/control.js), replaced by the DOM-free Pickup.
//   import Sail from './src/native/index.js';
//   const renderer = new Sail.Renderer({width: 1920, height: 1080});   // canvas -> size
//   renderer.update(scene); const pixels = renderer.render(scene);     // Uint8ClampedArray W*H*4, bottom row first
import {Renderer} from './renderer';
import {Scene} from './scene';
import {Cube, Sphere, Rectangle, Cone, Cylinder, Disk, Hyperboloid, Paraboloid, Cornellbox} from '../scene/geometry';
import {AreaLight, PointLight, SpotLight} from '../scene/light';
import {Camera} from '../scene/camera';
import {Matte, Mirror, Metal, Glass} from '../scene/material';
import {UniformColor, Checkerboard, Checkerboard2, Bilerp, Mix, Scale, UV} from '../scene/texture';
import {Matrix, Vector} from '../utils/matrix';
import {Color} from './color';
import {Pickup} from './pickup';

const Sail = {
    Renderer, Scene, Cube, Sphere, Rectangle, Cone, Cylinder, Disk, Hyperboloid, Paraboloid,
    AreaLight, PointLight, SpotLight, Cornellbox, Camera,
    Control: Pickup,
    Matte, Mirror, Metal, Glass,
    UniformColor, Checkerboard, Checkerboard2, Bilerp, Mix, Scale, UV,
    Color, Matrix, Vector
};

export default Sail;
export {Sail};

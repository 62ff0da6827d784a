// Color of the reference (src/core/color.js:7-30) plus the alias its README uses: Sail.Color.create (README.md:49; SURVEY Q13).
import {Color as ReferenceColor} from '../core/color';

class Color extends ReferenceColor {
    static create(color) { return ReferenceColor.createTexture(color); }
}

export {Color};

// The ONE place that knows where the N-API addon lives.  Layout: this directory is copied to <sail>/src/native/ of a checkout of
// the reference, the addon is built in <sail>/napi/ (node-gyp, napi/binding.gyp) -- hence ../../napi/...; SAIL_NAPI overrides.
// (A compiled addon cannot be `import`ed as an ES module; createRequire is the documented way to load it from one.)
import {createRequire} from 'module';

const require = createRequire(import.meta.url);
const native = require(process.env.SAIL_NAPI || '../../napi/build/Release/sail_napi.node');

export default native;

// sail_scene.h -- the packed, pre-resolved scene block the kernels read (host + device).
//
// Replaces the three R32F textures of the reference (objects 18 x n, texParams 16 x tn,
// lights 18 x ln; /root/reference/src/core/tracer.js:42-90) and the per-ray `texture()` decoding of
// every parseX() (e.g. src/shader/shape/sphere.glsl:18-27).  sail_pack_scene() (sail_pack.cpp)
// resolves, once per upload: integer categories, the material / texture row of each object
// (undoing the `/float(tn-1)` normalisation exactly as NEAREST+CLAMP sampling resolves it), the
// Cornell-box "material row 0" quirk (cornellbox.glsl:17), the light-category row quirk
// (shader.light.js:14-15) and every quantity that depends on scene constants only (bounding boxes,
// r*r, (r/h)^2, rectangle frames, reciprocals of constant divisors).  All of these are computed with
// the same FP32 operations, in the same order, as the shader would per ray, so hoisting them does
// not change a single bit of the result.
//
// One block is a few hundred bytes to a few KB; every CTA stages it into shared memory, where the
// object scan reads it with warp-uniform (broadcast) addresses.
#ifndef SAIL_SCENE_H
#define SAIL_SCENE_H
#ifndef __CUDACC_RTC__  // (run-time compilation, sail_jit.cpp: no host headers)
#include <stddef.h>
#include <stdint.h>
#endif

#define SAIL_MAX_OBJECTS 64
#define SAIL_MAX_LIGHTS 32

// const/define.glsl:18-44
enum { SAIL_CUBE = 1, SAIL_SPHERE = 2, SAIL_RECTANGLE = 3, SAIL_CONE = 4, SAIL_CYLINDER = 5, SAIL_DISK = 6,
       SAIL_HYPERBOLOID = 7, SAIL_PARABOLOID = 8, SAIL_CORNELLBOX = 9 };
enum { SAIL_AREA = 0, SAIL_POINT = 1, SAIL_SPOT = 2 };
enum { SAIL_MATTE = 1, SAIL_MIRROR = 2, SAIL_METAL = 3, SAIL_GLASS = 4 };
// Feature bits for kernel specialisation (see sail_shapes.cuh).
enum { SAIL_FEAT_QUADRICS = 1, SAIL_FEAT_GLOSSY = 2, SAIL_FEAT_TEXTURED = 4, SAIL_FEAT_NEE = 8 };
enum { SAIL_UNIFORM_COLOR = 0, SAIL_CHECKERBOARD = 5, SAIL_CHECKERBOARD2 = 7, SAIL_BILERP = 8, SAIL_MIXF = 9,
       SAIL_SCALE = 10, SAIL_UVF = 11 };

// Geometry of one object: what the nearest-hit scan touches.  48 B header + 160 B parameters = 208 B.
struct __attribute__((aligned(16))) SailShape {
    int category;
    float rev;        // reverseNormal ? -1 : +1   (Cornell box: +1, cornellbox.glsl:18)
    // The two corners, interleaved per axis {lo.x, hi.x, lo.y, hi.y, lo.z, hi.z} at an 8-byte aligned offset: the
    // slab test loads each axis as one 64-bit register pair for the packed FP32 pipe (FADD2 / FMUL2).
    // lo = cube/cornellbox/rectangle min or a bounding-box corner of the quadrics; hi = the max corner.
    float lohi[6];
    float emission[3];
    int matCategory;  // texParams[matRow][0]
    float q[40];      // shape parameters, layout per category (SQ_* below)
};
#ifndef __CUDACC_RTC__
static_assert(sizeof(SailShape) == 208 && offsetof(SailShape, lohi) == 8, "SailShape layout");
#endif
// q[] layouts
// cube / cornellbox: per face (x-, x+, y-, y+, z-, z+) the geometric normal intersectX() returns and its dpdu
enum { SQ_BOX_FACE = 0, SQ_BOX_FACE_STRIDE = 6 };
enum { SQ_SPH_C = 0, SQ_SPH_R = 3, SQ_SPH_RR = 4, SQ_SPH_INVR = 5, SQ_SPH_PDF = 6 };
enum { SQ_RECT_DPDU = 0, SQ_RECT_N = 3, SQ_RECT_SS = 6, SQ_RECT_TS = 9, SQ_RECT_MAXX = 12, SQ_RECT_MAXY = 13,
       SQ_RECT_INVMAXX = 14, SQ_RECT_INVMAXY = 15, SQ_RECT_DPDV = 16, SQ_RECT_PDF = 19 };
enum { SQ_CONE_P = 0, SQ_CONE_H = 3, SQ_CONE_K = 4, SQ_CONE_INVH = 5, SQ_CONE_R = 6, SQ_CONE_TANA = 7, SQ_CONE_INVTANA = 8 };
enum { SQ_CYL_P = 0, SQ_CYL_H = 3, SQ_CYL_RR = 4, SQ_CYL_INVH = 5 };
enum { SQ_DISK_P = 0, SQ_DISK_R = 3, SQ_DISK_INNER = 4, SQ_DISK_RR = 5, SQ_DISK_IRR = 6, SQ_DISK_INVDR = 7,
       SQ_DISK_DIFF = 8, SQ_DISK_PDF = 9 };
enum { SQ_HYP_P = 0, SQ_HYP_P1 = 3, SQ_HYP_P2 = 6, SQ_HYP_AH = 9, SQ_HYP_CH = 10, SQ_HYP_ZMIN = 11, SQ_HYP_ZMAX = 12,
       SQ_HYP_INVDZ = 13 };
enum { SQ_PAR_P = 0, SQ_PAR_ZMIN = 3, SQ_PAR_ZMAX = 4, SQ_PAR_K = 5, SQ_PAR_DZ = 6, SQ_PAR_INVDZ = 7 };

// Material + texture of one object (read only for the winning hit).  128 B.
struct __attribute__((aligned(16))) SailSurface {
    int matCategory;
    int texCategory;
    float m[14];  // material row columns 1..14 (texParams layout) + derived (SM_*)
    float t[16];  // texture row columns 1..13 + derived (ST_*)
};
enum { SM_MATTE_KD = 0, SM_MATTE_SIGMA = 1, SM_MATTE_A = 2, SM_MATTE_B = 3 };
enum { SM_MIRROR_KR = 0 };
enum { SM_METAL_UR = 0, SM_METAL_VR = 1, SM_METAL_ETA = 2, SM_METAL_K = 5 };
enum { SM_GLASS_KR = 0, SM_GLASS_KT = 1, SM_GLASS_ETA = 2, SM_GLASS_UR = 3, SM_GLASS_VR = 4 };
enum { ST_RGB = 0, ST_CHECK_SIZE = 0, ST_CHECK_LINE = 1, ST_C1 = 0, ST_C2 = 3, ST_CHECK2_SIZE = 6, ST_MIX_AMOUNT = 6,
       ST_B00 = 0, ST_B01 = 3, ST_B10 = 6, ST_B11 = 9 };

// One entry per value of the light index drawn by light_sample() (0..ln, the last one only reachable
// if fract() rounds up to 1.0).  64 B.
struct __attribute__((aligned(16))) SailLight {
    int category;   // from the quirky category row (index==0 ? row 0 : last row)
    int object;     // area: resolved row of the geometry in objects
    float emission[3];
    float from[3];
    float cosTotalWidth, cosFalloffStart, invFalloffRange;
    float pad[5];
};

struct __attribute__((aligned(16))) SailSceneBlock {
    int n, ln;      // object count, light count (uniforms n, ln; src/core/shader.js:41-43)
    int features;   // SAIL_F_* bits the scene content needs (kernel variant selection)
    unsigned emissiveMask;  // bit i: object i (i < 32) has a non-zero emission; objects >= 32 count as emissive
    SailShape shape[SAIL_MAX_OBJECTS];
    SailSurface surf[SAIL_MAX_OBJECTS];
    SailLight light[SAIL_MAX_LIGHTS + 1];
};

// Byte offsets of the three arrays inside the block (the staging copy works in 16-byte units).
#define SAIL_SCENE_SHAPE_OFF 16
#define SAIL_SCENE_SURF_OFF (SAIL_SCENE_SHAPE_OFF + SAIL_MAX_OBJECTS * (int)sizeof(SailShape))
#define SAIL_SCENE_LIGHT_OFF (SAIL_SCENE_SURF_OFF + SAIL_MAX_OBJECTS * (int)sizeof(SailSurface))

#ifndef __CUDACC_RTC__
static_assert(offsetof(SailSceneBlock, shape) == SAIL_SCENE_SHAPE_OFF && offsetof(SailSceneBlock, surf) == SAIL_SCENE_SURF_OFF &&
              offsetof(SailSceneBlock, light) == SAIL_SCENE_LIGHT_OFF, "SailSceneBlock layout");
// Host packer (sail_pack.cpp).  Returns 0 or a negative sail error code; `err` receives a message.
int sail_pack_scene(const float* objects, int n, const float* texparams, int tn, const float* lights, int ln,
                    SailSceneBlock* out, char* err, int errlen);
#endif

#endif

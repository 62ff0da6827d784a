// sail_pack.cpp -- host side of scene upload: the reference wire format -> SailSceneBlock.
//
// Replaces Tracer.update()'s texture uploads (/root/reference/src/core/tracer.js:42-90) plus the
// per-ray parseX()/readFloat() decoding in the shaders (src/shader/util/texhelper.glsl:1-36 and every
// shape/material/texture/light parse function).  Table addressing follows what NEAREST +
// CLAMP_TO_EDGE sampling of an R32F texture resolves to (src/core/webgl.js:146-158): a column is
// addressed as c/17 (objects, lights) or c/15 (texParams) of an 18/16 texel wide texture, a row as
// i/(rows-1); NaN coordinates (rows == 1) resolve to texel 0.
//
// Build with -ffp-contract=off: every derived constant must equal what the FP32 shader arithmetic
// (oracle/glsl.h model: a/b = a*rcp(b), correctly rounded sqrt) would compute per ray.
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "detmath.h"
#include "sail_scene.h"

namespace {

const float kObjectsLen = 17.0f, kLightsLen = 17.0f, kTexLen = 15.0f;  // const/define.glsl:1-3
const float kPI = 3.141592653589793f, kINVPI = 0.3183098861837907f;

struct Table {
    const float* d;
    int width, rows;
};

int texel(float coord, int size) {
    if (!(coord == coord)) return 0;
    float f = floorf(coord * (float)size);
    if (f < 0.0f) return 0;
    if (f > (float)(size - 1)) return size - 1;
    return (int)f;
}
float rowCoord(float v, int count) { return dm_div(v, (float)(count - 1)); }
int toInt(float a) {
    if (!(a > -2147483648.0f && a < 2147483648.0f)) return 0;
    return (int)a;
}
// readFloat(tex, vec2(c + k, row), glslWidth): texhelper.glsl:1-36 (readVec3 steps x by 1/width).
float fetch(const Table& t, float c, int k, float glslWidth, float row) {
    if (t.width <= 0 || t.rows <= 0) return 0.0f;
    float x = c / glslWidth;
    for (int i = 0; i < k; i++) x = x + 1.0f / glslWidth;
    return t.d[texel(row, t.rows) * t.width + texel(x, t.width)];
}
void fetch3(const Table& t, float c, float glslWidth, float row, float* out) {
    for (int k = 0; k < 3; k++) out[k] = fetch(t, c, k, glslWidth, row);
}

struct H3 {
    float x, y, z;
};
float hdot(H3 a, H3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
H3 hcross(H3 x, H3 y) { return H3{x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
float hlength(H3 a) { return dm_sqrt(hdot(a, a)); }
H3 hscale(H3 a, float s) { return H3{a.x * s, a.y * s, a.z * s}; }
H3 hnormalize(H3 a) { return hscale(a, dm_rcp(dm_sqrt(hdot(a, a)))); }
void st3(float* p, H3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

int fail(char* err, int errlen, const char* msg, int i) {
    if (err && errlen > 0) snprintf(err, errlen, msg, i);
    return -2;  // SAIL_ERR_INVALID
}

}  // namespace

int sail_pack_scene(const float* objects, int n, const float* texparams, int tn, const float* lights, int ln,
                    SailSceneBlock* out, char* err, int errlen) {
    if (n < 0 || tn < 0 || ln < 0) return fail(err, errlen, "negative table size (%d)", n < 0 ? n : (tn < 0 ? tn : ln));
    if (n > SAIL_MAX_OBJECTS) return fail(err, errlen, "too many objects (%d)", n);
    if (ln > SAIL_MAX_LIGHTS) return fail(err, errlen, "too many lights (%d)", ln);
    if ((n && !objects) || (tn && !texparams) || (ln && !lights)) return fail(err, errlen, "null table pointer (%d)", 0);
    memset(out, 0, sizeof(*out));
    out->n = n;
    out->ln = ln;
    Table O{objects, 18, n}, T{texparams, 16, tn}, L{lights, 18, ln};

    for (int i = 0; i < n; i++) {
        SailShape& s = out->shape[i];
        SailSurface& u = out->surf[i];
        float row = rowCoord((float)i, n);  // float(i)/float(n-1), shader.shape.js:34
        int cat = toInt(fetch(O, 0.0f, 0, 1.0f, row));  // texture(objects, vec2(0.0, row)).r
        s.category = cat;
        float colRev, colMat, colTex, colEm;
        switch (cat) {
            case SAIL_CUBE: case SAIL_RECTANGLE: case SAIL_PARABOLOID: colRev = 7; colMat = 8; colTex = 9; colEm = 10; break;
            case SAIL_SPHERE: colRev = 5; colMat = 6; colTex = 7; colEm = 8; break;
            case SAIL_CONE: case SAIL_CYLINDER: case SAIL_DISK: colRev = 6; colMat = 7; colTex = 8; colEm = 9; break;
            case SAIL_HYPERBOLOID: colRev = 12; colMat = 13; colTex = 14; colEm = 15; break;
            case SAIL_CORNELLBOX: colRev = -1; colMat = 7; colTex = -1; colEm = -1; break;  // cornellbox.glsl:13-21
            default: return fail(err, errlen, "object %d: unknown shape category", i);
        }
        bool reverse = (colRev >= 0) && toInt(fetch(O, colRev, 0, kObjectsLen, row)) == 1;
        s.rev = reverse ? -1.0f : 1.0f;
        if (colEm >= 0) fetch3(O, colEm, kObjectsLen, row, s.emission);
        float matIndex = rowCoord(fetch(O, colMat, 0, kObjectsLen, row), tn);
        // ---- material row (matte.glsl:1-6, mirror.glsl:1-3, metal.glsl:1-6, glass.glsl:1-8) ----
        u.matCategory = s.matCategory = toInt(fetch(T, 0.0f, 0, kTexLen, matIndex));
        for (int k = 0; k < 14; k++) u.m[k] = fetch(T, (float)(k + 1), 0, kTexLen, matIndex);
        // ---- texture row (shader.texture.js:22-29) ----
        u.texCategory = SAIL_UNIFORM_COLOR;
        if (colTex >= 0) {
            float texIndex = rowCoord(fetch(O, colTex, 0, kObjectsLen, row), tn);
            int tc = toInt(fetch(T, 0.0f, 0, kTexLen, texIndex));
            float c[13];
            for (int k = 0; k < 13; k++) c[k] = fetch(T, (float)(k + 1), 0, kTexLen, texIndex);
            // readVec3 walks x by 1/15 from the first column; same texels as the direct fetches above
            // for columns <= 13 (checked in tests/test_pack.py).
            memcpy(u.t, c, sizeof(c));
            u.texCategory = tc;
            switch (tc) {
                case SAIL_UNIFORM_COLOR: break;
                case SAIL_CHECKERBOARD:  // checkerboard.glsl:6-12: width = 0.5*lineWidth/size
                    u.t[13] = dm_rcp(c[ST_CHECK_SIZE]);
                    u.t[14] = dm_div(0.5f * c[ST_CHECK_LINE], c[ST_CHECK_SIZE]);
                    break;
                case SAIL_CHECKERBOARD2:  // checkerboard2.glsl:7-16
                    u.t[13] = dm_rcp(c[ST_CHECK2_SIZE]);
                    break;
                case SAIL_BILERP: case SAIL_UVF: break;
                case SAIL_MIXF: {  // mixf.glsl:7-13: (1-amount)*color1 + amount*color2 is scene constant
                    float a = c[ST_MIX_AMOUNT];
                    for (int k = 0; k < 3; k++) u.t[ST_RGB + k] = (1.0f - a) * c[ST_C1 + k] + a * c[ST_C2 + k];
                    break;
                }
                case SAIL_SCALE:  // scale.glsl:6-11
                    for (int k = 0; k < 3; k++) u.t[ST_RGB + k] = c[ST_C1 + k] * c[ST_C2 + k];
                    break;
                default:  // no arm in getSurfaceColor: BLACK
                    u.texCategory = SAIL_UNIFORM_COLOR;
                    u.t[0] = u.t[1] = u.t[2] = 0.0f;
            }
        }
        // ---- geometry ----
        float* q = s.q;
        float lo[3] = {0.0f, 0.0f, 0.0f}, hi[3] = {0.0f, 0.0f, 0.0f};
        switch (cat) {
            case SAIL_CUBE: case SAIL_CORNELLBOX: {
                fetch3(O, 1, kObjectsLen, row, lo);
                fetch3(O, 4, kObjectsLen, row, hi);
                // Per-face normal (cube.glsl:25-38 `c*vec3(..)`; cornellbox.glsl:39-51,83 `-vec3(..)`) and dpdu
                // (cube.glsl:40-48): scene constants, evaluated here with the shader's own operations.
                const float c = (cat == SAIL_CUBE) ? s.rev : 1.0f;
                const float axis[6][3] = {{-1, 0, 0}, {1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {0, 0, -1}, {0, 0, 1}};
                for (int f = 0; f < 6; f++) {
                    H3 nrm{c * axis[f][0], c * axis[f][1], c * axis[f][2]};
                    if (cat == SAIL_CORNELLBOX) nrm = H3{-nrm.x, -nrm.y, -nrm.z};
                    H3 dpdu = (dm_abs(nrm.x) < 0.5f) ? hcross(nrm, H3{1.0f, 0.0f, 0.0f}) : hcross(nrm, H3{0.0f, 1.0f, 0.0f});
                    st3(q + SQ_BOX_FACE + f * SQ_BOX_FACE_STRIDE, nrm);
                    st3(q + SQ_BOX_FACE + f * SQ_BOX_FACE_STRIDE + 3, dpdu);
                }
                break;
            }
            case SAIL_SPHERE: {  // sphere.glsl:10-16,45-92
                float c[3];
                fetch3(O, 1, kObjectsLen, row, c);
                float r = fetch(O, 4, 0, kObjectsLen, row);
                for (int k = 0; k < 3; k++) { lo[k] = c[k] - r; hi[k] = c[k] + r; q[SQ_SPH_C + k] = c[k]; }
                q[SQ_SPH_R] = r;
                q[SQ_SPH_RR] = r * r;
                q[SQ_SPH_INVR] = dm_rcp(r);
                q[SQ_SPH_PDF] = dm_div(kINVPI, r * r);
                break;
            }
            case SAIL_RECTANGLE: {  // rectangle.glsl:32-70
                float mn[3], mx[3];
                fetch3(O, 1, kObjectsLen, row, mn);
                fetch3(O, 4, kObjectsLen, row, mx);
                memcpy(lo, mn, sizeof(mn));
                memcpy(hi, mx, sizeof(mx));
                H3 dpdu{mx[0] - mn[0], 0.0f, 0.0f}, dpdv{0.0f, mx[1] - mn[1], mx[2] - mn[2]};
                H3 nrm = hnormalize(hcross(dpdu, dpdv));
                float maxX = hlength(dpdu), maxY = hlength(dpdv);
                H3 ss = hscale(dpdu, dm_rcp(maxX));
                H3 ts = hcross(nrm, ss);
                st3(q + SQ_RECT_DPDU, dpdu); st3(q + SQ_RECT_N, nrm); st3(q + SQ_RECT_SS, ss); st3(q + SQ_RECT_TS, ts);
                st3(q + SQ_RECT_DPDV, dpdv);
                q[SQ_RECT_MAXX] = maxX; q[SQ_RECT_MAXY] = maxY;
                q[SQ_RECT_INVMAXX] = dm_rcp(maxX); q[SQ_RECT_INVMAXY] = dm_rcp(maxY);
                q[SQ_RECT_PDF] = dm_rcp(maxX * maxY);
                break;
            }
            case SAIL_CONE: case SAIL_CYLINDER: {  // cone.glsl:11-17,48-99; cylinder.glsl:11-17,40-90
                float p[3];
                fetch3(O, 1, kObjectsLen, row, p);
                float h = fetch(O, 4, 0, kObjectsLen, row), r = fetch(O, 5, 0, kObjectsLen, row);
                lo[0] = p[0] - r; lo[1] = p[1] - 0.0f; lo[2] = p[2] - r;
                hi[0] = p[0] + r; hi[1] = p[1] + h; hi[2] = p[2] + r;
                for (int k = 0; k < 3; k++) q[k] = p[k];
                q[3] = h;
                if (cat == SAIL_CONE) {
                    float k = dm_div(r, h);
                    q[SQ_CONE_TANA] = k;
                    q[SQ_CONE_INVTANA] = dm_rcp(k);
                    q[SQ_CONE_K] = k * k;
                    q[SQ_CONE_INVH] = dm_rcp(h);
                    q[SQ_CONE_R] = r;
                } else {
                    q[SQ_CYL_RR] = r * r;
                    q[SQ_CYL_INVH] = dm_rcp(h);
                }
                break;
            }
            case SAIL_DISK: {  // disk.glsl:36-83
                fetch3(O, 1, kObjectsLen, row, q + SQ_DISK_P);
                float r = fetch(O, 4, 0, kObjectsLen, row), inner = fetch(O, 5, 0, kObjectsLen, row);
                q[SQ_DISK_R] = r; q[SQ_DISK_INNER] = inner;
                q[SQ_DISK_RR] = r * r; q[SQ_DISK_IRR] = inner * inner;
                q[SQ_DISK_INVDR] = dm_rcp(r - inner);
                q[SQ_DISK_DIFF] = inner - r;
                float area = 2.0f * kPI * 0.5f * (r * r - inner * inner);
                q[SQ_DISK_PDF] = dm_rcp(area);
                break;
            }
            case SAIL_HYPERBOLOID: {  // hyperboloid.glsl:13-24,60-111
                float p[3], p1[3], p2[3];
                fetch3(O, 1, kObjectsLen, row, p);
                fetch3(O, 4, kObjectsLen, row, p1);
                fetch3(O, 7, kObjectsLen, row, p2);
                float r1 = dm_sqrt(p1[0] * p1[0] + p1[1] * p1[1]), r2 = dm_sqrt(p2[0] * p2[0] + p2[1] * p2[1]);
                float rMax = dm_max(r1, r2), zMin = dm_min(p1[2], p2[2]), zMax = dm_max(p1[2], p2[2]);
                lo[0] = p[0] - rMax; lo[1] = p[1] - (-zMin); lo[2] = p[2] - rMax;
                hi[0] = p[0] + rMax; hi[1] = p[1] + zMax; hi[2] = p[2] + rMax;
                for (int k = 0; k < 3; k++) { q[SQ_HYP_P + k] = p[k]; q[SQ_HYP_P1 + k] = p1[k]; q[SQ_HYP_P2 + k] = p2[k]; }
                q[SQ_HYP_AH] = fetch(O, 10, 0, kObjectsLen, row);
                q[SQ_HYP_CH] = fetch(O, 11, 0, kObjectsLen, row);
                q[SQ_HYP_ZMIN] = zMin; q[SQ_HYP_ZMAX] = zMax;
                q[SQ_HYP_INVDZ] = dm_rcp(p2[2] - p1[2]);
                break;
            }
            case SAIL_PARABOLOID: {  // paraboloid.glsl:12-20,51-103
                float p[3];
                fetch3(O, 1, kObjectsLen, row, p);
                float z0 = fetch(O, 4, 0, kObjectsLen, row), z1 = fetch(O, 5, 0, kObjectsLen, row);
                float r = fetch(O, 6, 0, kObjectsLen, row);
                float zMin = dm_min(z0, z1), zMax = dm_max(z0, z1);
                lo[0] = p[0] - r; lo[1] = p[1] - (-zMin); lo[2] = p[2] - r;
                hi[0] = p[0] + r; hi[1] = p[1] + zMax; hi[2] = p[2] + r;
                for (int k = 0; k < 3; k++) q[SQ_PAR_P + k] = p[k];
                q[SQ_PAR_ZMIN] = zMin; q[SQ_PAR_ZMAX] = zMax;
                q[SQ_PAR_K] = dm_div(zMax, r * r);
                q[SQ_PAR_DZ] = zMax - zMin;
                q[SQ_PAR_INVDZ] = dm_rcp(zMax - zMin);
                break;
            }
        }
        for (int k = 0; k < 3; k++) { s.lohi[2 * k] = lo[k]; s.lohi[2 * k + 1] = hi[k]; }
    }

    // ---- kernel features actually needed by this scene ----
    int feat = (ln > 0) ? SAIL_FEAT_NEE : 0;
    for (int i = 0; i < n; i++) {
        int c = out->shape[i].category, m = out->surf[i].matCategory, tc = out->surf[i].texCategory;
        if (c == SAIL_CONE || c == SAIL_CYLINDER || c == SAIL_DISK || c == SAIL_HYPERBOLOID || c == SAIL_PARABOLOID) feat |= SAIL_FEAT_QUADRICS;
        if (m == SAIL_METAL || m == SAIL_GLASS) feat |= SAIL_FEAT_GLOSSY;
        if (tc == SAIL_CHECKERBOARD || tc == SAIL_CHECKERBOARD2 || tc == SAIL_BILERP || tc == SAIL_UVF) feat |= SAIL_FEAT_TEXTURED;
    }
    out->features = feat;
    unsigned em = 0;
    for (int i = 0; i < n && i < 32; i++) {
        const float* e = out->shape[i].emission;
        if (!(e[0] == 0.0f && e[1] == 0.0f && e[2] == 0.0f)) em |= 1u << i;  // NaN counts as emissive
    }
    out->emissiveMask = em;

    // ---- lights: one resolved entry per drawable index (shader.light.js:12-22) ----
    for (int index = 0; ln > 0 && index <= ln; index++) {
        SailLight& l = out->light[index];
        // category: readInt(lights, vec2(0.0, index), TEX_PARAMS_LENGTH) -- raw int as a row coordinate (Q4)
        l.category = toInt(fetch(L, 0.0f, 0, kTexLen, (float)index));
        float row = rowCoord((float)index, ln);
        l.object = 0;
        switch (l.category) {
            case SAIL_AREA: {  // area.glsl:6-11
                int obj = toInt(fetch(L, 1, 0, kLightsLen, row));
                fetch3(L, 2, kLightsLen, row, l.emission);
                l.object = (n > 0) ? texel(rowCoord((float)obj, n), n) : 0;
                break;
            }
            case SAIL_POINT:  // point.glsl:6-11
                fetch3(L, 1, kLightsLen, row, l.from);
                fetch3(L, 4, kLightsLen, row, l.emission);
                break;
            case SAIL_SPOT:  // spot.glsl:8-15
                l.cosTotalWidth = fetch(L, 1, 0, kLightsLen, row);
                l.cosFalloffStart = fetch(L, 2, 0, kLightsLen, row);
                fetch3(L, 3, kLightsLen, row, l.from);
                fetch3(L, 6, kLightsLen, row, l.emission);
                l.invFalloffRange = dm_rcp(l.cosFalloffStart - l.cosTotalWidth);
                break;
            default: break;  // no arm: light_sample returns BLACK
        }
    }
    return 0;
}

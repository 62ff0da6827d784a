// sail_present.cuh -- the display pass: RenderShader's full-screen quad
// (/root/reference/src/core/renderer.js:63, src/shader/main/vsrender.glsl:4-7, main/fsrender.glsl:4-6)
// with the filter plugins of src/shader/filter/: color.glsl, gamma.glsl, tonemapping.glsl, normal.glsl,
// position.glsl, window.glsl (box / gaussian / mitchell / sinc / triangle share it, the 16 weights come
// from the host) and wavelet.glsl.
//
// The canvas becomes a headless RGBA buffer.  colorMap / normalMap / positionMap are the FP32 device
// buffers of the trace pass, sampled like the reference's LINEAR + REPEAT textures
// (src/core/webgl.js:153-156) with 8-bit bilinear weights; 512 is generalised to W / H and 1280 to
// 2.5 W / 2.5 H.
#ifndef SAIL_PRESENT_CUH
#define SAIL_PRESENT_CUH
#include "sail_vec.cuh"

struct V4 { float x, y, z, w; };
SD V4 mk4(float x, float y, float z, float w) { V4 v; v.x = x; v.y = y; v.z = z; v.w = w; return v; }
SD V4 operator+(V4 a, V4 b) { return mk4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
SD V4 operator-(V4 a, V4 b) { return mk4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
SD V4 operator*(V4 a, float s) { return mk4(a.x * s, a.y * s, a.z * s, a.w * s); }
SD float dot4(V4 a, V4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }

struct PresentParams {
    int W, H;
    int filter_id;
    float gamma_c;
    float radius[2];
    float weights[16];
    const float* color;
    const float4* color4;     // the colour map padded to 16 bytes per texel (sail_pad_rgb_kernel): what the window filters and the wavelet fetch
    const float4* position4;  // the position map likewise (wavelet only)
    const float* normal;
    const float* position;
    float* out_f32;        // W*H*4 or NULL
    unsigned char* out_u8; // W*H*4 or NULL
};

SD int wrapIndex(int i, int n) {
    int m = i % n;
    return m < 0 ? m + n : m;
}
SD V4 texelRGB(const float* img, int W, int H, int x, int y) {
    const float* p = img + 3 * ((size_t)wrapIndex(y, H) * W + wrapIndex(x, W));
    return mk4(__ldg(p), __ldg(p + 1), __ldg(p + 2), 1.0f);
}
SD void linearCoord(float u, int size, int& i0, float& t) {
    float f = u * (float)size - 0.5f;
    if (!(f == f) || !(f > -1e9f && f < 1e9f)) { i0 = 0; t = 0.0f; return; }
    float fl = dm_floor(f);
    i0 = (int)fl;
    t = dm_floor((f - fl) * 256.0f + 0.5f) * (1.0f / 256.0f);
}
SD V4 textureLinear(const float* img, int W, int H, float u, float v) {
    int x0, y0;
    float tx, ty;
    linearCoord(u, W, x0, tx);
    linearCoord(v, H, y0, ty);
    V4 c00 = texelRGB(img, W, H, x0, y0), c10 = texelRGB(img, W, H, x0 + 1, y0);
    V4 c01 = texelRGB(img, W, H, x0, y0 + 1), c11 = texelRGB(img, W, H, x0 + 1, y0 + 1);
    V4 a = c00 * (1.0f - tx) + c10 * tx;
    V4 b = c01 * (1.0f - tx) + c11 * tx;
    return a * (1.0f - ty) + b * ty;
}

// window.glsl:1-8 windowSampler + the LINEAR fetch, split per axis.  The 64 taps of a window filter are the products of 8
// x coordinates (tx +- 4 offsets) and 8 y coordinates: the bounds test, the texel pair, its wrap-around and the 8-bit blend
// weight of a coordinate are computed once per axis value instead of once per tap (16 instead of 128 coordinate set-ups, 32
// instead of 512 integer modulos per pixel).  Every tap still evaluates exactly the expressions of textureLinear().
struct AxisTap {
    int i0, i1;  // wrapped texel indices (rowTap: offsets of the two rows' first texels)
    float t;     // blend weight in 1/256 steps; 1 - t is formed at the use (three registers per axis value, not five)
};
SD bool axisInside(float c) { return !(c < 0.0f || c > 1.0f); }  // windowSampler's test, window.glsl:2-3
SD AxisTap axisTap(float c, int size) {
    AxisTap a;
    int i0;
    linearCoord(c, size, i0, a.t);
    a.i0 = wrapIndex(i0, size);
    a.i1 = wrapIndex(i0 + 1, size);
    return a;
}
// Texel index = row * W + column as 32-bit arithmetic on a row offset prepared once per y value (rowTap): one add and one
// widening multiply-add per load instead of a 64-bit multiply per texel (a quarter of the kernel's instructions were address
// arithmetic).  W * H < 2^31 texels.
SD V3 texel3(const float4* img, unsigned idx) {
    const float4 c = __ldg(img + idx);
    return mk3(c.x, c.y, c.z);
}
SD AxisTap rowTap(AxisTap y, int W) {  // i0 / i1 become offsets of the rows' first texels
    y.i0 *= W;
    y.i1 *= W;
    return y;
}
SD V3 windowTap(const PresentParams& P, const AxisTap& X, const AxisTap& Yr, bool ok, int& count) {
    if (!ok) return splat3(0.0f);
    count++;
    const V3 c00 = texel3(P.color4, (unsigned)(Yr.i0 + X.i0)), c10 = texel3(P.color4, (unsigned)(Yr.i0 + X.i1));
    const V3 c01 = texel3(P.color4, (unsigned)(Yr.i1 + X.i0)), c11 = texel3(P.color4, (unsigned)(Yr.i1 + X.i1));
    // textureLinear(): a = c00*(1-tx) + c10*tx, b = c01*(1-tx) + c11*tx, a*(1-ty) + b*ty, per channel, plain scalar operations
    // (the red / green pair through the packed FP32 pipe was measured: 10 % fewer instructions, same 0.345 ms)
    const float xo = 1.0f - X.t, yo = 1.0f - Yr.t;
    const V3 a = mk3(c00.x * xo + c10.x * X.t, c00.y * xo + c10.y * X.t, c00.z * xo + c10.z * X.t);
    const V3 b = mk3(c01.x * xo + c11.x * X.t, c01.y * xo + c11.y * X.t, c01.z * xo + c11.z * X.t);
    return mk3(a.x * yo + b.x * Yr.t, a.y * yo + b.y * Yr.t, a.z * yo + b.z * Yr.t);
}
// window.glsl:26-44 (table lookup i*j+j as written)
SD V4 filterWindow(const PresentParams& P, float tx, float ty) {
    V3 color = splat3(0.0f);
    float weightSum = 0.0f;
    const float rW = dm_rcp((float)P.W), rH = dm_rcp((float)P.H);
    AxisTap Xp[4], Xm[4], Yp[4], Ym[4];
    unsigned inside = 0;  // bit k: Xp[k], 4 + k: Xm[k], 8 + k: Yp[k], 12 + k: Ym[k] inside [0, 1]
#pragma unroll
    for (int k = 0; k < 4; k++) {
        // window.glsl:10-24: x = vec2(i/W, 0), y = vec2(0, j/H); the shader's `+ 0.0` / `- 0.0` terms leave the sums unchanged
        const float xx = fdiv(((float)k + 0.5f) * P.radius[0], 4.0f) * rW;
        const float yy = fdiv(((float)k + 0.5f) * P.radius[1], 4.0f) * rH;
        const float xp = tx + xx + 0.0f, xm = tx - xx + 0.0f, yp = ty + 0.0f + yy, ym = ty + 0.0f - yy;
        Xp[k] = axisTap(xp, P.W);
        Xm[k] = axisTap(xm, P.W);
        Yp[k] = rowTap(axisTap(yp, P.H), P.W);
        Ym[k] = rowTap(axisTap(ym, P.H), P.W);
        inside |= (axisInside(xp) ? 1u : 0u) << k | (axisInside(xm) ? 1u : 0u) << (4 + k) | (axisInside(yp) ? 1u : 0u) << (8 + k) |
                  (axisInside(ym) ? 1u : 0u) << (12 + k);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int count = 0;
            V3 tmp = splat3(0.0f);
            const bool xpIn = (inside >> j) & 1u, xmIn = (inside >> (4 + j)) & 1u, ypIn = (inside >> (8 + i)) & 1u, ymIn = (inside >> (12 + i)) & 1u;
            tmp = tmp + windowTap(P, Xp[j], Yp[i], xpIn && ypIn, count);
            tmp = tmp + windowTap(P, Xp[j], Ym[i], xpIn && ymIn, count);
            tmp = tmp + windowTap(P, Xm[j], Yp[i], xmIn && ypIn, count);
            tmp = tmp + windowTap(P, Xm[j], Ym[i], xmIn && ymIn, count);
            float weight = P.weights[i * j + j];
            weightSum += weight * (float)count;
            color = color + tmp * weight;
        }
    }
    V3 c = vdivs(color, weightSum);
    return mk4(c.x, c.y, c.z, 1.0f);
}
// textureLinear() on prepared axis values (AxisTap: texel pair, wrap-around and 8-bit blend weight of one coordinate): the same
// V4 expressions on the same texels; the colour map comes from the padded copy, the position map from its 12-byte layout.
SD V4 bilinearOf(V4 c00, V4 c10, V4 c01, V4 c11, const AxisTap& X, const AxisTap& Y) {
    const float xo = 1.0f - X.t, yo = 1.0f - Y.t;
    V4 a = c00 * xo + c10 * X.t;
    V4 b = c01 * xo + c11 * X.t;
    return a * yo + b * Y.t;
}
SD V4 texel4(const float4* img, unsigned idx) {
    const float4 q = __ldg(img + idx);
    return mk4(q.x, q.y, q.z, 1.0f);
}
SD V4 bilinearAt(const float4* img, const AxisTap& X, const AxisTap& Yr) {  // Yr = rowTap(...)
    return bilinearOf(texel4(img, (unsigned)(Yr.i0 + X.i0)), texel4(img, (unsigned)(Yr.i0 + X.i1)), texel4(img, (unsigned)(Yr.i1 + X.i0)),
                      texel4(img, (unsigned)(Yr.i1 + X.i1)), X, Yr);
}
// wavelet.glsl:1-54.  The 5 x 5 tap positions do not depend on the level: their coordinates are set up once per axis value (10
// per pixel; the literal loop sets up 45 taps x 2 maps x 2 axes, each with two integer modulos), both maps are fetched from
// 16-byte-per-texel copies, and what levels 1 and 2 evaluate again for the taps they share with level 0 -- the fetched colour,
// the colour and position weights, the squared colour distance -- is kept from level 0 in shared memory (6 floats for each of the
// 13 taps with an even distance to the centre tap; level 2's 7 taps are among them).  Taps, weights, every expression and the
// order of the sums are those of the shader: 25 + 13 + 7 accumulations, level by level.
#define SAIL_WAVELET_CACHED_TAPS 13
#define SAIL_WAVELET_SMEM_BYTES (SAIL_WAVELET_CACHED_TAPS * 6 * 256 * sizeof(float))
#define SAIL_WAVELET_H(q) ((q) % 5 == 0 ? 0.375f : (((q) % 5 == 1 || (q) % 5 == 4) ? 0.25f : 0.0625f))  // h[q % 5], wavelet.glsl:23
SD V4 filterWavelet(const PresentParams& P, float tx, float ty, float* cache) {
    V4 color = mk4(0.0f, 0.0f, 0.0f, 0.0f);
    float weightSum = 0.0f;
    V4 cval = textureLinear(P.color, P.W, P.H, tx, ty);
    V4 pval = textureLinear(P.position, P.W, P.H, tx, ty);
    const float Rx = P.radius[0], Ry = P.radius[1];
    const float Wf = (float)P.W, Hf = (float)P.H;
    AxisTap X[5], Y[5];
#pragma unroll
    for (int k = 0; k < 5; k++) {
        X[k] = axisTap(tx - fdiv(Rx, Wf) + fdiv(((float)k + 0.5f) * Rx, 2.5f * Wf), P.W);
        Y[k] = rowTap(axisTap(ty - fdiv(Ry, Hf) + fdiv(((float)k + 0.5f) * Ry, 2.5f * Hf), P.H), P.W);
    }
    // ---- level 0: stepwidth 0, step 1, every tap ----
    {
        const float stepwidth = 0.0f;
#pragma unroll
        for (int count = 0; count < 25; count++) {
            const int i = count / 5, j = count % 5;
            const int delt = count > 12 ? count - 12 : 12 - count;
            const float _h = SAIL_WAVELET_H(delt);
            // W(), wavelet.glsl:5-20: the "normal" weight reuses the colour difference as written
            V4 ctmp = bilinearAt(P.color4, X[j], Y[i]);
            V4 t = cval - ctmp;
            float dist2 = dot4(t, t);
            const float d2c = dist2;
            float c_w = fmin2(dm_exp(fdiv(-dist2, 4.0f)), 1.0f);
            dist2 = fmax2(fdiv(d2c, stepwidth * stepwidth), 0.0f);
            float n_w = fmin2(dm_exp(fdiv(-dist2, 128.0f)), 1.0f);
            V4 ptmp = bilinearAt(P.position4, X[j], Y[i]);
            t = pval - ptmp;
            dist2 = dot4(t, t);
            float p_w = fmin2(dm_exp(fdiv(-dist2, 1.0f)), 1.0f);
            if ((count & 1) == 0) {  // delt even: levels 1 (and 2, delt % 4 == 0) take this tap again
                float* c = cache + (count / 2) * 6 * 256;
                c[0] = ctmp.x; c[256] = ctmp.y; c[512] = ctmp.z; c[768] = c_w; c[1024] = p_w; c[1280] = d2c;
            }
            float weight = c_w * n_w * p_w * _h;
            ctmp = ctmp * weight;
            weightSum += weight;
            color = color + ctmp;
        }
    }
    // ---- levels 1 and 2: stepwidth 1 / 3, step 2 / 4 ----
#pragma unroll 1
    for (int n = 1; n < 3; n++) {
        const float stepwidth = (float)(1 << n) - 1.0f;
#pragma unroll
        for (int count = 0; count < 25; count += 2) {
            const int delt = count > 12 ? count - 12 : 12 - count;
            if (delt & ((1 << n) - 1)) continue;  // uniform: delt % step != 0 -> _h = 0, the shader adds nothing
            const float _h = n == 1 ? SAIL_WAVELET_H(delt / 2) : SAIL_WAVELET_H(delt / 4);
            const float* c = cache + (count / 2) * 6 * 256;
            // the bilinear fetch of a colour map whose texels carry w = 1 has w = omt + t = 1 exactly (t is a multiple of 1/256)
            V4 ctmp = mk4(c[0], c[256], c[512], 1.0f);
            const float c_w = c[768], p_w = c[1024], d2c = c[1280];
            float dist2 = fmax2(fdiv(d2c, stepwidth * stepwidth), 0.0f);
            float n_w = fmin2(dm_exp(fdiv(-dist2, 128.0f)), 1.0f);
            float weight = c_w * n_w * p_w * _h;
            ctmp = ctmp * weight;
            weightSum += weight;
            color = color + ctmp;
        }
    }
    float r = dm_rcp(weightSum);
    return mk4(color.x * r, color.y * r, color.z * r, color.w * r);
}
#undef SAIL_WAVELET_H
// tonemapping.glsl:1-4
SD float tonemap1(float c) {
    float x = fmax2(0.0f, c - 0.004f);
    return fdiv(x * (6.2f * x + 0.5f), x * (6.2f * x + 1.7f) + 0.06f);
}

// KIND: 0 = the point filters (color / gamma / tonemapping / normal / position), 1 = window filters, 2 = wavelet -- three
// instantiations, so the 64-tap window filter's registers (112) do not set the occupancy of the one-fetch filters.
extern __shared__ float4 sail_present_smem[];  // KIND 2: SAIL_WAVELET_SMEM_BYTES (the wavelet's tap cache)
// CTAs per SM the register allocation aims at.  Measured at 1080p (tools/ubench/present_win.cu, three registers per prepared axis
// value): window filters 0.355 / 0.326 / 0.336 ms at 2 / 3 / 4 (85 registers at 3, spills at 4), wavelet 0.700 / 0.711 / 0.765 ms
// (its 80 KB tap cache allows two).
#ifndef SAIL_PRESENT_MIN_CTAS
#define SAIL_PRESENT_MIN_CTAS(KIND) ((KIND) == 0 ? 1 : ((KIND) == 1 ? 3 : 2))
#endif
template <int KIND>
__global__ void __launch_bounds__(256, SAIL_PRESENT_MIN_CTAS(KIND)) sail_present_kernel(const __grid_constant__ PresentParams P) {
    int x = blockIdx.x * 32 + (threadIdx.x & 31);
    int y = blockIdx.y * 8 + (threadIdx.x >> 5);
    if (x >= P.W || y >= P.H) return;
    // vsrender.glsl:5: texCoord = vertex.xy*0.5+0.5 interpolated to the pixel centre
    float tx = fdiv((float)x + 0.5f, (float)P.W), ty = fdiv((float)y + 0.5f, (float)P.H);
    V4 v = mk4(0.0f, 0.0f, 0.0f, 0.0f);
    switch (P.filter_id) {
        case 0: {  // color.glsl
            V4 c = textureLinear(P.color, P.W, P.H, tx, ty);
            v = mk4(c.x, c.y, c.z, 1.0f);
            break;
        }
        case 1: {  // gamma.glsl
            V4 c = textureLinear(P.color, P.W, P.H, tx, ty);
            float g = dm_rcp(P.gamma_c);
            v = mk4(dm_pow(c.x, g), dm_pow(c.y, g), dm_pow(c.z, g), 1.0f);
            break;
        }
        case 10: {  // tonemapping.glsl
            V4 c = textureLinear(P.color, P.W, P.H, tx, ty);
            v = mk4(tonemap1(c.x), tonemap1(c.y), tonemap1(c.z), 1.0f);
            break;
        }
        case 7: {  // normal.glsl
            V4 c = textureLinear(P.normal, P.W, P.H, tx, ty);
            v = mk4(c.x, c.y, c.z, 1.0f);
            break;
        }
        case 8: {  // position.glsl
            V4 c = textureLinear(P.position, P.W, P.H, tx, ty);
            v = mk4(c.x, c.y, c.z, 1.0f);
            break;
        }
        case 9: if (KIND == 2) v = filterWavelet(P, tx, ty, reinterpret_cast<float*>(sail_present_smem) + threadIdx.x); break;
        case 2: case 3: case 4: case 5: case 6: if (KIND == 1) v = filterWindow(P, tx, ty); break;
        default: v = mk4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    size_t i = (size_t)y * P.W + x;
    const float comp[4] = {v.x, v.y, v.z, v.w};
    if (P.out_f32) {
        reinterpret_cast<float4*>(P.out_f32)[i] = make_float4(v.x, v.y, v.z, v.w);
    }
    if (P.out_u8) {
        unsigned char b[4];
        for (int k = 0; k < 4; k++) {
            float f = comp[k];
            f = (f == f) ? dm_min(dm_max(f, 0.0f), 1.0f) : 0.0f;  // canvas: clamp, NaN -> 0
            b[k] = (unsigned char)dm_floor(f * 255.0f + 0.5f);
        }
        reinterpret_cast<uchar4*>(P.out_u8)[i] = make_uchar4(b[0], b[1], b[2], b[3]);
    }
}

// RGB (12 bytes per texel, the accumulator's layout) -> RGB1 (16 bytes): the window filters' fetch format
__global__ void __launch_bounds__(256) sail_pad_rgb_kernel(const float* __restrict__ rgb, float4* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i < n) out[i] = make_float4(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], 1.0f);
}

#ifndef __CUDACC_RTC__
static inline bool sailPresentWantsPadded(int filter_id) { return (filter_id >= 2 && filter_id <= 6) || filter_id == 9; }
// `pad`: W*H float4 of scratch, required for the window filters and the wavelet (sailPresentWantsPadded), unused otherwise;
// `pad2`: the same again for the wavelet (its position map)
static inline cudaError_t sailLaunchPresent(PresentParams P, float4* pad, float4* pad2, cudaStream_t stream) {
    dim3 grid((P.W + 31) / 32, (P.H + 7) / 8);
    P.color4 = nullptr;
    P.position4 = nullptr;
    const size_t n = (size_t)P.W * P.H;
    if (sailPresentWantsPadded(P.filter_id)) {
        sail_pad_rgb_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(P.color, pad, n);
        P.color4 = pad;
    }
    if (P.filter_id == 9) {
        // > 48 KB of dynamic shared memory is opt-in per device; the call is a table update, made on every launch
        cudaError_t e = cudaFuncSetAttribute(sail_present_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SAIL_WAVELET_SMEM_BYTES);
        if (e != cudaSuccess) return e;
        sail_pad_rgb_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(P.position, pad2, n);
        P.position4 = pad2;
        sail_present_kernel<2><<<grid, 256, SAIL_WAVELET_SMEM_BYTES, stream>>>(P);
    } else if (P.filter_id >= 2 && P.filter_id <= 6) sail_present_kernel<1><<<grid, 256, 0, stream>>>(P);
    else sail_present_kernel<0><<<grid, 256, 0, stream>>>(P);
    return cudaSuccess;
}
#endif

#endif

// sail_shade.cuh -- RNG, BSDF sampling, next-event estimation and the bounce loop.
//
// Follows /root/reference/src/shader/util/random.glsl:1-18, util/sampler.glsl:1-54,
// material/{ssutility,fresnel,microfacet,bsdf,matte,mirror,metal,glass}.glsl,
// material/shader.material.js:21-29, light/{area,point,spot}.glsl, light/shader.light.js:12-31,
// shape/shader.shape.js:53-67 (sampleGeometry) and trace/path.glsl:1-38.
//
// Differences in structure (not in values): the hash RNG pair is evaluated once per bounce and reused
// by the BSDF sample, the light pick and the light-point sample (the reference re-evaluates the same
// expression at each site); X_f() is only evaluated for matte hits because path.glsl:10
// uses `f` for matte, non-emissive hits only; out-parameters that the GLSL leaves unwritten read as 0.
#ifndef SAIL_SHADE_CUH
#define SAIL_SHADE_CUH
#include "sail_shapes.cuh"

// ---- util/random.glsl:5-10 ---------------------------------------------------------------------
// dm_sin (detmath.h) of two arguments at once: the same operations in the same order, with every FP32 step that
// both lanes share issued as one packed instruction (FMUL2 / FFMA2).  Each half rounds exactly like the scalar code,
// so sinPair(a, b) == (dm_sin(a), dm_sin(b)) bit for bit (tests/test_gpu_parity.py::test_device_math_matches_host
// checks the scalar path against the host, the frame tests check this one against the oracle).
SD V2 sinPair(float a, float b) {
    const F2 x = pk2(a, b);
    float ta, tb;
    upk2(mul2(x, splat2(DM_TWO_OVER_PI)), ta, tb);
    const float na = rintf(ta), nb = rintf(tb);
    const F2 nn = pk2(-na, -nb);
    F2 r = fma2(nn, splat2(DM_PIO2_HI), x);
    r = fma2(nn, splat2(DM_PIO2_MID), r);
    r = fma2(nn, splat2(DM_PIO2_LO), r);
    const F2 z = mul2(r, r);
    // dm_sin_kernel
    F2 ps = fma2(splat2(-1.9515295891e-4f), z, splat2(8.3321608736e-3f));
    ps = fma2(ps, z, splat2(-1.6666654611e-1f));
    const F2 sn = fma2(mul2(ps, z), r, r);
    // dm_cos_kernel
    F2 pc = fma2(splat2(2.443315711809948e-5f), z, splat2(-1.388731625493765e-3f));
    pc = fma2(pc, z, splat2(4.166664568298827e-2f));
    pc = mul2(mul2(pc, z), z);
    const F2 cs = add2(fma2(splat2(-0.5f), z, splat2(1.0f)), pc);
    float sa, sb, ca, cb;
    upk2(sn, sa, sb);
    upk2(cs, ca, cb);
    const int qa = (int)na, qb = (int)nb;
    float va = (qa & 1) ? ca : sa, vb = (qb & 1) ? cb : sb;
    va = (qa & 2) ? -va : va;
    vb = (qb & 2) ? -vb : vb;
#ifndef SAIL_RNG_SMALL_ARGS
    // Payne-Hanek (detmath.h): hours into a session.  A scene-specialised kernel is built without this arm while the host can
    // prove |argument| < 2^22 for every pixel, bounce and frame of the launch (sail_api.cu rngArgsSmall), and rebuilt with it after.
    if (!(dm_abs(a) < 4194304.0f)) va = dm_sin_large(a);
    if (!(dm_abs(b) < 4194304.0f)) vb = dm_sin_large(b);
#endif
    return mk2(va, vb);
}
static __device__ __noinline__ float2 sail_sin2(float a, float b) { V2 v = sinPair(a, b); return make_float2(v.x, v.y); }
template <int F>
SD V2 random2(float fx, float fy, float fz, float seed) {
    float ax = fx + seed, ay = fy + seed, az = fz + seed;
    float d1 = ax * 12.9898f + ay * 78.233f + az * 151.7182f;
    float d2 = ax * 63.7264f + ay * 10.873f + az * 623.6736f;
    V2 s2;
    if (SAIL_SMALL(F)) s2 = sinPair(d1, d2);
    else { float2 t = sail_sin2(d1, d2); s2 = mk2(t.x, t.y); }
    return mk2(ffract(s2.x * 43758.5453f + seed), ffract(s2.y * 43758.5453f + seed));
}

// Russian roulette -- no counterpart in the reference (fixed MAXBOUNCES, const/define.glsl:6); the opt-in extension the brief
// names, off unless sail_config.rr_start_bounce > 0 (SailLaunch::rrStart) and then defined by the oracle's trace()
// (oracle/shader_trace.h): from bounce rrStart on a path survives with probability q = clamp(max(fpdf), 0.05, 1), decided by one
// more draw of the shader's hash, random(scale, seed) of util/random.glsl:1-4 with a third scale vector; survivors are scaled
// by 1 / q.  Returns false when the path ends.  SAIL_HAS_RR: a scene-specialised kernel is built with or without it.
#ifdef SAIL_JIT_NRUNS
#define SAIL_HAS_RR (SAIL_JIT_RR)
#else
#define SAIL_HAS_RR 1
#endif
template <int F>
SD bool russianRoulette(int rrStart, int deepth, int maxBounces, float fx, float fy, float fz, float seed, V3& fpdf) {
    if (!SAIL_HAS_RR || rrStart <= 0 || deepth < rrStart || deepth >= maxBounces) return true;
    const float q = fclamp(fmax2(fpdf.x, fmax2(fpdf.y, fpdf.z)), 0.05f, 1.0f);
    const float ax = fx + seed, ay = fy + seed, az = fz + seed;
    const float d = ax * 37.1234f + ay * 91.4567f + az * 231.9876f;
    const float u = ffract(sail_sin(d) * 43758.5453f + seed);
    if (!(u < q)) return false;
    fpdf = fpdf * dm_rcp(q);
    return true;
}

// ---- util/sampler.glsl ---------------------------------------------------------------------------
SD V3 uniformSampleSphere(V2 u) {
    float z = 1.0f - 2.0f * u.x;
    float r = sqrtUnit(1.0f - z * z);  // z = 1 - 2u in [-1, 1]: the argument is 0 or >= 2^-24 (see sinTheta)
    float2 sc2 = sail_sincos(SAIL_TWO_PI * u.y);
    float s = sc2.x, c = sc2.y;
    return mk3(r * c, r * s, z);
}
template <int F>
SD V3 cosineSampleHemisphere(V2 u) {
    // u comes from random2(): fract() of a value that is 0 or at least 2^-24 in magnitude after the cancellation with a
    // seed >= 1 (sail_render_n rejects timeSinceStart < 0), so u.x, u.y and 1 - u.x lie in {0} u [2^-24, 1] (or are NaN):
    // the domain of sqrtUnit / sincosUnit.
    float r = sqrtUnit(u.x);
    float2 sc2 = sincosF<F>(SAIL_TWO_PI * u.y);
    float s = sc2.x, c = sc2.y;
    return mk3(r * c, r * s, sqrtUnit(1.0f - u.x));
}
SD V2 concentricSampleDisk(V2 u) {
    float uOffset = 2.0f * u.x - 1.0f;
    float vOffset = 2.0f * u.y - 1.0f;
    if (uOffset == 0.0f && vOffset == 0.0f) return mk2(0.0f, 0.0f);
    float theta, r;
    if (dm_abs(uOffset) > dm_abs(vOffset)) {
        r = uOffset;
        theta = fdiv(vOffset, uOffset) * SAIL_PIOVER4;
    } else {
        r = vOffset;
        theta = SAIL_PIOVER2 - fdiv(uOffset, vOffset) * SAIL_PIOVER4;
    }
    float2 sc2 = sail_sincos(theta);
    float s = sc2.x, c = sc2.y;
    return mk2(r * c, r * s);
}

// ---- material/ssutility.glsl:1-45 ------------------------------------------------------------------
SD float absCosTheta(V3 w) { return dm_abs(w.z); }
SD float cos2Theta(V3 w) { return w.z * w.z; }
SD float sin2Theta(V3 w) { return fmax2(0.0f, 1.0f - cos2Theta(w)); }
// max(0, 1 - c*c) is 0, or >= 2^-24: for c*c in [1/2, 1] the subtraction is exact, a multiple of 2^-24; below 1/2 the result
// exceeds 1/2; above 1, infinite or NaN it is clamped to 0 (FMNMX returns the non-NaN operand).  So sqrtUnit's domain holds.
SD float sinTheta(V3 w) { return sqrtUnit(sin2Theta(w)); }
SD bool equalZero(float x) { return x < 1e-3f && x > -1e-3f; }
SD float tan2Theta(V3 w) {
    float c2 = cos2Theta(w);
    if (c2 < SAIL_EPSILON) return SAIL_INF;
    return fdiv(sin2Theta(w), c2);
}
SD float cosPhi(V3 w) {
    float st = sinTheta(w);
    return equalZero(st) ? 1.0f : fclamp(fdiv(w.x, st), -1.0f, 1.0f);
}
SD float sinPhi(V3 w) {
    float st = sinTheta(w);
    return equalZero(st) ? 0.0f : fclamp(fdiv(w.y, st), -1.0f, 1.0f);
}
SD bool sameHemisphere(V3 w, V3 wp) { return w.z * wp.z > SAIL_EPSILON; }

// ---- material/fresnel.glsl:31-70 -------------------------------------------------------------------
SD float frDielectric(float cosThetaI, float etaI, float etaT) {
    cosThetaI = fclamp(cosThetaI, -1.0f, 1.0f);
    float sinThetaI = sqrtUnit(fmax2(0.0f, 1.0f - cosThetaI * cosThetaI));  // domain: see sinTheta()
    float sinThetaT = fdiv(etaI, etaT) * sinThetaI;
    if (sinThetaT >= 1.0f) return 1.0f;
    float cosThetaT = sqrtUnit(fmax2(0.0f, 1.0f - sinThetaT * sinThetaT));
    float TI = etaT * cosThetaI, IT = etaI * cosThetaT, II = etaI * cosThetaI, TT = etaT * cosThetaT;
    float Rparl = fdiv(TI - IT, TI + IT);
    float Rperp = fdiv(II - TT, II + TT);
    return fdiv(Rparl * Rparl + Rperp * Rperp, 2.0f);
}
SD float frConductor1(float cosThetaI, float cosThetaI2, float sinThetaI2, float etaI, float etaT, float k) {
    float eta = fdiv(etaT, etaI);
    float etak = fdiv(k, etaI);
    float eta2 = eta * eta;
    float etak2 = etak * etak;
    float t0 = eta2 - etak2 - sinThetaI2;
    float a2plusb2 = dm_sqrt(t0 * t0 + 4.0f * eta2 * etak2);
    float t1 = a2plusb2 + cosThetaI2;
    float a = dm_sqrt(0.5f * (a2plusb2 + t0));
    float t2 = 2.0f * cosThetaI * a;
    float Rs = fdiv(t1 - t2, t1 + t2);
    float t3 = cosThetaI2 * a2plusb2 + sinThetaI2 * sinThetaI2;
    float t4 = t2 * sinThetaI2;
    float Rp = fdiv(Rs * (t3 - t4), t3 + t4);
    return 0.5f * (Rp + Rs);
}
SD V3 frConductor(float cosThetaI, V3 etaI, V3 etaT, V3 k) {
    cosThetaI = fclamp(cosThetaI, -1.0f, 1.0f);
    float c2 = cosThetaI * cosThetaI;
    float s2 = 1.0f - c2;
    return mk3(frConductor1(cosThetaI, c2, s2, etaI.x, etaT.x, k.x), frConductor1(cosThetaI, c2, s2, etaI.y, etaT.y, k.y),
               frConductor1(cosThetaI, c2, s2, etaI.z, etaT.z, k.z));
}

// ---- material/microfacet.glsl:41-71 (Trowbridge-Reitz; Beckmann is never constructed) ----------------
SD V3 trowbridgeReitz_sample_wh(V2 u, float alphax, float alphay, V3 wo) {
    float cosTheta_ = 0.0f, phi = SAIL_TWO_PI * u.x;
    // SAIL_HAS_TR_ISO / SAIL_HAS_TR_ANISO: which arm the scene's metal / rough-glass roughness pairs select (sail_jit.cpp)
    if (SAIL_HAS_TR_ISO && (!SAIL_HAS_TR_ANISO || alphax == alphay)) {
        float tanTheta2 = fdiv(alphax * alphax * u.x, 1.0f - u.x);
        cosTheta_ = dm_rcp(dm_sqrt(1.0f + tanTheta2));
    } else if (SAIL_HAS_TR_ANISO) {
        phi = dm_atan(fdiv(alphay, alphax) * dm_tan(SAIL_PIOVER2 + SAIL_TWO_PI * u.x));
        if (u.x > 0.5f) phi += SAIL_PI;
        float2 sc2 = sail_sincos(phi);
        float sp = sc2.x, cp = sc2.y;
        float alphax2 = alphax * alphax, alphay2 = alphay * alphay;
        float alpha2 = dm_rcp(fdiv(cp * cp, alphax2) + fdiv(sp * sp, alphay2));
        float tanTheta2 = fdiv(alpha2 * u.x, 1.0f - u.x);
        cosTheta_ = dm_rcp(dm_sqrt(1.0f + tanTheta2));
    }
    float sinTheta_ = sqrtUnit(fmax2(0.0f, 1.0f - cosTheta_ * cosTheta_));
    float2 scw = sail_sincos(phi);
    float sp = scw.x, cp = scw.y;
    V3 wh = mk3(sinTheta_ * cp, sinTheta_ * sp, cosTheta_);  // sphericalDirection, utility.glsl:53-56
    if (!sameHemisphere(wo, wh)) wh = -wh;
    return wh;
}
SD float trowbridgeReitz_d(float alphax, float alphay, V3 wh) {
    float tan2 = tan2Theta(wh);
    if (tan2 >= SAIL_INF) return 0.001f;
    float cos4Theta = cos2Theta(wh) * cos2Theta(wh);
    float cp = cosPhi(wh), sp = sinPhi(wh);
    float e = (fdiv(cp * cp, alphax * alphax) + fdiv(sp * sp, alphay * alphay)) * tan2;
    return dm_rcp(SAIL_PI * alphax * alphay * cos4Theta * (1.0f + e) * (1.0f + e));
}
SD float trowbridgeReitz_pdf(float alphax, float alphay, V3 wh) { return trowbridgeReitz_d(alphax, alphay, wh) * absCosTheta(wh); }

// ---- material/bsdf.glsl:162-196 microfacet reflection ----------------------------------------------
// fresnelKind: 1 = conductor (etaI = WHITE, etaT, k), 2 = dielectric (etaI.x, etaT.x)
struct MicroR {
    V3 R;
    int fresnelKind;
    V3 etaI, etaT, k;
    float ax, ay;
};
SD V3 microfacet_r_f(const MicroR& mr, V3 wo, V3 wi) {
    float cosThetaO = absCosTheta(wo), cosThetaI = absCosTheta(wi);
    V3 wh = wi + wo;
    if (cosThetaI < SAIL_EPSILON || cosThetaO < SAIL_EPSILON) return splat3(0.0f);
    if (equalZero(wh.x) && equalZero(wh.y) && equalZero(wh.z)) return splat3(0.0f);
    wh = normalize3(wh);
    float c = dot3(wi, wh);
    V3 F = (mr.fresnelKind == 2) ? splat3(1.0f) * frDielectric(c, mr.etaI.x, mr.etaT.x) : frConductor(c, mr.etaI, mr.etaT, mr.k);
    return vdivs(mr.R * trowbridgeReitz_d(mr.ax, mr.ay, wh) * F, 4.0f * cosThetaI * cosThetaO);
}
SD V3 microfacet_r_sample_f(const MicroR& mr, V2 u, V3 wo, V3& wi, float& pdf) {
    if (wo.z < SAIL_EPSILON) return splat3(0.0f);  // wi, pdf stay unwritten (= 0)
    V3 wh = trowbridgeReitz_sample_wh(u, mr.ax, mr.ay, wo);
    wi = reflect3(-wo, wh);
    if (!sameHemisphere(wo, wi)) return splat3(0.0f);  // pdf stays unwritten
    pdf = fdiv(trowbridgeReitz_pdf(mr.ax, mr.ay, wh), 4.0f * dot3(wo, wh));
    return microfacet_r_f(mr, wo, wi);
}

// ---- material/bsdf.glsl:198-244 microfacet transmission --------------------------------------------
struct MicroT {
    V3 T;
    float etaA, etaB;
    bool into;
    float ax, ay;
};
SD V3 microfacet_t_f(const MicroT& mt, V3 wo, V3 wi) {
    if (sameHemisphere(wo, wi)) return splat3(0.0f);
    float cosThetaO = wo.z, cosThetaI = wi.z;
    if (equalZero(cosThetaI) || equalZero(cosThetaO)) return splat3(0.0f);
    float eta = mt.into ? fdiv(mt.etaB, mt.etaA) : fdiv(mt.etaA, mt.etaB);
    V3 wh = normalize3(wo + wi * eta);
    if (wh.z < -SAIL_EPSILON) wh = -wh;
    float F = frDielectric(dot3(wo, wh), mt.etaA, mt.etaB);
    float sqrtDenom = dot3(wo, wh) + eta * dot3(wi, wh);
    float s = dm_abs(fdiv(eta * eta * trowbridgeReitz_d(mt.ax, mt.ay, wh) * dm_abs(dot3(wi, wh)) * dm_abs(dot3(wo, wh)),
                          cosThetaI * cosThetaO * sqrtDenom * sqrtDenom));
    return (1.0f - F) * mt.T * s;
}
SD float microfacet_t_pdf(const MicroT& mt, V3 wo, V3 wi) {
    if (sameHemisphere(wo, wi)) return 0.001f;
    float eta = mt.into ? fdiv(mt.etaB, mt.etaA) : fdiv(mt.etaA, mt.etaB);
    V3 wh = normalize3(wo + wi * eta);
    float sqrtDenom = dot3(wo, wh) + eta * dot3(wi, wh);
    float dwh_dwi = dm_abs(fdiv(eta * eta * dot3(wi, wh), sqrtDenom * sqrtDenom));
    return trowbridgeReitz_pdf(mt.ax, mt.ay, wh) * dwh_dwi;
}
SD V3 microfacet_t_sample_f(const MicroT& mt, V2 u, V3 wo, V3& wi, float& pdf) {
    if (equalZero(wo.z)) return splat3(0.0f);
    V3 wh = trowbridgeReitz_sample_wh(u, mt.ax, mt.ay, wo);
    float eta = mt.into ? fdiv(mt.etaA, mt.etaB) : fdiv(mt.etaB, mt.etaA);
    wi = refract3(-wo, wh, eta);
    pdf = microfacet_t_pdf(mt, wo, wi);
    return microfacet_t_f(mt, wo, wi);
}

// ---- bsdf.glsl:45-76 Oren-Nayar ----------------------------------------------------------------------
SD V3 orenNayar_f(V3 R, float A, float B, V3 wo, V3 wi) {
    float sinThetaI = sinTheta(wi);
    float sinThetaO = sinTheta(wo);
    float maxCos = 0.0f;
    if (sinThetaI > SAIL_EPSILON && sinThetaO > SAIL_EPSILON) {
        float sinPhiI = sinPhi(wi), cosPhiI = cosPhi(wi);
        float sinPhiO = sinPhi(wo), cosPhiO = cosPhi(wo);
        float dCos = cosPhiI * cosPhiO + sinPhiI * sinPhiO;
        maxCos = fmax2(0.0f, dCos);
    }
    float sinAlpha, tanBeta;
    if (absCosTheta(wi) > absCosTheta(wo)) {
        sinAlpha = sinThetaO;
        tanBeta = fdiv(sinThetaI, absCosTheta(wi));
    } else {
        sinAlpha = sinThetaI;
        tanBeta = fdiv(sinThetaO, absCosTheta(wo));
    }
    return R * SAIL_INVPI * (A + B * maxCos * sinAlpha * tanBeta);
}

// ---- material(): shader.material.js:21-29 + matte/mirror/metal/glass.glsl ---------------------------
// Returns fpdf (unclamped); writes wi (local frame) and, for matte only, f = matte_f().
#define SAIL_MAT_ARM(c) SAIL_HAS_MAT(c)
template <int F>
SD V3 material(const SailSurface& su, int matCategory, V3 sc, bool into, V2 u, V3 wo, V3& wi, V3& f) {
    const float* m = su.m;
    wi = splat3(0.0f);
    f = splat3(0.0f);
    float pdf = 0.0f;
    V3 fs = splat3(0.0f);
    switch (matCategory) {
        case SAIL_MATTE: if (SAIL_MAT_ARM(SAIL_MATTE)) {  // matte.glsl:8-37
            V3 R = m[SM_MATTE_KD] * sc;
            wi = cosineSampleHemisphere<F>(u);
            pdf = sameHemisphere(wo, wi) ? absCosTheta(wi) * SAIL_INVPI : 0.0f;
            // SAIL_HAS_LAMBERT / SAIL_HAS_OREN_NAYAR: which of the two the scene's matte materials select (sail_jit.cpp)
            if (SAIL_HAS_LAMBERT && (!SAIL_HAS_OREN_NAYAR || m[SM_MATTE_SIGMA] < SAIL_EPSILON)) fs = R * SAIL_INVPI;  // bsdf.glsl:7-19
            else if (SAIL_HAS_OREN_NAYAR) fs = orenNayar_f(R, m[SM_MATTE_A], m[SM_MATTE_B], wo, wi);  // bsdf.glsl:45-76
            f = fs;  // matte_f() evaluates the same expression on the same (wo, wi)
            break;
        }
        case SAIL_MIRROR: if (SAIL_MAT_ARM(SAIL_MIRROR)) {  // mirror.glsl:5-16, bsdf.glsl:93-98 with the no-op Fresnel (= WHITE)
            V3 R = m[SM_MIRROR_KR] * sc;
            wi = mk3(-wo.x, -wo.y, wo.z);
            fs = vdivs(splat3(1.0f) * R, absCosTheta(wi));
            return fs * absCosTheta(wi);  // pdf = 1: `/ pdf` multiplies by rcp(1) = 1 exactly
        }
        case SAIL_METAL: if ((F & SAIL_F_GLOSSY) && SAIL_MAT_ARM(SAIL_METAL)) {  // metal.glsl:8-23
            MicroR mr;
            mr.R = sc;
            mr.fresnelKind = 1;
            mr.etaI = splat3(1.0f);
            mr.etaT = ld3(m + SM_METAL_ETA);
            mr.k = ld3(m + SM_METAL_K);
            mr.ax = m[SM_METAL_UR];
            mr.ay = m[SM_METAL_VR];
            fs = microfacet_r_sample_f(mr, u, wo, wi, pdf);
            break;
        }
        case SAIL_GLASS: if ((F & SAIL_F_GLOSSY) && SAIL_MAT_ARM(SAIL_GLASS)) {  // glass.glsl:10-35
            float kr = m[SM_GLASS_KR], kt = m[SM_GLASS_KT], eta = m[SM_GLASS_ETA];
            float ur = m[SM_GLASS_UR], vr = m[SM_GLASS_VR];
            const bool smooth = ur < SAIL_EPSILON && vr < SAIL_EPSILON;
            if (SAIL_HAS_SMOOTH_GLASS && (smooth || !SAIL_HAS_ROUGH_GLASS)) {  // bsdf.glsl:141-158 SpecularFr(1, eta)
                float Fr = frDielectric(wo.z, 1.0f, eta);
                if (u.x < Fr) {
                    wi = mk3(-wo.x, -wo.y, wo.z);
                    pdf = 1.0f;
                    fs = vdivs(kr * sc, absCosTheta(wi));
                } else {
                    float etaI = into ? 1.0f : eta;
                    float etaT = into ? eta : 1.0f;
                    wi = refract3(-wo, mk3(0.0f, 0.0f, 1.0f), fdiv(etaI, etaT));
                    V3 ft = (kt * sc) * (1.0f - Fr);
                    pdf = 1.0f;
                    fs = vdivs(ft, absCosTheta(wi));
                }
            } else if (SAIL_HAS_ROUGH_GLASS) {
                float p = u.x;
                u.x = fmin2(u.x * 2.0f - 1.0f, SAIL_ONEMINUSEPSILON);
                if (p < 0.5f) {
                    MicroR mr;
                    mr.R = kr * sc;
                    mr.fresnelKind = 2;
                    mr.etaI = splat3(1.0f);
                    mr.etaT = splat3(eta);
                    mr.k = splat3(0.0f);
                    mr.ax = ur;
                    mr.ay = vr;
                    fs = microfacet_r_sample_f(mr, u, wo, wi, pdf);
                } else {
                    MicroT mt;
                    mt.T = kt * sc;
                    mt.etaA = 1.0f;
                    mt.etaB = eta;
                    mt.into = into;
                    mt.ax = ur;
                    mt.ay = vr;
                    fs = microfacet_t_sample_f(mt, u, wo, wi, pdf);
                }
            }
            break;
        }
        default:  // no arm: fpdf is never written (reads as 0)
            return splat3(0.0f);
    }
    return vdivs(fs * absCosTheta(wi), pdf);  // `f * absCosTheta(wi) / pdf`
}

// ---- sampleGeometry(), shader.shape.js:53-67 ----------------------------------------------------------
// Only sphere / rectangle / disk have a sampler; the other shapes return BLACK and leave pdf unwritten
// (cube.glsl:50-52 ...), after which normalForX(BLACK) is still evaluated.
template <int F>
static __device__ __noinline__ V3 stubShapeNormal(const SailShape& s) {
    const float* q = s.q;
    const V3 hit = splat3(0.0f);
    switch (s.category) {
        case SAIL_CUBE: return boxFaceNormal(hit, loOf(s), hiOf(s), s.rev);       // cube.glsl:25-38
        case SAIL_CORNELLBOX: return boxFaceNormal(hit, loOf(s), hiOf(s), 1.0f);  // cornellbox.glsl:39-51
        case SAIL_CONE: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_LIGHT_SHAPE(SAIL_CONE)) {  // cone.glsl:38-46
            V3 hh = hit - ld3(q + SQ_CONE_P);
            float d = dm_sqrt(hh.x * hh.x + hh.y * hh.y);
            float x1 = d * q[SQ_CONE_INVTANA];
            float x2 = d * q[SQ_CONE_TANA];
            V3 no = mk3(0.0f, 0.0f, q[SQ_CONE_H] - x1 - x2);
            return s.rev * normalize3(hh - no);
        }
        case SAIL_CYLINDER: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_LIGHT_SHAPE(SAIL_CYLINDER))  // cylinder.glsl:36-38
            return s.rev * normalize3(mk3(hit.x - q[SQ_CYL_P], hit.y - q[SQ_CYL_P + 1], 0.0f));
        case SAIL_HYPERBOLOID: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_LIGHT_SHAPE(SAIL_HYPERBOLOID)) {  // hyperboloid.glsl:48-58
            V3 p1 = ld3(q + SQ_HYP_P1), p2 = ld3(q + SQ_HYP_P2);
            float v = (hit.z - p1.z) * q[SQ_HYP_INVDZ];
            V3 pr = (1.0f - v) * p1 + v * p2;
            float phi = sail_atan2(pr.x * hit.y - hit.x * pr.y, hit.x * pr.x + hit.y * pr.y);
            if (phi < 0.0f) phi += SAIL_TWO_PI;
            float2 sc2 = sail_sincos(phi);
            float sp = sc2.x, cp = sc2.y;
            V3 dpdu = dpduAround(hit.x, hit.y);
            V3 dpdv = mk3((p2.x - p1.x) * cp - (p2.y - p1.y) * sp, (p2.x - p1.x) * sp + (p2.y - p1.y) * cp, p2.z - p1.z);
            return s.rev * objToWorld(normalize3(cross3(dpdu, dpdv)));
        }
        case SAIL_PARABOLOID: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_LIGHT_SHAPE(SAIL_PARABOLOID)) {  // paraboloid.glsl:42-49
            V3 dpdu = dpduAround(hit.x, hit.y);
            float inv = dm_rcp(2.0f * hit.z);
            V3 dpdv = q[SQ_PAR_DZ] * mk3(hit.x * inv, hit.y * inv, 1.0f);
            return s.rev * objToWorld(normalize3(cross3(dpdu, dpdv)));
        }
    }
    return splat3(0.0f);
}
template <int F>
SD V3 sampleGeometry(const SceneView& sc, V2 u, int i, V3& normal, float& pdf) {
    normal = splat3(0.0f);
    pdf = 0.0f;
    if (sc.n <= 0) return splat3(0.0f);
    const SailShape& s = sc.shape[i];
    const float* q = s.q;
    switch (s.category) {
        case SAIL_SPHERE: if (SAIL_HAS_LIGHT_SHAPE(SAIL_SPHERE)) {  // sphere.glsl:29-31,88-92
            V3 c = ld3(q + SQ_SPH_C);
            V3 p = uniformSampleSphere(u) * q[SQ_SPH_R] + c;
            pdf = q[SQ_SPH_PDF];
            normal = (s.rev * (p - c)) * q[SQ_SPH_INVR];
            return p;
        }
        case SAIL_RECTANGLE: if (SAIL_HAS_LIGHT_SHAPE(SAIL_RECTANGLE)) {  // rectangle.glsl:25-30,65-70
            V3 p = loOf(s) + ld3(q + SQ_RECT_DPDU) * u.x + ld3(q + SQ_RECT_DPDV) * u.y;
            pdf = q[SQ_RECT_PDF];
            normal = s.rev * ld3(q + SQ_RECT_N);
            return p;
        }
        case SAIL_DISK: if ((F & SAIL_F_QUADRICS) && SAIL_HAS_LIGHT_SHAPE(SAIL_DISK)) {  // disk.glsl:32-34,77-83
            V2 pd = concentricSampleDisk(u);
            V3 p = mk3(pd.x * q[SQ_DISK_R] + q[SQ_DISK_P], q[SQ_DISK_P + 1], pd.y * q[SQ_DISK_R] + q[SQ_DISK_P + 2]);
            pdf = q[SQ_DISK_PDF];
            normal = s.rev * mk3(0.0f, 1.0f, 0.0f);
            return p;
        }
    }
    if (SAIL_HAS_STUB_LIGHT_SHAPE) normal = stubShapeNormal<F>(s);
    return splat3(0.0f);
}

// ---- light_sample(), shader.light.js:12-22 + area/point/spot.glsl -----------------------------------
// Three steps instead of the shader's one arm per light kind: (1) the point on the light (per kind, cheap), (2) ONE shadow
// ray for all lanes of the warp whatever kind of light each lane drew, (3) the radiance formula of the kind.  The shader
// calls testShadow() inside each arm (area.glsl:20, point.glsl:17, spot.glsl:27); with several kinds of light in a scene that
// would run the nearest-hit scan -- the most expensive function of the path -- once per kind with a third of the lanes each.
// Every value is computed by the same operations on the same operands; only the order of independent steps differs.
template <int F>
SD V3 light_sample(const SceneView& sc, V2 u, V3 hit, V3 insNormal) {
    const int ln = sc.ln;
    if (ln == 0) return splat3(0.0f);
    int index = toInt(u.x * (float)ln);  // randomInt(seed, 0, ln): same hash value as u.x
    if (index > ln) index = ln;
    const SailLight& l = sc.light[index];
    const int cat = l.category;
    V3 toLight = splat3(0.0f), normal = splat3(0.0f);
    float pdf = 0.0f;
    if (cat == SAIL_AREA && SAIL_HAS_LIGHT(SAIL_AREA)) {  // area.glsl:16-17
        V3 p = sampleGeometry<F>(sc, u, l.object, normal, pdf);
        toLight = p - hit;
    } else if (cat == SAIL_POINT && SAIL_HAS_LIGHT(SAIL_POINT)) {  // point.glsl:14-15
        V3 p = ld3(l.from) + uniformSampleSphere(u) * 0.1f;
        toLight = p - hit;
    } else if (cat == SAIL_SPOT && SAIL_HAS_LIGHT(SAIL_SPOT)) {  // spot.glsl:25
        toLight = ld3(l.from) - hit;
    } else {
        return splat3(0.0f);  // no arm for this category
    }
    if (testShadow<F>(sc, hit, toLight)) return splat3(0.0f);
    switch (cat) {
        case SAIL_AREA: if (SAIL_HAS_LIGHT(SAIL_AREA)) {  // area.glsl:18-22
            V3 normToLight = normalize3(toLight);
            return vdivs(ld3(l.emission) * fmax2(0.0f, dot3(normal, -normToLight)) * fmax2(0.0f, dot3(normToLight, insNormal)), pdf);
        }
        case SAIL_POINT: if (SAIL_HAS_LIGHT(SAIL_POINT))  // point.glsl:18-19
            return ld3(l.emission) * fmax2(0.0f, dot3(normalize3(toLight), insNormal));
        case SAIL_SPOT: if (SAIL_HAS_LIGHT(SAIL_SPOT)) {  // spot.glsl:17-39 (axis fixed to -y)
            V3 normToLight = normalize3(toLight);
            float d = length3(toLight);
            V3 w = -normToLight;
            float cosTheta_ = -w.y;
            float fall;
            if (cosTheta_ < l.cosTotalWidth) fall = 0.0f;
            else if (cosTheta_ >= l.cosFalloffStart) fall = 1.0f;
            else {
                float delta = (cosTheta_ - l.cosTotalWidth) * l.invFalloffRange;
                float delta2 = delta * delta;
                fall = delta2 * delta2;
            }
            return vdivs(ld3(l.emission) * fall * fmax2(0.0f, dot3(normToLight, insNormal)), d * d);
        }
    }
    return splat3(0.0f);
}

// ---- shade(), path.glsl:1-14 ----------------------------------------------------------------------------
template <int F>
SD V3 shade(const SceneView& sc, const Hit& h, V2 u, V3 wo, V3& wi, V3& fpdf) {
    V3 f;
    // box faces: |dpdu| is exactly 1 (components 0 / +-1), so normalize() multiplies by rcp(sqrt(1)) = 1
    V3 ss = h.unitDpdu ? h.dpdu : normalize3(h.dpdu), ts = cross3(h.n, ss);
    wo = worldToLocal(wo, h.n, ss, ts);
    fpdf = vclamp01(material<F>(sc.surf[h.index], h.matCategory, h.sc, h.into, u, wo, wi, f));
    wi = localToWorld(wi, h.n, ss, ts);
    V3 direct = splat3(0.0f);
    if ((F & SAIL_F_NEE) && isblack(h.emission) && h.matCategory == SAIL_MATTE) direct = direct + light_sample<F>(sc, u, h.p, h.n) * f;
    return h.emission + direct;
}

// First-hit AOVs of one path (fstrace.glsl:8,15-16 plus the hit id / depth this library adds).  The trace kernels do not
// write them: the reference overwrites its normal / position attachments every frame (webgl.js:65-73), which here would be
// 32 B per pixel per launch that nobody reads on most frames.  sail_aov_kernel (sail_combine.cuh) computes them from the last
// frame's uniforms when a reader or a filter asks (sail_api.cu ensureAovs).
struct AovSink {
    float* normal;
    float* position;
    int* hitId;
    float* depth;
    size_t pix;
    bool on;
};
SD void writeAov(const AovSink& a, V3 n, V3 p, int id, float t) {
    // fstrace.glsl:15-16: out_normal = n/2+0.5, out_position = normalize(p)
    V3 on = vdivs(n, 2.0f) + 0.5f;
    V3 op = normalize3(p);
    a.normal[3 * a.pix] = on.x; a.normal[3 * a.pix + 1] = on.y; a.normal[3 * a.pix + 2] = on.z;
    a.position[3 * a.pix] = op.x; a.position[3 * a.pix + 1] = op.y; a.position[3 * a.pix + 2] = op.z;
    a.hitId[a.pix] = id;
    a.depth[a.pix] = t;
}

// The last-bounce reductions of tracePath() hang on this test.  -DSAIL_LITERAL_LAST_BOUNCE (SAIL_B200_JIT_DEFINES) turns them
// off: every bounce is then evaluated in full like the reference's loop -- same image, bit for bit; bench.py reports that rate too.
#ifdef SAIL_LITERAL_LAST_BOUNCE
#define SAIL_IS_LAST_BOUNCE(d, maxB) false
#else
#define SAIL_IS_LAST_BOUNCE(d, maxB) ((d) == (maxB))
#endif

// True when shade() of this hit reduces to `emission + BLACK` (path.glsl:10: NEE only for non-emissive matte hits).
template <int F>
SD bool lastBounceIsEmissionOnly(const Hit& h) {
    if (!(F & SAIL_F_NEE)) return true;
    return !(isblack(h.emission) && h.matCategory == SAIL_MATTE);
}

// False when a last-bounce hit on this object cannot change the radiance sample: it emits nothing and gets no NEE.
template <int F>
SD bool lastBounceNeedsRecord(const SailShape& s) {
    if (!isblack(ld3(s.emission))) return true;
    return (F & SAIL_F_NEE) && s.matCategory == SAIL_MATTE;
}

// ---- trace(), path.glsl:16-38 -----------------------------------------------------------------------------
// fx, fy, fz = gl_FragCoord.xyz.  Returns the radiance sample e.
template <int F>
SD V3 tracePath(const SceneView& sc, Ray ray, int maxBounces, float fx, float fy, float fz, float timeSinceStart, int rrStart = 0) {
    V3 fpdf = splat3(1.0f), e = splat3(0.0f);
    int deepth = 0;
    while (deepth++ < maxBounces) {
        // Last bounce without NEE: only a hit on an emissive object adds anything (below).  If the ray does not even
        // touch one, the nearest-hit scan over the other objects is dead code as well.
        if (!(F & SAIL_F_NEE) && SAIL_IS_LAST_BOUNCE(deepth, maxBounces) && !touchesMasked<F>(sc, ray, sc.emissive)) break;
        const Nearest nr = nearestObject<F>(sc, ray);
        if (nr.id < 0) break;
        // Last bounce, object without emission (and no NEE for it): the hit adds (BLACK + BLACK) * fpdf = +0 to `e`
        // (fpdf is a product of values clamped to [0, 1]; `e` is never -0), so not even its hit record is needed.
        if (SAIL_IS_LAST_BOUNCE(deepth, maxBounces) && !lastBounceNeedsRecord<F>(sc.shape[nr.id])) break;
        Hit h;
        hitRecord<F>(sc, nr.id, nr.t, ray, h);
        // The last bounce only contributes `shade() * fpdf` (path.glsl:27): the sampled direction, the BSDF weight and the
        // respawned ray it also produces are never read again.  shade() = emission + direct, and direct is non-zero only
        // for non-emissive matte hits with lights present (path.glsl:10), so everywhere else the RNG pair, the shading
        // frame and the BSDF sample are dead code on this bounce and are not evaluated; what is added to `e` is the same
        // expression, `(emission + 0) * fpdf`.
        if (SAIL_IS_LAST_BOUNCE(deepth, maxBounces) && lastBounceIsEmissionOnly<F>(h)) {
            e = e + (h.emission + splat3(0.0f)) * fpdf;
            break;
        }
        float seed = timeSinceStart + (float)deepth;
        V2 u = random2<F>(fx, fy, fz, seed);
        V3 wi, _fpdf;
        e = e + shade<F>(sc, h, u, -ray.d, wi, _fpdf) * fpdf;
        if (SAIL_IS_LAST_BOUNCE(deepth, maxBounces)) break;
        fpdf = fpdf * _fpdf;
        if (!russianRoulette<F>(rrStart, deepth, maxBounces, fx, fy, fz, seed, fpdf)) break;
        float outdot = dot3(h.n, wi);
        ray.o = h.p + h.n * (outdot > SAIL_EPSILON ? 0.0001f : -0.0001f);
        ray.d = wi;
    }
    return e;
}

#endif

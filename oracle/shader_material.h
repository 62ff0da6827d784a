// oracle/shader_material.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/material/:
//   ssutility.glsl:1-45, fresnel.glsl:1-77, microfacet.glsl:41-92 (Trowbridge-Reitz; the Beckmann
//   branch :7-39 is unreachable because every material constructs TROWBRIDGEREITZ -- metal.glsl:16,
//   glass.glsl:22 -- and is not restated), bsdf.glsl:1-244, matte.glsl, mirror.glsl, metal.glsl,
//   glass.glsl and the dispatcher of shader.material.js:21-29.
#ifndef SAIL_ORACLE_SHADER_MATERIAL_H
#define SAIL_ORACLE_SHADER_MATERIAL_H
#include "shader_util.h"

// ---- ssutility.glsl:1-45 ---------------------------------------------------------------------
static inline real cosTheta(vec3 w) { return w.z; }
static inline real cos2Theta(vec3 w) { return w.z * w.z; }
static inline real absCosTheta(vec3 w) { return abs(w.z); }
static inline real sin2Theta(vec3 w) { return max(real(0.0f), real(1.0f) - cos2Theta(w)); }
static inline real sinTheta(vec3 w) { return sqrt(sin2Theta(w)); }
static inline real tan2Theta(vec3 w) {
    real cos2T = cos2Theta(w);
    if (cos2T < real(EPSILON)) return real(INF);
    return sin2Theta(w) / cos2T;
}
static inline real cosPhi(vec3 w) {
    real sinTheta_ = sinTheta(w);
    return (equalZero(sinTheta_)) ? real(1.0f) : clamp(w.x / sinTheta_, real(-1.0f), real(1.0f));
}
static inline real sinPhi(vec3 w) {
    real sinTheta_ = sinTheta(w);
    return (equalZero(sinTheta_)) ? real(0.0f) : clamp(w.y / sinTheta_, real(-1.0f), real(1.0f));
}
static inline real cos2Phi(vec3 w) { return cosPhi(w) * cosPhi(w); }
static inline real sin2Phi(vec3 w) { return sinPhi(w) * sinPhi(w); }
static inline bool sameHemisphere(vec3 w, vec3 wp) { return w.z * wp.z > real(EPSILON); }

// ---- fresnel.glsl:1-77 -----------------------------------------------------------------------
struct Fresnel {
    int type = NOOP;
    vec3 etaI, etaT, k;
};
static inline Fresnel createFresnelD(real etaI, real etaT) {
    Fresnel f;
    f.etaI = vec3(etaI);
    f.etaT = vec3(etaT);
    f.type = DIELECTRIC;
    return f;
}
static inline Fresnel createFresnelC(vec3 etaI, vec3 etaT, vec3 k) {
    Fresnel f;
    f.etaI = etaI;
    f.etaT = etaT;
    f.k = k;
    f.type = CONDUCTOR;
    return f;
}
static inline Fresnel createFresnelN() {
    Fresnel f;
    f.type = NOOP;
    return f;
}
static inline real frDielectric(real cosThetaI, real etaI, real etaT) {
    cosThetaI = clamp(cosThetaI, real(-1.0f), real(1.0f));
    real sinThetaI = sqrt(max(real(0.0f), real(1.0f) - cosThetaI * cosThetaI));
    real sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= real(1.0f)) return real(1.0f);
    real cosThetaT = sqrt(max(real(0.0f), real(1.0f) - sinThetaT * sinThetaT));
    real TI = etaT * cosThetaI, IT = etaI * cosThetaT, II = etaI * cosThetaI, TT = etaT * cosThetaT;
    real Rparl = (TI - IT) / (TI + IT);
    real Rperp = (II - TT) / (II + TT);
    return (Rparl * Rparl + Rperp * Rperp) / real(2.0f);
}
static inline vec3 frConductor(real cosThetaI, vec3 etaI, vec3 etaT, vec3 k) {
    cosThetaI = clamp(cosThetaI, real(-1.0f), real(1.0f));
    vec3 eta = etaT / etaI;
    vec3 etak = k / etaI;
    real cosThetaI2 = cosThetaI * cosThetaI;
    real sinThetaI2 = real(1.0f) - cosThetaI2;
    vec3 eta2 = eta * eta;
    vec3 etak2 = etak * etak;
    vec3 t0 = eta2 - etak2 - sinThetaI2;
    vec3 a2plusb2 = sqrt(t0 * t0 + real(4.0f) * eta2 * etak2);
    vec3 t1 = a2plusb2 + cosThetaI2;
    vec3 a = sqrt(real(0.5f) * (a2plusb2 + t0));
    vec3 t2 = real(2.0f) * cosThetaI * a;
    vec3 Rs = (t1 - t2) / (t1 + t2);
    vec3 t3 = cosThetaI2 * a2plusb2 + sinThetaI2 * sinThetaI2;
    vec3 t4 = t2 * sinThetaI2;
    vec3 Rp = Rs * (t3 - t4) / (t3 + t4);
    return real(0.5f) * (Rp + Rs);
}
static inline vec3 frEvaluate(const Fresnel& f, real cosThetaI) {
    if (f.type == DIELECTRIC) return WHITE * frDielectric(cosThetaI, f.etaI.x, f.etaT.x);
    else if (f.type == CONDUCTOR) return frConductor(cosThetaI, f.etaI, f.etaT, f.k);
    return WHITE;
}

// ---- microfacet.glsl:1-5,41-92 ---------------------------------------------------------------
struct MicrofacetDistribution {
    real alphax, alphay;
    int type = 0;
    MicrofacetDistribution() {}
    MicrofacetDistribution(real ax, real ay, int t) : alphax(ax), alphay(ay), type(t) {}
};
// microfacet.glsl:41-59.  Uses u.x for both the azimuth and the polar angle; u.y is unused (Q11).
static inline vec3 trowbridgeReitz_sample_wh(vec2 u, real alphax, real alphay, vec3 wo) {
    real cosTheta_ = 0.0f, phi = real(2.0f) * real(PI) * u.x;
    if (alphax == alphay) {
        real tanTheta2 = alphax * alphax * u.x / (real(1.0f) - u.x);
        cosTheta_ = real(1.0f) / sqrt(real(1.0f) + tanTheta2);
    } else {
        phi = atan(alphay / alphax * tan(real(PIOVER2) + real(2.0f) * real(PI) * u.x));
        if (u.x > real(0.5f)) phi += real(PI);
        real sinPhi_ = sin(phi), cosPhi_ = cos(phi);
        real alphax2 = alphax * alphax, alphay2 = alphay * alphay;
        real alpha2 = real(1.0f) / (cosPhi_ * cosPhi_ / alphax2 + sinPhi_ * sinPhi_ / alphay2);
        real tanTheta2 = alpha2 * u.x / (real(1.0f) - u.x);
        cosTheta_ = real(1.0f) / sqrt(real(1.0f) + tanTheta2);
    }
    real sinTheta_ = sqrt(max(real(0.0f), real(1.0f) - cosTheta_ * cosTheta_));
    vec3 wh = sphericalDirection(sinTheta_, cosTheta_, phi);
    if (!sameHemisphere(wo, wh)) wh = -wh;
    return wh;
}
// microfacet.glsl:61-67
static inline real trowbridgeReitz_d(real alphax, real alphay, vec3 wh) {
    real tan2Theta_ = tan2Theta(wh);
    if (tan2Theta_ >= real(INF)) return real(0.001f);
    real cos4Theta = cos2Theta(wh) * cos2Theta(wh);
    real e = (cos2Phi(wh) / (alphax * alphax) + sin2Phi(wh) / (alphay * alphay)) * tan2Theta_;
    return real(1.0f) / (real(PI) * alphax * alphay * cos4Theta * (real(1.0f) + e) * (real(1.0f) + e));
}
// microfacet.glsl:69-71
static inline real trowbridgeReitz_pdf(real alphax, real alphay, vec3 wo, vec3 wh) {
    return trowbridgeReitz_d(alphax, alphay, wh) * absCosTheta(wh);
}
// microfacet.glsl:73-92 (BECKMANN arms unreachable, see header)
static inline vec3 microfacet_sample_wh(const MicrofacetDistribution& md, vec2 u, vec3 wo) {
    if (md.type == TROWBRIDGEREITZ) return trowbridgeReitz_sample_wh(u, md.alphax, md.alphay, wo);
    return BLACK;
}
static inline real microfacet_d(const MicrofacetDistribution& md, vec3 wh) {
    if (md.type == TROWBRIDGEREITZ) return trowbridgeReitz_d(md.alphax, md.alphay, wh);
    return real(0.0f);
}
static inline real microfacet_pdf(const MicrofacetDistribution& md, vec3 wo, vec3 wh) {
    if (md.type == TROWBRIDGEREITZ) return trowbridgeReitz_pdf(md.alphax, md.alphay, wo, wh);
    return real(0.0f);
}

#include "shader_bsdf.h"
#endif

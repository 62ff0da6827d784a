// oracle/shader_bsdf.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/material/bsdf.glsl:1-244,
// matte.glsl:1-37, mirror.glsl:1-26, metal.glsl:1-34, glass.glsl:1-58, shader.material.js:21-29.
// LambertianT and SpecularT (bsdf.glsl:21-38,100-124) are never constructed by any material and are
// not restated.  Included from shader_material.h.
#ifndef SAIL_ORACLE_SHADER_BSDF_H
#define SAIL_ORACLE_SHADER_BSDF_H

// ---- bsdf.glsl:3-19 Lambertian reflection ----------------------------------------------------
struct LambertianR {
    vec3 R;
};
static inline vec3 lambertian_r_f(const LambertianR& lr, vec3 wo, vec3 wi) { return lr.R * real(INVPI); }
static inline real lambertian_r_pdf(const LambertianR& lr, vec3 wo, vec3 wi) {
    return sameHemisphere(wo, wi) ? absCosTheta(wi) * real(INVPI) : real(0.0f);
}
static inline vec3 lambertian_r_sample_f(const LambertianR& lr, vec2 u, vec3 wo, vec3& wi, real& pdf) {
    wi = cosineSampleHemisphere(u);
    pdf = lambertian_r_pdf(lr, wo, wi);
    return lambertian_r_f(lr, wo, wi);
}

// ---- bsdf.glsl:40-76 Oren-Nayar --------------------------------------------------------------
struct OrenNayar {
    vec3 R;
    real A, B;
};
static inline vec3 orenNayar_f(const OrenNayar& on, vec3 wo, vec3 wi) {
    real sinThetaI = sinTheta(wi);
    real sinThetaO = sinTheta(wo);
    real maxCos = 0.0f;
    if (sinThetaI > real(EPSILON) && sinThetaO > real(EPSILON)) {
        real sinPhiI = sinPhi(wi), cosPhiI = cosPhi(wi);
        real sinPhiO = sinPhi(wo), cosPhiO = cosPhi(wo);
        real dCos = cosPhiI * cosPhiO + sinPhiI * sinPhiO;
        maxCos = max(real(0.0f), dCos);
    }
    real sinAlpha, tanBeta;
    if (absCosTheta(wi) > absCosTheta(wo)) {
        sinAlpha = sinThetaO;
        tanBeta = sinThetaI / absCosTheta(wi);
    } else {
        sinAlpha = sinThetaI;
        tanBeta = sinThetaO / absCosTheta(wo);
    }
    return on.R * real(INVPI) * (on.A + on.B * maxCos * sinAlpha * tanBeta);
}
static inline real orenNayar_pdf(const OrenNayar& on, vec3 wo, vec3 wi) {
    return sameHemisphere(wo, wi) ? absCosTheta(wi) * real(INVPI) : real(0.0f);
}
static inline vec3 orenNayar_sample_f(const OrenNayar& on, vec2 u, vec3 wo, vec3& wi, real& pdf) {
    wi = cosineSampleHemisphere(u);
    pdf = orenNayar_pdf(on, wo, wi);
    return orenNayar_f(on, wo, wi);
}

// ---- bsdf.glsl:80-98 specular reflection -----------------------------------------------------
struct SpecularR {
    vec3 R;
    Fresnel f;
};
static inline vec3 specular_r_f(const SpecularR& sr, vec3 wo, vec3 wi) { return BLACK; }
static inline vec3 specular_r_sample_f(const SpecularR& sr, vec2 u, vec3 wo, vec3& wi, real& pdf) {
    wi = vec3(-wo.x, -wo.y, wo.z);
    pdf = 1.0f;
    return frEvaluate(sr.f, cosTheta(wi)) * sr.R / absCosTheta(wi);
}

// ---- bsdf.glsl:126-158 specular Fresnel (reflect or refract) ---------------------------------
struct SpecularFr {
    vec3 R, T;
    real etaA, etaB;
    bool into;
};
static inline vec3 specular_fr_f(const SpecularFr& sf, vec3 wo, vec3 wi) { return BLACK; }
static inline vec3 specular_fr_sample_f(const SpecularFr& sf, vec2 u, vec3 wo, vec3& wi, real& pdf) {
    real F = frDielectric(cosTheta(wo), sf.etaA, sf.etaB);
    if (u.x < F) {
        wi = vec3(-wo.x, -wo.y, wo.z);
        pdf = 1.0f;
        return sf.R / absCosTheta(wi);
    } else {
        real etaI = sf.into ? sf.etaA : sf.etaB;
        real etaT = sf.into ? sf.etaB : sf.etaA;
        wi = refract(-wo, vec3(0.0f, 0.0f, 1.0f), etaI / etaT);
        vec3 ft = sf.T * (real(1.0f) - F);
        pdf = 1.0f;
        return ft / absCosTheta(wi);
    }
}

// ---- bsdf.glsl:162-196 microfacet reflection -------------------------------------------------
struct MicrofacetR {
    vec3 R;
    Fresnel f;
    MicrofacetDistribution md;
};
static inline vec3 microfacet_r_f(const MicrofacetR& mr, vec3 wo, vec3 wi) {
    real cosThetaO = absCosTheta(wo), cosThetaI = absCosTheta(wi);
    vec3 wh = wi + wo;
    if (cosThetaI < real(EPSILON) || cosThetaO < real(EPSILON)) return BLACK * real(0.001f);
    if (equalZero(wh.x) && equalZero(wh.y) && equalZero(wh.z)) return BLACK * real(0.001f);
    wh = normalize(wh);
    vec3 F = frEvaluate(mr.f, dot(wi, wh));
    return mr.R * microfacet_d(mr.md, wh) * F / (real(4.0f) * cosThetaI * cosThetaO);
}
static inline vec3 microfacet_r_sample_f(const MicrofacetR& mr, vec2 u, vec3 wo, vec3& wi, real& pdf) {
    if (wo.z < real(EPSILON)) return BLACK * real(0.001f);  // wi, pdf left unwritten (Q7)
    vec3 wh = microfacet_sample_wh(mr.md, u, wo);
    wi = reflect(-wo, wh);
    if (!sameHemisphere(wo, wi)) return BLACK * real(0.001f);  // pdf left unwritten (Q7)
    pdf = microfacet_pdf(mr.md, wo, wh) / (real(4.0f) * dot(wo, wh));
    return microfacet_r_f(mr, wo, wi);
}

// ---- bsdf.glsl:198-244 microfacet transmission -----------------------------------------------
struct MicrofacetT {
    vec3 T;
    real etaA, etaB;
    bool into;
    MicrofacetDistribution md;
};
static inline vec3 microfacet_t_f(const MicrofacetT& mt, vec3 wo, vec3 wi) {
    if (sameHemisphere(wo, wi)) return BLACK * real(0.001f);
    real cosThetaO = cosTheta(wo);
    real cosThetaI = cosTheta(wi);
    if (equalZero(cosThetaI) || equalZero(cosThetaO)) return BLACK * real(0.001f);
    real eta = mt.into ? (mt.etaB / mt.etaA) : (mt.etaA / mt.etaB);
    vec3 wh = normalize(wo + wi * eta);
    if (wh.z < real(-EPSILON)) wh = -wh;
    real F = frDielectric(dot(wo, wh), mt.etaA, mt.etaB);
    real sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
    return (real(1.0f) - F) * mt.T *
           abs(eta * eta * microfacet_d(mt.md, wh) * abs(dot(wi, wh)) * abs(dot(wo, wh)) /
               (cosThetaI * cosThetaO * sqrtDenom * sqrtDenom));
}
static inline real microfacet_t_pdf(const MicrofacetT& mt, vec3 wo, vec3 wi) {
    if (sameHemisphere(wo, wi)) return real(0.001f);
    real eta = mt.into ? (mt.etaB / mt.etaA) : (mt.etaA / mt.etaB);
    vec3 wh = normalize(wo + wi * eta);
    real sqrtDenom = dot(wo, wh) + eta * dot(wi, wh);
    real dwh_dwi = abs((eta * eta * dot(wi, wh)) / (sqrtDenom * sqrtDenom));
    return microfacet_pdf(mt.md, wo, wh) * dwh_dwi;
}
static inline vec3 microfacet_t_sample_f(const MicrofacetT& mt, vec2 u, vec3 wo, vec3& wi, real& pdf) {
    if (equalZero(wo.z)) return BLACK * real(0.001f);  // wi, pdf left unwritten (Q7)
    vec3 wh = microfacet_sample_wh(mt.md, u, wo);
    real eta = mt.into ? (mt.etaA / mt.etaB) : (mt.etaB / mt.etaA);
    wi = refract(-wo, wh, eta);
    pdf = microfacet_t_pdf(mt, wo, wi);
    return microfacet_t_f(mt, wo, wi);
}

// ---- matte.glsl:1-37 -------------------------------------------------------------------------
static inline void matte_attr(real matIndex, real& kd, real& sigma, real& A, real& B) {
    kd = readFloat(texParams, vec2(1.0f, matIndex), TEX_PARAMS_LENGTH);
    sigma = readFloat(texParams, vec2(2.0f, matIndex), TEX_PARAMS_LENGTH);
    A = readFloat(texParams, vec2(3.0f, matIndex), TEX_PARAMS_LENGTH);
    B = readFloat(texParams, vec2(4.0f, matIndex), TEX_PARAMS_LENGTH);
}
static inline vec3 matte(vec2 u, real matIndex, vec3 sc, vec3 wo, vec3& wi, bool into) {
    vec3 f;
    real pdf;
    real kd, sigma, A, B;
    matte_attr(matIndex, kd, sigma, A, B);
    if (sigma < real(EPSILON)) {
        LambertianR diffuseR{kd * sc};
        f = lambertian_r_sample_f(diffuseR, u, wo, wi, pdf);
    } else {
        OrenNayar diffuseR{kd * sc, A, B};
        f = orenNayar_sample_f(diffuseR, u, wo, wi, pdf);
    }
    return f * absCosTheta(wi) / pdf;
}
static inline vec3 matte_f(real matIndex, vec3 sc, vec3 wo, vec3 wi, bool into) {
    real kd, sigma, A, B;
    matte_attr(matIndex, kd, sigma, A, B);
    if (sigma < real(EPSILON)) {
        LambertianR diffuseR{kd * sc};
        return lambertian_r_f(diffuseR, wo, wi);
    } else {
        OrenNayar diffuseR{kd * sc, A, B};
        return orenNayar_f(diffuseR, wo, wi);
    }
}

// ---- mirror.glsl:1-26 ------------------------------------------------------------------------
static inline vec3 mirror(vec2 u, real matIndex, vec3 sc, vec3 wo, vec3& wi, bool into) {
    vec3 f;
    real pdf;
    real kr = readFloat(texParams, vec2(1.0f, matIndex), TEX_PARAMS_LENGTH);
    SpecularR sr{kr * sc, createFresnelN()};
    f = specular_r_sample_f(sr, u, wo, wi, pdf);
    return f * absCosTheta(wi) / pdf;
}
static inline vec3 mirror_f(real matIndex, vec3 sc, vec3 wo, vec3 wi, bool into) {
    real kr = readFloat(texParams, vec2(1.0f, matIndex), TEX_PARAMS_LENGTH);
    SpecularR sr{kr * sc, createFresnelN()};
    return specular_r_f(sr, wo, wi);
}

// ---- metal.glsl:1-34 -------------------------------------------------------------------------
static inline void metal_attr(real matIndex, real& uroughness, real& vroughness, vec3& eta, vec3& k) {
    uroughness = readFloat(texParams, vec2(1.0f, matIndex), TEX_PARAMS_LENGTH);
    vroughness = readFloat(texParams, vec2(2.0f, matIndex), TEX_PARAMS_LENGTH);
    eta = readVec3(texParams, vec2(3.0f, matIndex), TEX_PARAMS_LENGTH);
    k = readVec3(texParams, vec2(6.0f, matIndex), TEX_PARAMS_LENGTH);
}
static inline vec3 metal(vec2 u, real matIndex, vec3 sc, vec3 wo, vec3& wi, bool into) {
    vec3 f;
    real pdf;
    real uroughness, vroughness;
    vec3 eta, k;
    metal_attr(matIndex, uroughness, vroughness, eta, k);
    MicrofacetDistribution md(uroughness, vroughness, TROWBRIDGEREITZ);
    Fresnel fresnel = createFresnelC(WHITE, eta, k);
    MicrofacetR mr{sc, fresnel, md};
    f = microfacet_r_sample_f(mr, u, wo, wi, pdf);
    return f * absCosTheta(wi) / pdf;
}
static inline vec3 metal_f(real matIndex, vec3 sc, vec3 wo, vec3 wi, bool into) {
    real uroughness, vroughness;
    vec3 eta, k;
    metal_attr(matIndex, uroughness, vroughness, eta, k);
    MicrofacetDistribution md(uroughness, vroughness, TROWBRIDGEREITZ);
    Fresnel fresnel = createFresnelC(WHITE, eta, k);
    MicrofacetR mr{sc, fresnel, md};
    return microfacet_r_f(mr, wo, wi);
}

// ---- glass.glsl:1-58 -------------------------------------------------------------------------
static inline void glass_attr(real matIndex, real& kr, real& kt, real& eta, real& uroughness, real& vroughness) {
    kr = readFloat(texParams, vec2(1.0f, matIndex), TEX_PARAMS_LENGTH);
    kt = readFloat(texParams, vec2(2.0f, matIndex), TEX_PARAMS_LENGTH);
    eta = readFloat(texParams, vec2(3.0f, matIndex), TEX_PARAMS_LENGTH);
    uroughness = readFloat(texParams, vec2(4.0f, matIndex), TEX_PARAMS_LENGTH);
    vroughness = readFloat(texParams, vec2(5.0f, matIndex), TEX_PARAMS_LENGTH);
}
static inline vec3 glass(vec2 u, real matIndex, vec3 sc, vec3 wo, vec3& wi, bool into) {
    vec3 f;
    real pdf;
    real kr, kt, eta, uroughness, vroughness;
    glass_attr(matIndex, kr, kt, eta, uroughness, vroughness);
    bool isSpecular = uroughness < real(EPSILON) && vroughness < real(EPSILON);
    if (isSpecular) {
        SpecularFr sf{kr * sc, kt * sc, 1.0f, eta, into};
        f = specular_fr_sample_f(sf, u, wo, wi, pdf);
    } else {
        MicrofacetDistribution md(uroughness, vroughness, TROWBRIDGEREITZ);
        real p = u.x;
        u.x = min(u.x * real(2.0f) - real(1.0f), real(ONEMINUSEPSILON));
        if (p < real(0.5f)) {
            Fresnel fresnel = createFresnelD(1.0f, eta);
            MicrofacetR mr{kr * sc, fresnel, md};
            f = microfacet_r_sample_f(mr, u, wo, wi, pdf);
        } else {
            MicrofacetT mt{kt * sc, 1.0f, eta, into, md};
            f = microfacet_t_sample_f(mt, u, wo, wi, pdf);
        }
    }
    return f * absCosTheta(wi) / pdf;
}
static inline vec3 glass_f(real matIndex, vec3 sc, vec3 wo, vec3 wi, bool into) {
    real kr, kt, eta, uroughness, vroughness;
    glass_attr(matIndex, kr, kt, eta, uroughness, vroughness);
    bool isSpecular = vroughness == real(0.0f) && uroughness == real(0.0f);
    if (isSpecular) {
        SpecularFr sf{kr * sc, kt * sc, 1.0f, eta, into};
        return specular_fr_f(sf, wo, wi);
    } else {
        MicrofacetDistribution md(uroughness, vroughness, TROWBRIDGEREITZ);
        if (sameHemisphere(wo, wi)) {
            Fresnel fresnel = createFresnelD(1.0f, eta);
            MicrofacetR mr{kr * sc, fresnel, md};
            return microfacet_r_f(mr, wo, wi);
        } else {
            MicrofacetT mt{kt * sc, 1.0f, eta, into, md};
            return microfacet_t_f(mt, wo, wi);
        }
    }
}

// ---- shader.material.js:21-29 ----------------------------------------------------------------
static inline vec3 material(const Intersect& ins, vec3 wo, vec3& wi, vec3& f) {
    f = BLACK;
    vec3 fpdf;
    if (false) {
    } else if (ins.matCategory == MATTE) {
        fpdf = matte(random2(ins.seed), ins.matIndex, ins.sc, wo, wi, ins.into);
        f = matte_f(ins.matIndex, ins.sc, wo, wi, ins.into);
    } else if (ins.matCategory == MIRROR) {
        fpdf = mirror(random2(ins.seed), ins.matIndex, ins.sc, wo, wi, ins.into);
        f = mirror_f(ins.matIndex, ins.sc, wo, wi, ins.into);
    } else if (ins.matCategory == METAL) {
        fpdf = metal(random2(ins.seed), ins.matIndex, ins.sc, wo, wi, ins.into);
        f = metal_f(ins.matIndex, ins.sc, wo, wi, ins.into);
    } else if (ins.matCategory == GLASS) {
        fpdf = glass(random2(ins.seed), ins.matIndex, ins.sc, wo, wi, ins.into);
        f = glass_f(ins.matIndex, ins.sc, wo, wi, ins.into);
    }
    return fpdf;
}
#endif

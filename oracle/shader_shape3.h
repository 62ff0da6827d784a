// oracle/shader_shape3.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/shape/hyperboloid.glsl:1-116,
// paraboloid.glsl:1-108 and the dispatchers of shader.shape.js:28-67.  Included from shader_shape2.h.
#ifndef SAIL_ORACLE_SHADER_SHAPE3_H
#define SAIL_ORACLE_SHADER_SHAPE3_H

// ---- hyperboloid.glsl:1-116 ------------------------------------------------------------------
struct Hyperboloid {
    vec3 p, p1, p2;
    real ah, ch;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForHyperboloid(const Ray& ray, const Hyperboloid& hy) {
    real r1 = sqrt(hy.p1.x * hy.p1.x + hy.p1.y * hy.p1.y);
    real r2 = sqrt(hy.p2.x * hy.p2.x + hy.p2.y * hy.p2.y);
    real rMax = max(r1, r2);
    real zMin = min(hy.p1.z, hy.p2.z);
    real zMax = max(hy.p1.z, hy.p2.z);
    Boundbox box(hy.p - vec3(rMax, -zMin, rMax), hy.p + vec3(rMax, zMax, rMax));
    return testBoundbox(ray, box);
}
static inline Hyperboloid parseHyperboloid(real index) {
    Hyperboloid hy;
    hy.p = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    hy.p1 = readVec3(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    hy.p2 = readVec3(objects, vec2(7.0f, index), OBJECTS_LENGTH);
    hy.ah = readFloat(objects, vec2(10.0f, index), OBJECTS_LENGTH);
    hy.ch = readFloat(objects, vec2(11.0f, index), OBJECTS_LENGTH);
    hy.reverseNormal = readBool(objects, vec2(12.0f, index), OBJECTS_LENGTH);
    hy.matIndex = rowCoordF(readFloat(objects, vec2(13.0f, index), OBJECTS_LENGTH), tn_);
    hy.texIndex = rowCoordF(readFloat(objects, vec2(14.0f, index), OBJECTS_LENGTH), tn_);
    hy.emission = readVec3(objects, vec2(15.0f, index), OBJECTS_LENGTH);
    return hy;
}
static inline void computeDpDForHyperboloid(vec3 hit, vec3 p1, vec3 p2, real phi, vec3& dpdu, vec3& dpdv) {
    real sinPhi_ = sin(phi), cosPhi_ = cos(phi);
    dpdu = vec3(real(-2.0f) * real(PI) * hit.y, real(2.0f) * real(PI) * hit.x, 0.0f);
    dpdv = vec3((p2.x - p1.x) * cosPhi_ - (p2.y - p1.y) * sinPhi_, (p2.x - p1.x) * sinPhi_ + (p2.y - p1.y) * cosPhi_,
                p2.z - p1.z);
}
static inline vec3 normalForHyperboloid(vec3 hit, const Hyperboloid& hy) {
    real v = (hit.z - hy.p1.z) / (hy.p2.z - hy.p1.z);
    vec3 pr = (real(1.0f) - v) * hy.p1 + v * hy.p2;
    real phi = atan(pr.x * hit.y - hit.x * pr.y, hit.x * pr.x + hit.y * pr.y);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    vec3 dpdu, dpdv;
    computeDpDForHyperboloid(hit, hy.p1, hy.p2, phi, dpdu, dpdv);
    vec3 normal = TO_WORLD(normalize(cross(dpdu, dpdv)));
    return real(hy.reverseNormal ? -1.0f : 1.0f) * normal;
}
static inline Intersect intersectHyperboloid(Ray ray, const Hyperboloid& hy) {
    Intersect result;
    result.d = MAX_DISTANCE;
    ray.dir = TO_LOCAL(ray.dir);
    ray.origin = TO_LOCAL(ray.origin - hy.p);
    real a = hy.ah * ray.dir.x * ray.dir.x + hy.ah * ray.dir.y * ray.dir.y - hy.ch * ray.dir.z * ray.dir.z;
    real b = real(2.0f) * (hy.ah * ray.dir.x * ray.origin.x + hy.ah * ray.dir.y * ray.origin.y - hy.ch * ray.dir.z * ray.origin.z);
    real c = hy.ah * ray.origin.x * ray.origin.x + hy.ah * ray.origin.y * ray.origin.y - hy.ch * ray.origin.z * ray.origin.z - real(1.0f);
    real t1, t2, t;
    if (!quadratic(a, b, c, t1, t2)) return result;
    if (t2 < real(-EPSILON)) return result;
    t = t1;
    if (t1 < real(EPSILON)) t = t2;
    vec3 hit = ray.origin + t * ray.dir;
    real zMin = min(hy.p1.z, hy.p2.z);
    real zMax = max(hy.p1.z, hy.p2.z);
    if (hit.z < zMin || hit.z > zMax) {
        if (t == t2) return result;
        t = t2;
        hit = ray.origin + t * ray.dir;
        if (hit.z < zMin || hit.z > zMax) return result;
    }
    if (t >= real(MAX_DISTANCE)) return result;
    real v = (hit.z - hy.p1.z) / (hy.p2.z - hy.p1.z);
    vec3 pr = (real(1.0f) - v) * hy.p1 + v * hy.p2;
    real phi = atan(pr.x * hit.y - hit.x * pr.y, hit.x * pr.x + hit.y * pr.y);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    real u = phi / (real(2.0f) * real(PI));
    result.d = t;
    computeDpDForHyperboloid(hit, hy.p1, hy.p2, phi, result.dpdu, result.dpdv);
    result.normal = normalize(cross(result.dpdu, result.dpdv));
    result.hit = hit;
    result.matIndex = hy.matIndex;
    result.sc = getSurfaceColor(result.hit, vec2(u, v), hy.texIndex);
    result.emission = hy.emission;
    result.hit = TO_WORLD(result.hit) + hy.p;
    result.normal = TO_WORLD(result.normal);
    result.dpdu = TO_WORLD(result.dpdu);
    result.dpdv = TO_WORLD(result.dpdv);
    return result;
}
static inline vec3 sampleHyperboloid(vec2 u, const Hyperboloid& hy, real& pdf) { return BLACK; }

// ---- paraboloid.glsl:1-108 -------------------------------------------------------------------
struct Paraboloid {
    vec3 p;
    real z0, z1, r;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForParaboloid(const Ray& ray, const Paraboloid& pa) {
    real zMin = min(pa.z0, pa.z1);
    real zMax = max(pa.z0, pa.z1);
    Boundbox box(pa.p - vec3(pa.r, -zMin, pa.r), pa.p + vec3(pa.r, zMax, pa.r));
    return testBoundbox(ray, box);
}
static inline Paraboloid parseParaboloid(real index) {
    Paraboloid pa;
    pa.p = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    pa.z0 = readFloat(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    pa.z1 = readFloat(objects, vec2(5.0f, index), OBJECTS_LENGTH);
    pa.r = readFloat(objects, vec2(6.0f, index), OBJECTS_LENGTH);
    pa.reverseNormal = readBool(objects, vec2(7.0f, index), OBJECTS_LENGTH);
    pa.matIndex = rowCoordF(readFloat(objects, vec2(8.0f, index), OBJECTS_LENGTH), tn_);
    pa.texIndex = rowCoordF(readFloat(objects, vec2(9.0f, index), OBJECTS_LENGTH), tn_);
    pa.emission = readVec3(objects, vec2(10.0f, index), OBJECTS_LENGTH);
    return pa;
}
static inline void computeDpDForParaboloid(vec3 hit, real zMax, real zMin, vec3& dpdu, vec3& dpdv) {
    dpdu = vec3(real(-2.0f) * real(PI) * hit.y, real(2.0f) * real(PI) * hit.x, 0.0f);
    dpdv = (zMax - zMin) * vec3(hit.x / (real(2.0f) * hit.z), hit.y / (real(2.0f) * hit.z), 1.0f);
}
static inline vec3 normalForParaboloid(vec3 hit, const Paraboloid& pa) {
    real zMin = min(pa.z0, pa.z1);
    real zMax = max(pa.z0, pa.z1);
    vec3 dpdu, dpdv;
    computeDpDForParaboloid(hit, zMax, zMin, dpdu, dpdv);
    vec3 normal = TO_WORLD(normalize(cross(dpdu, dpdv)));
    return real(pa.reverseNormal ? -1.0f : 1.0f) * normal;
}
static inline Intersect intersectParaboloid(Ray ray, const Paraboloid& pa) {
    Intersect result;
    result.d = MAX_DISTANCE;
    ray.dir = TO_LOCAL(ray.dir);
    ray.origin = TO_LOCAL(ray.origin - pa.p);
    real zMin = min(pa.z0, pa.z1);
    real zMax = max(pa.z0, pa.z1);
    real k = zMax / (pa.r * pa.r);
    real a = k * (ray.dir.x * ray.dir.x + ray.dir.y * ray.dir.y);
    real b = real(2.0f) * k * (ray.dir.x * ray.origin.x + ray.dir.y * ray.origin.y) - ray.dir.z;
    real c = k * (ray.origin.x * ray.origin.x + ray.origin.y * ray.origin.y) - ray.origin.z;
    real t1, t2, t;
    if (!quadratic(a, b, c, t1, t2)) return result;
    if (t2 < real(-EPSILON)) return result;
    t = t1;
    if (t1 < real(EPSILON)) t = t2;
    vec3 hit = ray.origin + t * ray.dir;
    if (hit.z < zMin || hit.z > zMax) {
        if (t == t2) return result;
        t = t2;
        hit = ray.origin + t * ray.dir;
        if (hit.z < zMin || hit.z > zMax) return result;
    }
    if (t >= real(MAX_DISTANCE)) return result;
    real phi = atan(hit.y, hit.x);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    real u = phi / (real(2.0f) * real(PI));
    real v = (hit.z - zMin) / (zMax - zMin);
    result.d = t;
    computeDpDForParaboloid(hit, zMax, zMin, result.dpdu, result.dpdv);
    result.normal = normalize(cross(result.dpdu, result.dpdv));
    result.hit = hit;
    result.matIndex = pa.matIndex;
    result.sc = getSurfaceColor(result.hit, vec2(u, v), pa.texIndex);
    result.emission = pa.emission;
    result.hit = TO_WORLD(result.hit) + pa.p;
    result.normal = TO_WORLD(result.normal);
    result.dpdu = TO_WORLD(result.dpdu);
    result.dpdv = TO_WORLD(result.dpdv);
    return result;
}
static inline vec3 sampleParaboloid(vec2 u, const Paraboloid& pa, real& pdf) { return BLACK; }

// ---- shader.shape.js:28-51 intersectObjects ---------------------------------------------------
// One arm per shape plugin; the reference only compiles the arms of the plugins present in the scene
// (scene.js:86-102), which is every category that occurs in `objects`, so all arms are present here.
#define SAIL_ARM(Type, name)                                                      \
    Type name = parse##Type(rowCoord(i, n_));                                     \
    if (!testBoundboxFor##Type(ray, name)) continue;                              \
    tmp = intersect##Type(ray, name);                                             \
    vec3 nn = real(name.reverseNormal ? -1.0f : 1.0f) * tmp.normal;               \
    bool faceObj = dot(nn, ray.dir) < real(-EPSILON);                             \
    tmp.emission = faceObj ? tmp.emission : BLACK;                                \
    tmp.index = i;

static inline Intersect intersectObjects(const Ray& ray) {
    Intersect ins;
    ins.d = MAX_DISTANCE;
    ins.index = -1;  // struct.glsl:11 leaves it undefined on a miss; -1 = "no hit" for the HIT_ID AOV (Q19)
    for (int i = 0; i < n_; i++) {
        Intersect tmp;
        tmp.d = MAX_DISTANCE;
        int category = toInt(texture(objects, vec2(0.0f, rowCoord(i, n_))));
        if (false) {
        } else if (category == CUBE) { SAIL_ARM(Cube, cube)
        } else if (category == SPHERE) { SAIL_ARM(Sphere, sphere)
        } else if (category == RECTANGLE) { SAIL_ARM(Rectangle, rectangle)
        } else if (category == CONE) { SAIL_ARM(Cone, cone)
        } else if (category == CYLINDER) { SAIL_ARM(Cylinder, cylinder)
        } else if (category == DISK) { SAIL_ARM(Disk, disk)
        } else if (category == HYPERBOLOID) { SAIL_ARM(Hyperboloid, hyperboloid)
        } else if (category == PARABOLOID) { SAIL_ARM(Paraboloid, paraboloid)
        } else if (category == CORNELLBOX) { SAIL_ARM(Cornellbox, cornellbox)
        }
        if (tmp.d < ins.d) ins = tmp;
    }
    ins.matCategory = readInt(texParams, vec2(0.0f, ins.matIndex), TEX_PARAMS_LENGTH);
    ins.into = dot(ins.normal, ray.dir) < real(-EPSILON);
    if (!ins.into) ins.normal = -ins.normal;
    return ins;
}

// ---- shader.shape.js:53-67 sampleGeometry ----------------------------------------------------
#define SAIL_SAMPLE_ARM(Type, name)               \
    Type name = parse##Type(rowCoord(i, n_));     \
    result = sample##Type(u, name, pdf);          \
    normal = normalFor##Type(result, name);

static inline vec3 sampleGeometry(vec2 u, int i, vec3& normal, real& pdf) {
    normal = BLACK;
    pdf = 0.0f;
    int category = toInt(texture(objects, vec2(0.0f, rowCoord(i, n_))));
    vec3 result = BLACK;
    if (false) {
    } else if (category == CUBE) { SAIL_SAMPLE_ARM(Cube, cube)
    } else if (category == SPHERE) { SAIL_SAMPLE_ARM(Sphere, sphere)
    } else if (category == RECTANGLE) { SAIL_SAMPLE_ARM(Rectangle, rectangle)
    } else if (category == CONE) { SAIL_SAMPLE_ARM(Cone, cone)
    } else if (category == CYLINDER) { SAIL_SAMPLE_ARM(Cylinder, cylinder)
    } else if (category == DISK) { SAIL_SAMPLE_ARM(Disk, disk)
    } else if (category == HYPERBOLOID) { SAIL_SAMPLE_ARM(Hyperboloid, hyperboloid)
    } else if (category == PARABOLOID) { SAIL_SAMPLE_ARM(Paraboloid, paraboloid)
    } else if (category == CORNELLBOX) { SAIL_SAMPLE_ARM(Cornellbox, cornellbox)
    }
    return result;
}
#endif

// oracle/shader_shape.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/shape/:
//   boundbox.glsl:1-17, cube.glsl:1-87, cornellbox.glsl:1-90, sphere.glsl:1-92, rectangle.glsl:1-70
//   (this file) and cone/cylinder/disk/hyperboloid/paraboloid + the dispatchers of
//   shader.shape.js:28-67 (shader_shape2.h).
#ifndef SAIL_ORACLE_SHADER_SHAPE_H
#define SAIL_ORACLE_SHADER_SHAPE_H
#include "shader_texture.h"

// ---- boundbox.glsl:1-17 ----------------------------------------------------------------------
struct Boundbox {
    vec3 max, min;
    Boundbox(vec3 a, vec3 b) : max(a), min(b) {}  // constructor order (max,min); callers pass (lo,hi)
};
static inline bool testBoundbox(const Ray& ray, const Boundbox& box) {
    vec3 tMin = (box.min - ray.origin) / ray.dir;
    vec3 tMax = (box.max - ray.origin) / ray.dir;
    vec3 t1 = min(tMin, tMax);
    vec3 t2 = max(tMin, tMax);
    real tNear = max(max(t1.x, t1.y), t1.z);
    real tFar = min(min(t2.x, t2.y), t2.z);
    if (tNear < real(0.0f) && tFar < real(0.0f)) return false;
    return tNear < tFar;
}

// ---- cube.glsl:1-87 --------------------------------------------------------------------------
struct Cube {
    vec3 min, max;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForCube(const Ray& ray, const Cube& cube) { return true; }
static inline Cube parseCube(real index) {
    Cube cube;
    cube.min = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    cube.max = readVec3(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    cube.reverseNormal = readBool(objects, vec2(7.0f, index), OBJECTS_LENGTH);
    cube.matIndex = rowCoordF(readFloat(objects, vec2(8.0f, index), OBJECTS_LENGTH), tn_);
    cube.texIndex = rowCoordF(readFloat(objects, vec2(9.0f, index), OBJECTS_LENGTH), tn_);
    cube.emission = readVec3(objects, vec2(10.0f, index), OBJECTS_LENGTH);
    return cube;
}
static inline vec3 normalForCube(vec3 hit, const Cube& cube) {
    real c = (cube.reverseNormal ? -1.0f : 1.0f);
    if (hit.x < cube.min.x + real(0.0001f)) return c * vec3(-1.0f, 0.0f, 0.0f);
    else if (hit.x > cube.max.x - real(0.0001f)) return c * vec3(1.0f, 0.0f, 0.0f);
    else if (hit.y < cube.min.y + real(0.0001f)) return c * vec3(0.0f, -1.0f, 0.0f);
    else if (hit.y > cube.max.y - real(0.0001f)) return c * vec3(0.0f, 1.0f, 0.0f);
    else if (hit.z < cube.min.z + real(0.0001f)) return c * vec3(0.0f, 0.0f, -1.0f);
    else return c * vec3(0.0f, 0.0f, 1.0f);
}
static inline void computeDpDForCube(vec3 normal, vec3& dpdu, vec3& dpdv) {
    if (abs(normal.x) < real(0.5f)) {
        dpdu = cross(normal, vec3(1.0f, 0.0f, 0.0f));
    } else {
        dpdu = cross(normal, vec3(0.0f, 1.0f, 0.0f));
    }
    dpdv = cross(normal, dpdu);
}
static inline vec3 sampleCube(vec2 u, const Cube& cube, real& pdf) { return BLACK; }  // pdf unwritten (Q7)
static inline vec2 getCubeUV(vec3 hit, const Cube& cube) {
    vec3 tr = cube.max - cube.min;
    hit = hit - cube.min;
    if (hit.x < cube.min.x + real(0.0001f) || hit.x > cube.max.x - real(0.0001f)) return hit.yz() / tr.yz();
    else if (hit.y < cube.min.y + real(0.0001f) || hit.y > cube.max.y - real(0.0001f)) return hit.xz() / tr.xz();
    else return hit.xy() / tr.xy();
}
static inline Intersect intersectCube(Ray ray, const Cube& cube) {
    Intersect result;
    result.d = MAX_DISTANCE;
    vec3 tMin = (cube.min - ray.origin) / ray.dir;
    vec3 tMax = (cube.max - ray.origin) / ray.dir;
    vec3 t1 = min(tMin, tMax);
    vec3 t2 = max(tMin, tMax);
    real tNear = max(max(t1.x, t1.y), t1.z);
    real tFar = min(min(t2.x, t2.y), t2.z);
    real t = -1.0f;
    if (tNear > real(EPSILON) && tNear < tFar) t = tNear;
    else if (tNear < tFar) t = tFar;
    if (t > real(EPSILON)) {
        result.d = t;
        result.hit = ray.origin + t * ray.dir;
        result.normal = normalForCube(ray.origin + t * ray.dir, cube);
        computeDpDForCube(result.normal, result.dpdu, result.dpdv);
        result.matIndex = cube.matIndex;
        result.sc = getSurfaceColor(result.hit, getCubeUV(result.hit, cube), cube.texIndex);
        result.emission = cube.emission;
    }
    return result;
}

// ---- cornellbox.glsl:1-90 --------------------------------------------------------------------
struct Cornellbox {
    vec3 min, max;
    real matIndex;
    bool reverseNormal = false;
    vec3 emission;
};
static inline bool testBoundboxForCornellbox(const Ray& ray, const Cornellbox& box) { return true; }
static inline Cornellbox parseCornellbox(real index) {
    Cornellbox box;
    box.min = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    box.max = readVec3(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    // cornellbox.glsl:17 reads column 7 (= reverseNormal = 0), so the material is texParams row 0 (Q3).
    box.matIndex = rowCoordF(readFloat(objects, vec2(7.0f, index), OBJECTS_LENGTH), tn_);
    box.reverseNormal = false;
    box.emission = BLACK;
    return box;
}
static inline vec3 getCornellboxColor(vec3 hit, vec3 mn, vec3 mx) {
    if (hit.x < mn.x + real(0.0001f)) return GREEN;
    else if (hit.x > mx.x - real(0.0001f)) return BLUE;
    else if (hit.y < mn.y + real(0.0001f)) return WHITE;
    else if (hit.y > mx.y - real(0.0001f)) return WHITE;
    else if (hit.z > mn.z + real(0.0001f)) return WHITE;
    return BLACK;
}
static inline vec3 normalForCornellbox(vec3 hit, const Cornellbox& box) {
    if (hit.x < box.min.x + real(0.0001f)) return vec3(-1.0f, 0.0f, 0.0f);
    else if (hit.x > box.max.x - real(0.0001f)) return vec3(1.0f, 0.0f, 0.0f);
    else if (hit.y < box.min.y + real(0.0001f)) return vec3(0.0f, -1.0f, 0.0f);
    else if (hit.y > box.max.y - real(0.0001f)) return vec3(0.0f, 1.0f, 0.0f);
    else if (hit.z < box.min.z + real(0.0001f)) return vec3(0.0f, 0.0f, -1.0f);
    else return vec3(0.0f, 0.0f, 1.0f);
}
static inline void computeDpDForCornellbox(vec3 normal, vec3& dpdu, vec3& dpdv) { computeDpDForCube(normal, dpdu, dpdv); }
static inline vec3 sampleCornellbox(vec2 u, const Cornellbox& box, real& pdf) { return BLACK; }
static inline Intersect intersectCornellbox(Ray ray, const Cornellbox& box) {
    Intersect result;
    result.d = MAX_DISTANCE;
    vec3 tMin = (box.min - ray.origin) / ray.dir;
    vec3 tMax = (box.max - ray.origin) / ray.dir;
    vec3 t1 = min(tMin, tMax);
    vec3 t2 = max(tMin, tMax);
    real tNear = max(max(t1.x, t1.y), t1.z);
    real tFar = min(min(t2.x, t2.y), t2.z);
    real t = -1.0f;
    if (tNear < tFar) t = tFar;
    if (t > real(EPSILON)) {
        result.d = t;
        result.hit = ray.origin + t * ray.dir;
        result.normal = -normalForCornellbox(ray.origin + t * ray.dir, box);
        computeDpDForCornellbox(result.normal, result.dpdu, result.dpdv);
        result.matIndex = box.matIndex;
        result.sc = getCornellboxColor(result.hit, box.min, box.max);
        result.emission = BLACK;
    }
    return result;
}

// worldToLocal / localToWorld with the constant object frame N=(0,1,0), S=(0,0,-1), T=(1,0,0)
// (define.glsl:62-64) are constant-folded the way a shader compiler folds x*0, x*1 and x*-1:
// they reduce to the swizzles below (value-identical for finite inputs; keeps OPS_alg honest).
static inline vec3 objToLocal(vec3 v) { return vec3(-v.z, v.x, v.y); }
static inline vec3 objToWorld(vec3 v) { return vec3(v.y, v.z, -v.x); }
#define TO_LOCAL(v) objToLocal(v)
#define TO_WORLD(v) objToWorld(v)

// ---- sphere.glsl:1-92 ------------------------------------------------------------------------
struct Sphere {
    vec3 c;
    real r;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForSphere(const Ray& ray, const Sphere& sphere) {
    Boundbox box(sphere.c - vec3(sphere.r), sphere.c + vec3(sphere.r));
    return testBoundbox(ray, box);
}
static inline Sphere parseSphere(real index) {
    Sphere sphere;
    sphere.c = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    sphere.r = readFloat(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    sphere.reverseNormal = readBool(objects, vec2(5.0f, index), OBJECTS_LENGTH);
    sphere.matIndex = rowCoordF(readFloat(objects, vec2(6.0f, index), OBJECTS_LENGTH), tn_);
    sphere.texIndex = rowCoordF(readFloat(objects, vec2(7.0f, index), OBJECTS_LENGTH), tn_);
    sphere.emission = readVec3(objects, vec2(8.0f, index), OBJECTS_LENGTH);
    return sphere;
}
static inline vec3 normalForSphere(vec3 hit, const Sphere& sphere) {
    return real(sphere.reverseNormal ? -1.0f : 1.0f) * (hit - sphere.c) / sphere.r;
}
static inline void computeDpDForSphere(vec3 hit, real radius, vec3& dpdu, vec3& dpdv) {
    real theta = acos(clamp(hit.z / radius, real(-1.0f), real(1.0f)));
    real zRadius = sqrt(hit.x * hit.x + hit.y * hit.y);
    real invZRadius = real(1.0f) / zRadius;
    real cosPhi_ = hit.x * invZRadius;
    real sinPhi_ = hit.y * invZRadius;
    dpdu = vec3(real(-2.0f) * real(PI) * hit.y, real(2.0f) * real(PI) * hit.x, 0.0f);
    dpdv = real(PI) * vec3(hit.z * cosPhi_, hit.z * sinPhi_, -radius * sin(theta));
}
static inline Intersect intersectSphere(Ray ray, const Sphere& sphere) {
    Intersect result;
    result.d = MAX_DISTANCE;
    ray.dir = TO_LOCAL(ray.dir);
    ray.origin = TO_LOCAL(ray.origin - sphere.c);
    real a = dot(ray.dir, ray.dir);
    real b = real(2.0f) * dot(ray.origin, ray.dir);
    real c = dot(ray.origin, ray.origin) - sphere.r * sphere.r;
    real t1, t2, t;
    if (!quadratic(a, b, c, t1, t2)) return result;
    if (t2 < real(EPSILON)) return result;
    t = t1;
    if (t1 < real(EPSILON)) t = t2;
    if (t >= real(MAX_DISTANCE)) return result;
    vec3 hit = ray.origin + t * ray.dir;
    if (hit.x == real(0.0f) && hit.y == real(0.0f)) hit.x = real(1e-5f) * sphere.r;
    real phi = atan(hit.y, hit.x);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    real u = phi / (real(2.0f) * real(PI));
    real theta = acos(clamp(hit.z / sphere.r, real(-1.0f), real(1.0f)));
    real v = theta / real(PI);
    result.d = t;
    result.hit = ray.origin + t * ray.dir;
    computeDpDForSphere(result.hit, sphere.r, result.dpdu, result.dpdv);
    result.normal = normalize(cross(result.dpdv, result.dpdu));
    result.matIndex = sphere.matIndex;
    result.sc = getSurfaceColor(result.hit, vec2(u, v), sphere.texIndex);
    result.emission = sphere.emission;
    result.hit = TO_WORLD(result.hit) + sphere.c;
    result.normal = TO_WORLD(result.normal);
    result.dpdu = TO_WORLD(result.dpdu);
    result.dpdv = TO_WORLD(result.dpdv);
    return result;
}
static inline vec3 sampleSphere(vec2 u, const Sphere& sphere, real& pdf) {
    vec3 p = uniformSampleSphere(u);
    pdf = real(INVPI) / (sphere.r * sphere.r);
    return p * sphere.r + sphere.c;
}

// ---- rectangle.glsl:1-70 ---------------------------------------------------------------------
struct Rectangle {
    vec3 min, max;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForRectangle(const Ray& ray, const Rectangle& r) { return true; }
static inline Rectangle parseRectangle(real index) {
    Rectangle r;
    r.min = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    r.max = readVec3(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    r.reverseNormal = readBool(objects, vec2(7.0f, index), OBJECTS_LENGTH);
    r.matIndex = rowCoordF(readFloat(objects, vec2(8.0f, index), OBJECTS_LENGTH), tn_);
    r.texIndex = rowCoordF(readFloat(objects, vec2(9.0f, index), OBJECTS_LENGTH), tn_);
    r.emission = readVec3(objects, vec2(10.0f, index), OBJECTS_LENGTH);
    return r;
}
static inline vec3 normalForRectangle(vec3 hit, const Rectangle& r) {
    vec3 x = vec3(r.max.x - r.min.x, 0.0f, 0.0f);
    vec3 y = vec3(0.0f, (r.max - r.min).yz());
    vec3 normal = normalize(cross(x, y));
    return real(r.reverseNormal ? -1.0f : 1.0f) * normal;
}
static inline Intersect intersectRectangle(Ray ray, const Rectangle& r) {
    Intersect result;
    result.d = MAX_DISTANCE;
    result.dpdu = vec3(r.max.x - r.min.x, 0.0f, 0.0f);
    result.dpdv = vec3(0.0f, (r.max - r.min).yz());
    result.normal = normalize(cross(result.dpdu, result.dpdv));
    real maxX = length(result.dpdu);
    real maxY = length(result.dpdv);
    vec3 ss = result.dpdu / maxX, ts = cross(result.normal, ss);
    ray.dir = worldToLocal(ray.dir, result.normal, ss, ts);
    ray.origin = worldToLocal(ray.origin - r.min, result.normal, ss, ts);
    if (ray.dir.z == real(0.0f)) return result;
    real t = -ray.origin.z / ray.dir.z;
    if (t < real(EPSILON)) return result;
    vec3 hit = ray.origin + t * ray.dir;
    if (hit.x > maxX || hit.y > maxY || hit.x < real(-EPSILON) || hit.y < real(-EPSILON)) return result;
    result.d = t;
    result.matIndex = r.matIndex;
    result.sc = getSurfaceColor(hit, vec2(hit.x / maxX, hit.y / maxY), r.texIndex);
    result.emission = r.emission;
    result.hit = localToWorld(hit, result.normal, ss, ts) + r.min;
    return result;
}
static inline vec3 sampleRectangle(vec2 u, const Rectangle& r, real& pdf) {
    vec3 x = vec3(r.max.x - r.min.x, 0.0f, 0.0f);
    vec3 y = vec3(0.0f, (r.max - r.min).yz());
    pdf = real(1.0f) / (length(x) * length(y));
    return r.min + x * u.x + y * u.y;
}

#include "shader_shape2.h"
#endif

// oracle/shader_shape2.h -- TEST INFRASTRUCTURE ONLY.  Restates src/shader/shape/cone.glsl:1-104,
// cylinder.glsl:1-95, disk.glsl:1-83, hyperboloid.glsl:1-116, paraboloid.glsl:1-108 and the
// intersectObjects / sampleGeometry dispatchers of shader.shape.js:28-67.  Included from shader_shape.h.
#ifndef SAIL_ORACLE_SHADER_SHAPE2_H
#define SAIL_ORACLE_SHADER_SHAPE2_H


// ---- cone.glsl:1-104 -------------------------------------------------------------------------
struct Cone {
    vec3 p;
    real h, r;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForCone(const Ray& ray, const Cone& cone) {
    Boundbox box(cone.p - vec3(cone.r, 0.0f, cone.r), cone.p + vec3(cone.r, cone.h, cone.r));
    return testBoundbox(ray, box);
}
static inline Cone parseCone(real index) {
    Cone cone;
    cone.p = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    cone.h = readFloat(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    cone.r = readFloat(objects, vec2(5.0f, index), OBJECTS_LENGTH);
    cone.reverseNormal = readBool(objects, vec2(6.0f, index), OBJECTS_LENGTH);
    cone.matIndex = rowCoordF(readFloat(objects, vec2(7.0f, index), OBJECTS_LENGTH), tn_);
    cone.texIndex = rowCoordF(readFloat(objects, vec2(8.0f, index), OBJECTS_LENGTH), tn_);
    cone.emission = readVec3(objects, vec2(9.0f, index), OBJECTS_LENGTH);
    return cone;
}
static inline void computeDpDForCone(vec3 hit, real h, vec3& dpdu, vec3& dpdv) {
    real v = hit.z / h;
    dpdu = vec3(real(-2.0f) * real(PI) * hit.y, real(2.0f) * real(PI) * hit.x, 0.0f);
    dpdv = vec3(-hit.x / (real(1.0f) - v), -hit.y / (real(1.0f) - v), h);
}
static inline vec3 normalForCone(vec3 hit, const Cone& cone) {
    hit = hit - cone.p;
    real tana = cone.r / cone.h;
    real d = sqrt(hit.x * hit.x + hit.y * hit.y);
    real x1 = d / tana;
    real x2 = d * tana;
    vec3 no = vec3(0.0f, 0.0f, cone.h - x1 - x2);
    return real(cone.reverseNormal ? -1.0f : 1.0f) * normalize(hit - no);
}
static inline Intersect intersectCone(Ray ray, const Cone& cone) {
    Intersect result;
    result.d = MAX_DISTANCE;
    ray.dir = TO_LOCAL(ray.dir);
    ray.origin = TO_LOCAL(ray.origin - cone.p);
    real k = cone.r / cone.h;
    k = k * k;
    real a = ray.dir.x * ray.dir.x + ray.dir.y * ray.dir.y - k * ray.dir.z * ray.dir.z;
    real b = real(2.0f) * (ray.dir.x * ray.origin.x + ray.dir.y * ray.origin.y - k * ray.dir.z * (ray.origin.z - cone.h));
    real c = ray.origin.x * ray.origin.x + ray.origin.y * ray.origin.y - k * (ray.origin.z - cone.h) * (ray.origin.z - cone.h);
    real t1, t2, t;
    if (!quadratic(a, b, c, t1, t2)) return result;
    if (t2 < real(-EPSILON)) return result;
    t = t1;
    if (t1 < real(EPSILON)) t = t2;
    vec3 hit = ray.origin + t * ray.dir;
    if (hit.z < real(-EPSILON) || hit.z > cone.h) {
        if (t == t2) return result;
        t = t2;
        hit = ray.origin + t * ray.dir;
        if (hit.z < real(-EPSILON) || hit.z > cone.h) return result;
    }
    if (t >= real(MAX_DISTANCE)) return result;
    real phi = atan(hit.y, hit.x);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    real u = phi / (real(2.0f) * real(PI));
    real v = hit.z / cone.h;
    result.d = t;
    computeDpDForCone(hit, cone.h, result.dpdu, result.dpdv);
    result.normal = normalize(cross(result.dpdu, result.dpdv));
    result.hit = hit;
    result.matIndex = cone.matIndex;
    result.sc = getSurfaceColor(result.hit, vec2(u, v), cone.texIndex);
    result.emission = cone.emission;
    result.hit = TO_WORLD(result.hit) + cone.p;
    result.normal = TO_WORLD(result.normal);
    result.dpdu = TO_WORLD(result.dpdu);
    result.dpdv = TO_WORLD(result.dpdv);
    return result;
}
static inline vec3 sampleCone(vec2 u, const Cone& cone, real& pdf) { return BLACK; }

// ---- cylinder.glsl:1-95 ----------------------------------------------------------------------
struct Cylinder {
    vec3 p;
    real h, r;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForCylinder(const Ray& ray, const Cylinder& cy) {
    Boundbox box(cy.p - vec3(cy.r, 0.0f, cy.r), cy.p + vec3(cy.r, cy.h, cy.r));
    return testBoundbox(ray, box);
}
static inline Cylinder parseCylinder(real index) {
    Cylinder cy;
    cy.p = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    cy.h = readFloat(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    cy.r = readFloat(objects, vec2(5.0f, index), OBJECTS_LENGTH);
    cy.reverseNormal = readBool(objects, vec2(6.0f, index), OBJECTS_LENGTH);
    cy.matIndex = rowCoordF(readFloat(objects, vec2(7.0f, index), OBJECTS_LENGTH), tn_);
    cy.texIndex = rowCoordF(readFloat(objects, vec2(8.0f, index), OBJECTS_LENGTH), tn_);
    cy.emission = readVec3(objects, vec2(9.0f, index), OBJECTS_LENGTH);
    return cy;
}
static inline void computeDpDForCylinder(vec3 hit, real h, vec3& dpdu, vec3& dpdv) {
    dpdu = vec3(real(-2.0f) * real(PI) * hit.y, real(2.0f) * real(PI) * hit.x, 0.0f);
    dpdv = vec3(0.0f, 0.0f, h);
}
static inline vec3 normalForCylinder(vec3 hit, const Cylinder& cy) {
    return real(cy.reverseNormal ? -1.0f : 1.0f) * normalize(vec3(hit.x - cy.p.x, hit.y - cy.p.y, 0.0f));
}
static inline Intersect intersectCylinder(Ray ray, const Cylinder& cy) {
    Intersect result;
    result.d = MAX_DISTANCE;
    ray.dir = TO_LOCAL(ray.dir);
    ray.origin = TO_LOCAL(ray.origin - cy.p);
    real a = ray.dir.x * ray.dir.x + ray.dir.y * ray.dir.y;
    real b = real(2.0f) * (ray.dir.x * ray.origin.x + ray.dir.y * ray.origin.y);
    real c = ray.origin.x * ray.origin.x + ray.origin.y * ray.origin.y - cy.r * cy.r;
    real t1, t2, t;
    if (!quadratic(a, b, c, t1, t2)) return result;
    if (t2 < real(-EPSILON)) return result;
    t = t1;
    if (t1 < real(EPSILON)) t = t2;
    vec3 hit = ray.origin + t * ray.dir;
    if (hit.z < real(-EPSILON) || hit.z > cy.h) {
        if (t == t2) return result;
        t = t2;
        hit = ray.origin + t * ray.dir;
        if (hit.z < real(-EPSILON) || hit.z > cy.h) return result;
    }
    if (t >= real(MAX_DISTANCE)) return result;
    real phi = atan(hit.y, hit.x);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    real u = phi / (real(2.0f) * real(PI));
    real v = hit.z / cy.h;
    result.d = t;
    computeDpDForCylinder(hit, cy.h, result.dpdu, result.dpdv);
    result.normal = normalize(cross(result.dpdu, result.dpdv));
    result.hit = hit;
    result.matIndex = cy.matIndex;
    result.sc = getSurfaceColor(result.hit, vec2(u, v), cy.texIndex);
    result.emission = cy.emission;
    result.hit = TO_WORLD(result.hit) + cy.p;
    result.normal = TO_WORLD(result.normal);
    result.dpdu = TO_WORLD(result.dpdu);
    result.dpdv = TO_WORLD(result.dpdv);
    return result;
}
static inline vec3 sampleCylinder(vec2 u, const Cylinder& cy, real& pdf) { return BLACK; }

// ---- disk.glsl:1-83 --------------------------------------------------------------------------
struct Disk {
    vec3 p;
    real r, innerR;
    real matIndex, texIndex;
    vec3 emission;
    bool reverseNormal = false;
};
static inline bool testBoundboxForDisk(const Ray& ray, const Disk& disk) { return true; }
static inline Disk parseDisk(real index) {
    Disk disk;
    disk.p = readVec3(objects, vec2(1.0f, index), OBJECTS_LENGTH);
    disk.r = readFloat(objects, vec2(4.0f, index), OBJECTS_LENGTH);
    disk.innerR = readFloat(objects, vec2(5.0f, index), OBJECTS_LENGTH);
    disk.reverseNormal = readBool(objects, vec2(6.0f, index), OBJECTS_LENGTH);
    disk.matIndex = rowCoordF(readFloat(objects, vec2(7.0f, index), OBJECTS_LENGTH), tn_);
    disk.texIndex = rowCoordF(readFloat(objects, vec2(8.0f, index), OBJECTS_LENGTH), tn_);
    disk.emission = readVec3(objects, vec2(9.0f, index), OBJECTS_LENGTH);
    return disk;
}
static inline void computeDpDForDisk(vec3 hit, real r, real innerR, real dist2, vec3& dpdu, vec3& dpdv) {
    dpdu = vec3(real(-2.0f) * real(PI) * hit.y, real(2.0f) * real(PI) * hit.x, 0.0f);
    dpdv = vec3(hit.x, hit.y, 0.0f) * (innerR - r) / sqrt(dist2);
}
static inline vec3 normalForDisk(vec3 hit, const Disk& disk) {
    return real(disk.reverseNormal ? -1.0f : 1.0f) * vec3(0.0f, 1.0f, 0.0f);
}
static inline Intersect intersectDisk(Ray ray, const Disk& disk) {
    Intersect result;
    result.d = MAX_DISTANCE;
    ray.dir = TO_LOCAL(ray.dir);
    ray.origin = TO_LOCAL(ray.origin - disk.p);
    if (ray.dir.z == real(0.0f)) return result;
    real t = -ray.origin.z / ray.dir.z;
    if (t <= real(0.0f)) return result;
    vec3 hit = ray.origin + t * ray.dir;
    real dist2 = hit.x * hit.x + hit.y * hit.y;
    if (dist2 > disk.r * disk.r || dist2 < disk.innerR * disk.innerR) return result;
    if (t >= real(MAX_DISTANCE)) return result;
    real phi = atan(hit.y, hit.x);
    if (phi < real(0.0f)) phi += real(2.0f) * real(PI);
    real u = phi / (real(2.0f) * real(PI));
    real rHit = sqrt(dist2);
    real oneMinusV = ((rHit - disk.innerR) / (disk.r - disk.innerR));
    real v = real(1.0f) - oneMinusV;
    result.d = t;
    computeDpDForDisk(hit, disk.r, disk.innerR, dist2, result.dpdu, result.dpdv);
    result.normal = normalize(cross(result.dpdu, result.dpdv));
    result.hit = hit;
    result.matIndex = disk.matIndex;
    result.sc = getSurfaceColor(result.hit, vec2(u, v), disk.texIndex);
    result.emission = disk.emission;
    result.hit = TO_WORLD(result.hit) + disk.p;
    result.normal = TO_WORLD(result.normal);
    result.dpdu = TO_WORLD(result.dpdu);
    result.dpdv = TO_WORLD(result.dpdv);
    return result;
}
static inline vec3 sampleDisk(vec2 u, const Disk& disk, real& pdf) {
    vec2 pd = concentricSampleDisk(u);
    vec3 p = vec3(pd.x * disk.r + disk.p.x, disk.p.y, pd.y * disk.r + disk.p.z);
    real area = real(2.0f) * real(PI) * real(0.5f) * (disk.r * disk.r - disk.innerR * disk.innerR);
    pdf = real(1.0f) / area;
    return p;
}

#include "shader_shape3.h"
#endif

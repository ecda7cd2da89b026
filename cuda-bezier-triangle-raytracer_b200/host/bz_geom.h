// bz_geom.h -- FP32 3-vector / 3x3 arithmetic of the host preprocessing, with the evaluation order
// the reference gets from Eigen's fixed-size kernels (reference/3dGeomUtil.h:23-27 typedefs):
// three-term reductions are a0 + (a1 + a2), normalisation divides by sqrt(squaredNorm) and leaves a
// zero vector alone, the 3x3 inverse is the cofactor form scaled by 1/det.  Built with
// -ffp-contract=off, so the tables come out bit-identical to the reference's -O2 x86 build.
#ifndef BZ_GEOM_H
#define BZ_GEOM_H

#include <array>
#include <cmath>
#include <cstdint>

namespace bz {

struct Vec3 {
  float x, y, z;
  float operator[](int aI) const { return aI == 0 ? x : (aI == 1 ? y : z); }
  float &operator[](int aI) { return aI == 0 ? x : (aI == 1 ? y : z); }
};

using Vertex = Vec3;
using Vector = Vec3;
using Triangle = std::array<Vec3, 3u>;

inline Vec3 vec(float aX, float aY, float aZ) { return Vec3{aX, aY, aZ}; }
inline Vec3 operator+(Vec3 const &aA, Vec3 const &aB) { return Vec3{aA.x + aB.x, aA.y + aB.y, aA.z + aB.z}; }
inline Vec3 operator-(Vec3 const &aA, Vec3 const &aB) { return Vec3{aA.x - aB.x, aA.y - aB.y, aA.z - aB.z}; }
inline Vec3 operator-(Vec3 const &aA) { return Vec3{-aA.x, -aA.y, -aA.z}; }
inline Vec3 operator*(Vec3 const &aA, float aF) { return Vec3{aA.x * aF, aA.y * aF, aA.z * aF}; }
inline Vec3 operator*(float aF, Vec3 const &aA) { return Vec3{aF * aA.x, aF * aA.y, aF * aA.z}; }
inline Vec3 operator/(Vec3 const &aA, float aF) { return Vec3{aA.x / aF, aA.y / aF, aA.z / aF}; }
inline bool operator==(Vec3 const &aA, Vec3 const &aB) { return aA.x == aB.x && aA.y == aB.y && aA.z == aB.z; }
inline bool operator!=(Vec3 const &aA, Vec3 const &aB) { return !(aA == aB); }

inline float dot(Vec3 const &aA, Vec3 const &aB) { return aA.x * aB.x + (aA.y * aB.y + aA.z * aB.z); }
inline float squaredNorm(Vec3 const &aA) { return dot(aA, aA); }
inline float norm(Vec3 const &aA) { return std::sqrt(squaredNorm(aA)); }
inline Vec3 cross(Vec3 const &aA, Vec3 const &aB) {
  return Vec3{aA.y * aB.z - aA.z * aB.y, aA.z * aB.x - aA.x * aB.z, aA.x * aB.y - aA.y * aB.x};
}
inline Vec3 normalized(Vec3 const &aA) {
  float const z = squaredNorm(aA);
  return z > 0.0f ? aA / std::sqrt(z) : aA;
}

// row-major 3x3
struct Mat3 {
  float m[3][3];
  static Mat3 identity() { return Mat3{{{1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}}}; }
};

inline Vec3 operator*(Mat3 const &aM, Vec3 const &aV) {
  return Vec3{aM.m[0][0] * aV.x + (aM.m[0][1] * aV.y + aM.m[0][2] * aV.z),
              aM.m[1][0] * aV.x + (aM.m[1][1] * aV.y + aM.m[1][2] * aV.z),
              aM.m[2][0] * aV.x + (aM.m[2][1] * aV.y + aM.m[2][2] * aV.z)};
}

inline Mat3 operator*(Mat3 const &aM, float aF) {
  Mat3 result;
  for(int r = 0; r < 3; ++r) { for(int c = 0; c < 3; ++c) { result.m[r][c] = aM.m[r][c] * aF; } }
  return result;
}

// signed cofactor with cyclic indices, inverse(r, c) = cofactor(c, r) / det
inline Mat3 inverse(Mat3 const &aM) {
  auto cof = [&aM](int aI, int aJ) {
    int const i1 = (aI + 1) % 3, i2 = (aI + 2) % 3, j1 = (aJ + 1) % 3, j2 = (aJ + 2) % 3;
    return aM.m[i1][j1] * aM.m[i2][j2] - aM.m[i1][j2] * aM.m[i2][j1];
  };
  float const c00 = cof(0, 0), c10 = cof(1, 0), c20 = cof(2, 0);
  float const det = c00 * aM.m[0][0] + (c10 * aM.m[1][0] + c20 * aM.m[2][0]);
  float const inv = 1.0f / det;
  Mat3 result;
  result.m[0][0] = c00 * inv;       result.m[0][1] = c10 * inv;       result.m[0][2] = c20 * inv;
  result.m[1][0] = cof(0, 1) * inv; result.m[1][1] = cof(1, 1) * inv; result.m[1][2] = cof(2, 1) * inv;
  result.m[2][0] = cof(0, 2) * inv; result.m[2][1] = cof(1, 2) * inv; result.m[2][2] = cof(2, 2) * inv;
  return result;
}

// Plane: point . normal == constant  (reference/3dGeomUtil.h:218-334)
struct Plane {
  Vec3  normal;
  float constant;

  float distance(Vec3 const &aPoint) const { return dot(aPoint, normal) - constant; }          // 3dGeomUtil.h:307
  void makeDistancePositive(Vec3 const &aPoint) {                                                // 3dGeomUtil.h:310-317
    if(distance(aPoint) < 0.0f) { normal = -normal; constant = -constant; }
  }
  // 3dGeomUtil.h:231-236: normal along p0 -> p1, through p1 * proportion + p0 * (1 - proportion)
  static Plane from1proportion2points(float aProportion, Vec3 const &aP0, Vec3 const &aP1) {
    Plane result;
    result.normal = normalized(aP1 - aP0);
    result.constant = dot(result.normal, aP1 * aProportion + aP0 * (1.0f - aProportion));
    return result;
  }
  static Plane from3points(Vec3 const &aP0, Vec3 const &aP1, Vec3 const &aP2) {                  // 3dGeomUtil.h:239-244
    Plane result;
    result.normal = normalized(cross(aP1 - aP0, aP2 - aP0));
    result.constant = dot(result.normal, aP0);
    return result;
  }
  static Plane from1vector2points(Vec3 const &aDirection, Vec3 const &aP0, Vec3 const &aP1) {    // 3dGeomUtil.h:250-255
    Plane result;
    result.normal = normalized(cross(aDirection, aP1 - aP0));
    result.constant = dot(result.normal, aP0);
    return result;
  }
  // common point of three planes (3dGeomUtil.h:266-275): rows = normals, inverse * constants
  static Vec3 meet(Plane const &aA, Plane const &aB, Plane const &aC) {
    Mat3 const rows{{{aA.normal.x, aA.normal.y, aA.normal.z}, {aB.normal.x, aB.normal.y, aB.normal.z}, {aC.normal.x, aC.normal.y, aC.normal.z}}};
    return inverse(rows) * Vec3{aA.constant, aB.constant, aC.constant};
  }
};

inline Vec3 faceNormal(Vec3 const &aV0, Vec3 const &aV1, Vec3 const &aV2) { return cross(aV1 - aV0, aV2 - aV0); }   // util::getNormal, 3dGeomUtil.h:33-40
inline Vec3 faceNormal(Triangle const &aFace) { return faceNormal(aFace[0], aFace[1], aFace[2]); }

// util::getAltitude (3dGeomUtil.h:123-128)
inline Vec3 altitude(Vec3 const &aCommon1, Vec3 const &aCommon2, Vec3 const &aIndependent) {
  Vec3 const common = aCommon2 - aCommon1;
  Vec3 const independent = aIndependent - aCommon1;
  float const foot = dot(common, independent) / squaredNorm(common);
  return independent - common * foot;
}

}

#endif

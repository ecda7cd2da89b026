// TEST INFRASTRUCTURE ONLY -- CPU restatement of the renderer stages that have NO reference
// counterpart (reference/rayTracer.cpp is an empty file): primary-ray generation (K1) and
// background shading (K3).  The definitions are the ones DESIGN.md section "Renderer definitions"
// states; the CUDA kernels implement the same formulas independently.  FP32 throughout, no fused
// multiply-add (build without -mfma), 3-term sums as a0 + (a1 + a2) like the reference's Eigen
// reductions, direction normalised like Ray::Ray (reference/3dGeomUtil.h:176-178: v / sqrt(v.v)).
#ifndef BZ_ORACLE_RENDER_DEF_H
#define BZ_ORACLE_RENDER_DEF_H

#include <cmath>
#include <cstdint>
#include "bz_capi.h"

namespace bzdef {

inline float dot3(float const *aA, float const *aB) {
  return aA[0] * aB[0] + (aA[1] * aB[1] + aA[2] * aB[2]);
}

inline void primaryRay(bz_camera const &aCamera, uint32_t const aPx, uint32_t const aPy, uint32_t const aSx, uint32_t const aSy,
                       float *aStart, float *aDir) {
  float const strata = static_cast<float>(aCamera.strata);
  float const fu = static_cast<float>(aPx) + (static_cast<float>(aSx) + 0.5f) / strata;
  float const fv = static_cast<float>(aPy) + (static_cast<float>(aSy) + 0.5f) / strata;
  float const a = (fu * (2.0f / static_cast<float>(aCamera.width)) - 1.0f) * aCamera.half_w;
  float const b = (1.0f - fv * (2.0f / static_cast<float>(aCamera.height))) * aCamera.half_h;
  float offset[3];
  for(int k = 0; k < 3; ++k) { offset[k] = aCamera.right[k] * a + aCamera.up[k] * b; }
  float raw[3];
  if(aCamera.ortho != 0u) {
    for(int k = 0; k < 3; ++k) {
      aStart[k] = aCamera.eye[k] + offset[k];
      raw[k] = aCamera.forward[k];
    }
  }
  else {
    for(int k = 0; k < 3; ++k) {
      aStart[k] = aCamera.eye[k];
      raw[k] = aCamera.forward[k] + offset[k];
    }
  }
  float const z = dot3(raw, raw);
  if(z > 0.0f) {
    float const norm = std::sqrt(z);
    for(int k = 0; k < 3; ++k) { aDir[k] = raw[k] / norm; }
  }
  else {
    for(int k = 0; k < 3; ++k) { aDir[k] = raw[k]; }
  }
}

// Plane::intersect semantics (reference/3dGeomUtil.h:279-296): |cos| >= 1e-5, distance > 0.
inline uint32_t shadeBackground(bz_scene const &aScene, float const *aStart, float const *aDir) {
  float const cosIncidence = dot3(aDir, aScene.bg_plane);
  if(!(std::fabs(cosIncidence) >= 0.00001f)) { return aScene.value_lost; }
  float const distance = (aScene.bg_plane[3] - dot3(aScene.bg_plane, aStart)) / cosIncidence;
  if(!(distance > 0.0f)) { return aScene.value_lost; }
  float point[3];
  for(int k = 0; k < 3; ++k) { point[k] = aStart[k] + distance * aDir[k]; }
  float const ca = std::floor(dot3(point, aScene.bg_axis_a) / aScene.cell);
  float const cb = std::floor(dot3(point, aScene.bg_axis_b) / aScene.cell);
  // parity of floor(a) + floor(b) without leaving float range: fmod on exact integers is exact
  float const pa = std::fabs(std::fmod(ca, 2.0f));
  float const pb = std::fabs(std::fmod(cb, 2.0f));
  return (pa != pb) ? aScene.value_bright : aScene.value_dark;
}

}

#endif

// TEST INFRASTRUCTURE ONLY -- C wrapper around the REAL reference classes.
//
// Compiled by oracle/build_ref.sh together with the reference's own mesh.cpp, bezierTriangle.cpp,
// bezierMesh.cpp and bezierLens.cpp, read from where they lie under /root/reference, against
// oracle/eigen_shim (Eigen is absent here) into oracle/_ref/libbzref*.so.  Nothing of the product
// (cuda-bezier-triangle-raytracer_b200/) links or loads this; only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs use it, as the checker / the timed CPU arm.
//
// Everything that is arithmetic on the hot path is done by the reference's own code:
// Mesh::standardizeVertices/standardizeNormals, BezierMesh::BezierMesh, BezierTriangle::intersect,
// BezierMesh::intersect, BezierLens::refract, UniformHemisphere (belt table and getRandom).  What this file adds is
//   * table dumping (private BezierTriangle fields -> bz_patch),
//   * bzref_intersect_batch: BezierMesh::intersect does not report WHICH patch was hit
//     (bezierMesh.cpp:206-227 returns no index), so the 20-line loop is repeated here around the
//     public per-patch BezierTriangle::intersect to recover it, and cross-checked against the
//     real BezierMesh::intersect result on every ray,
//   * the CPU restatement of the renderer stages the reference never wrote (ray generation,
//     bounce loop shaped like test.cpp:376-398, background shading), std::thread over rows.
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <cmath>
#include <string>
#include <vector>
#include <array>
#include <map>
#include <list>
#include <deque>
#include <tuple>
#include <thread>
#include <atomic>
#include <fstream>
#include <sstream>
#include <iostream>
#include <numeric>
#include <algorithm>
#include <exception>
#include <stdexcept>
#include <functional>
#include <unordered_map>
#include <unordered_set>
#include <random>
#include "stl_reader.h"

// Expose BezierTriangle's private fields to the table dump below (layout is unaffected).
#define private public
#include "bezierLens.h"
#include "hostUtil.h"
#undef private

#include "bz_capi.h"
#include "render_def.h"

namespace {

thread_local std::string gError;

void dumpPatch(BezierTriangle const &aIn, bz_patch &aOut) {
  std::memset(&aOut, 0, sizeof(aOut));
  for(int i = 0; i < 3; ++i) { aOut.plane[i] = aIn.mUnderlyingPlane.mNormal(i); }
  aOut.plane[3] = aIn.mUnderlyingPlane.mConstant;
  for(int p = 0; p < 3; ++p) {
    for(int i = 0; i < 3; ++i) { aOut.divider[p][i] = aIn.mNeighbourDividerPlanes[p].mNormal(i); }
    aOut.divider[p][3] = aIn.mNeighbourDividerPlanes[p].mConstant;
    aOut.neighbours[p] = aIn.mNeighbours[p];
  }
  for(int c = 0; c < 10; ++c) {
    for(int i = 0; i < 3; ++i) { aOut.control[c][i] = aIn.mControlPoints[c](i); }
  }
  for(int r = 0; r < 3; ++r) {
    for(int c = 0; c < 3; ++c) { aOut.bary_inv[r][c] = aIn.mBarycentricInverse(r, c); }
  }
  aOut.height_inside = aIn.mHeightInside;
  aOut.height_outside = aIn.mHeightOutside;
  for(int i = 0; i < 3; ++i) { aOut.dir_b[i] = aIn.mBezierDerivativeDirectionVectorB(i); }
}

Ray makeRay(float const *aStart, float const *aDir) {
  Ray ray;                              // no normalisation: the batch carries unit directions already
  ray.mStart = Vertex(aStart[0], aStart[1], aStart[2]);
  ray.mDirection = Vector(aDir[0], aDir[1], aDir[2]);
  return ray;
}

template<typename tLambda>
void parallelFor(uint64_t const aCount, int aThreads, tLambda &&aBody) {
  if(aThreads < 1) { aThreads = 1; }
  if(aThreads == 1 || aCount < 2u) {
    for(uint64_t i = 0u; i < aCount; ++i) { aBody(i); }
    return;
  }
  std::atomic<uint64_t> next{0u};
  uint64_t const chunk = std::max<uint64_t>(1u, std::min<uint64_t>(64u, aCount / (static_cast<uint64_t>(aThreads) * 4u)));
  std::vector<std::thread> pool;
  for(int t = 0; t < aThreads; ++t) {
    pool.emplace_back([&]() {
      for(;;) {
        uint64_t const begin = next.fetch_add(chunk);
        if(begin >= aCount) { break; }
        uint64_t const end = std::min(aCount, begin + chunk);
        for(uint64_t i = begin; i < end; ++i) { aBody(i); }
      }
    });
  }
  for(auto &thread : pool) { thread.join(); }
}

// bezierMesh.cpp:206-227 repeated around the public per-patch API to learn the patch index.
BezierIntersection intersectWithIndex(BezierMesh const &aMesh, Ray const &aRay, uint32_t &aPatch) {
  BezierIntersection result;
  result.mIntersection.mDistance = std::numeric_limits<float>::max();
  result.mWhat = BezierIntersection::What::cNone;
  aPatch = BZ_NO_PATCH;
  uint32_t const count = static_cast<uint32_t>(aMesh.size());
  for(uint32_t i = 0u; i < count; ++i) {
    auto const &bezier = aMesh[i];
    uint32_t evaluated = i;
    auto candidate = bezier.intersect(aRay, BezierTriangle::LimitPlaneIntersection::cThis);
    if(candidate.mWhat == BezierIntersection::What::cFollowSide0 ||
       candidate.mWhat == BezierIntersection::What::cFollowSide1 ||
       candidate.mWhat == BezierIntersection::What::cFollowSide2) {
      evaluated = bezier.getNeighbours()[static_cast<uint32_t>(candidate.mWhat)];
      candidate = aMesh[evaluated].intersect(aRay, BezierTriangle::LimitPlaneIntersection::cNone);
    }
    if(candidate.mWhat == BezierIntersection::What::cIntersect && candidate.mIntersection.mDistance < result.mIntersection.mDistance) {
      result = candidate;
      aPatch = evaluated;
    }
  }
  return result;
}

struct RenderCounters {
  std::atomic<uint64_t> mPrimary{0u}, mCalls{0u}, mEntered{0u}, mExited{0u}, mLost{0u};
};

}

extern "C" {

const char *bzref_last_error(void) { return gError.c_str(); }

// 1 = built with the one-line mPoint hoist, 0 = as written upstream.
int bzref_is_hoisted(void) {
#ifdef BZREF_HOISTED
  return 1;
#else
  return 0;
#endif
}

void *bzref_mesh_new(void) { return new Mesh(); }
void bzref_mesh_free(void *aMesh) { delete static_cast<Mesh*>(aMesh); }
void *bzref_mesh_clone(void *aMesh) { return new Mesh(*static_cast<Mesh*>(aMesh)); }

void bzref_mesh_make_ellipsoid(void *aMesh, int32_t aSectors, int32_t aBelts, float aX, float aY, float aZ) {
  static_cast<Mesh*>(aMesh)->makeEllipsoid(aSectors, aBelts, Vector(aX, aY, aZ));
}

void bzref_mesh_make_unit_sphere(void *aMesh, int32_t aSectors, int32_t aBelts) {
  static_cast<Mesh*>(aMesh)->makeUnitSphere(aSectors, aBelts);
}

int bzref_mesh_read_stl(void *aMesh, char const *aPath) {
  try {
    static_cast<Mesh*>(aMesh)->readMesh(aPath);
    return 0;
  }
  catch(std::exception const &e) {
    gError = e.what();
    return -1;
  }
}

void bzref_mesh_write_stl(void *aMesh, char const *aPath) { static_cast<Mesh*>(aMesh)->writeMesh(aPath); }

void bzref_mesh_set_triangles(void *aMesh, float const *aTriangles, uint32_t aCount) {
  auto &mesh = *static_cast<Mesh*>(aMesh);
  mesh.clear();
  mesh.reserve(aCount);
  for(uint32_t i = 0u; i < aCount; ++i) {
    Triangle triangle;
    for(uint32_t v = 0u; v < 3u; ++v) {
      triangle[v] = Vertex(aTriangles[i * 9u + v * 3u], aTriangles[i * 9u + v * 3u + 1u], aTriangles[i * 9u + v * 3u + 2u]);
    }
    mesh.push_back(triangle);
  }
}

uint32_t bzref_mesh_size(void *aMesh) { return static_cast<uint32_t>(static_cast<Mesh*>(aMesh)->size()); }

void bzref_mesh_get_triangles(void *aMesh, float *aOut) {
  auto const &mesh = *static_cast<Mesh*>(aMesh);
  for(uint32_t i = 0u; i < mesh.size(); ++i) {
    for(uint32_t v = 0u; v < 3u; ++v) {
      for(uint32_t k = 0u; k < 3u; ++k) { aOut[i * 9u + v * 3u + k] = mesh[i][v](k); }
    }
  }
}

void bzref_mesh_translate(void *aMesh, float aX, float aY, float aZ) { *static_cast<Mesh*>(aMesh) += Vector(aX, aY, aZ); }
void bzref_mesh_scale(void *aMesh, float aFactor) { *static_cast<Mesh*>(aMesh) *= aFactor; }
void bzref_mesh_transform(void *aMesh, float const *aRowMajor3x3, float const *aDisplacement) {
  Transform transform;
  for(int r = 0; r < 3; ++r) {
    for(int c = 0; c < 3; ++c) { transform(r, c) = aRowMajor3x3[r * 3 + c]; }
  }
  static_cast<Mesh*>(aMesh)->transform(transform, Vertex(aDisplacement[0], aDisplacement[1], aDisplacement[2]));
}

void bzref_mesh_standardize_vertices(void *aMesh) { static_cast<Mesh*>(aMesh)->standardizeVertices(); }

int bzref_mesh_standardize_normals(void *aMesh) {
  try {
    static_cast<Mesh*>(aMesh)->standardizeNormals();
    return 0;
  }
  catch(char const *text) {           // mesh.cpp:204 throws a char const*
    gError = text;
    return -4;
  }
  catch(std::exception const &e) {
    gError = e.what();
    return -4;
  }
}

// 3 fellow-triangle indices + 3 common-side starts per face (mesh.h:28-32)
void bzref_mesh_get_neighbours(void *aMesh, uint32_t *aFellows, uint8_t *aStarts) {
  auto const &neighbours = static_cast<Mesh*>(aMesh)->getFace2neighbours();
  for(uint32_t i = 0u; i < neighbours.size(); ++i) {
    for(uint32_t k = 0u; k < 3u; ++k) {
      aFellows[i * 3u + k] = neighbours[i].mFellowTriangles[k];
      aStarts[i * 3u + k] = neighbours[i].mFellowCommonSideStarts[k];
    }
  }
}

uint32_t bzref_mesh_neighbour_count(void *aMesh) { return static_cast<uint32_t>(static_cast<Mesh*>(aMesh)->getFace2neighbours().size()); }

// average normal of every triangle corner (3 x 3 floats per face), via getVertex2averageNormals (mesh.h:69)
int bzref_mesh_get_corner_normals(void *aMesh, float *aOut) {
  auto const &mesh = *static_cast<Mesh*>(aMesh);
  auto const &normals = mesh.getVertex2averageNormals();
  try {
    for(uint32_t i = 0u; i < mesh.size(); ++i) {
      for(uint32_t v = 0u; v < 3u; ++v) {
        auto const &normal = normals.at(mesh[i][v]);
        for(uint32_t k = 0u; k < 3u; ++k) { aOut[i * 9u + v * 3u + k] = normal(k); }
      }
    }
    return 0;
  }
  catch(std::exception const &e) {
    gError = e.what();
    return -1;
  }
}

void *bzref_bezier_new(void *aMesh) {
  try {
    return new BezierMesh(*static_cast<Mesh*>(aMesh));
  }
  catch(std::exception const &e) {     // std::out_of_range from .at() when standardizeNormals was skipped (bezierMesh.cpp:17-18)
    gError = e.what();
    return nullptr;
  }
}
void bzref_bezier_free(void *aBezier) { delete static_cast<BezierMesh*>(aBezier); }
uint32_t bzref_bezier_size(void *aBezier) { return static_cast<uint32_t>(static_cast<BezierMesh*>(aBezier)->size()); }

void bzref_bezier_dump(void *aBezier, bz_patch *aOut) {
  auto const &bezier = *static_cast<BezierMesh*>(aBezier);
  for(uint32_t i = 0u; i < bezier.size(); ++i) { dumpPatch(bezier[i], aOut[i]); }
}

uint32_t bzref_sizeof_bezier_triangle(void) { return static_cast<uint32_t>(sizeof(BezierTriangle)); }

// BezierTriangle::interpolate for table/interpolation checks (bezierTriangle.cpp:105-121)
void bzref_bezier_interpolate(void *aBezier, uint32_t aPatch, float aB0, float aB1, float aB2, float *aOut) {
  auto const point = (*static_cast<BezierMesh*>(aBezier))[aPatch].interpolate(aB0, aB1, aB2);
  for(int k = 0; k < 3; ++k) { aOut[k] = point(k); }
}

// Mesh BezierMesh::splitThickBezierTriangles() (bezierMesh.cpp:79-130)
void *bzref_bezier_split_thick(void *aBezier) { return new Mesh(static_cast<BezierMesh*>(aBezier)->splitThickBezierTriangles()); }
// Mesh BezierMesh::interpolate(divisor) (bezierMesh.cpp:55-66)
void *bzref_bezier_tessellate(void *aBezier, int32_t aDivisor) { return new Mesh(static_cast<BezierMesh*>(aBezier)->interpolate(aDivisor)); }

// one patch: BezierTriangle::intersect(ray, limit); returns What
uint32_t bzref_patch_intersect(void *aBezier, uint32_t aPatch, float const *aStart, float const *aDir, uint32_t aLimit,
                               float *aDistance, float *aPoint, float *aNormal, float *aCos, float *aBary) {
  auto const &bezier = *static_cast<BezierMesh*>(aBezier);
  auto const result = bezier[aPatch].intersect(makeRay(aStart, aDir), static_cast<BezierTriangle::LimitPlaneIntersection>(aLimit));
  if(result.mWhat != BezierIntersection::What::cNone) {
    *aDistance = result.mIntersection.mDistance;
    for(int k = 0; k < 3; ++k) {
      aPoint[k] = result.mIntersection.mPoint(k);
      aNormal[k] = result.mNormal(k);
      aBary[k] = result.mBarycentric(k);
    }
    *aCos = (result.mWhat == BezierIntersection::What::cIntersect ? result.mIntersection.mCosIncidence : 0.0f);
  }
  return static_cast<uint32_t>(result.mWhat);
}

// BezierMesh::intersect over a batch + the index of the hit patch; returns the number of rays on
// which the index-recovering loop and the real BezierMesh::intersect disagreed (must be 0).
uint64_t bzref_intersect_batch(void *aBezier, float const *aStart, float const *aDir, uint64_t aCount,
                               uint32_t *aWhat, uint32_t *aPatch, float *aDistance, float *aPoint, float *aNormal,
                               float *aCos, float *aBary, int aThreads) {
  auto const &bezier = *static_cast<BezierMesh*>(aBezier);
  std::atomic<uint64_t> mismatches{0u};
  parallelFor(aCount, aThreads, [&](uint64_t const i) {
    Ray const ray = makeRay(aStart + i * 3u, aDir + i * 3u);
    uint32_t patch;
    auto const mine = intersectWithIndex(bezier, ray, patch);
    auto const real = bezier.intersect(ray);
    bool same = (mine.mWhat == real.mWhat);
    if(same && real.mWhat == BezierIntersection::What::cIntersect) {
      same = std::memcmp(&mine.mIntersection.mDistance, &real.mIntersection.mDistance, sizeof(float)) == 0 &&
             std::memcmp(mine.mIntersection.mPoint.data(), real.mIntersection.mPoint.data(), 3u * sizeof(float)) == 0 &&
             std::memcmp(mine.mNormal.data(), real.mNormal.data(), 3u * sizeof(float)) == 0;
    }
    if(!same) { ++mismatches; }
    bool const hit = (real.mWhat == BezierIntersection::What::cIntersect);
    aWhat[i] = hit ? BZ_WHAT_INTERSECT : BZ_WHAT_NONE;
    if(aPatch != nullptr) { aPatch[i] = hit ? patch : BZ_NO_PATCH; }
    if(aDistance != nullptr) { aDistance[i] = hit ? real.mIntersection.mDistance : 0.0f; }
    if(aCos != nullptr) { aCos[i] = hit ? real.mIntersection.mCosIncidence : 0.0f; }
    for(int k = 0; k < 3; ++k) {
      if(aPoint != nullptr) { aPoint[i * 3u + k] = hit ? real.mIntersection.mPoint(k) : 0.0f; }
      if(aNormal != nullptr) { aNormal[i * 3u + k] = hit ? real.mNormal(k) : 0.0f; }
      if(aBary != nullptr) { aBary[i * 3u + k] = hit ? real.mBarycentric(k) : 0.0f; }
    }
  });
  return mismatches.load();
}

void *bzref_lens_new(float aRi, void *aBezier) { return new BezierLens(aRi, *static_cast<BezierMesh*>(aBezier)); }
void bzref_lens_free(void *aLens) { delete static_cast<BezierLens*>(aLens); }

// BezierLens::refract over a batch (bezierLens.cpp:4-34).  On cNone the output ray repeats the input
// (the reference leaves it partly uninitialised).
void bzref_refract_batch(void *aLens, float const *aStart, float const *aDir, uint32_t const *aExpected, uint64_t aCount,
                         float *aOutStart, float *aOutDir, uint32_t *aOutStatus, int aThreads) {
  auto const &lens = *static_cast<BezierLens*>(aLens);
  parallelFor(aCount, aThreads, [&](uint64_t const i) {
    Ray const ray = makeRay(aStart + i * 3u, aDir + i * 3u);
    auto const [refracted, status] = lens.refract(ray, static_cast<RefractionResult>(aExpected[i]));
    aOutStatus[i] = static_cast<uint32_t>(status);
    for(int k = 0; k < 3; ++k) {
      aOutStart[i * 3u + k] = (status == RefractionResult::cNone ? ray.mStart(k) : refracted.mStart(k));
      aOutDir[i * 3u + k] = (status == RefractionResult::cNone ? ray.mDirection(k) : refracted.mDirection(k));
    }
  });
}

// Primary rays exactly as DESIGN.md defines them (CPU restatement, render_def.h).
void bzref_generate_rays(bz_camera const *aCamera, uint32_t aRow0, uint32_t aRow1, float *aStart, float *aDir) {
  uint32_t const spp = aCamera->strata * aCamera->strata;
  uint64_t i = 0u;
  for(uint32_t py = aRow0; py < aRow1; ++py) {
    for(uint32_t px = 0u; px < aCamera->width; ++px) {
      for(uint32_t s = 0u; s < spp; ++s) {
        bzdef::primaryRay(*aCamera, px, py, s % aCamera->strata, s / aCamera->strata, aStart + i * 3u, aDir + i * 3u);
        ++i;
      }
    }
  }
}

// CPU renderer: ray generation + bounce loop around the REAL BezierLens::refract + shading.
// aPixelStep > 1 renders only pixels with px % step == 0 && py % step == 0 (bounded CPU samples);
// the other pixels are left untouched.  Returns the number of primary rays traced.
uint64_t bzref_render(void *aLens, bz_camera const *aCamera, bz_scene const *aScene, uint32_t aRow0, uint32_t aRow1,
                      uint32_t aPixelStep, uint8_t *aRgba, bz_stats *aStats, int aThreads) {
  auto const &lens = *static_cast<BezierLens*>(aLens);
  RenderCounters counters;
  uint32_t const strata = aCamera->strata;
  uint32_t const spp = strata * strata;
  uint32_t const step = (aPixelStep == 0u ? 1u : aPixelStep);
  std::vector<std::pair<uint32_t, uint32_t>> pixels;
  for(uint32_t py = aRow0; py < aRow1; ++py) {
    if(py % step != 0u) { continue; }
    for(uint32_t px = 0u; px < aCamera->width; px += step) { pixels.emplace_back(px, py); }
  }
  parallelFor(pixels.size(), aThreads, [&](uint64_t const aIndex) {
    auto const [px, py] = pixels[aIndex];
    uint32_t sum = 0u;
    uint64_t calls = 0u, entered = 0u, exited = 0u, lost = 0u;
    for(uint32_t s = 0u; s < spp; ++s) {
      float start[3], dir[3];
      bzdef::primaryRay(*aCamera, px, py, s % strata, s / strata, start, dir);
      Ray current = makeRay(start, dir);
      RefractionResult expected = RefractionResult::cInside;
      for(uint32_t crossing = 0u; crossing < aScene->max_crossings; ++crossing) {   // test.cpp:376-398 generalised
        auto const [refracted, status] = lens.refract(current, expected);
        ++calls;
        if(status == RefractionResult::cNone) { break; }
        current = refracted;
        if(status == RefractionResult::cInside) { ++entered; } else { ++exited; }
        expected = (expected == RefractionResult::cInside ? RefractionResult::cOutside : RefractionResult::cInside);
      }
      uint32_t value;
      if(expected == RefractionResult::cInside) {
        float const s3[3] = { current.mStart(0), current.mStart(1), current.mStart(2) };
        float const d3[3] = { current.mDirection(0), current.mDirection(1), current.mDirection(2) };
        value = bzdef::shadeBackground(*aScene, s3, d3);
      }
      else {
        value = aScene->value_lost;
        ++lost;
      }
      sum += value;
    }
    uint8_t const pixel = static_cast<uint8_t>((sum + spp / 2u) / spp);
    uint8_t *out = aRgba + (static_cast<uint64_t>(py - aRow0) * aCamera->width + px) * 4u;
    out[0] = pixel; out[1] = pixel; out[2] = pixel; out[3] = 255u;
    counters.mPrimary += spp;
    counters.mCalls += calls;
    counters.mEntered += entered;
    counters.mExited += exited;
    counters.mLost += lost;
  });
  if(aStats != nullptr) {
    aStats->primary_rays += counters.mPrimary.load();
    aStats->refract_calls += counters.mCalls.load();
    aStats->entered += counters.mEntered.load();
    aStats->exited += counters.mExited.load();
    aStats->lost += counters.mLost.load();
  }
  return counters.mPrimary.load();
}

// ---- UniformHemisphere, reference/hostUtil.h:8-24, hostUtil.cpp:3-29 (illumination mode) ----

// belt table of the REAL constructor: (patch width, patches before the belt) per belt -> patch count
uint32_t bzref_hemisphere(uint32_t aBelts, float *aWidth, uint32_t *aFirst) {
  UniformHemisphere hemisphere(aBelts);
  for(uint32_t i = 0u; i < aBelts; ++i) {
    if(aWidth != nullptr) { aWidth[i] = hemisphere.mBelts[i].first; }
    if(aFirst != nullptr) { aFirst[i] = hemisphere.mBelts[i].second; }
  }
  return hemisphere.getPatchCount();
}

// The REAL getRandom() fed with chosen draws.  std::ranlux24_base is subtract_with_carry_engine<uint_fast32_t, 24, 10, 24>: with
// index 0 and no carry its next two outputs are x[14] - x[0] and x[15] - x[1], and std::uniform_real_distribution<float>
// turns a 24-bit output o into o * 2^-24 * (b - a) + a (one draw, generate_canonical<float, 24>).  The engine state is set
// through its public stream extractor (24 words, carry, index), so the two draws of a call are aDraw1[i], aDraw2[i] (< 2^24):
// u1 = aDraw1 * 2^-24 for the incidence, u2 = aDraw2 * 2^-24 for the turn -- the map the C ABI defines for its own generator.
// -> 0, or the number of calls whose engine did not deliver the chosen draws (self-check)
uint64_t bzref_hemisphere_random(uint32_t aBelts, uint32_t const *aDraw1, uint32_t const *aDraw2, uint64_t aCount, float *aDirection, uint32_t *aPatch) {
  UniformHemisphere hemisphere(aBelts);
  uint64_t wrong = 0u;
  for(uint64_t i = 0u; i < aCount; ++i) {
    std::ostringstream text;
    for(int word = 0; word < 24; ++word) { text << (word == 14 ? aDraw1[i] : (word == 15 ? aDraw2[i] : 0u)) << ' '; }
    text << "0 0";
    std::istringstream state(text.str());
    state >> hemisphere.mRandomGenerator;
    std::ranlux24_base probe = hemisphere.mRandomGenerator;
    if(probe() != aDraw1[i] || probe() != aDraw2[i]) { ++wrong; }
    auto [direction, index] = hemisphere.getRandom();
    aDirection[i * 3u] = direction(0); aDirection[i * 3u + 1u] = direction(1); aDirection[i * 3u + 2u] = direction(2);
    aPatch[i] = index;
  }
  return wrong;
}

}

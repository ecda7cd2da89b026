// bz_bezier_lens.h -- host C++ face of the refraction path: the reference's BezierLens (reference/bezierLens.h:7-28)
// with the same names and argument meaning, implemented on top of the C ABI (include/bz_capi.h).  The lens owns the
// device tables (SoA patch streams + cull hierarchy) of one CUDA device; every query runs in the sm_100a kernels.
// There is no host arithmetic here and no CPU fallback: without a CUDA device the constructor throws.
//
// Single-ray methods keep API parity with the reference (they are batches of one); the batch methods are the real
// entry points.  Header-only; link against libbz_b200.so.
#ifndef BZ_BEZIER_LENS_H
#define BZ_BEZIER_LENS_H

#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>
#include "bz_capi.h"
#include "bz_bezier_mesh.h"
#include "bz_geom.h"

namespace bz {

// reference/3dGeomUtil.h:168-206: the constructor normalises the direction
struct Ray final {
  Vertex mStart;
  Vector mDirection;
  Ray() = default;
  Ray(Vertex const &aStart, Vector const &aDirection) : mStart(aStart), mDirection(normalized(aDirection)) {}
};

// reference/bezierLens.h:7-11
enum class RefractionResult : uint32_t { cNone = BZ_STATUS_NONE, cInside = BZ_STATUS_INSIDE, cOutside = BZ_STATUS_OUTSIDE };

// reference/bezierTriangle.h:7-20 (+ the index of the hit patch, which the reference does not report)
struct BezierIntersection final {
  enum class What : uint32_t { cFollowSide0 = 0u, cFollowSide1 = 1u, cFollowSide2 = 2u, cNone = BZ_WHAT_NONE, cIntersect = BZ_WHAT_INTERSECT };
  bool     mValid;
  Vertex   mPoint;
  float    mCosIncidence;
  float    mDistance;
  Vertex   mBarycentric;
  Vector   mNormal;
  What     mWhat;
  uint32_t mPatch;
};

class BezierLens final {
public:
  // BezierLens(float aRi, BezierMesh const&) and BezierLens(float aRi, BezierMesh &&), reference/bezierLens.h:23-24.  The reference's
  // lens owns a copy of the mesh; here it owns the device tables made from it, and copies of a lens SHARE them (they are
  // read-only, and every call works in a per-stream / per-call context of the handle: include/bz_capi.h "Threads and streams").
  BezierLens(float const aRi, BezierMesh const &aMesh, int const aDevice = 0) : mRefractiveIndex(aRi), mPatchCount(static_cast<uint32_t>(aMesh.size())) {
    bz_lens *lens = nullptr;
    check(bz_lens_create(aMesh.records(), mPatchCount, aRi, aDevice, &lens));
    mShared = std::shared_ptr<bz_lens>(lens, [](bz_lens *aLens) { bz_lens_destroy(aLens); });
  }
  BezierLens(float const aRi, BezierMesh &&aMesh, int const aDevice = 0) : BezierLens(aRi, static_cast<BezierMesh const &>(aMesh), aDevice) {}
  BezierLens(BezierLens const &) = default;
  BezierLens(BezierLens &&) noexcept = default;
  BezierLens &operator=(BezierLens const &) = default;
  BezierLens &operator=(BezierLens &&) noexcept = default;
  ~BezierLens() = default;

  // std::pair<Ray, RefractionResult> refract(Ray const&, RefractionResult aExpected) const, reference/bezierLens.h:27.
  // On cNone the returned ray is the incoming one (the reference leaves it partly uninitialised).
  std::pair<Ray, RefractionResult> refract(Ray const &aRay, RefractionResult const aExpected) const {
    float const start[3] = { aRay.mStart.x, aRay.mStart.y, aRay.mStart.z }, dir[3] = { aRay.mDirection.x, aRay.mDirection.y, aRay.mDirection.z };
    uint32_t const expected = static_cast<uint32_t>(aExpected);
    float outStart[3], outDir[3];
    uint32_t status = 0u;
    check(bz_refract_batch_host(mShared.get(), start, dir, &expected, 1u, outStart, outDir, &status, nullptr));
    Ray result;
    result.mStart = Vec3{outStart[0], outStart[1], outStart[2]};
    result.mDirection = Vec3{outDir[0], outDir[1], outDir[2]};
    return std::make_pair(result, static_cast<RefractionResult>(status));
  }

  // the same over n host rays (3 floats each); results in caller-owned arrays, aOutPatch may be null
  void refract(float const *aStart, float const *aDirection, uint32_t const *aExpected, uint64_t const aCount,
               float *aOutStart, float *aOutDirection, uint32_t *aOutStatus, uint32_t *aOutPatch = nullptr) const {
    check(bz_refract_batch_host(mShared.get(), aStart, aDirection, aExpected, aCount, aOutStart, aOutDirection, aOutStatus, aOutPatch));
  }

  // BezierIntersection BezierMesh::intersect(Ray const&) const, reference/bezierMesh.h:37
  BezierIntersection intersect(Ray const &aRay) const {
    float const start[3] = { aRay.mStart.x, aRay.mStart.y, aRay.mStart.z }, dir[3] = { aRay.mDirection.x, aRay.mDirection.y, aRay.mDirection.z };
    uint32_t what = 0u, patch = 0u;
    float distance = 0.0f, cosine = 0.0f, point[3], normal[3], bary[3];
    check(bz_intersect_batch_host(mShared.get(), start, dir, 1u, &what, &patch, &distance, point, normal, &cosine, bary));
    BezierIntersection result;
    result.mWhat = static_cast<BezierIntersection::What>(what);
    result.mValid = (what == BZ_WHAT_INTERSECT);
    result.mPatch = patch;
    result.mDistance = distance;
    result.mCosIncidence = cosine;
    result.mPoint = Vec3{point[0], point[1], point[2]};
    result.mNormal = Vec3{normal[0], normal[1], normal[2]};
    result.mBarycentric = Vec3{bary[0], bary[1], bary[2]};
    return result;
  }

  // headless render of rows [aRow0, aRow1) into a host RGBA8 buffer (definitions in include/bz_capi.h)
  void render(bz_camera const &aCamera, bz_scene const &aScene, uint32_t const aRow0, uint32_t const aRow1, uint8_t *aRgba, bz_stats *aStats = nullptr) const {
    check(bz_render_host(mShared.get(), &aCamera, &aScene, aRow0, aRow1, aRgba, aStats));
  }

  // illumination mode (README.md:161-194): emitter -> lens -> counting bins of a target rectangle, host bins are accumulated into
  void illuminate(bz_emitter const &aEmitter, bz_target const &aTarget, uint32_t const aMaxCrossings, uint32_t *aBins, bz_stats *aStats = nullptr) const {
    check(bz_illuminate_host(mShared.get(), &aEmitter, &aTarget, aMaxCrossings, aBins, aStats));
  }

  float getRefractiveIndex() const { return mRefractiveIndex; }
  uint32_t size() const { return mPatchCount; }
  bz_lens *handle() const { return mShared.get(); }       // for the device-pointer entry points of the C ABI

private:
  float    mRefractiveIndex;
  uint32_t mPatchCount;
  std::shared_ptr<bz_lens> mShared;      // copies of a lens share the device tables

  static void check(int const aCode) {
    if(aCode != BZ_OK) { throw std::runtime_error("libbz_b200 error " + std::to_string(aCode) + ": " + bz_last_error()); }
  }
};

}

#endif

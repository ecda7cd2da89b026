// bz_bezier_mesh.h -- Clough-Tocher split of a standardized Mesh into cubic Bezier patches: host C++ API with
// the names of the reference's BezierTriangle / BezierMesh (reference/bezierTriangle.h:22-111,
// bezierMesh.h:10-45).  A patch IS its bz_patch record (include/bz_capi.h): the table this class builds is
// what bz_lens_create uploads, so there is no second representation to keep in step.
#ifndef BZ_BEZIER_MESH_H
#define BZ_BEZIER_MESH_H

#include <array>
#include <cstdint>
#include <vector>
#include "bz_capi.h"
#include "bz_geom.h"
#include "bz_mesh.h"

namespace bz {

class BezierTriangle final {
public:
  static constexpr uint32_t csControlPointsSize = 10u;

  BezierTriangle() = default;
  // bezierTriangle.cpp:4-43: control points 0, 1, 3, 4, 5, 8, 9 and divider plane 0
  BezierTriangle(Vertex const &aOriginalCommonVertex0, Vertex const &aOriginalCommonVertex1, Vertex const &aOriginalCentroid,
                 Vector const &aAverageNormal0, Vector const &aAverageNormal1, Plane const &aPlaneBetweenOriginalNeighbours,
                 std::array<uint32_t, 3u> const &aNeighbourIndices);
  void setMissingFields1(Vertex const &aOriginalCentroid, BezierTriangle const &aTriangleNext, BezierTriangle const &aTrianglePrevious);
  void setMissingFields2(Vertex const &, BezierTriangle const &aTriangleNext, BezierTriangle const &);
  void setMissingFields3(Vertex const &, BezierTriangle const &aTriangleNext, BezierTriangle const &aTrianglePrevious);

  Vertex getControlPoint(uint32_t const aI) const { return Vec3{mRecord.control[aI][0], mRecord.control[aI][1], mRecord.control[aI][2]}; }
  std::array<uint32_t, 3u> getNeighbours() const { return {mRecord.neighbours[0], mRecord.neighbours[1], mRecord.neighbours[2]}; }
  Plane getUnderlyingPlane() const { return Plane{Vec3{mRecord.plane[0], mRecord.plane[1], mRecord.plane[2]}, mRecord.plane[3]}; }
  float getHeightInside() const { return mRecord.height_inside; }
  float getHeightOutside() const { return mRecord.height_outside; }

  Vertex interpolateLinear(float const aBary0, float const aBary1, float const aBary2) const;
  Vertex interpolate(float const aBary0, float const aBary1, float const aBary2) const;
  Vertex interpolate(Vertex const &aBary) const { return interpolate(aBary.x, aBary.y, aBary.z); }
  Vertex interpolateAboveOriginalCentroid() const { return getControlPoint(2u); }

  bz_patch const &record() const { return mRecord; }

private:
  static constexpr float   csProportionControlOnOriginalSide           = 0.291f;
  static constexpr float   csProportionControlOnOriginalVertexCentroid = 0.304f;
  static constexpr float   csProportionControlOnOriginalMedian         = 0.2f;
  static constexpr float   csHeightSafetyFactor                        = 1.33333333f;
  static constexpr int32_t csHeightSampleDivisor                       = 5;

  bz_patch mRecord;

  void setControl(uint32_t const aI, Vec3 const &aPoint) { mRecord.control[aI][0] = aPoint.x; mRecord.control[aI][1] = aPoint.y; mRecord.control[aI][2] = aPoint.z; }
  void setDivider(uint32_t const aI, Plane const &aPlane) {
    mRecord.divider[aI][0] = aPlane.normal.x; mRecord.divider[aI][1] = aPlane.normal.y; mRecord.divider[aI][2] = aPlane.normal.z; mRecord.divider[aI][3] = aPlane.constant;
  }
};

static_assert(sizeof(BezierTriangle) == sizeof(bz_patch), "a BezierMesh is a contiguous bz_patch table");

class BezierMesh final {
public:
  // bezierMesh.cpp:4-34; throws std::out_of_range when aMesh has not been through standardizeNormals()
  BezierMesh(Mesh const &aMesh);

  auto size() const { return mMesh.size(); }
  auto cbegin() const { return mMesh.cbegin(); }
  auto cend() const { return mMesh.cend(); }
  auto const &operator[](uint32_t const aI) const { return mMesh[aI]; }

  Mesh interpolate(int32_t const aDivisor) const;             // bezierMesh.cpp:55-66
  std::vector<Vertex> dumpControlPoints() const;              // bezierMesh.cpp:68-77
  Mesh splitThickBezierTriangles() const;                     // bezierMesh.cpp:79-198

  // the table bz_lens_create takes
  bz_patch const *records() const { return reinterpret_cast<bz_patch const *>(mMesh.data()); }

private:
  static constexpr std::array<float, 3u> csSampleRatiosOriginalSide      = { 0.25f, 0.5f, 0.75f };
  static constexpr float                 csBezierHeightPerPerimeterLimit = 0.03f;
  static constexpr float                 csSplitBezierInterpolateFactor  = 0.7f;

  std::vector<BezierTriangle> mMesh;
  Mesh::Face2neighbours       mOriginalNeighbours;

  Vertex interpolate(uint32_t const aIndex, float const aBary0, float const aBary1, float const aBary2) const;
};

}

#endif

// bz_bezier_mesh.cpp -- builds the bz_patch table (control points, underlying plane, divider planes,
// barycentric inverse, heights, derivative direction, neighbours) from a standardized Mesh.
// Follows reference/bezierTriangle.cpp:4-121 and reference/bezierMesh.cpp:4-198; the arithmetic order is the
// reference's, so the table is bit-identical to the reference's BezierTriangle fields.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include "bz_bezier_mesh.h"

namespace bz {

namespace {

// control point roles (bezierTriangle.h:31-51): 0 = 300, 1 = 030, 2 = 003 (above the original centroid),
// 3 = 210, 4 = 120 (on the original side), 5 = 021, 6 = 012 (towards the centroid), 7 = 102, 8 = 201 (back from it), 9 = 111
enum : uint32_t { c300 = 0u, c030 = 1u, c003 = 2u, c210 = 3u, c120 = 4u, c021 = 5u, c012 = 6u, c102 = 7u, c201 = 8u, c111 = 9u };

}

BezierTriangle::BezierTriangle(Vertex const &aOriginalCommonVertex0, Vertex const &aOriginalCommonVertex1, Vertex const &aOriginalCentroid,
                               Vector const &aAverageNormal0, Vector const &aAverageNormal1, Plane const &aPlaneBetweenOriginalNeighbours,
                               std::array<uint32_t, 3u> const &aNeighbourIndices) {
  std::memset(&mRecord, 0, sizeof(mRecord));
  for(uint32_t i = 0u; i < 3u; ++i) { mRecord.neighbours[i] = aNeighbourIndices[i]; }
  Vec3 const &v0 = aOriginalCommonVertex0, &v1 = aOriginalCommonVertex1;
  setControl(c300, v0);
  setControl(c030, v1);

  // tangent planes at the two original vertices: the first plane of every triple below keeps the surface C1 across patches
  Plane const tangent0{aAverageNormal0, dot(v0, aAverageNormal0)};
  Plane const tangent1{aAverageNormal1, dot(v1, aAverageNormal1)};
  Vec3 const on210 = Plane::meet(tangent0, aPlaneBetweenOriginalNeighbours, Plane::from1proportion2points(csProportionControlOnOriginalSide, v0, v1));
  Vec3 const on120 = Plane::meet(tangent1, aPlaneBetweenOriginalNeighbours, Plane::from1proportion2points(csProportionControlOnOriginalSide, v1, v0));
  setControl(c210, on210);
  setControl(c120, on120);

  Vec3 const originalNormal = faceNormal(v0, v1, aOriginalCentroid);
  setControl(c201, Plane::meet(tangent0, Plane::from1vector2points(originalNormal, v0, aOriginalCentroid),
                               Plane::from1proportion2points(csProportionControlOnOriginalVertexCentroid, v0, aOriginalCentroid)));
  setControl(c021, Plane::meet(tangent1, Plane::from1vector2points(originalNormal, v1, aOriginalCentroid),
                               Plane::from1proportion2points(csProportionControlOnOriginalVertexCentroid, v1, aOriginalCentroid)));

  Vec3 const middle = Plane::meet(Plane::from1vector2points(aPlaneBetweenOriginalNeighbours.normal, on210, on120),
                                  Plane::from1proportion2points(0.5f, on210, on120),
                                  Plane::from1proportion2points(csProportionControlOnOriginalMedian, (v0 + v1) / 2.0f, aOriginalCentroid));
  setControl(c111, middle);

  Plane divider = aPlaneBetweenOriginalNeighbours;
  divider.makeDistancePositive(middle);
  setDivider(0u, divider);
}

void BezierTriangle::setMissingFields1(Vertex const &aOriginalCentroid, BezierTriangle const &aTriangleNext, BezierTriangle const &aTrianglePrevious) {
  Vec3 const v0 = getControlPoint(c300), v1 = getControlPoint(c030), middle = getControlPoint(c111);
  Vec3 const originalNormal = faceNormal(v0, v1, aOriginalCentroid);
  Plane const throughMiddles0 = Plane::from3points(getControlPoint(c201), middle, aTrianglePrevious.getControlPoint(c111));
  Plane const throughMiddles1 = Plane::from3points(getControlPoint(c021), aTriangleNext.getControlPoint(c111), middle);
  setControl(c102, Plane::meet(throughMiddles0, Plane::from1vector2points(originalNormal, v0, aOriginalCentroid),
                               Plane::from1proportion2points(csProportionControlOnOriginalVertexCentroid, aOriginalCentroid, v0)));
  setControl(c012, Plane::meet(throughMiddles1, Plane::from1vector2points(originalNormal, v1, aOriginalCentroid),
                               Plane::from1proportion2points(csProportionControlOnOriginalVertexCentroid, aOriginalCentroid, v1)));
}

void BezierTriangle::setMissingFields2(Vertex const &, BezierTriangle const &aTriangleNext, BezierTriangle const &) {
  Vec3 const v0 = getControlPoint(c300), v1 = getControlPoint(c030);
  Vec3 const top = (getControlPoint(c102) + getControlPoint(c012) + aTriangleNext.getControlPoint(c012)) / 3.0f;
  setControl(c003, top);

  Plane const plane = Plane::from3points(v0, v1, top);
  mRecord.plane[0] = plane.normal.x; mRecord.plane[1] = plane.normal.y; mRecord.plane[2] = plane.normal.z; mRecord.plane[3] = plane.constant;

  // b = inverse(columns v0 v1 top) * point  (util::getBarycentricInverse, 3dGeomUtil.h:64-72)
  Mat3 const inv = inverse(Mat3{{{v0.x, v1.x, top.x}, {v0.y, v1.y, top.y}, {v0.z, v1.z, top.z}}});
  for(int r = 0; r < 3; ++r) { for(int c = 0; c < 3; ++c) { mRecord.bary_inv[r][c] = inv.m[r][c]; } }

  float inside = 0.0f, outside = 0.0f;
  Triangle const corners{Vec3{1.0f, 0.0f, 0.0f}, Vec3{0.0f, 1.0f, 0.0f}, Vec3{0.0f, 0.0f, 1.0f}};
  Mesh::divide(corners, csHeightSampleDivisor, [&](Triangle const &aBary) {
    for(uint32_t i = 0u; i < 3u; ++i) {
      float const distance = plane.distance(interpolate(aBary[i]));
      inside = std::min(inside, distance);
      outside = std::max(outside, distance);
    }
  });
  mRecord.height_inside = inside * csHeightSafetyFactor;
  mRecord.height_outside = outside * csHeightSafetyFactor;

  // direction A is the constant (1, 0, -1) and is not stored; B is perpendicular to it in the patch plane
  Vec3 const dirB = inv * cross(top - v0, plane.normal);
  mRecord.dir_b[0] = dirB.x; mRecord.dir_b[1] = dirB.y; mRecord.dir_b[2] = dirB.z;
}

void BezierTriangle::setMissingFields3(Vertex const &, BezierTriangle const &aTriangleNext, BezierTriangle const &aTrianglePrevious) {
  Vec3 const normal = getUnderlyingPlane().normal, top = getControlPoint(c003), middle = getControlPoint(c111);
  Plane towardsNext = Plane::from1vector2points(normal + aTriangleNext.getUnderlyingPlane().normal, getControlPoint(c030), top);
  Plane towardsPrevious = Plane::from1vector2points(normal + aTrianglePrevious.getUnderlyingPlane().normal, getControlPoint(c300), top);
  towardsNext.makeDistancePositive(middle);
  towardsPrevious.makeDistancePositive(middle);
  setDivider(1u, towardsNext);
  setDivider(2u, towardsPrevious);
}

Vertex BezierTriangle::interpolateLinear(float const aBary0, float const aBary1, float const aBary2) const {
  return getControlPoint(c300) * aBary0 + getControlPoint(c030) * aBary1 + getControlPoint(c003) * aBary2;
}

// cubic Bernstein form, summation order of bezierTriangle.cpp:105-121
Vertex BezierTriangle::interpolate(float const aBary0, float const aBary1, float const aBary2) const {
  float const square0 = aBary0 * aBary0, square1 = aBary1 * aBary1, square2 = aBary2 * aBary2;
  auto cp = [this](uint32_t aI) { return getControlPoint(aI); };
  Vec3 const corners = cp(c300) * aBary0 * square0 + cp(c030) * aBary1 * square1 + cp(c003) * aBary2 * square2;
  Vec3 const sides = cp(c210) * aBary1 * square0 + cp(c120) * aBary0 * square1 + cp(c021) * aBary2 * square1
                   + cp(c012) * aBary1 * square2 + cp(c102) * aBary0 * square2 + cp(c201) * aBary2 * square0;
  return corners + 3.0f * sides + cp(c111) * aBary0 * aBary1 * aBary2 * 6.0f;
}

BezierMesh::BezierMesh(Mesh const &aMesh) : mOriginalNeighbours(aMesh.getFace2neighbours()) {
  if(mOriginalNeighbours.size() != aMesh.size()) {
    throw std::out_of_range("BezierMesh: the mesh has no neighbour table (standardizeNormals not run?)");
  }
  uint32_t const faces = static_cast<uint32_t>(aMesh.size());
  mMesh.reserve(static_cast<size_t>(faces) * 3u);
  std::vector<Vec3> centroids(faces);
  for(uint32_t f = 0u; f < faces; ++f) {
    auto const &neigh = mOriginalNeighbours[f];
    Triangle const &original = aMesh[f];
    centroids[f] = (original[0u] + original[1u] + original[2u]) / 3.0f;
    Vec3 const normal = normalized(faceNormal(original));
    for(uint32_t corner = 0u; corner < 3u; ++corner) {       // Clough-Tocher: one patch per original side
      Vec3 const &v0 = original[corner];
      Vec3 const &v1 = original[(corner + 1u) % 3u];
      Plane const between = Plane::from1vector2points(normal + normalized(faceNormal(aMesh[neigh.mFellowTriangles[corner]])), v0, v1);
      std::array<uint32_t, 3u> const neighbours{3u * neigh.mFellowTriangles[corner] + neigh.mFellowCommonSideStarts[corner],
                                                f * 3u + (corner + 1u) % 3u, f * 3u + (corner + 2u) % 3u};
      mMesh.emplace_back(v0, v1, centroids[f], aMesh.getAverageNormal(v0), aMesh.getAverageNormal(v1), between, neighbours);
    }
  }
  using Pass = void (BezierTriangle::*)(Vertex const &, BezierTriangle const &, BezierTriangle const &);
  for(Pass const pass : {&BezierTriangle::setMissingFields1, &BezierTriangle::setMissingFields2, &BezierTriangle::setMissingFields3}) {
    for(uint32_t i = 0u; i < mMesh.size(); ++i) {
      uint32_t const sub = i % 3u, base = i - sub;
      (mMesh[i].*pass)(centroids[base / 3u], mMesh[base + (sub + 1u) % 3u], mMesh[base + (sub + 2u) % 3u]);
    }
  }
}

Mesh BezierMesh::interpolate(int32_t const aDivisor) const {
  Mesh result;
  Triangle const corners{Vec3{1.0f, 0.0f, 0.0f}, Vec3{0.0f, 1.0f, 0.0f}, Vec3{0.0f, 0.0f, 1.0f}};
  Mesh::divide(corners, aDivisor, [&result, this](Triangle const &aBary) {
    for(auto const &bezier : mMesh) {
      result.push_back(Triangle{bezier.interpolate(aBary[0u]), bezier.interpolate(aBary[1u]), bezier.interpolate(aBary[2u])});
    }
  });
  return result;
}

std::vector<Vertex> BezierMesh::dumpControlPoints() const {
  std::vector<Vertex> result;
  result.reserve(mMesh.size() * BezierTriangle::csControlPointsSize);
  for(auto const &bezier : mMesh) {
    for(uint32_t i = 0u; i < BezierTriangle::csControlPointsSize; ++i) { result.push_back(bezier.getControlPoint(i)); }
  }
  return result;
}

Vertex BezierMesh::interpolate(uint32_t const aIndex, float const aBary0, float const aBary1, float const aBary2) const {
  auto const &triangle = mMesh[aIndex];
  return triangle.interpolate(aBary0, aBary1, aBary2) * csSplitBezierInterpolateFactor
       + triangle.interpolateLinear(aBary0, aBary1, aBary2) * (1.0f - csSplitBezierInterpolateFactor);
}

// bezierMesh.cpp:79-198.  An original triangle whose Bezier surface rises more than 3 % of its perimeter
// above its plane gets all three sides halved, and so does the matching side of each neighbour; a triangle
// then falls apart into 1, 2, 3 or 4 pieces depending on how many of its sides were halved.  The midpoints sit
// 70 % of the way from the flat side midpoint to the Bezier surface.
Mesh BezierMesh::splitThickBezierTriangles() const {
  uint32_t const faces = static_cast<uint32_t>(mOriginalNeighbours.size());
  std::vector<uint8_t> halved(faces, 0u);      // bit i: side (i, i + 1) is halved
  auto originalOf = [this](uint32_t aFace) {
    return Triangle{mMesh[aFace * 3u].getControlPoint(0u), mMesh[aFace * 3u + 1u].getControlPoint(0u), mMesh[aFace * 3u + 2u].getControlPoint(0u)};
  };
  for(uint32_t f = 0u; f < faces; ++f) {
    Triangle const original = originalOf(f);
    Plane const plane = Plane::from3points(original[0u], original[1u], original[2u]);
    float rise = std::abs(plane.distance(mMesh[f * 3u].interpolateAboveOriginalCentroid()));
    for(uint32_t i = 0u; i < 3u; ++i) {
      for(float const ratio : csSampleRatiosOriginalSide) {
        rise = std::max(rise, std::abs(plane.distance(mMesh[f * 3u + i].interpolate(ratio, 1.0f - ratio, 0.0f))));
      }
    }
    float const perimeter = norm(original[0u] - original[1u]) + norm(original[1u] - original[2u]) + norm(original[2u] - original[0u]);
    if(rise / perimeter > csBezierHeightPerPerimeterLimit) {
      halved[f] = 7u;
      auto const &neigh = mOriginalNeighbours[f];
      for(uint32_t side = 0u; side < 3u; ++side) {
        halved[neigh.mFellowTriangles[side]] |= static_cast<uint8_t>(1u << neigh.mFellowCommonSideStarts[side]);
      }
    }
  }
  Mesh result;
  for(uint32_t f = 0u; f < faces; ++f) {
    Triangle const t = originalOf(f);
    auto mid = [this, f](uint32_t aSide) { return interpolate(f * 3u + aSide, 0.5f, 0.5f, 0.0f); };
    uint8_t const mask = halved[f];
    int const count = (mask & 1u) + ((mask >> 1u) & 1u) + ((mask >> 2u) & 1u);
    if(count == 0) {
      result.push_back(t);
    }
    else if(count == 1) {                        // fan from the corner opposite the halved side
      uint32_t const side = (mask == 1u ? 0u : (mask == 2u ? 1u : 2u));
      uint32_t const after = (side + 1u) % 3u, before = (side + 2u) % 3u;
      Vec3 const m = mid(side);
      result.push_back(Triangle{t[after], t[before], m});
      result.push_back(Triangle{t[before], t[side], m});
    }
    else if(count == 2) {                        // corner triangle + quadrilateral cut along its shorter diagonal
      uint32_t const whole = (mask == 3u ? 2u : (mask == 5u ? 1u : 0u));        // the side that stays whole starts here
      uint32_t const after = (whole + 1u) % 3u, before = (whole + 2u) % 3u;
      Vec3 const mBefore = mid(before), mAfter = mid(after);
      result.push_back(Triangle{t[before], mBefore, mAfter});
      if(norm(t[after] - mBefore) < norm(t[whole] - mAfter)) {
        result.push_back(Triangle{t[after], mAfter, mBefore});
        result.push_back(Triangle{t[whole], t[after], mBefore});
      }
      else {
        result.push_back(Triangle{t[after], mAfter, t[whole]});
        result.push_back(Triangle{t[whole], mAfter, mBefore});
      }
    }
    else {                                       // middle triangle + three corner triangles
      Triangle const middle{mid(0u), mid(1u), mid(2u)};
      result.push_back(middle);
      for(uint32_t i = 0u; i < 3u; ++i) { result.push_back(Triangle{t[i], middle[i], middle[(i + 2u) % 3u]}); }
    }
  }
  return result;
}

}

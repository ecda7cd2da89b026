// bz_mesh.h -- triangle-mesh preprocessing that feeds the Bezier tables: host C++ API with the names,
// argument meaning and error behaviour of the reference's Mesh (reference/mesh.h:18-133).
//
// The implementation is flat-array based (sorted index runs instead of multimaps, an open vertex index
// instead of node-based hash containers) but every floating-point result -- snapped vertices, winding,
// neighbour tables, angle-weighted vertex normals -- is bit-identical to the reference's, including the
// places where the reference's result depends on the iteration order of its hash containers
// (bz_mesh.cpp says where).
#ifndef BZ_MESH_H
#define BZ_MESH_H

#include <array>
#include <cstdint>
#include <functional>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>
#include "bz_geom.h"

namespace bz {

class Mesh final {
public:
  using TheMesh = std::vector<Triangle>;
  using value_type = Triangle;

  // reference/mesh.h:28-32
  struct Neighbours final {
    std::array<uint32_t, 3u> mFellowTriangles;         // neighbour across edge (i, i + 1)
    std::array<uint8_t, 3u>  mFellowCommonSideStarts;  // corner of the neighbour where the common side starts
  };
  using Face2neighbours = std::vector<Neighbours>;
  // exact-value vertex identity (reference/mesh.h:34-38 hashes the three floats; -0 and +0 are one vertex)
  struct VertexHash {
    std::size_t operator()(Vec3 const &aV) const {
      auto bits = [](float aF) { uint32_t u; float const f = (aF == 0.0f ? 0.0f : aF); std::memcpy(&u, &f, 4u); return static_cast<std::size_t>(u); };
      return (bits(aV.x) * 0x9e3779b97f4a7c15ull) ^ (bits(aV.y) << 21u) ^ (bits(aV.z) << 42u) ^ bits(aV.z);
    }
  };
  using CornerNormals = std::vector<std::array<Vector, 3u>>;   // average normal of every corner's vertex

  Mesh() = default;

  auto size() const { return mMesh.size(); }
  auto begin() const { return mMesh.begin(); }
  auto end() const { return mMesh.end(); }
  auto cbegin() const { return mMesh.cbegin(); }
  auto cend() const { return mMesh.cend(); }
  void reserve(uint32_t const aSize) { mMesh.reserve(aSize); }
  void push_back(Triangle const &aTriangle) { mMesh.push_back(aTriangle); }
  void clear() { mMesh.clear(); }
  auto &operator[](uint32_t const aI) { return mMesh[aI]; }
  auto const &operator[](uint32_t const aI) const { return mMesh[aI]; }

  TheMesh const &getMesh() const { return mMesh; }
  Face2neighbours const &getFace2neighbours() const { return mFace2neighbours; }
  // The reference keeps an unordered_map<Vertex, Vector> (mesh.h:69); the same values are held here
  // per vertex index, with a lookup by vertex value.  Throws std::out_of_range like .at() does
  // (bezierMesh.cpp:17-18) when standardizeNormals() has not run or the vertex is unknown.
  Vector const &getAverageNormal(Vertex const &aVertex) const;
  // Vertex2averageNormals const &getVertex2averageNormals() const, reference/mesh.h:69: a read-only map view with the calls the
  // reference's users make (bezierMesh.cpp:17-18: .at(vertex)); iteration goes over (vertex, normal) in first-seen order
  class Vertex2averageNormals final {
  public:
    explicit Vertex2averageNormals(Mesh const &aMesh) : mMesh(aMesh) {}
    Vector const &at(Vertex const &aVertex) const { return mMesh.getAverageNormal(aVertex); }      // throws std::out_of_range like unordered_map::at
    size_t count(Vertex const &aVertex) const { return mMesh.mNormalIndex.count(aVertex); }
    size_t size() const { return mMesh.mVertexNormals.size(); }
    bool empty() const { return mMesh.mVertexNormals.empty(); }
    std::pair<Vertex const &, Vector const &> operator[](size_t const aI) const { return { mMesh.mUniqueVertices[aI], mMesh.mVertexNormals[aI] }; }
  private:
    Mesh const &mMesh;
  };
  Vertex2averageNormals getVertex2averageNormals() const { return Vertex2averageNormals(*this); }
  std::vector<Vertex> const &getUniqueVertices() const { return mUniqueVertices; }   // first-seen order
  std::vector<Vector> const &getVertexNormals() const { return mVertexNormals; }     // same order

  void standardizeVertices();                                    // reference/mesh.cpp:72-91
  std::vector<Vertex> getVertices() const;                       // distinct vertices, first-seen order
  void standardizeNormals();                                     // reference/mesh.cpp:310-357; throws char const* "Vertex on edge detected."

  void transform(Mat3 const &aTransform, Vertex const aDisplacement);   // mesh.cpp:361-367
  Mesh &operator+=(Vector const aDisplacement) { transform(Mat3::identity(), aDisplacement); return *this; }
  Mesh &operator*=(Mat3 const &aTransform) { transform(aTransform, Vec3{0.0f, 0.0f, 0.0f}); return *this; }
  Mesh &operator*=(float const &aFactor) { transform(Mat3::identity() * aFactor, Vec3{0.0f, 0.0f, 0.0f}); return *this; }

  void splitTriangles(float const aMaxTriangleSide);             // mesh.cpp:375-385
  void splitTriangles(int32_t const aDivisor);                   // mesh.cpp:389-395

  void readMesh(std::string const &aFilename);                   // STL, ASCII or binary; throws std::runtime_error
  void writeMesh(std::string const &aFilename) const;            // ASCII STL, mesh.cpp:419-430

  void makeSolidOfRevolution(int32_t const aSectors, int32_t const aBelts, std::function<float(float)> aEnvelope, Vector const &aSize);
  void makeEllipsoid(int32_t const aSectors, int32_t const aBelts, Vector const &aSize);
  void makeUnitSphere(int32_t const aSectors, int32_t const aBelts) { makeEllipsoid(aSectors, aBelts, Vec3{1.0f, 1.0f, 1.0f}); }

  // triangle subdivision pattern shared with BezierMesh (util::divide, 3dGeomUtil.h:96-121)
  static void divide(Triangle const &aTriangle, int32_t const aDivisor, std::function<void(Triangle const &)> const &aCollector);

private:
  static constexpr float csStandardizeVerticesEpsilonFactor        = 0.2f;
  static constexpr float csStandardizeNormalsEpsilon               = 0.01f;
  static constexpr float csStandardizeNormalsIndependentMoveFactor = 0.2f;

  TheMesh             mMesh;
  Face2neighbours     mFace2neighbours;
  std::vector<Vertex> mUniqueVertices;
  std::vector<Vector> mVertexNormals;
  std::unordered_map<Vec3, uint32_t, VertexHash> mNormalIndex;   // vertex -> slot of mUniqueVertices / mVertexNormals;
                                                                  // never cleared, first value wins (the reference
                                                                  // emplaces into a map it never clears, mesh.cpp:305)
  std::vector<std::array<uint32_t, 3u>> mFace2vertex;   // as of the last indexFaces()

  float getSmallestSide() const;
  void indexFaces();                                             // vertex indices in first-seen order
  void createFace2neighbour();
  static uint32_t getIndependentFrom(Triangle const &aFaceTarget, Triangle const &aFaceOther);
  static void orientLike(Triangle const &aFaceKnown, Triangle &aFaceUnknown);
  void calculateNormalAverages4vertices();
};

// iteration order of a libstdc++ std::unordered_set<uint32_t> filled by inserting aInserted in sequence
// (decides the seed face of standardizeNormals when several faces tie, see bz_mesh.cpp)
std::vector<uint32_t> hashSetIterationOrder(std::vector<uint32_t> const &aInserted);

}

#endif

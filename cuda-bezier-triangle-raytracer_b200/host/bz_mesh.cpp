// bz_mesh.cpp -- host preprocessing of the triangle mesh (see bz_mesh.h).  Follows the behaviour of
// reference/mesh.cpp; citations name the reference lines a block reproduces.
//
// Three results of the reference depend on the iteration order of libstdc++ hash containers; they are
// reproduced here explicitly instead of by using the same containers:
//   (1) faces sharing an edge are met newest-first (unordered_multimap::equal_range, mesh.cpp:198-208);
//   (2) the seed face of the winding flood is the first maximum in unordered_set<uint32_t> iteration
//       order (mesh.cpp:224-239) -- bz::hashSetIterationOrder() replays that order;
//   (3) the angle-weighted normal of a vertex sums its faces newest-first (mesh.cpp:292-303), which fixes
//       the FP32 rounding of the sum.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <unordered_map>
#include "bz_mesh.h"

namespace bz {

namespace {

using VertexKeyHash = Mesh::VertexHash;

using VertexIndex = std::unordered_map<Vec3, uint32_t, VertexKeyHash>;

bool lexicographicLess(Vec3 const &aA, Vec3 const &aB) {
  return aA.x < aB.x || (aA.x == aB.x && aA.y < aB.y) || (aA.x == aB.x && aA.y == aB.y && aA.z < aB.z);
}

}

// Iteration order of a libstdc++ std::unordered_set<uint32_t> after inserting aInserted in sequence
// (GCC's _Hashtable with identity hash, prime bucket counts, max load factor 1): a singly linked node
// list where a node entering an empty bucket goes to the list head and a node entering an occupied
// bucket goes to the front of that bucket's run; a rehash re-threads the list in its current order.
std::vector<uint32_t> hashSetIterationOrder(std::vector<uint32_t> const &aInserted) {
  static const uint32_t csGrowth[] = { 13u, 29u, 59u, 127u, 257u, 541u, 1109u, 2357u, 5087u, 10273u, 20753u };
  constexpr int cNil = -1, cBefore = -2, cNone = -3;
  std::vector<uint32_t> value;
  std::vector<int> next;
  int head = cNil;
  std::vector<int> bucket(1u, cNone);
  size_t nextResize = 0u;
  size_t growth = 0u;
  auto getNext = [&](int aNode) { return aNode == cBefore ? head : next[aNode]; };
  auto setNext = [&](int aNode, int aTo) { if(aNode == cBefore) { head = aTo; } else { next[aNode] = aTo; } };
  auto rethread = [&](size_t aBuckets) {
    std::vector<int> fresh(aBuckets, cNone);
    int p = head;
    head = cNil;
    size_t beginBucket = 0u;
    while(p != cNil) {
      int const following = next[p];
      size_t const b = value[p] % aBuckets;
      if(fresh[b] == cNone) {
        next[p] = head;
        head = p;
        fresh[b] = cBefore;
        if(next[p] != cNil) { fresh[beginBucket] = p; }
        beginBucket = b;
      }
      else {
        next[p] = getNext(fresh[b]);
        setNext(fresh[b], p);
      }
      p = following;
    }
    bucket.swap(fresh);
  };
  for(uint32_t const v : aInserted) {
    if(std::find(value.begin(), value.end(), v) != value.end()) { continue; }
    if(value.size() + 1u > nextResize) {
      if(growth >= sizeof(csGrowth) / sizeof(csGrowth[0])) { return aInserted; }   // far beyond any vertex valence
      rethread(csGrowth[growth]);
      nextResize = csGrowth[growth];
      ++growth;
    }
    int const node = static_cast<int>(value.size());
    value.push_back(v);
    next.push_back(cNil);
    size_t const b = v % bucket.size();
    if(bucket[b] != cNone) {
      next[node] = getNext(bucket[b]);
      setNext(bucket[b], node);
    }
    else {
      next[node] = head;
      head = node;
      if(next[node] != cNil) { bucket[value[next[node]] % bucket.size()] = node; }
      bucket[b] = cBefore;
    }
  }
  std::vector<uint32_t> result;
  for(int p = head; p != cNil; p = next[p]) { result.push_back(value[p]); }
  return result;
}

float Mesh::getSmallestSide() const {
  float smallest = std::numeric_limits<float>::max();
  for(auto const &triangle : mMesh) {
    for(uint32_t i = 0u; i < 3u; ++i) { smallest = std::min(smallest, norm(triangle[i] - triangle[(i + 1u) % 3u])); }
  }
  return smallest;
}

// mesh.cpp:14-91.  Per axis: corners sorted by that coordinate (stable: equal coordinates keep face /
// corner order, as multimap insertion does), cut into runs whose span from the run start stays below
// epsilon; the axis with the shortest longest run is used, and inside every run each corner is snapped
// to the lexicographically largest corner closer than epsilon.
void Mesh::standardizeVertices() {
  float const epsilon = getSmallestSide() * csStandardizeVerticesEpsilonFactor;
  uint32_t const corners = static_cast<uint32_t>(mMesh.size()) * 3u;
  std::array<std::vector<uint32_t>, 3u> sorted;        // corner ids (3 * face + corner)
  std::array<std::vector<uint32_t>, 3u> runEnds;       // exclusive end position of every run
  std::array<uint32_t, 3u> longest;
  for(int axis = 0; axis < 3; ++axis) {
    auto &order = sorted[axis];
    order.resize(corners);
    for(uint32_t i = 0u; i < corners; ++i) { order[i] = i; }
    std::stable_sort(order.begin(), order.end(), [this, axis](uint32_t aA, uint32_t aB) {
      return mMesh[aA / 3u][aA % 3u][axis] < mMesh[aB / 3u][aB % 3u][axis];
    });
    uint32_t maximum = 0u, counter = 0u;
    float startValue = 0.0f;
    for(uint32_t i = 0u; i < corners; ++i) {
      float const now = mMesh[order[i] / 3u][order[i] % 3u][axis];
      if(i == 0u) {
        startValue = now;
        counter = 1u;
      }
      else if(now - startValue >= epsilon) {
        runEnds[axis].push_back(i);
        startValue = now;
        maximum = std::max(maximum, counter);
        counter = 1u;
      }
      else {
        ++counter;
      }
    }
    runEnds[axis].push_back(corners);
    maximum = std::max(maximum, counter);
    longest[axis] = static_cast<uint32_t>(static_cast<float>(maximum));   // the reference returns the count as float (mesh.cpp:24)
  }
  int const best = static_cast<int>(std::min_element(longest.cbegin(), longest.cend()) - longest.cbegin());
  float const epsilonSquared = epsilon * epsilon;
  uint32_t begin = 0u;
  for(uint32_t const end : runEnds[best]) {
    for(uint32_t i = begin; i < end; ++i) {
      Vec3 &v1 = mMesh[sorted[best][i] / 3u][sorted[best][i] % 3u];
      for(uint32_t j = begin; j < end; ++j) {
        Vec3 const &v2 = mMesh[sorted[best][j] / 3u][sorted[best][j] % 3u];
        if(squaredNorm(v1 - v2) < epsilonSquared && lexicographicLess(v1, v2)) { v1 = v2; }
      }
    }
    begin = end;
  }
}

std::vector<Vertex> Mesh::getVertices() const {
  std::vector<Vertex> result;
  VertexIndex seen;
  for(auto const &triangle : mMesh) {
    for(auto const &vertex : triangle) {
      if(seen.emplace(vertex, static_cast<uint32_t>(result.size())).second) { result.push_back(vertex); }
    }
  }
  return result;
}

Vector const &Mesh::getAverageNormal(Vertex const &aVertex) const {
  auto const found = mNormalIndex.find(aVertex);
  if(found == mNormalIndex.end()) {
    throw std::out_of_range("Mesh::getAverageNormal: vertex has no average normal (standardizeNormals not run?)");
  }
  return mVertexNormals[found->second];
}

// mesh.cpp:118-153: vertex indices in first-seen order (faces, then corners)
void Mesh::indexFaces() {
  mFace2vertex.clear();
  mFace2vertex.reserve(mMesh.size());
  VertexIndex seen;
  seen.reserve(mMesh.size());
  for(auto const &face : mMesh) {
    std::array<uint32_t, 3u> indices;
    for(uint32_t corner = 0u; corner < 3u; ++corner) {
      auto const inserted = seen.emplace(face[corner], static_cast<uint32_t>(seen.size()));
      indices[corner] = inserted.first->second;
    }
    mFace2vertex.push_back(indices);
  }
}

// mesh.cpp:185-222.  Every undirected edge lists the faces using it; the reference looks them up newest-first
// and takes the first that is not the asking face.
void Mesh::createFace2neighbour() {
  uint32_t const faces = static_cast<uint32_t>(mMesh.size());
  struct EdgeUse { uint64_t edge; uint32_t sequence; uint32_t face; };
  std::vector<EdgeUse> uses;
  uses.reserve(static_cast<size_t>(faces) * 3u);
  auto edgeKey = [](uint32_t aA, uint32_t aB) { return (static_cast<uint64_t>(std::min(aA, aB)) << 32u) | std::max(aA, aB); };
  for(uint32_t f = 0u; f < faces; ++f) {
    for(uint32_t i = 0u; i < 3u; ++i) {
      uses.push_back({edgeKey(mFace2vertex[f][i], mFace2vertex[f][(i + 1u) % 3u]), static_cast<uint32_t>(uses.size()), f});
    }
  }
  std::sort(uses.begin(), uses.end(), [](EdgeUse const &aA, EdgeUse const &aB) {
    return aA.edge < aB.edge || (aA.edge == aB.edge && aA.sequence > aB.sequence);     // newest first inside an edge
  });
  mFace2neighbours.clear();
  mFace2neighbours.reserve(faces);
  static const uint8_t csResolve[4u][4u] = {{3u, 0u, 2u, 3u}, {0u, 3u, 1u, 3u}, {2u, 1u, 3u, 3u}, {3u, 3u, 3u, 3u}};
  for(uint32_t f = 0u; f < faces; ++f) {
    auto const &face = mFace2vertex[f];
    Neighbours neighbours;
    for(uint32_t i = 0u; i < 3u; ++i) {
      uint32_t const v0 = face[i], v1 = face[(i + 1u) % 3u];
      uint64_t const key = edgeKey(v0, v1);
      auto at = std::lower_bound(uses.begin(), uses.end(), key, [](EdgeUse const &aUse, uint64_t aKey) { return aUse.edge < aKey; });
      if(at->face == f) {
        ++at;
        if(at == uses.end() || at->edge != key) { throw "Vertex on edge detected."; }
      }
      uint32_t const other = at->face;
      neighbours.mFellowTriangles[i] = other;
      auto const &otherFace = mFace2vertex[other];
      uint32_t const raw0 = static_cast<uint32_t>(std::find(otherFace.cbegin(), otherFace.cend(), v0) - otherFace.cbegin());
      uint32_t const raw1 = static_cast<uint32_t>(std::find(otherFace.cbegin(), otherFace.cend(), v1) - otherFace.cbegin());
      neighbours.mFellowCommonSideStarts[i] = csResolve[raw0][raw1];
    }
    mFace2neighbours.push_back(neighbours);
  }
}

uint32_t Mesh::getIndependentFrom(Triangle const &aFaceTarget, Triangle const &aFaceOther) {
  for(uint32_t i = 0u; i < 3u; ++i) {
    if(std::find(aFaceOther.cbegin(), aFaceOther.cend(), aFaceTarget[i]) == aFaceOther.cend()) { return i; }
  }
  return 3u;
}

// mesh.cpp:250-282: give aFaceUnknown the winding of its neighbour aFaceKnown.  The two altitudes onto the
// common edge point to opposite sides exactly when the faces are unfolded; with the face normals that tells
// whether the unknown face has to be flipped.  Nearly perpendicular faces are judged after tilting the
// unknown face's free corner towards the known one.
void Mesh::orientLike(Triangle const &aFaceKnown, Triangle &aFaceUnknown) {
  uint32_t const freeKnown = getIndependentFrom(aFaceKnown, aFaceUnknown);
  uint32_t const freeUnknown = getIndependentFrom(aFaceUnknown, aFaceKnown);
  if(freeKnown == 3u || freeUnknown == 3u) { return; }    // coincident faces: the reference reads out of bounds here
  uint32_t const k1 = (freeKnown + 1u) % 3u, k2 = (freeKnown + 2u) % 3u;
  uint32_t const u1 = (freeUnknown + 1u) % 3u, u2 = (freeUnknown + 2u) % 3u;
  Vec3 const altitudeKnown = altitude(aFaceKnown[k1], aFaceKnown[k2], aFaceKnown[freeKnown]);
  Vec3 altitudeUnknown = altitude(aFaceUnknown[u1], aFaceUnknown[u2], aFaceUnknown[freeUnknown]);
  float altitudes = dot(altitudeKnown, altitudeUnknown);
  Vec3 const normalKnown = faceNormal(aFaceKnown);
  Vec3 normalUnknown = faceNormal(aFaceUnknown);
  float normals = dot(normalKnown, normalUnknown);
  if(std::abs(normals / (norm(normalKnown) * norm(normalUnknown))) < csStandardizeNormalsEpsilon) {
    Vec3 const moved = aFaceUnknown[freeUnknown]
                     + csStandardizeNormalsIndependentMoveFactor * (aFaceKnown[freeKnown] - (aFaceKnown[k1] + aFaceKnown[k2]) / 2.0f);
    Triangle tilted = aFaceUnknown;
    tilted[freeUnknown] = moved;
    altitudeUnknown = altitude(aFaceUnknown[u1], aFaceUnknown[u2], moved);
    altitudes = dot(altitudeKnown, altitudeUnknown);
    normalUnknown = faceNormal(tilted);
    normals = dot(normalKnown, normalUnknown);
  }
  if(altitudes * normals > 0.0f) { std::swap(aFaceUnknown[u1], aFaceUnknown[u2]); }
}

// mesh.cpp:284-308: angle-weighted average of the face normals around every vertex, faces taken newest-first
void Mesh::calculateNormalAverages4vertices() {
  uint32_t const faces = static_cast<uint32_t>(mMesh.size());
  // corner uses grouped by vertex: mFace2vertex is current (createFace2neighbour ran on the final winding)
  uint32_t vertices = 0u;
  for(auto const &face : mFace2vertex) { for(uint32_t const v : face) { vertices = std::max(vertices, v + 1u); } }
  std::vector<uint32_t> first(vertices + 1u, 0u);
  for(auto const &face : mFace2vertex) { for(uint32_t const v : face) { ++first[v + 1u]; } }
  for(uint32_t v = 0u; v < vertices; ++v) { first[v + 1u] += first[v]; }
  std::vector<uint32_t> uses(static_cast<size_t>(faces) * 3u);
  std::vector<uint32_t> fill(first.begin(), first.end() - 1);
  for(uint32_t f = 0u; f < faces; ++f) { for(uint32_t i = 0u; i < 3u; ++i) { uses[fill[mFace2vertex[f][i]]++] = f; } }
  std::vector<uint8_t> done(vertices, 0u);
  for(uint32_t f = 0u; f < faces; ++f) {
    for(uint32_t corner = 0u; corner < 3u; ++corner) {
      uint32_t const v = mFace2vertex[f][corner];
      if(done[v] != 0u) { continue; }
      done[v] = 1u;
      Vec3 const vertex = mMesh[f][corner];
      Vec3 sum{0.0f, 0.0f, 0.0f};
      for(uint32_t k = first[v + 1u]; k-- > first[v]; ) {
        Triangle const &triangle = mMesh[uses[k]];
        uint32_t const which = static_cast<uint32_t>(std::find(triangle.cbegin(), triangle.cend(), vertex) - triangle.cbegin());
        Vec3 const sideA = triangle[(which + 1u) % 3u] - triangle[which];
        Vec3 const sideB = triangle[(which + 2u) % 3u] - triangle[which];
        float const cosAngle = dot(sideA, sideB) / (norm(sideA) * norm(sideB));
        // the reference's unqualified acos() resolves to the double overload; the product is taken in float
        sum = sum + normalized(faceNormal(triangle)) * static_cast<float>(::acos(static_cast<double>(cosAngle)));
      }
      sum = normalized(sum);
      if(mNormalIndex.emplace(vertex, static_cast<uint32_t>(mUniqueVertices.size())).second) {
        mUniqueVertices.push_back(vertex);
        mVertexNormals.push_back(sum);
      }
    }
  }
}

// mesh.cpp:310-357
void Mesh::standardizeNormals() {
  indexFaces();
  // vertex with the smallest x (first strict minimum) and the faces around it (mesh.cpp:155-183)
  float smallestX = std::numeric_limits<float>::max();
  uint32_t smallestVertex = 0xffffffffu;
  for(uint32_t f = 0u; f < mMesh.size(); ++f) {
    for(uint32_t i = 0u; i < 3u; ++i) {
      if(mMesh[f][i].x < smallestX) { smallestX = mMesh[f][i].x; smallestVertex = mFace2vertex[f][i]; }
    }
  }
  std::vector<uint32_t> around;
  for(uint32_t f = 0u; f < mMesh.size(); ++f) {
    for(uint32_t i = 0u; i < 3u; ++i) { if(mFace2vertex[f][i] == smallestVertex) { around.push_back(f); } }
  }
  createFace2neighbour();
  if(mMesh.empty()) { return; }

  // seed: the face around that vertex facing -x most squarely (mesh.cpp:224-239), ties by hash-set order
  Vec3 const desired{-1.0f, 0.0f, 0.0f};
  float best = -std::numeric_limits<float>::max();
  uint32_t seed = 0u;
  for(uint32_t const f : hashSetIterationOrder(around)) {
    float const facing = std::abs(dot(desired, normalized(faceNormal(mMesh[f]))));
    if(facing > best) { best = facing; seed = f; }
  }
  if(dot(desired, faceNormal(mMesh[seed])) < 0.0f) { std::swap(mMesh[seed][0u], mMesh[seed][1u]); }    // mesh.cpp:241-248

  // depth-first flood over the neighbour graph (mesh.cpp:325-350)
  struct Pair { uint32_t known, unknown; };
  std::vector<Pair> stack;
  std::vector<uint8_t> remaining(mFace2neighbours.size(), 1u);
  for(uint32_t const f : mFace2neighbours[seed].mFellowTriangles) { stack.push_back({seed, f}); }
  remaining[seed] = 0u;
  while(!stack.empty()) {
    Pair const now = stack.back();
    stack.pop_back();
    if(remaining[now.unknown] != 0u) { orientLike(mMesh[now.known], mMesh[now.unknown]); }
    remaining[now.unknown] = 0u;
    for(uint32_t const f : mFace2neighbours[now.unknown].mFellowTriangles) {
      if(remaining[f] != 0u && now.unknown != f) { stack.push_back({now.unknown, f}); }
    }
  }
  indexFaces();                // corners were swapped
  createFace2neighbour();
  calculateNormalAverages4vertices();
}

void Mesh::transform(Mat3 const &aTransform, Vertex const aDisplacement) {
  for(auto &triangle : mMesh) {
    for(auto &vertex : triangle) { vertex = aTransform * vertex + aDisplacement; }
  }
}

// util::divide (3dGeomUtil.h:96-121): aDivisor^2 small triangles, emitted strip by strip starting at corner 0
// and walking towards corner 1; inside a strip upright and inverted triangles alternate along the 0 -> 2
// direction.  Points are accumulated by repeated addition of the two step vectors, as the reference does.
void Mesh::divide(Triangle const &aTriangle, int32_t const aDivisor, std::function<void(Triangle const &)> const &aCollector) {
  float const divisor = static_cast<float>(aDivisor);
  Vec3 const step01 = (aTriangle[1] - aTriangle[0]) / divisor;
  Vec3 const step02 = (aTriangle[2] - aTriangle[0]) / divisor;
  Vec3 stripStart = aTriangle[0];
  Vec3 apex = stripStart;
  Vec3 towards1 = aDivisor > 1 ? apex + step01 : aTriangle[1];
  Vec3 towards2 = aDivisor > 1 ? apex + step02 : aTriangle[2];
  for(int32_t strip = 0; strip + 1 < aDivisor; ++strip) {
    for(int32_t cell = 0; cell + strip + 1 < aDivisor; ++cell) {
      aCollector(Triangle{apex, towards1, towards2});
      Vec3 const advanced = towards1 + step02;
      aCollector(Triangle{towards1, advanced, towards2});
      towards1 = advanced;
      apex = towards2;
      towards2 = towards2 + step02;
    }
    aCollector(Triangle{apex, towards1, towards2});
    stripStart = stripStart + step01;
    apex = stripStart;
    towards1 = apex + step01;
    towards2 = apex + step02;
  }
  aCollector(Triangle{apex, aTriangle[1], towards2});
}

void Mesh::splitTriangles(float const aMaxTriangleSide) {
  TheMesh result;
  auto collect = [&result](Triangle const &aNew) { result.push_back(aNew); };
  for(auto const &triangle : mMesh) {
    float longest = norm(triangle[0] - triangle[1]);
    longest = std::max(longest, norm(triangle[0] - triangle[2]));
    longest = std::max(longest, norm(triangle[1] - triangle[2]));
    divide(triangle, static_cast<int32_t>(std::ceil(longest / aMaxTriangleSide)), collect);
  }
  mMesh = result;
}

void Mesh::splitTriangles(int32_t const aDivisor) {
  TheMesh result;
  auto collect = [&result](Triangle const &aNew) { result.push_back(aNew); };
  for(auto const &triangle : mMesh) { divide(triangle, aDivisor, collect); }
  mMesh = result;
}

// STL input.  Behaviour of the reader the reference uses (repos/stl_reader, stl_reader.h:514-721) as far as
// Mesh::readMesh (mesh.cpp:399-415) can observe it: format sniffing on the first 256 bytes, "vertex x y z"
// lines parsed with atof, triangles closed by "endfacet", binary records of 50 bytes, and triangles with two
// equal corners dropped.
void Mesh::readMesh(std::string const &aFilename) {
  mMesh.clear();
  mFace2neighbours.clear();
  std::ifstream probe(aFilename, std::ios::binary);
  if(!probe) { throw std::runtime_error("Couldn't open file " + aFilename); }
  char head[256];
  probe.read(head, sizeof(head));
  std::string start(head, static_cast<size_t>(probe.gcount()));
  std::transform(start.begin(), start.end(), start.begin(), [](unsigned char aC) { return static_cast<char>(::tolower(aC)); });
  bool const ascii = start.find("solid") != std::string::npos && start.find("\n") != std::string::npos
                  && start.find("facet") != std::string::npos && start.find("normal") != std::string::npos;
  probe.close();
  TheMesh raw;
  if(ascii) {
    std::ifstream in(aFilename);
    std::string line, token;
    std::vector<Vec3> pending;
    int lineCount = 1;
    while(std::getline(in, line)) {
      std::istringstream words(line);
      if(words >> token) {
        if(token == "vertex") {
          std::string cx, cy, cz;
          if(!(words >> cx >> cy >> cz)) {
            throw std::runtime_error("ERROR while reading from " + aFilename + ": vertex not specified correctly in line " + std::to_string(lineCount));
          }
          pending.push_back(Vec3{static_cast<float>(::atof(cx.c_str())), static_cast<float>(::atof(cy.c_str())), static_cast<float>(::atof(cz.c_str()))});
        }
        else if(token == "facet") {
          pending.clear();
        }
        else if(token == "endfacet") {
          if(pending.size() != 3u) {
            throw std::runtime_error("ERROR while reading from " + aFilename + ": bad number of vertices specified for face in line " + std::to_string(lineCount));
          }
          raw.push_back(Triangle{pending[0], pending[1], pending[2]});
        }
      }
      ++lineCount;
    }
  }
  else {
    std::ifstream in(aFilename, std::ios::binary);
    char header[80];
    uint32_t count = 0u;
    in.read(header, sizeof(header));
    if(!in) { throw std::runtime_error("Error while parsing binary stl header in file " + aFilename); }
    in.read(reinterpret_cast<char*>(&count), 4);
    if(!in) { throw std::runtime_error("Couldnt determine number of triangles in binary stl file " + aFilename); }
    for(uint32_t t = 0u; t < count; ++t) {
      char record[50];
      in.read(record, sizeof(record));
      if(!in) { throw std::runtime_error("Error while parsing trianlge in binary stl file " + aFilename); }
      float values[12];
      std::memcpy(values, record, sizeof(values));
      raw.push_back(Triangle{Vec3{values[3], values[4], values[5]}, Vec3{values[6], values[7], values[8]}, Vec3{values[9], values[10], values[11]}});
    }
  }
  // The reader merges corners that compare equal (so -0 and +0 are one coordinate) onto the first of each run of
  // a std::sort by (x, y, z), and drops triangles that end up with two identical corners (stl_reader.h:449-506).
  // The same sort on the same sequence picks the same representative, signed zeros included.
  struct Corner {
    float data[3];
    int32_t index;
    bool operator<(Corner const &aC) const {
      return data[0] < aC.data[0] || (data[0] == aC.data[0] && data[1] < aC.data[1]) || (data[0] == aC.data[0] && data[1] == aC.data[1] && data[2] < aC.data[2]);
    }
    bool same(Corner const &aC) const { return data[0] == aC.data[0] && data[1] == aC.data[1] && data[2] == aC.data[2]; }
  };
  std::vector<Corner> corners;
  corners.reserve(raw.size() * 3u);
  for(auto const &triangle : raw) {
    for(auto const &vertex : triangle) { corners.push_back(Corner{{vertex.x, vertex.y, vertex.z}, static_cast<int32_t>(corners.size())}); }
  }
  if(corners.empty()) { return; }
  std::sort(corners.begin(), corners.end());
  std::vector<uint32_t> merged(corners.size());       // corner in file order -> position of its representative
  size_t runStart = 0u;
  for(size_t i = 0u; i < corners.size(); ++i) {
    if(i > 0u && !corners[i].same(corners[i - 1u])) { runStart = i; }
    merged[static_cast<size_t>(corners[i].index)] = static_cast<uint32_t>(runStart);
  }
  for(size_t t = 0u; t < raw.size(); ++t) {
    uint32_t const a = merged[3u * t], b = merged[3u * t + 1u], c = merged[3u * t + 2u];
    if(a != b && a != c && b != c) {
      auto at = [&corners](uint32_t aI) { return Vec3{corners[aI].data[0], corners[aI].data[1], corners[aI].data[2]}; };
      mMesh.push_back(Triangle{at(a), at(b), at(c)});
    }
  }
}

void Mesh::writeMesh(std::string const &aFilename) const {
  std::ofstream out(aFilename);
  out << "solid Exported from Blender-2.82 (sub 7)\n";
  for(auto const &triangle : mMesh) {
    out << "facet normal 0.000000 0.000000 0.000000\nouter loop\n";
    for(auto const &vertex : triangle) { out << "vertex " << vertex.x << ' ' << vertex.y << ' ' << vertex.z << '\n'; }
    out << "endloop\nendfacet\n";
  }
  out << "endsolid Exported from Blender-2.82 (sub 7)\n";
}

// mesh.cpp:434-477.  aBelts rings of 2 * aSectors triangles between aBelts + 2 latitudes; ring vertices are
// offset by half a sector from one latitude to the next.  The quirk that the lower corner of the second
// triangle is scaled by aSize(0) once more than the ring vertices is kept (harmless for aSize(0) == 1).
void Mesh::makeSolidOfRevolution(int32_t const aSectors, int32_t const aBelts, std::function<float(float)> aEnvelope, Vector const &aSize) {
  constexpr float cPi = 3.14159265358979323846;
  mMesh.clear();
  mMesh.reserve(static_cast<uint32_t>(2 * aSectors * aBelts));
  mFace2neighbours.clear();
  float const halfSector = cPi / aSectors;
  float const fullSector = halfSector * 2.0f;
  float const beltAngle = cPi / (aBelts + 1.0f);
  float latitude[3] = { 0.0f, beltAngle, 2.0f * beltAngle };                     // up, middle, down
  float radius[3] = { 0.0f, aSize.x * aEnvelope(std::cos(latitude[1])), aSize.x * aEnvelope(std::cos(latitude[2])) };
  float height[3] = { aSize.z, aSize.z * std::cos(latitude[1]), aSize.z * std::cos(latitude[2]) };
  float bias = 0.0f;
  for(int32_t belt = 0; belt < aBelts; ++belt) {
    float angleApex = bias + halfSector;
    float angleLeft = bias + 0.0f;
    float angleRight = bias + fullSector;
    for(int32_t sector = 0; sector < aSectors; ++sector) {
      Vec3 const upper{radius[0] * std::sin(angleApex), aSize.y * radius[0] * std::cos(angleApex), height[0]};
      Vec3 const left{radius[1] * std::sin(angleLeft), aSize.y * radius[1] * std::cos(angleLeft), height[1]};
      Vec3 const right{radius[1] * std::sin(angleRight), aSize.y * radius[1] * std::cos(angleRight), height[1]};
      mMesh.push_back(Triangle{upper, left, right});
      Vec3 const lower{aSize.x * radius[2] * std::sin(angleApex), aSize.y * radius[2] * std::cos(angleApex), height[2]};
      mMesh.push_back(Triangle{left, right, lower});
      angleApex += fullSector;
      angleLeft = angleRight;
      angleRight += fullSector;
    }
    latitude[0] = latitude[1];
    latitude[1] = latitude[2];
    latitude[2] += beltAngle;
    radius[0] = radius[1];
    radius[1] = radius[2];
    radius[2] = aSize.x * aEnvelope(std::cos(latitude[2]));
    height[0] = height[1];
    height[1] = height[2];
    height[2] = aSize.z * std::cos(latitude[2]);
    bias += halfSector;
  }
}

void Mesh::makeEllipsoid(int32_t const aSectors, int32_t const aBelts, Vector const &aSize) {
  // the reference's envelope takes ::sqrt, which resolves to the double overload (mesh.h:99)
  makeSolidOfRevolution(aSectors, aBelts, [](float const aX) { return static_cast<float>(std::sqrt(static_cast<double>(1.0f - aX * aX))); }, aSize);
}

}

// Test program for the host C++ API (bz::Mesh -> bz::BezierMesh -> bz::BezierLens), compiled and run by
// tests/test_cpp_api.py.  Usage: cpp_api_check [gpu].  Without a CUDA device the lens constructor must throw
// (there is no CPU path); with "gpu" it refracts the SURVEY.md section 4 anchor ray through the unit sphere.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <vector>
#include <cuda_runtime.h>
#include "bz_bezier_lens.h"
#include "bz_hemisphere.h"

int main(int argc, char **argv) {
  bool const gpu = (argc > 1 && std::strcmp(argv[1], "gpu") == 0);
  bz::Mesh mesh;
  mesh.makeUnitSphere(16, 8);
  mesh.standardizeVertices();
  mesh.standardizeNormals();
  if(mesh.size() != 256u || mesh.getFace2neighbours().size() != 256u) { std::printf("bad mesh\n"); return 1; }
  // Mesh::getVertex2averageNormals(), reference/mesh.h:69: a map view; .at() of a corner is what BezierMesh's constructor reads
  auto const normals = mesh.getVertex2averageNormals();
  if(normals.size() != mesh.getUniqueVertices().size() || normals.count(mesh[0u][0u]) != 1u) { std::printf("bad normal map\n"); return 1; }
  bz::Vector const &corner = normals.at(mesh[0u][1u]);
  if(std::fabs(corner.x * corner.x + corner.y * corner.y + corner.z * corner.z - 1.0f) > 1e-5f) { std::printf("normal not unit\n"); return 1; }
  bz::BezierMesh bezier(mesh);
  if(bezier.size() != 768u) { std::printf("bad bezier mesh\n"); return 1; }
  auto const neighbours = bezier[0u].getNeighbours();               // patch 0 neighbours {91, 1, 2}
  if(neighbours[0] != 91u || neighbours[1] != 1u || neighbours[2] != 2u) { std::printf("bad neighbours\n"); return 1; }
  try {                                                              // skipping standardizeNormals: std::out_of_range
    bz::Mesh raw;
    raw.makeUnitSphere(8, 4);
    bz::BezierMesh broken(raw);
    std::printf("missing out_of_range\n");
    return 1;
  }
  catch(std::out_of_range const &) {}
  try {
    bz::Mesh bad;
    bad.makeEllipsoid(182, 91, bz::Vec3{1.0f, 4.0f, 2.0f});
    bad.standardizeVertices();
    bad.standardizeNormals();
    std::printf("missing throw\n");
    return 1;
  }
  catch(char const *text) { if(std::strcmp(text, "Vertex on edge detected.") != 0) { return 1; } }   // mesh.cpp:204
  bz::UniformHemisphere hemisphere(16u);                             // host part of the emitter sampler (hostUtil.cpp:3-14)
  if(hemisphere.getPatchCount() != 660u || hemisphere.getBelt(0u).second != 0u || hemisphere.getBelt(1u).second != 4u) {
    std::printf("bad hemisphere %u\n", hemisphere.getPatchCount());
    return 1;
  }
  try {
    bz::BezierLens original(1.3f, bz::BezierMesh(mesh));             // BezierLens(float, BezierMesh&&), reference/bezierLens.h:24
    if(!gpu) { std::printf("lens built without a device?\n"); return 1; }
    bz::BezierLens lens(original);                                    // copies share the device tables (the reference copies the mesh)
    { bz::BezierLens moved(std::move(original)); if(moved.handle() != lens.handle()) { std::printf("copy does not share\n"); return 1; } }
    bz::Ray ray(bz::Vec3{-10.0f, 0.1f, 0.2f}, bz::Vec3{2.0f, 0.0f, 0.0f});
    auto const hit = lens.intersect(ray);
    if(hit.mWhat != bz::BezierIntersection::What::cIntersect || hit.mPatch != 350u) { std::printf("bad hit %u\n", hit.mPatch); return 1; }
    auto const in = lens.refract(ray, bz::RefractionResult::cInside);
    if(in.second != bz::RefractionResult::cInside) { std::printf("no entry\n"); return 1; }
    auto const out = lens.refract(in.first, bz::RefractionResult::cOutside);
    if(out.second != bz::RefractionResult::cOutside || std::fabs(out.first.mDirection.x - 0.99463236f) > 2e-6f) { std::printf("bad exit\n"); return 1; }
    auto const wrong = lens.refract(ray, bz::RefractionResult::cOutside);
    if(wrong.second != bz::RefractionResult::cNone) { std::printf("expected-side check failed\n"); return 1; }
    // illumination mode: a small emitter in front of the sphere, a target behind it
    bz_emitter emitter = { {-3.0f, -0.1f, -0.1f}, {0.0f, 0.2f, 0.0f}, {0.0f, 0.0f, 0.2f}, {1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f},
                           2u, 2u, 2048u, 16u, 7u };
    bz_target target = { {4.0f, -2.0f, -2.0f}, {0.0f, 4.0f, 0.0f}, {0.0f, 0.0f, 4.0f}, 8u, 8u };
    uint32_t bins[64] = {};
    bz_stats stats = {};
    lens.illuminate(emitter, target, 2u, bins, &stats);
    uint64_t landed = 0u;
    for(uint32_t b : bins) { landed += b; }
    if(stats.primary_rays != 8192u || stats.entered == 0u || landed == 0u || landed > 8192u) { std::printf("bad illumination %llu\n", (unsigned long long)landed); return 1; }
    // row-sharded render without PyTorch: communicator of one rank, this rank's stripes, gather + assembly on the root
    {
      bz_camera camera = { {-12.0f, 0.0f, 0.0f}, {1.0f, 0.0f, 0.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}, 0.14f, 0.105f, 96u, 72u, 2u, 0u };
      bz_scene scene = { {1.0f, 0.0f, 0.0f, 12.0f}, {0.0f, 1.0f, 0.0f}, {0.0f, 0.0f, 1.0f}, 1.0f, 25u, 230u, 0u, 2u };
      std::vector<uint8_t> whole(96u * 72u * 4u), gathered(96u * 72u * 4u);
      lens.render(camera, scene, 0u, 72u, whole.data());
      uint8_t id[BZ_COMM_ID_BYTES];
      bz_comm *comm = nullptr;
      if(bz_comm_unique_id(id) != BZ_OK || bz_comm_create(id, 0u, 1u, 0, &comm) != BZ_OK) { std::printf("no communicator: %s\n", bz_last_error()); return 1; }
      uint8_t *dBand = nullptr, *dImage = nullptr;
      if(cudaMalloc((void**)&dBand, whole.size()) != cudaSuccess || cudaMalloc((void**)&dImage, whole.size()) != cudaSuccess) { std::printf("cudaMalloc failed\n"); return 1; }
      int rc = bz_render_striped(lens.handle(), &camera, &scene, 8u, 0u, 1u, dBand, nullptr, nullptr);
      if(rc == BZ_OK) { rc = bz_gather_rows(comm, dBand, camera.width, camera.height, 8u, 0u, dImage, nullptr); }
      if(rc != BZ_OK || cudaMemcpy(gathered.data(), dImage, gathered.size(), cudaMemcpyDeviceToHost) != cudaSuccess) { std::printf("gather failed: %s\n", bz_last_error()); return 1; }
      if(gathered != whole) { std::printf("gathered frame differs\n"); return 1; }
      cudaFree(dBand); cudaFree(dImage);
      bz_comm_destroy(comm);
    }
    std::printf("ok gpu\n");
  }
  catch(std::runtime_error const &e) {
    if(gpu) { std::printf("lens failed: %s\n", e.what()); return 1; }
    std::printf("ok cpu (%s)\n", e.what());
  }
  return 0;
}

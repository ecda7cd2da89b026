// bz_host_capi.cpp -- C ABI over the host preprocessing classes (bz::Mesh, bz::BezierMesh): the bz_mesh_* and
// bz_bezier_* entry points of include/bz_capi.h.  No exception crosses the boundary: the reference's
// `throw "Vertex on edge detected."` (char const*), std::runtime_error and std::out_of_range all become
// BZ_ERR_PREPROCESS with the message kept for bz_last_error().
#include <cstring>
#include <exception>
#include <new>
#include <string>
#include "bz_capi.h"
#include "bz_bezier_mesh.h"
#include "bz_error.h"
#include "bz_mesh.h"

struct bz_mesh { bz::Mesh mesh; };
struct bz_bezier { bz::BezierMesh bezier; };

namespace {

template<typename tBody>
int guarded(tBody &&aBody) {
  try {
    return aBody();
  }
  catch(char const *text) { return bz_fail(BZ_ERR_PREPROCESS, text); }
  catch(std::bad_alloc const &) { return bz_fail(BZ_ERR_PREPROCESS, "out of host memory"); }
  catch(std::exception const &e) { return bz_fail(BZ_ERR_PREPROCESS, e.what()); }
  catch(...) { return bz_fail(BZ_ERR_PREPROCESS, "unknown exception in host preprocessing"); }
}

int nullArgument() { return bz_fail(BZ_ERR_INVALID, "null argument"); }

}

extern "C" {

int bz_mesh_create(bz_mesh **out_mesh) {
  if(out_mesh == nullptr) { return nullArgument(); }
  return guarded([&] { *out_mesh = new bz_mesh(); return BZ_OK; });
}

int bz_mesh_destroy(bz_mesh *mesh) {
  delete mesh;
  return BZ_OK;
}

int bz_mesh_clone(const bz_mesh *mesh, bz_mesh **out_mesh) {
  if(mesh == nullptr || out_mesh == nullptr) { return nullArgument(); }
  return guarded([&] { *out_mesh = new bz_mesh(*mesh); return BZ_OK; });
}

int bz_mesh_make_ellipsoid(bz_mesh *mesh, int32_t sectors, int32_t belts, const float size[3]) {
  if(mesh == nullptr || size == nullptr) { return nullArgument(); }
  if(sectors <= 0 || belts <= 0) { return bz_fail(BZ_ERR_INVALID, "sectors and belts must be positive"); }
  return guarded([&] { mesh->mesh.makeEllipsoid(sectors, belts, bz::Vec3{size[0], size[1], size[2]}); return BZ_OK; });
}

int bz_mesh_make_unit_sphere(bz_mesh *mesh, int32_t sectors, int32_t belts) {
  float const one[3] = { 1.0f, 1.0f, 1.0f };
  return bz_mesh_make_ellipsoid(mesh, sectors, belts, one);
}

int bz_mesh_read_stl(bz_mesh *mesh, const char *path) {
  if(mesh == nullptr || path == nullptr) { return nullArgument(); }
  return guarded([&] { mesh->mesh.readMesh(path); return BZ_OK; });
}

int bz_mesh_write_stl(const bz_mesh *mesh, const char *path) {
  if(mesh == nullptr || path == nullptr) { return nullArgument(); }
  return guarded([&] { mesh->mesh.writeMesh(path); return BZ_OK; });
}

int bz_mesh_set_triangles(bz_mesh *mesh, const float *triangles, uint32_t count) {
  if(mesh == nullptr || (triangles == nullptr && count != 0u)) { return nullArgument(); }
  return guarded([&] {
    mesh->mesh.clear();
    mesh->mesh.reserve(count);
    for(uint32_t i = 0u; i < count; ++i) {
      float const *t = triangles + static_cast<size_t>(i) * 9u;
      mesh->mesh.push_back(bz::Triangle{bz::Vec3{t[0], t[1], t[2]}, bz::Vec3{t[3], t[4], t[5]}, bz::Vec3{t[6], t[7], t[8]}});
    }
    return BZ_OK;
  });
}

int bz_mesh_size(const bz_mesh *mesh, uint32_t *out_count) {
  if(mesh == nullptr || out_count == nullptr) { return nullArgument(); }
  *out_count = static_cast<uint32_t>(mesh->mesh.size());
  return BZ_OK;
}

int bz_mesh_get_triangles(const bz_mesh *mesh, float *out_triangles) {
  if(mesh == nullptr || (out_triangles == nullptr && mesh->mesh.size() != 0u)) { return nullArgument(); }
  float *out = out_triangles;
  for(auto const &triangle : mesh->mesh) {
    for(auto const &vertex : triangle) { *out++ = vertex.x; *out++ = vertex.y; *out++ = vertex.z; }
  }
  return BZ_OK;
}

int bz_mesh_transform(bz_mesh *mesh, const float matrix_row_major[9], const float displacement[3]) {
  if(mesh == nullptr || matrix_row_major == nullptr || displacement == nullptr) { return nullArgument(); }
  bz::Mat3 m;
  std::memcpy(m.m, matrix_row_major, sizeof(m.m));
  mesh->mesh.transform(m, bz::Vec3{displacement[0], displacement[1], displacement[2]});
  return BZ_OK;
}

int bz_mesh_split_triangles(bz_mesh *mesh, int32_t divisor) {
  if(mesh == nullptr) { return nullArgument(); }
  if(divisor <= 0) { return bz_fail(BZ_ERR_INVALID, "divisor must be positive"); }
  return guarded([&] { mesh->mesh.splitTriangles(divisor); return BZ_OK; });
}

int bz_mesh_split_triangles_by_side(bz_mesh *mesh, float max_side) {
  if(mesh == nullptr) { return nullArgument(); }
  if(!(max_side > 0.0f)) { return bz_fail(BZ_ERR_INVALID, "max_side must be positive"); }
  return guarded([&] { mesh->mesh.splitTriangles(max_side); return BZ_OK; });
}

int bz_mesh_standardize_vertices(bz_mesh *mesh) {
  if(mesh == nullptr) { return nullArgument(); }
  return guarded([&] { mesh->mesh.standardizeVertices(); return BZ_OK; });
}

int bz_mesh_standardize_normals(bz_mesh *mesh) {
  if(mesh == nullptr) { return nullArgument(); }
  return guarded([&] { mesh->mesh.standardizeNormals(); return BZ_OK; });
}

int bz_mesh_neighbour_count(const bz_mesh *mesh, uint32_t *out_count) {
  if(mesh == nullptr || out_count == nullptr) { return nullArgument(); }
  *out_count = static_cast<uint32_t>(mesh->mesh.getFace2neighbours().size());
  return BZ_OK;
}

int bz_mesh_get_face2neighbours(const bz_mesh *mesh, uint32_t *out_fellow, uint8_t *out_starts) {
  if(mesh == nullptr) { return nullArgument(); }
  auto const &table = mesh->mesh.getFace2neighbours();
  if(!table.empty() && (out_fellow == nullptr || out_starts == nullptr)) { return nullArgument(); }
  for(size_t f = 0u; f < table.size(); ++f) {
    for(uint32_t i = 0u; i < 3u; ++i) {
      out_fellow[f * 3u + i] = table[f].mFellowTriangles[i];
      out_starts[f * 3u + i] = table[f].mFellowCommonSideStarts[i];
    }
  }
  return BZ_OK;
}

int bz_mesh_get_corner_normals(const bz_mesh *mesh, float *out_normals) {
  if(mesh == nullptr || (out_normals == nullptr && mesh->mesh.size() != 0u)) { return nullArgument(); }
  return guarded([&] {
    float *out = out_normals;
    for(auto const &triangle : mesh->mesh) {
      for(auto const &vertex : triangle) {
        bz::Vec3 const &n = mesh->mesh.getAverageNormal(vertex);
        *out++ = n.x; *out++ = n.y; *out++ = n.z;
      }
    }
    return BZ_OK;
  });
}

int bz_bezier_create(const bz_mesh *mesh, bz_bezier **out_bezier) {
  if(mesh == nullptr || out_bezier == nullptr) { return nullArgument(); }
  *out_bezier = nullptr;
  return guarded([&] { *out_bezier = new bz_bezier{bz::BezierMesh(mesh->mesh)}; return BZ_OK; });
}

int bz_bezier_destroy(bz_bezier *bezier) {
  delete bezier;
  return BZ_OK;
}

int bz_bezier_size(const bz_bezier *bezier, uint32_t *out_count) {
  if(bezier == nullptr || out_count == nullptr) { return nullArgument(); }
  *out_count = static_cast<uint32_t>(bezier->bezier.size());
  return BZ_OK;
}

int bz_bezier_get_patches(const bz_bezier *bezier, bz_patch *out_patches) {
  if(bezier == nullptr || (out_patches == nullptr && bezier->bezier.size() != 0u)) { return nullArgument(); }
  if(bezier->bezier.size() != 0u) { std::memcpy(out_patches, bezier->bezier.records(), bezier->bezier.size() * sizeof(bz_patch)); }
  return BZ_OK;
}

int bz_bezier_interpolate(const bz_bezier *bezier, int32_t divisor, bz_mesh **out_mesh) {
  if(bezier == nullptr || out_mesh == nullptr) { return nullArgument(); }
  if(divisor <= 0) { return bz_fail(BZ_ERR_INVALID, "divisor must be positive"); }
  return guarded([&] { *out_mesh = new bz_mesh{bezier->bezier.interpolate(divisor)}; return BZ_OK; });
}

int bz_bezier_split_thick(const bz_bezier *bezier, bz_mesh **out_mesh) {
  if(bezier == nullptr || out_mesh == nullptr) { return nullArgument(); }
  return guarded([&] { *out_mesh = new bz_mesh{bezier->bezier.splitThickBezierTriangles()}; return BZ_OK; });
}

int bz_lens_create_from_bezier(const bz_bezier *bezier, float refractive_index, int device, bz_lens **out_lens) {
  if(bezier == nullptr || out_lens == nullptr) { return nullArgument(); }
  return bz_lens_create(bezier->bezier.records(), static_cast<uint32_t>(bezier->bezier.size()), refractive_index, device, out_lens);
}

}

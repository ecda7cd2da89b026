// TEST INFRASTRUCTURE ONLY -- runs the product's device arithmetic (csrc/bz_math.cuh) and the
// cull-hierarchy logic (host/bz_bvh.cpp + the slab test the kernel uses) on the CPU, so that the
// `-m "not gpu"` suite can compare them with the oracle before GPU time is spent.
// It is compiled by tests/conftest.py into tests/_build/, is not part of libbz_b200.so and is not
// a CPU fallback: the product library has no host compute path.
// Build: g++ -O2 -ffp-contract=off (no FMA contraction, like nvcc -fmad=false on the device).
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>
#include "bz_capi.h"
#include "../cuda-bezier-triangle-raytracer_b200/csrc/bz_math.cuh"
#include "../cuda-bezier-triangle-raytracer_b200/host/bz_bvh.h"

namespace {

void unpack(const bz_patch &p, bz_cull_rec &c, bz_newton_rec &t) {
  c.n = bz_v(p.plane[0], p.plane[1], p.plane[2]); c.k = p.plane[3];
  c.b0 = bz_v(p.bary_inv[0][0], p.bary_inv[0][1], p.bary_inv[0][2]);
  c.b1 = bz_v(p.bary_inv[1][0], p.bary_inv[1][1], p.bary_inv[1][2]);
  c.b2 = bz_v(p.bary_inv[2][0], p.bary_inv[2][1], p.bary_inv[2][2]);
  c.h_in = p.height_inside; c.h_out = p.height_outside;
  for(int i = 0; i < 10; ++i) { t.cp[i] = bz_v(p.control[i][0], p.control[i][1], p.control[i][2]); }
  t.dir_b = bz_v(p.dir_b[0], p.dir_b[1], p.dir_b[2]);
  for(int i = 0; i < 3; ++i) {
    t.div_n[i] = bz_v(p.divider[i][0], p.divider[i][1], p.divider[i][2]);
    t.div_k[i] = p.divider[i][3];
    t.nb[i] = p.neighbours[i];
  }
}

struct Best {
  unsigned long long key;
  bz_hit hit;
  uint32_t patch;
};

// one loop iteration of bezierMesh.cpp:211-225 for patch `index`, given it passed stage A
void candidate(const bz_patch *patches, uint32_t index, const bzv3 &s, const bzv3 &d, Best &best) {
  bz_cull_rec c; bz_newton_rec t;
  unpack(patches[index], c, t);
  float cosine, distance;
  if(!bz_stage_a(c, s, d, true, cosine, distance)) { return; }
  bz_hit hit;
  bz_stage_b(c, t, s, d, cosine, distance, hit);
  uint32_t evaluated = index;
  if(hit.what <= BZ_FOLLOW2) {
    evaluated = t.nb[hit.what];
    unpack(patches[evaluated], c, t);
    hit.what = BZ_MISS;
    if(bz_stage_a(c, s, d, false, cosine, distance)) { bz_stage_b(c, t, s, d, cosine, distance, hit); }
  }
  if(hit.what == BZ_HIT && hit.distance < 3.402823466e+38f) {
    unsigned long long key = ((unsigned long long)bz_float_key(hit.distance) << 32) | index;
    if(key < best.key) { best.key = key; best.hit = hit; best.patch = evaluated; }
  }
}

}

extern "C" {

// mode 0: every patch; mode 1: through the cull hierarchy.  visited[0] += stage-A tests, visited[1] += nodes.
void shim_intersect_batch(const bz_patch *patches, uint32_t count, const float *start, const float *dir, uint64_t n, int mode,
                          uint32_t *out_what, uint32_t *out_patch, float *out_distance, float *out_point, float *out_normal,
                          float *out_cos, float *out_bary, uint64_t *visited) {
  bz_bvh bvh;
  if(mode == 1) { bz_build_bvh(patches, count, bvh); }
  for(uint64_t r = 0u; r < n; ++r) {
    bzv3 s = bz_v(start[r * 3u], start[r * 3u + 1u], start[r * 3u + 2u]);
    bzv3 d = bz_v(dir[r * 3u], dir[r * 3u + 1u], dir[r * 3u + 2u]);
    Best best;
    best.key = ~0ull;
    best.patch = BZ_NO_PATCH;
    if(mode == 0) {
      for(uint32_t i = 0u; i < count; ++i) { candidate(patches, i, s, d, best); }
      visited[0] += count;
    }
    else {
      float slack = 3.0517578125e-05f * (std::fabs(s.x) + std::fabs(s.y) + std::fabs(s.z) + 2.0f * bvh.extent);
      float o[3] = { s.x, s.y, s.z }, dd[3] = { d.x, d.y, d.z }, inv[3], alo[3], ahi[3];
      for(int k = 0; k < 3; ++k) {
        inv[k] = (std::fabs(dd[k]) > 1e-20f) ? 1.0f / dd[k] : std::copysign(1e20f, dd[k]);
        alo[k] = -(o[k] + slack) * inv[k];
        ahi[k] = -(o[k] - slack) * inv[k];
      }
      uint32_t const nodeCount = (uint32_t)(bvh.nodes.size() / 8u);
      uint32_t node = 0u;
      while(node < nodeCount) {                     // depth-first order with skip links, as the kernel walks it
        ++visited[1];
        const float *p = bvh.nodes.data() + (size_t)node * 8u;
        uint32_t link, code;
        std::memcpy(&link, p + 3, sizeof(link));
        std::memcpy(&code, p + 7, sizeof(code));
        float tnear = 0.0f, tfar = 3.0e38f;
        for(int a = 0; a < 3; ++a) {
          float t0 = std::fma(p[a], inv[a], alo[a]), t1 = std::fma(p[4 + a], inv[a], ahi[a]);
          tnear = std::fmax(tnear, std::fmin(t0, t1));
          tfar = std::fmin(tfar, std::fmax(t0, t1));
        }
        bool const hit = (tnear <= tfar * 1.000001f);
        if(hit && code != 0u) {
          for(uint32_t j = 0u; j < (code & 7u); ++j) { candidate(patches, bvh.order[(code >> 3) + j], s, d, best); }
          visited[0] += (code & 7u);
        }
        node = (hit && code == 0u) ? node + 1u : link;
      }
      if(!(std::fabs(s.x) + std::fabs(s.y) + std::fabs(s.z) <= bvh.origin_limit)) {
        for(uint32_t i : bvh.wide) { candidate(patches, i, s, d, best); }
        visited[0] += bvh.wide.size();
      }
      for(uint32_t i : bvh.always) { candidate(patches, i, s, d, best); }
      visited[0] += bvh.always.size();
    }
    bool ok = (best.key != ~0ull);
    out_what[r] = ok ? BZ_WHAT_INTERSECT : BZ_WHAT_NONE;
    out_patch[r] = ok ? best.patch : BZ_NO_PATCH;
    out_distance[r] = ok ? best.hit.distance : 0.0f;
    out_cos[r] = ok ? best.hit.cos_incidence : 0.0f;
    const float pt[3] = { best.hit.point.x, best.hit.point.y, best.hit.point.z };
    const float nr[3] = { best.hit.normal.x, best.hit.normal.y, best.hit.normal.z };
    const float ba[3] = { best.hit.bary.x, best.hit.bary.y, best.hit.bary.z };
    for(int k = 0; k < 3; ++k) {
      out_point[r * 3u + k] = ok ? pt[k] : 0.0f;
      out_normal[r * 3u + k] = ok ? nr[k] : 0.0f;
      out_bary[r * 3u + k] = ok ? ba[k] : 0.0f;
    }
  }
}

// Snell block alone: returns status before the expected-side check, dir updated in place
uint32_t shim_snell(float ri, float cos_incidence, const float *normal, float *dir) {
  bzv3 d = bz_v(dir[0], dir[1], dir[2]);
  uint32_t status = bz_snell(ri, cos_incidence, bz_v(normal[0], normal[1], normal[2]), d);
  dir[0] = d.x; dir[1] = d.y; dir[2] = d.z;
  return status;
}

void shim_bvh_info(const bz_patch *patches, uint32_t count, uint32_t *out) {
  bz_bvh bvh;
  bz_build_bvh(patches, count, bvh);
  out[0] = (uint32_t)(bvh.nodes.size() / 8u);
  out[1] = bvh.depth;
  out[2] = (uint32_t)bvh.always.size();
  out[3] = (uint32_t)bvh.order.size();
  out[4] = (uint32_t)bvh.wide.size();
}

// indices of the wide patches (hierarchy members with grown boxes); returns their count
uint32_t shim_bvh_wide(const bz_patch *patches, uint32_t count, uint32_t *out, uint32_t capacity, float *out_origin_limit) {
  bz_bvh bvh;
  bz_build_bvh(patches, count, bvh);
  for(uint32_t i = 0u; i < bvh.wide.size() && i < capacity; ++i) { out[i] = bvh.wide[i]; }
  *out_origin_limit = bvh.origin_limit;
  return (uint32_t)bvh.wide.size();
}

}

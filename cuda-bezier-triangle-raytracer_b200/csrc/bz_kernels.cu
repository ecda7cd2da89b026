// bz_kernels.cu -- sm_100a kernels of the Bezier-triangle intersect/refract path.
//
// One persistent kernel, bz_trace_kernel<JOB, BRUTE>, serves the three entry points of the C ABI:
//   JOB_RENDER     K1 ray generation + bounce loop around BezierLens::refract + K3 background shading
//   JOB_REFRACT    BezierLens::refract over a ray batch   (reference/bezierLens.cpp:4-34)
//   JOB_INTERSECT  BezierMesh::intersect over a ray batch (reference/bezierMesh.cpp:206-227)
// Every CTA owns a pool of BZ_NT ray slots (one per thread).  One pool iteration is one refract()
// call for every live slot:
//   1. refill   empty slots take the next primary-ray indices from a global counter (one atomic per
//               warp) and generate / load their ray -- live rays stay compacted across bounces
//   2. stage A  BRUTE: every patch, reference loop order, cull records staged through shared memory
//               culled: 4-wide BVH over the flat triangles, exact stage A at the leaves
//               survivors (slot, patch) are appended to a shared-memory queue
//   3. stage B  the queue is evaluated with full warps (bracket + 4 root-search steps); follow-side
//               results go to a second queue and are evaluated against the neighbour patch
//               the nearest hit per slot is a 64-bit atomicMin on (distance, loop index), which is the
//               reference's "first strictly smaller distance wins" rule (bezierMesh.cpp:220)
//   4. Snell    per slot: refraction, expected-side check, termination / shading
// Compiled with -fmad=false: see bz_math.cuh.  The box tests of the cull use explicit FMAs; they
// only have to be conservative.
#include <cuda_runtime.h>
#include <stdint.h>
#include <algorithm>
#include "bz_device.h"
#include "bz_math.cuh"

#define BZ_NT        256
#define BZ_QA        1024
#define BZ_QB        256
#define BZ_PAY       11
#define BZ_TILE      256
#define BZ_KEY_NONE  0xffffffffffffffffull
#define BZ_NSTAT     9

#define ST_PRIMARY 0
#define ST_CALLS   1
#define ST_A       2
#define ST_B       3
#define ST_SNELL   4
#define ST_ENTER   5
#define ST_EXIT    6
#define ST_LOST    7
#define ST_NODES   8

__device__ __forceinline__ void bz_unpack_cull(const float4 &r0, const float4 &r1, const float4 &r2, const float4 &r3, bz_cull_rec &c) {
  c.n = bz_v(r0.x, r0.y, r0.z); c.k = r0.w;
  c.b0 = bz_v(r1.x, r1.y, r1.z); c.h_in = r1.w;
  c.b1 = bz_v(r2.x, r2.y, r2.z); c.h_out = r2.w;
  c.b2 = bz_v(r3.x, r3.y, r3.z);
}

__device__ __forceinline__ uint32_t bz_load_cull(const float4 *base, uint32_t index, bz_cull_rec &c) {
  const float4 *p = base + (size_t)index * BZ_CULL_F4;
  float4 r0 = __ldg(p), r1 = __ldg(p + 1), r2 = __ldg(p + 2), r3 = __ldg(p + 3);
  bz_unpack_cull(r0, r1, r2, r3, c);
  return __float_as_uint(r3.w);
}

__device__ __forceinline__ void bz_load_newton(const float4 *base, uint32_t index, bz_newton_rec &t) {
  const float4 *p = base + (size_t)index * BZ_NEWTON_F4;
  float w[48];
#pragma unroll
  for(int i = 0; i < BZ_NEWTON_F4; ++i) {
    float4 r = __ldg(p + i);
    w[4 * i] = r.x; w[4 * i + 1] = r.y; w[4 * i + 2] = r.z; w[4 * i + 3] = r.w;
  }
#pragma unroll
  for(int i = 0; i < 10; ++i) { t.cp[i] = bz_v(w[2 * i], w[2 * i + 1], w[20 + i]); }
  t.dir_b = bz_v(w[30], w[31], w[32]);
#pragma unroll
  for(int i = 0; i < 3; ++i) {
    t.div_n[i] = bz_v(w[33 + 4 * i], w[34 + 4 * i], w[35 + 4 * i]);
    t.div_k[i] = w[36 + 4 * i];
    t.nb[i] = __float_as_uint(w[45 + i]);
  }
}

// K1: primary ray of sample (px, py, sx, sy); definitions in include/bz_capi.h (bz_camera).
__device__ __forceinline__ void bz_primary_ray(const bz_camera &cam, uint32_t px, uint32_t py, uint32_t sx, uint32_t sy, bzv3 &start, bzv3 &dir) {
  float strata = (float)cam.strata;
  float fu = (float)px + ((float)sx + 0.5f) / strata;
  float fv = (float)py + ((float)sy + 0.5f) / strata;
  float a = (fu * (2.0f / (float)cam.width) - 1.0f) * cam.half_w;
  float b = (1.0f - fv * (2.0f / (float)cam.height)) * cam.half_h;
  bzv3 offset = bz_v(cam.right[0] * a + cam.up[0] * b, cam.right[1] * a + cam.up[1] * b, cam.right[2] * a + cam.up[2] * b);
  bzv3 raw;
  if(cam.ortho != 0u) {
    start = bz_v(cam.eye[0] + offset.x, cam.eye[1] + offset.y, cam.eye[2] + offset.z);
    raw = bz_v(cam.forward[0], cam.forward[1], cam.forward[2]);
  }
  else {
    start = bz_v(cam.eye[0], cam.eye[1], cam.eye[2]);
    raw = bz_v(cam.forward[0] + offset.x, cam.forward[1] + offset.y, cam.forward[2] + offset.z);
  }
  dir = bz_normalized(raw);    // Ray::Ray, 3dGeomUtil.h:176-178
}

// K3: checker on the background plane, Plane::intersect semantics (3dGeomUtil.h:279-296).
__device__ __forceinline__ uint32_t bz_shade_background(const bz_scene &sc, const bzv3 &start, const bzv3 &dir) {
  bzv3 n = bz_v(sc.bg_plane[0], sc.bg_plane[1], sc.bg_plane[2]);
  float cosine = bz_dot(dir, n);
  if(!(fabsf(cosine) >= BZ_PLANE_EPSILON)) { return sc.value_lost; }
  float distance = (sc.bg_plane[3] - bz_dot(n, start)) / cosine;
  if(!(distance > 0.0f)) { return sc.value_lost; }
  bzv3 p = bz_v(start.x + distance * dir.x, start.y + distance * dir.y, start.z + distance * dir.z);
  float ca = floorf(bz_dot(p, bz_v(sc.bg_axis_a[0], sc.bg_axis_a[1], sc.bg_axis_a[2])) / sc.cell);
  float cb = floorf(bz_dot(p, bz_v(sc.bg_axis_b[0], sc.bg_axis_b[1], sc.bg_axis_b[2])) / sc.cell);
  float pa = fabsf(fmodf(ca, 2.0f));
  float pb = fabsf(fmodf(cb, 2.0f));
  return (pa != pb) ? sc.value_bright : sc.value_dark;
}

// BezierTriangle::intersect(ray, limit) for one (ray, patch) pair; the single out-of-line copy of
// stage B in the kernel image.  Returns the number of stage-B evaluations done (0 or 1).
__device__ __noinline__ uint32_t bz_eval_pair(const float4 *cull, const float4 *newton, float sx, float sy, float sz,
                                              float dx, float dy, float dz, uint32_t patch, uint32_t limit_this, bz_hit *out, uint32_t *out_nb) {
  bz_cull_rec c;
  bz_load_cull(cull, patch, c);
  bzv3 s = bz_v(sx, sy, sz), d = bz_v(dx, dy, dz);
  float cosine, distance;
  bz_hit hit;
  hit.what = BZ_MISS;
  uint32_t evaluated = 0u;
  if(bz_stage_a(c, s, d, limit_this != 0u, cosine, distance)) {
    bz_newton_rec t;
    bz_load_newton(newton, patch, t);
    bz_stage_b(c, t, s, d, cosine, distance, hit);
    evaluated = 1u;
    if(hit.what <= BZ_FOLLOW2) { *out_nb = t.nb[hit.what]; }
  }
  *out = hit;
  return evaluated;
}

__device__ __forceinline__ unsigned long long bz_make_key(float distance, uint32_t loop_index) {
  return ((unsigned long long)bz_float_key(distance) << 32) | (unsigned long long)loop_index;
}

__device__ __forceinline__ float bz_key_distance(unsigned long long key) {
  uint32_t k = (uint32_t)(key >> 32);
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// a candidate takes part in the nearest-hit selection iff it is cIntersect and its distance is below the
// initial numeric_limits<float>::max() (bezierMesh.cpp:209, 220)
__device__ __forceinline__ bool bz_selectable(const bz_hit &hit) {
  return hit.what == BZ_HIT && hit.distance < 3.402823466e+38f;
}

struct bz_pool {
  float              ray[6][BZ_NT];        // start xyz, dir xyz of every slot
  unsigned long long best[BZ_NT];          // (distance key, loop index) of the nearest hit so far
  float              pay[BZ_PAY][BZ_NT];   // point 3, normal 3, cos, bary 3, evaluated patch
  uint2              qa[BZ_QA];            // stage-A survivors: (slot, patch)
  uint32_t           qb_slot[BZ_QB];       // follow-side hops: slot, neighbour patch, loop index of the originating patch
  uint32_t           qb_patch[BZ_QB];
  uint32_t           qb_loop[BZ_QB];
  uint32_t           na, nb;
};

__device__ __forceinline__ void bz_store_payload(bz_pool &sm, uint32_t slot, const bz_hit &hit, uint32_t patch) {
  sm.pay[0][slot] = hit.point.x;  sm.pay[1][slot] = hit.point.y;  sm.pay[2][slot] = hit.point.z;
  sm.pay[3][slot] = hit.normal.x; sm.pay[4][slot] = hit.normal.y; sm.pay[5][slot] = hit.normal.z;
  sm.pay[6][slot] = hit.cos_incidence;
  sm.pay[7][slot] = hit.bary.x;   sm.pay[8][slot] = hit.bary.y;   sm.pay[9][slot] = hit.bary.z;
  sm.pay[10][slot] = __uint_as_float(patch);
}

// One stage-A survivor of slot `slot`: queue it for the dense stage-B pass, or -- queue full --
// evaluate it right here (follow hop included) into the owner's private best.
__device__ __forceinline__ void bz_push_candidate(bz_pool &sm, const bz_tables &T, uint32_t slot, uint32_t patch, const bzv3 &s, const bzv3 &d,
                                                  unsigned long long &ikey, bz_hit &ihit, uint32_t &ipatch, unsigned long long &count_a, uint32_t &count_b) {
  uint32_t at = atomicAdd(&sm.na, 1u);
  if(at < BZ_QA) {
    sm.qa[at] = make_uint2(slot, patch);
    return;
  }
  bz_hit hit;
  uint32_t nbp = 0u, evaluated = patch;
  count_b += bz_eval_pair(T.cull, T.newton, s.x, s.y, s.z, d.x, d.y, d.z, patch, 1u, &hit, &nbp);
  if(hit.what <= BZ_FOLLOW2) {
    evaluated = nbp;
    uint32_t unused = 0u;
    count_a += 1u;
    count_b += bz_eval_pair(T.cull, T.newton, s.x, s.y, s.z, d.x, d.y, d.z, nbp, 0u, &hit, &unused);
  }
  if(bz_selectable(hit)) {
    unsigned long long key = bz_make_key(hit.distance, patch);
    if(key < ikey) { ikey = key; ihit = hit; ipatch = evaluated; }
  }
}

template<int JOB, bool BRUTE>
__global__ void __launch_bounds__(BZ_NT, 2) bz_trace_kernel(const bz_tables T, const bz_job J) {
  __shared__ bz_pool sm;
  const uint32_t t = threadIdx.x;
  const uint32_t lane = t & 31u;
  bool live = false;
  unsigned long long ray_id = 0ull;
  bzv3 s = bz_v(0.0f, 0.0f, 0.0f), d = s;
  uint32_t expected = BZ_STATUS_INSIDE, crossings = 0u, pixel = 0u;
  unsigned long long c_primary = 0ull, c_calls = 0ull, c_a = 0ull, c_nodes = 0ull;
  uint32_t c_b = 0u, c_snell = 0u, c_enter = 0u, c_exit = 0u, c_lost = 0u;
  const uint32_t spp = J.camera.strata * J.camera.strata;

  for(;;) {
    // ---- 1. refill: dead slots take the next primary rays, live rays stay where they are ----
    uint32_t need = __ballot_sync(0xffffffffu, !live);
    if(need != 0u) {
      unsigned long long base = 0ull;
      int leader = __ffs(need) - 1;
      if((int)lane == leader) { base = atomicAdd(J.next_ray, (unsigned long long)__popc(need)); }
      base = __shfl_sync(0xffffffffu, base, leader);
      if(!live) {
        unsigned long long id = base + (unsigned long long)__popc(need & ((1u << lane) - 1u));
        if(id < J.ray_count) {
          ray_id = id;
          live = true;
          crossings = 0u;
          ++c_primary;
          if(JOB == BZ_JOB_RENDER) {
            pixel = (uint32_t)(id / spp);
            uint32_t sample = (uint32_t)(id - (unsigned long long)pixel * spp);
            uint32_t row = pixel / J.camera.width;
            uint32_t px = pixel - row * J.camera.width;
            uint32_t py = (J.stripe_rows == 0u) ? J.row0 + row
                        : ((row / J.stripe_rows) * J.stripe_ranks + J.stripe_rank) * J.stripe_rows + row % J.stripe_rows;
            bz_primary_ray(J.camera, px, py, sample % J.camera.strata, sample / J.camera.strata, s, d);
            expected = BZ_STATUS_INSIDE;
          }
          else {
            s = bz_v(J.in_start[id * 3u], J.in_start[id * 3u + 1u], J.in_start[id * 3u + 2u]);
            d = bz_v(J.in_dir[id * 3u], J.in_dir[id * 3u + 1u], J.in_dir[id * 3u + 2u]);
            expected = (JOB == BZ_JOB_REFRACT) ? J.in_expected[id] : 0u;
          }
        }
      }
    }
    sm.ray[0][t] = s.x; sm.ray[1][t] = s.y; sm.ray[2][t] = s.z;
    sm.ray[3][t] = d.x; sm.ray[4][t] = d.y; sm.ray[5][t] = d.z;
    sm.best[t] = BZ_KEY_NONE;
    if(t == 0u) { sm.na = 0u; sm.nb = 0u; }
    unsigned long long ikey = BZ_KEY_NONE;
    bz_hit ihit;
    ihit.what = BZ_MISS;
    uint32_t ipatch = BZ_NO_PATCH;
    if(__syncthreads_or(live ? 1 : 0) == 0) { break; }

    // ---- 2. stage A over the patches ----
    if constexpr (BRUTE) {
      __shared__ float4 tile[BZ_TILE * BZ_CULL_F4];
      for(uint32_t base = 0u; base < T.patch_count; base += BZ_TILE) {
        uint32_t n = min((uint32_t)BZ_TILE, T.patch_count - base);
        __syncthreads();
        if(t < n) {
          const float4 *p = T.cull + (size_t)(base + t) * BZ_CULL_F4;
          tile[t * 4u] = __ldg(p); tile[t * 4u + 1u] = __ldg(p + 1); tile[t * 4u + 2u] = __ldg(p + 2); tile[t * 4u + 3u] = __ldg(p + 3);
        }
        __syncthreads();
        if(live) {
          for(uint32_t j = 0u; j < n; ++j) {
            bz_cull_rec c;
            bz_unpack_cull(tile[j * 4u], tile[j * 4u + 1u], tile[j * 4u + 2u], tile[j * 4u + 3u], c);
            float cosine, distance;
            if(bz_stage_a(c, s, d, true, cosine, distance)) {
              bz_push_candidate(sm, T, t, base + j, s, d, ikey, ihit, ipatch, c_a, c_b);
            }
          }
          c_a += n;
        }
      }
    }
    __syncthreads();

    // ---- 3. stage B: dense passes over the survivor queue, then over the follow-hop queue ----
    const uint32_t na = min(sm.na, (uint32_t)BZ_QA);
    for(uint32_t base = 0u; base < na; base += BZ_NT) {
      uint32_t q = base + t, slot = 0u, evaluated = 0u;
      unsigned long long key = BZ_KEY_NONE;
      bz_hit hit;
      if(q < na) {
        uint2 entry = sm.qa[q];
        slot = entry.x;
        evaluated = entry.y;
        float rx = sm.ray[0][slot], ry = sm.ray[1][slot], rz = sm.ray[2][slot];
        float ux = sm.ray[3][slot], uy = sm.ray[4][slot], uz = sm.ray[5][slot];
        uint32_t nbp = 0u;
        c_b += bz_eval_pair(T.cull, T.newton, rx, ry, rz, ux, uy, uz, entry.y, 1u, &hit, &nbp);
        if(hit.what <= BZ_FOLLOW2) {
          uint32_t at = atomicAdd(&sm.nb, 1u);
          if(at < BZ_QB) {
            sm.qb_slot[at] = slot; sm.qb_patch[at] = nbp; sm.qb_loop[at] = entry.y;
          }
          else {
            uint32_t unused = 0u;
            evaluated = nbp;
            c_a += 1u;
            c_b += bz_eval_pair(T.cull, T.newton, rx, ry, rz, ux, uy, uz, nbp, 0u, &hit, &unused);
          }
        }
        if(bz_selectable(hit)) {
          key = bz_make_key(hit.distance, entry.y);
          atomicMin(&sm.best[slot], key);
        }
      }
      __syncthreads();
      if(key != BZ_KEY_NONE && sm.best[slot] == key) { bz_store_payload(sm, slot, hit, evaluated); }
    }
    __syncthreads();
    const uint32_t nb = min(sm.nb, (uint32_t)BZ_QB);
    for(uint32_t base = 0u; base < nb; base += BZ_NT) {
      uint32_t q = base + t, slot = 0u, evaluated = 0u;
      unsigned long long key = BZ_KEY_NONE;
      bz_hit hit;
      if(q < nb) {
        slot = sm.qb_slot[q];
        evaluated = sm.qb_patch[q];
        uint32_t unused = 0u;
        c_a += 1u;
        c_b += bz_eval_pair(T.cull, T.newton, sm.ray[0][slot], sm.ray[1][slot], sm.ray[2][slot],
                            sm.ray[3][slot], sm.ray[4][slot], sm.ray[5][slot], evaluated, 0u, &hit, &unused);
        if(bz_selectable(hit)) {
          key = bz_make_key(hit.distance, sm.qb_loop[q]);
          atomicMin(&sm.best[slot], key);
        }
      }
      __syncthreads();
      if(key != BZ_KEY_NONE && sm.best[slot] == key) { bz_store_payload(sm, slot, hit, evaluated); }
    }
    __syncthreads();

    // ---- 4. per slot: nearest hit -> Snell -> expected-side check -> continue or finish ----
    if(live) {
      unsigned long long key = sm.best[t];
      bz_hit hit;
      uint32_t patch;
      if(ikey < key) {
        key = ikey; hit = ihit; patch = ipatch;
      }
      else {
        hit.point = bz_v(sm.pay[0][t], sm.pay[1][t], sm.pay[2][t]);
        hit.normal = bz_v(sm.pay[3][t], sm.pay[4][t], sm.pay[5][t]);
        hit.cos_incidence = sm.pay[6][t];
        hit.bary = bz_v(sm.pay[7][t], sm.pay[8][t], sm.pay[9][t]);
        patch = __float_as_uint(sm.pay[10][t]);
      }
      const bool found = (key != BZ_KEY_NONE);
      ++c_calls;
      if(JOB == BZ_JOB_INTERSECT) {
        J.out_status[ray_id] = found ? BZ_WHAT_INTERSECT : BZ_WHAT_NONE;
        if(J.out_patch != nullptr) { J.out_patch[ray_id] = found ? patch : BZ_NO_PATCH; }
        if(J.out_distance != nullptr) { J.out_distance[ray_id] = found ? bz_key_distance(key) : 0.0f; }
        if(J.out_cos != nullptr) { J.out_cos[ray_id] = found ? hit.cos_incidence : 0.0f; }
        if(J.out_point != nullptr) {
          J.out_point[ray_id * 3u] = found ? hit.point.x : 0.0f; J.out_point[ray_id * 3u + 1u] = found ? hit.point.y : 0.0f; J.out_point[ray_id * 3u + 2u] = found ? hit.point.z : 0.0f;
        }
        if(J.out_normal != nullptr) {
          J.out_normal[ray_id * 3u] = found ? hit.normal.x : 0.0f; J.out_normal[ray_id * 3u + 1u] = found ? hit.normal.y : 0.0f; J.out_normal[ray_id * 3u + 2u] = found ? hit.normal.z : 0.0f;
        }
        if(J.out_bary != nullptr) {
          J.out_bary[ray_id * 3u] = found ? hit.bary.x : 0.0f; J.out_bary[ray_id * 3u + 1u] = found ? hit.bary.y : 0.0f; J.out_bary[ray_id * 3u + 2u] = found ? hit.bary.z : 0.0f;
        }
        live = false;
      }
      else {
        uint32_t status = BZ_STATUS_NONE;
        bzv3 refracted = d;
        if(found) {
          status = bz_snell(T.ri, hit.cos_incidence, hit.normal, refracted);
          ++c_snell;
        }
        status = (status == expected) ? status : BZ_STATUS_NONE;
        if(status != BZ_STATUS_NONE) {
          s = hit.point;
          d = refracted;
          if(status == BZ_STATUS_INSIDE) { ++c_enter; } else { ++c_exit; }
        }
        if(JOB == BZ_JOB_REFRACT) {
          J.out_start[ray_id * 3u] = s.x; J.out_start[ray_id * 3u + 1u] = s.y; J.out_start[ray_id * 3u + 2u] = s.z;
          J.out_dir[ray_id * 3u] = d.x; J.out_dir[ray_id * 3u + 1u] = d.y; J.out_dir[ray_id * 3u + 2u] = d.z;
          J.out_status[ray_id] = status;
          if(J.out_patch != nullptr) { J.out_patch[ray_id] = found ? patch : BZ_NO_PATCH; }
          live = false;
        }
        else {
          bool done = (status == BZ_STATUS_NONE);
          if(!done) {
            expected = (expected == BZ_STATUS_INSIDE) ? BZ_STATUS_OUTSIDE : BZ_STATUS_INSIDE;
            ++crossings;
            done = (crossings >= J.scene.max_crossings);
          }
          if(done) {
            uint32_t value;
            if(expected == BZ_STATUS_INSIDE) {
              value = bz_shade_background(J.scene, s, d);
            }
            else {
              value = J.scene.value_lost;
              ++c_lost;
            }
            atomicAdd(J.accum + pixel, value);
            live = false;
          }
        }
      }
    }
  }

  // ---- work counters: warp reduce, one atomic per warp and counter ----
  if(J.stats != nullptr) {
    unsigned long long v[BZ_NSTAT] = { c_primary, c_calls, c_a, (unsigned long long)c_b, (unsigned long long)c_snell,
                                       (unsigned long long)c_enter, (unsigned long long)c_exit, (unsigned long long)c_lost, c_nodes };
#pragma unroll
    for(int i = 0; i < BZ_NSTAT; ++i) {
      unsigned long long x = v[i];
      for(int o = 16; o > 0; o >>= 1) { x += __shfl_down_sync(0xffffffffu, x, o); }
      if(lane == 0u && x != 0ull) { atomicAdd(J.stats + i, x); }
    }
  }
}

// K3, second half: samples -> pixel, integer resolve (sum + spp/2) / spp, R = G = B = value, A = 255.
__global__ void bz_resolve_kernel(const uint32_t *accum, uchar4 *rgba, uint64_t pixels, uint32_t spp) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if(i < pixels) {
    uint32_t v = (accum[i] + spp / 2u) / spp;
    rgba[i] = make_uchar4((unsigned char)v, (unsigned char)v, (unsigned char)v, 255u);
  }
}

// Multi-GPU image assembly: `bands` holds, per rank, that rank's stripes in image order (band_rows rows reserved per
// rank, as an all-gather of equally sized buffers leaves them); every image row is copied to its place.  One uint
// (one RGBA pixel) per thread, rows are contiguous so the copies coalesce.
__global__ void bz_assemble_kernel(const uint32_t *bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                                   uint32_t ranks, uint32_t *image) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if(i < (uint64_t)width * height) {
    uint32_t row = (uint32_t)(i / width), px = (uint32_t)(i - (uint64_t)row * width);
    uint32_t stripe = row / stripe_rows;
    uint32_t rank = stripe % ranks;
    uint32_t local = (stripe / ranks) * stripe_rows + row % stripe_rows;
    image[i] = bands[((uint64_t)rank * band_rows + local) * width + px];
  }
}

// One rank's stripe band -> its rows of the image (bz_gather_rows, peer mode: `image` is the ROOT's buffer, mapped into this process, and
// the stores travel over NVLink).  One uint4 (four RGBA pixels) per thread when the rows allow it, rows are contiguous so the stores coalesce.
template<typename WORD>
__global__ void bz_scatter_band_kernel(const WORD *band, uint32_t rows, uint32_t row_words, uint32_t stripe_rows, uint32_t rank, uint32_t ranks, WORD *image) {
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if(i < (uint64_t)rows * row_words) {
    uint32_t local = (uint32_t)(i / row_words), at = (uint32_t)(i - (uint64_t)local * row_words);
    uint32_t row = ((local / stripe_rows) * ranks + rank) * stripe_rows + local % stripe_rows;
    image[(uint64_t)row * row_words + at] = band[i];
  }
}

// ---- flags of the peer gather (csrc/bz_comm.cu): generation counters in the ROOT's buffer, read and written across NVLink ----
__device__ __forceinline__ uint32_t bz_load_acquire_sys(const uint32_t *word) {
  uint32_t value;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(value) : "l"(word) : "memory");
  return value;
}

__device__ __forceinline__ void bz_store_release_sys(uint32_t *word, uint32_t value) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(word), "r"(value) : "memory");
}

// waits until *word >= value (counters only grow); a rank that never answers ends the wait after two minutes with a trap, which the
// next CUDA call of this process reports -- a lost rank must not leave the GPU spinning for ever
__device__ __forceinline__ void bz_wait_word(const uint32_t *word, uint32_t value) {
  unsigned long long start = 0ull;
  for(uint32_t spins = 0u; (int32_t)(bz_load_acquire_sys(word) - value) < 0; ++spins) {
    if((spins & 1023u) == 1023u) {
      unsigned long long now;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      if(start == 0ull) { start = now; }
      else if(now - start > 120000000000ull) { __trap(); }
    }
    __nanosleep(200);
  }
}

// The scatter of a non-root rank: waits until the root has taken the previous frame out of its buffer (free_word >= need_free), stores the
// band's rows into their place, and the CTA that finishes last publishes "rows of call `done_value` are there" in the root's buffer.
template<typename WORD>
__global__ void bz_scatter_band_flagged_kernel(const WORD *band, uint32_t rows, uint32_t row_words, uint32_t stripe_rows, uint32_t rank, uint32_t ranks, WORD *image,
                                               const uint32_t *free_word, uint32_t need_free, uint32_t *done_word, uint32_t done_value, uint32_t *finished) {
  if(threadIdx.x == 0u) { bz_wait_word(free_word, need_free); }
  __syncthreads();
  uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if(i < (uint64_t)rows * row_words) {
    uint32_t local = (uint32_t)(i / row_words), at = (uint32_t)(i - (uint64_t)local * row_words);
    uint32_t row = ((local / stripe_rows) * ranks + rank) * stripe_rows + local % stripe_rows;
    image[(uint64_t)row * row_words + at] = band[i];
  }
  __threadfence_system();                   // this thread's stores are visible on the root before anything that follows
  __syncthreads();
  if(threadIdx.x == 0u) {
    if(atomicAdd(finished, 1u) == gridDim.x - 1u) {
      *finished = 0u;                       // (for the next call; calls on one communicator do not overlap)
      __threadfence_system();
      bz_store_release_sys(done_word, done_value);
    }
  }
}

// root: one thread per rank waits for that rank's "rows of call `value` are there"
__global__ void bz_wait_ranks_kernel(const uint32_t *done_words, uint32_t ranks, uint32_t root, uint32_t value) {
  if(threadIdx.x < ranks && threadIdx.x != root) { bz_wait_word(done_words + threadIdx.x, value); }
}

__global__ void bz_publish_word_kernel(uint32_t *word, uint32_t value) {
  __threadfence_system();
  bz_store_release_sys(word, value);
}

// K1 alone: primary rays of a row band into n x 3 float arrays (bz_generate_rays).
__global__ void bz_raygen_kernel(const bz_camera cam, uint32_t row0, uint64_t count, float *start, float *dir) {
  uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if(id < count) {
    uint32_t spp = cam.strata * cam.strata;
    uint32_t pixel = (uint32_t)(id / spp);
    uint32_t sample = (uint32_t)(id - (uint64_t)pixel * spp);
    uint32_t row = pixel / cam.width;
    bzv3 s, d;
    bz_primary_ray(cam, pixel - row * cam.width, row0 + row, sample % cam.strata, sample / cam.strata, s, d);
    start[id * 3u] = s.x; start[id * 3u + 1u] = s.y; start[id * 3u + 2u] = s.z;
    dir[id * 3u] = d.x; dir[id * 3u + 1u] = d.y; dir[id * 3u + 2u] = d.z;
  }
}

// FP32 issue-rate probe: BZ_PROBE_CHAINS independent FMA chains per thread (enough to cover the 4-cycle latency of the pipe
// at any occupancy), `iters` dependent steps each, the loop unrolled 16 x (256 FMAs per trip) so that loop control is ~1 % of the issue slots.
#define BZ_PROBE_CHAINS 16
__global__ void __launch_bounds__(256) bz_fma_probe_kernel(float *sink, uint32_t iters) {
  float a[BZ_PROBE_CHAINS];
#pragma unroll
  for(int k = 0; k < BZ_PROBE_CHAINS; ++k) { a[k] = threadIdx.x * 1e-3f + (float)k; }
  const float m = 0.999f, c = 1e-4f;
#pragma unroll 16
  for(uint32_t i = 0u; i < iters; ++i) {
#pragma unroll
    for(int k = 0; k < BZ_PROBE_CHAINS; ++k) { a[k] = __fmaf_rn(a[k], m, c); }
  }
  float r = 0.0f;
#pragma unroll
  for(int k = 0; k < BZ_PROBE_CHAINS; ++k) { r += a[k]; }
  if(r == 123.456f) { sink[0] = r; }
}

extern "C++" cudaError_t bz_launch_trace(int job, bool brute, const bz_tables &tables, const bz_job &desc, int grid, cudaStream_t stream) {
#define BZ_LAUNCH(JOBV) \
  if(brute) { bz_trace_kernel<JOBV, true><<<grid, BZ_NT, 0, stream>>>(tables, desc); } \
  else { return cudaErrorInvalidValue; }     /* BZ_MODE_CULLED runs bz_warp_trace_kernel (bz_trace.cu) */
  if(job == BZ_JOB_RENDER) { BZ_LAUNCH(BZ_JOB_RENDER) }
  else if(job == BZ_JOB_REFRACT) { BZ_LAUNCH(BZ_JOB_REFRACT) }
  else { BZ_LAUNCH(BZ_JOB_INTERSECT) }
#undef BZ_LAUNCH
  return cudaGetLastError();
}

cudaError_t bz_launch_resolve(const uint32_t *accum, uint8_t *rgba, uint64_t pixels, uint32_t spp, cudaStream_t stream) {
  if(pixels == 0u) { return cudaSuccess; }
  bz_resolve_kernel<<<(unsigned)((pixels + 255u) / 256u), 256, 0, stream>>>(accum, (uchar4*)rgba, pixels, spp);
  return cudaGetLastError();
}

cudaError_t bz_launch_raygen(const bz_camera &cam, uint32_t row0, uint64_t count, float *start, float *dir, cudaStream_t stream) {
  if(count == 0u) { return cudaSuccess; }
  bz_raygen_kernel<<<(unsigned)((count + 255u) / 256u), 256, 0, stream>>>(cam, row0, count, start, dir);
  return cudaGetLastError();
}

cudaError_t bz_launch_assemble(const uint8_t *bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                               uint32_t ranks, uint8_t *image, cudaStream_t stream) {
  uint64_t const pixels = (uint64_t)width * height;
  if(pixels == 0u) { return cudaSuccess; }
  bz_assemble_kernel<<<(unsigned)((pixels + 255u) / 256u), 256, 0, stream>>>((const uint32_t*)bands, band_rows, width, height, stripe_rows, ranks, (uint32_t*)image);
  return cudaGetLastError();
}

cudaError_t bz_launch_scatter_band(const uint8_t *band, uint32_t rows, uint32_t width, uint32_t stripe_rows, uint32_t rank, uint32_t ranks,
                                   uint8_t *image, cudaStream_t stream) {
  if((uint64_t)rows * width == 0u) { return cudaSuccess; }
  if(width % 4u == 0u && ((uintptr_t)band | (uintptr_t)image) % 16u == 0u) {
    uint64_t const words = (uint64_t)rows * (width / 4u);
    bz_scatter_band_kernel<uint4><<<(unsigned)((words + 255u) / 256u), 256, 0, stream>>>((const uint4*)band, rows, width / 4u, stripe_rows, rank, ranks, (uint4*)image);
  }
  else {
    uint64_t const words = (uint64_t)rows * width;
    bz_scatter_band_kernel<uint32_t><<<(unsigned)((words + 255u) / 256u), 256, 0, stream>>>((const uint32_t*)band, rows, width, stripe_rows, rank, ranks, (uint32_t*)image);
  }
  return cudaGetLastError();
}

cudaError_t bz_launch_scatter_band_flagged(const uint8_t *band, uint32_t rows, uint32_t width, uint32_t stripe_rows, uint32_t rank, uint32_t ranks, uint8_t *image,
                                           const uint32_t *free_word, uint32_t need_free, uint32_t *done_word, uint32_t done_value, uint32_t *finished, cudaStream_t stream) {
  if(width % 4u == 0u && ((uintptr_t)band | (uintptr_t)image) % 16u == 0u) {
    uint64_t const words = (uint64_t)rows * (width / 4u);
    bz_scatter_band_flagged_kernel<uint4><<<(unsigned)std::max<uint64_t>(1u, (words + 255u) / 256u), 256, 0, stream>>>((const uint4*)band, rows, width / 4u, stripe_rows, rank, ranks,
                                                                                                         (uint4*)image, free_word, need_free, done_word, done_value, finished);
  }
  else {
    uint64_t const words = (uint64_t)rows * width;
    bz_scatter_band_flagged_kernel<uint32_t><<<(unsigned)std::max<uint64_t>(1u, (words + 255u) / 256u), 256, 0, stream>>>((const uint32_t*)band, rows, width, stripe_rows, rank, ranks,
                                                                                                            (uint32_t*)image, free_word, need_free, done_word, done_value, finished);
  }
  return cudaGetLastError();
}

cudaError_t bz_launch_wait_ranks(const uint32_t *done_words, uint32_t ranks, uint32_t root, uint32_t value, cudaStream_t stream) {
  bz_wait_ranks_kernel<<<1, 1024, 0, stream>>>(done_words, ranks, root, value);
  return cudaGetLastError();
}

cudaError_t bz_launch_publish_word(uint32_t *word, uint32_t value, cudaStream_t stream) {
  bz_publish_word_kernel<<<1, 1, 0, stream>>>(word, value);
  return cudaGetLastError();
}

cudaError_t bz_launch_fma_probe(float *sink, uint32_t iters, int grid, cudaStream_t stream) {
  bz_fma_probe_kernel<<<grid, 256, 0, stream>>>(sink, iters);
  return cudaGetLastError();
}

int bz_trace_threads(void) { return BZ_NT; }

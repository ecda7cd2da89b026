// bz_wave.cu -- the wavefront form of the RENDER job: three small kernels per surface crossing instead of one persistent
// kernel that does everything (csrc/bz_trace.cu).
//
// Why: the persistent kernel runs walk, stage A, stage B and the resolve step interleaved on the 16 warps of an SM.  Its
// 58 KB of code do not fit the instruction caches (ncu: `no_instruction` 0.65 .. 1.33 stall cycles per issued instruction,
// depending on nothing but code layout), stage B sees 29 of 32 lanes, and 128 registers leave 4 warps per scheduler.  Here
// every kernel is one phase:
//   bz_wave_walk_kernel     packets of 32 rays: K1 ray generation (first crossing) or reload, the warp-cooperative packet
//                           walk of the cull hierarchy, stage A (bezierTriangle.cpp:124-131) on record pairs; the survivors
//                           go to a global candidate list, chained per ray.  Per-lane walk for incoherent packets, as in
//                           the persistent kernel.
//   bz_wave_eval_kernel     one candidate per lane, all 32 lanes busy: stage B (bezierTriangle.cpp:132-194), the follow-side
//                           hop (bezierMesh.cpp:213-217) through a small per-warp queue; key + hit record per candidate.
//                           Its code is the root search and nothing else.
//   bz_wave_resolve_kernel  one ray per thread: the nearest result of its chain (strict <, first patch wins:
//                           bezierMesh.cpp:220), Snell (bezierLens.cpp:8-33), expected-side check, next ray or K3 shading.
// The arithmetic is the persistent kernel's, function for function (csrc/bz_math.cuh, csrc/bz_trace_common.cuh): the two
// pipelines give the same bits, and tests/ hold both to the oracle.  The lists live in HBM, which this path leaves idle
// otherwise (19 GB of traffic per 4K frame = 1.25 TB/s averaged; the persistent kernel: 65 MB).
// The host side (csrc/bz_capi.cu: bz_run_wave) cuts the lens tiles of a frame into at least two chunks that interleave over the
// weight-sorted tile order and runs them alternately on two buffer sets / two internal streams, so that one chunk's kernels fill
// the drained boundaries of the other's; the walk kernel hands the heaviest tiles of a chunk out in parts (bzv_take_packet).
// Measured on the 4K frame: 15.3 ms against the persistent kernel's 17.6.
// Compiled twice, like csrc/bz_trace.cu: parity build (-fmad=false) and -DBZ_FAST.
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include <algorithm>
#include "bz_device.h"
#if defined(BZ_FAST)
#define BZ_NS bz_fast_wave
#define BZ_SYM(name) name##_fast
#else
#define BZ_NS bz_exact_wave
#define BZ_SYM(name) name
#endif
namespace BZ_NS {
#include "bz_math.cuh"
#include "bz_trace_common.cuh"

#define BZV_WALK_WARPS   8
#ifndef BZV_WALK_CTAS
#define BZV_WALK_CTAS    4       // resident CTAs per SM the walk kernel is compiled for (64 registers, 16 B of spills; 3 CTAs at 72 registers measured 0.7 % slower)
#endif
#ifndef BZV_EVAL_WARPS
#define BZV_EVAL_WARPS   4
#endif
#ifndef BZV_EVAL_CTAS
#define BZV_EVAL_CTAS    4
#endif
#define BZV_WALK_BURST   2
#define BZV_QA_CAP       256
#define BZV_QB_CAP       160
#define BZV_CAND_CAP     160
#define BZV_STACK_CAP    96
#define BZV_HOP_CAP      64
#define BZV_ROOT bz_stage_b_root_unfolded      // the evaluation kernel is nothing but stage B: the written-out form (17.81 -> 17.33 ms per 4K frame)

struct bzv_walk_pool {                     // one per warp, shared memory
  float              ray[6][32];           // the packet's rays by lane (the per-lane walk reads them by slot)
  uint32_t           head[32];             // 1 + newest candidate of the ray, 0 = none
  uint32_t           qa_leaf[BZV_QA_CAP];  // per-lane walk: (slot, leaf record) pairs waiting for stage A
  uint32_t           qb_patch[BZV_QB_CAP]; // stage-A survivors waiting for the flush to the global list
  uint8_t            qa_slot[BZV_QA_CAP];
  uint8_t            qb_slot[BZV_QB_CAP];
  uint32_t           cand[BZV_CAND_CAP];   // packet walk: leaf codes the packet may meet
  uint32_t           stack[BZV_STACK_CAP]; // packet walk: 32-wide nodes still to visit
  uint32_t           qa_n, qb_n;
  unsigned long long count[BZ_NSTAT];
};

static_assert(BZV_QA_CAP >= 32 * BZ_LEAF_MAX * BZV_WALK_BURST + 64 && BZV_QB_CAP >= 96, "queue thresholds");

#define BZV_COUNT(index, amount) do { const unsigned long long amount_ = (unsigned long long)(amount); if(lane == 0u) { P.count[index] += amount_; } } while(0)

// the t-th tile of the job's hand-out order (csrc/bz_trace.cu: bzw_fetch_tile)
__device__ __forceinline__ uint32_t bzv_tile_of(const bz_job &J, uint32_t t) {
  if(J.tile_list != nullptr) { return __ldg(J.tile_list + t); }
  const uint32_t turn = t / J.tiles_x, tile_x = t - turn * J.tiles_x;
  const uint32_t tile_y = ((turn & 1u) == 0u) ? (turn >> 1) : (J.tiles_y - 1u - (turn >> 1));
  return tile_y * J.tiles_x + tile_x;
}

// stage-A survivors of the warp -> the global candidate list; every entry is pushed onto the chain of its ray
__device__ __forceinline__ void bzv_flush(bzv_walk_pool &P, const bz_wave &W, uint32_t slot_base, uint32_t lane) {
  __syncwarp();
  const uint32_t n = P.qb_n;
  if(n == 0u) { return; }
  uint32_t base = 0u;
  if(lane == 0u) { base = atomicAdd(W.counters + 1, n); }
  base = __shfl_sync(0xffffffffu, base, 0);
  if(base + n > W.cand_cap || base + n < base) {
    if(lane == 0u) { atomicOr(W.counters + 3, 1u); }          // sticky: reported by the next call on the context
  }
  else {
    for(uint32_t i = lane; i < n; i += 32u) {
      const uint32_t slot = P.qb_slot[i], index = base + i;
      __stcs(W.cand + index, make_uint2(slot_base + slot, P.qb_patch[i]));      // (the lists are written once and read once: streaming)
      __stcs(W.cand_next + index, atomicExch(&P.head[slot], index + 1u));
    }
  }
  __syncwarp();
  if(lane == 0u) { P.qb_n = 0u; }
  __syncwarp();
}

// Work distribution of the walk kernel: a CTA takes whole tiles of the chunk from a global counter, its warps take the packets
// of the tile from a shared-memory word (tile << 32 | packets handed out) -- the warps of an SM work in one image
// neighbourhood and meet the same hierarchy nodes and patch records in L1 (as in csrc/bz_trace.cu).  The warp whose packet
// is the first one past the end of the tile fetches the next tile and takes its first packet; the others wait for it.
__device__ __forceinline__ unsigned long long bzv_fetch_tile(const bz_wave &W) {
  const uint32_t next = atomicAdd(W.counters, 1u);
  return (unsigned long long)(next < W.unit_count ? next : BZ_END) << 32;
}

// The unit of the hand-out: the first W.split_tiles tiles of the chunk (the heaviest: the fan of thin patches around a pole of the body
// costs a hundred times an ordinary tile) go out in parts of W.unit_packets consecutive packets, so that several SMs share them instead
// of one CTA being the tail of the launch; the others go out whole.  Units [0, split_tiles * units_per_tile) are the parts.
__device__ __forceinline__ uint32_t bzv_unit_packets(const bz_wave &W, uint32_t unit) {
  if(unit >= W.split_tiles * W.units_per_tile) { return W.packets_per_tile; }
  return min(W.unit_packets, W.packets_per_tile - (unit % W.units_per_tile) * W.unit_packets);
}

__device__ __forceinline__ unsigned long long bzv_take_packet(unsigned long long *cta_work, const bz_wave &W) {
  for(;;) {
    const unsigned long long w = atomicAdd(cta_work, 1ull);
    const uint32_t unit = (uint32_t)(w >> 32), at = (uint32_t)w;
    if(unit == BZ_END) { return w; }
    const uint32_t limit = bzv_unit_packets(W, unit);
    if(at < limit) { return w; }
    if(at == limit) {
      const unsigned long long fresh = bzv_fetch_tile(W);
      atomicExch(cta_work, fresh | 1ull);
      return fresh;
    }
    while((uint32_t)(*(volatile unsigned long long*)cta_work >> 32) == unit) { __nanosleep(64); }
  }
}

// ------------------------------------------------------------------ walk + stage A ----
__global__ void __launch_bounds__(32 * BZV_WALK_WARPS, BZV_WALK_CTAS) bz_wave_walk_kernel(const bz_tables T, const bz_job J, const bz_wave W) {
  extern __shared__ __align__(16) unsigned char bzv_shared[];
  bzv_walk_pool &P = reinterpret_cast<bzv_walk_pool*>(bzv_shared)[threadIdx.x >> 5];
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t spp = J.camera.strata * J.camera.strata;
  const uint32_t direct_end = T.always_first + T.always_count;
  __shared__ unsigned long long cta_work;
  const uint32_t below = (1u << lane) - 1u;
  const float one = J.one;
  if(lane == 0u) { P.qa_n = 0u; P.qb_n = 0u; }
  if(lane < BZ_NSTAT) { P.count[lane] = 0ull; }
  if(threadIdx.x == 0u) { cta_work = bzv_fetch_tile(W); }
  __syncthreads();                           // the only CTA barrier of the kernel

  for(;;) {
    unsigned long long work = 0ull;
    if(lane == 0u) { work = bzv_take_packet(&cta_work, W); }
    work = __shfl_sync(0xffffffffu, work, 0);
    const uint32_t unit = (uint32_t)(work >> 32);
    if(unit == BZ_END) { break; }
    uint32_t in_chunk, in_tile = (uint32_t)work;      // tile of the chunk, packet of the tile
    if(unit >= W.split_tiles * W.units_per_tile) { in_chunk = unit - W.split_tiles * (W.units_per_tile - 1u); }
    else {
      in_chunk = unit / W.units_per_tile;
      in_tile += (unit - in_chunk * W.units_per_tile) * W.unit_packets;
    }
    const uint32_t packet = in_chunk * W.packets_per_tile + in_tile;
    const uint32_t slot_base = packet * 32u, ray_slot = slot_base + lane;
    // ---- the packet's rays: generated (first crossing) or reloaded ----
    bzv3 s = bz_v(0.0f, 0.0f, 0.0f), d = s;
    bool live = false;
    if(W.level == 0u) {
      const uint32_t tile = bzv_tile_of(J, W.first_tile + in_chunk * W.tile_stride);
      const uint32_t at = in_tile * 32u + lane;
      const uint32_t inside = bzw_div(at, J.div_spp), sample = at - inside * spp;      // pixel of the tile in Morton order, sample of the pixel
      const uint32_t tile_y = tile / J.tiles_x, tile_x = tile - tile_y * J.tiles_x;
      const uint32_t px = tile_x * J.tile_side + bzw_even_bits(inside), row = tile_y * J.tile_side + bzw_even_bits(inside >> 1);
      live = at < J.tile_rays && px < J.camera.width && row < J.rows;
      uint32_t pixel = 0u;
      if(live) {
        const uint32_t sy = bzw_div(sample, J.div_strata);
        bzw_primary_ray(J.camera, J.two_over_w, J.two_over_h, px, bzw_frame_row(J, row), sample - sy * J.camera.strata, sy, s, d);
        pixel = row * J.camera.width + px;
      }
      __stcs(W.ray_a + ray_slot, make_float4(s.x, s.y, s.z, d.x));
      __stcs(W.ray_b + ray_slot, make_float4(d.y, d.z, __uint_as_float(pixel), __uint_as_float(live ? (1u | (BZ_STATUS_INSIDE << 1)) : 0u)));
    }
    else {
      const float4 b = W.ray_b[ray_slot];
      live = (__float_as_uint(b.w) & 1u) != 0u;
      if(__ballot_sync(0xffffffffu, live) == 0u) { continue; }
      const float4 a = W.ray_a[ray_slot];
      s = bz_v(a.x, a.y, a.z);
      d = bz_v(a.w, b.x, b.y);
    }
    const uint32_t live_mask = __ballot_sync(0xffffffffu, live);
    if(W.level == 0u) { BZV_COUNT(BZC_PRIMARY, __popc(live_mask)); }
    BZV_COUNT(BZC_CALLS, __popc(live_mask));
    if(live_mask == 0u) { continue; }
    P.head[lane] = 0u;
    P.ray[0][lane] = s.x; P.ray[1][lane] = s.y; P.ray[2][lane] = s.z;
    P.ray[3][lane] = d.x; P.ray[4][lane] = d.y; P.ray[5][lane] = d.z;
    __syncwarp();
    const float l1 = fabsf(s.x) + fabsf(s.y) + fabsf(s.z);

    // ---- packet walk: one lane per child box of the 32-wide hierarchy (as csrc/bz_trace.cu) ----
    bool packet_done = false;
    uint32_t found = 0u;
    if(T.packet_count != 0u) {
      const float inf = __int_as_float(0x7f800000);
      const float l1_max = bzw_warp_max(live ? l1 : 0.0f);
      const float margin = 2.0f * 3.0517578125e-05f * (l1_max + 2.0f * T.extent);
      const bzw_axis ax = bzw_make_axis(bzw_warp_min(live ? s.x : inf), bzw_warp_max(live ? s.x : -inf), bzw_warp_min(live ? d.x : inf), bzw_warp_max(live ? d.x : -inf), margin);
      const bzw_axis ay = bzw_make_axis(bzw_warp_min(live ? s.y : inf), bzw_warp_max(live ? s.y : -inf), bzw_warp_min(live ? d.y : inf), bzw_warp_max(live ? d.y : -inf), margin);
      const bzw_axis az = bzw_make_axis(bzw_warp_min(live ? s.z : inf), bzw_warp_max(live ? s.z : -inf), bzw_warp_min(live ? d.z : inf), bzw_warp_max(live ? d.z : -inf), margin);
      const uint32_t direct_first = (l1_max <= T.origin_limit) ? T.always_first : T.wide_first;
      const uint32_t direct_n = direct_end - direct_first;
      uint32_t sp = 1u;
      found = (direct_n + 6u) / 7u;
      packet_done = (found <= 32u);      // one seed entry per lane; more records than that outside the hierarchy: per-lane walk
      if(packet_done) {
        if(lane < found) { P.cand[lane] = ((direct_first + 7u * lane) << 3) | min(7u, direct_n - 7u * lane); }
        if(lane == 0u) { P.stack[0] = 0u; }
        __syncwarp();
      }
      uint32_t visits = 0u;
      while(packet_done && sp != 0u) {
        --sp;
        const uint32_t node = P.stack[sp] * 224u;
        const uint32_t code = __float_as_uint(__ldg(T.packet + lane + 192u + node));
        float tnear = 0.0f, tfar = inf;
        bzw_axis_bounds(ax, T.packet + lane, node, tnear, tfar);
        bzw_axis_bounds(ay, T.packet + lane + 32u, node, tnear, tfar);
        bzw_axis_bounds(az, T.packet + lane + 64u, node, tnear, tfar);
        const bool hit = (code != 0u) && (tnear <= tfar);
        const bool inner = (code & 0x80000000u) != 0u;
        const uint32_t hit_inner = __ballot_sync(0xffffffffu, hit && inner);
        const uint32_t hit_leaf = __ballot_sync(0xffffffffu, hit && !inner);
        ++visits;
        if(sp + __popc(hit_inner) > BZV_STACK_CAP || found + __popc(hit_leaf) > BZV_CAND_CAP) { packet_done = false; break; }
        if(hit && inner) { P.stack[sp + __popc(hit_inner & below)] = code & 0x7fffffffu; }
        if(hit && !inner) { P.cand[found + __popc(hit_leaf & below)] = code; }
        sp += __popc(hit_inner);
        found += __popc(hit_leaf);
        __syncwarp();
      }
      BZV_COUNT(BZC_PNODES, visits);
      __syncwarp();
    }

    if(packet_done) {
      // ---- stage A, packet form: two records per trip on the packed FP32 pipe, every live ray of the packet tests them ----
      BZV_COUNT(BZC_PWALKS, 1u);
      uint32_t level = 0u, tested = 0u, cand_at = 0u, cand_sub = 0u;
      while(cand_at < found) {
        const uint32_t code = P.cand[cand_at];
        const uint32_t left = (code & 7u) - cand_sub;
        const float4 *rec = T.pairs + (size_t)((code >> 3) + cand_sub) * (2u * BZ_CULL_F4);
        const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3);
        const float4 r4 = __ldg(rec + 4), r5 = __ldg(rec + 5), r6 = __ldg(rec + 6), r7 = __ldg(rec + 7);
        cand_sub += min(left, 2u);
        if(left <= 2u) { cand_sub = 0u; ++cand_at; }
        tested += min(left, 2u);
        const float2 nx = make_float2(r0.x, r0.y), ny = make_float2(r0.z, r0.w), nz = make_float2(r1.x, r1.y), kk = make_float2(r1.z, r1.w);
        const float2 cosine = bzw_add2(bzw_mul2(nx, d.x), bzw_add2(bzw_mul2(ny, d.y), bzw_mul2(nz, d.z), one), one);
        const float2 along = bzw_add2(bzw_mul2(nx, s.x), bzw_add2(bzw_mul2(ny, s.y), bzw_mul2(nz, s.z), one), one);
#if defined(BZ_FAST)
        const float2 above = __fadd2_rn(kk, make_float2(-along.x, -along.y));
#else
        const float2 above = __ffma2_rn(along, make_float2(-one, -one), kk);
#endif
        bool pass_a = live && fabsf(cosine.x) >= BZ_PLANE_EPSILON;
        bool pass_b = live && left >= 2u && fabsf(cosine.y) >= BZ_PLANE_EPSILON;
        float2 distance = make_float2(0.0f, 0.0f);
        if(pass_a) { distance.x = BZ_DIV(above.x, cosine.x); }
        if(pass_b) { distance.y = BZ_DIV(above.y, cosine.y); }
        pass_a = pass_a && distance.x > 0.0f && fabsf(distance.x) > -r3.z && fabsf(distance.x) > r5.z;
        pass_b = pass_b && distance.y > 0.0f && fabsf(distance.y) > -r3.w && fabsf(distance.y) > r5.w;
        if(__ballot_sync(0xffffffffu, pass_a || pass_b) != 0u) {
          const float2 px = __ffma2_rn(bzw_mul2(distance, d.x), make_float2(one, one), make_float2(s.x, s.x));
          const float2 py = __ffma2_rn(bzw_mul2(distance, d.y), make_float2(one, one), make_float2(s.y, s.y));
          const float2 pz = __ffma2_rn(bzw_mul2(distance, d.z), make_float2(one, one), make_float2(s.z, s.z));
#define BZV_ROW(x, y, z) bzw_add2(__fmul2_rn(x, px), bzw_add2(__fmul2_rn(y, py), __fmul2_rn(z, pz), one), one)
          const float2 b0 = BZV_ROW(make_float2(r2.x, r2.y), make_float2(r2.z, r2.w), make_float2(r3.x, r3.y));
          const float2 b1 = BZV_ROW(make_float2(r4.x, r4.y), make_float2(r4.z, r4.w), make_float2(r5.x, r5.y));
          const float2 b2 = BZV_ROW(make_float2(r6.x, r6.y), make_float2(r6.z, r6.w), make_float2(r7.x, r7.y));
#undef BZV_ROW
          pass_a = pass_a && b0.x >= 0.0f && b0.x <= 1.0f && b1.x >= 0.0f && b1.x <= 1.0f && b2.x >= 0.0f && b2.x <= 1.0f;
          pass_b = pass_b && b0.y >= 0.0f && b0.y <= 1.0f && b1.y >= 0.0f && b1.y <= 1.0f && b2.y >= 0.0f && b2.y <= 1.0f;
          const uint32_t passed_a = __ballot_sync(0xffffffffu, pass_a), passed_b = __ballot_sync(0xffffffffu, pass_b);
          if((passed_a | passed_b) != 0u) {
            if(pass_a) {
              const uint32_t at = level + __popc(passed_a & below);
              P.qb_patch[at] = __float_as_uint(r7.z);
              P.qb_slot[at] = (uint8_t)lane;
            }
            level += __popc(passed_a);
            if(pass_b) {
              const uint32_t at = level + __popc(passed_b & below);
              P.qb_patch[at] = __float_as_uint(r7.w);
              P.qb_slot[at] = (uint8_t)lane;
            }
            level += __popc(passed_b);
            if(level > BZV_QB_CAP - 64u) {
              if(lane == 0u) { P.qb_n = level; }
              bzv_flush(P, W, slot_base, lane);
              level = 0u;
            }
          }
        }
      }
      if(lane == 0u) { P.qb_n = level; }
      BZV_COUNT(BZC_STAGE_A, tested * __popc(live_mask));
      bzv_flush(P, W, slot_base, lane);
    }
    else {
      // ---- per-lane walk (no packet hierarchy, or the packet is not coherent enough): skip-link hierarchy, (slot, record)
      //      pairs through queue A, scalar stage A 32 pairs at a time ----
      BZV_COUNT(BZC_LWALKS, __popc(live_mask));
      uint32_t direct = BZ_END, node = BZ_END;
      if(live) {
        direct = (l1 <= T.origin_limit) ? T.always_first : T.wide_first;
        if(direct >= direct_end) { direct = BZ_END; }
        node = (T.node_count != 0u) ? 0u : BZ_END;
      }
      const float slack = 3.0517578125e-05f * (l1 + 2.0f * T.extent);
      const float ix = (fabsf(d.x) > 1e-20f) ? 1.0f / d.x : copysignf(1e20f, d.x);
      const float iy = (fabsf(d.y) > 1e-20f) ? 1.0f / d.y : copysignf(1e20f, d.y);
      const float iz = (fabsf(d.z) > 1e-20f) ? 1.0f / d.z : copysignf(1e20f, d.z);
      const float ax_lo = -(s.x + slack) * ix, ax_hi = -(s.x - slack) * ix;
      const float ay_lo = -(s.y + slack) * iy, ay_hi = -(s.y - slack) * iy;
      const float az_lo = -(s.z + slack) * iz, az_hi = -(s.z - slack) * iz;
      uint32_t steps = 0u;
      bool walked = false;
      for(;;) {
        __syncwarp();
        uint32_t qa = P.qa_n;
        __syncwarp();
        // walk while queue A has room for a whole burst
        while(!walked && qa <= BZV_QA_CAP - 32u * BZ_LEAF_MAX * BZV_WALK_BURST) {
          if(__ballot_sync(0xffffffffu, direct != BZ_END) != 0u) {
            if(direct != BZ_END) {                       // one direct record per lane and turn
              const uint32_t at = atomicAdd(&P.qa_n, 1u);
              P.qa_leaf[at] = direct;
              P.qa_slot[at] = (uint8_t)lane;
              ++direct;
              if(direct >= direct_end) { direct = BZ_END; }
            }
          }
          else {
#pragma unroll
            for(int u = 0; u < BZV_WALK_BURST; ++u) {
              if(node != BZ_END) {
                const float4 lo = __ldg(T.nodes + (size_t)node * BZ_NODE_F4);
                const float4 hi = __ldg(T.nodes + (size_t)node * BZ_NODE_F4 + 1);
                ++steps;
                const float x0 = __fmaf_rn(lo.x, ix, ax_lo), x1 = __fmaf_rn(hi.x, ix, ax_hi);
                const float y0 = __fmaf_rn(lo.y, iy, ay_lo), y1 = __fmaf_rn(hi.y, iy, ay_hi);
                const float z0 = __fmaf_rn(lo.z, iz, az_lo), z1 = __fmaf_rn(hi.z, iz, az_hi);
                const float tnear = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
                const float tfar = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
                const bool hit = (tnear <= tfar * 1.000001f);
                const uint32_t code = __float_as_uint(hi.w);
                if(hit && code != 0u) {
                  const uint32_t count = code & 7u;
                  const uint32_t at = atomicAdd(&P.qa_n, count);
                  for(uint32_t j = 0u; j < count; ++j) {
                    P.qa_leaf[at + j] = (code >> 3) + j;
                    P.qa_slot[at + j] = (uint8_t)lane;
                  }
                }
                node = (hit && code == 0u) ? node + 1u : __float_as_uint(lo.w);
              }
            }
            if(__ballot_sync(0xffffffffu, node != BZ_END) == 0u) { walked = true; }
          }
          __syncwarp();
          qa = P.qa_n;
          __syncwarp();
        }
        if(qa == 0u) {
          if(walked) { break; }
          continue;
        }
        // scalar stage A on (up to) 32 pairs from the tail of queue A
        const uint32_t n = min(qa, 32u);
        const uint32_t q = qa - n + lane;
        if(lane == 0u) { P.qa_n = qa - n; }
        BZV_COUNT(BZC_STAGE_A, n);
        bool pass = false;
        uint32_t patch = 0u, slot = 0u;
        if(lane < n) {
          slot = P.qa_slot[q];
          bz_cull_rec c;
          bzw_unpack_cull(T.leaf + (size_t)P.qa_leaf[q] * BZ_CULL_F4, c, patch);
          const bzv3 rs = bz_v(P.ray[0][slot], P.ray[1][slot], P.ray[2][slot]);
          const bzv3 rd = bz_v(P.ray[3][slot], P.ray[4][slot], P.ray[5][slot]);
          float cosine, distance;
          pass = bz_stage_a(c, rs, rd, true, cosine, distance);
        }
        __syncwarp();
        const uint32_t passed = __ballot_sync(0xffffffffu, pass);
        if(passed != 0u) {
          const uint32_t level = P.qb_n;
          __syncwarp();
          if(pass) {
            const uint32_t at = level + __popc(passed & below);
            P.qb_patch[at] = patch;
            P.qb_slot[at] = (uint8_t)slot;
          }
          __syncwarp();
          if(lane == 0u) { P.qb_n = level + __popc(passed); }
          __syncwarp();
          if(level + __popc(passed) > BZV_QB_CAP - 32u) { bzv_flush(P, W, slot_base, lane); }
        }
      }
      steps = __reduce_add_sync(0xffffffffu, steps);
      BZV_COUNT(BZC_NODES, steps);
      bzv_flush(P, W, slot_base, lane);
    }
    if(live) { __stcs(W.head + ray_slot, P.head[lane]); }
    __syncwarp();
  }
  __syncwarp();
  if(J.stats != nullptr && lane < BZ_NSTAT && P.count[lane] != 0ull) { atomicAdd(J.stats + lane, P.count[lane]); }
}

// ------------------------------------------------------------------ stage B ----
#define BZV_STAGE_F4 15                    // float4 per candidate in the staging buffer: ray 2, stage-A record 4, control points + dirB 9

struct bzv_eval_pool {                     // one per warp, shared memory
  float4   stage[BZV_STAGE_F4][32];        // the inputs of the warp's NEXT batch, copied in by cp.async while the current one is evaluated
  uint32_t hop_index[BZV_HOP_CAP];         // follow-side hops waiting for their evaluation: the candidate the hop belongs to
  uint32_t hop_patch[BZV_HOP_CAP];         //                                                the neighbour patch to evaluate
  uint32_t hop_n;
};

__device__ __forceinline__ void bzv_copy16(float4 *to_shared, const float4 *from_global) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"((uint32_t)__cvta_generic_to_shared(to_shared)), "l"(from_global));
}

// ray, stage-A record and stage-B record (without the divider planes) of one candidate -> this lane's column of the staging buffer
__device__ __forceinline__ void bzv_stage(bzv_eval_pool &P, const bz_tables &T, const bz_wave &W, uint2 entry, uint32_t lane) {
  bzv_copy16(&P.stage[0][lane], W.ray_a + entry.x);
  bzv_copy16(&P.stage[1][lane], W.ray_b + entry.x);
  const float4 *cull = T.cull + (size_t)entry.y * BZ_CULL_F4, *newton = T.newton + (size_t)entry.y * BZ_NEWTON_F4;
#pragma unroll
  for(int i = 0; i < BZ_CULL_F4; ++i) { bzv_copy16(&P.stage[2 + i][lane], cull + i); }
#pragma unroll
  for(int i = 0; i < 9; ++i) { bzv_copy16(&P.stage[6 + i][lane], newton + i); }
}

// One candidate per lane.  Batches of 32 candidates are dealt to the warps of the grid round-robin (every batch costs the
// same: a fixed six-trip loop).  The list is read two batches ahead and the inputs of the next batch -- its rays (a gather from
// the ray buffer in HBM) and its patch records (L2) -- are staged into shared memory by cp.async while the current batch is
// in its root search, so a batch starts on data that is already on the SM (north_star subsystem 1: tables staged per block
// through shared memory).  Hops are evaluated from a small per-warp queue, 32 at a time, with direct loads.
__global__ void __launch_bounds__(32 * BZV_EVAL_WARPS, BZV_EVAL_CTAS) bz_wave_eval_kernel(const bz_tables T, const bz_wave W, const float one, unsigned long long *stats) {
  __shared__ __align__(16) bzv_eval_pool pools[BZV_EVAL_WARPS];
  bzv_eval_pool &P = pools[threadIdx.x >> 5];
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t total = (W.counters[3] != 0u) ? 0u : min(W.counters[1], W.cand_cap);      // (after an overflow the list has holes: nothing is evaluated, bz_lens_check reports it)
  if(blockIdx.x == 0u && threadIdx.x == 0u) { W.counters[0] = 0u; }                        // the walk kernel of this pass is done: its tile counter, for the next pass
  uint32_t evaluations = 0u, hop_tests = 0u;      // per lane
  const uint32_t stride = gridDim.x * BZV_EVAL_WARPS * 32u;
  uint32_t first = (blockIdx.x * BZV_EVAL_WARPS + (threadIdx.x >> 5)) * 32u;               // the warp's next batch of the list
  // entry_now: the batch at `first`, its inputs staged (or in flight); entry_next: the batch after it
  uint2 entry_now = make_uint2(0u, 0u), entry_next = make_uint2(0u, 0u);
  if(first + lane < total) { entry_now = __ldcs(W.cand + first + lane); }
  if(first + stride + lane < total) { entry_next = __ldcs(W.cand + first + stride + lane); }
  if(first + lane < total) { bzv_stage(P, T, W, entry_now, lane); }
  asm volatile("cp.async.commit_group;");
  if(lane == 0u) { P.hop_n = 0u; }
  __syncwarp();
  for(;;) {
    // the next batch: 32 hops when there are that many (or nothing else is left), else the warp's next 32 candidates of the list
    __syncwarp();
    const uint32_t hops = P.hop_n;
    __syncwarp();
    uint32_t index = 0u, patch = 0u;
    uint2 entry = make_uint2(0u, 0u), entry_ahead = make_uint2(0u, 0u);
    bool active = false, first_level = true, reloaded = false;
    float4 ra = make_float4(0.0f, 0.0f, 0.0f, 0.0f), rb = ra;
    bz_cull_rec c;
    uint32_t own;
    bz_newton_rec t;
#if !defined(BZ_FAST)
    bzw_newton2 t2;
#endif
    if(hops >= 32u || (first >= total && hops != 0u)) {
      const uint32_t n = min(hops, 32u);
      if(lane < n) {
        index = P.hop_index[hops - n + lane];
        patch = P.hop_patch[hops - n + lane];
        entry = __ldcs(W.cand + index);
        active = true;
        ra = W.ray_a[entry.x];
        rb = W.ray_b[entry.x];
        bzw_unpack_cull(T.cull + (size_t)patch * BZ_CULL_F4, c, own);
#if defined(BZ_FAST)
        bzw_load_newton_root(T.newton + (size_t)patch * BZ_NEWTON_F4, t);
#else
        bzw_load_newton2(T.newton + (size_t)patch * BZ_NEWTON_F4, one, t2);
#endif
      }
      first_level = false;
      __syncwarp();
      if(lane == 0u) { P.hop_n = hops - n; }
    }
    else if(first < total) {
      index = first + lane;
      active = index < total;
      entry = entry_now;
      patch = entry.y;
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
      if(active) {
        ra = P.stage[0][lane];
        rb = P.stage[1][lane];
        bzw_unpack_cull<true>(&P.stage[2][lane], c, own);
#if defined(BZ_FAST)
        bzw_load_newton_root<true>(&P.stage[6][lane], t);
#else
        bzw_load_newton2<true>(&P.stage[6][lane], one, t2);
#endif
      }
      __syncwarp();
      // the buffer is free again: the inputs of the batch after this one, and the list two batches ahead
      first += stride;
      entry_now = entry_next;
      if(first + lane < total) { bzv_stage(P, T, W, entry_now, lane); }
      asm volatile("cp.async.commit_group;");
      if(first + stride + lane < total) { entry_ahead = __ldcs(W.cand + first + stride + lane); }
      reloaded = true;
    }
    else { break; }
    if(active) {
      const bzv3 s = bz_v(ra.x, ra.y, ra.z), d = bz_v(ra.w, rb.x, rb.y);
      unsigned long long key = BZ_KEY_NONE;
      float cosine, distance;
      if(!first_level) { ++hop_tests; }
      if(bz_stage_a(c, s, d, first_level, cosine, distance)) {
        bz_hit hit;
        hit.what = BZ_MISS;
#if defined(BZ_FAST)
        if(BZV_ROOT(c, t, s, d, cosine, distance, hit)) {
#else
        if(BZV_ROOT(c, t2, s, d, cosine, distance, hit)) {
#endif
          bzw_load_newton_side(T.newton + (size_t)patch * BZ_NEWTON_F4, t);
          bz_stage_b_side(t, d, hit);
        }
        ++evaluations;
        if(hit.what <= BZ_FOLLOW2) {
          if(first_level) {                  // one hop, no recursion (bezierMesh.cpp:213-217)
            const uint32_t at = atomicAdd(&P.hop_n, 1u);
            P.hop_index[at] = index;
            P.hop_patch[at] = (hit.what == BZ_FOLLOW0) ? t.nb[0] : ((hit.what == BZ_FOLLOW1) ? t.nb[1] : t.nb[2]);
          }
        }
        else if(hit.what == BZ_HIT && hit.distance < 3.402823466e+38f) {      // bezierMesh.cpp:209, 220
          key = bzw_make_key(hit.distance, entry.y);                           // loop index = the patch the loop stood at
          __stcs(W.cand_pay + 2u * (size_t)index, make_float4(hit.point.x, hit.point.y, hit.point.z, hit.cos_incidence));
          __stcs(W.cand_pay + 2u * (size_t)index + 1u, make_float4(hit.normal.x, hit.normal.y, hit.normal.z, __uint_as_float(patch)));
        }
      }
      if(first_level || key != BZ_KEY_NONE) { __stcs(W.cand_key + index, key); }
    }
    // the list entries requested before the root search are taken over only now: touching the destination of a load in
    // flight (even to copy it into the loop-carried register) would park the warp on it for a whole DRAM round trip
    asm volatile("" : "+r"(entry_ahead.x), "+r"(entry_ahead.y));
    if(reloaded) { entry_next = entry_ahead; }
  }
  evaluations = __reduce_add_sync(0xffffffffu, evaluations);
  hop_tests = __reduce_add_sync(0xffffffffu, hop_tests);
  if(stats != nullptr && lane == 0u) {
    if(evaluations != 0u) { atomicAdd(stats + BZC_STAGE_B, (unsigned long long)evaluations); }
    if(hop_tests != 0u) { atomicAdd(stats + BZC_STAGE_A, (unsigned long long)hop_tests); }
  }
}

// ------------------------------------------------------------------ resolve ----
__global__ void __launch_bounds__(256) bz_wave_resolve_kernel(const bz_tables T, const bz_job J, const bz_wave W) {
  const uint32_t rays = W.tile_count * W.packets_per_tile * 32u;
  const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if(slot == 0u) { W.counters[1] = 0u; }      // the evaluation of this pass is done: the candidate count, for the next pass
  const bool last = (W.level + 1u >= max(1u, J.scene.max_crossings));      // every ray ends here: nothing is kept
  uint32_t events = 0u;                    // byte 0 Snell block reached, 1 entered, 2 exited, 3 lost
  if(slot < rays) {
    // the three loads of a ray are issued together (the kernel is latency-bound: one dependent step less per ray)
    const float4 rb = __ldcs(W.ray_b + slot);
    const float4 ra = __ldcs(W.ray_a + slot);
    uint32_t link = __ldcs(W.head + slot);
    const uint32_t st = __float_as_uint(rb.w);
    if((st & 1u) != 0u) {
      uint32_t expected = (st >> 1) & 3u, crossings = st >> 4;
      // the nearest result of the ray's chain: (distance, loop index) keys are distinct, strict < of bezierMesh.cpp:220
      unsigned long long best = BZ_KEY_NONE;
      uint32_t winner = 0u;
      while(link != 0u) {
        const unsigned long long key = __ldcs(W.cand_key + (link - 1u));
        const uint32_t next = __ldcs(W.cand_next + (link - 1u));
        if(key < best) { best = key; winner = link - 1u; }
        link = next;
      }
      bzv3 s = bz_v(ra.x, ra.y, ra.z), d = bz_v(ra.w, rb.x, rb.y);
      const bool found = (best != BZ_KEY_NONE);
      uint32_t status = BZ_STATUS_NONE;
      bzv3 refracted = d, point = s;
      if(found) {
        const float4 p0 = __ldcs(W.cand_pay + 2u * (size_t)winner), p1 = __ldcs(W.cand_pay + 2u * (size_t)winner + 1u);
        point = bz_v(p0.x, p0.y, p0.z);
        status = bz_snell(T.ri, p0.w, bz_v(p1.x, p1.y, p1.z), refracted);
        events |= 1u;
      }
      status = (status == expected) ? status : BZ_STATUS_NONE;
      if(status != BZ_STATUS_NONE) {
        s = point;
        d = refracted;
        events |= (status == BZ_STATUS_INSIDE) ? 0x100u : 0x10000u;
      }
      bool done = (status == BZ_STATUS_NONE);
      if(!done) {
        expected = (expected == BZ_STATUS_INSIDE) ? BZ_STATUS_OUTSIDE : BZ_STATUS_INSIDE;
        ++crossings;
        done = (crossings >= J.scene.max_crossings);
      }
      if(done) {
        uint32_t value;
        if(expected == BZ_STATUS_INSIDE) { value = bzw_shade_background(J.scene, s, d); }      // the ray is in air: it reaches the background
        else { value = J.scene.value_lost; events |= 0x1000000u; }                             // entered and never left
        atomicAdd(J.accum + __float_as_uint(rb.z), value);
        if(!last) { W.ray_b[slot] = make_float4(rb.x, rb.y, rb.z, __uint_as_float(0u)); }
      }
      else {
        W.ray_a[slot] = make_float4(s.x, s.y, s.z, d.x);
        W.ray_b[slot] = make_float4(d.y, d.z, rb.z, __uint_as_float(1u | (expected << 1) | (crossings << 4)));
      }
    }
  }
  if(J.stats != nullptr) {
    // event counts: warp totals into shared memory, one atomic per counter and CTA (millions of rays end here per pass)
    __shared__ uint32_t tally[4];
    if(threadIdx.x < 4u) { tally[threadIdx.x] = 0u; }
    __syncthreads();
    const uint32_t snell = __reduce_add_sync(0xffffffffu, events & 0xffu), entered = __reduce_add_sync(0xffffffffu, (events >> 8) & 0xffu);
    const uint32_t exited = __reduce_add_sync(0xffffffffu, (events >> 16) & 0xffu), lost = __reduce_add_sync(0xffffffffu, events >> 24);
    if((threadIdx.x & 31u) == 0u) {
      if(snell != 0u) { atomicAdd(&tally[0], snell); }
      if(entered != 0u) { atomicAdd(&tally[1], entered); }
      if(exited != 0u) { atomicAdd(&tally[2], exited); }
      if(lost != 0u) { atomicAdd(&tally[3], lost); }
    }
    __syncthreads();
    if(threadIdx.x < 4u && tally[threadIdx.x] != 0u) { atomicAdd(J.stats + BZC_SNELL + threadIdx.x, (unsigned long long)tally[threadIdx.x]); }
  }
}

}  // namespace BZ_NS
using namespace BZ_NS;

static size_t bzv_walk_shared() { return sizeof(bzv_walk_pool) * BZV_WALK_WARPS; }

// One crossing of one chunk: walk + stage A, stage B, resolve, stream-ordered.  `counters` [0..2] are zero on entry (the handle
// zeroes them once per frame) and are left zero: each kernel resets what the kernel before it consumed.
cudaError_t BZ_SYM(bz_launch_wave_pass)(const bz_tables &tables, const bz_job &desc, const bz_wave &wave, int sm_count, cudaStream_t stream) {
  static int walk_ctas[64] = {}, eval_ctas[64] = {};          // resident CTAs per SM, per device (0: not asked yet)
  int device = 0;
  cudaError_t e = cudaGetDevice(&device);
  if(e != cudaSuccess) { return e; }
  device &= 63;
  if(walk_ctas[device] == 0) {
    e = cudaFuncSetAttribute(bz_wave_walk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bzv_walk_shared());
    if(e != cudaSuccess) { return e; }
    int walk = 0, eval = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&walk, bz_wave_walk_kernel, 32 * BZV_WALK_WARPS, bzv_walk_shared());
    if(e != cudaSuccess) { return e; }
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&eval, bz_wave_eval_kernel, 32 * BZV_EVAL_WARPS, 0);
    if(e != cudaSuccess) { return e; }
    eval_ctas[device] = eval > 0 ? eval : 1;
    walk_ctas[device] = walk > 0 ? walk : 1;
  }
  const uint32_t packets = wave.tile_count * wave.packets_per_tile;
  if(packets == 0u) { return cudaSuccess; }
  const uint32_t walk_grid = std::min<uint32_t>((packets + BZV_WALK_WARPS - 1u) / BZV_WALK_WARPS, (uint32_t)sm_count * (uint32_t)walk_ctas[device]);
  bz_wave_walk_kernel<<<walk_grid, 32 * BZV_WALK_WARPS, bzv_walk_shared(), stream>>>(tables, desc, wave);
  bz_wave_eval_kernel<<<(uint32_t)sm_count * (uint32_t)eval_ctas[device], 32 * BZV_EVAL_WARPS, 0, stream>>>(tables, wave, desc.one, desc.stats);
  bz_wave_resolve_kernel<<<(packets * 32u + 255u) / 256u, 256, 0, stream>>>(tables, desc, wave);
  return cudaGetLastError();
}

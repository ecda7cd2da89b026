// bz_trace.cu -- the persistent, warp-scheduled intersect/refract kernel (BZ_MODE_CULLED) for sm_100a.
//
// One kernel, bz_warp_trace_kernel<JOB>, serves the three entry points of the C ABI:
//   JOB_RENDER     K1 ray generation + bounce loop around BezierLens::refract + K3 background shading
//   JOB_REFRACT    BezierLens::refract over a ray batch   (reference/bezierLens.cpp:4-34)
//   JOB_INTERSECT  BezierMesh::intersect over a ray batch (reference/bezierMesh.cpp:206-227)
//
// Every WARP is its own scheduler: it owns a pool of BZ_SLOTS = 32 * BZ_GROUPS ray slots and two work queues in
// shared memory and never meets another warp at a barrier.  One pass of its loop is one refract() call for every live
// slot:
//   refill     a group whose 32 rays are all finished takes the next 32 primary-ray indices from a global counter (one
//              atomic per warp and group): consecutive rays are neighbours on the image, and the rays of a group stay
//              together across bounces, so a group is a coherent packet
//   walk       PACKET walk first: the 32 rays of a group are bounded by an origin box and a direction box (CREDUX.F32
//              warp reductions), and the warp walks the 32-children-per-node form of the cull hierarchy cooperatively,
//              ONE LANE PER CHILD BOX, with a conservative interval slab test and a node stack in shared memory; the
//              leaves the packet may meet become a candidate list.  A packet that is not coherent enough (candidate
//              list or stack overflow) falls back to the LANE walk: per lane, the binary hierarchy in depth-first
//              order with skip links (no stack), leaf hits push (slot, leaf record) pairs to queue A
//   stage A    plane hit, self-hit gate, barycentric inside test (bezierTriangle.cpp:124-131); survivors go to queue B.
//              Packet candidates are tested record by record by all rays of the group at once (one broadcast record
//              load per patch); queue A pairs of the lane walk are consumed 32 at a time
//   stage B    queue B is consumed 32 candidates at a time, whatever slot they belong to: bracket + 4 root-search steps
//              (cpp:132-194); follow-side results are appended to queue B as one-hop candidates (bezierMesh.cpp:213-217);
//              the nearest hit per slot is a 64-bit shared-memory atomicMin on (distance, loop index) -- the reference's
//              "first strictly smaller distance wins" (bezierMesh.cpp:220)
//   resolve    per slot: Snell refraction, expected-side check, next bounce or shading
// Which of walk / stage A / stage B runs next is decided by queue levels, so the two stages always run on full
// 32-lane batches except when a pass runs dry.
// Compiled with -fmad=false (see bz_math.cuh); the box tests use explicit FMAs, they only have to be conservative.
// The file is compiled twice (Makefile): as the parity build (-fmad=false, IEEE division and square root, bit-exact with
// the reference) and, with -DBZ_FAST -fmad=true, as the fast build behind BZ_MODE_FAST (csrc/bz_math.cuh says what
// changes).  Each build lives in its own namespace; the launchers of the fast build carry the suffix _fast.
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "bz_device.h"
#if defined(BZ_FAST)
#define BZ_NS bz_fast
#define BZ_SYM(name) name##_fast
#else
#define BZ_NS bz_exact
#define BZ_SYM(name) name
#endif
namespace BZ_NS {
#include "bz_math.cuh"
#include "bz_trace_common.cuh"

#ifndef BZ_GROUPS
#define BZ_GROUPS    2
#endif
#ifndef BZ_WARPS
#define BZ_WARPS     16      // one CTA per SM: its warps share tiles of the job (L1 locality), never a barrier
#endif
#ifndef BZ_MIN_CTAS
#define BZ_MIN_CTAS  1
#endif
#define BZ_SLOTS     (32 * BZ_GROUPS)
#ifndef BZ_WALK_BURST
#define BZ_WALK_BURST 2
#endif
//        // hierarchy steps per lane between two looks at the level of queue A
#ifndef BZ_QA_CAP
#define BZ_QA_CAP    256
#endif
#ifndef BZ_CAND_CAP
#define BZ_CAND_CAP  160     // leaves a packet may collect before it is declared incoherent
#endif
#define BZ_STACK_CAP 96
#define BZ_QB_CAP    160
#define BZ_LEAF_PUSH BZ_LEAF_MAX // upper bound of pairs one lane pushes in one walk step

struct bz_warp_pool {
  float              ray[6][BZ_SLOTS];       // start xyz, dir xyz
  unsigned long long best[BZ_SLOTS];         // (distance key, loop index) of the nearest hit of this pass
  float              pay[11][BZ_SLOTS];      // point 3, normal 3, cos, bary 3, evaluated patch
  uint32_t           id_lo[BZ_SLOTS];        // RENDER: local pixel index; batch jobs: ray index
  uint32_t           id_hi[BZ_SLOTS];
  uint32_t           state[BZ_SLOTS];        // bit 0 live, bits 1-2 expected status, bits 4.. crossings so far
  uint32_t           qa_leaf[BZ_QA_CAP];     // queue A: record index in the `leaf` table
  uint32_t           qb_patch[BZ_QB_CAP];    // queue B: patch to evaluate
  uint32_t           qb_loop[BZ_QB_CAP];     //          loop index of the candidate (the originating patch for a hop)
  uint8_t            qa_slot[BZ_QA_CAP];
  uint8_t            qb_slot[BZ_QB_CAP];     //          slot | 0x80 for a first-level candidate (inside test applies, may hop)
  uint32_t           cand[BZ_CAND_CAP];      // packet walk: leaf codes (first slot << 3 | count) the packet may meet
  uint32_t           stack[BZ_STACK_CAP];    // packet walk: 32-wide nodes still to visit
  uint32_t           qa_n, qb_n;
  unsigned long long count[BZ_NSTAT];        // work counters of this warp (bz_stats order), kept out of the register file
};

// work counters live in shared memory and are bumped by lane 0 with warp-level totals
// (the amount is evaluated by every lane: it may contain a warp vote)
#define BZW_COUNT(index, amount) do { const unsigned long long amount_ = (unsigned long long)(amount); if(lane == 0u) { P.count[index] += amount_; } } while(0)

static_assert(BZ_SLOTS <= 128, "queue entries keep the slot in 7 bits");
static_assert(BZ_QA_CAP >= 32 * BZ_LEAF_PUSH * BZ_WALK_BURST + 64 && BZ_QB_CAP >= 96, "queue thresholds");

// ---- work distribution: tiles of the job from a global counter per CTA, packets of a tile from shared memory per warp ----
// cta_work = tile << 32 | rays of the tile handed out so far.  The warp whose packet is the first one past the end of the
// tile fetches the next tile (and takes its first packet); warps arriving in between wait for the tile number to change.
// RENDER: tile = tile_y * tiles_x + tile_x.  With a tile list (the tiles that may meet the lens, csrc/bz_capi.cu) the order
// is the list's; without one the rows are taken from the outside in (top, bottom, second, second to last, ...): the costliest
// packets of a frame sit on the silhouette of the body (grazing rays meet many leaf boxes and walk per lane), the centre
// rows are uniform -- ending on them keeps the tail of the persistent grid short, which matters for the short per-GPU
// launches of a row-sharded frame.
__device__ __forceinline__ unsigned long long bzw_fetch_tile(const bz_job &J) {
  const unsigned long long next = atomicAdd(J.next_ray, 1ull);
  if(next >= J.tile_count) { return (unsigned long long)BZ_END << 32; }
  uint32_t tile = (uint32_t)next;
  if(J.tile_list != nullptr) { tile = __ldg(J.tile_list + next); }
  else if(J.tile_side != 0u) {
    const uint32_t turn = tile / J.tiles_x, tile_x = tile - turn * J.tiles_x;
    const uint32_t tile_y = ((turn & 1u) == 0u) ? (turn >> 1) : (J.tiles_y - 1u - (turn >> 1));
    tile = tile_y * J.tiles_x + tile_x;
  }
  return (unsigned long long)tile << 32;
}

__device__ __forceinline__ unsigned long long bzw_take_packet(unsigned long long *cta_work, const bz_job &J) {
  for(;;) {
    const unsigned long long w = atomicAdd(cta_work, 32ull);
    const uint32_t tile = (uint32_t)(w >> 32), at = (uint32_t)w;
    if(tile == BZ_END || at < J.tile_rays) { return w; }
    if(at < J.tile_rays + 32u) {
      const unsigned long long fresh = bzw_fetch_tile(J);
      atomicExch(cta_work, fresh | 32ull);
      return fresh;
    }
    while((uint32_t)(*(volatile unsigned long long*)cta_work >> 32) == tile) { __nanosleep(64); }
  }
}

template<int JOB>
__global__ void __launch_bounds__(32 * BZ_WARPS, BZ_MIN_CTAS) bz_warp_trace_kernel(const bz_tables T, const bz_job J) {
  extern __shared__ __align__(16) unsigned char bzw_shared[];
  bz_warp_pool &P = reinterpret_cast<bz_warp_pool*>(bzw_shared)[threadIdx.x >> 5];
  __shared__ unsigned long long cta_work;    // tile of the job this CTA is working on << 32 | rays of it handed out
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t spp = J.camera.strata * J.camera.strata;
  const uint32_t direct_end = T.always_first + T.always_count;
  bool exhausted = false;                    // warp-uniform: the global ray counter has run past the end
  // batch jobs: rays of the call (BZ_JOB_TARGET after bz_pass_kernel: the count is known on the device only)
  const unsigned long long ray_count = (JOB == BZ_JOB_TARGET && J.ray_count_dev != nullptr) ? (unsigned long long)__ldg(J.ray_count_dev) : J.ray_count;
  uint32_t tally = 0u, tally_turns = 0u;     // per lane: packed event counters of the resolve step; warp-uniform: turns since their last fold

#pragma unroll
  for(int g = 0; g < BZ_GROUPS; ++g) { P.state[g * 32 + lane] = 0u; }
  if(lane == 0u) { P.qa_n = 0u; P.qb_n = 0u; }
  if(lane < BZ_NSTAT) { P.count[lane] = 0ull; }
  if(threadIdx.x == 0u) { cta_work = bzw_fetch_tile(J); }
  __syncthreads();                           // the only CTA barrier of the kernel
  __syncwarp();

  for(;;) {
    // ================= one pass: one refract() call for every live slot =================
    int group = 0;                           // warp-uniform: group whose walk is in progress; BZ_GROUPS = all walked
    bool group_ready = false;                // warp-uniform: refill of `group` done, walk state valid
    bool any_live = false;                   // warp-uniform
    uint32_t node = BZ_END, direct = BZ_END; // per lane: lane-walk state of the current group
    uint32_t cand_n = 0u, cand_at = 0u, cand_sub = 0u, cand_group = 0u;   // warp-uniform: packet candidates still to be tested by stage A

    for(;;) {
      __syncwarp();
      const uint32_t qa = P.qa_n, qb = P.qb_n;
      __syncwarp();
      const bool walked = (group == BZ_GROUPS);
      int action;                            // warp-uniform: 0 walk, 1 stage A batch, 2 stage B batch, 3 pass complete, 4 stage A on packet candidates
      if(qb > BZ_QB_CAP - 64u) { action = 2; }
      else if(cand_at < cand_n) { action = 4; }
      else if(qa > BZ_QA_CAP - 32u * BZ_LEAF_PUSH * BZ_WALK_BURST || (walked && qa != 0u)) { action = 1; }
      else if(!walked) { action = 0; }
      else if(qb != 0u) { action = 2; }
      else { action = 3; }
      if(action == 3) { break; }

      if(action == 0) {
        const uint32_t slot = (uint32_t)group * 32u + lane;
        if(!group_ready) {
          // ---- refill: a group whose rays are all finished takes the next 32 primary rays ----
          uint32_t st = P.state[slot];
          bool live = (st & 1u) != 0u;
          const uint32_t dead = __ballot_sync(0xffffffffu, !live);
          const uint32_t need = (dead == 0xffffffffu) ? dead : 0u;
          while(need != 0u && !exhausted) {
            // the CTA's warps share one tile of the job (a square block of pixels, or 1024 consecutive batch rays), so
            // the 16 warps of an SM work next to each other and meet the same hierarchy nodes and patch records in L1
            unsigned long long w = 0ull;
            if(lane == 0u) { w = bzw_take_packet(&cta_work, J); }
            w = __shfl_sync(0xffffffffu, w, 0);
            const uint32_t tile = (uint32_t)(w >> 32), at = (uint32_t)w + lane;
            if(tile == BZ_END) { exhausted = true; break; }
            bool valid = at < J.tile_rays;
            bzv3 s, d;
            uint32_t expected = BZ_STATUS_INSIDE;
            if(JOB == BZ_JOB_RENDER) {
              // (divisions by per-job constants are multiplications by the host's reciprocals, bz_job::div_*)
              const uint32_t inside = bzw_div(at, J.div_spp), sample = at - inside * spp;   // pixel of the tile in Morton order, sample of the pixel
              const uint32_t tile_y = tile / J.tiles_x, tile_x = tile - tile_y * J.tiles_x;
              const uint32_t px = tile_x * J.tile_side + bzw_even_bits(inside), row = tile_y * J.tile_side + bzw_even_bits(inside >> 1);
              valid = valid && px < J.camera.width && row < J.rows;
              if(valid) {
                const uint32_t sy = bzw_div(sample, J.div_strata);
                bzw_primary_ray(J.camera, J.two_over_w, J.two_over_h, px, bzw_frame_row(J, row), sample - sy * J.camera.strata, sy, s, d);
                P.id_lo[slot] = row * J.camera.width + px;
              }
            }
            else {
              const unsigned long long id = (unsigned long long)tile * J.tile_rays + at;
              valid = valid && id < ray_count;
              if(valid) {
                s = bz_v(J.in_start[id * 3u], J.in_start[id * 3u + 1u], J.in_start[id * 3u + 2u]);
                d = bz_v(J.in_dir[id * 3u], J.in_dir[id * 3u + 1u], J.in_dir[id * 3u + 2u]);
                expected = (JOB == BZ_JOB_REFRACT) ? (J.in_expected[id] & 3u) : ((JOB == BZ_JOB_TARGET) ? BZ_STATUS_INSIDE : 0u);
                P.id_lo[slot] = (uint32_t)id;
                P.id_hi[slot] = (uint32_t)(id >> 32);
              }
            }
            if(valid) {
              P.ray[0][slot] = s.x; P.ray[1][slot] = s.y; P.ray[2][slot] = s.z;
              P.ray[3][slot] = d.x; P.ray[4][slot] = d.y; P.ray[5][slot] = d.z;
              st = 1u | (expected << 1);
              P.state[slot] = st;
              live = true;
            }
            if(__ballot_sync(0xffffffffu, valid) != 0u) { break; }      // (a packet of an edge tile may lie outside the image)
          }
          P.best[slot] = BZ_KEY_NONE;
          const uint32_t live_mask = __ballot_sync(0xffffffffu, live);
          any_live = any_live || (live_mask != 0u);
          BZW_COUNT(BZC_PRIMARY, __popc(live_mask & need));
          BZW_COUNT(BZC_CALLS, __popc(live_mask));
          if(live_mask == 0u) { ++group; continue; }      // nothing to walk in this group
          const float sx = P.ray[0][slot], sy = P.ray[1][slot], sz = P.ray[2][slot];
          const float l1 = fabsf(sx) + fabsf(sy) + fabsf(sz);
          // ---- packet walk: one lane per child box of the 32-wide hierarchy ----
          bool packet_done = false;
          if(T.packet_count != 0u) {
            const float inf = __int_as_float(0x7f800000);
            const float dx = P.ray[3][slot], dy = P.ray[4][slot], dz = P.ray[5][slot];
            const float l1_max = bzw_warp_max(live ? l1 : 0.0f);
            // twice the largest per-ray slack of the lane walk: one share is the cull's own guarantee (host/bz_bvh.cpp),
            // the other covers the rounding of the interval arithmetic below (a few ulp of |box| + |origin|)
            const float margin = 2.0f * 3.0517578125e-05f * (l1_max + 2.0f * T.extent);
            const bzw_axis ax = bzw_make_axis(bzw_warp_min(live ? sx : inf), bzw_warp_max(live ? sx : -inf),
                                              bzw_warp_min(live ? dx : inf), bzw_warp_max(live ? dx : -inf), margin);
            const bzw_axis ay = bzw_make_axis(bzw_warp_min(live ? sy : inf), bzw_warp_max(live ? sy : -inf),
                                              bzw_warp_min(live ? dy : inf), bzw_warp_max(live ? dy : -inf), margin);
            const bzw_axis az = bzw_make_axis(bzw_warp_min(live ? sz : inf), bzw_warp_max(live ? sz : -inf),
                                              bzw_warp_min(live ? dz : inf), bzw_warp_max(live ? dz : -inf), margin);
            const uint32_t below = (1u << lane) - 1u;
            // records outside the hierarchy: the always segment, and the wide segment when a ray starts beyond origin_limit
            const uint32_t direct_first = (l1_max <= T.origin_limit) ? T.always_first : T.wide_first;
            const uint32_t direct_n = direct_end - direct_first;
            uint32_t sp = 1u, found = (direct_n + 6u) / 7u;
            packet_done = (found <= 32u);      // one seed entry per lane; more records than that outside the hierarchy: per-lane walk
            if(packet_done) {
              if(lane < found) { P.cand[lane] = ((direct_first + 7u * lane) << 3) | min(7u, direct_n - 7u * lane); }
              if(lane == 0u) { P.stack[0] = 0u; }
              __syncwarp();
            }
            uint32_t visits = 0u;
            while(packet_done && sp != 0u) {
              --sp;
              const uint32_t node = P.stack[sp] * 224u;             // (the table has < 2^32 words: bz_lens_create checks)
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
              if(sp + __popc(hit_inner) > BZ_STACK_CAP || found + __popc(hit_leaf) > BZ_CAND_CAP) { packet_done = false; break; }
              if(hit && inner) { P.stack[sp + __popc(hit_inner & below)] = code & 0x7fffffffu; }
              if(hit && !inner) { P.cand[found + __popc(hit_leaf & below)] = code; }
              sp += __popc(hit_inner);
              found += __popc(hit_leaf);
              __syncwarp();
            }
            BZW_COUNT(BZC_PNODES, visits);
            if(packet_done) {
              cand_n = found; cand_at = 0u; cand_sub = 0u; cand_group = (uint32_t)group;
              BZW_COUNT(BZC_PWALKS, 1u);
              ++group;
              continue;
            }
            __syncwarp();
          }
          // ---- lane walk (no packet hierarchy, or the packet was not coherent enough) ----
          BZW_COUNT(BZC_LWALKS, __popc(live_mask));
          if(live) {
            direct = (l1 <= T.origin_limit) ? T.always_first : T.wide_first;     // far origins also test the wide segment
            if(direct >= direct_end) { direct = BZ_END; }
            node = (T.node_count != 0u) ? 0u : BZ_END;
          }
          else {
            direct = BZ_END;
            node = BZ_END;
          }
          group_ready = true;
        }
        // ---- walk: records outside the hierarchy first, then the hierarchy, until done or queue A fills up ----
        {
          const float sx = P.ray[0][slot], sy = P.ray[1][slot], sz = P.ray[2][slot];
          const float dx = P.ray[3][slot], dy = P.ray[4][slot], dz = P.ray[5][slot];
          // per-ray slack of the box tests: covers the rounding of stage A's in-plane point and barycentrics for
          // patches whose conditioning is <= 16 (host/bz_bvh.cpp)
          const float slack = 3.0517578125e-05f * (fabsf(sx) + fabsf(sy) + fabsf(sz) + 2.0f * T.extent);
          const float ix = (fabsf(dx) > 1e-20f) ? 1.0f / dx : copysignf(1e20f, dx);
          const float iy = (fabsf(dy) > 1e-20f) ? 1.0f / dy : copysignf(1e20f, dy);
          const float iz = (fabsf(dz) > 1e-20f) ? 1.0f / dz : copysignf(1e20f, dz);
          const float ax_lo = -(sx + slack) * ix, ax_hi = -(sx - slack) * ix;
          const float ay_lo = -(sy + slack) * iy, ay_hi = -(sy - slack) * iy;
          const float az_lo = -(sz + slack) * iz, az_hi = -(sz - slack) * iz;
          // records outside the hierarchy (always segment; wide segment for far origins): one pair per lane and turn
          uint32_t room = BZ_QA_CAP - qa;      // warp-uniform lower bound of the free entries of queue A
          bool paused = false;
          while(__ballot_sync(0xffffffffu, direct != BZ_END) != 0u) {
            if(room < 32u) { paused = true; break; }
            if(direct != BZ_END) {
              const uint32_t at = atomicAdd(&P.qa_n, 1u);
              P.qa_leaf[at] = direct;
              P.qa_slot[at] = (uint8_t)slot;
              ++direct;
              if(direct >= direct_end) { direct = BZ_END; }
            }
            room -= 32u;
          }
          // a burst of the hierarchy walk pushes up to 32 * BZ_LEAF_PUSH * BZ_WALK_BURST pairs before it looks at the level
          // again: the direct records above may have used that room up
          if(room < 32u * BZ_LEAF_PUSH * BZ_WALK_BURST) { paused = true; }
          // the hierarchy: BZ_WALK_BURST steps per lane between two looks at the queue level
          uint32_t steps = 0u;
          while(!paused) {
#pragma unroll
            for(int u = 0; u < BZ_WALK_BURST; ++u) {
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
                if(hit && code != 0u) {            // leaf: its records go to queue A
                  const uint32_t count = code & 7u;
                  const uint32_t at = atomicAdd(&P.qa_n, count);
                  for(uint32_t j = 0u; j < count; ++j) {
                    P.qa_leaf[at + j] = (code >> 3) + j;
                    P.qa_slot[at + j] = (uint8_t)slot;
                  }
                }
                // inner node that was hit: its first child is the next node; otherwise skip the subtree (BZ_END after the last)
                node = (hit && code == 0u) ? node + 1u : __float_as_uint(lo.w);
              }
            }
            if(__ballot_sync(0xffffffffu, node != BZ_END) == 0u) {
              group_ready = false;
              ++group;
              break;
            }
            __syncwarp();
            if(P.qa_n > BZ_QA_CAP - 32u * BZ_LEAF_PUSH * BZ_WALK_BURST) { break; }
          }
          steps = __reduce_add_sync(0xffffffffu, steps);
          BZW_COUNT(BZC_NODES, steps);
        }
        continue;
      }

      if(action == 4) {
        // ---- stage A, packet form: every candidate record is tested by all live rays of its group at once ----
        const uint32_t slot = cand_group * 32u + lane;
        const bool live = (P.state[slot] & 1u) != 0u;
        const bzv3 s = bz_v(P.ray[0][slot], P.ray[1][slot], P.ray[2][slot]);
        const bzv3 d = bz_v(P.ray[3][slot], P.ray[4][slot], P.ray[5][slot]);
        const uint32_t below = (1u << lane) - 1u;
        uint32_t level = qb;                  // warp-uniform copy of the level of queue B
        uint32_t tested = 0u;
        // Two records per trip on the packed FP32 pipe: record j of the `pairs` table holds leaf records j and j + 1
        // interleaved word by word, so every float4 is two aligned (A, B) register pairs.  All addresses are the same for
        // the whole warp (broadcast loads).  Each half sees the operations of bz_stage_a in their order; the two IEEE
        // divisions stay scalar.
        const float one = J.one;
        while(cand_at < cand_n && level <= BZ_QB_CAP - 64u) {
          const uint32_t code = P.cand[cand_at];
          const uint32_t left = (code & 7u) - cand_sub;             // records of this entry still to be tested
          const float4 *rec = T.pairs + (size_t)((code >> 3) + cand_sub) * (2u * BZ_CULL_F4);
          const float4 r0 = __ldg(rec), r1 = __ldg(rec + 1), r2 = __ldg(rec + 2), r3 = __ldg(rec + 3);
          const float4 r4 = __ldg(rec + 4), r5 = __ldg(rec + 5), r6 = __ldg(rec + 6), r7 = __ldg(rec + 7);
          cand_sub += min(left, 2u);
          if(left <= 2u) { cand_sub = 0u; ++cand_at; }
          tested += min(left, 2u);
          // words of a cull record: 0-2 n, 3 k, 4-6 Binv row 0, 7 hIn, 8-10 row 1, 11 hOut, 12-14 row 2, 15 patch
          const float2 nx = make_float2(r0.x, r0.y), ny = make_float2(r0.z, r0.w), nz = make_float2(r1.x, r1.y), kk = make_float2(r1.z, r1.w);
          const float2 cosine = bzw_add2(bzw_mul2(nx, d.x), bzw_add2(bzw_mul2(ny, d.y), bzw_mul2(nz, d.z), one), one);         // bz_dot(d, n)
          const float2 along = bzw_add2(bzw_mul2(nx, s.x), bzw_add2(bzw_mul2(ny, s.y), bzw_mul2(nz, s.z), one), one);          // bz_dot(n, s)
#if defined(BZ_FAST)
          const float2 above = __fadd2_rn(kk, make_float2(-along.x, -along.y));
#else
          const float2 above = __ffma2_rn(along, make_float2(-one, -one), kk);                                                  // k - bz_dot(n, s)
#endif
          bool pass_a = live && fabsf(cosine.x) >= BZ_PLANE_EPSILON;
          bool pass_b = live && left >= 2u && fabsf(cosine.y) >= BZ_PLANE_EPSILON;
          float2 distance = make_float2(0.0f, 0.0f);
          if(pass_a) { distance.x = BZ_DIV(above.x, cosine.x); }
          if(pass_b) { distance.y = BZ_DIV(above.y, cosine.y); }
          pass_a = pass_a && distance.x > 0.0f && fabsf(distance.x) > -r3.z && fabsf(distance.x) > r5.z;     // hIn (word 7), hOut (word 11)
          pass_b = pass_b && distance.y > 0.0f && fabsf(distance.y) > -r3.w && fabsf(distance.y) > r5.w;
          if(__ballot_sync(0xffffffffu, pass_a || pass_b) != 0u) {
            const float2 px = __ffma2_rn(bzw_mul2(distance, d.x), make_float2(one, one), make_float2(s.x, s.x));               // s + distance * d
            const float2 py = __ffma2_rn(bzw_mul2(distance, d.y), make_float2(one, one), make_float2(s.y, s.y));
            const float2 pz = __ffma2_rn(bzw_mul2(distance, d.z), make_float2(one, one), make_float2(s.z, s.z));
#define BZW_ROW(x, y, z) bzw_add2(__fmul2_rn(x, px), bzw_add2(__fmul2_rn(y, py), __fmul2_rn(z, pz), one), one)             /* bz_dot(row, p) */
            const float2 b0 = BZW_ROW(make_float2(r2.x, r2.y), make_float2(r2.z, r2.w), make_float2(r3.x, r3.y));
            const float2 b1 = BZW_ROW(make_float2(r4.x, r4.y), make_float2(r4.z, r4.w), make_float2(r5.x, r5.y));
            const float2 b2 = BZW_ROW(make_float2(r6.x, r6.y), make_float2(r6.z, r6.w), make_float2(r7.x, r7.y));
#undef BZW_ROW
            pass_a = pass_a && b0.x >= 0.0f && b0.x <= 1.0f && b1.x >= 0.0f && b1.x <= 1.0f && b2.x >= 0.0f && b2.x <= 1.0f;
            pass_b = pass_b && b0.y >= 0.0f && b0.y <= 1.0f && b1.y >= 0.0f && b1.y <= 1.0f && b2.y >= 0.0f && b2.y <= 1.0f;
            const uint32_t passed_a = __ballot_sync(0xffffffffu, pass_a), passed_b = __ballot_sync(0xffffffffu, pass_b);
            if((passed_a | passed_b) != 0u) {
              if(pass_a) {
                const uint32_t at = level + __popc(passed_a & below);
                P.qb_patch[at] = __float_as_uint(r7.z);
                P.qb_loop[at] = __float_as_uint(r7.z);
                P.qb_slot[at] = (uint8_t)(slot | 0x80u);
              }
              level += __popc(passed_a);
              if(pass_b) {
                const uint32_t at = level + __popc(passed_b & below);
                P.qb_patch[at] = __float_as_uint(r7.w);
                P.qb_loop[at] = __float_as_uint(r7.w);
                P.qb_slot[at] = (uint8_t)(slot | 0x80u);
              }
              level += __popc(passed_b);
            }
          }
        }
        if(lane == 0u) { P.qb_n = level; }
        BZW_COUNT(BZC_STAGE_A, tested * __popc(__ballot_sync(0xffffffffu, live)));
        continue;
      }

      if(action == 1) {
        // ---- stage A on (up to) 32 pairs from the tail of queue A ----
        const uint32_t n = min(qa, 32u);
        const uint32_t q = qa - n + lane;
        if(lane == 0u) { P.qa_n = qa - n; }
        BZW_COUNT(BZC_STAGE_A, n);
        if(lane < n) {
          const uint32_t slot = P.qa_slot[q];
          bz_cull_rec c;
          uint32_t patch;
          bzw_unpack_cull(T.leaf + (size_t)P.qa_leaf[q] * BZ_CULL_F4, c, patch);
          const bzv3 s = bz_v(P.ray[0][slot], P.ray[1][slot], P.ray[2][slot]);
          const bzv3 d = bz_v(P.ray[3][slot], P.ray[4][slot], P.ray[5][slot]);
          float cosine, distance;
          if(bz_stage_a(c, s, d, true, cosine, distance)) {
            const uint32_t at = atomicAdd(&P.qb_n, 1u);
            P.qb_patch[at] = patch;
            P.qb_loop[at] = patch;
            P.qb_slot[at] = (uint8_t)(slot | 0x80u);
          }
        }
        continue;
      }

      // ---- action 2: stage B on (up to) 32 candidates from the tail of queue B ----
      {
        const uint32_t n = min(qb, 32u);
        const uint32_t q = qb - n + lane;
        if(lane == 0u) { P.qb_n = qb - n; }
        uint32_t slot = 0u, patch = 0u, loop = 0u;
        bool first_level = false;
        if(lane < n) {
          const uint32_t packed = P.qb_slot[q];
          slot = packed & 0x7fu;
          first_level = (packed & 0x80u) != 0u;
          patch = P.qb_patch[q];
          loop = P.qb_loop[q];
        }
        __syncwarp();                          // the entries are read before anyone appends a hop
        unsigned long long key = BZ_KEY_NONE;
        bz_hit hit;
        hit.what = BZ_MISS;
        bool evaluated = false;
        if(lane < n) {
          bz_cull_rec c;
          uint32_t index;
          bzw_unpack_cull(T.cull + (size_t)patch * BZ_CULL_F4, c, index);
          const bzv3 s = bz_v(P.ray[0][slot], P.ray[1][slot], P.ray[2][slot]);
          const bzv3 d = bz_v(P.ray[3][slot], P.ray[4][slot], P.ray[5][slot]);
          float cosine, distance;
          if(bz_stage_a(c, s, d, first_level, cosine, distance)) {
            bz_newton_rec t;
#if defined(BZ_FAST)
            bzw_load_newton_root(T.newton + (size_t)patch * BZ_NEWTON_F4, t);
            if(bz_stage_b_root(c, t, s, d, cosine, distance, hit)) {
#else
            bzw_newton2 t2;
            bzw_load_newton2(T.newton + (size_t)patch * BZ_NEWTON_F4, J.one, t2);
            if(bz_stage_b_root(c, t2, s, d, cosine, distance, hit)) {
#endif
              bzw_load_newton_side(T.newton + (size_t)patch * BZ_NEWTON_F4, t);
              bz_stage_b_side(t, d, hit);
            }
            evaluated = true;
            if(hit.what <= BZ_FOLLOW2) {
              if(first_level) {                // one hop, no recursion (bezierMesh.cpp:213-217)
                const uint32_t at = atomicAdd(&P.qb_n, 1u);
                P.qb_patch[at] = (hit.what == BZ_FOLLOW0) ? t.nb[0] : ((hit.what == BZ_FOLLOW1) ? t.nb[1] : t.nb[2]);     // (no dynamic index: keeps the record in registers)
                P.qb_loop[at] = loop;
                P.qb_slot[at] = (uint8_t)slot;
              }
            }
            else if(hit.what == BZ_HIT && hit.distance < 3.402823466e+38f) {      // bezierMesh.cpp:209, 220
              key = bzw_make_key(hit.distance, loop);
              atomicMin(&P.best[slot], key);
            }
          }
        }
        __syncwarp();
        BZW_COUNT(BZC_STAGE_A, __popc(__ballot_sync(0xffffffffu, lane < n && !first_level)));
        BZW_COUNT(BZC_STAGE_B, __popc(__ballot_sync(0xffffffffu, evaluated)));
        if(key != BZ_KEY_NONE && P.best[slot] == key) {
          P.pay[0][slot] = hit.point.x;  P.pay[1][slot] = hit.point.y;  P.pay[2][slot] = hit.point.z;
          P.pay[3][slot] = hit.normal.x; P.pay[4][slot] = hit.normal.y; P.pay[5][slot] = hit.normal.z;
          P.pay[6][slot] = hit.cos_incidence;
          P.pay[7][slot] = hit.bary.x;   P.pay[8][slot] = hit.bary.y;   P.pay[9][slot] = hit.bary.z;
          P.pay[10][slot] = __uint_as_float(patch);
        }
      }
    }
    if(!any_live) { break; }

    // ================= resolve: nearest hit -> Snell -> expected-side check -> next bounce or finish =================
    __syncwarp();
#pragma unroll 1
    for(int g = 0; g < BZ_GROUPS; ++g) {
      const uint32_t slot = (uint32_t)g * 32u + lane;
      const uint32_t st = P.state[slot];
      uint32_t events = 0u;                  // byte 0 Snell block reached, 1 entered, 2 exited, 3 lost
      if((st & 1u) != 0u) {
      uint32_t expected = (st >> 1) & 3u;
      uint32_t crossings = st >> 4;
      const unsigned long long key = P.best[slot];
      const bool found = (key != BZ_KEY_NONE);
      bzv3 s = bz_v(P.ray[0][slot], P.ray[1][slot], P.ray[2][slot]);
      bzv3 d = bz_v(P.ray[3][slot], P.ray[4][slot], P.ray[5][slot]);
      bz_hit hit;
      hit.point = bz_v(P.pay[0][slot], P.pay[1][slot], P.pay[2][slot]);
      hit.normal = bz_v(P.pay[3][slot], P.pay[4][slot], P.pay[5][slot]);
      hit.cos_incidence = P.pay[6][slot];
      hit.bary = bz_v(P.pay[7][slot], P.pay[8][slot], P.pay[9][slot]);
      const uint32_t patch = __float_as_uint(P.pay[10][slot]);
      bool live = true;
      if(JOB == BZ_JOB_INTERSECT) {
        const unsigned long long ray_id = ((unsigned long long)P.id_hi[slot] << 32) | P.id_lo[slot];
        J.out_status[ray_id] = found ? BZ_WHAT_INTERSECT : BZ_WHAT_NONE;
        if(J.out_patch != nullptr) { J.out_patch[ray_id] = found ? patch : BZ_NO_PATCH; }
        if(J.out_distance != nullptr) { J.out_distance[ray_id] = found ? bzw_key_distance(key) : 0.0f; }
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
          events |= 1u;
        }
        status = (status == expected) ? status : BZ_STATUS_NONE;
        if(status != BZ_STATUS_NONE) {
          s = hit.point;
          d = refracted;
          events |= (status == BZ_STATUS_INSIDE) ? 0x100u : 0x10000u;
        }
        if(JOB == BZ_JOB_REFRACT) {
          const unsigned long long ray_id = ((unsigned long long)P.id_hi[slot] << 32) | P.id_lo[slot];
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
            if(expected == BZ_STATUS_INSIDE) {      // the ray is in air: it reaches the background (RENDER) / the target (TARGET)
              value = (JOB == BZ_JOB_TARGET) ? bzw_land_on_target(J, s, d) : bzw_shade_background(J.scene, s, d);
            }
            else {                                  // entered and never left
              value = (JOB == BZ_JOB_TARGET) ? BZ_LAND_LOST : J.scene.value_lost;
              events |= 0x1000000u;
            }
            if(JOB == BZ_JOB_TARGET) {
              if(value < BZ_LAND_MISSED) { atomicAdd(J.bins + value, 1u); }
              if(J.out_landing != nullptr) { J.out_landing[((unsigned long long)P.id_hi[slot] << 32) | P.id_lo[slot]] = value; }
            }
            else {
              atomicAdd(J.accum + P.id_lo[slot], value);
            }
            live = false;
          }
          else {
            P.ray[0][slot] = s.x; P.ray[1][slot] = s.y; P.ray[2][slot] = s.z;
            P.ray[3][slot] = d.x; P.ray[4][slot] = d.y; P.ray[5][slot] = d.z;
          }
        }
      }
      P.state[slot] = live ? (1u | (expected << 1) | (crossings << 4)) : 0u;
      }
      if(JOB != BZ_JOB_INTERSECT) {
        // four event counters per lane in the bytes of one register, folded into the warp's counters every 255 turns
        tally += events;
        if(++tally_turns == 255u) {
          BZW_COUNT(BZC_SNELL, __reduce_add_sync(0xffffffffu, tally & 0xffu));
          BZW_COUNT(BZC_ENTER, __reduce_add_sync(0xffffffffu, (tally >> 8) & 0xffu));
          BZW_COUNT(BZC_EXIT, __reduce_add_sync(0xffffffffu, (tally >> 16) & 0xffu));
          BZW_COUNT(BZC_LOST, __reduce_add_sync(0xffffffffu, tally >> 24));
          tally = 0u;
          tally_turns = 0u;
        }
      }
    }
  }

  // ---- work counters: one atomic per warp and counter ----
  BZW_COUNT(BZC_SNELL, __reduce_add_sync(0xffffffffu, tally & 0xffu));
  BZW_COUNT(BZC_ENTER, __reduce_add_sync(0xffffffffu, (tally >> 8) & 0xffu));
  BZW_COUNT(BZC_EXIT, __reduce_add_sync(0xffffffffu, (tally >> 16) & 0xffu));
  BZW_COUNT(BZC_LOST, __reduce_add_sync(0xffffffffu, tally >> 24));
  __syncwarp();
  if(J.stats != nullptr && lane < BZ_NSTAT && P.count[lane] != 0ull) { atomicAdd(J.stats + lane, P.count[lane]); }
}

}  // namespace BZ_NS
using namespace BZ_NS;

static size_t bzw_shared_bytes() { return sizeof(bz_warp_pool) * BZ_WARPS; }

static cudaError_t bzw_configure() {
  static bool done[64] = {};                 // the attribute is per device
  int device = 0;
  cudaError_t e = cudaGetDevice(&device);
  if(e != cudaSuccess) { return e; }
  bool &configured = done[device & 63];
  if(configured) { return cudaSuccess; }
  e = cudaFuncSetAttribute(bz_warp_trace_kernel<BZ_JOB_RENDER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bzw_shared_bytes());
  if(e == cudaSuccess) { e = cudaFuncSetAttribute(bz_warp_trace_kernel<BZ_JOB_REFRACT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bzw_shared_bytes()); }
  if(e == cudaSuccess) { e = cudaFuncSetAttribute(bz_warp_trace_kernel<BZ_JOB_INTERSECT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bzw_shared_bytes()); }
  if(e == cudaSuccess) { e = cudaFuncSetAttribute(bz_warp_trace_kernel<BZ_JOB_TARGET>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bzw_shared_bytes()); }
  configured = (e == cudaSuccess);
  return e;
}

cudaError_t BZ_SYM(bz_launch_warp_trace)(int job, const bz_tables &tables, const bz_job &desc, int grid, cudaStream_t stream) {
  cudaError_t e = bzw_configure();
  if(e != cudaSuccess) { return e; }
  if(job == BZ_JOB_RENDER) { bz_warp_trace_kernel<BZ_JOB_RENDER><<<grid, 32 * BZ_WARPS, bzw_shared_bytes(), stream>>>(tables, desc); }
  else if(job == BZ_JOB_REFRACT) { bz_warp_trace_kernel<BZ_JOB_REFRACT><<<grid, 32 * BZ_WARPS, bzw_shared_bytes(), stream>>>(tables, desc); }
  else if(job == BZ_JOB_TARGET) { bz_warp_trace_kernel<BZ_JOB_TARGET><<<grid, 32 * BZ_WARPS, bzw_shared_bytes(), stream>>>(tables, desc); }
  else { bz_warp_trace_kernel<BZ_JOB_INTERSECT><<<grid, 32 * BZ_WARPS, bzw_shared_bytes(), stream>>>(tables, desc); }
  return cudaGetLastError();
}

// resident CTAs per SM of the render instantiation (the persistent grid is sized from it)
cudaError_t BZ_SYM(bz_warp_trace_occupancy)(int *out_ctas_per_sm, int *out_rays_per_cta) {
  *out_rays_per_cta = BZ_WARPS * BZ_SLOTS;
  cudaError_t e = bzw_configure();
  if(e != cudaSuccess) { return e; }
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(out_ctas_per_sm, bz_warp_trace_kernel<BZ_JOB_RENDER>, 32 * BZ_WARPS, bzw_shared_bytes());
}

#if !defined(BZ_FAST)      // the kernels below do none of stage A / B: one copy, in the parity build
// ---- sky tiles: the tiles of a RENDER job none of whose rays can meet the lens (csrc/bz_capi.cu: bz_plan_tiles) ----
// Every refract() of such a ray is cNone (one call each, as in the bounce loop), so the pixel is the sum of the background
// values of its samples: one CTA per tile, one thread per pixel, a plain store into the accumulator (no other kernel
// writes these pixels).  The same ray generation and shading functions as the trace kernel, the same bits.
__global__ void __launch_bounds__(1024) bz_sky_kernel(const bz_job J, const uint32_t *tiles) {
  const uint32_t tile = __ldg(tiles + blockIdx.x);
  const uint32_t tile_y = tile / J.tiles_x, tile_x = tile - tile_y * J.tiles_x;
  const uint32_t in_y = threadIdx.x / J.tile_side, in_x = threadIdx.x - in_y * J.tile_side;
  const uint32_t px = tile_x * J.tile_side + in_x, row = tile_y * J.tile_side + in_y;
  const bool valid = px < J.camera.width && row < J.rows;
  const uint32_t spp = J.camera.strata * J.camera.strata;
  if(valid) {
    const uint32_t py = bzw_frame_row(J, row);
    uint32_t sum = 0u;
#pragma unroll 1
    for(uint32_t sy = 0u; sy < J.camera.strata; ++sy) {
#pragma unroll 1
      for(uint32_t sx = 0u; sx < J.camera.strata; ++sx) {
        bzv3 s, d;
        bzw_primary_ray(J.camera, J.two_over_w, J.two_over_h, px, py, sx, sy, s, d);
        sum += bzw_shade_background(J.scene, s, d);
      }
    }
    J.accum[row * J.camera.width + px] = sum;
  }
  const unsigned long long rays = (unsigned long long)__syncthreads_count(valid) * spp;
  if(threadIdx.x == 0u && J.stats != nullptr && rays != 0ull) {
    atomicAdd(J.stats + BZC_PRIMARY, rays);
    atomicAdd(J.stats + BZC_CALLS, rays);
  }
}

cudaError_t bz_launch_sky(const bz_job &desc, const uint32_t *d_tiles, uint32_t count, cudaStream_t stream) {
  if(count == 0u) { return cudaSuccess; }
  bz_sky_kernel<<<count, desc.tile_side * desc.tile_side, 0, stream>>>(desc, d_tiles);
  return cudaGetLastError();
}

// ---- illumination mode: emitter rays that pass beside the lens ----
// Most rays of a hemisphere emitter never come near the lens.  A ray that misses the root box of the cull hierarchy grown
// by TWICE the lane walk's slack fails stage A of every patch (the hierarchy's guarantee, with a share to spare for the
// rounding of this test), so its only refract() is cNone: it lands on the target here, and its patch key becomes
// BZ_RAY_PASSED so that the grouping (csrc/bz_illum.cu) leaves it out of the rays the trace kernel gets.
__global__ void __launch_bounds__(256) bz_pass_kernel(const bz_tables T, const bz_job J, const float *start, const float *dir, uint32_t *patch, uint64_t count) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool passed = false;
  if(i < count) {
    const bzv3 s = bz_v(start[i * 3u], start[i * 3u + 1u], start[i * 3u + 2u]);
    const bzv3 d = bz_v(dir[i * 3u], dir[i * 3u + 1u], dir[i * 3u + 2u]);
    const float l1 = fabsf(s.x) + fabsf(s.y) + fabsf(s.z);
    // records outside the hierarchy are tested by every ray (always segment) or by rays from far away (wide segment)
    const bool direct = (T.always_count != 0u) || !(l1 <= T.origin_limit);
    bool misses = (T.node_count == 0u);
    if(T.node_count != 0u) {
      const float4 lo = __ldg(T.nodes), hi = __ldg(T.nodes + 1);
      const float slack = 2.0f * 3.0517578125e-05f * (l1 + 2.0f * T.extent);
      const float ix = (fabsf(d.x) > 1e-20f) ? 1.0f / d.x : copysignf(1e20f, d.x);
      const float iy = (fabsf(d.y) > 1e-20f) ? 1.0f / d.y : copysignf(1e20f, d.y);
      const float iz = (fabsf(d.z) > 1e-20f) ? 1.0f / d.z : copysignf(1e20f, d.z);
      const float x0 = (lo.x - slack - s.x) * ix, x1 = (hi.x + slack - s.x) * ix;
      const float y0 = (lo.y - slack - s.y) * iy, y1 = (hi.y + slack - s.y) * iy;
      const float z0 = (lo.z - slack - s.z) * iz, z1 = (hi.z + slack - s.z) * iz;
      const float tnear = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), 0.0f));
      const float tfar = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fmaxf(z0, z1));
      misses = !(tnear <= tfar * 1.000001f);      // (NaN: not a miss)
    }
    passed = misses && !direct;
    if(passed) {
      const uint32_t value = bzw_land_on_target(J, s, d);
      if(value < BZ_LAND_MISSED) { atomicAdd(J.bins + value, 1u); }
      patch[i] = BZ_RAY_PASSED;
    }
  }
  const unsigned long long rays = (unsigned long long)__syncthreads_count(passed);
  if(threadIdx.x == 0u && J.stats != nullptr && rays != 0ull) {
    atomicAdd(J.stats + BZC_PRIMARY, rays);
    atomicAdd(J.stats + BZC_CALLS, rays);
  }
}

cudaError_t bz_launch_pass(const bz_tables &tables, const bz_job &desc, const float *start, const float *dir, uint32_t *patch, uint64_t count, cudaStream_t stream) {
  if(count == 0u) { return cudaSuccess; }
  bz_pass_kernel<<<(unsigned)((count + 255u) / 256u), 256, 0, stream>>>(tables, desc, start, dir, patch, count);
  return cudaGetLastError();
}
#endif

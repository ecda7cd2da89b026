// bz_device.h -- layout of the patch tables in HBM and the job descriptors the kernels take.
//
// Tables (all 16-byte aligned float4 streams, built once per lens by bz_lens_create):
//   cull     4 float4 per patch, ORIGINAL patch order   stage-A record  (bezierTriangle.h:64,75,77-78)
//            [0] plane n.xyz, k   [1] Binv row0.xyz, hIn   [2] Binv row1.xyz, hOut   [3] Binv row2.xyz, patch index bits
//   newton  12 float4 per patch, ORIGINAL patch order   stage-B record  (bezierTriangle.h:65,68,73,80)
//            words 0..19 (x, y) of the 10 control points (aligned pairs for FMUL2 / FFMA2), 20..29 their z, 30..32 dirB,
//            33..44 divider planes (n.xyz,k) x3, 45..47 neighbours
//   leaf     4 float4 per leaf slot: the cull records again, in BVH leaf order, followed by the "wide"
//            segment (tested directly only by rays starting beyond origin_limit) and the "always"
//            segment (patches the cull cannot bound, tested for every ray)
//   packet   7 x 32 words per node of the same hierarchy collapsed to 32 children per node (one lane per child):
//            lo.x[32] lo.y[32] lo.z[32] hi.x[32] hi.y[32] hi.z[32] code[32]; code 0 = empty child, bit 31 = inner child
//            (low bits: node index), else first_slot << 3 | count
//   pairs    8 float4 per leaf slot j: the leaf records j and j + 1 interleaved word by word, so that the packet form of
//            stage A tests two records per trip with FMUL2 / FFMA2 (every float4 = two aligned (A, B) register pairs)
//   nodes    2 float4 per node of the binary cull hierarchy, depth-first order:
//            [0] box lo.xyz, skip link (index of the first node after this subtree, taken when the box is missed)
//            [1] box hi.xyz, leaf code (0: inner node, its first child is the next node; else first_slot << 3 | count)
#ifndef BZ_DEVICE_H
#define BZ_DEVICE_H

#include <stdint.h>
#include <cuda_runtime.h>
#include "bz_capi.h"

#define BZ_CULL_F4    4
#define BZ_NEWTON_F4  12
#define BZ_NODE_F4    2
#define BZ_LEAF_MAX   2     // patches per leaf of the cull hierarchy (host/bz_bvh.cpp builds to this)

struct bz_tables {
  const float4 *cull;
  const float4 *newton;
  const float4 *leaf;
  const float4 *pairs;     // 8 float4 per leaf slot j: records j and j + 1 of `leaf` interleaved word by word (packet stage A)
  const float4 *nodes;
  const float  *packet;    // the hierarchy again, 32 children per node, 7 x 32 words per node (host/bz_bvh.h)
  uint32_t packet_count;
  uint32_t patch_count;
  uint32_t node_count;
  uint32_t wide_first;     // slots [wide_first, always_first): copies of the ill-conditioned patches whose grown boxes are
                           // only valid for ray origins with |s|_1 <= origin_limit; farther rays test them directly
  uint32_t always_first;   // slot range [always_first, always_first + always_count) of `leaf`
  uint32_t always_count;
  float    origin_limit;
  float    ri;
  float    extent;         // max |coordinate| over the flat triangles: sizes the per-ray cull slack
};

#define BZ_JOB_RENDER     0
#define BZ_JOB_REFRACT    1
#define BZ_JOB_INTERSECT  2
#define BZ_JOB_TARGET     3     // bounce loop over a ray batch, then the target rectangle of the illumination mode

struct bz_job {
  // BZ_JOB_RENDER
  bz_camera camera;
  bz_scene  scene;
  uint32_t  row0;            // first image row of this call
  uint32_t  stripe_rows;     // 0: rows [row0,row0+rows) contiguous; else striped: local row r is image row
  uint32_t  stripe_rank;     //    ((r / stripe_rows) * stripe_ranks + stripe_rank) * stripe_rows + r % stripe_rows
  uint32_t  stripe_ranks;
  uint32_t *accum;           // one u32 per local pixel: sum of the sample values
  // BZ_JOB_REFRACT / BZ_JOB_INTERSECT
  const float    *in_start;
  const float    *in_dir;
  const uint32_t *in_expected;
  float    *out_start;
  float    *out_dir;
  uint32_t *out_status;      // refract: status; intersect: what
  uint32_t *out_patch;
  float    *out_distance;
  float    *out_point;
  float    *out_normal;
  float    *out_cos;
  float    *out_bary;
  // BZ_JOB_TARGET (rays from in_start / in_dir, bounce limit from scene.max_crossings)
  bz_target target;
  float     target_plane[4];   // unit normal of the target rectangle and its constant
  uint32_t *bins;              // bins_v x bins_u counters, accumulated
  uint32_t *out_landing;       // nullable: per ray the bin, BZ_LAND_LOST or BZ_LAND_MISSED
  // all
  uint64_t  ray_count;       // primary rays of this call
  const uint32_t *ray_count_dev;   // BZ_JOB_TARGET, nullable: the rays actually there (<= ray_count), known on the device only
  float     one;             // 1.0f, as a run-time value (csrc/bz_trace.cu: packed sums)
  // work distribution: the job is cut into tiles handed to CTAs; RENDER: square blocks of tile_side x tile_side pixels
  // (tile_side a power of two, pixels of a tile in Morton order, tile_rays = tile_side^2 * spp); batch jobs: tile_rays
  // consecutive rays
  uint32_t  rows;            // RENDER: image rows of this call (local rows of the band)
  uint32_t  tile_side, tile_rays, tiles_x, tiles_y;
  uint32_t  div_spp[2], div_strata[2], div_stripe[2];   // reciprocals for n / spp, n / strata, n / stripe_rows (bzw_div)
  float     two_over_w, two_over_h;                      // 2.0f / (float)width, 2.0f / (float)height
  // RENDER: tiles whose rays may meet the lens, in the order they are handed out (tile_y * tiles_x + tile_x), or null = every
  // tile of the band, rows from the outside in.  The other tiles of the band are shaded by bz_sky_kernel (csrc/bz_trace.cu).
  const uint32_t *tile_list;
  uint64_t  tile_count;
  unsigned long long *next_ray;   // tile counter, zeroed before launch
  unsigned long long *stats;      // 9 counters in bz_stats order, accumulated; nullable
};

// ---- wavefront pipeline of a RENDER job (csrc/bz_wave.cu) ----
// The tiles of the job that can meet the lens are cut into chunks; a chunk goes, once per surface crossing, through three
// kernels that talk through these buffers (sized per context, L2 / HBM resident):
//   bz_wave_walk_kernel     per packet of 32 rays: ray generation (first crossing) or reload, packet walk, stage A; the
//                           survivors become candidates (ray slot, patch) appended to `cand`, chained per ray through
//                           `head` / `cand_next`
//   bz_wave_eval_kernel     per candidate: stage B (and the one follow-side hop); result key + hit record per candidate
//   bz_wave_resolve_kernel  per ray: nearest result of its chain, Snell, next ray or shading
struct bz_wave {
  float4   *ray_a;          // [ray_cap]      start.xyz, dir.x
  float4   *ray_b;          // [ray_cap]      dir.y, dir.z, local pixel index (bits), state (bits): bit 0 live, 1-2 expected status, 4.. crossings
  uint32_t *head;           // [ray_cap]      1 + index of the newest candidate of the ray's current call, 0 = none
  uint2    *cand;           // [cand_cap]     (ray slot, patch)
  uint32_t *cand_next;      // [cand_cap]     1 + index of the ray's next candidate, 0 = end of the chain
  unsigned long long *cand_key;   // [cand_cap]   (distance key, loop index) of the candidate's hit, BZ_KEY_NONE without one
  float4   *cand_pay;       // [2 * cand_cap] point.xyz, cos incidence | normal.xyz, evaluated patch (bits)
  uint32_t *counters;       // [0] next packet of the walk  [1] candidates of this pass  [2] next batch of the evaluation
                            // [3] sticky: the candidate buffer overflowed (the frame is incomplete, the next call reports it)
  uint32_t  ray_cap, cand_cap;
  uint32_t  first_tile;     // chunk = tiles first_tile + i * tile_stride, i < tile_count, of the job's tile order (heaviest first:
  uint32_t  tile_stride;    // interleaved chunks carry equal shares of the heavy tiles)
  uint32_t  tile_count;
  uint32_t  packets_per_tile;
  uint32_t  split_tiles;    // the walk kernel hands the first split_tiles tiles of the chunk to its CTAs in parts of unit_packets
  uint32_t  unit_packets;   // consecutive packets (units_per_tile = ceil(packets_per_tile / unit_packets) parts per tile), the
  uint32_t  units_per_tile; // other tiles whole; unit_count = split_tiles * units_per_tile + (tile_count - split_tiles)
  uint32_t  unit_count;
  uint32_t  level;          // 0: the walk kernel generates the primary rays
};

#define BZ_RAY_PASSED 0xffffffffu     // illumination mode: patch key of an emitter ray that passed beside the lens (bz_pass_kernel)

// UniformHemisphere belt table (reference/hostUtil.cpp:8-13), passed to the emission kernel by value
#define BZ_MAX_BELTS 128
struct bz_belts {
  float    width[BZ_MAX_BELTS];     // patch width of the belt: 2 pi / patches of the belt
  uint32_t first[BZ_MAX_BELTS];     // patches before the belt
  float    belt_width;              // pi / 2 / belts
  uint32_t belts, patches;
};

#endif

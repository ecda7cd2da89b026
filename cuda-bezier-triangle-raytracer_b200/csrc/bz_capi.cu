// bz_capi.cu -- the C ABI of include/bz_capi.h: lens handles, table upload, kernel launches.
// No CPU path: every computing entry point needs a CUDA device and fails with BZ_ERR_NO_DEVICE /
// BZ_ERR_CUDA otherwise.
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <mutex>
#include <string>
#include <vector>
#include "bz_capi.h"
#include "bz_device.h"
#include "../host/bz_bvh.h"
#include "../host/bz_error.h"

cudaError_t bz_launch_trace(int job, bool brute, const bz_tables &tables, const bz_job &desc, int grid, cudaStream_t stream);
cudaError_t bz_launch_resolve(const uint32_t *accum, uint8_t *rgba, uint64_t pixels, uint32_t spp, cudaStream_t stream);
cudaError_t bz_launch_raygen(const bz_camera &cam, uint32_t row0, uint64_t count, float *start, float *dir, cudaStream_t stream);
cudaError_t bz_launch_assemble(const uint8_t *bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                               uint32_t ranks, uint8_t *image, cudaStream_t stream);
cudaError_t bz_launch_fma_probe(float *sink, uint32_t iters, int grid, cudaStream_t stream);
int bz_trace_threads(void);
cudaError_t bz_launch_warp_trace(int job, const bz_tables &tables, const bz_job &desc, int grid, cudaStream_t stream);
cudaError_t bz_launch_pass(const bz_tables &tables, const bz_job &desc, const float *start, const float *dir, uint32_t *patch, uint64_t count, cudaStream_t stream);
cudaError_t bz_launch_sky(const bz_job &desc, const uint32_t *d_tiles, uint32_t count, cudaStream_t stream);
cudaError_t bz_warp_trace_occupancy(int *out_ctas_per_sm, int *out_rays_per_cta);
cudaError_t bz_launch_warp_trace_fast(int job, const bz_tables &tables, const bz_job &desc, int grid, cudaStream_t stream);      // csrc/bz_trace.cu built with -DBZ_FAST
cudaError_t bz_warp_trace_occupancy_fast(int *out_ctas_per_sm, int *out_rays_per_cta);
cudaError_t bz_launch_wave_pass(const bz_tables &tables, const bz_job &desc, const bz_wave &wave, int sm_count, cudaStream_t stream);           // csrc/bz_wave.cu
cudaError_t bz_launch_wave_pass_fast(const bz_tables &tables, const bz_job &desc, const bz_wave &wave, int sm_count, cudaStream_t stream);
cudaError_t bz_launch_emit(const bz_emitter &em, const bz_belts &belts, uint64_t first, uint64_t count, float *start, float *dir, uint32_t *patch,
                           cudaStream_t stream);
cudaError_t bz_launch_group_by_patch(const uint32_t *patch, uint64_t count, uint32_t patches, uint32_t ways, uint32_t *counts,
                                     const float *start, const float *dir, float *out_start, float *out_dir, int sm_count, cudaStream_t stream);

static thread_local std::string gError;

// shared with host/bz_host_capi.cpp (declared in host/bz_error.h)
int bz_fail(int code, const std::string &text) {
  gError = text;
  return code;
}

#define BZ_CUDA(call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) { \
  return bz_fail(BZ_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while(0)

#ifndef BZ_WAVE_LANES
#define BZ_WAVE_LANES 2       // buffer sets / internal streams the chunks of a frame alternate between
#endif
#define BZ_WAVE_RAYS   (64u << 20)     // default chunk: every kernel boundary costs ~40 us of drained grid, so a frame is one chunk when it fits
#define BZ_WAVE_CANDS  3u              // candidates reserved per ray slot (the 4K headline frame needs 1.45)
#define BZ_WAVE_AUTO_RAYS (1ull << 19) // BZ_PIPELINE_AUTO: frames with at least this many ray slots on lens tiles (cfg2: 0.37 vs 0.41 ms; cfg1, below it: 0.16 vs 0.13) ...
#define BZ_WAVE_AUTO_CROSSINGS 4u      // ... and at most this many crossings (three kernel launches each) take the wavefront pipeline

struct bz_lens;

// Working set of the calls a lens gets on ONE stream: everything a call writes besides the caller's buffers.  The tables of
// the lens are read-only, so calls on different streams (each with its own context) may overlap -- BezierLens::refract is
// const and re-entrant in the reference (bezierLens.h:27).  A call holds `busy` while it enqueues its work.
struct bz_context {
  bz_lens     *lens;
  cudaStream_t stream;                // the stream this context serves; host contexts: a non-blocking stream of their own
  bool         owns_stream;
  std::mutex   busy;
  unsigned long long *d_counter;      // work counter of the persistent kernel
  uint32_t  *d_accum;                 // render scratch: one u32 per pixel of the largest band so far
  size_t     accum_pixels;
  void      *d_illum;                 // illumination mode: emitted and grouped rays
  size_t     illum_bytes;
  void      *d_scratch;               // *_host staging
  size_t     scratch_bytes;
  cudaEvent_t ev_begin, ev_end;
  bool       timed;
  // RENDER: which tiles of a frame can meet the lens (bz_plan_tiles); the plan of the last frame is kept
  struct {
    bool      valid;
    uint32_t  mode;
    bz_camera camera;
    uint32_t  row0, rows, stripe_rows, rank, ranks, side;
    uint32_t  lens_count, sky_count;  // d_tiles = lens tiles in hand-out order, then sky tiles
  } plan;
  uint32_t  *d_tiles;
  size_t     tiles_capacity;
  uint32_t  *h_tiles;                 // pinned staging of the tile list (the upload is a true asynchronous copy)
  size_t     h_tiles_capacity;
  cudaEvent_t ev_tiles;               // the last upload out of h_tiles has been consumed
  double     plan_ms;                 // host time of the most recent call's bz_plan_tiles + upload (0: the kept plan was used)
  uint32_t   launches;                // kernels of the most recent call
  // RENDER, wavefront pipeline (csrc/bz_wave.cu): two sets of ray slots + candidate list, each with a stream of its own -- the chunks of a
  // frame alternate between them, so that one chunk's kernels fill the drained grid at the other's kernel boundaries
  bz_wave    wave[BZ_WAVE_LANES];
  void      *d_wave;                  // one allocation behind the pointers of both sets
  size_t     wave_bytes;
  cudaStream_t lane[BZ_WAVE_LANES];
  cudaEvent_t ev_fork, ev_join[BZ_WAVE_LANES];
};

struct bz_lens {
  int        device;
  int        sm_count;
  uint32_t   mode;
  uint32_t   pipeline;                // BZ_PIPELINE_*
  uint32_t   wave_chunk;              // ray slots per chunk of the wavefront pipeline
  uint32_t   wave_unit;               // packets per hand-out unit of the walk kernel
  bz_tables  tables;
  float4    *d_cull, *d_newton, *d_leaf, *d_pairs, *d_nodes, *d_packet;
  uint32_t   bvh_depth;
  std::vector<float> outline;         // up to BZ_OUTLINE_BOXES boxes of the cull hierarchy that cover every patch: lo.xyz hi.xyz each
  std::mutex lock;                    // guards the two lists below
  std::vector<bz_context*> contexts;  // one per stream seen so far (bz_lens_release_stream drops one)
  std::vector<bz_context*> idle_host; // contexts of finished *_host calls, waiting for the next one
};

// the context of the calling thread's most recent call (bz_last_kernel_ms / bz_last_kernel_launches)
static thread_local bz_context *gLastContext = nullptr;

static void bz_outline_boxes(const bz_bvh &bvh, std::vector<float> &out);
static void bz_context_free(bz_context *ctx);
static int bz_wave_overflowed(bz_context *ctx, bool *out);

static int bz_check_device(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if(e != cudaSuccess || count == 0) {
    cudaGetLastError();
    return bz_fail(BZ_ERR_NO_DEVICE, "no usable CUDA device (there is no CPU path)");
  }
  if(device < 0 || device >= count) { return bz_fail(BZ_ERR_INVALID, "device index out of range"); }
  return BZ_OK;
}

// ---- contexts ----
static void bz_context_free(bz_context *ctx) {
  if(ctx == nullptr) { return; }
  cudaFree(ctx->d_counter); cudaFree(ctx->d_accum); cudaFree(ctx->d_scratch); cudaFree(ctx->d_illum); cudaFree(ctx->d_tiles); cudaFree(ctx->d_wave);
  if(ctx->h_tiles != nullptr) { cudaFreeHost(ctx->h_tiles); }
  if(ctx->ev_begin) { cudaEventDestroy(ctx->ev_begin); }
  if(ctx->ev_end) { cudaEventDestroy(ctx->ev_end); }
  if(ctx->ev_tiles) { cudaEventDestroy(ctx->ev_tiles); }
  if(ctx->ev_fork) { cudaEventDestroy(ctx->ev_fork); }
  for(int k = 0; k < BZ_WAVE_LANES; ++k) {
    if(ctx->ev_join[k]) { cudaEventDestroy(ctx->ev_join[k]); }
    if(ctx->lane[k] != nullptr) { cudaStreamDestroy(ctx->lane[k]); }
  }
  if(ctx->owns_stream && ctx->stream != nullptr) { cudaStreamDestroy(ctx->stream); }
  delete ctx;
}

static int bz_context_new(bz_lens *lens, cudaStream_t stream, bool own_stream, bz_context **out) {
  BZ_CUDA(cudaSetDevice(lens->device));
  bz_context *ctx = new bz_context();      // (value-initialised: every plain member is zero)
  ctx->lens = lens;
  ctx->stream = stream;
  cudaError_t e = cudaSuccess;
  if(own_stream) {
    e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
    ctx->owns_stream = (e == cudaSuccess);
  }
  if(e == cudaSuccess) { e = cudaMalloc((void**)&ctx->d_counter, 64u); }
  if(e == cudaSuccess) { e = cudaEventCreate(&ctx->ev_begin); }
  if(e == cudaSuccess) { e = cudaEventCreate(&ctx->ev_end); }
  if(e == cudaSuccess) { e = cudaEventCreateWithFlags(&ctx->ev_tiles, cudaEventDisableTiming); }
  if(e != cudaSuccess) {
    std::string const text = cudaGetErrorString(e);
    bz_context_free(ctx);
    return bz_fail(BZ_ERR_CUDA, "bz_context_new: " + text);
  }
  *out = ctx;
  return BZ_OK;
}

// the context that serves `stream` on this lens (made on first use)
static int bz_context_for(bz_lens *lens, cudaStream_t stream, bz_context **out) {
  std::lock_guard<std::mutex> hold(lens->lock);
  for(bz_context *ctx : lens->contexts) {
    if(ctx->stream == stream) { *out = ctx; return BZ_OK; }
  }
  int rc = bz_context_new(lens, stream, false, out);
  if(rc == BZ_OK) { lens->contexts.push_back(*out); }
  return rc;
}

// a *_host call: takes an idle host context (own non-blocking stream) or makes one, gives it back when it ends, so that
// host calls from any number of threads run side by side on one lens
struct bz_host_call {
  bz_lens *lens;
  bz_context *ctx;
  int rc;
  explicit bz_host_call(bz_lens *aLens) : lens(aLens), ctx(nullptr), rc(BZ_OK) {
    {
      std::lock_guard<std::mutex> hold(lens->lock);
      if(!lens->idle_host.empty()) { ctx = lens->idle_host.back(); lens->idle_host.pop_back(); }
    }
    if(ctx == nullptr) { rc = bz_context_new(lens, nullptr, true, &ctx); }
  }
  ~bz_host_call() {
    if(ctx != nullptr) {
      std::lock_guard<std::mutex> hold(lens->lock);
      lens->idle_host.push_back(ctx);
    }
  }
};

extern "C" {

const char *bz_last_error(void) { return gError.c_str(); }

int bz_device_count(int *out_count) {
  if(out_count == nullptr) { return bz_fail(BZ_ERR_INVALID, "out_count is null"); }
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if(e != cudaSuccess) { cudaGetLastError(); count = 0; }
  *out_count = count;
  return BZ_OK;
}

uint32_t bz_sizeof_patch(void) { return (uint32_t)sizeof(bz_patch); }
uint32_t bz_sizeof_camera(void) { return (uint32_t)sizeof(bz_camera); }
uint32_t bz_sizeof_scene(void) { return (uint32_t)sizeof(bz_scene); }
uint32_t bz_sizeof_stats(void) { return (uint32_t)sizeof(bz_stats); }

int bz_lens_destroy(bz_lens *lens) {
  if(lens == nullptr) { return BZ_OK; }
  cudaSetDevice(lens->device);
  for(bz_context *ctx : lens->contexts) { bz_context_free(ctx); }
  for(bz_context *ctx : lens->idle_host) { bz_context_free(ctx); }
  if(gLastContext != nullptr && gLastContext->lens == lens) { gLastContext = nullptr; }
  cudaFree(lens->d_cull); cudaFree(lens->d_newton); cudaFree(lens->d_leaf); cudaFree(lens->d_pairs); cudaFree(lens->d_nodes); cudaFree(lens->d_packet);
  delete lens;
  return BZ_OK;
}

static void bz_pack_cull(const bz_patch &p, uint32_t index, float *out) {
  out[0] = p.plane[0]; out[1] = p.plane[1]; out[2] = p.plane[2]; out[3] = p.plane[3];
  out[4] = p.bary_inv[0][0]; out[5] = p.bary_inv[0][1]; out[6] = p.bary_inv[0][2]; out[7] = p.height_inside;
  out[8] = p.bary_inv[1][0]; out[9] = p.bary_inv[1][1]; out[10] = p.bary_inv[1][2]; out[11] = p.height_outside;
  out[12] = p.bary_inv[2][0]; out[13] = p.bary_inv[2][1]; out[14] = p.bary_inv[2][2];
  std::memcpy(out + 15, &index, sizeof(index));
}

int bz_lens_create(const bz_patch *h_patches, uint32_t patch_count, float refractive_index, int device, bz_lens **out_lens) {
  if(out_lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "out_lens is null"); }
  *out_lens = nullptr;
  if(h_patches == nullptr && patch_count != 0u) { return bz_fail(BZ_ERR_INVALID, "h_patches is null"); }
  if(patch_count >= (1u << 28)) { return bz_fail(BZ_ERR_INVALID, "too many patches"); }
  // the follow hop indexes the tables with the neighbour indices (bezierMesh.cpp:213-217): a bad one would be a wild device read
  for(uint32_t i = 0u; i < patch_count; ++i) {
    for(int k = 0; k < 3; ++k) {
      if(h_patches[i].neighbours[k] >= patch_count) { return bz_fail(BZ_ERR_INVALID, "patch " + std::to_string(i) + ": neighbour index out of range"); }
    }
  }
  int rc = bz_check_device(device);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaSetDevice(device));

  bz_bvh bvh;
  bz_build_bvh(h_patches, patch_count, bvh);
  static_assert(BZ_LEAF_MAX <= 7, "leaf codes keep the count in 3 bits");
  std::vector<float> cull((size_t)patch_count * 16u), newton((size_t)patch_count * 48u);
  for(uint32_t i = 0u; i < patch_count; ++i) {
    const bz_patch &p = h_patches[i];
    bz_pack_cull(p, i, cull.data() + (size_t)i * 16u);
    float *w = newton.data() + (size_t)i * 48u;
    for(int k = 0; k < 10; ++k) { w[2 * k] = p.control[k][0]; w[2 * k + 1] = p.control[k][1]; w[20 + k] = p.control[k][2]; }    // (x, y) pairs on even words: the packed FP32 pipe reads them as aligned register pairs
    std::memcpy(w + 30, p.dir_b, 3u * sizeof(float));
    std::memcpy(w + 33, p.divider, 12u * sizeof(float));
    std::memcpy(w + 45, p.neighbours, 3u * sizeof(uint32_t));
  }
  std::vector<uint32_t> slot2patch(bvh.order);
  slot2patch.insert(slot2patch.end(), bvh.wide.begin(), bvh.wide.end());
  slot2patch.insert(slot2patch.end(), bvh.always.begin(), bvh.always.end());
  std::vector<float> leaf(slot2patch.size() * 16u);
  for(size_t k = 0u; k < slot2patch.size(); ++k) { bz_pack_cull(h_patches[slot2patch[k]], slot2patch[k], leaf.data() + k * 16u); }

  bz_lens *lens = new bz_lens();      // (value-initialised: every plain member is zero)
  lens->device = device;
  lens->mode = BZ_MODE_CULLED;
  lens->wave_chunk = BZ_WAVE_RAYS;
  lens->wave_unit = 8u;                   // (one packet per warp of a walk CTA)
  lens->bvh_depth = bvh.depth;
  bz_outline_boxes(bvh, lens->outline);
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if(e != cudaSuccess) { delete lens; return bz_fail(BZ_ERR_CUDA, cudaGetErrorString(e)); }
  lens->sm_count = prop.multiProcessorCount;
  // every table gets 256 zeroed bytes of tail padding: a record fetched one past the end (a load the compiler moved
  // above its loop test) stays inside the allocation and decodes as a patch no ray can meet (zero normal)
  auto upload = [&](const std::vector<float> &src, float4 **dst) -> cudaError_t {
    size_t bytes = src.size() * sizeof(float) + 256u;
    cudaError_t r = cudaMalloc((void**)dst, bytes);
    if(r != cudaSuccess) { return r; }
    r = cudaMemset(*dst, 0, bytes);
    if(r != cudaSuccess) { return r; }
    if(!src.empty()) { r = cudaMemcpy(*dst, src.data(), src.size() * sizeof(float), cudaMemcpyHostToDevice); }
    return r;
  };
  e = upload(cull, &lens->d_cull);
  if(e == cudaSuccess) { e = upload(newton, &lens->d_newton); }
  if(e == cudaSuccess) { e = upload(leaf, &lens->d_leaf); }
  {
    // slot j of `pairs` = records j and j + 1 of `leaf`, word by word (the last slot pairs with a record no ray can meet)
    std::vector<float> pairs(slot2patch.size() * 32u, 0.0f);
    for(size_t k = 0u; k < slot2patch.size(); ++k) {
      for(size_t w = 0u; w < 16u; ++w) {
        pairs[k * 32u + 2u * w] = leaf[k * 16u + w];
        pairs[k * 32u + 2u * w + 1u] = (k + 1u < slot2patch.size()) ? leaf[(k + 1u) * 16u + w] : 0.0f;
      }
    }
    if(e == cudaSuccess) { e = upload(pairs, &lens->d_pairs); }
  }
  if(e == cudaSuccess) { e = upload(bvh.nodes, &lens->d_nodes); }
  if(e == cudaSuccess) { e = upload(bvh.packet_nodes, &lens->d_packet); }
  if(e != cudaSuccess) {
    std::string text = cudaGetErrorString(e);
    bz_lens_destroy(lens);
    return bz_fail(BZ_ERR_CUDA, "bz_lens_create: " + text);
  }
  lens->tables.cull = lens->d_cull;
  lens->tables.newton = lens->d_newton;
  lens->tables.leaf = lens->d_leaf;
  lens->tables.pairs = lens->d_pairs;
  lens->tables.nodes = lens->d_nodes;
  lens->tables.packet = (const float*)lens->d_packet;
  lens->tables.packet_count = (uint32_t)(bvh.packet_nodes.size() / 224u);
  lens->tables.patch_count = patch_count;
  lens->tables.node_count = (uint32_t)(bvh.nodes.size() / 8u);
  lens->tables.wide_first = (uint32_t)bvh.order.size();
  lens->tables.always_first = (uint32_t)(bvh.order.size() + bvh.wide.size());
  lens->tables.origin_limit = bvh.origin_limit;
  lens->tables.always_count = (uint32_t)bvh.always.size();
  lens->tables.ri = refractive_index;
  lens->tables.extent = bvh.extent;
  *out_lens = lens;
  return BZ_OK;
}

int bz_lens_set_mode(bz_lens *lens, uint32_t mode) {
  if(lens == nullptr || mode > BZ_MODE_FAST) { return bz_fail(BZ_ERR_INVALID, "bad lens or mode"); }
  lens->mode = mode;
  return BZ_OK;
}

int bz_lens_set_pipeline(bz_lens *lens, uint32_t pipeline) {
  if(lens == nullptr || pipeline > BZ_PIPELINE_WAVEFRONT) { return bz_fail(BZ_ERR_INVALID, "bad lens or pipeline"); }
  lens->pipeline = pipeline;
  return BZ_OK;
}

int bz_lens_set_wave_chunk(bz_lens *lens, uint32_t ray_slots) {
  if(lens == nullptr || ray_slots < 1024u || ray_slots > (1u << 26)) { return bz_fail(BZ_ERR_INVALID, "bad lens or chunk size (1024 .. 2^26 ray slots)"); }
  lens->wave_chunk = ray_slots;
  return BZ_OK;
}

int bz_lens_check(bz_lens *lens) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  BZ_CUDA(cudaSetDevice(lens->device));
  std::vector<bz_context*> all;
  {
    std::lock_guard<std::mutex> hold(lens->lock);
    all = lens->contexts;
    all.insert(all.end(), lens->idle_host.begin(), lens->idle_host.end());
  }
  bool overflowed = false;
  for(bz_context *ctx : all) {
    std::lock_guard<std::mutex> hold(ctx->busy);
    bool here = false;
    int rc = bz_wave_overflowed(ctx, &here);
    if(rc != BZ_OK) { return rc; }
    overflowed = overflowed || here;
  }
  if(overflowed) {
    return bz_fail(BZ_ERR_OVERFLOW, "candidate list of the wavefront pipeline overflowed (more than 3 stage-A survivors per ray over a whole chunk): the last frames are incomplete; "
                                    "render with bz_lens_set_pipeline(lens, BZ_PIPELINE_PERSISTENT)");
  }
  return BZ_OK;
}

int bz_lens_release_stream(bz_lens *lens, void *stream) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  bz_context *found = nullptr;
  {
    std::lock_guard<std::mutex> hold(lens->lock);
    for(size_t i = 0u; i < lens->contexts.size(); ++i) {
      if(lens->contexts[i]->stream == (cudaStream_t)stream) { found = lens->contexts[i]; lens->contexts.erase(lens->contexts.begin() + (long)i); break; }
    }
  }
  if(found == nullptr) { return BZ_OK; }
  BZ_CUDA(cudaSetDevice(lens->device));
  BZ_CUDA(cudaStreamSynchronize(found->stream));
  if(gLastContext == found) { gLastContext = nullptr; }
  bz_context_free(found);
  return BZ_OK;
}

int bz_lens_patch_count(const bz_lens *lens, uint32_t *out_count) {
  if(lens == nullptr || out_count == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  *out_count = lens->tables.patch_count;
  return BZ_OK;
}

int bz_lens_info(const bz_lens *lens, uint32_t out[4]) {
  if(lens == nullptr || out == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  out[0] = lens->tables.node_count;
  out[1] = lens->bvh_depth;
  out[2] = lens->tables.always_count;
  out[3] = (uint32_t)lens->sm_count;
  return BZ_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ launches ----

// ---- RENDER: which tiles of the frame can meet the lens ----
// A ray that misses a box of the cull hierarchy (grown by the walk's slack) fails stage A of every patch below it -- that
// is the hierarchy's own guarantee.  bz_plan_tiles projects a few hundred boxes that together cover every patch (the
// "outline", picked at bz_lens_create) onto the image; a tile no projected box touches has no ray that can meet the lens:
// every refract() of its rays is cNone and bz_sky_kernel shades it, the trace kernel gets the other tiles only.
//   pinhole: ray = eye + t M (a, b, 1), M = [right | up | forward], t > 0: it meets a box only if (a, b) lies in the
//            projection (x / z, y / z) of Q = M^-1 (box - eye), a convex body -- within the bounding rectangle of its 8
//            projected corners when all of them have z > 0; a box with all z <= 0 lies behind every ray;
//   orthographic: ray = eye + M (a, b, t): (a, b) within the bounding rectangle of the (x, y) of Q's corners.
// The boxes are grown by twice the packet walk's own margin, the rectangles by two whole pixels (the float rounding of
// (a, b) and of the normalised direction is below 1e-2 pixel).  Anything unusual -> no plan: every tile is traced.
#define BZ_OUTLINE_BOXES 1024u

// boxes of the hierarchy covering every patch: the root, repeatedly replacing the largest inner box by its two children
static void bz_outline_boxes(const bz_bvh &bvh, std::vector<float> &out) {
  out.clear();
  size_t const count = bvh.nodes.size() / 8u;
  if(count == 0u) { return; }
  auto word = [&](size_t node, int k) { uint32_t w; std::memcpy(&w, &bvh.nodes[node * 8u + k], sizeof(w)); return w; };
  auto area = [&](size_t node) {
    const float *p = &bvh.nodes[node * 8u];
    double const dx = (double)p[4] - p[0], dy = (double)p[5] - p[1], dz = (double)p[6] - p[2];
    return dx * dy + dy * dz + dz * dx;
  };
  std::vector<std::pair<double, size_t>> heap;      // (area, node) of the inner boxes still splittable
  std::vector<size_t> kept;
  if(word(0u, 7) == 0u) { heap.push_back({ area(0u), 0u }); } else { kept.push_back(0u); }
  while(!heap.empty() && kept.size() + heap.size() < BZ_OUTLINE_BOXES) {
    std::pop_heap(heap.begin(), heap.end());
    size_t const node = heap.back().second;
    heap.pop_back();
    size_t const left = node + 1u, right = word(left, 3);      // depth-first order: first child follows, its skip link is the sibling
    for(size_t child : { left, right }) {
      if(child >= count) { out.clear(); return; }              // (cannot happen for a table bz_build_bvh made)
      if(word(child, 7) == 0u) { heap.push_back({ area(child), child }); std::push_heap(heap.begin(), heap.end()); }
      else { kept.push_back(child); }
    }
  }
  for(auto const &entry : heap) { kept.push_back(entry.second); }
  for(size_t node : kept) {
    const float *p = &bvh.nodes[node * 8u];
    out.insert(out.end(), { p[0], p[1], p[2], p[4], p[5], p[6] });
  }
}

// -> false: no plan (trace every tile).  true: lens_tiles (hand-out order: rows from the outside in) and sky_tiles.
static bool bz_plan_tiles(const bz_lens *lens, const bz_job &desc, std::vector<uint32_t> &lens_tiles, std::vector<uint32_t> &sky_tiles) {
  bz_camera const &cam = desc.camera;
  uint32_t const side = desc.tile_side;
  if((lens->mode != BZ_MODE_CULLED && lens->mode != BZ_MODE_FAST) || lens->outline.empty() || lens->tables.always_count != 0u || side * side > 1024u) { return false; }
  double const m[3][3] = { { cam.right[0], cam.up[0], cam.forward[0] }, { cam.right[1], cam.up[1], cam.forward[1] }, { cam.right[2], cam.up[2], cam.forward[2] } };
  double const det = m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) + m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
  double scale = 0.0;
  for(int r = 0; r < 3; ++r) { for(int c = 0; c < 3; ++c) { scale = std::max(scale, std::fabs(m[r][c])); } }
  if(!(std::fabs(det) > 1e-9 * scale * scale * scale) || !(std::fabs((double)cam.half_w) > 0.0) || !(std::fabs((double)cam.half_h) > 0.0)) { return false; }
  double inv[3][3];
  for(int r = 0; r < 3; ++r) {
    for(int c = 0; c < 3; ++c) {
      int const r1 = (c + 1) % 3, r2 = (c + 2) % 3, c1 = (r + 1) % 3, c2 = (r + 2) % 3;
      inv[r][c] = (m[r1][c1] * m[r2][c2] - m[r1][c2] * m[r2][c1]) / det;             // cofactor of m[c][r]
    }
  }
  // the largest |origin|_1 of a ray of this camera; beyond origin_limit the wide segment is tested directly (no plan)
  double origin = std::fabs((double)cam.eye[0]) + std::fabs((double)cam.eye[1]) + std::fabs((double)cam.eye[2]);
  if(cam.ortho != 0u) {
    for(int k = 0; k < 3; ++k) { origin += std::fabs((double)cam.half_w * cam.right[k]) + std::fabs((double)cam.half_h * cam.up[k]); }
  }
  if(!(origin * 1.001 <= (double)lens->tables.origin_limit)) { return false; }
  double const grow = 4.0 * 3.0517578125e-05 * (origin + 2.0 * (double)lens->tables.extent);
  // cells of side x side pixels of the WHOLE frame a projected box touches
  uint32_t const cells_x = (cam.width + side - 1u) / side, cells_y = (cam.height + side - 1u) / side;
  std::vector<uint16_t> touched((size_t)cells_x * cells_y, 0u);     // how many outline boxes project onto the cell (saturating)
  for(size_t box = 0u; box < lens->outline.size(); box += 6u) {
    const float *lo = &lens->outline[box], *hi = lo + 3;
    double lo_u = 1e300, hi_u = -1e300, lo_v = 1e300, hi_v = -1e300;
    int behind = 0;
    for(int corner = 0; corner < 8; ++corner) {
      double p[3];
      for(int k = 0; k < 3; ++k) { p[k] = (((corner >> k) & 1) != 0 ? (double)hi[k] + grow : (double)lo[k] - grow) - (double)cam.eye[k]; }
      double const x = inv[0][0] * p[0] + inv[0][1] * p[1] + inv[0][2] * p[2];
      double const y = inv[1][0] * p[0] + inv[1][1] * p[1] + inv[1][2] * p[2];
      double const z = inv[2][0] * p[0] + inv[2][1] * p[1] + inv[2][2] * p[2];
      double a = x, b = y;
      if(cam.ortho == 0u) {
        if(!(z > 1e-6 * (std::fabs(x) + std::fabs(y) + std::fabs(z)))) { ++behind; continue; }
        a = x / z; b = y / z;
      }
      double const fu = (a / (double)cam.half_w + 1.0) * 0.5 * (double)cam.width;    // csrc/bz_trace.cu: bzw_primary_ray
      double const fv = (1.0 - b / (double)cam.half_h) * 0.5 * (double)cam.height;
      if(!std::isfinite(fu) || !std::isfinite(fv)) { return false; }
      lo_u = std::min(lo_u, fu); hi_u = std::max(hi_u, fu); lo_v = std::min(lo_v, fv); hi_v = std::max(hi_v, fv);
    }
    if(behind == 8) { continue; }                       // behind every ray
    if(behind != 0) { return false; }                   // the box reaches beside or behind the eye
    double const u0 = std::floor(lo_u) - 2.0, u1 = std::floor(hi_u) + 2.0, v0 = std::floor(lo_v) - 2.0, v1 = std::floor(hi_v) + 2.0;
    if(u1 < 0.0 || v1 < 0.0 || u0 > (double)cam.width - 1.0 || v0 > (double)cam.height - 1.0) { continue; }      // off the image
    uint32_t const cx0 = (uint32_t)std::max(u0, 0.0) / side, cx1 = (uint32_t)std::min(u1, (double)cam.width - 1.0) / side;
    uint32_t const cy0 = (uint32_t)std::max(v0, 0.0) / side, cy1 = (uint32_t)std::min(v1, (double)cam.height - 1.0) / side;
    for(uint32_t cy = cy0; cy <= cy1; ++cy) {
      uint16_t *cell = &touched[(size_t)cy * cells_x];
      for(uint32_t cx = cx0; cx <= cx1; ++cx) { cell[cx] = (uint16_t)std::min<uint32_t>(cell[cx] + 1u, 0xffffu); }
    }
  }
  // tiles of the band: a tile's pixel columns are one cell column; its rows of the whole frame span one or more cell rows
  auto frame_row = [&](uint32_t row) {                  // csrc/bz_trace.cu: bzw_frame_row
    if(desc.stripe_rows == 0u) { return desc.row0 + row; }
    uint32_t const stripe = row / desc.stripe_rows;
    return (stripe * desc.stripe_ranks + desc.stripe_rank) * desc.stripe_rows + (row - stripe * desc.stripe_rows);
  };
  lens_tiles.clear();
  sky_tiles.clear();
  lens_tiles.reserve((size_t)desc.tiles_x * desc.tiles_y);
  sky_tiles.reserve((size_t)desc.tiles_x * desc.tiles_y);
  std::vector<uint32_t> weight;                         // per lens tile: outline boxes that project onto it
  weight.reserve((size_t)desc.tiles_x * desc.tiles_y);
  for(uint32_t turn = 0u; turn < desc.tiles_y; ++turn) {
    uint32_t const ty = ((turn & 1u) == 0u) ? (turn >> 1) : (desc.tiles_y - 1u - (turn >> 1));     // rows from the outside in
    uint32_t const row_a = ty * side, row_b = std::min(row_a + side, desc.rows) - 1u;
    uint32_t const cy0 = std::min(frame_row(row_a), cam.height - 1u) / side, cy1 = std::min(frame_row(row_b), cam.height - 1u) / side;
    for(uint32_t tx = 0u; tx < desc.tiles_x; ++tx) {
      uint32_t boxes = 0u;
      for(uint32_t cy = cy0; cy <= cy1; ++cy) { boxes = std::max<uint32_t>(boxes, touched[(size_t)cy * cells_x + tx]); }
      if(boxes != 0u) { lens_tiles.push_back(ty * desc.tiles_x + tx); weight.push_back(boxes); }
      else { sky_tiles.push_back(ty * desc.tiles_x + tx); }
    }
  }
  // the tiles many boxes project onto (silhouette of the body: thick projections, grazing rays, the costliest packets of the
  // frame) are handed out first, so the persistent grid ends on cheap tiles; within a weight the outside-in order stays
  {
    // (a stable counting sort by descending weight: the plan is rebuilt for every frame of a moving camera)
    uint32_t heaviest = 0u;
    for(uint32_t w : weight) { heaviest = std::max(heaviest, w); }
    std::vector<uint32_t> first(heaviest + 2u, 0u);
    for(uint32_t w : weight) { ++first[heaviest - w + 1u]; }
    for(uint32_t k = 1u; k < first.size(); ++k) { first[k] += first[k - 1u]; }
    std::vector<uint32_t> sorted(lens_tiles.size());
    for(size_t i = 0u; i < lens_tiles.size(); ++i) { sorted[first[heaviest - weight[i]]++] = lens_tiles[i]; }
    lens_tiles.swap(sorted);
  }
  return true;
}

// ---- RENDER through the wavefront pipeline (csrc/bz_wave.cu) ----
// The tiles of the job (desc.tile_count of them, in hand-out order) are cut into chunks of at most BZ_WAVE_RAYS ray slots; a
// chunk goes through walk + stage A, stage B and resolve once per surface crossing.  The buffers belong to the handle and are
// sized for the largest chunk so far: 36 bytes per ray slot, 52 bytes per candidate, BZ_WAVE_CANDS candidates per ray slot.

static int bz_wave_reserve(bz_context *ctx, uint32_t ray_cap) {
  uint64_t const cand_cap = std::max<uint64_t>((uint64_t)ray_cap * BZ_WAVE_CANDS, 1u << 16);
  auto round = [](uint64_t bytes) { return (bytes + 255u) / 256u * 256u; };
  uint64_t const sizes[8] = { round((uint64_t)ray_cap * 16u), round((uint64_t)ray_cap * 16u), round((uint64_t)ray_cap * 4u), round(cand_cap * 8u), round(cand_cap * 4u),
                              round(cand_cap * 8u), round(cand_cap * 32u), 256u };
  uint64_t one_set = 0u;
  for(uint64_t size : sizes) { one_set += size; }
  if(ctx->lane[0] == nullptr) {
    for(int k = 0; k < BZ_WAVE_LANES; ++k) {
      BZ_CUDA(cudaStreamCreateWithFlags(&ctx->lane[k], cudaStreamNonBlocking));
      BZ_CUDA(cudaEventCreateWithFlags(&ctx->ev_join[k], cudaEventDisableTiming));
    }
    BZ_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
  }
  if(ctx->d_wave == nullptr || ctx->wave_bytes < (uint64_t)BZ_WAVE_LANES * one_set || ctx->wave[0].ray_cap < ray_cap) {
    if(ctx->d_wave != nullptr) {
      for(int k = 0; k < BZ_WAVE_LANES; ++k) { BZ_CUDA(cudaStreamSynchronize(ctx->lane[k])); }
      BZ_CUDA(cudaFree(ctx->d_wave));
      ctx->d_wave = nullptr;
      ctx->wave_bytes = 0u;
    }
    BZ_CUDA(cudaMalloc(&ctx->d_wave, (uint64_t)BZ_WAVE_LANES * one_set));
    BZ_CUDA(cudaMemset(ctx->d_wave, 0, (uint64_t)BZ_WAVE_LANES * one_set));
    ctx->wave_bytes = (uint64_t)BZ_WAVE_LANES * one_set;
    for(int k = 0; k < BZ_WAVE_LANES; ++k) {
      char *at = (char*)ctx->d_wave + (size_t)k * one_set;
      bz_wave &wave = ctx->wave[k];
      wave.ray_a = (float4*)at; at += sizes[0];
      wave.ray_b = (float4*)at; at += sizes[1];
      wave.head = (uint32_t*)at; at += sizes[2];
      wave.cand = (uint2*)at; at += sizes[3];
      wave.cand_next = (uint32_t*)at; at += sizes[4];
      wave.cand_key = (unsigned long long*)at; at += sizes[5];
      wave.cand_pay = (float4*)at; at += sizes[6];
      wave.counters = (uint32_t*)at;
      wave.ray_cap = ray_cap;
      wave.cand_cap = (uint32_t)std::min<uint64_t>(cand_cap, 0xfffffff0ull);
    }
  }
  return BZ_OK;
}

// device memory the wavefront pipeline takes for a job of `slots` ray slots
static uint64_t bz_wave_bytes(const bz_lens *lens, uint64_t slots) {
  uint64_t const chunks = std::max<uint64_t>(BZ_WAVE_LANES, (slots + lens->wave_chunk - 1u) / lens->wave_chunk);
  return (uint64_t)BZ_WAVE_LANES * ((slots + chunks - 1u) / chunks) * (36u + 52u * BZ_WAVE_CANDS);
}

// sky_tiles / sky_count: the tiles beside the lens; bz_sky_kernel is launched on the call's stream AFTER the lanes have their work, so it
// runs beside the wavefront kernels and fills the SMs they leave idle at their boundaries (nobody else writes those pixels)
static int bz_run_wave(bz_context *ctx, const bz_job &desc, const uint32_t *sky_tiles, uint32_t sky_count) {
  bz_lens *lens = ctx->lens;
  cudaStream_t const stream = ctx->stream;
  uint32_t const packets_per_tile = (desc.tile_rays + 31u) / 32u;
  uint32_t const tile_slots = packets_per_tile * 32u;
  uint32_t const tiles = (uint32_t)desc.tile_count;
  // at least two chunks (one per lane), more when a chunk would exceed the configured size
  uint32_t const chunks = std::max<uint32_t>(std::min<uint32_t>(tiles, BZ_WAVE_LANES), (uint32_t)(((uint64_t)tiles * tile_slots + lens->wave_chunk - 1u) / lens->wave_chunk));
  uint32_t const chunk_tiles = (tiles + chunks - 1u) / chunks;
  int rc = bz_wave_reserve(ctx, chunk_tiles * tile_slots);
  if(rc != BZ_OK) { return rc; }
  uint32_t const levels = std::max<uint32_t>(1u, desc.scene.max_crossings);
  BZ_CUDA(cudaEventRecord(ctx->ev_fork, stream));
  for(int k = 0; k < BZ_WAVE_LANES; ++k) {
    BZ_CUDA(cudaStreamWaitEvent(ctx->lane[k], ctx->ev_fork, 0));
    BZ_CUDA(cudaMemsetAsync(ctx->wave[k].counters, 0, 3u * sizeof(uint32_t), ctx->lane[k]));      // (the kernels leave them zero; a failed launch may not)
  }
  // chunk c takes the tiles c, c + chunks, c + 2 chunks, ... of the hand-out order (heaviest first), so all chunks weigh the same
  for(uint32_t turn = 0u; turn < chunks && turn < tiles; ++turn) {
    int const k = (int)(turn % BZ_WAVE_LANES);
    bz_wave wave = ctx->wave[k];
    wave.first_tile = turn;
    wave.tile_stride = chunks;
    wave.tile_count = (tiles - turn + chunks - 1u) / chunks;
    wave.packets_per_tile = packets_per_tile;
    wave.unit_packets = std::min(packets_per_tile, lens->wave_unit);
    wave.units_per_tile = (packets_per_tile + wave.unit_packets - 1u) / wave.unit_packets;
    // the heaviest sixteenth of the chunk's tiles, and at least eight tiles per SM, go out in parts (measured, cfg4: the walk launches of the
    // two ranks of eight that own the poles of the body 2.30 -> 2.13 ms per frame share, the whole frame on one GPU +0.3 %; all tiles in
    // parts: 2.12 ms and +3.4 % -- the parts of a tile no longer meet in one SM's L1)
    wave.split_tiles = std::min(wave.tile_count, std::max((wave.tile_count + 15u) / 16u, 8u * (uint32_t)lens->sm_count));
    wave.unit_count = wave.split_tiles * wave.units_per_tile + (wave.tile_count - wave.split_tiles);
    for(uint32_t level = 0u; level < levels; ++level) {
      wave.level = level;
      if(lens->mode == BZ_MODE_FAST) { BZ_CUDA(bz_launch_wave_pass_fast(lens->tables, desc, wave, lens->sm_count, ctx->lane[k])); }
      else { BZ_CUDA(bz_launch_wave_pass(lens->tables, desc, wave, lens->sm_count, ctx->lane[k])); }
      ctx->launches += 3u;
    }
  }
  if(sky_count != 0u) { BZ_CUDA(bz_launch_sky(desc, sky_tiles, sky_count, stream)); ++ctx->launches; }
  for(int k = 0; k < BZ_WAVE_LANES; ++k) {
    BZ_CUDA(cudaEventRecord(ctx->ev_join[k], ctx->lane[k]));
    BZ_CUDA(cudaStreamWaitEvent(stream, ctx->ev_join[k], 0));
  }
  return BZ_OK;
}

// the sticky overflow flags of the context's two candidate lists -> out (synchronises `stream` with the host); cleared when set
static int bz_wave_overflowed(bz_context *ctx, bool *out) {
  *out = false;
  if(ctx->d_wave == nullptr) { return BZ_OK; }
  uint32_t flags[BZ_WAVE_LANES] = {};
  for(int k = 0; k < BZ_WAVE_LANES; ++k) { BZ_CUDA(cudaMemcpyAsync(&flags[k], ctx->wave[k].counters + 3, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream)); }
  BZ_CUDA(cudaStreamSynchronize(ctx->stream));      // (the lanes were joined into the stream by the call that used them)
  for(int k = 0; k < BZ_WAVE_LANES; ++k) {
    if(flags[k] != 0u) {
      BZ_CUDA(cudaMemsetAsync(ctx->wave[k].counters + 3, 0, sizeof(uint32_t), ctx->stream));
      *out = true;
    }
  }
  return BZ_OK;
}

static int bz_run(bz_context *ctx, int job, bz_job &desc) {
  bz_lens *lens = ctx->lens;
  cudaStream_t const stream = ctx->stream;
  gLastContext = ctx;
  ctx->plan_ms = 0.0;
  BZ_CUDA(cudaSetDevice(lens->device));
  ctx->timed = false;
  ctx->launches = 0u;
  if(desc.ray_count == 0u) { return BZ_OK; }
  BZ_CUDA(cudaMemsetAsync(ctx->d_counter, 0, sizeof(unsigned long long), stream));
  desc.next_ray = ctx->d_counter;
  desc.one = 1.0f;
  bool const brute = (lens->mode == BZ_MODE_BRUTE);
  int grid;
  uint32_t sky_count = 0u;                // RENDER: tiles shaded by bz_sky_kernel (after the lens tiles in ctx->d_tiles)
  if(brute) {
    uint64_t const threads = (uint64_t)bz_trace_threads();
    uint64_t const wanted = (desc.ray_count + threads - 1u) / threads;
    grid = (int)std::min<uint64_t>((uint64_t)lens->sm_count * 2u, wanted);
  }
  else {
    // tiles of the job: RENDER: square pixel blocks of about 1024 rays (side a power of two); batches: 1024 consecutive rays
    if(job == BZ_JOB_RENDER) {
      uint32_t const spp = desc.camera.strata * desc.camera.strata;
      uint32_t side = 32u;
      while(side > 2u && (uint64_t)side * side * spp > 1536u) { side /= 2u; }
      // small frames: fewer than ~16 tiles per SM leave the launch waiting on the last heavy tiles (1080p, 8 crossings:
      // 0.58 -> 0.48 ms with 16 x 16 tiles), so the tiles shrink, down to 8 packets, until there are enough of them
      auto tiles_of = [&](uint32_t s) { return (uint64_t)((desc.camera.width + s - 1u) / s) * ((desc.rows + s - 1u) / s); };
      while(side > 4u && (uint64_t)side * side * spp >= 1024u && tiles_of(side) < 16ull * (uint64_t)lens->sm_count) { side /= 2u; }
      if((uint64_t)side * side * spp >= (1ull << 31)) { return bz_fail(BZ_ERR_INVALID, "too many samples per pixel"); }
      // divisions by job constants become multiplications (exact while n * d < 2^32: n < tile_rays + 32 for spp,
      // n < spp for strata, n < rows for stripe_rows)
      auto reciprocal = [](uint32_t d, uint32_t out[2]) { out[0] = (d <= 1u) ? 0u : (uint32_t)(((1ull << 32) + d - 1u) / d); out[1] = (d == 1u) ? 1u : 0u; };
      if(desc.camera.strata > 128u) { return bz_fail(BZ_ERR_INVALID, "more than 128 x 128 samples per pixel"); }
      if((uint64_t)desc.rows * std::max(desc.stripe_rows, 1u) >= (1ull << 32)) { return bz_fail(BZ_ERR_INVALID, "band too large for the stripe layout"); }
      reciprocal(spp, desc.div_spp);
      reciprocal(desc.camera.strata, desc.div_strata);
      reciprocal(desc.stripe_rows, desc.div_stripe);
      desc.two_over_w = 2.0f / (float)desc.camera.width;
      desc.two_over_h = 2.0f / (float)desc.camera.height;
      desc.tile_side = side;
      desc.tile_rays = side * side * spp;
      desc.tiles_x = (desc.camera.width + side - 1u) / side;
      desc.tiles_y = (desc.rows + side - 1u) / side;
      desc.tile_count = (uint64_t)desc.tiles_x * desc.tiles_y;
      if(desc.tile_count >= 0xffffffffull) { return bz_fail(BZ_ERR_INVALID, "job too large"); }
      // the tiles that can meet the lens go to the trace kernel, the others to bz_sky_kernel; the plan of the last frame
      // is kept (same camera, same band: nothing to do)
      auto &plan = ctx->plan;
      bool const same = (lens->mode == BZ_MODE_CULLED || lens->mode == BZ_MODE_FAST) && plan.valid && std::memcmp(&plan.camera, &desc.camera, sizeof(bz_camera)) == 0 && plan.row0 == desc.row0 && plan.rows == desc.rows
                     && plan.stripe_rows == desc.stripe_rows && plan.rank == desc.stripe_rank && plan.ranks == desc.stripe_ranks && plan.side == side;
      if(!same) {
        auto const t0 = std::chrono::steady_clock::now();
        plan.valid = false;
        std::vector<uint32_t> lens_tiles, sky_tiles;
        if(bz_plan_tiles(lens, desc, lens_tiles, sky_tiles)) {
          size_t const total = lens_tiles.size() + sky_tiles.size();
          if(total > ctx->tiles_capacity) {
            if(ctx->d_tiles != nullptr) { BZ_CUDA(cudaFree(ctx->d_tiles)); ctx->d_tiles = nullptr; ctx->tiles_capacity = 0u; }
            BZ_CUDA(cudaMalloc((void**)&ctx->d_tiles, total * sizeof(uint32_t)));
            ctx->tiles_capacity = total;
          }
          if(total > ctx->h_tiles_capacity) {
            if(ctx->h_tiles != nullptr) { BZ_CUDA(cudaEventSynchronize(ctx->ev_tiles)); BZ_CUDA(cudaFreeHost(ctx->h_tiles)); ctx->h_tiles = nullptr; ctx->h_tiles_capacity = 0u; }
            BZ_CUDA(cudaMallocHost((void**)&ctx->h_tiles, total * sizeof(uint32_t)));
            ctx->h_tiles_capacity = total;
          }
          else { BZ_CUDA(cudaEventSynchronize(ctx->ev_tiles)); }      // the previous upload out of the staging buffer is done (a moving camera plans every frame)
          std::memcpy(ctx->h_tiles, lens_tiles.data(), lens_tiles.size() * sizeof(uint32_t));
          std::memcpy(ctx->h_tiles + lens_tiles.size(), sky_tiles.data(), sky_tiles.size() * sizeof(uint32_t));
          BZ_CUDA(cudaMemcpyAsync(ctx->d_tiles, ctx->h_tiles, total * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
          BZ_CUDA(cudaEventRecord(ctx->ev_tiles, stream));
          plan.valid = true;
          plan.camera = desc.camera;
          plan.row0 = desc.row0; plan.rows = desc.rows; plan.stripe_rows = desc.stripe_rows; plan.rank = desc.stripe_rank; plan.ranks = desc.stripe_ranks; plan.side = side;
          plan.sky_count = (uint32_t)sky_tiles.size();
          plan.lens_count = (uint32_t)lens_tiles.size();
        }
        ctx->plan_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
      }
      if(plan.valid) {
        desc.tile_list = ctx->d_tiles;
        desc.tile_count = plan.lens_count;
        sky_count = plan.sky_count;
      }
    }
    else {
      desc.tile_side = 0u;
      desc.tile_rays = 1024u;
      desc.tiles_x = 0u;
      desc.tile_count = (desc.ray_count + 1023u) / 1024u;
    }
    if(desc.tile_count >= 0xffffffffull) { return bz_fail(BZ_ERR_INVALID, "job too large"); }
    // persistent grid: every resident CTA slot of the device, or fewer when the job has fewer tiles
    int perSm = 0, raysPerCta = 0;
    if(lens->mode == BZ_MODE_FAST) { BZ_CUDA(bz_warp_trace_occupancy_fast(&perSm, &raysPerCta)); }
    else { BZ_CUDA(bz_warp_trace_occupancy(&perSm, &raysPerCta)); }
    grid = (int)std::min<uint64_t>((uint64_t)lens->sm_count * (uint64_t)std::max(perSm, 1), desc.tile_count);
  }
  BZ_CUDA(cudaEventRecord(ctx->ev_begin, stream));
  if(brute) { BZ_CUDA(bz_launch_trace(job, true, lens->tables, desc, grid, stream)); ++ctx->launches; }
  else {
    bz_tables tables = lens->tables;
    if(lens->mode == BZ_MODE_LANES) { tables.packet_count = 0u; }
    // RENDER: the wavefront pipeline when asked for.  BZ_PIPELINE_AUTO takes it where it measured faster: the
    // parity build, large frames, few crossings -- and only while its ray / candidate lists fit comfortably into free device memory.
    bool wavefront = job == BZ_JOB_RENDER && lens->mode != BZ_MODE_LANES && grid != 0 && lens->pipeline == BZ_PIPELINE_WAVEFRONT;
    if(job == BZ_JOB_RENDER && lens->pipeline == BZ_PIPELINE_AUTO && lens->mode == BZ_MODE_CULLED && grid != 0) {
      uint64_t const slots = desc.tile_count * (uint64_t)(((desc.tile_rays + 31u) / 32u) * 32u);
      if(slots >= BZ_WAVE_AUTO_RAYS && desc.scene.max_crossings <= BZ_WAVE_AUTO_CROSSINGS) {
        uint64_t const need = bz_wave_bytes(lens, slots);
        size_t free_bytes = 0u, total_bytes = 0u;
        wavefront = ctx->wave_bytes >= need || (cudaMemGetInfo(&free_bytes, &total_bytes) == cudaSuccess && need <= free_bytes / 4u);
      }
    }
    if(wavefront) {
      int rc = bz_run_wave(ctx, desc, ctx->d_tiles + desc.tile_count, sky_count);
      if(rc != BZ_OK) { return rc; }
      sky_count = 0u;
    }
    if(sky_count != 0u) { BZ_CUDA(bz_launch_sky(desc, ctx->d_tiles + desc.tile_count, sky_count, stream)); ++ctx->launches; }
    if(!wavefront && grid != 0) {
      if(lens->mode == BZ_MODE_FAST) { BZ_CUDA(bz_launch_warp_trace_fast(job, tables, desc, grid, stream)); }
      else { BZ_CUDA(bz_launch_warp_trace(job, tables, desc, grid, stream)); }
      ++ctx->launches;
    }
  }
  BZ_CUDA(cudaEventRecord(ctx->ev_end, stream));
  ctx->timed = true;
  return BZ_OK;
}

static int bz_scratch(bz_context *ctx, size_t bytes) {
  if(bytes <= ctx->scratch_bytes) { return BZ_OK; }
  BZ_CUDA(cudaSetDevice(ctx->lens->device));
  if(ctx->d_scratch != nullptr) { BZ_CUDA(cudaFree(ctx->d_scratch)); ctx->d_scratch = nullptr; ctx->scratch_bytes = 0u; }
  BZ_CUDA(cudaMalloc(&ctx->d_scratch, bytes));
  ctx->scratch_bytes = bytes;
  return BZ_OK;
}

static int bz_render_common(bz_context *ctx, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t rows,
                            uint32_t stripe_rows, uint32_t rank, uint32_t ranks, uint8_t *d_rgba, bz_stats *d_stats) {
  bz_lens *lens = ctx->lens;
  cudaStream_t const stream = ctx->stream;
  if(camera->width == 0u || camera->height == 0u || camera->strata == 0u) { return bz_fail(BZ_ERR_INVALID, "empty camera"); }
  if(rows == 0u) { ctx->timed = false; ctx->launches = 0u; gLastContext = ctx; return BZ_OK; }
  if(d_rgba == nullptr) { return bz_fail(BZ_ERR_INVALID, "d_rgba is null"); }
  BZ_CUDA(cudaSetDevice(lens->device));
  size_t const pixels = (size_t)rows * camera->width;
  if(pixels >= (1ull << 32)) { return bz_fail(BZ_ERR_INVALID, "band too large"); }
  if(pixels > ctx->accum_pixels) {
    if(ctx->d_accum != nullptr) { BZ_CUDA(cudaFree(ctx->d_accum)); ctx->d_accum = nullptr; ctx->accum_pixels = 0u; }
    BZ_CUDA(cudaMalloc((void**)&ctx->d_accum, pixels * sizeof(uint32_t)));
    ctx->accum_pixels = pixels;
  }
  BZ_CUDA(cudaMemsetAsync(ctx->d_accum, 0, pixels * sizeof(uint32_t), stream));
  bz_job desc;
  std::memset(&desc, 0, sizeof(desc));
  desc.camera = *camera;
  desc.scene = *scene;
  desc.row0 = row0;
  desc.rows = rows;
  desc.stripe_rows = stripe_rows;
  desc.stripe_rank = rank;
  desc.stripe_ranks = ranks;
  desc.accum = ctx->d_accum;
  uint32_t const spp = camera->strata * camera->strata;
  desc.ray_count = (uint64_t)pixels * spp;
  desc.stats = (unsigned long long*)d_stats;
  int rc = bz_run(ctx, BZ_JOB_RENDER, desc);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(bz_launch_resolve(ctx->d_accum, d_rgba, pixels, spp, stream));
  ++ctx->launches;
  return BZ_OK;
}

extern "C" {

// the job descriptions of the two batch entry points (everything but the launch)
static int bz_refract_job(const float *d_start, const float *d_dir, const uint32_t *d_expected, uint64_t n,
                          float *d_out_start, float *d_out_dir, uint32_t *d_out_status, uint32_t *d_out_patch, bz_job &desc) {
  if(n != 0u && (d_start == nullptr || d_dir == nullptr || d_expected == nullptr || d_out_start == nullptr || d_out_dir == nullptr || d_out_status == nullptr)) {
    return bz_fail(BZ_ERR_INVALID, "null ray buffer");
  }
  std::memset(&desc, 0, sizeof(desc));
  desc.in_start = d_start; desc.in_dir = d_dir; desc.in_expected = d_expected;
  desc.out_start = d_out_start; desc.out_dir = d_out_dir; desc.out_status = d_out_status; desc.out_patch = d_out_patch;
  desc.ray_count = n;
  return BZ_OK;
}

static int bz_intersect_job(const float *d_start, const float *d_dir, uint64_t n, uint32_t *d_out_what, uint32_t *d_out_patch, float *d_out_distance,
                            float *d_out_point, float *d_out_normal, float *d_out_cos, float *d_out_bary, bz_job &desc) {
  if(n != 0u && (d_start == nullptr || d_dir == nullptr || d_out_what == nullptr)) { return bz_fail(BZ_ERR_INVALID, "null ray buffer"); }
  std::memset(&desc, 0, sizeof(desc));
  desc.in_start = d_start; desc.in_dir = d_dir;
  desc.out_status = d_out_what; desc.out_patch = d_out_patch; desc.out_distance = d_out_distance; desc.out_point = d_out_point;
  desc.out_normal = d_out_normal; desc.out_cos = d_out_cos; desc.out_bary = d_out_bary;
  desc.ray_count = n;
  return BZ_OK;
}

int bz_refract_batch(bz_lens *lens, const float *d_start, const float *d_dir, const uint32_t *d_expected, uint64_t n,
                     float *d_out_start, float *d_out_dir, uint32_t *d_out_status, uint32_t *d_out_patch, void *stream) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  bz_job desc;
  int rc = bz_refract_job(d_start, d_dir, d_expected, n, d_out_start, d_out_dir, d_out_status, d_out_patch, desc);
  if(rc != BZ_OK) { return rc; }
  bz_context *ctx = nullptr;
  rc = bz_context_for(lens, (cudaStream_t)stream, &ctx);
  if(rc != BZ_OK) { return rc; }
  std::lock_guard<std::mutex> hold(ctx->busy);
  return bz_run(ctx, BZ_JOB_REFRACT, desc);
}

int bz_intersect_batch(bz_lens *lens, const float *d_start, const float *d_dir, uint64_t n,
                       uint32_t *d_out_what, uint32_t *d_out_patch, float *d_out_distance, float *d_out_point,
                       float *d_out_normal, float *d_out_cos, float *d_out_bary, void *stream) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  bz_job desc;
  int rc = bz_intersect_job(d_start, d_dir, n, d_out_what, d_out_patch, d_out_distance, d_out_point, d_out_normal, d_out_cos, d_out_bary, desc);
  if(rc != BZ_OK) { return rc; }
  bz_context *ctx = nullptr;
  rc = bz_context_for(lens, (cudaStream_t)stream, &ctx);
  if(rc != BZ_OK) { return rc; }
  std::lock_guard<std::mutex> hold(ctx->busy);
  return bz_run(ctx, BZ_JOB_INTERSECT, desc);
}

int bz_refract_batch_host(bz_lens *lens, const float *h_start, const float *h_dir, const uint32_t *h_expected, uint64_t n,
                          float *h_out_start, float *h_out_dir, uint32_t *h_out_status, uint32_t *h_out_patch) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  if(n == 0u) { return BZ_OK; }
  if(h_start == nullptr || h_dir == nullptr || h_expected == nullptr || h_out_start == nullptr || h_out_dir == nullptr || h_out_status == nullptr) {
    return bz_fail(BZ_ERR_INVALID, "null ray buffer");
  }
  bz_host_call call(lens);
  if(call.rc != BZ_OK) { return call.rc; }
  bz_context *ctx = call.ctx;
  cudaStream_t const st = ctx->stream;
  int rc = bz_scratch(ctx, (size_t)n * 60u);
  if(rc != BZ_OK) { return rc; }
  float *ds = (float*)ctx->d_scratch, *dd = ds + n * 3u, *os = dd + n * 3u, *od = os + n * 3u;
  uint32_t *de = (uint32_t*)(od + n * 3u), *ost = de + n, *op = ost + n;
  BZ_CUDA(cudaMemcpyAsync(ds, h_start, n * 12u, cudaMemcpyHostToDevice, st));
  BZ_CUDA(cudaMemcpyAsync(dd, h_dir, n * 12u, cudaMemcpyHostToDevice, st));
  BZ_CUDA(cudaMemcpyAsync(de, h_expected, n * 4u, cudaMemcpyHostToDevice, st));
  bz_job desc;
  rc = bz_refract_job(ds, dd, de, n, os, od, ost, op, desc);
  if(rc == BZ_OK) { rc = bz_run(ctx, BZ_JOB_REFRACT, desc); }
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaMemcpyAsync(h_out_start, os, n * 12u, cudaMemcpyDeviceToHost, st));
  BZ_CUDA(cudaMemcpyAsync(h_out_dir, od, n * 12u, cudaMemcpyDeviceToHost, st));
  BZ_CUDA(cudaMemcpyAsync(h_out_status, ost, n * 4u, cudaMemcpyDeviceToHost, st));
  if(h_out_patch != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_patch, op, n * 4u, cudaMemcpyDeviceToHost, st)); }
  BZ_CUDA(cudaStreamSynchronize(st));
  return BZ_OK;
}

int bz_intersect_batch_host(bz_lens *lens, const float *h_start, const float *h_dir, uint64_t n,
                            uint32_t *h_out_what, uint32_t *h_out_patch, float *h_out_distance, float *h_out_point,
                            float *h_out_normal, float *h_out_cos, float *h_out_bary) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  if(n == 0u) { return BZ_OK; }
  if(h_start == nullptr || h_dir == nullptr || h_out_what == nullptr) { return bz_fail(BZ_ERR_INVALID, "null ray buffer"); }
  bz_host_call call(lens);
  if(call.rc != BZ_OK) { return call.rc; }
  bz_context *ctx = call.ctx;
  cudaStream_t const st = ctx->stream;
  int rc = bz_scratch(ctx, (size_t)n * 76u);
  if(rc != BZ_OK) { return rc; }
  float *ds = (float*)ctx->d_scratch, *dd = ds + n * 3u, *opt = dd + n * 3u, *onr = opt + n * 3u, *oba = onr + n * 3u;
  float *odi = oba + n * 3u, *oco = odi + n;
  uint32_t *owh = (uint32_t*)(oco + n), *opa = owh + n;
  BZ_CUDA(cudaMemcpyAsync(ds, h_start, n * 12u, cudaMemcpyHostToDevice, st));
  BZ_CUDA(cudaMemcpyAsync(dd, h_dir, n * 12u, cudaMemcpyHostToDevice, st));
  bz_job desc;
  rc = bz_intersect_job(ds, dd, n, owh, opa, odi, opt, onr, oco, oba, desc);
  if(rc == BZ_OK) { rc = bz_run(ctx, BZ_JOB_INTERSECT, desc); }
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaMemcpyAsync(h_out_what, owh, n * 4u, cudaMemcpyDeviceToHost, st));
  if(h_out_patch != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_patch, opa, n * 4u, cudaMemcpyDeviceToHost, st)); }
  if(h_out_distance != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_distance, odi, n * 4u, cudaMemcpyDeviceToHost, st)); }
  if(h_out_cos != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_cos, oco, n * 4u, cudaMemcpyDeviceToHost, st)); }
  if(h_out_point != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_point, opt, n * 12u, cudaMemcpyDeviceToHost, st)); }
  if(h_out_normal != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_normal, onr, n * 12u, cudaMemcpyDeviceToHost, st)); }
  if(h_out_bary != nullptr) { BZ_CUDA(cudaMemcpyAsync(h_out_bary, oba, n * 12u, cudaMemcpyDeviceToHost, st)); }
  BZ_CUDA(cudaStreamSynchronize(st));
  return BZ_OK;
}

int bz_generate_rays(const bz_camera *camera, uint32_t row0, uint32_t row1, float *d_start, float *d_dir, void *stream) {
  if(camera == nullptr || row1 < row0 || row1 > camera->height) { return bz_fail(BZ_ERR_INVALID, "bad camera or rows"); }
  uint64_t const count = (uint64_t)(row1 - row0) * camera->width * camera->strata * camera->strata;
  if(count != 0u && (d_start == nullptr || d_dir == nullptr)) { return bz_fail(BZ_ERR_INVALID, "null ray buffer"); }
  int devices = 0;
  if(cudaGetDeviceCount(&devices) != cudaSuccess || devices == 0) { cudaGetLastError(); return bz_fail(BZ_ERR_NO_DEVICE, "no usable CUDA device (there is no CPU path)"); }
  BZ_CUDA(bz_launch_raygen(*camera, row0, count, d_start, d_dir, (cudaStream_t)stream));
  return BZ_OK;
}

int bz_render(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t row1,
              uint8_t *d_rgba, bz_stats *d_stats, void *stream) {
  if(lens == nullptr || camera == nullptr || scene == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  if(row1 < row0 || row1 > camera->height) { return bz_fail(BZ_ERR_INVALID, "bad rows"); }
  bz_context *ctx = nullptr;
  int rc = bz_context_for(lens, (cudaStream_t)stream, &ctx);
  if(rc != BZ_OK) { return rc; }
  std::lock_guard<std::mutex> hold(ctx->busy);
  return bz_render_common(ctx, camera, scene, row0, row1 - row0, 0u, 0u, 1u, d_rgba, d_stats);
}

// ---- bz_render_static: bz_render for callers that may not allocate (no counterpart in the reference).
// The scratch of the render -- the work counter of the persistent kernel and one u32 accumulator per pixel -- lives in a
// file-scope device array instead of the context's cudaMalloc'ed buffers; nothing on this path allocates or frees.  Every GPU that loads the
// library holds the array (8 MiB + 256 B) until exit; all calls share it, so calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char bz_static_scratch[256u + (size_t)BZ_STATIC_MAX_PIXELS * sizeof(uint32_t)];

// the array's device address on `device` (which is current); asked once per device -- in host code the array's NAME is a host-side
// stand-in, not a device address
static int bz_static_scratch_address(int device, unsigned char **out) {
  static std::mutex guard;
  static unsigned char *address[64] = {};
  if(device < 0 || device >= 64) { return bz_fail(BZ_ERR_INVALID, "device index out of range"); }
  std::lock_guard<std::mutex> hold(guard);
  if(address[device] == nullptr) {
    void *p = nullptr;
    cudaError_t const e = cudaGetSymbolAddress(&p, bz_static_scratch);
    if(e != cudaSuccess || p == nullptr) { return bz_fail(BZ_ERR_CUDA, std::string("cudaGetSymbolAddress(bz_static_scratch): ") + cudaGetErrorString(e)); }
    address[device] = (unsigned char*)p;
  }
  *out = address[device];
  return BZ_OK;
}

int bz_render_static(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t row1,
                     uint8_t *d_rgba, bz_stats *d_stats, void *stream) {
  if(lens == nullptr || camera == nullptr || scene == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  if(row1 < row0 || row1 > camera->height) { return bz_fail(BZ_ERR_INVALID, "bad rows"); }
  if(camera->width == 0u || camera->height == 0u || camera->strata == 0u) { return bz_fail(BZ_ERR_INVALID, "empty camera"); }
  uint32_t const rows = row1 - row0;
  if(rows == 0u) { return BZ_OK; }
  if(d_rgba == nullptr) { return bz_fail(BZ_ERR_INVALID, "d_rgba is null"); }
  uint64_t const pixels = (uint64_t)rows * camera->width;
  if(pixels > BZ_STATIC_MAX_PIXELS) {
    return bz_fail(BZ_ERR_INVALID, "bz_render_static: " + std::to_string(pixels) + " pixels exceed the static scratch of " + std::to_string(BZ_STATIC_MAX_PIXELS)
                                   + " (BZ_STATIC_MAX_PIXELS); render the frame in row bands or use bz_render");
  }
  if(camera->strata > 128u) { return bz_fail(BZ_ERR_INVALID, "more than 128 x 128 samples per pixel"); }
  cudaStream_t const st = (cudaStream_t)stream;
  BZ_CUDA(cudaSetDevice(lens->device));
  unsigned char *scratch = nullptr;
  int rc = bz_static_scratch_address(lens->device, &scratch);
  if(rc != BZ_OK) { return rc; }
  unsigned long long *counter = (unsigned long long*)scratch;
  uint32_t *accum = (uint32_t*)(scratch + 256u);
  // the array keeps the last call's contents: the counter and the accumulators of this band start at zero
  BZ_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), st));
  BZ_CUDA(cudaMemsetAsync(accum, 0, (size_t)pixels * sizeof(uint32_t), st));
  uint32_t const spp = camera->strata * camera->strata;
  bz_job desc;
  std::memset(&desc, 0, sizeof(desc));
  desc.camera = *camera;
  desc.scene = *scene;
  desc.row0 = row0;
  desc.rows = rows;
  desc.stripe_ranks = 1u;
  desc.accum = accum;
  desc.next_ray = counter;
  desc.one = 1.0f;
  desc.ray_count = pixels * spp;
  desc.stats = (unsigned long long*)d_stats;
  if(lens->mode == BZ_MODE_BRUTE) {
    uint64_t const threads = (uint64_t)bz_trace_threads();
    int const grid = (int)std::min<uint64_t>((uint64_t)lens->sm_count * 2u, (desc.ray_count + threads - 1u) / threads);
    BZ_CUDA(bz_launch_trace(BZ_JOB_RENDER, true, lens->tables, desc, grid, st));
  }
  else {
    // the tiles of bz_run, all of them traced by the persistent kernel: no tile plan (its list is a per-frame upload into an allocated
    // buffer), so the rays beside the lens go through the root test of the walk instead of bz_sky_kernel -- same pixels, more time
    uint32_t side = 32u;
    while(side > 2u && (uint64_t)side * side * spp > 1536u) { side /= 2u; }
    auto tiles_of = [&](uint32_t s) { return (uint64_t)((camera->width + s - 1u) / s) * ((rows + s - 1u) / s); };
    while(side > 4u && (uint64_t)side * side * spp >= 1024u && tiles_of(side) < 16ull * (uint64_t)lens->sm_count) { side /= 2u; }
    auto reciprocal = [](uint32_t d, uint32_t out[2]) { out[0] = (d <= 1u) ? 0u : (uint32_t)(((1ull << 32) + d - 1u) / d); out[1] = (d == 1u) ? 1u : 0u; };
    reciprocal(spp, desc.div_spp);
    reciprocal(camera->strata, desc.div_strata);
    reciprocal(0u, desc.div_stripe);
    desc.two_over_w = 2.0f / (float)camera->width;
    desc.two_over_h = 2.0f / (float)camera->height;
    desc.tile_side = side;
    desc.tile_rays = side * side * spp;
    desc.tiles_x = (camera->width + side - 1u) / side;
    desc.tiles_y = (rows + side - 1u) / side;
    desc.tile_count = (uint64_t)desc.tiles_x * desc.tiles_y;
    int perSm = 0, raysPerCta = 0;
    if(lens->mode == BZ_MODE_FAST) { BZ_CUDA(bz_warp_trace_occupancy_fast(&perSm, &raysPerCta)); }
    else { BZ_CUDA(bz_warp_trace_occupancy(&perSm, &raysPerCta)); }
    int const grid = (int)std::min<uint64_t>((uint64_t)lens->sm_count * (uint64_t)std::max(perSm, 1), desc.tile_count);
    bz_tables tables = lens->tables;
    if(lens->mode == BZ_MODE_LANES) { tables.packet_count = 0u; }
    if(lens->mode == BZ_MODE_FAST) { BZ_CUDA(bz_launch_warp_trace_fast(BZ_JOB_RENDER, tables, desc, grid, st)); }
    else { BZ_CUDA(bz_launch_warp_trace(BZ_JOB_RENDER, tables, desc, grid, st)); }
  }
  BZ_CUDA(bz_launch_resolve(accum, d_rgba, (size_t)pixels, spp, st));
  return BZ_OK;
}

uint32_t bz_striped_rows(uint32_t height, uint32_t stripe_rows, uint32_t rank, uint32_t ranks) {
  if(stripe_rows == 0u || ranks == 0u || rank >= ranks) { return 0u; }
  uint32_t rows = 0u;
  for(uint32_t first = rank * stripe_rows; first < height; first += ranks * stripe_rows) { rows += std::min(stripe_rows, height - first); }
  return rows;
}

int bz_render_striped(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t stripe_rows, uint32_t rank,
                      uint32_t ranks, uint8_t *d_rgba, bz_stats *d_stats, void *stream) {
  if(lens == nullptr || camera == nullptr || scene == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  if(stripe_rows == 0u || ranks == 0u || rank >= ranks) { return bz_fail(BZ_ERR_INVALID, "bad stripe layout"); }
  uint32_t const rows = bz_striped_rows(camera->height, stripe_rows, rank, ranks);
  bz_context *ctx = nullptr;
  int rc = bz_context_for(lens, (cudaStream_t)stream, &ctx);
  if(rc != BZ_OK) { return rc; }
  std::lock_guard<std::mutex> hold(ctx->busy);
  return bz_render_common(ctx, camera, scene, 0u, rows, stripe_rows, rank, ranks, d_rgba, d_stats);
}

int bz_assemble_stripes(const uint8_t *d_bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                        uint32_t ranks, uint8_t *d_rgba, void *stream) {
  if(stripe_rows == 0u || ranks == 0u) { return bz_fail(BZ_ERR_INVALID, "bad stripe layout"); }
  if((uint64_t)width * height != 0u && (d_bands == nullptr || d_rgba == nullptr)) { return bz_fail(BZ_ERR_INVALID, "null image buffer"); }
  for(uint32_t rank = 0u; rank < ranks; ++rank) {
    if(bz_striped_rows(height, stripe_rows, rank, ranks) > band_rows) { return bz_fail(BZ_ERR_INVALID, "band_rows too small for the stripe layout"); }
  }
  BZ_CUDA(bz_launch_assemble(d_bands, band_rows, width, height, stripe_rows, ranks, d_rgba, (cudaStream_t)stream));
  return BZ_OK;
}

int bz_render_host(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t row1,
                   uint8_t *h_rgba, bz_stats *h_stats) {
  if(lens == nullptr || camera == nullptr || scene == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  if(row1 < row0 || row1 > camera->height) { return bz_fail(BZ_ERR_INVALID, "bad rows"); }
  size_t const bytes = (size_t)(row1 - row0) * camera->width * 4u;
  if(bytes == 0u) { return BZ_OK; }
  if(h_rgba == nullptr) { return bz_fail(BZ_ERR_INVALID, "h_rgba is null"); }
  bz_host_call call(lens);
  if(call.rc != BZ_OK) { return call.rc; }
  bz_context *ctx = call.ctx;
  cudaStream_t const st = ctx->stream;
  int rc = bz_scratch(ctx, bytes + 256u);
  if(rc != BZ_OK) { return rc; }
  bz_stats *d_stats = (bz_stats*)ctx->d_scratch;
  uint8_t *d_rgba = (uint8_t*)ctx->d_scratch + 256u;
  BZ_CUDA(cudaMemsetAsync(d_stats, 0, sizeof(bz_stats), st));
  rc = bz_render_common(ctx, camera, scene, row0, row1 - row0, 0u, 0u, 1u, d_rgba, h_stats != nullptr ? d_stats : nullptr);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaMemcpyAsync(h_rgba, d_rgba, bytes, cudaMemcpyDeviceToHost, st));
  bz_stats fresh;
  if(h_stats != nullptr) { BZ_CUDA(cudaMemcpyAsync(&fresh, d_stats, sizeof(bz_stats), cudaMemcpyDeviceToHost, st)); }
  bool overflow = false;
  rc = bz_wave_overflowed(ctx, &overflow);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaStreamSynchronize(st));
  if(h_stats != nullptr) {
    uint64_t *acc = (uint64_t*)h_stats;
    const uint64_t *add = (const uint64_t*)&fresh;
    for(size_t i = 0u; i < sizeof(bz_stats) / sizeof(uint64_t); ++i) { acc[i] += add[i]; }
  }
  if(overflow) {
    return bz_fail(BZ_ERR_OVERFLOW, "candidate list of the wavefront pipeline overflowed: the frame is incomplete; render with bz_lens_set_pipeline(lens, BZ_PIPELINE_PERSISTENT)");
  }
  return BZ_OK;
}

// the context of the calling thread's most recent call, if it (still) belongs to `lens`
static bz_context *bz_last_context(const bz_lens *lens) {
  bz_context *ctx = gLastContext;
  return (ctx != nullptr && ctx->lens == lens) ? ctx : nullptr;
}

int bz_last_kernel_ms(bz_lens *lens, float *out_ms) {
  if(lens == nullptr || out_ms == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  *out_ms = 0.0f;
  bz_context *ctx = bz_last_context(lens);
  if(ctx == nullptr || !ctx->timed) { return BZ_OK; }
  BZ_CUDA(cudaSetDevice(lens->device));
  BZ_CUDA(cudaEventSynchronize(ctx->ev_end));
  BZ_CUDA(cudaEventElapsedTime(out_ms, ctx->ev_begin, ctx->ev_end));
  return BZ_OK;
}

int bz_last_kernel_launches(const bz_lens *lens, uint32_t *out_count) {
  if(lens == nullptr || out_count == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  bz_context *ctx = bz_last_context(lens);
  *out_count = (ctx != nullptr) ? ctx->launches : 0u;
  return BZ_OK;
}

int bz_last_plan_ms(const bz_lens *lens, float *out_ms) {
  if(lens == nullptr || out_ms == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  bz_context *ctx = bz_last_context(lens);
  *out_ms = (ctx != nullptr) ? (float)ctx->plan_ms : 0.0f;
  return BZ_OK;
}

int bz_fp32_peak_probe(int device, float *out_tflops, float *out_ms) {
  if(out_tflops == nullptr) { return bz_fail(BZ_ERR_INVALID, "out_tflops is null"); }
  int rc = bz_check_device(device);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  BZ_CUDA(cudaGetDeviceProperties(&prop, device));
  float *sink = nullptr;
  BZ_CUDA(cudaMalloc((void**)&sink, 16u));
  cudaEvent_t a, b;
  BZ_CUDA(cudaEventCreate(&a));
  BZ_CUDA(cudaEventCreate(&b));
  int const grid = prop.multiProcessorCount * 8;
  uint32_t const iters = 1u << 15;
  float best = 1e30f;
  for(int rep = 0; rep < 4; ++rep) {
    BZ_CUDA(cudaEventRecord(a, 0));
    BZ_CUDA(bz_launch_fma_probe(sink, iters, grid, 0));
    BZ_CUDA(cudaEventRecord(b, 0));
    BZ_CUDA(cudaEventSynchronize(b));
    float ms = 0.0f;
    BZ_CUDA(cudaEventElapsedTime(&ms, a, b));
    if(rep > 0) { best = std::min(best, ms); }
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  cudaFree(sink);
  double const flop = 2.0 * 16.0 * (double)iters * 256.0 * (double)grid;      // BZ_PROBE_CHAINS chains (csrc/bz_kernels.cu)
  *out_tflops = (float)(flop / ((double)best * 1e-3) * 1e-12);
  if(out_ms != nullptr) { *out_ms = best; }
  return BZ_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ illumination mode ----

// UniformHemisphere::UniformHemisphere, reference/hostUtil.cpp:3-14: sin and ceil in double (the reference's ::sin(float) resolves
// to the double overload), patch width and belt width in float
static int bz_make_belts(uint32_t count, bz_belts &out) {
  if(count == 0u || count > BZ_MAX_BELTS) { return bz_fail(BZ_ERR_INVALID, "belts must be 1..128"); }
  const float pi = 3.14159265358979323846f;
  std::memset(&out, 0, sizeof(out));
  out.belts = count;
  out.belt_width = pi / 2.0f / (float)count;
  uint32_t total = 0u;
  for(uint32_t i = 0u; i < count; ++i) {
    float const angle = (2.0f * (float)i + 1.0f) / (4.0f * (float)count) * pi;
    uint32_t const now = (uint32_t)std::ceil((double)(4.0f * (float)count) * std::sin((double)angle));
    out.width[i] = pi * 2.0f / (float)now;
    out.first[i] = total;
    total += now;
  }
  out.patches = total;
  return BZ_OK;
}

static int bz_check_emitter(const bz_emitter *em, uint64_t *out_rays) {
  if(em == nullptr) { return bz_fail(BZ_ERR_INVALID, "emitter is null"); }
  if(em->points_u == 0u || em->points_v == 0u || em->rays_per_point == 0u) { return bz_fail(BZ_ERR_INVALID, "empty emitter"); }
  *out_rays = (uint64_t)em->points_u * em->points_v * em->rays_per_point;
  return BZ_OK;
}

extern "C" {

uint32_t bz_sizeof_emitter(void) { return (uint32_t)sizeof(bz_emitter); }
uint32_t bz_sizeof_target(void) { return (uint32_t)sizeof(bz_target); }

int bz_hemisphere_patches(uint32_t belts, uint32_t *out_patch_count, float *out_width, uint32_t *out_first) {
  bz_belts table;
  int rc = bz_make_belts(belts, table);
  if(rc != BZ_OK) { return rc; }
  if(out_patch_count != nullptr) { *out_patch_count = table.patches; }
  for(uint32_t i = 0u; i < belts; ++i) {
    if(out_width != nullptr) { out_width[i] = table.width[i]; }
    if(out_first != nullptr) { out_first[i] = table.first[i]; }
  }
  return BZ_OK;
}

int bz_emit_rays(const bz_emitter *emitter, uint64_t first, uint64_t count, float *d_start, float *d_dir, uint32_t *d_patch, void *stream) {
  uint64_t rays = 0u;
  int rc = bz_check_emitter(emitter, &rays);
  if(rc != BZ_OK) { return rc; }
  if(first > rays || count > rays - first) { return bz_fail(BZ_ERR_INVALID, "ray range beyond the emission"); }
  if(count != 0u && (d_start == nullptr || d_dir == nullptr)) { return bz_fail(BZ_ERR_INVALID, "null ray buffer"); }
  bz_belts table;
  rc = bz_make_belts(emitter->belts, table);
  if(rc != BZ_OK) { return rc; }
  int devices = 0;
  if(cudaGetDeviceCount(&devices) != cudaSuccess || devices == 0) { cudaGetLastError(); return bz_fail(BZ_ERR_NO_DEVICE, "no usable CUDA device (there is no CPU path)"); }
  BZ_CUDA(bz_launch_emit(*emitter, table, first, count, d_start, d_dir, d_patch, (cudaStream_t)stream));
  return BZ_OK;
}

// the BZ_JOB_TARGET description of a call (everything but the launch)
static int bz_target_job(bz_lens *lens, const float *d_start, const float *d_dir, uint64_t n, const bz_target *target, uint32_t max_crossings,
                         uint32_t *d_bins, uint32_t *d_landing, bz_stats *d_stats, bz_job &desc) {
  if(lens == nullptr || target == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  if(target->bins_u == 0u || target->bins_v == 0u || (uint64_t)target->bins_u * target->bins_v >= BZ_LAND_MISSED) { return bz_fail(BZ_ERR_INVALID, "bad bin counts"); }
  if(n != 0u && (d_start == nullptr || d_dir == nullptr || d_bins == nullptr)) { return bz_fail(BZ_ERR_INVALID, "null buffer"); }
  if(lens->mode == BZ_MODE_BRUTE) { return bz_fail(BZ_ERR_INVALID, "the illumination mode runs on the culled kernel only"); }
  std::memset(&desc, 0, sizeof(desc));
  desc.in_start = d_start; desc.in_dir = d_dir;
  desc.target = *target;
  // plane of the rectangle: unit normal of edge_u x edge_v (normalised like Eigen's normalized()), constant = n . origin
  const float *u = target->edge_u, *v = target->edge_v;
  float const cx = u[1] * v[2] - u[2] * v[1], cy = u[2] * v[0] - u[0] * v[2], cz = u[0] * v[1] - u[1] * v[0];
  float const length = std::sqrt(cx * cx + (cy * cy + cz * cz));
  if(!(length > 0.0f)) { return bz_fail(BZ_ERR_INVALID, "degenerate target rectangle"); }
  desc.target_plane[0] = cx / length; desc.target_plane[1] = cy / length; desc.target_plane[2] = cz / length;
  desc.target_plane[3] = desc.target_plane[0] * target->origin[0] + (desc.target_plane[1] * target->origin[1] + desc.target_plane[2] * target->origin[2]);
  desc.bins = d_bins;
  desc.out_landing = d_landing;
  desc.scene.max_crossings = max_crossings;
  desc.ray_count = n;
  desc.stats = (unsigned long long*)d_stats;
  return BZ_OK;
}

int bz_trace_to_target(bz_lens *lens, const float *d_start, const float *d_dir, uint64_t n, const bz_target *target, uint32_t max_crossings,
                       uint32_t *d_bins, uint32_t *d_landing, bz_stats *d_stats, void *stream) {
  bz_job desc;
  int rc = bz_target_job(lens, d_start, d_dir, n, target, max_crossings, d_bins, d_landing, d_stats, desc);
  if(rc != BZ_OK) { return rc; }
  bz_context *ctx = nullptr;
  rc = bz_context_for(lens, (cudaStream_t)stream, &ctx);
  if(rc != BZ_OK) { return rc; }
  std::lock_guard<std::mutex> hold(ctx->busy);
  return bz_run(ctx, BZ_JOB_TARGET, desc);
}

// emit -> pass -> group by hemisphere patch -> trace, on the context's stream
static int bz_illuminate_on(bz_context *ctx, const bz_emitter *emitter, const bz_target *target, uint32_t max_crossings, uint32_t *d_bins, bz_stats *d_stats) {
  bz_lens *lens = ctx->lens;
  uint64_t rays = 0u;
  int rc = bz_check_emitter(emitter, &rays);
  if(rc != BZ_OK) { return rc; }
  if(rays >= (1ull << 32)) { return bz_fail(BZ_ERR_INVALID, "more than 2^32 - 1 emitter rays in one call (the grouping cursors are 32-bit)"); }
  bz_belts table;
  rc = bz_make_belts(emitter->belts, table);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaSetDevice(lens->device));
  // scratch of the context: emitted rays, patch keys, group offsets, grouped rays
  uint32_t const ways = 64u;              // counters per patch (csrc/bz_illum.cu)
  size_t const groups = (size_t)table.patches * ways;
  size_t const keys = ((size_t)rays * 4u + 255u) / 256u * 256u, counts = ((size_t)(groups + 1u) * 4u + 255u) / 256u * 256u;
  if(ctx->d_illum == nullptr || ctx->illum_bytes < (size_t)rays * 48u + keys + counts) {
    if(ctx->d_illum != nullptr) { BZ_CUDA(cudaFree(ctx->d_illum)); ctx->d_illum = nullptr; ctx->illum_bytes = 0u; }
    BZ_CUDA(cudaMalloc(&ctx->d_illum, (size_t)rays * 48u + keys + counts));
    ctx->illum_bytes = (size_t)rays * 48u + keys + counts;
  }
  cudaStream_t const st = ctx->stream;
  float *start = (float*)ctx->d_illum, *dir = start + rays * 3u, *sorted_start = dir + rays * 3u, *sorted_dir = sorted_start + rays * 3u;
  uint32_t *patch = (uint32_t*)(sorted_dir + rays * 3u), *group = (uint32_t*)((char*)patch + keys);
  bz_job desc;
  rc = bz_target_job(lens, sorted_start, sorted_dir, rays, target, max_crossings, d_bins, nullptr, d_stats, desc);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(bz_launch_emit(*emitter, table, 0u, rays, start, dir, patch, st));
  // rays that pass beside the lens land on the target at once; the others are grouped by patch and traced
  BZ_CUDA(bz_launch_pass(lens->tables, desc, start, dir, patch, rays, st));
  BZ_CUDA(cudaMemsetAsync(group, 0, counts, st));
  BZ_CUDA(bz_launch_group_by_patch(patch, rays, table.patches, ways, group, start, dir, sorted_start, sorted_dir, lens->sm_count, st));
  desc.ray_count_dev = group + groups;
  rc = bz_run(ctx, BZ_JOB_TARGET, desc);
  ctx->launches += 5u;                   // emit, pass, count, scan, scatter
  return rc;
}

int bz_illuminate(bz_lens *lens, const bz_emitter *emitter, const bz_target *target, uint32_t max_crossings,
                  uint32_t *d_bins, bz_stats *d_stats, void *stream) {
  if(lens == nullptr) { return bz_fail(BZ_ERR_INVALID, "lens is null"); }
  bz_context *ctx = nullptr;
  int rc = bz_context_for(lens, (cudaStream_t)stream, &ctx);
  if(rc != BZ_OK) { return rc; }
  std::lock_guard<std::mutex> hold(ctx->busy);
  return bz_illuminate_on(ctx, emitter, target, max_crossings, d_bins, d_stats);
}

int bz_illuminate_host(bz_lens *lens, const bz_emitter *emitter, const bz_target *target, uint32_t max_crossings,
                       uint32_t *h_bins, bz_stats *h_stats) {
  if(lens == nullptr || target == nullptr || h_bins == nullptr) { return bz_fail(BZ_ERR_INVALID, "null argument"); }
  size_t const bytes = (size_t)target->bins_u * target->bins_v * 4u;
  if(bytes == 0u) { return bz_fail(BZ_ERR_INVALID, "bad bin counts"); }
  bz_host_call call(lens);
  if(call.rc != BZ_OK) { return call.rc; }
  bz_context *ctx = call.ctx;
  cudaStream_t const st = ctx->stream;
  int rc = bz_scratch(ctx, bytes + 128u);
  if(rc != BZ_OK) { return rc; }
  bz_stats *d_stats = (bz_stats*)ctx->d_scratch;
  uint32_t *d_bins = (uint32_t*)((uint8_t*)ctx->d_scratch + 128u);
  BZ_CUDA(cudaMemsetAsync(d_stats, 0, sizeof(bz_stats), st));
  BZ_CUDA(cudaMemcpyAsync(d_bins, h_bins, bytes, cudaMemcpyHostToDevice, st));      // accumulate into the caller's counts
  rc = bz_illuminate_on(ctx, emitter, target, max_crossings, d_bins, h_stats != nullptr ? d_stats : nullptr);
  if(rc != BZ_OK) { return rc; }
  BZ_CUDA(cudaMemcpyAsync(h_bins, d_bins, bytes, cudaMemcpyDeviceToHost, st));
  bz_stats fresh;
  if(h_stats != nullptr) { BZ_CUDA(cudaMemcpyAsync(&fresh, d_stats, sizeof(bz_stats), cudaMemcpyDeviceToHost, st)); }
  BZ_CUDA(cudaStreamSynchronize(st));
  if(h_stats != nullptr) {
    uint64_t *acc = (uint64_t*)h_stats;
    const uint64_t *add = (const uint64_t*)&fresh;
    for(size_t i = 0u; i < sizeof(bz_stats) / sizeof(uint64_t); ++i) { acc[i] += add[i]; }
  }
  return BZ_OK;
}

}  // extern "C"

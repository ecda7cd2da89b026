// bz_illum.cu -- illumination mode: emitter sampling and ray grouping.
//
//   bz_emit_kernel       UniformHemisphere::getRandom (reference/hostUtil.cpp:16-29) for a batch of emitter rays.  The
//                        reference evaluates acos / sin / cos in double (its ::acos(float) etc. resolve to the double
//                        overloads) and rounds to float; the kernel does the same with the CUDA FP64 functions, so the
//                        directions agree with a CPU restatement to the last float bit except where two correctly
//                        working double libraries differ in their last bit right at a float rounding boundary.
//                        The random numbers come from a counter-based generator (include/bz_capi.h), not ranlux24.
//   bz_patch_count / bz_patch_scan / bz_patch_scatter
//                        counting sort of the rays by hemisphere patch (README.md:176-186: one ray vector per patch, so
//                        that the rays traced together have similar directions) -- hand-written, no Thrust.
// The tracing itself is BZ_JOB_TARGET of csrc/bz_trace.cu.
#include <cuda_runtime.h>
#include <stdint.h>
#include "bz_device.h"
#include "bz_math.cuh"

#define BZ_PI 3.14159265358979323846f      // cgPi, reference/3dGeomUtil.h

__host__ __device__ __forceinline__ uint64_t bz_mix64(uint64_t z) {      // the output function of splitmix64
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

__device__ __forceinline__ float bz_uniform(uint64_t seed, uint64_t ray, uint32_t draw) {
  const uint64_t bits = bz_mix64(seed + (2ull * ray + draw + 1ull) * 0x9E3779B97F4A7C15ull);
  return (float)(bits >> 40) * 5.9604644775390625e-08f;      // 24 bits -> [0, 1), exact
}

__global__ void bz_emit_kernel(const bz_emitter em, const bz_belts belts, uint64_t first, uint64_t count, float *start, float *dir, uint32_t *patch) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if(i >= count) { return; }
  const uint64_t ray = first + i;
  const uint64_t point = ray / em.rays_per_point;
  const uint32_t pv = (uint32_t)(point / em.points_u), pu = (uint32_t)(point - (uint64_t)pv * em.points_u);
  const float fu = ((float)pu + 0.5f) / (float)em.points_u, fv = ((float)pv + 0.5f) / (float)em.points_v;
  const bzv3 s = bz_v(em.origin[0] + (em.edge_u[0] * fu + em.edge_v[0] * fv),
                      em.origin[1] + (em.edge_u[1] * fu + em.edge_v[1] * fv),
                      em.origin[2] + (em.edge_u[2] * fu + em.edge_v[2] * fv));
  // hostUtil.cpp:19-25
  const float u1 = bz_uniform(em.seed, ray, 0u), u2 = bz_uniform(em.seed, ray, 1u);
  const float incidence = (float)acos((double)u1);
  const float belt_radius = (float)sin((double)incidence);
  const float turn = u2 * (BZ_PI * 2.0f);
  bzv3 local = bz_v((float)cos((double)incidence), (float)((double)belt_radius * cos((double)turn)), (float)((double)belt_radius * sin((double)turn)));
  local = bz_normalized(local);
  // hostUtil.cpp:26-27 (the belt index is clamped: u1 == 0 would index one past the table in the reference)
  const uint32_t belt = min((uint32_t)(incidence / belts.belt_width), belts.belts - 1u);
  const uint32_t index = min(belts.first[belt] + (uint32_t)(turn / belts.width[belt]), belts.patches - 1u);
  // into the emitter's frame, then the Ray constructor's normalisation (3dGeomUtil.h:176-178)
  const bzv3 world = bz_v(em.normal[0] * local.x + (em.tangent_a[0] * local.y + em.tangent_b[0] * local.z),
                          em.normal[1] * local.x + (em.tangent_a[1] * local.y + em.tangent_b[1] * local.z),
                          em.normal[2] * local.x + (em.tangent_a[2] * local.y + em.tangent_b[2] * local.z));
  const bzv3 d = bz_normalized(world);
  start[i * 3u] = s.x; start[i * 3u + 1u] = s.y; start[i * 3u + 2u] = s.z;
  dir[i * 3u] = d.x; dir[i * 3u + 1u] = d.y; dir[i * 3u + 2u] = d.z;
  if(patch != nullptr) { patch[i] = index; }
}

// ---- counting sort by hemisphere patch ----
// README.md:176-186: one ray vector per patch, so that the rays traced together have similar directions.  Every patch has
// `ways` counters (key = patch * ways + ray mod ways; the order inside a patch does not matter): 4.2 M atomics on 660
// addresses took 0.46 ms, on 660 x 64 they take 0.05 ms.  Hand-written, no Thrust.
__device__ __forceinline__ uint32_t bz_group_key(const uint32_t *patch, uint64_t i, uint32_t ways) {
  return patch[i] * ways + (uint32_t)(i % ways);
}

__global__ void bz_patch_count_kernel(const uint32_t *patch, uint64_t count, uint32_t ways, uint32_t *counts) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for(uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    if(patch[i] != BZ_RAY_PASSED) { atomicAdd(counts + bz_group_key(patch, i, ways), 1u); }      // (bz_pass_kernel dealt with the others)
  }
}

// one CTA: counts[0..n) -> exclusive prefix sums in place, counts[n] = their total
__global__ void bz_patch_scan_kernel(uint32_t *counts, uint32_t n) {
  __shared__ uint32_t partial[1024];
  const uint32_t per = (n + blockDim.x - 1u) / blockDim.x;
  const uint32_t begin = min(threadIdx.x * per, n), end = min(begin + per, n);
  uint32_t sum = 0u;
  for(uint32_t i = begin; i < end; ++i) { sum += counts[i]; }
  partial[threadIdx.x] = sum;
  __syncthreads();
  if(threadIdx.x == 0u) {
    uint32_t run = 0u;
    for(uint32_t t = 0u; t < blockDim.x; ++t) { const uint32_t v = partial[t]; partial[t] = run; run += v; }
  }
  __syncthreads();
  uint32_t run = partial[threadIdx.x];
  for(uint32_t i = begin; i < end; ++i) { const uint32_t v = counts[i]; counts[i] = run; run += v; }
  if(threadIdx.x == blockDim.x - 1u) { counts[n] = run; }
}

__global__ void bz_patch_scatter_kernel(const uint32_t *patch, uint64_t count, uint32_t ways, uint32_t *cursor,
                                        const float *start, const float *dir, float *out_start, float *out_dir) {
  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for(uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    if(patch[i] == BZ_RAY_PASSED) { continue; }
    const uint64_t at = atomicAdd(cursor + bz_group_key(patch, i, ways), 1u);
    out_start[at * 3u] = start[i * 3u]; out_start[at * 3u + 1u] = start[i * 3u + 1u]; out_start[at * 3u + 2u] = start[i * 3u + 2u];
    out_dir[at * 3u] = dir[i * 3u]; out_dir[at * 3u + 1u] = dir[i * 3u + 1u]; out_dir[at * 3u + 2u] = dir[i * 3u + 2u];
  }
}

cudaError_t bz_launch_emit(const bz_emitter &em, const bz_belts &belts, uint64_t first, uint64_t count, float *start, float *dir, uint32_t *patch,
                           cudaStream_t stream) {
  if(count == 0u) { return cudaSuccess; }
  bz_emit_kernel<<<(unsigned)((count + 255u) / 256u), 256, 0, stream>>>(em, belts, first, count, start, dir, patch);
  return cudaGetLastError();
}

// rays grouped by patch: counts must hold patches * ways + 1 zeroed counters; they end up as the end offsets of the groups,
// the last one as the number of rays grouped (rays with the key BZ_RAY_PASSED are left out)
cudaError_t bz_launch_group_by_patch(const uint32_t *patch, uint64_t count, uint32_t patches, uint32_t ways, uint32_t *counts,
                                     const float *start, const float *dir, float *out_start, float *out_dir, int sm_count, cudaStream_t stream) {
  if(count == 0u) { return cudaSuccess; }
  const unsigned grid = (unsigned)sm_count * 8u;
  bz_patch_count_kernel<<<grid, 256, 0, stream>>>(patch, count, ways, counts);
  bz_patch_scan_kernel<<<1, 1024, 0, stream>>>(counts, patches * ways);
  bz_patch_scatter_kernel<<<grid, 256, 0, stream>>>(patch, count, ways, counts, start, dir, out_start, out_dir);
  return cudaGetLastError();
}

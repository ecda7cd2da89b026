/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the illumination mode's two non-tracing stages, used as the checker of
 * bz_emit_rays / bz_trace_to_target (include/bz_capi.h).
 * Pinned to the reference for the sampler: oracle/_ref holds the real UniformHemisphere (reference/hostUtil.cpp compiled from
 * /root/reference); tests/test_illumination.py feeds its getRandom() the draws of the generator below and compares belt
 * table, patch index and direction with this file bit for bit (golden copy: tests/golden/golden_hemisphere.npz).  The
 * reference's random stream (std::ranlux24_base) is replaced by the counter-based generator the C ABI defines -- both turn a
 * 24-bit draw o into o * 2^-24 -- and the reference has no target rectangle (README.md:161-194 was never written): the
 * landing stage is this repository's definition, unpinned.  This file restates
 *   UniformHemisphere::UniformHemisphere   reference/hostUtil.cpp:3-14
 *   UniformHemisphere::getRandom           reference/hostUtil.cpp:16-29   (acos / sin / cos in double, as ::acos(float) resolves)
 *   Plane::intersect                       reference/3dGeomUtil.h:279-296 (for the landing point on the target)
 * and the definitions of bz_emitter / bz_target.  Built with -ffp-contract=off.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include "bz_capi.h"

#define BZO_PI 3.14159265358979323846f     /* cgPi */

static uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

static float uniform01(uint64_t seed, uint64_t ray, uint32_t draw) {
  uint64_t bits = mix64(seed + (2ull * ray + draw + 1ull) * 0x9E3779B97F4A7C15ull);
  return (float)(bits >> 40) * 5.9604644775390625e-08f;
}

static float dot3(const float *a, const float *b) { return a[0] * b[0] + (a[1] * b[1] + a[2] * b[2]); }

static void normalize3(float *v) {     /* Eigen normalized(): z > 0 ? v / sqrt(z) : v */
  float z = dot3(v, v);
  if(z > 0.0f) {
    float n = sqrtf(z);
    v[0] = v[0] / n; v[1] = v[1] / n; v[2] = v[2] / n;
  }
}

/* hostUtil.cpp:3-14 -> patch count; width / first hold `belts` entries */
uint32_t bzo_hemisphere(uint32_t belts, float *width, uint32_t *first) {
  uint32_t total = 0u;
  for(uint32_t i = 0u; i < belts; ++i) {
    float angle = (2.0f * (float)i + 1.0f) / (4.0f * (float)belts) * BZO_PI;
    uint32_t now = (uint32_t)ceil((double)(4.0f * (float)belts) * sin((double)angle));
    if(width) { width[i] = BZO_PI * 2.0f / (float)now; }
    if(first) { first[i] = total; }
    total += now;
  }
  return total;
}

/* hostUtil.cpp:16-29 for rays [first, first + count) of the emission order */
void bzo_emit(const bz_emitter *em, uint64_t first_ray, uint64_t count, float *start, float *dir, uint32_t *patch) {
  float width[128];
  uint32_t first[128];
  uint32_t patches = bzo_hemisphere(em->belts, width, first);
  float belt_width = BZO_PI / 2.0f / (float)em->belts;
  for(uint64_t i = 0u; i < count; ++i) {
    uint64_t ray = first_ray + i;
    uint64_t point = ray / em->rays_per_point;
    uint32_t pv = (uint32_t)(point / em->points_u), pu = (uint32_t)(point - (uint64_t)pv * em->points_u);
    float fu = ((float)pu + 0.5f) / (float)em->points_u, fv = ((float)pv + 0.5f) / (float)em->points_v;
    for(int k = 0; k < 3; ++k) { start[i * 3u + k] = em->origin[k] + (em->edge_u[k] * fu + em->edge_v[k] * fv); }
    float u1 = uniform01(em->seed, ray, 0u), u2 = uniform01(em->seed, ray, 1u);
    float incidence = (float)acos((double)u1);
    float belt_radius = (float)sin((double)incidence);
    float turn = u2 * (BZO_PI * 2.0f);
    float local[3] = { (float)cos((double)incidence), (float)((double)belt_radius * cos((double)turn)), (float)((double)belt_radius * sin((double)turn)) };
    normalize3(local);
    uint32_t belt = (uint32_t)(incidence / belt_width);
    if(belt > em->belts - 1u) { belt = em->belts - 1u; }
    uint32_t index = first[belt] + (uint32_t)(turn / width[belt]);
    if(index > patches - 1u) { index = patches - 1u; }
    float world[3];
    for(int k = 0; k < 3; ++k) { world[k] = em->normal[k] * local[0] + (em->tangent_a[k] * local[1] + em->tangent_b[k] * local[2]); }
    normalize3(world);
    for(int k = 0; k < 3; ++k) { dir[i * 3u + k] = world[k]; }
    if(patch) { patch[i] = index; }
  }
}

/* landing bin of rays that ended in air (in_air[i] != 0), BZ_LAND_LOST otherwise; bins (nullable) is accumulated into */
void bzo_land(const bz_target *tg, const float *start, const float *dir, const uint8_t *in_air, uint64_t n, uint32_t *landing, uint32_t *bins) {
  const float *u = tg->edge_u, *v = tg->edge_v;
  float cx = u[1] * v[2] - u[2] * v[1], cy = u[2] * v[0] - u[0] * v[2], cz = u[0] * v[1] - u[1] * v[0];
  float length = sqrtf(cx * cx + (cy * cy + cz * cz));
  float plane[4] = { cx / length, cy / length, cz / length, 0.0f };
  plane[3] = plane[0] * tg->origin[0] + (plane[1] * tg->origin[1] + plane[2] * tg->origin[2]);
  for(uint64_t i = 0u; i < n; ++i) {
    uint32_t where = BZ_LAND_MISSED;
    const float *s = start + i * 3u, *d = dir + i * 3u;
    if(!in_air[i]) { where = BZ_LAND_LOST; }
    else {
      float cosine = dot3(d, plane);
      if(fabsf(cosine) >= 0.00001f) {
        float distance = (plane[3] - dot3(plane, s)) / cosine;
        if(distance > 0.0f) {
          float rel[3];
          for(int k = 0; k < 3; ++k) { rel[k] = s[k] + distance * d[k] - tg->origin[k]; }
          float su = dot3(rel, u) / dot3(u, u), sv = dot3(rel, v) / dot3(v, v);
          if(su >= 0.0f && su < 1.0f && sv >= 0.0f && sv < 1.0f) {
            uint32_t bu = (uint32_t)(su * (float)tg->bins_u), bv = (uint32_t)(sv * (float)tg->bins_v);
            if(bu > tg->bins_u - 1u) { bu = tg->bins_u - 1u; }
            if(bv > tg->bins_v - 1u) { bv = tg->bins_v - 1u; }
            where = bv * tg->bins_u + bu;
          }
        }
      }
    }
    landing[i] = where;
    if(bins && where < BZ_LAND_MISSED) { bins[where] += 1u; }
  }
}

/* TEST INFRASTRUCTURE ONLY -- plain-C CPU restatement of the reference's per-ray hot path.
 *
 * Parity status: PINNED against the real reference compiled here (oracle/_ref/libbzref.so, built by
 * oracle/build_ref.sh from /root/reference + oracle/eigen_shim + the one documented mPoint hoist):
 * tests/test_oracle.py requires bit-identical outputs on every config.  The reference's own tests
 * hold no golden vector for this path, so that build is the pin.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call this file; the
 * product (cuda-bezier-triangle-raytracer_b200/) never does.
 *
 * Input is the bz_patch table of include/bz_capi.h.  FP32, no fused multiply-add (build without
 * -mfma / -march=native), Eigen 3.3 evaluation orders (3-term reductions a0 + (a1 + a2)).
 * Each function cites the reference lines it follows (paths relative to /root/reference/reference).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <pthread.h>
#include "bz_capi.h"

#define BZO_LIMIT_THIS 0u   /* BezierTriangle::LimitPlaneIntersection::cThis, bezierTriangle.h:25-28 */
#define BZO_LIMIT_NONE 1u

#define BZO_FOLLOW0    0u   /* BezierIntersection::What, bezierTriangle.h:8-14 */
#define BZO_FOLLOW1    1u
#define BZO_FOLLOW2    2u
#define BZO_NONE       3u
#define BZO_INTERSECT  4u

typedef struct bzo_hit {
  uint32_t what;
  uint32_t patch;
  float    distance;
  float    point[3];
  float    normal[3];
  float    bary[3];
  float    cos_incidence;
  float    runner_up;      /* mesh_intersect: distance of the nearest accepted hit that did not win (FLT_MAX without one) */
} bzo_hit;

static const float cPlaneEpsilon      = 0.00001f;    /* Plane::csRayPlaneIntersectionEpsilon, 3dGeomUtil.h:219 */
static const float cMaxDistanceFromRay = 0.01f;      /* bezierTriangle.h:60 */
static const float cMinimalRayDistance = 1.0f;       /* bezierTriangle.h:61 */
static const float cEstimationEpsilon  = 0.000001f;  /* bezierTriangle.h:62 */
static const float cMaxSin2            = 0.99f;      /* bezierLens.h:16 */
static const float cMinSin2            = 1e-12f;     /* bezierLens.h:17 */

static float dot3(const float *a, const float *b) { return a[0] * b[0] + (a[1] * b[1] + a[2] * b[2]); }

static void mat_vec(const float m[3][3], const float *v, float *out) {
  for(int r = 0; r < 3; ++r) { out[r] = m[r][0] * v[0] + (m[r][1] * v[1] + m[r][2] * v[2]); }
}

static void normalize3(float *v) {               /* Eigen normalized(): z > 0 ? v / sqrt(z) : v */
  float z = dot3(v, v);
  if(z > 0.0f) {
    float n = sqrtf(z);
    v[0] = v[0] / n; v[1] = v[1] / n; v[2] = v[2] / n;
  }
}

/* Plane::distance, 3dGeomUtil.h:307 */
static float plane_distance(const float *plane, const float *p) { return dot3(p, plane) - plane[3]; }

/* Plane::project, 3dGeomUtil.h:303 */
static void plane_project(const float *plane, const float *p, float *out) {
  float d = dot3(p, plane) - plane[3];
  for(int k = 0; k < 3; ++k) { out[k] = p[k] - plane[k] * d; }
}

/* BezierTriangle::interpolate, bezierTriangle.cpp:105-121 */
static void interpolate(const bz_patch *t, const float *b, float *out) {
  float b0 = b[0], b1 = b[1], b2 = b[2];
  float b0_2 = b0 * b0, b1_2 = b1 * b1, b2_2 = b2 * b2;
  for(int k = 0; k < 3; ++k) {
    float inner = t->control[3][k] * b1 * b0_2 +
                  t->control[4][k] * b0 * b1_2 +
                  t->control[5][k] * b2 * b1_2 +
                  t->control[6][k] * b1 * b2_2 +
                  t->control[7][k] * b0 * b2_2 +
                  t->control[8][k] * b2 * b0_2;
    out[k] = t->control[0][k] * b0 * b0_2 +
             t->control[1][k] * b1 * b1_2 +
             t->control[2][k] * b2 * b2_2 +
             3.0f * inner +
             t->control[9][k] * b0 * b1 * b2 * 6.0f;
  }
}

/* BezierTriangle::getNormal, bezierTriangle.cpp:197-233; dirA = (1,0,-1), bezierTriangle.cpp:83 */
static void get_normal(const bz_patch *t, const float *b, float *out) {
  float b0 = b[0], b1 = b[1], b2 = b[2];
  float b0_2 = b0 * b0, b1_2 = b1 * b1, b2_2 = b2 * b2;
  float ca[3], cb[3];
  for(int k = 0; k < 3; ++k) {
    float c0 = t->control[0][k] * b0_2 + t->control[7][k] * b2_2 + t->control[4][k] * b1_2 +
               2.0f * (t->control[8][k] * b0 * b2 + t->control[3][k] * b0 * b1 + t->control[9][k] * b2 * b1);
    float c1 = t->control[1][k] * b1_2 + t->control[6][k] * b2_2 + t->control[3][k] * b0_2 +
               2.0f * (t->control[9][k] * b0 * b2 + t->control[4][k] * b0 * b1 + t->control[5][k] * b1 * b2);
    float c2 = t->control[2][k] * b2_2 + t->control[8][k] * b0_2 + t->control[5][k] * b1_2 +
               2.0f * (t->control[7][k] * b0 * b2 + t->control[6][k] * b1 * b2 + t->control[9][k] * b0 * b1);
    ca[k] = 1.0f * c0 + 0.0f * c1 + -1.0f * c2;
    cb[k] = t->dir_b[0] * c0 + t->dir_b[1] * c1 + t->dir_b[2] * c2;
  }
  out[0] = ca[1] * cb[2] - ca[2] * cb[1];
  out[1] = ca[2] * cb[0] - ca[0] * cb[2];
  out[2] = ca[0] * cb[1] - ca[1] * cb[0];
  normalize3(out);
}

/* |distance(pointOnRay)| - |distance(interpolate(Binv * project(pointOnRay)))|, bezierTriangle.cpp:137-143 */
static float bracket_diff(const bz_patch *t, const float *start, const float *dir, float at) {
  float on_ray[3], projected[3], bary[3], surface[3];
  for(int k = 0; k < 3; ++k) { on_ray[k] = start[k] + dir[k] * at; }
  plane_project(t->plane, on_ray, projected);
  mat_vec(t->bary_inv, projected, bary);
  interpolate(t, bary, surface);
  return fabsf(plane_distance(t->plane, on_ray)) - fabsf(plane_distance(t->plane, surface));
}

/* BezierTriangle::intersect, bezierTriangle.cpp:123-195 (Plane::intersect 3dGeomUtil.h:279-296, hoisted).
 * *stale is incremented when an iteration of the Newton loop meets |proj . n| < 1e-5, where the
 * reference reads an unassigned mPoint; this restatement then keeps the
 * previous iteration's projected point. */
static void patch_intersect(const bz_patch *t, const float *start, const float *dir, uint32_t limit, bzo_hit *result, uint64_t *stale) {
  result->what = BZO_NONE;
  float cosine = dot3(dir, t->plane);
  if(!(fabsf(cosine) >= cPlaneEpsilon)) { return; }
  float distance = (t->plane[3] - dot3(t->plane, start)) / cosine;
  if(!(distance > 0.0f)) { return; }
  float in_plane[3];
  for(int k = 0; k < 3; ++k) { in_plane[k] = start[k] + distance * dir[k]; }
  if(!(fabsf(distance) > -t->height_inside && fabsf(distance) > t->height_outside)) { return; }
  float bary[3];
  mat_vec(t->bary_inv, in_plane, bary);
  if(!(limit == BZO_LIMIT_NONE ||
       (bary[0] >= 0.0f && bary[0] <= 1.0f && bary[1] >= 0.0f && bary[1] <= 1.0f && bary[2] >= 0.0f && bary[2] <= 1.0f))) { return; }

  float distance_inside = t->height_inside / cosine;
  float distance_outside = t->height_outside / cosine;
  float closer = distance + (cosine > 0.0f ? distance_inside : distance_outside);
  float further = distance + (cosine > 0.0f ? distance_outside : distance_inside);
  float diff_closer = bracket_diff(t, start, dir, closer);
  float diff_further = bracket_diff(t, start, dir, further);
  float middle;
  float denom = diff_closer - diff_further;
  if(fabsf(denom) < cEstimationEpsilon) {
    middle = (closer + further) / 2.0f;
  }
  else {
    middle = (diff_closer * further - diff_further * closer) / denom;
  }
  float projection[3] = { t->plane[0], t->plane[1], t->plane[2] };
  float projected[3] = { 0.0f, 0.0f, 0.0f };
  float surface[3], normal[3];
  for(int i = 0; i < 4; ++i) {                                    /* csRootSearchIterations, bezierTriangle.h:58 */
    result->distance = middle;
    float on_ray[3];
    for(int k = 0; k < 3; ++k) { on_ray[k] = start[k] + dir[k] * middle; }
    float cos_projection = dot3(projection, t->plane);
    if(fabsf(cos_projection) >= cPlaneEpsilon) {
      float along = (t->plane[3] - dot3(t->plane, on_ray)) / cos_projection;
      for(int k = 0; k < 3; ++k) { projected[k] = on_ray[k] + along * projection[k]; }   /* sign-agnostic: the hoist */
    }
    else {
      ++*stale;
    }
    mat_vec(t->bary_inv, projected, result->bary);
    get_normal(t, result->bary, normal);
    interpolate(t, result->bary, surface);
    for(int k = 0; k < 3; ++k) { projection[k] = surface[k] - projected[k]; }
    normalize3(projection);
    float to_surface[3];
    for(int k = 0; k < 3; ++k) { to_surface[k] = surface[k] - start[k]; }
    middle = dot3(to_surface, normal) / dot3(dir, normal);
  }
  /* Ray::getDistance, 3dGeomUtil.h:182-190 */
  float to_surface[3], perpendicular[3];
  for(int k = 0; k < 3; ++k) { to_surface[k] = surface[k] - start[k]; }
  float along_ray = dot3(to_surface, dir);
  for(int k = 0; k < 3; ++k) { perpendicular[k] = to_surface[k] - along_ray * dir[k]; }
  float from_ray = sqrtf(dot3(perpendicular, perpendicular));
  if(from_ray > cMaxDistanceFromRay || result->distance < (further - closer) * cMinimalRayDistance) { return; }
  for(int k = 0; k < 3; ++k) {
    result->point[k] = surface[k];
    result->normal[k] = normal[k];
  }
  uint32_t outside = (plane_distance(t->divider[0], surface) < 0.0f ? 1u : 0u);
  outside |= (plane_distance(t->divider[1], surface) < 0.0f ? 2u : 0u);
  outside |= (plane_distance(t->divider[2], surface) < 0.0f ? 4u : 0u);
  if(outside == 1u) { result->what = BZO_FOLLOW0; }
  else if(outside == 2u) { result->what = BZO_FOLLOW1; }
  else if(outside == 4u) { result->what = BZO_FOLLOW2; }
  else {
    result->what = BZO_INTERSECT;
    result->cos_incidence = dot3(dir, normal);
  }
}

/* BezierMesh::intersect, bezierMesh.cpp:206-227.  work[0] += stage-A tests, work[1] += stage-B
 * evaluations (patches that passed the gate at bezierTriangle.cpp:126-131), work[2] += stale. */
static void mesh_intersect(const bz_patch *patches, uint32_t count, const float *start, const float *dir, bzo_hit *result, uint64_t *work) {
  result->what = BZO_NONE;
  result->patch = BZ_NO_PATCH;
  result->distance = 3.402823466e+38f;
  float runner_up = 3.402823466e+38f;      /* not part of the reference: lets the tests COUNT edge-ambiguous rays (SURVEY.md section 4 iii) */
  for(uint32_t i = 0u; i < count; ++i) {
    bzo_hit candidate;
    uint32_t evaluated = i;
    patch_intersect(patches + i, start, dir, BZO_LIMIT_THIS, &candidate, work + 2);
    if(candidate.what == BZO_FOLLOW0 || candidate.what == BZO_FOLLOW1 || candidate.what == BZO_FOLLOW2) {
      evaluated = patches[i].neighbours[candidate.what];
      patch_intersect(patches + evaluated, start, dir, BZO_LIMIT_NONE, &candidate, work + 2);
    }
    if(candidate.what == BZO_INTERSECT && candidate.distance < result->distance) {
      if(result->distance < runner_up) { runner_up = result->distance; }
      *result = candidate;
      result->patch = evaluated;
    }
    else if(candidate.what == BZO_INTERSECT && candidate.distance < runner_up) { runner_up = candidate.distance; }
  }
  result->runner_up = runner_up;
  work[0] += count;
}

/* BezierLens::refract, bezierLens.cpp:4-34.  Returns the status; on cNone the out ray repeats the in ray. */
static uint32_t lens_refract(const bz_patch *patches, uint32_t count, float ri, const float *start, const float *dir, uint32_t expected,
                             float *out_start, float *out_dir, uint32_t *out_patch, uint64_t *work) {
  bzo_hit hit;
  uint32_t status = BZ_STATUS_NONE;
  float new_dir[3] = { dir[0], dir[1], dir[2] };
  mesh_intersect(patches, count, start, dir, &hit, work);
  if(hit.what == BZO_INTERSECT) {
    status = (hit.cos_incidence < 0.0f ? BZ_STATUS_INSIDE : BZ_STATUS_OUTSIDE);
    float factor = (status == BZ_STATUS_INSIDE ? 1.0f / ri : ri);
    float sin2 = factor * factor * (1.0f - hit.cos_incidence * hit.cos_incidence);
    if(sin2 < cMaxSin2) {
      if(sin2 > cMinSin2) {
        float direction_factor = (status == BZ_STATUS_INSIDE ? 1.0f : -1.0f);
        float cos1 = fabsf(hit.cos_incidence);
        double cos2 = sqrt((double)(1.0f - sin2));          /* ::sqrt(float) resolves to double sqrt(double), SURVEY.md fact 6 */
        float along = (float)((double)(factor * cos1) - cos2);
        for(int k = 0; k < 3; ++k) { new_dir[k] = dir[k] * factor + (hit.normal[k] * direction_factor) * along; }
        normalize3(new_dir);
      }
    }
    else {
      status = BZ_STATUS_NONE;                              /* total reflection or close to it: ignored, bezierLens.cpp:26 */
    }
  }
  status = (status == expected ? status : BZ_STATUS_NONE);
  for(int k = 0; k < 3; ++k) {
    out_start[k] = (status == BZ_STATUS_NONE ? start[k] : hit.point[k]);
    out_dir[k] = (status == BZ_STATUS_NONE ? dir[k] : new_dir[k]);
  }
  if(out_patch != 0) { *out_patch = (hit.what == BZO_INTERSECT ? hit.patch : BZ_NO_PATCH); }
  return status;
}

/* ---------------------------------------------------------------- batch entry points ---- */

typedef struct bzo_job {
  const bz_patch *patches; uint32_t count; float ri;
  const float *start; const float *dir; const uint32_t *expected; uint64_t n;
  float *out_start; float *out_dir; uint32_t *out_status; uint32_t *out_patch;
  uint32_t *out_what; float *out_distance; float *out_point; float *out_normal; float *out_cos; float *out_bary; float *out_runner_up;
  int refract; int thread; int threads; uint64_t work[3];
} bzo_job;

static void *bzo_worker(void *arg) {
  bzo_job *job = (bzo_job*)arg;
  uint64_t per = (job->n + (uint64_t)job->threads - 1u) / (uint64_t)job->threads;
  uint64_t begin = per * (uint64_t)job->thread;
  uint64_t end = begin + per < job->n ? begin + per : job->n;
  for(uint64_t i = begin; i < end; ++i) {
    if(job->refract) {
      job->out_status[i] = lens_refract(job->patches, job->count, job->ri, job->start + i * 3u, job->dir + i * 3u, job->expected[i],
                                        job->out_start + i * 3u, job->out_dir + i * 3u, job->out_patch ? job->out_patch + i : 0, job->work);
    }
    else {
      bzo_hit hit;
      mesh_intersect(job->patches, job->count, job->start + i * 3u, job->dir + i * 3u, &hit, job->work);
      int ok = (hit.what == BZO_INTERSECT);
      job->out_what[i] = ok ? BZ_WHAT_INTERSECT : BZ_WHAT_NONE;
      if(job->out_patch) { job->out_patch[i] = ok ? hit.patch : BZ_NO_PATCH; }
      if(job->out_distance) { job->out_distance[i] = ok ? hit.distance : 0.0f; }
      if(job->out_cos) { job->out_cos[i] = ok ? hit.cos_incidence : 0.0f; }
      if(job->out_runner_up) { job->out_runner_up[i] = hit.runner_up; }
      for(int k = 0; k < 3; ++k) {
        if(job->out_point) { job->out_point[i * 3u + k] = ok ? hit.point[k] : 0.0f; }
        if(job->out_normal) { job->out_normal[i * 3u + k] = ok ? hit.normal[k] : 0.0f; }
        if(job->out_bary) { job->out_bary[i * 3u + k] = ok ? hit.bary[k] : 0.0f; }
      }
    }
  }
  return 0;
}

static void bzo_run(bzo_job *proto, int threads, uint64_t *work) {
  if(threads < 1) { threads = 1; }
  if(threads > 256) { threads = 256; }
  bzo_job jobs[256];
  pthread_t ids[256];
  for(int t = 0; t < threads; ++t) {
    jobs[t] = *proto;
    jobs[t].thread = t;
    jobs[t].threads = threads;
    memset(jobs[t].work, 0, sizeof(jobs[t].work));
    if(threads > 1) { pthread_create(&ids[t], 0, bzo_worker, &jobs[t]); } else { bzo_worker(&jobs[t]); }
  }
  for(int t = 0; t < threads; ++t) {
    if(threads > 1) { pthread_join(ids[t], 0); }
    if(work) { for(int k = 0; k < 3; ++k) { work[k] += jobs[t].work[k]; } }
  }
}

/* work (nullable): [0] += stage-A tests, [1] unused, [2] += stale projections (see patch_intersect) */
void bzo_refract_batch(const bz_patch *patches, uint32_t count, float ri, const float *start, const float *dir, const uint32_t *expected,
                       uint64_t n, float *out_start, float *out_dir, uint32_t *out_status, uint32_t *out_patch, int threads, uint64_t *work) {
  bzo_job job;
  memset(&job, 0, sizeof(job));
  job.patches = patches; job.count = count; job.ri = ri; job.start = start; job.dir = dir; job.expected = expected; job.n = n;
  job.out_start = out_start; job.out_dir = out_dir; job.out_status = out_status; job.out_patch = out_patch; job.refract = 1;
  bzo_run(&job, threads, work);
}

void bzo_intersect_batch(const bz_patch *patches, uint32_t count, const float *start, const float *dir, uint64_t n,
                         uint32_t *out_what, uint32_t *out_patch, float *out_distance, float *out_point, float *out_normal,
                         float *out_cos, float *out_bary, int threads, uint64_t *work) {
  bzo_job job;
  memset(&job, 0, sizeof(job));
  job.patches = patches; job.count = count; job.start = start; job.dir = dir; job.n = n;
  job.out_what = out_what; job.out_patch = out_patch; job.out_distance = out_distance; job.out_point = out_point;
  job.out_normal = out_normal; job.out_cos = out_cos; job.out_bary = out_bary; job.refract = 0;
  bzo_run(&job, threads, work);
}

/* bzo_intersect_batch plus, per ray, the distance of the nearest accepted hit that lost (FLT_MAX without one): a ray whose two
 * best hits lie within 1e-3 of the shape size of each other sits on a patch edge (SURVEY.md section 4 iii) */
void bzo_intersect_batch_runner_up(const bz_patch *patches, uint32_t count, const float *start, const float *dir, uint64_t n,
                                   uint32_t *out_what, uint32_t *out_patch, float *out_distance, float *out_runner_up, int threads) {
  bzo_job job;
  memset(&job, 0, sizeof(job));
  job.patches = patches; job.count = count; job.start = start; job.dir = dir; job.n = n;
  job.out_what = out_what; job.out_patch = out_patch; job.out_distance = out_distance; job.out_runner_up = out_runner_up; job.refract = 0;
  bzo_run(&job, threads, 0);
}

/* one patch, one ray: returns What (0..4) */
uint32_t bzo_patch_intersect(const bz_patch *patch, const float *start, const float *dir, uint32_t limit,
                             float *out_distance, float *out_point, float *out_normal, float *out_cos, float *out_bary) {
  bzo_hit hit;
  uint64_t stale = 0u;
  memset(&hit, 0, sizeof(hit));
  patch_intersect(patch, start, dir, limit, &hit, &stale);
  if(hit.what != BZO_NONE) {
    *out_distance = hit.distance;
    *out_cos = (hit.what == BZO_INTERSECT ? hit.cos_incidence : 0.0f);
    for(int k = 0; k < 3; ++k) { out_point[k] = hit.point[k]; out_normal[k] = hit.normal[k]; out_bary[k] = hit.bary[k]; }
  }
  return hit.what;
}

void bzo_interpolate(const bz_patch *patch, float b0, float b1, float b2, float *out) {
  float b[3] = { b0, b1, b2 };
  interpolate(patch, b, out);
}

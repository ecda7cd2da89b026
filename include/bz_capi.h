/* bz_capi.h -- C ABI of the B200-native Bezier-triangle intersect/refract path.
 *
 * This is the drop-in boundary for ONE hot path of balazs-bamer/cuda-bezier-triangle-raytracer:
 *   BezierLens::refract          reference/bezierLens.h:27,  bezierLens.cpp:4-34
 *   -> BezierMesh::intersect     reference/bezierMesh.h:37,  bezierMesh.cpp:206-227
 *   -> BezierTriangle::intersect reference/bezierTriangle.h:105, bezierTriangle.cpp:123-195
 * plus the ray-generation / bounce-loop / shading stages the reference never wrote
 * (reference/rayTracer.cpp is empty; the loop shape follows reference/test.cpp:376-398).
 *
 * Plain pointers and sizes only.  Every function returns BZ_OK (0) or a negative error code;
 * bz_last_error() gives the text for the calling thread.  There is no CPU fallback: every entry
 * point that computes needs a CUDA device and fails loudly without one.
 *
 * Pointers named d_* are DEVICE pointers, h_* are HOST pointers.  `stream` is a cudaStream_t
 * passed as void* (NULL = the legacy default stream).  Calls taking a stream are asynchronous
 * on it; *_host calls are synchronous and include the host<->device copies.
 */
#ifndef BZ_CAPI_H
#define BZ_CAPI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BZ_OK                 0
#define BZ_ERR_INVALID       -1   /* bad argument */
#define BZ_ERR_CUDA          -2   /* CUDA runtime error, see bz_last_error() */
#define BZ_ERR_NO_DEVICE     -3   /* no usable CUDA device: there is no CPU path */
#define BZ_ERR_PREPROCESS    -4   /* mesh preprocessing threw (e.g. "Vertex on edge detected.", reference/mesh.cpp:204) */
#define BZ_ERR_OVERFLOW      -5   /* a device-side work list overflowed (bz_lens_check) */
#define BZ_ERR_COMM          -6   /* NCCL missing or failing (bz_comm_*, bz_gather_rows) */

/* RefractionResult, reference/bezierLens.h:7-11 */
#define BZ_STATUS_NONE        0u
#define BZ_STATUS_INSIDE      1u
#define BZ_STATUS_OUTSIDE     2u

/* BezierIntersection::What, reference/bezierTriangle.h:8-14 (batch intersect reports NONE or INTERSECT) */
#define BZ_WHAT_NONE          3u
#define BZ_WHAT_INTERSECT     4u

#define BZ_NO_PATCH  0xffffffffu

/* POD mirror of the per-patch record BezierTriangle holds (reference/bezierTriangle.h:64-80).
 * mBezierDerivativeDirectionVectorA is the constant (1,0,-1) (bezierTriangle.cpp:83) and is not
 * stored.  256 bytes, 16-byte aligned. */
typedef struct bz_patch {
  float    plane[4];         /* mUnderlyingPlane: unit normal xyz, constant        (h:64) */
  float    divider[3][4];    /* mNeighbourDividerPlanes[i]: normal xyz, constant   (h:65) */
  uint32_t neighbours[3];    /* mNeighbours, patch indices after Clough-Tocher     (h:68) */
  float    height_inside;    /* mHeightInside  (< 0, already x1.3333)              (h:77) */
  float    control[10][3];   /* mControlPoints, index map h:31-51                  (h:73) */
  float    height_outside;   /* mHeightOutside (> 0, already x1.3333)              (h:78) */
  float    dir_b[3];         /* mBezierDerivativeDirectionVectorB                  (h:80) */
  float    bary_inv[3][3];   /* mBarycentricInverse, row-major: b = bary_inv * P   (h:75) */
  float    reserved[1];
} bz_patch;

/* Camera of the headless renderer (no reference counterpart; definitions in DESIGN.md).
 * Sample (px,py,sx,sy), strata S:  fu = px + (sx+0.5)/S, fv = py + (sy+0.5)/S,
 *   a = (fu*(2/width) - 1)*half_w,  b = (1 - fv*(2/height))*half_h
 * pinhole: start = eye,                     dir = normalize(forward + (right*a + up*b))
 * ortho:   start = eye + (right*a + up*b),  dir = normalize(forward)                      */
typedef struct bz_camera {
  float    eye[3];
  float    forward[3];
  float    right[3];
  float    up[3];
  float    half_w;
  float    half_h;
  uint32_t width;
  uint32_t height;
  uint32_t strata;           /* samples per pixel = strata*strata, fixed stratified grid, no jitter */
  uint32_t ortho;            /* 0 = pinhole, 1 = orthographic */
} bz_camera;

/* Scene behind the lens + bounce-loop limit. */
typedef struct bz_scene {
  float    bg_plane[4];      /* background plane: unit normal xyz, constant (Plane semantics, 3dGeomUtil.h:279-296) */
  float    bg_axis_a[3];     /* checker coordinate axes: ia = floor((P . axis_a)/cell) */
  float    bg_axis_b[3];
  float    cell;
  uint32_t value_dark;       /* checker values, 0..255 */
  uint32_t value_bright;
  uint32_t value_lost;       /* ray entered and never left, or never reached the background */
  uint32_t max_crossings;    /* bounce loop: at most this many refract() calls per primary ray */
} bz_scene;

/* Work counters: executed work, never brute-force-equivalent work. */
typedef struct bz_stats {
  uint64_t primary_rays;
  uint64_t refract_calls;    /* surface-crossing attempts = BezierLens::refract calls */
  uint64_t stage_a;          /* plane + barycentric tests executed  (bezierTriangle.cpp:124-131) */
  uint64_t stage_b;          /* Newton evaluations executed incl. follow hops (cpp:132-194) */
  uint64_t snell;            /* calls that reached the Snell block (bezierLens.cpp:8-27) */
  uint64_t entered;          /* refract() returned cInside */
  uint64_t exited;           /* refract() returned cOutside */
  uint64_t lost;             /* primary rays shaded value_lost */
  uint64_t node_visits;      /* binary cull-hierarchy nodes visited by the per-lane walk (0 in BZ_MODE_BRUTE) */
  uint64_t packet_nodes;     /* 32-children nodes visited by the warp-cooperative packet walk */
  uint64_t packet_walks;     /* packet walks that completed (one per group of <= 32 rays and refract() call) */
  uint64_t lane_walk_calls;  /* refract() calls whose packet was not coherent enough and walked per lane */
} bz_stats;

typedef struct bz_lens bz_lens;
typedef struct bz_mesh bz_mesh;       /* host object: Mesh,       reference/mesh.h:18-133 */
typedef struct bz_bezier bz_bezier;   /* host object: BezierMesh, reference/bezierMesh.h:10-45 */

/* intersect strategy of a lens handle */
#define BZ_MODE_BRUTE   0u   /* every ray tests every patch, reference loop order (bezierMesh.cpp:211) */
#define BZ_MODE_CULLED  1u   /* exact candidate-set-preserving cull (warp-cooperative packet walk, per-lane walk for incoherent packets), then the same stage A/B */
#define BZ_MODE_LANES   2u   /* as BZ_MODE_CULLED with the packet walk switched off: every ray walks the hierarchy on its own lane (test / comparison arm) */
#define BZ_MODE_FAST    3u   /* BZ_MODE_CULLED's kernel built with FMA contraction, MUFU reciprocal / reciprocal square root and FP32 cos2: NOT bit-exact;
                                held to north_star's tolerance instead (status exact, patch index exact away from patch edges, points and directions
                                within 1e-5 relative, >= 90 % of the pixels identical).  Ray generation and shading are the same bits in every mode. */

const char *bz_last_error(void);
int  bz_device_count(int *out_count);
/* size/offset self-check so a foreign-language binding can verify its struct mirrors */
uint32_t bz_sizeof_patch(void);
uint32_t bz_sizeof_camera(void);
uint32_t bz_sizeof_scene(void);
uint32_t bz_sizeof_stats(void);

/* ---- host preprocessing (C++ classes bz::Mesh / bz::BezierMesh in cuda-bezier-triangle-raytracer_b200/host) ----
 * Mesh, reference/mesh.h:50-100.  Triangles are n x 3 corners x 3 floats.  Preprocessing errors
 * (e.g. the reference's `throw "Vertex on edge detected."`, mesh.cpp:204; std::runtime_error from the
 * STL reader; std::out_of_range when standardizeNormals was skipped, bezierMesh.cpp:17-18) come back
 * as BZ_ERR_PREPROCESS with the text in bz_last_error(). */
int  bz_mesh_create(bz_mesh **out_mesh);
int  bz_mesh_destroy(bz_mesh *mesh);
int  bz_mesh_clone(const bz_mesh *mesh, bz_mesh **out_mesh);
int  bz_mesh_make_ellipsoid(bz_mesh *mesh, int32_t sectors, int32_t belts, const float size[3]);   /* mesh.h:99 */
int  bz_mesh_make_unit_sphere(bz_mesh *mesh, int32_t sectors, int32_t belts);                        /* mesh.h:100 */
int  bz_mesh_read_stl(bz_mesh *mesh, const char *path);                                              /* mesh.cpp:399-415 */
int  bz_mesh_write_stl(const bz_mesh *mesh, const char *path);                                       /* mesh.cpp:419-430 */
int  bz_mesh_set_triangles(bz_mesh *mesh, const float *triangles, uint32_t count);
int  bz_mesh_size(const bz_mesh *mesh, uint32_t *out_count);
int  bz_mesh_get_triangles(const bz_mesh *mesh, float *out_triangles);
int  bz_mesh_transform(bz_mesh *mesh, const float matrix_row_major[9], const float displacement[3]); /* mesh.cpp:361-367 */
int  bz_mesh_split_triangles(bz_mesh *mesh, int32_t divisor);                                        /* mesh.cpp:389-395 */
int  bz_mesh_split_triangles_by_side(bz_mesh *mesh, float max_side);                                 /* mesh.cpp:375-385 */
int  bz_mesh_standardize_vertices(bz_mesh *mesh);                                                    /* mesh.cpp:72-91 */
int  bz_mesh_standardize_normals(bz_mesh *mesh);                                                     /* mesh.cpp:310-357 */
/* getFace2neighbours (mesh.h:68): out_fellow n x 3 u32, out_starts n x 3 u8; the count is 0 before standardizeNormals */
int  bz_mesh_neighbour_count(const bz_mesh *mesh, uint32_t *out_count);
int  bz_mesh_get_face2neighbours(const bz_mesh *mesh, uint32_t *out_fellow, uint8_t *out_starts);
/* getVertex2averageNormals (mesh.h:69) looked up for every corner: n x 3 x 3 floats */
int  bz_mesh_get_corner_normals(const bz_mesh *mesh, float *out_normals);

/* BezierMesh(Mesh const&), reference/bezierMesh.h:24, bezierMesh.cpp:4-34 */
int  bz_bezier_create(const bz_mesh *mesh, bz_bezier **out_bezier);
int  bz_bezier_destroy(bz_bezier *bezier);
int  bz_bezier_size(const bz_bezier *bezier, uint32_t *out_count);
int  bz_bezier_get_patches(const bz_bezier *bezier, bz_patch *out_patches);
int  bz_bezier_interpolate(const bz_bezier *bezier, int32_t divisor, bz_mesh **out_mesh);           /* bezierMesh.cpp:55-66 */
int  bz_bezier_split_thick(const bz_bezier *bezier, bz_mesh **out_mesh);                            /* bezierMesh.cpp:79-198 */
int  bz_lens_create_from_bezier(const bz_bezier *bezier, float refractive_index, int device, bz_lens **out_lens);

/* BezierLens(float aRi, BezierMesh const&)  -- reference/bezierLens.h:23-24.
 * Uploads the patch records to `device`, repacks them into SoA float4 streams and builds the
 * cull hierarchy.  h_patches is read during the call only. */
int  bz_lens_create(const bz_patch *h_patches, uint32_t patch_count, float refractive_index, int device, bz_lens **out_lens);
int  bz_lens_destroy(bz_lens *lens);
int  bz_lens_set_mode(bz_lens *lens, uint32_t mode);
int  bz_lens_patch_count(const bz_lens *lens, uint32_t *out_count);
/* How bz_render* runs the tiles that can meet the lens (BZ_MODE_CULLED / BZ_MODE_FAST; the results are the same bits):
 *   BZ_PIPELINE_PERSISTENT  one persistent warp-scheduled kernel does walk, stage A, stage B and resolve (csrc/bz_trace.cu)
 *   BZ_PIPELINE_WAVEFRONT   three kernels per surface crossing over global ray / candidate lists (csrc/bz_wave.cu)
 *   BZ_PIPELINE_AUTO        the default: the wavefront kernels for frames of the parity build with at least 2^19 ray slots on lens tiles and
 *                           at most 4 crossings, while their lists (36 bytes per ray slot + 3 x 52 bytes of candidate list per slot; 8.6 GB
 *                           for the 4K x 16 spp headline frame) fit into a quarter of the free device memory; else the persistent kernel
 * The batch entry points (bz_refract_batch, bz_intersect_batch, bz_trace_to_target) always use the persistent kernel. */
#define BZ_PIPELINE_AUTO        0u
#define BZ_PIPELINE_PERSISTENT  1u
#define BZ_PIPELINE_WAVEFRONT   2u
int  bz_lens_set_pipeline(bz_lens *lens, uint32_t pipeline);
/* Ray slots per chunk of the wavefront pipeline (default 64 Mi: a frame is one chunk when it fits; device memory: 36 bytes per slot +
 * 3 x 52 bytes of candidate list per slot, allocated for the largest chunk a context has seen). */
int  bz_lens_set_wave_chunk(bz_lens *lens, uint32_t ray_slots);
/* Synchronises with the work in flight on the handle and reports (once) a device-side overflow of the wavefront pipeline's
 * candidate list as BZ_ERR_OVERFLOW; the *_host render calls do this themselves. */
int  bz_lens_check(bz_lens *lens);
/* Threads and streams.  The tables of a lens are read-only after bz_lens_create; everything a call writes besides the caller's
 * buffers (pixel accumulator, work counters, tile plan, wavefront lists, staging, timing events) lives in a CONTEXT the lens keeps per
 * stream, made on the first call that names the stream.  So, like BezierLens::refract (const and re-entrant, reference/bezierLens.h:27):
 *   - calls on DIFFERENT streams of one lens may overlap, from one host thread or several;
 *   - calls on the same stream are ordered by the stream (two threads enqueueing on one stream take turns);
 *   - the *_host calls each borrow a private context with a non-blocking stream of its own, so any number of threads may call them
 *     side by side;
 *   - bz_lens_set_mode / bz_lens_set_pipeline / bz_lens_set_wave_chunk change the lens for all later calls: do not race them with calls.
 * bz_lens_release_stream waits for the stream's work and frees its context (call it before destroying a stream the lens has seen). */
int  bz_lens_release_stream(bz_lens *lens, void *stream);

/* BezierLens::refract over a batch  -- reference/bezierLens.h:27, bezierLens.cpp:4-34.
 * start/dir: n x 3 floats (dir must be unit length, as Ray::Ray leaves it, 3dGeomUtil.h:176-178).
 * expected: n x u32 (BZ_STATUS_*).  out_status == BZ_STATUS_NONE leaves out_start/out_dir of that
 * ray equal to the input ray.  d_out_patch (nullable): index of the hit patch or BZ_NO_PATCH. */
int  bz_refract_batch(bz_lens *lens, const float *d_start, const float *d_dir, const uint32_t *d_expected, uint64_t n,
                      float *d_out_start, float *d_out_dir, uint32_t *d_out_status, uint32_t *d_out_patch, void *stream);
int  bz_refract_batch_host(bz_lens *lens, const float *h_start, const float *h_dir, const uint32_t *h_expected, uint64_t n,
                           float *h_out_start, float *h_out_dir, uint32_t *h_out_status, uint32_t *h_out_patch);

/* BezierMesh::intersect over a batch  -- reference/bezierMesh.h:37, bezierMesh.cpp:206-227.
 * Outputs (all nullable except d_out_what): what (BZ_WHAT_*), patch, distance, point[3],
 * normal[3], cos incidence, barycentric[3]. */
int  bz_intersect_batch(bz_lens *lens, const float *d_start, const float *d_dir, uint64_t n,
                        uint32_t *d_out_what, uint32_t *d_out_patch, float *d_out_distance, float *d_out_point,
                        float *d_out_normal, float *d_out_cos, float *d_out_bary, void *stream);
int  bz_intersect_batch_host(bz_lens *lens, const float *h_start, const float *h_dir, uint64_t n,
                             uint32_t *h_out_what, uint32_t *h_out_patch, float *h_out_distance, float *h_out_point,
                             float *h_out_normal, float *h_out_cos, float *h_out_bary);

/* K1: primary rays of image rows [row0,row1), pixel-major then sample-major:
 * ray index = ((py-row0)*width + px)*spp + (sy*strata + sx).  Outputs n x 3 floats each. */
int  bz_generate_rays(const bz_camera *camera, uint32_t row0, uint32_t row1, float *d_start, float *d_dir, void *stream);

/* K1+K2+K3: render rows [row0,row1) into d_rgba ((row1-row0) x width x 4 bytes, row-major,
 * R=G=B=value, A=255).  d_stats (nullable) is ACCUMULATED into (device memory, caller zeroes).
 * In BZ_MODE_CULLED the call plans the frame: tiles of pixels onto which no box of the cull hierarchy projects
 * are shaded by a separate kernel (their rays count in primary_rays and refract_calls, they do no walk), the
 * other tiles are traced.  The plan of the last frame is kept in the stream's context (see "Threads and streams"). */
int  bz_render(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t row1,
               uint8_t *d_rgba, bz_stats *d_stats, void *stream);
int  bz_render_host(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t row1,
                    uint8_t *h_rgba, bz_stats *h_stats);

/* bz_render without any allocation.  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same pixels and
 * same d_stats counts of stage A / B / Snell work as bz_render on the same lens (any mode; the pipeline setting is ignored: always the
 * persistent kernel, and no tile plan, so rays beside the lens are traced instead of going to the sky kernel -- slower on frames that are
 * mostly background, and node_visits / packet counters are larger).  The render's scratch (work counter + one u32 accumulator per pixel)
 * is a static device array of 256 + 4 * BZ_STATIC_MAX_PIXELS bytes (8 MiB: a 1920 x 1080 band) that every GPU loading the library holds
 * until exit; (row1 - row0) * width beyond BZ_STATIC_MAX_PIXELS is BZ_ERR_INVALID (render in row bands, or use bz_render).  The call issues
 * memsets and kernels on `stream` and calls no cudaMalloc / cudaFree; it creates no context for the stream and leaves bz_last_kernel_ms /
 * bz_last_kernel_launches untouched.  ALL calls, from any lens, thread or stream, share the one array per GPU: calls that overlap on a
 * GPU race (serialise them on one stream, or with events); the array keeps the last call's contents and each call clears what it uses. */
#define BZ_STATIC_MAX_PIXELS 2097152u
int  bz_render_static(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t row0, uint32_t row1,
                      uint8_t *d_rgba, bz_stats *d_stats, void *stream);

/* Row bands for N GPUs (north_star subsystem 4): the image is cut into stripes of `stripe_rows` rows
 * dealt round-robin to `ranks` ranks; rank r renders stripes r, r+ranks, ... into a compact local
 * buffer (its rows in image order).  bz_striped_rows gives that buffer's row count. */
uint32_t bz_striped_rows(uint32_t height, uint32_t stripe_rows, uint32_t rank, uint32_t ranks);
int  bz_render_striped(bz_lens *lens, const bz_camera *camera, const bz_scene *scene, uint32_t stripe_rows, uint32_t rank,
                       uint32_t ranks, uint8_t *d_rgba, bz_stats *d_stats, void *stream);

/* The image gather of the row-sharded render (north_star subsystem 4: "NCCL gather over NVLink"; SURVEY.md section 8(b): bz_gather_rows).
 * One process per GPU.  Rank 0 makes an id (bz_comm_unique_id) and ships its 128 bytes to the other ranks by any means (MPI, a socket,
 * a file, torch.distributed); every rank then calls bz_comm_create with its rank -- a collective call (ncclCommInitRank).  NCCL is
 * loaded at run time; without it these calls return BZ_ERR_COMM and everything else still works.
 * bz_gather_rows: every rank passes its band of bz_render_striped (bz_striped_rows(height, stripe_rows, rank, ranks) rows); the root
 * ends up with d_rgba = height x width x 4 bytes in image order; d_rgba is ignored on the other ranks.  Asynchronous on `stream` on
 * every rank; every rank of the communicator calls it with the same width, height, stripe_rows and root.  Two transports:
 *   BZ_GATHER_PEER  (default) the root keeps an image buffer that the other ranks map (CUDA IPC); each rank's rows are STORED straight into
 *                   their place in it over NVLink / NVSwitch by one kernel per rank -- transfer and assembly fused, all ranks in parallel.
 *                   "Rows of call k are there" and "frame k has been taken out of the buffer" are generation counters in that buffer,
 *                   written and polled across NVLink by the same kernels: no NCCL call per frame.  The first call with a new frame size
 *                   or root sets the mapping up (collective: NCCL broadcast of the IPC handle, synchronises the stream); if IPC mapping
 *                   fails on any rank, all ranks fall back to BZ_GATHER_NCCL together (bz_comm_gather_mode tells which is in use).  A rank
 *                   that waits two minutes for another rank's flag traps (the next CUDA call reports it) instead of spinning for ever.
 *   BZ_GATHER_NCCL  bands by ncclSend / ncclRecv (one group) into a staging buffer on the root, then one pass that puts the rows in place.
 * bz_comm_set_gather: same mode on every rank, between calls.  bz_comm_destroy: collective like bz_comm_create -- every rank calls it, after
 * its last gather has completed (ncclCommDestroy waits for the other ranks; a rank that waits alone for the others before destroying
 * deadlocks).  The other ranks unmap the root's buffer before that meeting point, the root frees it after. */
#define BZ_COMM_ID_BYTES 128
typedef struct bz_comm bz_comm;
int  bz_comm_unique_id(uint8_t out_id[BZ_COMM_ID_BYTES]);
int  bz_comm_create(const uint8_t id[BZ_COMM_ID_BYTES], uint32_t rank, uint32_t ranks, int device, bz_comm **out_comm);
int  bz_comm_destroy(bz_comm *comm);
int  bz_gather_rows(bz_comm *comm, const uint8_t *d_band, uint32_t width, uint32_t height, uint32_t stripe_rows, uint32_t root,
                    uint8_t *d_rgba, void *stream);
#define BZ_GATHER_PEER 0u
#define BZ_GATHER_NCCL 1u
int  bz_comm_set_gather(bz_comm *comm, uint32_t mode);
int  bz_comm_gather_mode(const bz_comm *comm, uint32_t *out_mode);

/* After an all-gather of the ranks' stripe buffers (each padded to band_rows rows; NCCL over NVLink in bench.py):
 * d_bands = ranks x band_rows x width x 4 bytes -> d_rgba = height x width x 4 bytes in image order. */
int  bz_assemble_stripes(const uint8_t *d_bands, uint32_t band_rows, uint32_t width, uint32_t height, uint32_t stripe_rows,
                         uint32_t ranks, uint8_t *d_rgba, void *stream);

/* ---- illumination mode (reference/hostUtil.h:8-28, hostUtil.cpp:3-29, README.md:161-194) ----
 * A rectangular emitter behind the lens casts rays into the hemisphere around its normal; the rays cross the lens
 * (the same bounce loop as the renderer) and are counted where they land on a rectangular target.
 *
 * UniformHemisphere (hostUtil.cpp:3-29): belt i of `belts` has ceil(4 b sin((2i+1)/(4b) pi)) patches (double sin, as the
 * reference's ::sin resolves); a draw is incidence = acos(u1), turn = u2 * 2pi, direction (cos inc, sin inc cos turn,
 * sin inc sin turn) normalised, patch = first[belt] + turn / patch_width[belt], belt = incidence / (pi / 2 / belts).
 * The reference draws u1, u2 from std::ranlux24_base; here they come from a counter-based generator (statistical, not
 * bit, parity with the reference's stream): u = (splitmix64(seed + (2 * ray + draw) * 0x9E3779B97F4A7C15) >> 40) * 2^-24.
 * Emission order: ray = (point_v * points_u + point_u) * rays_per_point + k; the point is the centre of its grid cell. */
typedef struct bz_emitter {
  float    origin[3];        /* corner of the emitter rectangle */
  float    edge_u[3];        /* its two edges */
  float    edge_v[3];
  float    normal[3];        /* unit axis of the hemisphere (direction(0) of hostUtil.cpp:22) */
  float    tangent_a[3];     /* unit, direction(1) */
  float    tangent_b[3];     /* unit, direction(2) */
  uint32_t points_u, points_v;
  uint32_t rays_per_point;
  uint32_t belts;
  uint64_t seed;
} bz_emitter;

/* Target rectangle origin + s*edge_u + t*edge_v (edges orthogonal), bins_u x bins_v counting bins, row-major in v.
 * A ray in air meets the plane with Plane::intersect semantics (3dGeomUtil.h:279-296); s = ((P-origin).edge_u)/(edge_u.edge_u),
 * likewise t; it counts for bin (floor(t*bins_v), floor(s*bins_u)) when 0 <= s,t < 1. */
typedef struct bz_target {
  float    origin[3];
  float    edge_u[3];
  float    edge_v[3];
  uint32_t bins_u, bins_v;
} bz_target;

#define BZ_LAND_LOST    0xffffffffu   /* the ray entered the lens and never left (or was still inside after max_crossings) */
#define BZ_LAND_MISSED  0xfffffffeu   /* the ray ended in air but does not meet the target rectangle */

uint32_t bz_sizeof_emitter(void);
uint32_t bz_sizeof_target(void);
/* UniformHemisphere::getPatchCount and the belt table (hostUtil.cpp:8-13): out_width / out_first hold `belts` entries (nullable) */
int  bz_hemisphere_patches(uint32_t belts, uint32_t *out_patch_count, float *out_width, uint32_t *out_first);
/* UniformHemisphere::getRandom for rays [first, first + count) of the emission order: start, direction (n x 3 floats) and
 * hemisphere patch index per ray (d_patch nullable). */
int  bz_emit_rays(const bz_emitter *emitter, uint64_t first, uint64_t count, float *d_start, float *d_dir, uint32_t *d_patch, void *stream);
/* Bounce loop (as bz_render: expected = inside first, at most max_crossings refract() calls) over a ray batch, then the target:
 * d_bins (bins_v x bins_u u32) is ACCUMULATED into; d_landing (nullable) gets the bin index, BZ_LAND_LOST or BZ_LAND_MISSED
 * per ray.  d_stats as bz_render. */
int  bz_trace_to_target(bz_lens *lens, const float *d_start, const float *d_dir, uint64_t n, const bz_target *target, uint32_t max_crossings,
                        uint32_t *d_bins, uint32_t *d_landing, bz_stats *d_stats, void *stream);
/* The whole simulation: emit -> group the rays by hemisphere patch (counting sort: rays of one patch have similar directions,
 * README.md:176-186) -> bz_trace_to_target.  d_bins is accumulated into. */
int  bz_illuminate(bz_lens *lens, const bz_emitter *emitter, const bz_target *target, uint32_t max_crossings,
                   uint32_t *d_bins, bz_stats *d_stats, void *stream);
int  bz_illuminate_host(bz_lens *lens, const bz_emitter *emitter, const bz_target *target, uint32_t max_crossings,
                        uint32_t *h_bins, bz_stats *h_stats);

/* Cull-hierarchy facts of a lens: out[0] = 4-wide nodes, out[1] = levels, out[2] = patches in the
 * always-tested segment, out[3] = SM count of the device. */
int  bz_lens_info(const bz_lens *lens, uint32_t out[4]);

/* Measured FP32 issue rate of `device` (dependent-FMA chains, 2 flop per lane and instruction), in
 * TFLOP/s: the denominator of the FP32 roofline fraction. */
int  bz_fp32_peak_probe(int device, float *out_tflops, float *out_ms);

/* Device time (ms) of the intersect/refract kernels of the calling thread's most recent bz_render* / bz_*_batch call on this lens,
 * measured with CUDA events on the stream the call used; waits for that call's kernels. */
int  bz_last_kernel_ms(bz_lens *lens, float *out_ms);

/* Host time (ms) the most recent bz_render* call of the calling thread spent planning the frame (bz_plan_tiles: projecting the outline
 * boxes of the cull hierarchy, classifying and ordering the tiles) and staging the tile list for upload; 0 when the kept plan of the last frame was reused. */
int  bz_last_plan_ms(const bz_lens *lens, float *out_ms);

/* Kernels the most recent bz_render* / bz_refract_batch / bz_intersect_batch / bz_trace_to_target call on this lens launched
 * (RENDER: bz_sky_kernel when the frame has tiles beside the lens, the trace kernel when it has tiles on it, the resolve). */
int  bz_last_kernel_launches(const bz_lens *lens, uint32_t *out_count);

#ifdef __cplusplus
}
#endif

#endif

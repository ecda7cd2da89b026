// bz_trace_common.cuh -- device helpers shared by the persistent kernel (csrc/bz_trace.cu) and the wavefront kernels
// (csrc/bz_wave.cu): record loads, stage B on the packed FP32 pipe, K1 ray generation, K3 shading, the packet's interval
// slab test, tile arithmetic.  Included INSIDE the build's namespace (bz_exact / bz_fast), after bz_math.cuh.
#ifndef BZ_TRACE_COMMON_CUH
#define BZ_TRACE_COMMON_CUH

#define BZ_KEY_NONE  0xffffffffffffffffull
#define BZ_END       0xffffffffu
enum { BZC_PRIMARY = 0, BZC_CALLS, BZC_STAGE_A, BZC_STAGE_B, BZC_SNELL, BZC_ENTER, BZC_EXIT, BZC_LOST, BZC_NODES, BZC_PNODES, BZC_PWALKS, BZC_LWALKS };
#define BZ_NSTAT     12

// STAGED: the record sits in a shared-memory staging buffer, float4 i of this lane at p[i * 32] (csrc/bz_wave.cu); else in the
// global table at p[i]
template<bool STAGED>
__device__ __forceinline__ float4 bzw_word4(const float4 *p, int i) { return STAGED ? p[i * 32] : __ldg(p + i); }

template<bool STAGED = false>
__device__ __forceinline__ void bzw_unpack_cull(const float4 *p, bz_cull_rec &c, uint32_t &index) {
  float4 r0 = bzw_word4<STAGED>(p, 0), r1 = bzw_word4<STAGED>(p, 1), r2 = bzw_word4<STAGED>(p, 2), r3 = bzw_word4<STAGED>(p, 3);
  c.n = bz_v(r0.x, r0.y, r0.z); c.k = r0.w;
  c.b0 = bz_v(r1.x, r1.y, r1.z); c.h_in = r1.w;
  c.b1 = bz_v(r2.x, r2.y, r2.z); c.h_out = r2.w;
  c.b2 = bz_v(r3.x, r3.y, r3.z);
  index = __float_as_uint(r3.w);
}

// ---- stage B on the packed FP32 pipe (sm_100a FMUL2 / FFMA2: two FP32 operations per issue slot) ----
// The control points are kept as (x, y) pairs + z, the x and y components of bz_interpolate / bz_normal go through
// __fmul2_rn / __ffma2_rn, z stays scalar.  Every component sees the same IEEE operations in the same order as in
// bz_math.cuh, so the results are bit-identical (tests compare with the oracle).  A packed sum a + b is written
// fma(b, one, a) with `one` = 1.0f read from the job descriptor: exactly a + b, and a form ptxas cannot contract -- it
// fuses __fadd2_rn(__fmul2_rn(p, q), a) into one FFMA2 (a single rounding) even under -fmad=false.
__device__ __forceinline__ float2 bzw_mul2(float2 a, float s) { return __fmul2_rn(a, make_float2(s, s)); }
#if defined(BZ_FAST)
__device__ __forceinline__ float2 bzw_add2(float2 a, float2 b, float) { return __fadd2_rn(a, b); }      // (ptxas fuses a product feeding it into one FFMA2)
#else
__device__ __forceinline__ float2 bzw_add2(float2 a, float2 b, float one) { return __ffma2_rn(b, make_float2(one, one), a); }
#endif

// K1 and K3 are the same bits in both builds: IEEE operations that never contract (__fmul_rn / __fadd_rn / __fdiv_rn /
// __fsqrt_rn), three-term sums as a0 + (a1 + a2)
__device__ __forceinline__ float bzw_dot_rn(const bzv3 &a, const bzv3 &b) { return __fadd_rn(__fmul_rn(a.x, b.x), __fadd_rn(__fmul_rn(a.y, b.y), __fmul_rn(a.z, b.z))); }
__device__ __forceinline__ bzv3 bzw_normalized_rn(const bzv3 &v) {
  const float z = bzw_dot_rn(v, v);
  if(z > 0.0f) {
    const float n = __fsqrt_rn(z);
    return bz_v(__fdiv_rn(v.x, n), __fdiv_rn(v.y, n), __fdiv_rn(v.z, n));
  }
  return v;
}

#if defined(BZ_FAST)
// control points and dirB: words 0..32 of the stage-B record ((x, y) of the ten control points, their z, dirB)
template<bool STAGED = false>
__device__ __forceinline__ void bzw_load_newton_root(const float4 *p, bz_newton_rec &t) {
  float w[36];
#pragma unroll
  for(int i = 0; i < 9; ++i) {
    float4 r = bzw_word4<STAGED>(p, i);
    w[4 * i] = r.x; w[4 * i + 1] = r.y; w[4 * i + 2] = r.z; w[4 * i + 3] = r.w;
  }
#pragma unroll
  for(int i = 0; i < 10; ++i) { t.cp[i] = bz_v(w[2 * i], w[2 * i + 1], w[20 + i]); }
  t.dir_b = bz_v(w[30], w[31], w[32]);
}
#else
struct bzw_newton2 {
  float2 xy[10];
  float  z[10];
  bzv3   dir_b;
  float  one;
};

template<bool STAGED = false>
__device__ __forceinline__ void bzw_load_newton2(const float4 *p, float one, bzw_newton2 &t) {
  float w[36];
#pragma unroll
  for(int i = 0; i < 9; ++i) {
    float4 r = bzw_word4<STAGED>(p, i);
    w[4 * i] = r.x; w[4 * i + 1] = r.y; w[4 * i + 2] = r.z; w[4 * i + 3] = r.w;
  }
#pragma unroll
  for(int i = 0; i < 10; ++i) { t.xy[i] = make_float2(w[2 * i], w[2 * i + 1]); t.z[i] = w[20 + i]; }
  t.dir_b = bz_v(w[30], w[31], w[32]);
  t.one = one;
}

// bezierTriangle.cpp:105-121, operation for operation as bz_interpolate(const bz_newton_rec&, ...)
__device__ __forceinline__ bzv3 bz_interpolate(const bzw_newton2 &t, const bzv3 &b) {
  const float b0 = b.x, b1 = b.y, b2 = b.z;
  const float q0 = b0 * b0, q1 = b1 * b1, q2 = b2 * b2;
  const float one = t.one;
#define BZW_T2(i, m, n) bzw_mul2(bzw_mul2(t.xy[i], m), n)
  float2 a = bzw_add2(BZW_T2(0, b0, q0), BZW_T2(1, b1, q1), one);
  a = bzw_add2(a, BZW_T2(2, b2, q2), one);
  float2 g = bzw_add2(BZW_T2(3, b1, q0), BZW_T2(4, b0, q1), one);
  g = bzw_add2(g, BZW_T2(5, b2, q1), one);
  g = bzw_add2(g, BZW_T2(6, b1, q2), one);
  g = bzw_add2(g, BZW_T2(7, b0, q2), one);
  g = bzw_add2(g, BZW_T2(8, b2, q0), one);
  a = bzw_add2(a, bzw_mul2(g, 3.0f), one);
  a = bzw_add2(a, bzw_mul2(bzw_mul2(bzw_mul2(bzw_mul2(t.xy[9], b0), b1), b2), 6.0f), one);
#undef BZW_T2
  bzv3 r;
  r.x = a.x;
  r.y = a.y;
  r.z = t.z[0] * b0 * q0 + t.z[1] * b1 * q1 + t.z[2] * b2 * q2 +
        3.0f * (t.z[3] * b1 * q0 + t.z[4] * b0 * q1 + t.z[5] * b2 * q1 + t.z[6] * b1 * q2 + t.z[7] * b0 * q2 + t.z[8] * b2 * q0) +
        t.z[9] * b0 * b1 * b2 * 6.0f;
  return r;
}

// bezierTriangle.cpp:197-233, operation for operation as bz_normal(const bz_newton_rec&, ...)
__device__ __forceinline__ bzv3 bz_normal(const bzw_newton2 &t, const bzv3 &b) {
  const float b0 = b.x, b1 = b.y, b2 = b.z;
  const float q0 = b0 * b0, q1 = b1 * b1, q2 = b2 * b2;
  const float one = t.one;
  // c = cp[i]*qa + cp[j]*qb + cp[k]*qc + 2*(cp[l]*m0*n0 + cp[m]*m1*n1 + cp[n]*m2*n2)
  // (X + 2*S is one fma(S, 2, X): doubling is exact, so the single rounding is the reference's rounding of the sum)
#define BZW_C2(i, qa, j, qb, k, qc, l, m0, n0, m, m1, n1, n, m2, n2) \
  __ffma2_rn(bzw_add2(bzw_add2(bzw_mul2(bzw_mul2(t.xy[l], m0), n0), bzw_mul2(bzw_mul2(t.xy[m], m1), n1), one), \
                      bzw_mul2(bzw_mul2(t.xy[n], m2), n2), one), make_float2(2.0f, 2.0f), \
             bzw_add2(bzw_add2(bzw_mul2(t.xy[i], qa), bzw_mul2(t.xy[j], qb), one), bzw_mul2(t.xy[k], qc), one))
#define BZW_C1(i, qa, j, qb, k, qc, l, m0, n0, m, m1, n1, n, m2, n2) \
  __fmaf_rn(t.z[l] * m0 * n0 + t.z[m] * m1 * n1 + t.z[n] * m2 * n2, 2.0f, t.z[i] * qa + t.z[j] * qb + t.z[k] * qc)
  const float2 c0 = BZW_C2(0, q0, 7, q2, 4, q1, 8, b0, b2, 3, b0, b1, 9, b2, b1);
  const float2 c1 = BZW_C2(1, q1, 6, q2, 3, q0, 9, b0, b2, 4, b0, b1, 5, b1, b2);
  const float2 c2 = BZW_C2(2, q2, 8, q0, 5, q1, 7, b0, b2, 6, b1, b2, 9, b0, b1);
  const float z0 = BZW_C1(0, q0, 7, q2, 4, q1, 8, b0, b2, 3, b0, b1, 9, b2, b1);
  const float z1 = BZW_C1(1, q1, 6, q2, 3, q0, 9, b0, b2, 4, b0, b1, 5, b1, b2);
  const float z2 = BZW_C1(2, q2, 8, q0, 5, q1, 7, b0, b2, 6, b1, b2, 9, b0, b1);
#undef BZW_C2
#undef BZW_C1
  // a = 1*c0 + 0*c1 + -1*c2 (dirA = (1, 0, -1), bezierTriangle.cpp:83); e = dirB . (c0, c1, c2)
  // (1*c0 is c0; c0 + 0*c1 is one fma(c1, 0, c0): the product is an exact signed zero, or NaN; adding -1*c2 is subtracting c2)
  const float2 a2 = __ffma2_rn(c2, make_float2(-one, -one), __ffma2_rn(c1, make_float2(0.0f, 0.0f), c0));
  const float2 e2 = bzw_add2(bzw_add2(bzw_mul2(c0, t.dir_b.x), bzw_mul2(c1, t.dir_b.y), one), bzw_mul2(c2, t.dir_b.z), one);
  const bzv3 a = bz_v(a2.x, a2.y, __fmaf_rn(z1, 0.0f, z0) - z2);
  const bzv3 e = bz_v(e2.x, e2.y, t.dir_b.x * z0 + t.dir_b.y * z1 + t.dir_b.z * z2);
  const bzv3 n = bz_v(a.y * e.z - a.z * e.y, a.z * e.x - a.x * e.z, a.x * e.y - a.y * e.x);
  return bz_normalized(n);
}
#endif

// divider planes and neighbours: words 33..47, fetched after the root search (they are not live across its loop)
__device__ __forceinline__ void bzw_load_newton_side(const float4 *p, bz_newton_rec &t) {
  const float4 carry = __ldg(p + 8), r9 = __ldg(p + 9), r10 = __ldg(p + 10), r11 = __ldg(p + 11);
  t.div_n[0] = bz_v(carry.y, carry.z, carry.w); t.div_k[0] = r9.x;
  t.div_n[1] = bz_v(r9.y, r9.z, r9.w);          t.div_k[1] = r10.x;
  t.div_n[2] = bz_v(r10.y, r10.z, r10.w);       t.div_k[2] = r11.x;
  t.nb[0] = __float_as_uint(r11.y); t.nb[1] = __float_as_uint(r11.z); t.nb[2] = __float_as_uint(r11.w);
}

// K1: primary ray of sample (px, py, sx, sy); definitions in include/bz_capi.h (bz_camera).
// two_over_w / two_over_h are the quotients 2.0f / (float)width and 2.0f / (float)height, formed once on the host (the same
// IEEE division the per-ray code would do).
__device__ __forceinline__ void bzw_primary_ray(const bz_camera &cam, float two_over_w, float two_over_h, uint32_t px, uint32_t py, uint32_t sx, uint32_t sy,
                                                bzv3 &start, bzv3 &dir) {
  const float strata = (float)cam.strata;
  const float fu = __fadd_rn((float)px, __fdiv_rn(__fadd_rn((float)sx, 0.5f), strata));
  const float fv = __fadd_rn((float)py, __fdiv_rn(__fadd_rn((float)sy, 0.5f), strata));
  const float a = __fmul_rn(__fadd_rn(__fmul_rn(fu, two_over_w), -1.0f), cam.half_w);
  const float b = __fmul_rn(__fadd_rn(1.0f, -__fmul_rn(fv, two_over_h)), cam.half_h);
  const bzv3 offset = bz_v(__fadd_rn(__fmul_rn(cam.right[0], a), __fmul_rn(cam.up[0], b)), __fadd_rn(__fmul_rn(cam.right[1], a), __fmul_rn(cam.up[1], b)),
                           __fadd_rn(__fmul_rn(cam.right[2], a), __fmul_rn(cam.up[2], b)));
  bzv3 raw;
  if(cam.ortho != 0u) {
    start = bz_v(__fadd_rn(cam.eye[0], offset.x), __fadd_rn(cam.eye[1], offset.y), __fadd_rn(cam.eye[2], offset.z));
    raw = bz_v(cam.forward[0], cam.forward[1], cam.forward[2]);
  }
  else {
    start = bz_v(cam.eye[0], cam.eye[1], cam.eye[2]);
    raw = bz_v(__fadd_rn(cam.forward[0], offset.x), __fadd_rn(cam.forward[1], offset.y), __fadd_rn(cam.forward[2], offset.z));
  }
  dir = bzw_normalized_rn(raw);    // Ray::Ray, 3dGeomUtil.h:176-178
}

// K3: checker on the background plane, Plane::intersect semantics (3dGeomUtil.h:279-296).
__device__ __forceinline__ uint32_t bzw_shade_background(const bz_scene &sc, const bzv3 &start, const bzv3 &dir) {
  const bzv3 n = bz_v(sc.bg_plane[0], sc.bg_plane[1], sc.bg_plane[2]);
  const float cosine = bzw_dot_rn(dir, n);
  if(!(fabsf(cosine) >= BZ_PLANE_EPSILON)) { return sc.value_lost; }
  const float distance = __fdiv_rn(__fadd_rn(sc.bg_plane[3], -bzw_dot_rn(n, start)), cosine);
  if(!(distance > 0.0f)) { return sc.value_lost; }
  const bzv3 p = bz_v(__fadd_rn(start.x, __fmul_rn(distance, dir.x)), __fadd_rn(start.y, __fmul_rn(distance, dir.y)), __fadd_rn(start.z, __fmul_rn(distance, dir.z)));
  const float ca = floorf(__fdiv_rn(bzw_dot_rn(p, bz_v(sc.bg_axis_a[0], sc.bg_axis_a[1], sc.bg_axis_a[2])), sc.cell));
  const float cb = floorf(__fdiv_rn(bzw_dot_rn(p, bz_v(sc.bg_axis_b[0], sc.bg_axis_b[1], sc.bg_axis_b[2])), sc.cell));
  // |fmodf(c, 2)| for an integral-valued c, computed exactly without the library routine: c/2, floor, the doubling and
  // the difference are all exact in FP32 (and an infinite or NaN c gives NaN either way)
  const float pa = fabsf(__fadd_rn(ca, -__fmul_rn(2.0f, floorf(__fmul_rn(ca, 0.5f)))));
  const float pb = fabsf(__fadd_rn(cb, -__fmul_rn(2.0f, floorf(__fmul_rn(cb, 0.5f)))));
  return (pa != pb) ? sc.value_bright : sc.value_dark;
}

// Illumination mode: where a ray in air lands on the target rectangle (include/bz_capi.h, bz_target): the bin index,
// or BZ_LAND_MISSED.  Plane::intersect semantics (3dGeomUtil.h:279-296).
__device__ __forceinline__ uint32_t bzw_land_on_target(const bz_job &J, const bzv3 &start, const bzv3 &dir) {
  const bzv3 n = bz_v(J.target_plane[0], J.target_plane[1], J.target_plane[2]);
  const float cosine = bzw_dot_rn(dir, n);
  if(!(fabsf(cosine) >= BZ_PLANE_EPSILON)) { return BZ_LAND_MISSED; }
  const float distance = __fdiv_rn(__fadd_rn(J.target_plane[3], -bzw_dot_rn(n, start)), cosine);
  if(!(distance > 0.0f)) { return BZ_LAND_MISSED; }
  const bzv3 rel = bz_v(__fadd_rn(__fadd_rn(start.x, __fmul_rn(distance, dir.x)), -J.target.origin[0]), __fadd_rn(__fadd_rn(start.y, __fmul_rn(distance, dir.y)), -J.target.origin[1]),
                        __fadd_rn(__fadd_rn(start.z, __fmul_rn(distance, dir.z)), -J.target.origin[2]));
  const bzv3 eu = bz_v(J.target.edge_u[0], J.target.edge_u[1], J.target.edge_u[2]);
  const bzv3 ev = bz_v(J.target.edge_v[0], J.target.edge_v[1], J.target.edge_v[2]);
  const float su = __fdiv_rn(bzw_dot_rn(rel, eu), bzw_dot_rn(eu, eu)), sv = __fdiv_rn(bzw_dot_rn(rel, ev), bzw_dot_rn(ev, ev));
  if(!(su >= 0.0f && su < 1.0f && sv >= 0.0f && sv < 1.0f)) { return BZ_LAND_MISSED; }
  const uint32_t bu = min((uint32_t)__fmul_rn(su, (float)J.target.bins_u), J.target.bins_u - 1u);
  const uint32_t bv = min((uint32_t)__fmul_rn(sv, (float)J.target.bins_v), J.target.bins_v - 1u);
  return bv * J.target.bins_u + bu;
}

__device__ __forceinline__ unsigned long long bzw_make_key(float distance, uint32_t loop_index) {
  return ((unsigned long long)bz_float_key(distance) << 32) | (unsigned long long)loop_index;
}

__device__ __forceinline__ float bzw_key_distance(unsigned long long key) {
  uint32_t k = (uint32_t)(key >> 32);
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// CREDUX.{MIN,MAX}.F32: warp-wide float reduction into a uniform register, one instruction (sm_100a)
__device__ __forceinline__ float bzw_warp_min(float v) { float r; asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float bzw_warp_max(float v) { float r; asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v)); return r; }

// One axis of the packet's interval slab test.  Every ray of the packet has origin o in [o_lo, o_hi] and direction d in
// [d_lo, d_hi] on this axis, so for t >= 0 its coordinate lies in [o_lo + t d_lo, o_hi + t d_hi]; it can be inside the
// slab [lo - m, hi + m] only if   (1) o_lo + t d_lo <= hi + m   and   (2) o_hi + t d_hi >= lo - m.   Each is one bound on
// t whose kind follows from the sign of d_lo resp. d_hi, the same for all boxes:
//   d_lo > 0 (all rays ascend):   (2) bounds t from below, (1) from above
//   d_hi < 0 (all rays descend):  (1) bounds t from below, (2) from above
//   d_lo < 0 < d_hi (straddle):   both bound t from below, nothing from above
// The three cases become ONE select-free formula by choosing per packet which box face each term reads (an address
// offset) and by neutral coefficients (0, -inf) / (0, +inf):
//   near = max(face_n * in + cn, face_f * ix + cx),  far = face_f * jf + cf
struct bzw_axis {
  float in, cn, ix, cx, jf, cf;
  uint32_t off_n, off_f;        // word offsets of the two faces from the axis's lo[] row of a packet node: 0 (lo) or 96 (hi)
};

__device__ __forceinline__ bzw_axis bzw_make_axis(float o_lo, float o_hi, float d_lo, float d_hi, float margin) {
  const float inf = __int_as_float(0x7f800000);
  // a vanishing slope is pushed to the permissive side (the lane walk treats |d| < 1e-20 the same way)
  if(fabsf(d_lo) < 1e-20f) { d_lo = -1e-20f; }
  if(fabsf(d_hi) < 1e-20f) { d_hi = 1e-20f; }
  // MUFU.RCP (<= 2 ulp): the error is of the size of the rounding the margin already covers
  float inv1, inv2;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv1) : "f"(d_lo));      // 1e-20 <= |d| <= 1: no range fix-up needed
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv2) : "f"(d_hi));
  const float c1 = (margin - o_lo) * inv1;          // (1): t1 = hi * inv1 + c1
  const float c2 = (-margin - o_hi) * inv2;         // (2): t2 = lo * inv2 + c2
  bzw_axis a;
  if(d_lo > 0.0f)      { a.off_n = 0u;  a.in = inv2; a.cn = c2; a.off_f = 96u; a.jf = inv1; a.cf = c1;  a.ix = 0.0f; a.cx = -inf; }
  else if(d_hi < 0.0f) { a.off_n = 96u; a.in = inv1; a.cn = c1; a.off_f = 0u;  a.jf = inv2; a.cf = c2;  a.ix = 0.0f; a.cx = -inf; }
  else                 { a.off_n = 0u;  a.in = inv2; a.cn = c2; a.off_f = 96u; a.jf = 0.0f; a.cf = inf; a.ix = inv1; a.cx = c1; }
  return a;
}

// `row` = this lane's entry in the axis's lo[] row of node 0; `node` = word offset of the node
__device__ __forceinline__ void bzw_axis_bounds(const bzw_axis &a, const float *row, uint32_t node, float &tnear, float &tfar) {
  // one pointer per packet and face, pinned in a register pair, so that a node costs one IMAD.WIDE per load
  const float *near_face = row + a.off_n, *far_face = row + a.off_f;
  asm volatile("" : "+l"(near_face), "+l"(far_face));
  const float face_n = __ldg(near_face + node), face_f = __ldg(far_face + node);
  tnear = fmaxf(tnear, fmaxf(__fmaf_rn(face_n, a.in, a.cn), __fmaf_rn(face_f, a.ix, a.cx)));
  tfar = fminf(tfar, __fmaf_rn(face_f, a.jf, a.cf));
}

// n / d for a per-job divisor d: div = { ceil(2^32 / d) mod 2^32, d == 1 }, exact while n * d < 2^32 (the host checks)
__device__ __forceinline__ uint32_t bzw_div(uint32_t n, const uint32_t div[2]) { return __umulhi(n, div[0]) + n * div[1]; }

// RENDER: row of the whole frame behind a local row of the band (contiguous band from row0, or round-robin stripes)
__device__ __forceinline__ uint32_t bzw_frame_row(const bz_job &J, uint32_t row) {
  if(J.stripe_rows == 0u) { return J.row0 + row; }
  const uint32_t stripe = bzw_div(row, J.div_stripe);
  return (stripe * J.stripe_ranks + J.stripe_rank) * J.stripe_rows + (row - stripe * J.stripe_rows);
}

// the even bits of v, compacted: x of a Morton index (y = bzw_even_bits(v >> 1))
__device__ __forceinline__ uint32_t bzw_even_bits(uint32_t v) {
  v &= 0x55555555u;
  v = (v | (v >> 1)) & 0x33333333u;
  v = (v | (v >> 2)) & 0x0f0f0f0fu;
  v = (v | (v >> 4)) & 0x00ff00ffu;
  v = (v | (v >> 8)) & 0x0000ffffu;
  return v;
}

#endif

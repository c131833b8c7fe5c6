// traverse.h — per-ray traversal, host+device (the device kernels in kernels_*.cu wrap these; tests/hostemu compiles the
// very same functions for the CPU so the logic can be unit-tested without a GPU — that harness is test-only and is
// never linked into libptb.so).
//
//  * wide_closest / wide_any   : TLAS::intersect / TLAS::any_intersect (tlas.rs:66-144) over the 8-wide layout.
//  * bin_closest / bin_any     : the same two functions walking the reference's own binary trees in the reference's
//                                visit order (push_to_stack, blas.rs:133-162) — bit-exact including ties.
//
// Exactness argument for the wide kernels: a triangle's (t,u,v) is a pure function of the BLAS-space ray,
// the triangle planes and the entry distance of the *reference leaf box* (primitive.rs:147-155, boundingbox.rs:129), all
// of which are evaluated here with the reference's operation order (exact_math.h).  Quantised child boxes only cull;
// they are conservative by construction (host_wide.cpp) plus an explicit rounding slack below, and slab results are
// monotone under box inclusion, so every leaf the reference can reach with a hit is reached here.  The winner is the
// minimum t; only exact ties may resolve differently (the reference keeps the last-visited of equal hits).
#pragma once
#include "exact_math.h"
#include "layout.h"

// Development hook: a host-side tool (tools/stagecount.cpp) defines PTB_STAT to count the stages a ray runs through.
#ifndef PTB_STAT
#define PTB_STAT(stage)
#endif

namespace ptb {

using ex::V3;

#if defined(__CUDA_ARCH__)
#define PTB_LDG4U(p) __ldg(p)
#define PTB_LDG4F(p) __ldg(p)
#define PTB_LDGU(p) __ldg(p)
#define PTB_LDGF(p) __ldg(p)
#else
#define PTB_LDG4U(p) (*(p))
#define PTB_LDG4F(p) (*(p))
#define PTB_LDGU(p) (*(p))
#define PTB_LDGF(p) (*(p))
#endif

struct Hit
{
    float t, u, v;
    uint32_t slot_or_prim;  // wide: triangle slot; bin: BLAS-local prim id (resolved to prim id by finish_hit)
    uint32_t inst;          // InstanceRow index, 0xffffffff = miss
};

struct TravRay
{
    V3 o, d;     // current space (world or BLAS)
    V3 inv;      // exact 1/d (ray.rs:16), may be +-inf
    V3 idq;      // finite reciprocal for the conservative quantised tests
    uint32_t oi; // 7 ^ octant
};

// Finite stand-in for 1/d in the conservative quantised tests: 1/d clamped to +-2^60 (|d| < 2^-60, including +-0 whose
// exact reciprocal is +-inf, is treated as +-2^-60).  Derived from the exact reciprocal, so it costs no second division.
PTB_HD float finite_recip_of(float inv)
{
    const float big = 1.152921504606846976e18f;  // 2^60
    return fabsf(inv) > big ? copysignf(big, inv) : inv;
}
PTB_HD void set_ray(TravRay& r, V3 o, V3 d)
{
    r.o = o;
    r.d = d;
    r.inv = ex::recip(d);
    r.idq = ex::v3(finite_recip_of(r.inv.x), finite_recip_of(r.inv.y), finite_recip_of(r.inv.z));
    uint32_t oct = (d.x < 0.0f ? 4u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 1u : 0u);
    r.oi = 7u ^ oct;
}
// Ray::transform (ray.rs:22-28)
PTB_HD void transform_ray(TravRay& out, V3 o, V3 d, const float* inv12)
{
    V3 c0 = ex::v3(inv12[0], inv12[1], inv12[2]), c1 = ex::v3(inv12[3], inv12[4], inv12[5]), c2 = ex::v3(inv12[6], inv12[7], inv12[8]);
    V3 t = ex::v3(inv12[9], inv12[10], inv12[11]);
    set_ray(out, ex::add(ex::m3_mul(c0, c1, c2, o), t), ex::m3_mul(c0, c1, c2, d));
}

// Byte k of w as the float 2^23 + 256*byte: one byte-permute places the byte in mantissa bits 8..15 of 0x4B000000
// (no integer-to-float conversion); the 2^23 bias and the factor 256 are folded into the plane equations below.
// Ray::transform into an instance when the world ray's exact reciprocal is known: a component of the transformed direction
// that is bit-identical to the world direction's (identity / translation-only instances, or an axis a rotation leaves
// alone) reuses the world reciprocal instead of dividing again — same bits, since 1/x is a function of x's bits.
PTB_HD void transform_ray_reuse(TravRay& out, V3 o, V3 d, V3 inv_d, const float* inv12)
{
    V3 c0 = ex::v3(inv12[0], inv12[1], inv12[2]), c1 = ex::v3(inv12[3], inv12[4], inv12[5]), c2 = ex::v3(inv12[6], inv12[7], inv12[8]);
    V3 t = ex::v3(inv12[9], inv12[10], inv12[11]);
    out.o = ex::add(ex::m3_mul(c0, c1, c2, o), t);
    const V3 dd = ex::m3_mul(c0, c1, c2, d);
    out.d = dd;
    out.inv.x = ex::bits(dd.x) == ex::bits(d.x) ? inv_d.x : ex::rcp(dd.x);
    out.inv.y = ex::bits(dd.y) == ex::bits(d.y) ? inv_d.y : ex::rcp(dd.y);
    out.inv.z = ex::bits(dd.z) == ex::bits(d.z) ? inv_d.z : ex::rcp(dd.z);
    out.idq = ex::v3(finite_recip_of(out.inv.x), finite_recip_of(out.inv.y), finite_recip_of(out.inv.z));
    uint32_t oct = (dd.x < 0.0f ? 4u : 0u) | (dd.y < 0.0f ? 2u : 0u) | (dd.z < 0.0f ? 1u : 0u);
    out.oi = 7u ^ oct;
}

// `kc` = 0x4B000000 read from constant memory (qbyte_const), which the compiler cannot see through: PRMT then takes the byte selector as its
// immediate.  Given the literal, ptxas puts the LITERAL into the immediate slot and the selector into a register that the result
// overwrites, i.e. one extra MOV per byte — 51 of the 292 instructions of the node stage (profiles/r2a_c3_extend_sass.txt).
#if defined(__CUDACC__)
static __constant__ uint32_t g_qbyte_kc = 0x4B000000u;  // read through the constant bank: ptxas cannot fold it back into a literal
#endif
PTB_HD uint32_t qbyte_const()
{
#if defined(__CUDA_ARCH__)
    return g_qbyte_kc;
#else
    return 0x4B000000u;
#endif
}
PTB_HD float qbyte_biased(uint32_t w, int k, uint32_t kc)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(w, kc, 0x7504u | ((uint32_t)k << 4)));
#else
    return ex::from_bits(kc | (((w >> (8 * k)) & 0xffu) << 8));
#endif
}

// XOR-permute two packed 8-bit masks (bits 0..7 and 8..15): bit (s ^ x) of the result = bit s of the input
PTB_HD uint32_t xor_permute16(uint32_t m, uint32_t x)
{
    if (x & 1u) m = ((m & 0x5555u) << 1) | ((m >> 1) & 0x5555u);
    if (x & 2u) m = ((m & 0x3333u) << 2) | ((m >> 2) & 0x3333u);
    if (x & 4u) m = ((m & 0x0f0fu) << 4) | ((m >> 4) & 0x0f0fu);
    return m;
}

// Conservative test of the 8 quantised child boxes of one node; returns a hit mask by slot (bits of empty slots are
// meaningless: the caller masks with imask | lmask).
// Plane distance t = (p + q*2^e - o) * idq = q*s + b with s = 2^e*idq, b = (p - o)*idq, evaluated as one FMA per plane:
// t = v*s8 + c where v = 2^23 + 256*q (qbyte_biased), s8 = s/256, c = b - 2^15*s.
PTB_HD uint32_t wide_node_test(const uint4& q0, const uint4& q2, const uint4& q3, const uint4& q4, const TravRay& r, float t_max)
{
    const float px = ex::from_bits(q0.x), py = ex::from_bits(q0.y), pz = ex::from_bits(q0.z);
    const float sx0 = ex::from_bits((q0.w & 0xffu) << 23), sy0 = ex::from_bits(((q0.w >> 8) & 0xffu) << 23), sz0 = ex::from_bits(((q0.w >> 16) & 0xffu) << 23);
    const float sx = sx0 * r.idq.x, sy = sy0 * r.idq.y, sz = sz0 * r.idq.z;
    const float bx = (px - r.o.x) * r.idq.x, by = (py - r.o.y) * r.idq.y, bz = (pz - r.o.z) * r.idq.z;
    // rounding slack: the fused evaluation differs from the exact slab arithmetic by a few ulps of |b| + 255|s|, plus the
    // rounding of the bias term c (<= 2^-9 |s|)
    const float k = 9.5367432e-7f, kb = 7.8125e-3f;  // 2^-20, 2^-7
    const float ex_ = fmaf(fabsf(bx) + 255.0f * fabsf(sx), k, fabsf(sx) * kb), ey_ = fmaf(fabsf(by) + 255.0f * fabsf(sy), k, fabsf(sy) * kb),
                ez_ = fmaf(fabsf(bz) + 255.0f * fabsf(sz), k, fabsf(sz) * kb);
    const float bias = -32768.0f, inv256 = 0.00390625f;
    const float cnx = fmaf(bias, sx, bx - ex_), cfx = fmaf(bias, sx, bx + ex_), cny = fmaf(bias, sy, by - ey_), cfy = fmaf(bias, sy, by + ey_),
                cnz = fmaf(bias, sz, bz - ez_), cfz = fmaf(bias, sz, bz + ez_);
    const float s8x = sx * inv256, s8y = sy * inv256, s8z = sz * inv256;
    // near/far byte planes by direction sign
    const bool nx = r.idq.x < 0.0f, ny = r.idq.y < 0.0f, nz = r.idq.z < 0.0f;
    const uint32_t lox0 = q2.x, lox1 = q2.y, loy0 = q2.z, loy1 = q2.w, loz0 = q3.x, loz1 = q3.y;
    const uint32_t hix0 = q3.z, hix1 = q3.w, hiy0 = q4.x, hiy1 = q4.y, hiz0 = q4.z, hiz1 = q4.w;
    const uint32_t nx0 = nx ? hix0 : lox0, nx1 = nx ? hix1 : lox1, fx0 = nx ? lox0 : hix0, fx1 = nx ? lox1 : hix1;
    const uint32_t ny0 = ny ? hiy0 : loy0, ny1 = ny ? hiy1 : loy1, fy0 = ny ? loy0 : hiy0, fy1 = ny ? loy1 : hiy1;
    const uint32_t nz0 = nz ? hiz0 : loz0, nz1 = nz ? hiz1 : loz1, fz0 = nz ? loz0 : hiz0, fz1 = nz ? loz1 : hiz1;
    uint32_t mask = 0;
    const uint32_t kc = qbyte_const();
#pragma unroll
    for (int s = 0; s < 8; ++s)
    {
        const int b4 = s & 3;
        const uint32_t wnx = s < 4 ? nx0 : nx1, wfx = s < 4 ? fx0 : fx1, wny = s < 4 ? ny0 : ny1, wfy = s < 4 ? fy0 : fy1, wnz = s < 4 ? nz0 : nz1, wfz = s < 4 ? fz0 : fz1;
        const float tnx = fmaf(qbyte_biased(wnx, b4, kc), s8x, cnx), tfx = fmaf(qbyte_biased(wfx, b4, kc), s8x, cfx);
        const float tny = fmaf(qbyte_biased(wny, b4, kc), s8y, cny), tfy = fmaf(qbyte_biased(wfy, b4, kc), s8y, cfy);
        const float tnz = fmaf(qbyte_biased(wnz, b4, kc), s8z, cnz), tfz = fmaf(qbyte_biased(wfz, b4, kc), s8z, cfz);
        const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, ex::EPS));
        const float tf = fminf(fminf(tfx, tfy), fminf(tfz, t_max));
        if (tn <= tf) mask |= (1u << s);
    }
    return mask;
}

// hit record resolution shared by the kernels: normal, front-facing (primitive.rs:158-165, tlas.rs:103-105)
PTB_HD void finish_hit(const SceneDev& S, V3 d_world, uint32_t inst, uint32_t prim, float u, float v, V3* normal, bool* front)
{
    const InstanceRow& R = S.inst[inst];
    const float* nn = S.tri_normals + (size_t)(R.tri_base + prim) * 9;
    V3 n = ex::tri_normal(ex::v3(PTB_LDGF(nn + 0), PTB_LDGF(nn + 1), PTB_LDGF(nn + 2)), ex::v3(PTB_LDGF(nn + 3), PTB_LDGF(nn + 4), PTB_LDGF(nn + 5)),
                          ex::v3(PTB_LDGF(nn + 6), PTB_LDGF(nn + 7), PTB_LDGF(nn + 8)), u, v);
    // BLAS-space direction = inv_matrix3 * d (Ray::transform)
    V3 dl = ex::m3_mul(ex::v3(R.inv[0], R.inv[1], R.inv[2]), ex::v3(R.inv[3], R.inv[4], R.inv[5]), ex::v3(R.inv[6], R.inv[7], R.inv[8]), d_world);
    bool ff = ex::dot3(dl, n) < 0.0f;
    if (!ff) n = ex::neg(n);
    *normal = ex::m3_mul(ex::v3(R.fwd[0], R.fwd[1], R.fwd[2]), ex::v3(R.fwd[3], R.fwd[4], R.fwd[5]), ex::v3(R.fwd[6], R.fwd[7], R.fwd[8]), n);
    *front = ff;
}

// ------------------------------------------------------------------------------------------------------------------------
// 8-wide traversal as a resumable per-ray state machine.  Stack entries (uint2): x = child_base | leaf header index,
// y = kind/payload:
//   group   : y = hits_permuted (bits 0..7) | imask << 8            (bit 31 clear)
//   instance: y = 0x80000000                                          x = leaf header index of a TLAS leaf
// One wide_step() lets a ray flow through up to four stages, each at a single code location so that the lanes of a warp
// that need the same stage execute it together:  P (pop / descend / enter instance) -> N (fetch + test one wide node)
// -> L (one reference leaf: exact box, exact t_enter) -> T (one triangle).  Leaves of a BLAS node are finished (front to
// back) before its inner children are visited, TLAS leaves are stacked nearest-on-top: the visit order is that of the
// nested-loop formulation this replaces.
// ------------------------------------------------------------------------------------------------------------------------
// Traversal stack access: a plain array (local memory on the device), or HybridStack — the first NS entries of every thread in
// shared memory ([entry][thread], conflict-free 8-byte accesses), the rest in local memory (persist.cuh, `stack_shared` option).
PTB_HD uint2 stk_get(const uint2* a, int i) { return a[i]; }
PTB_HD void stk_set(uint2* a, int i, uint2 v) { a[i] = v; }
template <int NS, int STRIDE>
struct HybridStack
{
    uint2* sh;   // &shared[0][thread]
    uint2* loc;  // entries NS.. in local memory
};
template <int NS, int STRIDE>
PTB_HD uint2 stk_get(const HybridStack<NS, STRIDE>& h, int i) { return i < NS ? h.sh[i * STRIDE] : h.loc[i - NS]; }
template <int NS, int STRIDE>
PTB_HD void stk_set(const HybridStack<NS, STRIDE>& h, int i, uint2 v)
{
    if (i < NS) h.sh[i * STRIDE] = v;
    else h.loc[i - NS] = v;
}

constexpr uint32_t TRAV_NONE = 0xffffffffu;
constexpr uint32_t TILE_NODES_SHIFT_ = 28;  // tile-list instance entries: leaf-header index | number of node entries << 28 (wavefront.h)
struct WideTrav
{
    V3 ow, dw, invw;    // world-space ray and its exact inv_direction (ray.rs:1-8)
    uint32_t oiw;
    TravRay ray;        // current space: world, or BLAS space after Ray::transform
    float t_max;        // closest-hit: current nearest t (tlas.rs:96 / blas.rs:236,247)
    uint32_t node;      // wide node to fetch next, or TRAV_NONE
    uint2 G;            // inner children of the current node still to visit
    int sp, blas_sp;
    uint32_t cur_inst;  // InstanceRow index while inside a BLAS, TRAV_NONE in the TLAS
    uint32_t lh;        // BLAS node: hit leaves still to process (permuted order, highest bit = nearest)
    uint32_t lmask, leaf_base;
    uint32_t tri_i, tri_end;  // triangle slots of the current leaf still to test
    float t_enter;
    float hit_ud, hit_vd, hit_det;  // winner's undivided barycentrics (u = ud / det, v = vd / det: wide_hit_u / wide_hit_v)
    uint32_t hit_slot, hit_inst;  // hit_inst == TRAV_NONE: nothing found yet
    uint32_t any;       // any-hit query (TLAS::any_intersect) — used when the step is instantiated with MODE == 2
    uint32_t cand, cand_end;  // LIST traversals (camera rays of a tile, wide_begin_list): next / end entry of the tile's list in S.tile_lists
};

// Conservative pre-test against the widened world bounds of TLAS `which`: false only if no point of the ray (t >= 0) lies in
// them, so no triangle of that TLAS can be hit.  Approximate reciprocals are enough (the margin absorbs the rounding); an
// axis with a (nearly) zero direction component constrains nothing.  Used for the BSDF-sampled light probes, most of which
// point nowhere near the lights: they end here instead of after a ray set-up and a node test.
PTB_HD bool tlas_may_hit(const SceneDev& S, int which, V3 o, V3 d)
{
    const TlasDev& T = S.tlas[which];
    float tn = 0.0f, tf = ex::inf();
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
#pragma unroll
    for (int k = 0; k < 3; ++k)
        if (fabsf(dd[k]) > 1e-30f)
        {
#if defined(__CUDA_ARCH__)
            const float inv = __fdividef(1.0f, dd[k]);
#else
            const float inv = 1.0f / dd[k];
#endif
            const float t0 = (T.box_lo[k] - oo[k]) * inv, t1 = (T.box_hi[k] - oo[k]) * inv;
            tn = fmaxf(tn, fminf(t0, t1));
            tf = fminf(tf, fmaxf(t0, t1));
        }
    return tn <= tf * 1.00001f + 1e-30f;
}
PTB_HD float wide_hit_u(const WideTrav& s) { return ex::div(s.hit_ud, s.hit_det); }
PTB_HD float wide_hit_v(const WideTrav& s) { return ex::div(s.hit_vd, s.hit_det); }
PTB_HD void wide_begin(WideTrav& s, const SceneDev& S, int which, V3 o, V3 d, float t_max, bool any)
{
    s.ow = o;
    s.dw = d;
    set_ray(s.ray, o, d);
    s.invw = s.ray.inv;
    s.oiw = s.ray.oi;
    s.t_max = t_max;
    s.node = S.tlas[which].wide_root;
    s.G.x = 0;
    s.G.y = 0;
    s.sp = 0;
    s.blas_sp = 0;
    s.cur_inst = TRAV_NONE;
    s.lh = 0;
    s.lmask = 0;
    s.leaf_base = 0;
    s.tri_i = 0;
    s.tri_end = 0;
    s.t_enter = 0.0f;
    s.hit_ud = s.hit_vd = 0.0f;
    s.hit_det = 1.0f;
    s.hit_slot = 0;
    s.hit_inst = TRAV_NONE;
    s.any = any ? 1u : 0u;
    s.cand = s.cand_end = 0u;
}
// closest hit over a precomputed list of TLAS leaves instead of the TLAS itself: the list holds every leaf the ray can reach (a superset,
// nearest first, each with a lower bound of its entry distance), the leaves are entered through the same exact box test, so the nearest
// hit is the one the TLAS walk finds (equal hits in different instances may resolve differently: the visit order differs)
PTB_HD void wide_begin_list(WideTrav& s, const SceneDev& S, V3 o, V3 d, float t_max, uint32_t first, uint32_t end)
{
    wide_begin(s, S, 0, o, d, t_max, false);
    s.node = TRAV_NONE;
    s.cand = first;
    s.cand_end = end;
}

PTB_HD int hi_bit(uint32_t m)
{
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)m);
#else
    return 31 - __builtin_clz(m);
#endif
}
PTB_HD int lo_bit(uint32_t m)
{
#if defined(__CUDA_ARCH__)
    return __ffs((int)m) - 1;
#else
    return __builtin_ctz(m);
#endif
}
PTB_HD uint32_t popc32(uint32_t m)
{
#if defined(__CUDA_ARCH__)
    return (uint32_t)__popc(m);
#else
    return (uint32_t)__builtin_popcount(m);
#endif
}

// MODE 0: closest hit, 1: any hit, 2: s.any decides at run time.  Returns false when the traversal is complete
// (closest: s.hit_inst != TRAV_NONE tells whether something was hit; any: s.hit_inst != TRAV_NONE == occluded).
// wide_step_pop: stage P.  Returns false when the traversal is complete (stack empty, nothing pending).
template <bool LIST = false, class Stack>
PTB_HD bool wide_step_pop(WideTrav& s, const SceneDev& S, Stack stack)
{
    // ---- P: nothing pending -> next node from the current group, or pop ----------------------------------------------
    if (s.tri_i >= s.tri_end && s.lh == 0u && s.node == TRAV_NONE)
    {
        if (s.G.y == 0u)
        {
            if (s.cur_inst != TRAV_NONE && s.sp == s.blas_sp)
            {
                s.cur_inst = TRAV_NONE;  // leave the instance: back to the world ray
                s.ray.o = s.ow;
                s.ray.d = s.dw;
                s.ray.inv = s.invw;
                s.ray.idq = ex::v3(finite_recip_of(s.invw.x), finite_recip_of(s.invw.y), finite_recip_of(s.invw.z));
                s.ray.oi = s.oiw;
            }
            uint2 e;
            uint32_t list_nodes = 0u;  // LIST: BLAS nodes the tile's frustum stays within, listed after the instance's entry
            if (s.sp == 0)
            {
                if (!LIST || s.cand >= s.cand_end) return false;
                const uint2 c = S.tile_lists[s.cand++];
                if (ex::from_bits(c.y) > s.t_max) { s.cand = s.cand_end; return false; }  // nearest first: nothing further can be nearer than the hit
                e.x = c.x & ((1u << TILE_NODES_SHIFT_) - 1u);
                e.y = 0x80000000u;
                list_nodes = c.x >> TILE_NODES_SHIFT_;
                s.cand += list_nodes;  // the instance's node entries (read below when its box is hit)
            }
            else e = stk_get(stack, --s.sp);
            if (e.y & 0x80000000u)
            {
                // TLAS leaf (tlas.rs:88-100 / :132-139): exact leaf box with the current t_max, then Ray::transform
                const float4* lp = S.leaf_headers + (size_t)e.x * LEAF_F4;
                const float4 h0 = PTB_LDG4F(lp), h1 = PTB_LDG4F(lp + 1), i0 = PTB_LDG4F(lp + 2), i1 = PTB_LDG4F(lp + 3), i2 = PTB_LDG4F(lp + 4);
                float t_enter;
                PTB_STAT(3);
                if (ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), s.ray.o, s.ray.inv, s.t_max, &t_enter))
                {
                    PTB_STAT(4);
                    s.cur_inst = ex::bits(h1.w);
                    const uint32_t cw = ex::bits(h0.w);
                    // inv_matrix travels in the leaf record (no dependent InstanceRow fetch); the arithmetic of Ray::transform is
                    // carried out even for the identity (x*1 + y*0 + ... decides the sign of zeros)
                    const float inv12[12] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w, i2.x, i2.y, i2.z, i2.w};
                    transform_ray_reuse(s.ray, s.ow, s.dw, s.invw, inv12);
                    s.blas_sp = s.sp;
                    if (LIST && list_nodes)
                    {
                        // start below the BLAS root: the listed nodes as one-child groups, farthest first on the stack, the nearest at hand
                        const uint32_t one = 1u | ((1u << s.ray.oi) << 8);  // descends to G.x: hit bit 0 -> slot oi, the only bit of the mask
                        for (uint32_t q = list_nodes - 1u; q > 0u; --q)
                        {
                            uint2 g;
                            g.x = S.tile_lists[s.cand - list_nodes + q].x;
                            g.y = one;
                            stk_set(stack, s.sp++, g);
                        }
                        s.G.x = S.tile_lists[s.cand - list_nodes].x;
                        s.G.y = one;
                    }
                    else if (cw & TLAS_DIRECT_FLAG)
                    {
                        // a BLAS of one or two leaves: straight to the exact leaf tests (stage L of this very step)
                        s.lmask = (cw >> TLAS_DIRECT_BASE_BITS) & 0xffu;
                        s.leaf_base = cw & ((1u << TLAS_DIRECT_BASE_BITS) - 1u);
                        s.lh = xor_permute16(s.lmask, s.ray.oi) & 0xffu;
                    }
                    else s.node = cw;
                }
            }
            else s.G = e;
        }
        if (s.node == TRAV_NONE && s.G.y)
        {
            // descend into the nearest remaining child of the current group
            const uint32_t ih = s.G.y & 0xffu, imask = (s.G.y >> 8) & 0xffu;
            const int p = hi_bit(ih);
            const uint32_t sl = (uint32_t)p ^ s.ray.oi;
            s.node = s.G.x + popc32(imask & ((1u << sl) - 1u));
            s.G.y &= ~(1u << p);
            if (s.G.y & 0xffu) stk_set(stack, s.sp++, s.G);
            s.G.y = 0u;
        }
    }
    return true;
}
PTB_HD bool wide_wants_node(const WideTrav& s) { return s.node != TRAV_NONE; }
PTB_HD bool wide_has_leaf_work(const WideTrav& s) { return s.lh != 0u || s.tri_i < s.tri_end; }

// wide_step_work: stages N, L, T.  `allow_node` lets a warp postpone the (expensive) node stage until enough of its lanes
// want it (persist.cuh); a lane whose node is postponed simply keeps it pending.  Returns false when an any-hit query is
// decided.
template <int MODE, class Stack>
PTB_HD bool wide_step_work(WideTrav& s, const SceneDev& S, Stack stack, bool allow_node)
{
    const bool ANY = MODE == 2 ? (s.any != 0u) : (MODE == 1);
    // ---- N: fetch and test one wide node ---------------------------------------------------------------------------------
    if (allow_node && s.node != TRAV_NONE)
    {
        const uint4* np = S.wide + (size_t)s.node * 5;
        const uint4 q0 = PTB_LDG4U(np), q1 = PTB_LDG4U(np + 1), q2 = PTB_LDG4U(np + 2), q3 = PTB_LDG4U(np + 3), q4 = PTB_LDG4U(np + 4);
        s.node = TRAV_NONE;
        PTB_STAT(s.cur_inst != TRAV_NONE ? 0 : 5);
        const uint32_t hits = wide_node_test(q0, q2, q3, q4, s.ray, s.t_max);
        const uint32_t imask = q0.w >> 24, lmask = q1.z & 0xffu;
        const uint32_t perm = xor_permute16((hits & imask) | ((hits & lmask) << 8), s.ray.oi);
        const uint32_t ih = perm & 0xffu, lh = (perm >> 8) & 0xffu;
        s.G.x = q1.x;
        s.G.y = ih ? (ih | (imask << 8)) : 0u;
        if (s.cur_inst != TRAV_NONE)
        {
            s.lh = lh;  // triangle leaves: processed front to back before any inner child
            s.lmask = lmask;
            s.leaf_base = q1.y;
        }
        else
        {
            // TLAS leaves: push the inner group first, then the instances far to near (nearest popped first)
            if (s.G.y) { stk_set(stack, s.sp++, s.G); s.G.y = 0u; }
            uint32_t m = lh;
            while (m)
            {
                const int p = lo_bit(m);
                m &= m - 1u;
                const uint32_t sl = (uint32_t)p ^ s.ray.oi;
                uint2 e;
                e.x = q1.y + popc32(lmask & ((1u << sl) - 1u));
                e.y = 0x80000000u;
                stk_set(stack, s.sp++, e);
            }
        }
    }
    // ---- L: one reference leaf of the current BLAS node -------------------------------------------------------------------
    float4 pa = {0.0f, 0.0f, 0.0f, 0.0f}, pb = pa, pc = pa;  // planes of the leaf's first triangle, fetched with its header
    bool fresh = false;
    if (s.tri_i >= s.tri_end && s.lh)
    {
        const int p = hi_bit(s.lh);
        s.lh &= ~(1u << p);
        const uint32_t sl = (uint32_t)p ^ s.ray.oi;
        const uint32_t hidx = s.leaf_base + popc32(s.lmask & ((1u << sl) - 1u));
        const float4* lp = S.leaf_headers + (size_t)hidx * LEAF_F4;
        const float4 h0 = PTB_LDG4F(lp), h1 = PTB_LDG4F(lp + 1);
        pa = PTB_LDG4F(lp + 2); pb = PTB_LDG4F(lp + 3); pc = PTB_LDG4F(lp + 4);
        const uint32_t cw = ex::bits(h0.w);
        float t_enter = 0.0f;
        bool go = true;
        PTB_STAT(1);
        // blas.rs:137-138 / :264 — exact reference leaf box, exact entry distance (the BLAS root is entered untested with
        // t_estimate = 0 by the closest-hit query, blas.rs:217)
        if (ANY || !(cw & LEAF_ROOT_FLAG)) go = ex::slab(ex::v3(h0.x, h0.y, h0.z), ex::v3(h1.x, h1.y, h1.z), s.ray.o, s.ray.inv, s.t_max, &t_enter);
        if (go)
        {
            s.tri_i = ex::bits(h1.w);
            s.tri_end = s.tri_i + (cw & ~LEAF_ROOT_FLAG);
            s.t_enter = t_enter;
            fresh = true;
        }
    }
    // ---- T: one triangle ---------------------------------------------------------------------------------------------------
    if (s.tri_i < s.tri_end)
    {
        PTB_STAT(2);
        float4 a = pa, b = pb, c = pc;
        if (!fresh)
        {
            const float4* pp = S.tri_planes + (size_t)s.tri_i * 3;
            a = PTB_LDG4F(pp); b = PTB_LDG4F(pp + 1); c = PTB_LDG4F(pp + 2);
        }
        const float n0[4] = {a.x, a.y, a.z, a.w}, n1[4] = {b.x, b.y, b.z, b.w}, n2[4] = {c.x, c.y, c.z, c.w};
        if (ANY)
        {
            if (ex::tri_hit_bool(n0, n1, n2, s.ray.o, s.ray.d, s.t_max, s.t_enter))
            {
                s.hit_inst = s.cur_inst;
                return false;
            }
        }
        else
        {
            float t, raw[3];
            if (ex::tri_hit_raw(n0, n1, n2, s.ray.o, s.ray.d, s.t_max, s.t_enter, &t, raw))
            {
                s.t_max = t;
                s.hit_ud = raw[0]; s.hit_vd = raw[1]; s.hit_det = raw[2]; s.hit_slot = s.tri_i; s.hit_inst = s.cur_inst;
            }
        }
        ++s.tri_i;
    }
    return true;
}

template <int MODE>
PTB_HD bool wide_step(WideTrav& s, const SceneDev& S, uint2* stack)
{
    return wide_step_pop(s, S, stack) && wide_step_work<MODE>(s, S, stack, true);
}

template <bool ANY>
PTB_HD bool wide_traverse(const SceneDev& S, int which, V3 o_world, V3 d_world, float t_max, Hit* hit)
{
    uint2 stack[WIDE_STACK_CAP];
    WideTrav s;
    wide_begin(s, S, which, o_world, d_world, t_max, ANY);
    while (wide_step<ANY ? 1 : 0>(s, S, stack)) {}
    if (s.hit_inst == TRAV_NONE) return false;
    if (!ANY) { hit->t = s.t_max; hit->u = wide_hit_u(s); hit->v = wide_hit_v(s); hit->slot_or_prim = s.hit_slot; hit->inst = s.hit_inst; }
    return true;
}

// ------------------------------------------------------------------------------------------------------------------------
// Reference-order traversal over the binary trees (BinNode): a literal walk of tlas.rs:66-144 / blas.rs:214-294.
// ------------------------------------------------------------------------------------------------------------------------
struct BinEntry
{
    uint32_t id;
    float t;
};

PTB_HD void bin_push2(const BinNode* nodes, uint32_t base, uint32_t l, uint32_t r, const TravRay& ray, float t_max, BinEntry* st, int& sp)  // blas.rs:133-162
{
    const BinNode& L = nodes[base + l];
    const BinNode& Rn = nodes[base + r];
    float tl, tr;
    const bool hl = ex::slab(ex::v3(PTB_LDGF(&L.lo[0]), PTB_LDGF(&L.lo[1]), PTB_LDGF(&L.lo[2])), ex::v3(PTB_LDGF(&L.hi[0]), PTB_LDGF(&L.hi[1]), PTB_LDGF(&L.hi[2])), ray.o, ray.inv, t_max, &tl);
    const bool hr = ex::slab(ex::v3(PTB_LDGF(&Rn.lo[0]), PTB_LDGF(&Rn.lo[1]), PTB_LDGF(&Rn.lo[2])), ex::v3(PTB_LDGF(&Rn.hi[0]), PTB_LDGF(&Rn.hi[1]), PTB_LDGF(&Rn.hi[2])), ray.o, ray.inv, t_max, &tr);
    if (hl && hr)
    {
        if (tl < tr) { st[sp].id = r; st[sp++].t = tr; st[sp].id = l; st[sp++].t = tl; }
        else { st[sp].id = l; st[sp++].t = tl; st[sp].id = r; st[sp++].t = tr; }
    }
    else if (hl) { st[sp].id = l; st[sp++].t = tl; }
    else if (hr) { st[sp].id = r; st[sp++].t = tr; }
}

PTB_HD void load_planes(const SceneDev& S, uint32_t gtri, float n0[4], float n1[4], float n2[4])
{
    const float4* pp = S.planes_by_prim + (size_t)gtri * 3;
    const float4 a = PTB_LDG4F(pp), b = PTB_LDG4F(pp + 1), c = PTB_LDG4F(pp + 2);
    n0[0] = a.x; n0[1] = a.y; n0[2] = a.z; n0[3] = a.w;
    n1[0] = b.x; n1[1] = b.y; n1[2] = b.z; n1[3] = b.w;
    n2[0] = c.x; n2[1] = c.y; n2[2] = c.z; n2[3] = c.w;
}

template <bool ANY>
PTB_HD bool bin_traverse(const SceneDev& S, int which, V3 o_world, V3 d_world, float t_max, Hit* hit)
{
    BinEntry st[BIN_STACK_CAP];
    int sp = 0;
    TravRay wray;
    set_ray(wray, o_world, d_world);
    const TlasDev& T = S.tlas[which];
    const BinNode* nodes = S.bin;
    bool found = false;
    if (!ANY)
    {
        const BinNode& rt = nodes[T.bin_base + T.bin_root];  // tlas.rs:68
        float te;
        if (!ex::slab(ex::v3(rt.lo[0], rt.lo[1], rt.lo[2]), ex::v3(rt.hi[0], rt.hi[1], rt.hi[2]), wray.o, wray.inv, t_max, &te)) return false;
    }
    st[sp].id = T.bin_root;
    st[sp++].t = 0.0f;
    while (sp > 0)
    {
        const BinEntry cur = st[--sp];
        const BinNode& n = nodes[T.bin_base + cur.id];
        const uint32_t na = PTB_LDGU(&n.a), nb = PTB_LDGU(&n.b);
        if (ANY)
        {
            float te;  // tlas.rs:118
            if (!ex::slab(ex::v3(n.lo[0], n.lo[1], n.lo[2]), ex::v3(n.hi[0], n.hi[1], n.hi[2]), wray.o, wray.inv, t_max, &te)) continue;
        }
        else if (cur.t > t_max) continue;  // tlas.rs:80-83
        if ((na >> BIN_KIND_SHIFT) == 0)
        {
            if (ANY) { st[sp].id = na; st[sp++].t = 0.0f; st[sp].id = nb; st[sp++].t = 0.0f; }  // tlas.rs:125-130
            else bin_push2(nodes, T.bin_base, na, nb, wray, t_max, st, sp);
            continue;
        }
        // TLAS leaf: transform and walk the BLAS with its own stack region above sp
        const InstanceRow& R = S.inst[nb];
        TravRay ray;
        float inv12[12];
        for (int k = 0; k < 12; ++k) inv12[k] = PTB_LDGF(R.inv + k);
        transform_ray(ray, o_world, d_world, inv12);
        const uint32_t bbase = PTB_LDGU(&R.bin_base), tri_base = PTB_LDGU(&R.tri_base), lp_base = PTB_LDGU(&R.leafprim_base);
        const int floor_sp = sp;
        st[sp].id = PTB_LDGU(&R.bin_root);
        st[sp++].t = 0.0f;  // blas.rs:217: root pushed with 0.0, not box-tested (closest); any-hit tests every popped node (:264)
        while (sp > floor_sp)
        {
            const BinEntry c = st[--sp];
            const BinNode& bn = nodes[bbase + c.id];
            const uint32_t a = PTB_LDGU(&bn.a), b = PTB_LDGU(&bn.b);
            float t_enter = c.t;
            if (ANY)
            {
                if (!ex::slab(ex::v3(bn.lo[0], bn.lo[1], bn.lo[2]), ex::v3(bn.hi[0], bn.hi[1], bn.hi[2]), ray.o, ray.inv, t_max, &t_enter)) continue;
            }
            else if (c.t > t_max) continue;
            const uint32_t kind = a >> BIN_KIND_SHIFT, idx = a & BIN_INDEX_MASK;
            if (kind == BIN_BRANCH)
            {
                if (ANY) { st[sp].id = idx; st[sp++].t = 0.0f; st[sp].id = b; st[sp++].t = 0.0f; }
                else bin_push2(nodes, bbase, idx, b, ray, t_max, st, sp);
                continue;
            }
            const uint32_t cnt = kind == BIN_LEAF_SINGLE ? 1u : b;
            for (uint32_t k = 0; k < cnt; ++k)
            {
                const uint32_t prim = kind == BIN_LEAF_SINGLE ? idx : PTB_LDGU(S.leaf_prims + lp_base + idx + k);
                float n0[4], n1[4], n2[4];
                load_planes(S, tri_base + prim, n0, n1, n2);
                if (ANY)
                {
                    if (ex::tri_hit_bool(n0, n1, n2, ray.o, ray.d, t_max, t_enter)) return true;
                }
                else
                {
                    float t, u, v;
                    if (ex::tri_hit(n0, n1, n2, ray.o, ray.d, t_max, t_enter, &t, &u, &v))
                    {
                        t_max = t;
                        hit->t = t; hit->u = u; hit->v = v; hit->slot_or_prim = prim; hit->inst = nb;
                        found = true;
                    }
                }
            }
        }
    }
    return found;
}

}  // namespace ptb

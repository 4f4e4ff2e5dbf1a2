// bvh8_traverse.cuh -- ray vs BVH8: the 8-child quantised slab test and the traversal loop.
// Replaces the while-while BVH2 loop of detail/bvh/intersect.inl:35-134 on the fast path.
#pragma once

#include "bvh8.cuh"

namespace vsb
{

struct WideScene
{
    const Node80* nodes;      // BVH8 nodes, root = 0
    const void*   prims;      // TriRec[] or SphRec[] in leaf order
    int           prim_kind;  // 0 triangles, 1 spheres
    const uint32_t* rec_of_list = nullptr;   // primitive_list_index -> leaf-order record (the queue kernel's keys carry the list index)
};

struct RayCtx
{
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;          // 1/dir with |dir| clamped away from 0 (box test only)
    uint32_t octinv4;          // (7 - octant) replicated in 4 bytes
};

VSB_HD void ray_ctx_init(RayCtx& r, float ox, float oy, float oz, float dx, float dy, float dz)
{
    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz;
    const float eps = 8.2718061e-25f;   // 2^-80: keeps (p - o) * (1/eps) finite
    float cx = fabsf(dx) > eps ? dx : (f2u(dx) >> 31 ? -eps : eps);
    float cy = fabsf(dy) > eps ? dy : (f2u(dy) >> 31 ? -eps : eps);
    float cz = fabsf(dz) > eps ? dz : (f2u(dz) >> 31 ? -eps : eps);
    r.ix = 1.0f / cx; r.iy = 1.0f / cy; r.iz = 1.0f / cz;
    uint32_t oct = (cx < 0.0f ? 1u : 0u) | (cy < 0.0f ? 2u : 0u) | (cz < 0.0f ? 4u : 0u);
    r.octinv4 = (7u - oct) * 0x01010101u;
}

struct NodeRegs { uint32_t w[20]; };

// Loads n16 (2, 3 or 5) 128-bit words of a 16-byte aligned record: a BVH8 node (5), a leaf-order
// triangle (3) or sphere (2).  One code path for nodes and primitives, so a warp whose lanes are
// in different phases still issues all its loads together.
#ifndef VSB_L2_HINT
#define VSB_L2_HINT 0            // 1: L2 eviction priority by record kind (nodes evict_last, primitives evict_first)
#endif

#if defined(__CUDA_ARCH__) && VSB_L2_HINT
__device__ __forceinline__ uint4 ldg_hint(const uint4* p, unsigned long long policy)
{
    uint4 v;
    asm("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(policy));
    return v;
}
__device__ __forceinline__ unsigned long long l2_policy(bool keep)
{
    unsigned long long pol;
    if (keep) asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    else      asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
#endif

VSB_HD void load_record(const void* __restrict__ addr, int n16, NodeRegs& n)
{
#if defined(__CUDA_ARCH__)
    const uint4* p = reinterpret_cast<const uint4*>(addr);
#if VSB_L2_HINT
    // the node array (10 % of the scene bytes, re-visited by every ray) should outlive the
    // leaf-order primitives (visited once in a while) in L2
    const unsigned long long pol = n16 > 3 ? l2_policy(true) : l2_policy(false);
    const uint4 a = ldg_hint(p, pol), b = ldg_hint(p + 1, pol);
    uint4 c = make_uint4(0u, 0u, 0u, 0u), d = c, e = c;
    if (n16 > 2) c = ldg_hint(p + 2, pol);
    if (n16 > 3) { d = ldg_hint(p + 3, pol); e = ldg_hint(p + 4, pol); }
#else
    const uint4 a = __ldg(p), b = __ldg(p + 1);
    uint4 c = make_uint4(0u, 0u, 0u, 0u), d = c, e = c;
    if (n16 > 2) c = __ldg(p + 2);
    if (n16 > 3) { d = __ldg(p + 3); e = __ldg(p + 4); }
#endif
    n.w[0] = a.x; n.w[1] = a.y; n.w[2] = a.z; n.w[3] = a.w;
    n.w[4] = b.x; n.w[5] = b.y; n.w[6] = b.z; n.w[7] = b.w;
    n.w[8] = c.x; n.w[9] = c.y; n.w[10] = c.z; n.w[11] = c.w;
    n.w[12] = d.x; n.w[13] = d.y; n.w[14] = d.z; n.w[15] = d.w;
    n.w[16] = e.x; n.w[17] = e.y; n.w[18] = e.z; n.w[19] = e.w;
#else
    for (int i = 0; i < 20; ++i) n.w[i] = 0u;
    memcpy(n.w, addr, (size_t)n16 * 16u);
#endif
}

// bit i of the result = bit (i ^ x) of h, for an 8-bit mask h and x in 0..7
VSB_HD uint32_t perm_xor8(uint32_t h, uint32_t x)
{
    if (x & 1u) h = ((h & 0x55u) << 1) | ((h >> 1) & 0x55u);
    if (x & 2u) h = ((h & 0x33u) << 2) | ((h >> 2) & 0x33u);
    if (x & 4u) h = ((h & 0x0fu) << 4) | ((h >> 4) & 0x0fu);
    return h;
}

// 8 child bits -> 8 nibbles (bit i -> 0xF << 4 i).  The shifts are multiplications (IMAD: the fma pipe, which has room;
// the alu pipe is the one the kernel is bound by); no two partial products overlap, so + is |.
VSB_HD uint32_t expand_nibbles(uint32_t h)
{
    uint32_t t = (h * 4097u) & 0x000f000fu;       // bits 0..3 stay, bits 4..7 -> 16..19
    t = (t * 65u) & 0x03030303u;                  // pairs
    t = (t * 9u) & 0x11111111u;                   // singles, one per nibble
    return t * 15u;
}

// (acc << 1) | sign(d)
VSB_HD uint32_t funnel_shift_in_sign(uint32_t acc, float d)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(__float_as_uint(d), acc, 1);
#else
    return (acc << 1) | (f2u(d) >> 31);
#endif
}

// leaf-order record of bit `bit` of a primitive mask (slot space: bit 4 s + j = primitive j of the leaf in slot s)
VSB_HD uint32_t prim_record(uint32_t prim_base, uint32_t valid, int bit)
{
    return prim_base + (uint32_t)pop_count(valid & ((1u << bit) - 1u));
}

// Tests the 8 children of a node against the ray.  Returns the node hit mask: bits 24..31 = inner
// children (bit = 24 + (slot ^ octinv), so the highest set bit is the nearest slot); prim_hits
// gets one bit per leaf primitive of every hit leaf child, in SLOT space (bit 4 s + j; n.w[6] = the node's
// `valid` mask maps a bit to its record: prim_record()).
#ifndef VSB_NODE_FHFMA
#define VSB_NODE_FHFMA 0       // 1: plane * scale + offset as ONE mixed-precision FMA (f16 x f16 + f32, sm_100 FHFMA) -- an experiment, see below
#endif

#if VSB_NODE_FHFMA
// d = a.half[AH] * b.half[BH] + c: fma.rn.f32.f16 (SASS FHFMA; the half selectors are operand modifiers).  The product of two
// halves is exact in f32, so the only rounding is the final one, as in an FFMA.  Host: emulated (tests/hostsim counts
// node visits with it; the parity tests keep the all-f32 form on the host).
#if defined(__CUDA_ARCH__)
template <int AH, int BH>
__device__ __forceinline__ float fhfma(uint32_t a, uint32_t b, float c)
{
    float d;
    if (AH == 0 && BH == 0) asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.f16 %0, al, bl, %3; }" : "=f"(d) : "r"(a), "r"(b), "f"(c));
    if (AH == 0 && BH == 1) asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.f16 %0, al, bh, %3; }" : "=f"(d) : "r"(a), "r"(b), "f"(c));
    if (AH == 1 && BH == 0) asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.f16 %0, ah, bl, %3; }" : "=f"(d) : "r"(a), "r"(b), "f"(c));
    if (AH == 1 && BH == 1) asm("{ .reg .b16 al, ah, bl, bh; mov.b32 {al, ah}, %1; mov.b32 {bl, bh}, %2; fma.rn.f32.f16 %0, ah, bh, %3; }" : "=f"(d) : "r"(a), "r"(b), "f"(c));
    return d;
}
__device__ __forceinline__ uint32_t pack_half2(float lo, float hi)
{
    uint32_t d;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
#elif defined(VSB_HOSTSIM)
inline float half_bits_to_float(uint32_t h) { _Float16 x; uint16_t b = (uint16_t)h; memcpy(&x, &b, 2); return (float)x; }
template <int AH, int BH>
inline float fhfma(uint32_t a, uint32_t b, float c)
{
    const double pa = half_bits_to_float(AH ? a >> 16 : a & 0xffffu), pb = half_bits_to_float(BH ? b >> 16 : b & 0xffffu);
    return (float)(pa * pb + (double)c);
}
inline uint32_t pack_half2(float lo, float hi)
{
    const _Float16 l = (_Float16)lo, h = (_Float16)hi;
    uint16_t bl, bh; memcpy(&bl, &l, 2); memcpy(&bh, &h, 2);
    return (uint32_t)bl | ((uint32_t)bh << 16);
}
#else      // nvcc's host pass: never called
template <int AH, int BH> inline float fhfma(uint32_t, uint32_t, float c) { return c; }
inline uint32_t pack_half2(float, float) { return 0u; }
#endif

// The node test with the plane evaluation in FHFMA.  The byte b is the half 1024 + b (one PRMT per two children, no widening
// conversion any more: 48 fewer instructions per visit), the scale S = s * idir becomes a half too.  To keep S inside the
// half's normal range whatever the scene's units and the ray's direction, the node's t axis is rescaled by a power of two
// K so that the largest |S| lands in [2^7, 2^8): t' = K t, exact, and every comparison of the test is invariant under it.
// What the half costs: S_h = S' (1 + d), |d| <= 2^-11 (or an absolute 2^-25 where an axis is > 2^21 times faster than the
// slowest and S' falls into the subnormals).  The offset is formed with the ROUNDED scale, A = O' - 1024 S_h, so the
// evaluation returns O' + b S_h and the error is |b (S_h - S')| <= 255 * 2^-11 |S'| + 255 * 2^-25: an eighth of a grid cell,
// added to the margin that already covers the reference's own slab arithmetic.  NaNs (inf - inf in absurd ranges) are
// ignored by FMNMX, i.e. drop a constraint: conservative.  The host instantiation keeps the all-f32 form; both accept a
// superset of the reference's exact slab test, and hit records do not depend on which superset.
VSB_HD uint32_t test_node_slots_fhfma(const NodeRegs& n, const RayCtx& r, float limit, uint32_t& prim_hits)
{
    const float px = u2f(n.w[0]), py = u2f(n.w[1]), pz = u2f(n.w[2]);
    const uint32_t e = n.w[3];
    const float sx = u2f((e & 0xffu) << 23) * r.ix, sy = u2f(((e >> 8) & 0xffu) << 23) * r.iy, sz = u2f(((e >> 16) & 0xffu) << 23) * r.iz;
    const float ms = fmaxf(fmaxf(fabsf(sx), fabsf(sy)), fmaxf(fabsf(sz), 7.888609052210118e-31f));     // >= 2^-100
    const float K = u2f(0x82800000u - (f2u(ms) & 0x7f800000u));                                          // 2^(7 - floor(log2 ms))
    const float sxs = sx * K, sys = sy * K, szs = sz * K;
    const float oxs = (px - r.ox) * r.ix * K, oys = (py - r.oy) * r.iy * K, ozs = (pz - r.oz) * r.iz * K;
    const float lim = limit * K;
    const uint32_t sxy = pack_half2(sxs, sys), szz = pack_half2(szs, szs);
    const uint32_t m1024 = 0xe400e400u;                                                                   // half2(-1024, -1024)
    const float bx = fhfma<0, 0>(m1024, sxy, oxs), by = fhfma<0, 1>(m1024, sxy, oys), bz = fhfma<0, 0>(m1024, szz, ozs);
    const float k1 = 9.5367431640625e-07f + 2.384185791015625e-07f;                                       // 2^-20 + 2^-22
    const float k2 = 255.0f * 9.5367431640625e-07f + 0.000244140625f + 0.1246f;                           // + 255 * 2^-11 (+ slack)
    const float c0 = 7.62939453125e-06f;                                                                  // 256 * 2^-25
    const float mx = fma_f(fabsf(sxs), k2, fma_f(fabsf(oxs), k1, c0));
    const float my = fma_f(fabsf(sys), k2, fma_f(fabsf(oys), k1, c0));
    const float mz = fma_f(fabsf(szs), k2, fma_f(fabsf(ozs), k1, c0));
    const float onx = bx - mx, ofx = bx + mx;
    const float ony = by - my, ofy = by + my;
    const float onz = bz - mz, ofz = bz + mz;
    const bool nx = r.ix < 0.0f, ny = r.iy < 0.0f, nz = r.iz < 0.0f;
    const uint32_t bias = half_bias_bits();
    uint32_t miss = 0u;
#pragma unroll
    for (int half = 1; half >= 0; --half)
    {
        const uint32_t qlox = n.w[8 + half], qloy = n.w[10 + half], qloz = n.w[12 + half];
        const uint32_t qhix = n.w[14 + half], qhiy = n.w[16 + half], qhiz = n.w[18 + half];
        const uint32_t qnx = nx ? qhix : qlox, qfx = nx ? qlox : qhix;
        const uint32_t qny = ny ? qhiy : qloy, qfy = ny ? qloy : qhiy;
        const uint32_t qnz = nz ? qhiz : qloz, qfz = nz ? qloz : qhiz;
#pragma unroll
        for (int j = 1; j >= 0; --j)
        {
            const uint32_t sel = j ? 0x4342u : 0x4140u;
            const uint32_t pnx = byte_perm(qnx, bias, sel), pny = byte_perm(qny, bias, sel), pnz = byte_perm(qnz, bias, sel);
            const uint32_t pfx = byte_perm(qfx, bias, sel), pfy = byte_perm(qfy, bias, sel), pfz = byte_perm(qfz, bias, sel);
            {
                const float tn = fmaxf(fmaxf(fhfma<1, 0>(pnx, sxy, onx), fhfma<1, 1>(pny, sxy, ony)), fhfma<1, 0>(pnz, szz, onz));
                const float tf = fminf(fminf(fhfma<1, 0>(pfx, sxy, ofx), fhfma<1, 1>(pfy, sxy, ofy)), fhfma<1, 0>(pfz, szz, ofz));
                miss = funnel_shift_in_sign(miss, fminf(fminf(tf - tn, lim - tn), tf));
            }
            {
                const float tn = fmaxf(fmaxf(fhfma<0, 0>(pnx, sxy, onx), fhfma<0, 1>(pny, sxy, ony)), fhfma<0, 0>(pnz, szz, onz));
                const float tf = fminf(fminf(fhfma<0, 0>(pfx, sxy, ofx), fhfma<0, 1>(pfy, sxy, ofy)), fhfma<0, 0>(pfz, szz, ofz));
                miss = funnel_shift_in_sign(miss, fminf(fminf(tf - tn, lim - tn), tf));
            }
        }
    }
    const uint32_t child_hits = ~miss & 0xffu;
    prim_hits = expand_nibbles(child_hits) & n.w[6];
    return child_hits & (n.w[3] >> 24);
}
#endif

#ifndef VSB_HOST_NODE_FHFMA
#define VSB_HOST_NODE_FHFMA 0     // hostsim: 1 = the host also uses the FHFMA form (emulated), for visit counts
#endif

// test_node_slots: the inner-child hits in SLOT space (8 bits); test_node: permuted into traversal order, bits 24..31
VSB_HD uint32_t test_node_slots(const NodeRegs& n, const RayCtx& r, float limit, uint32_t& prim_hits)
{
#if VSB_NODE_FHFMA && (defined(__CUDA_ARCH__) || VSB_HOST_NODE_FHFMA)
    return test_node_slots_fhfma(n, r, limit, prim_hits);
#endif
    const float px = u2f(n.w[0]), py = u2f(n.w[1]), pz = u2f(n.w[2]);
    const uint32_t e = n.w[3];
#if VSB_NODE_HALFCVT
    // the byte b enters as f = 1024 + b (byte_as_biased_float), so t = f * S + O with S = s * idir,
    // O = (p - o) * idir - 1024 S
    const float sxs = u2f((e & 0xffu) << 23) * r.ix;
    const float sys = u2f(((e >> 8) & 0xffu) << 23) * r.iy;
    const float szs = u2f(((e >> 16) & 0xffu) << 23) * r.iz;
    const float oxs = (px - r.ox) * r.ix;
    const float oys = (py - r.oy) * r.iy;
    const float ozs = (pz - r.oz) * r.iz;
    // error bound of this evaluation + of the reference's own slab arithmetic (see bvh8.cuh):
    // 2^-20 * (|O| + 255 |s idir|) + 2^-12 |s idir| (the rounding of O - 1024 S is 2^-14 |S|)
    const float k1 = 9.5367431640625e-07f;                 // 2^-20
    const float k2 = 255.0f * 9.5367431640625e-07f + 0.000244140625f;
    const float mx = fma_f(fabsf(sxs), k2, fabsf(oxs) * k1);
    const float my = fma_f(fabsf(sys), k2, fabsf(oys) * k1);
    const float mz = fma_f(fabsf(szs), k2, fabsf(ozs) * k1);
    const float bx = fma_f(sxs, -1024.0f, oxs), by = fma_f(sys, -1024.0f, oys), bz = fma_f(szs, -1024.0f, ozs);
    const float onx = bx - mx, ofx = bx + mx;
    const float ony = by - my, ofy = by + my;
    const float onz = bz - mz, ofz = bz + mz;
#else
    // grid scale per axis, pre-multiplied by 2^15 for the byte->float trick: the byte b enters
    // as f = 1 + b * 2^-15, so t = f * S + O with S = s * idir * 2^15, O = (p - o) * idir - S
    const float sxs = u2f(((e & 0xffu) + 15u) << 23) * r.ix;
    const float sys = u2f((((e >> 8) & 0xffu) + 15u) << 23) * r.iy;
    const float szs = u2f((((e >> 16) & 0xffu) + 15u) << 23) * r.iz;
    const float oxs = (px - r.ox) * r.ix;
    const float oys = (py - r.oy) * r.iy;
    const float ozs = (pz - r.oz) * r.iz;
    // error bound of this evaluation + of the reference's own slab arithmetic (see bvh8.cuh):
    // 2^-20 * (|O| + 255 |s idir|) + 2^-8 |s idir|, with |s idir| = |S| * 2^-15
    const float k1 = 9.5367431640625e-07f;                 // 2^-20
    const float k2 = (255.0f * 9.5367431640625e-07f + 0.00390625f) * 3.0517578125e-05f;   // per |S|
    const float mx = fma_f(fabsf(sxs), k2, fabsf(oxs) * k1);
    const float my = fma_f(fabsf(sys), k2, fabsf(oys) * k1);
    const float mz = fma_f(fabsf(szs), k2, fabsf(ozs) * k1);
    const float onx = (oxs - sxs) - mx, ofx = (oxs - sxs) + mx;
    const float ony = (oys - sys) - my, ofy = (oys - sys) + my;
    const float onz = (ozs - szs) - mz, ofz = (ozs - szs) + mz;

#endif

    // near/far planes by direction sign (selected once per node, not per child)
    const bool nx = r.ix < 0.0f, ny = r.iy < 0.0f, nz = r.iz < 0.0f;
#if VSB_NODE_HALFCVT
    const uint32_t one = half_bias_bits();
#define VSB_PLANE_COORD byte_as_biased_float
#else
    const uint32_t one = unit_float_bits();
#define VSB_PLANE_COORD byte_as_unit_float
#endif
    // children 7 .. 0: the verdict of a child is a sign (1 = miss; equal floats differ by +0, and a difference never
    // rounds across 0); it is shifted into `miss` from the right, so child c ends at bit c -- one funnel shift per child
    // instead of a compare, a select and an OR (the alu pipe is the kernel's limiter)
    uint32_t miss = 0u;
#pragma unroll
    for (int half = 1; half >= 0; --half)
    {
        const uint32_t qlox = n.w[8 + half], qloy = n.w[10 + half], qloz = n.w[12 + half];
        const uint32_t qhix = n.w[14 + half], qhiy = n.w[16 + half], qhiz = n.w[18 + half];
        const uint32_t qnx = nx ? qhix : qlox, qfx = nx ? qlox : qhix;
        const uint32_t qny = ny ? qhiy : qloy, qfy = ny ? qloy : qhiy;
        const uint32_t qnz = nz ? qhiz : qloz, qfz = nz ? qloz : qhiz;
#pragma unroll
        for (int k = 3; k >= 0; --k)
        {
            const float tnx = fma_f(VSB_PLANE_COORD(qnx, k, one), sxs, onx);
            const float tny = fma_f(VSB_PLANE_COORD(qny, k, one), sys, ony);
            const float tnz = fma_f(VSB_PLANE_COORD(qnz, k, one), szs, onz);
            const float tfx = fma_f(VSB_PLANE_COORD(qfx, k, one), sxs, ofx);
            const float tfy = fma_f(VSB_PLANE_COORD(qfy, k, one), sys, ofy);
            const float tfz = fma_f(VSB_PLANE_COORD(qfz, k, one), szs, ofz);
            // max(tn, 0) <= min(tf, limit)  <=>  tn <= tf, tn <= limit, 0 <= tf (limit >= 0): three differences whose
            // minimum carries the verdict in its sign -- 3 alu-pipe operations per child instead of 4 (the subtractions
            // run on the fma pipe).  tf is never -0.0: the offsets are sums with a non-negative margin.
            const float tn = fmaxf(fmaxf(tnx, tny), tnz);
            const float tf = fminf(fminf(tfx, tfy), tfz);
            miss = funnel_shift_in_sign(miss, fminf(fminf(tf - tn, limit - tn), tf));
        }
    }
    const uint32_t child_hits = ~miss & 0xffu;
#undef VSB_PLANE_COORD
    prim_hits = expand_nibbles(child_hits) & n.w[6];
    return child_hits & (n.w[3] >> 24);
}

VSB_HD uint32_t test_node(const NodeRegs& n, const RayCtx& r, float limit, uint32_t& prim_hits)
{
    return perm_xor8(test_node_slots(n, r, limit, prim_hits), r.octinv4 & 7u) << 24;
}

#if defined(__CUDACC__)
// perm_xor8 as a table in shared memory (8 permutations x 256 masks = 2 KiB per CTA): one LOP3 + one LDS instead of
// ~10 alu-pipe instructions per node visit.  The CTA fills the table once (perm_lut_fill, then __syncthreads).
__device__ __forceinline__ void perm_lut_fill(uint8_t* lut)
{
    for (unsigned i = threadIdx.x; i < 2048u; i += blockDim.x) lut[i] = (uint8_t)perm_xor8(i & 0xffu, i >> 8);
}
__device__ __forceinline__ uint32_t test_node_lut(const NodeRegs& n, const RayCtx& r, float limit, uint32_t& prim_hits, const uint8_t* lut)
{
    const uint32_t h = test_node_slots(n, r, limit, prim_hits);
    return (uint32_t)lut[(r.octinv4 & 0x700u) | h] * 0x1000000u;
}
#endif

// Primitive test on a loaded leaf-order record + is_closer(hr, result, max_t)
// (update_if.h:50-56,75-79).  `limit` = min(result.t, max_t).
VSB_HD void test_prim(int prim_kind, const NodeRegs& n, const RayCtx& r, HitReg& best, float& limit)
{
    if (prim_kind == 0 || (prim_kind == 2 && !(n.w[11] & VSB_GENERIC_SPHERE_TAG)))
    {
        // TriRec: v1 | prim_id, e1 | geom_id, e2 | list_index
        float t = -1.0f, u = 0.0f, v = 0.0f;      // initialised: left undefined, u / v were kept in local memory across the early-outs
        if (isect_triangle(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, u2f(n.w[0]), u2f(n.w[1]), u2f(n.w[2]),
                           u2f(n.w[4]), u2f(n.w[5]), u2f(n.w[6]), u2f(n.w[8]), u2f(n.w[9]), u2f(n.w[10]), t, u, v)
            && t >= 0.0f && t < limit)
        {
            best.t = t; best.u = u; best.v = v; best.hit = 1;
            best.prim_id = n.w[3]; best.geom_id = n.w[7]; best.list_index = n.w[11];
            limit = t;
        }
    }
    else
    {
        // SphRec: center, radius | prim_id, geom_id, list_index, pad
        float t;
        if (isect_sphere(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, u2f(n.w[0]), u2f(n.w[1]), u2f(n.w[2]), u2f(n.w[3]), t)
            && t >= 0.0f && t < limit)
        {
            best.t = t; best.u = 0.0f; best.v = 0.0f; best.hit = 1;
            best.prim_id = n.w[4]; best.geom_id = n.w[5]; best.list_index = n.w[6];
            limit = t;
        }
    }
}

// The primitive test on a loaded leaf-order record as a pure function (the queue kernel tests a
// primitive for ANOTHER lane's ray): true when the reference's test hits, t = its distance
// (no range test).  Same arithmetic as test_prim().
VSB_HD bool prim_candidate(int prim_kind, const NodeRegs& n, float ox, float oy, float oz, float dx, float dy, float dz, float& t)
{
    if (prim_kind == 0 || (prim_kind == 2 && !(n.w[11] & VSB_GENERIC_SPHERE_TAG)))
    {
        float u = 0.0f, v = 0.0f;
        t = -1.0f;
        return isect_triangle(ox, oy, oz, dx, dy, dz, u2f(n.w[0]), u2f(n.w[1]), u2f(n.w[2]),
                              u2f(n.w[4]), u2f(n.w[5]), u2f(n.w[6]), u2f(n.w[8]), u2f(n.w[9]), u2f(n.w[10]), t, u, v);
    }
    return isect_sphere(ox, oy, oz, dx, dy, dz, u2f(n.w[0]), u2f(n.w[1]), u2f(n.w[2]), u2f(n.w[3]), t);
}

// primitive_list_index (position in the reference's indices[]) of a loaded leaf-order record
VSB_HD uint32_t record_list_index(int prim_kind, const NodeRegs& n)
{
    if (prim_kind == 1 || (prim_kind == 2 && (n.w[11] & VSB_GENERIC_SPHERE_TAG))) return n.w[6];
    return n.w[11];
}

// The hit record of leaf-order record `rec` for a ray: the test is evaluated again (bit-identical
// t / u / v), ids come from the record.  Used once per ray, when the queue kernel hands the
// winner of its (t, record) key to the program.
VSB_HD void hit_from_record(const WideScene& s, uint32_t rec, float ox, float oy, float oz, float dx, float dy, float dz, HitReg& best)
{
    NodeRegs n;
    const int n16 = wide_rec_words16(s.prim_kind);
    load_record(reinterpret_cast<const char*>(s.prims) + (size_t)rec * (size_t)(n16 * 16), n16, n);
    hit_reset(best);
    RayCtx r;
    r.ox = ox; r.oy = oy; r.oz = oz; r.dx = dx; r.dy = dy; r.dz = dz; r.ix = r.iy = r.iz = 0.0f; r.octinv4 = 0u;
    float limit = FLT_MAX;
    test_prim(s.prim_kind, n, r, best, limit);   // accepted candidates have 0 <= t < min(max_t, FLT_MAX): always a hit here
}

#define VSB_STACK_SIZE 48

// Per-ray traversal state.  Stack entries are node GROUPS (base index, mask): hit bits in
// 24..31 (already permuted by the ray octant: highest set bit = nearest child) and the node's
// inner mask in 0..7.  The primitive group (base + one bit per primitive, 32 bits) is never stacked.
#ifndef VSB_SHARED_STACK
#define VSB_SHARED_STACK 16     // entries per thread kept in shared memory (device); deeper ones spill to local
#endif
#ifndef VSB_STEP_MODE
#define VSB_STEP_MODE 0     // measured: chaining costs a second exposed latency per step (cfg5 +29 % time)
#endif
#ifndef VSB_PREFETCH_PUSH
#define VSB_PREFETCH_PUSH 0   // 1 / 2: prefetch (L2 / L1) the next sibling of a node group when it is pushed
#endif

// The deep part of the stack lives in its own (local-memory) object: kept inside TravState, its
// dynamically indexed arrays dragged the whole struct -- node group, primitive group, stack pointer
// -- into local memory, one LDL/STL per field access in the traversal loop.
#ifndef VSB_Q_SHARED_STACK
#define VSB_Q_SHARED_STACK 12   // shared entries per thread in the queue kernel (trace_kernel_q): leaves room for its ray / key / queue arrays
#endif
#define VSB_MIN_SHARED_STACK (VSB_Q_SHARED_STACK < VSB_SHARED_STACK ? VSB_Q_SHARED_STACK : VSB_SHARED_STACK)

struct TravSpill
{
#if defined(__CUDACC__)
    uint32_t b[VSB_STACK_SIZE - VSB_MIN_SHARED_STACK], m[VSB_STACK_SIZE - VSB_MIN_SHARED_STACK];
#else
    uint32_t b[VSB_STACK_SIZE], m[VSB_STACK_SIZE];
#endif
};

struct TravState
{
    uint32_t ng_base, ng_mask;      // current node group
    uint32_t pg_base, pg_mask;      // current primitive group (mask in slot space, see test_node)
    uint32_t pg_valid;              // the group's node's `valid` mask: prim_record(pg_base, pg_valid, bit)
    int      sp;
#if defined(__CUDACC__)
    uint32_t sstack;                // shared-window byte address of this thread's column of the block's stack:
                                    // entry k at words [2k*128], [(2k+1)*128] (one 32-bit register, no generic pointer)
#endif
    TravSpill* spill;               // set by the owner of the state
};

template <int SS = VSB_SHARED_STACK>
VSB_HD void stack_push(TravState& t, uint32_t b, uint32_t m)
{
#if defined(__CUDACC__)
    if (t.sp < SS)
    {
        const uint32_t a = t.sstack + (uint32_t)t.sp * 1024u;
        asm volatile("st.shared.u32 [%0], %1;" :: "r"(a), "r"(b) : "memory");
        asm volatile("st.shared.u32 [%0+512], %1;" :: "r"(a), "r"(m) : "memory");
    }
    else { t.spill->b[t.sp - SS] = b; t.spill->m[t.sp - SS] = m; }
#else
    t.spill->b[t.sp] = b; t.spill->m[t.sp] = m;
#endif
    ++t.sp;
}

template <int SS = VSB_SHARED_STACK>
VSB_HD void stack_pop(TravState& t, uint32_t& b, uint32_t& m)
{
    --t.sp;
#if defined(__CUDACC__)
    if (t.sp < SS)
    {
        const uint32_t a = t.sstack + (uint32_t)t.sp * 1024u;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(b) : "r"(a) : "memory");
        asm volatile("ld.shared.u32 %0, [%1+512];" : "=r"(m) : "r"(a) : "memory");
    }
    else { b = t.spill->b[t.sp - SS]; m = t.spill->m[t.sp - SS]; }
#else
    b = t.spill->b[t.sp]; m = t.spill->m[t.sp];
#endif
}

VSB_HD void trav_begin(TravState& t)
{
    // the root is "inner child at bit 31" of a virtual group with child_base 0 and imask 0
    t.ng_base = 0u; t.ng_mask = 0x80000000u;
    t.pg_base = 0u; t.pg_mask = 0u; t.pg_valid = 0u;
    t.sp = 0;
}

// One traversal step of one ray: EITHER one primitive test (while the lane has primitives
// pending) OR one node visit.  A warp whose lanes are in different phases issues all record
// loads together (load_record) and then runs the primitive test and the node test once each,
// every lane doing useful work in one of them -- instead of a per-lane loop over a leaf's
// primitives that runs at the occupancy of its longest leaf (measured: 3.5 of 32 lanes).
// Invariant on entry: pg_mask != 0 or the node group has hit bits.  Returns false when the ray
// is finished (stack empty, or `any` and a hit was accepted: exit_traversal.h:49-57).
VSB_HD bool trav_prim_mode(const TravState& t) { return t.pg_mask != 0u; }

// (A "paired" mode -- a lane down to its last pending primitive takes its next node visit in the same
// step, both record loads issued together -- was measured at 0.6-0.7x: two live records per lane cost
// more registers / spills than the saved iterations give back.)
VSB_HD bool trav_step(const WideScene& s, const RayCtx& r, TravState& t, HitReg& best, float& limit, bool any)
{
    NodeRegs n;
#if VSB_STEP_MODE == 1
    // Mode 1 ("chained"): the primitive part runs first for the lanes with pending primitives; a lane
    // that tests its LAST pending primitive falls through into the node part of the same step, so it
    // does not spend a whole extra warp iteration before its next node visit.
    if (t.pg_mask)
    {
        const int n16 = wide_rec_words16(s.prim_kind);
#if defined(__CUDA_ARCH__)
        const int n_enter = __popc(__activemask());
#endif
        do
        {
            const int bit = find_msb(t.pg_mask);
            t.pg_mask &= ~(1u << bit);
            load_record(reinterpret_cast<const char*>(s.prims) + (size_t)prim_record(t.pg_base, t.pg_valid, bit) * (size_t)(n16 * 16), n16, n);
            test_prim(s.prim_kind, n, r, best, limit);
            if (any && best.hit) return false;
        }
#if defined(__CUDA_ARCH__)
        while (t.pg_mask && __popc(__activemask()) * 8 >= n_enter * 7);     // coherent leaves: tight loop
#else
        while (false);
#endif
        if (t.pg_mask) return true;
        if ((t.ng_mask & 0xff000000u) == 0u)
        {
            if (t.sp == 0) return false;
            stack_pop(t, t.ng_base, t.ng_mask);
        }
    }
    {
        const int bit = find_msb(t.ng_mask);
        t.ng_mask &= ~(1u << bit);
        const uint32_t imask = t.ng_mask & 0xffu;
        const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xffu)) & 7u;
        const uint32_t rel = (uint32_t)pop_count(imask & ~(0xffffffffu << slot));
        const char* addr = reinterpret_cast<const char*>(s.nodes + (t.ng_base + rel));
        if (t.ng_mask & 0xff000000u) stack_push(t, t.ng_base, t.ng_mask);
        load_record(addr, 5, n);
        uint32_t ph;
        const uint32_t nh = test_node(n, r, limit, ph);
        t.ng_base = n.w[4]; t.ng_mask = nh | (n.w[3] >> 24);
        t.pg_base = n.w[5]; t.pg_mask = ph; t.pg_valid = n.w[6];
        if (t.pg_mask) return true;
    }
#else
    // Mode 0 ("unified loads"): a lane is in primitive phase or node phase for the whole step; all
    // lanes issue their record loads together, then each test runs once.
    const bool prim_mode = t.pg_mask != 0u;
    const char* addr;
    int n16;
    if (prim_mode)
    {
        const int bit = find_msb(t.pg_mask);
        t.pg_mask &= ~(1u << bit);
        const uint32_t index = prim_record(t.pg_base, t.pg_valid, bit);
        n16 = wide_rec_words16(s.prim_kind);
        addr = reinterpret_cast<const char*>(s.prims) + (size_t)index * (size_t)(n16 * 16);
    }
    else
    {
        const int bit = find_msb(t.ng_mask);
        t.ng_mask &= ~(1u << bit);
        const uint32_t imask = t.ng_mask & 0xffu;
        const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xffu)) & 7u;
        const uint32_t rel = (uint32_t)pop_count(imask & ~(0xffffffffu << slot));
        n16 = 5;
        addr = reinterpret_cast<const char*>(s.nodes + (t.ng_base + rel));
        if (t.ng_mask & 0xff000000u)
        {
            stack_push(t, t.ng_base, t.ng_mask);
#if defined(__CUDA_ARCH__) && VSB_PREFETCH_PUSH
            // the sibling this ray visits after the subtree it is about to enter
            const int bit2 = find_msb(t.ng_mask);
            const uint32_t slot2 = ((uint32_t)(bit2 - 24) ^ (r.octinv4 & 0xffu)) & 7u;
            const char* a2 = reinterpret_cast<const char*>(s.nodes + (t.ng_base + (uint32_t)pop_count(imask & ~(0xffffffffu << slot2))));
#if VSB_PREFETCH_PUSH == 1
            asm volatile("prefetch.global.L2 [%0];" :: "l"(a2));
            asm volatile("prefetch.global.L2 [%0];" :: "l"(a2 + 64));
#else
            asm volatile("prefetch.global.L1 [%0];" :: "l"(a2));
            asm volatile("prefetch.global.L1 [%0];" :: "l"(a2 + 64));
#endif
#endif
        }
    }

    load_record(addr, n16, n);

    if (prim_mode)
    {
        test_prim(s.prim_kind, n, r, best, limit);
        if (any && best.hit) return false;
#if defined(__CUDA_ARCH__)
        // coherent leaves: while (nearly) all lanes that entered this primitive phase still have
        // primitives, keep testing in a tight loop instead of paying a full step per primitive
        const int n_enter = __popc(__activemask());
        while (t.pg_mask && __popc(__activemask()) * 8 >= n_enter * 7)
        {
            const int bit = find_msb(t.pg_mask);
            t.pg_mask &= ~(1u << bit);
            load_record(reinterpret_cast<const char*>(s.prims) + (size_t)prim_record(t.pg_base, t.pg_valid, bit) * (size_t)(n16 * 16), n16, n);
            test_prim(s.prim_kind, n, r, best, limit);
            if (any && best.hit) return false;
        }
#endif
        if (t.pg_mask) return true;
    }
    else
    {
        uint32_t ph;
        const uint32_t nh = test_node(n, r, limit, ph);
        t.ng_base = n.w[4]; t.ng_mask = nh | (n.w[3] >> 24);
        t.pg_base = n.w[5]; t.pg_mask = ph; t.pg_valid = n.w[6];
        if (t.pg_mask) return true;
    }
#endif

    // no primitives pending: the next step is a node visit; pop when this group is used up
    if ((t.ng_mask & 0xff000000u) == 0u)
    {
        if (t.sp == 0) return false;
        stack_pop(t, t.ng_base, t.ng_mask);
    }
    return true;
}

// Single-ray traversal (sequential form of the loop the kernels run warp-synchronously).
template <bool ANY>
VSB_HD void trace_wide(const WideScene& s, const RayCtx& r, float max_t, HitReg& best)
{
    hit_reset(best);
    float limit = max_t;
    TravSpill spill;
    TravState t;
    t.spill = &spill;
    trav_begin(t);
    while (trav_step(s, r, t, best, limit, ANY)) { }
}

} // namespace vsb

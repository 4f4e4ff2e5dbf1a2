// bvh8.cuh -- the wide quantised BVH8: node record, BVH2 -> BVH8 collapse of one node, the
// 8-child slab test on quantised boxes, and the single-ray traversal loop.
//
// The BVH2 that comes in is the reference's binned-SAH tree (bvh.h:52-119 node layout,
// detail/bvh/build.inl tree shape).  Each BVH8 node is built from a BVH2 subtree root by
// greedily opening the child with the largest surface area until 8 children exist; child
// boxes are stored as 8-bit offsets from the node origin on a power-of-two grid, rounded
// outward, so every quantised box CONTAINS the reference box it came from.  The traversal's
// slab test is additionally inflated by a bound on its own rounding error, so the set of
// (ray, box) pairs it accepts is a superset of what the reference's exact-order slab test
// (math/intersect.h:52-70 + update_if.h:62-88) accepts: no reference hit can be culled.
//
// Host/device source: the CUDA kernels include it; tests/hostsim compiles it for the CPU to
// check the collapse and the traversal logic in the no-GPU test tier.
#pragma once

#include "vsb_common.cuh"
#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

namespace vsb
{

// ---- small portability layer ---------------------------------------------------------------
VSB_HD uint32_t byte_perm(uint32_t a, uint32_t b, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, sel);
#else
    uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i)
    {
        uint32_t s = (sel >> (4 * i)) & 0x7;
        r |= (uint32_t)((v >> (8 * s)) & 0xff) << (8 * i);
    }
    return r;
#endif
}

// ((1 << width) - 1) << pos in one instruction (PTX bmsk, sm_70+)
VSB_HD uint32_t bit_mask(uint32_t pos, uint32_t width)
{
#if defined(__CUDA_ARCH__)
    uint32_t r;
    asm("bmsk.clamp.b32 %0, %1, %2;" : "=r"(r) : "r"(pos), "r"(width));
    return r;
#else
    return (width >= 32 ? 0xffffffffu : ((1u << width) - 1u)) << pos;
#endif
}

VSB_HD int find_msb(uint32_t v)   // index of the highest set bit, v != 0
{
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int)v);
#else
    return 31 - __builtin_clz(v);
#endif
}

VSB_HD int pop_count(uint32_t v)
{
#if defined(__CUDA_ARCH__)
    return __popc(v);
#else
    return __builtin_popcount(v);
#endif
}

VSB_HD float fma_f(float a, float b, float c)
{
#if defined(__CUDA_ARCH__)
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}

// byte k of `w` as the float 1 + b * 2^-15 (exact): one PRMT, no int->float conversion.  `one` is
// 0x3F800000 read from the constant bank (unit_float_bits()): PRMT has one immediate slot, and with the
// constant as the immediate ptxas put the four selectors into registers and re-materialised them with a
// MOV per use (~48 of the node test's ~280 instructions, found in the SASS).
#if defined(__CUDACC__)
__constant__ uint32_t c_unit_float_bits = 0x3F800000u;        // a constant-bank value: opaque to ptxas's constant propagation
__constant__ uint32_t c_half_bias_bits = 0x64646464u;
#endif
VSB_HD uint32_t unit_float_bits()
{
#if defined(__CUDA_ARCH__)
    return c_unit_float_bits;
#else
    return 0x3F800000u;
#endif
}
VSB_HD float byte_as_unit_float(uint32_t w, int k, uint32_t one)
{
    return u2f(byte_perm(w, one, 0x7604u | ((uint32_t)k << 4)));
}

// VSB_NODE_HALFCVT = 1: a quantised plane byte enters the slab test as the float 1024 + b instead: ONE PRMT
// builds the half2 (0x6400 | b_k, 0x6400 | b_k+1) = (1024 + b_k, 1024 + b_k+1) for two children, two HADD2.F32
// widen it.  Per byte: half a PRMT on the alu pipe (the saturated one: 80 % of its peak on the cfg2 frame,
// profiles/r2_cfg2_q1) and one conversion on the fma pipe (21 %), against one PRMT.
#ifndef VSB_NODE_HALFCVT
#define VSB_NODE_HALFCVT 1      // measured: cfg2 6306 -> 6394, cfg5/4 1219 -> 1245 Mrays/s (gpurun_out/r2_ab3.log)
#endif
VSB_HD uint32_t half_bias_bits()
{
#if defined(__CUDA_ARCH__)
    return c_half_bias_bits;
#else
    return 0x64646464u;
#endif
}
VSB_HD float byte_as_biased_float(uint32_t w, int k, uint32_t bias)
{
#if defined(__CUDA_ARCH__)
    const uint32_t h2 = byte_perm(w, bias, (k & 2) ? 0x4342u : 0x4140u);
    const __half2 h = *reinterpret_cast<const __half2*>(&h2);
    return (k & 1) ? __high2float(h) : __low2float(h);
#else
    (void)bias;
    return 1024.0f + (float)((w >> (8 * k)) & 0xffu);
#endif
}

// ---- collapse of one BVH2 subtree root into one BVH8 node ----------------------------------

struct WideTask
{
    uint32_t bvh2_node;    // BVH2 subtree root this wide node is built from
    uint32_t wide_index;   // where to write the Node80
};

VSB_HD float half_area(const Node32& n)
{
    float sx = n.max_x - n.min_x, sy = n.max_y - n.min_y, sz = n.max_z - n.min_z;
    return sx * sy + sy * sz + sz * sx;
}

// Biased exponent e such that 255 * 2^(e-127) >= extent (0 for a flat extent).
VSB_HD uint32_t grid_exponent(float extent)
{
    if (!(extent > 0.0f)) return 0u;
    // smallest power of two s with 255 * s >= extent; start from the float's own exponent
    float s = extent / 255.0f;
    uint32_t e = (f2u(s) >> 23) & 0xffu;
    // u2f(e << 23) <= s < 2 * that; bump until the grid really covers the extent
    if (e == 0u) e = 1u;
    while (255.0f * u2f(e << 23) < extent && e < 254u) ++e;
    return e;
}

// Outward-rounded quantisation of [lo, hi] relative to origin p on grid scale s.
// Canonical dequantisation (what "contains" is checked against): p + q * s in float.
VSB_HD void quantise_axis(float p, float s, float lo, float hi, uint32_t& qlo, uint32_t& qhi)
{
    if (!(s > 0.0f)) { qlo = 0u; qhi = 0u; return; }
    float fl = floorf((lo - p) / s);
    float fh = ceilf((hi - p) / s);
    int a = fl < 0.0f ? 0 : (fl > 255.0f ? 255 : (int)fl);
    int b = fh < 0.0f ? 0 : (fh > 255.0f ? 255 : (int)fh);
    // enforce containment of the REAL-valued grid planes (double: p + q*s is exact there)
    while (a > 0 && (double)p + (double)a * (double)s > (double)lo) --a;
    while (b < 255 && (double)p + (double)b * (double)s < (double)hi) ++b;
    qlo = (uint32_t)a; qhi = (uint32_t)b;
}

// ---- optimal collapse (dynamic programming over the BVH2, bottom-up) -----------------------------
//
// cost[n][i-1], i = 1..7: the smallest SAH cost of representing the subtree of BVH2 node n by at
// most i roots (each root a BVH2 leaf, kept as a leaf child, or an internal BVH8 node):
//   leaf n:   C(n, i) = A_n * P_n * c_prim
//   inner n:  C(n, 1) = A_n * c_node + D(n, 8)                 (n becomes a BVH8 node)
//             C(n, i) = min(D(n, i), C(n, i-1))
//             D(n, j) = min over k of C(left, k) + C(right, j - k)   (distribute j roots)
// (the formulation of Ylitie, Karras, Laine: "Efficient incoherent ray traversal on GPUs through
// compressed wide BVHs", HPG 2017, section 4.1, with the reference's BVH2 leaves kept whole).
// dec[n]: bytes 0..6 = argmin k of D(n, j) for j = 2..8; bytes 8..13 = the effective root count
// behind C(n, i) for i = 2..7 (1 = "n is one root").  A greedy largest-area-first collapse
// leaves most bottom nodes half empty (measured: 3.95 children per node on the 20 M soup).
struct CollapseDec { uint8_t b[16]; };
struct CollapseCost { float c[8]; };      // c[7] unused (32-byte records)

#define VSB_COST_NODE 1.0f
#define VSB_COST_PRIM 0.3f

VSB_HD void dp_leaf(const Node32& n, CollapseCost& out, CollapseDec& dec)
{
    const float c = half_area(n) * (float)n.num_prims * VSB_COST_PRIM;
    for (int i = 0; i < 8; ++i) out.c[i] = c;
    for (int i = 0; i < 16; ++i) dec.b[i] = 1;
}

VSB_HD void dp_inner(const Node32& n, const CollapseCost& l, const CollapseCost& r, CollapseCost& out, CollapseDec& dec)
{
    float d[9];
    for (int j = 2; j <= 8; ++j)
    {
        float best = FLT_MAX; int bk = 1;
        for (int k = 1; k < j; ++k)
        {
            if (k > 7 || j - k > 7) continue;
            const float c = l.c[k - 1] + r.c[j - k - 1];
            if (c < best) { best = c; bk = k; }
        }
        d[j] = best; dec.b[j - 2] = (uint8_t)bk;
    }
    out.c[0] = half_area(n) * VSB_COST_NODE + d[8];
    uint8_t eff = 1;
    for (int i = 2; i <= 7; ++i)
    {
        if (d[i] < out.c[i - 2]) { out.c[i - 1] = d[i]; eff = (uint8_t)i; }
        else out.c[i - 1] = out.c[i - 2];
        dec.b[8 + i - 2] = eff;
    }
    out.c[7] = 0.0f;
    dec.b[7] = 0; dec.b[14] = 0; dec.b[15] = 0;
}

// children of the BVH8 node rooted at BVH2 node `root` according to the DP decisions
VSB_HD int expand_optimal(uint32_t root, const Node32* __restrict__ nodes2, const CollapseDec* __restrict__ dec, uint32_t* child)
{
    uint32_t st_node[16]; uint8_t st_budget[16];
    int sp = 0, n = 0;
    const Node32 r = nodes2[root];
    if (r.num_prims != 0) { child[0] = root; return 1; }
    // the top is forced to distribute 8 roots
    {
        const int k = dec[root].b[6];
        st_node[sp] = r.first + 1; st_budget[sp] = (uint8_t)(8 - k); ++sp;
        st_node[sp] = r.first;     st_budget[sp] = (uint8_t)k;       ++sp;
    }
    while (sp > 0)
    {
        --sp;
        const uint32_t m = st_node[sp]; const int j = st_budget[sp];
        const Node32 c = nodes2[m];
        int eff = 1;
        if (c.num_prims == 0 && j >= 2) eff = dec[m].b[8 + j - 2];
        if (c.num_prims != 0 || eff <= 1) { child[n++] = m; continue; }
        const int k = dec[m].b[eff - 2];
        st_node[sp] = c.first + 1; st_budget[sp] = (uint8_t)(eff - k); ++sp;
        st_node[sp] = c.first;     st_budget[sp] = (uint8_t)k;         ++sp;
    }
    return n;
}

// Collapses BVH2 subtree `task.bvh2_node` into out_nodes[task.wide_index]; appends the leaf
// primitives' indices (positions in indices[]) of this node to leaf_list at *prim_alloc, and the
// inner children as new tasks.  alloc_* are atomic fetch-add functors (device: atomicAdd).
// dec != NULL: optimal collapse (DP decisions); NULL: greedy largest-area-first.
// Returns the number of inner children.
template <typename AllocU32>
VSB_HD int collapse_node(const WideTask task, const Node32* __restrict__ nodes2, Node80* out_nodes,
                         uint32_t* leaf_list, WideTask* next_tasks,
                         AllocU32& alloc_node, AllocU32& alloc_prim, AllocU32& alloc_task,
                         const CollapseDec* __restrict__ dec = nullptr)
{
    uint32_t child[VSB_WIDE];
    int n = 0;
    uint32_t leaf_prims = 0;
    if (dec)
    {
        n = expand_optimal(task.bvh2_node, nodes2, dec, child);
        for (int i = 0; i < n; ++i) leaf_prims += nodes2[child[i]].num_prims;
    }
    else
    {
        const Node32 root = nodes2[task.bvh2_node];
        if (root.num_prims != 0) { child[n++] = task.bvh2_node; }          // single-leaf tree
        else { child[n++] = root.first; child[n++] = root.first + 1; }
        for (int i = 0; i < n; ++i) leaf_prims += nodes2[child[i]].num_prims;
        // greedily open the inner child with the largest surface area
        while (n < VSB_WIDE)
        {
            int best = -1; float best_area = -1.0f;
            for (int i = 0; i < n; ++i)
            {
                const Node32& c = nodes2[child[i]];
                if (c.num_prims != 0) continue;
                float a = half_area(c);
                if (a > best_area) { best_area = a; best = i; }
            }
            if (best < 0) break;
            const Node32 c = nodes2[child[best]];
            leaf_prims += nodes2[c.first].num_prims + nodes2[c.first + 1].num_prims;
            child[best] = c.first;
            child[n++] = c.first + 1;
        }
    }

    // node bounds = union of children
    float lox = FLT_MAX, loy = FLT_MAX, loz = FLT_MAX, hix = -FLT_MAX, hiy = -FLT_MAX, hiz = -FLT_MAX;
    for (int i = 0; i < n; ++i)
    {
        const Node32& c = nodes2[child[i]];
        lox = fminf(lox, c.min_x); loy = fminf(loy, c.min_y); loz = fminf(loz, c.min_z);
        hix = fmaxf(hix, c.max_x); hiy = fmaxf(hiy, c.max_y); hiz = fmaxf(hiz, c.max_z);
    }
    const uint32_t ex = grid_exponent(hix - lox), ey = grid_exponent(hiy - loy), ez = grid_exponent(hiz - loz);
    const float sx = u2f(ex << 23), sy = u2f(ey << 23), sz = u2f(ez << 23);

    // slot assignment: slot s is "the child furthest toward direction D_s",
    // D_s = (s&1 ? + : -, s&2 ? + : -, s&4 ? + : -); greedy max over the 8x8 cost matrix.
    // The traversal visits slots in decreasing (s ^ (7 - octant)), i.e. front to back.
    const float cx = 0.5f * (lox + hix), cy = 0.5f * (loy + hiy), cz = 0.5f * (loz + hiz);
    int slot_of[VSB_WIDE];
    int child_in_slot[VSB_WIDE];
    for (int i = 0; i < VSB_WIDE; ++i) { slot_of[i] = -1; child_in_slot[i] = -1; }
    for (int round = 0; round < n; ++round)
    {
        float best_cost = -FLT_MAX; int bc = -1, bs = -1;
        for (int i = 0; i < n; ++i)
        {
            if (slot_of[i] >= 0) continue;
            const Node32& c = nodes2[child[i]];
            float px = 0.5f * (c.min_x + c.max_x) - cx, py = 0.5f * (c.min_y + c.max_y) - cy, pz = 0.5f * (c.min_z + c.max_z) - cz;
            for (int s = 0; s < VSB_WIDE; ++s)
            {
                if (child_in_slot[s] >= 0) continue;
                float cost = ((s & 1) ? px : -px) + ((s & 2) ? py : -py) + ((s & 4) ? pz : -pz);
                if (cost > best_cost) { best_cost = cost; bc = i; bs = s; }
            }
        }
        slot_of[bc] = bs; child_in_slot[bs] = bc;
    }

    // allocate children / primitive ranges
    int n_inner = 0;
    for (int i = 0; i < n; ++i) n_inner += nodes2[child[i]].num_prims == 0 ? 1 : 0;
    const uint32_t child_base = n_inner ? alloc_node((uint32_t)n_inner) : 0u;
    const uint32_t prim_base = leaf_prims ? alloc_prim(leaf_prims) : 0u;
    const uint32_t task_base = n_inner ? alloc_task((uint32_t)n_inner) : 0u;

    Node80 out;
    for (int i = 0; i < 20; ++i) out.w[i] = 0u;
    out.w[0] = f2u(lox); out.w[1] = f2u(loy); out.w[2] = f2u(loz);
    uint32_t imask = 0, inner_rank = 0, prim_ofs = 0, valid = 0;
    uint8_t q[6][8];
    for (int s = 0; s < VSB_WIDE; ++s)
    {
        for (int k = 0; k < 6; ++k) q[k][s] = 0;
        // an empty slot keeps lo = 0 / hi = 0 and an empty nibble in `valid`: its bits never enter the hit mask
        const int ci = child_in_slot[s];
        if (ci < 0) continue;
        const Node32& c = nodes2[child[ci]];
        uint32_t a, b;
        quantise_axis(lox, sx, c.min_x, c.max_x, a, b); q[0][s] = (uint8_t)a; q[3][s] = (uint8_t)b;
        quantise_axis(loy, sy, c.min_y, c.max_y, a, b); q[1][s] = (uint8_t)a; q[4][s] = (uint8_t)b;
        quantise_axis(loz, sz, c.min_z, c.max_z, a, b); q[2][s] = (uint8_t)a; q[5][s] = (uint8_t)b;
        if (c.num_prims == 0)
        {
            imask |= 1u << s;          // inner children are identified by imask (their nibble of `valid` stays 0)
            next_tasks[task_base + inner_rank].bvh2_node = child[ci];
            next_tasks[task_base + inner_rank].wide_index = child_base + inner_rank;
            ++inner_rank;
        }
        else
        {
            valid |= ((1u << c.num_prims) - 1u) << (4 * s);      // records in slot order: index = prim_base + rank of the bit in `valid`
            for (uint32_t k = 0; k < c.num_prims; ++k) leaf_list[prim_base + prim_ofs + k] = c.first + k;
            prim_ofs += c.num_prims;
        }
    }
    out.w[3] = ex | (ey << 8) | (ez << 16) | (imask << 24);
    out.w[4] = child_base;
    out.w[5] = prim_base;
    out.w[6] = valid;
    out.w[7] = 0u;
    // w[8..9] qlo_x, w[10..11] qlo_y, w[12..13] qlo_z, w[14..15] qhi_x, w[16..17] qhi_y, w[18..19] qhi_z
    for (int k = 0; k < 6; ++k)
    {
        out.w[8 + 2 * k]     = (uint32_t)q[k][0] | ((uint32_t)q[k][1] << 8) | ((uint32_t)q[k][2] << 16) | ((uint32_t)q[k][3] << 24);
        out.w[8 + 2 * k + 1] = (uint32_t)q[k][4] | ((uint32_t)q[k][5] << 8) | ((uint32_t)q[k][6] << 16) | ((uint32_t)q[k][7] << 24);
    }
    out_nodes[task.wide_index] = out;
    return n_inner;
}

// Leaf-order primitive record for list position `li` (position in indices[], which is the
// reference's primitive_list_index, detail/bvh/hit_record.h:44-46).
VSB_HD void make_prim_record(int prim_kind, const void* prims, const uint32_t* indices, uint32_t li, void* out, uint32_t slot)
{
    const uint32_t p = indices ? indices[li] : li;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(prims) + (size_t)p * prim_stride(prim_kind));
    if (prim_kind == 2 && !prim_is_triangle(prim_kind, src))
    {
        // a sphere among 48-byte records: sphere layout in words 0..7, tag in word 11
        uint32_t* o = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(out) + (size_t)slot * 48u);
        o[0] = src[4]; o[1] = src[5]; o[2] = src[6]; o[3] = src[8];
        o[4] = src[1]; o[5] = src[0]; o[6] = li; o[7] = 0u;
        o[8] = 0u; o[9] = 0u; o[10] = 0u; o[11] = VSB_GENERIC_SPHERE_TAG;
        return;
    }
    if (prim_kind != 1)
    {
        TriRec r;
        r.v1x = u2f(src[4]);  r.v1y = u2f(src[5]);  r.v1z = u2f(src[6]);  r.prim_id = src[1];
        r.e1x = u2f(src[8]);  r.e1y = u2f(src[9]);  r.e1z = u2f(src[10]); r.geom_id = src[0];
        r.e2x = u2f(src[12]); r.e2y = u2f(src[13]); r.e2z = u2f(src[14]); r.list_index = li;
        reinterpret_cast<TriRec*>(out)[slot] = r;
    }
    else
    {
        SphRec r;
        r.cx = u2f(src[4]); r.cy = u2f(src[5]); r.cz = u2f(src[6]); r.radius = u2f(src[8]);
        r.prim_id = src[1]; r.geom_id = src[0]; r.list_index = li; r.pad = 0u;
        reinterpret_cast<SphRec*>(out)[slot] = r;
    }
}

} // namespace vsb

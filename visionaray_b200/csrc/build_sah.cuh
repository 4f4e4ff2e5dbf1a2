// build_sah.cuh -- the REFERENCE's binned-SAH tree, built on the GPU (second builder).
//
// build<index_bvh<P>>(first, last) of the reference (bvh.h:455 -> detail/bvh/build.inl:28-140 -> binned_sah_builder,
// detail/bvh/sah.h:150-762) is a recursion on one host thread: bin the centroids of a node's references into 16 bins
// along the widest axis of the centroid bounds, sweep the 15 bin boundaries for the cheapest split, std::partition the
// references, recurse (right subtree first).  Everything that DECIDES that tree is order-independent: bounds are
// minima / maxima, counts are integers, the cost of a candidate is a fixed expression of those.  So the same tree can be
// built breadth-first, all nodes of a level at once, and come out identical -- same nodes in the same places, same
// bounds, same indices[] -- as long as three things are reproduced exactly:
//
//   * the arithmetic (single-rounding rn_* operations in the reference's order, its float -> int conversions),
//   * the order std::partition leaves the references in (libstdc++'s bidirectional version is Hoare's scheme: the k-th
//     misplaced element from the left swaps with the k-th misplaced element from the right; that is a closed form over
//     a prefix sum of the predicate, so it parallelises without changing the result),
//   * the numbering: the recursion allocates a node's two children when the node is popped, right subtree first, so the
//     k-th inner node in right-first pre-order has its children at nodes[1 + 2k], nodes[2 + 2k], and a leaf covering
//     reference positions [first, end) of n is emitted after everything to its right: indices[n - end ...].
//     Right-first pre-order of nodes over position ranges is the order of the key (n - end, first); one radix sort.
//
// One caveat: where +0.0 and -0.0 both occur among the coordinates that meet in a minimum / maximum, the reference keeps
// the one it saw last and this builder keeps -0.0 in minima and +0.0 in maxima.  No decision depends on the sign of a
// zero; tests compare bounds as floats (-0 == +0) and everything else bit for bit.  NaN coordinates are not supported.
//
// Per level (all of it asynchronous; the host reads the number of open nodes back every 4 levels):
//   sah_bin_kernel        a block bins 128 consecutive reference positions into shared-memory bins, one set per open
//                         node ("segment") that touches its chunk; a segment that lies inside the chunk gets its split
//                         found right there by a 16-lane group; one that spans chunks is flushed to a global slot
//   sah_span_split_kernel the split of the spanning segments from their global slots
//   cub::DeviceScan over the predicate (SahFlag, evaluated inside the scan) + sah_mislist_kernel + sah_swap_kernel   the partition
//   sah_child_count_kernel + cub::DeviceScan + sah_emit_children_kernel   next level's segments (positions move to
//                         them in the next level's sah_bin_kernel)
// Spatial splits (sah.h:428-615) stay on the host builder.
#pragma once

#include "build_ploc.cuh"

namespace vsb
{

#define VSB_SAH_CHUNK 128
#define VSB_SAH_BINS 16
#define VSB_SAH_FIELDS 13          // prim min xyz, prim max xyz, centroid min xyz, centroid max xyz, count
#define VSB_SAH_MAX_LOCAL 28       // open segments hold >= 5 references: at most 25 inside a chunk of 128 + 2 partial ones
#define VSB_SAH_NONE 0xffffffffu
#define VSB_SAH_MAX_LEAF 4         // build<>() passes max_leaf_size = 4 (detail/bvh/build.inl:165-178)
#define VSB_SAH_SLOT_WORDS (VSB_SAH_FIELDS * VSB_SAH_BINS)

struct SahRef   { float lox, loy, loz; uint32_t index; float hix, hiy, hiz; uint32_t pad; };            // prim_ref, sah.h
struct SahSeg   { float pb[6]; float cb[6]; uint32_t first, end, node, pad; };                           // leaf_info + range
struct SahSplit { float lpb[6], lcb[6], rpb[6], rcb[6]; float k0, k1; int axis, index; uint32_t do_split, pad[3]; };
struct SahNode  { float pb[6]; uint32_t first, end, left, parent; };                                    // left == NONE: leaf
struct SahChild { uint32_t lseg, rseg, pivot, pad; };

// order-preserving float <-> uint32 (for atomicMin / atomicMax): -0.0 sorts below +0.0
VSB_HD uint32_t f2o(float f) { const uint32_t b = f2u(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
VSB_HD float o2f(uint32_t o) { return u2f((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o); }

#define VSB_SAH_FLT_MAX 3.4028234663852886e38f

VSB_HD uint32_t sah_field_init(int field)       // bin::clear(): bounds invalidated, counters 0
{
    if (field == 12) return 0u;
    return ((field % 6) < 3) ? f2o(VSB_SAH_FLT_MAX) : f2o(-VSB_SAH_FLT_MAX);
}

// aabb::safe_size + safe_half_surface_area (math/aabb.inl)
VSB_HD float sah_half_area(const float* b)
{
    float sx = rn_sub(b[3], b[0]), sy = rn_sub(b[4], b[1]), sz = rn_sub(b[5], b[2]);
    sx = 0.0f < sx ? sx : 0.0f; sy = 0.0f < sy ? sy : 0.0f; sz = 0.0f < sz ? sz : 0.0f;
    return rn_add(rn_add(rn_mul(sx, sy), rn_mul(sy, sz)), rn_mul(sz, sx));
}

// the widest axis of the centroid bounds and the projection onto the bins (sah.h:230-275, max_index of
// math/detail/vector.inl:811-821); false: no extent, the node stays a leaf
VSB_HD bool sah_projection(const float* cb, float& k0, float& k1, int& axis)
{
    const float s0 = rn_sub(cb[3], cb[0]), s1 = rn_sub(cb[4], cb[1]), s2 = rn_sub(cb[5], cb[2]);
    const int a = s1 < s0 ? 0 : 1;
    const float sa = a == 0 ? s0 : s1;
    axis = s2 < sa ? a : 2;
    const float ext = axis == 0 ? s0 : (axis == 1 ? s1 : s2);
    if (ext <= 0.0f) return false;
    k0 = cb[axis];
    k1 = rn_div((float)VSB_SAH_BINS, rn_sub(cb[3 + axis], k0));
    return true;
}

// static_cast<int>(k1 * (c - k0)) with the x86 result for values an int cannot hold (cvttss2si: INT_MIN)
VSB_HD int sah_project_unsafe(float k0, float k1, float c)
{
    const float x = rn_mul(k1, rn_sub(c, k0));
    if (!(x > -2147483648.0f && x < 2147483648.0f)) return (int)0x80000000;
    return (int)x;
}

VSB_HD int sah_project(float k0, float k1, float c)
{
    const int i = sah_project_unsafe(k0, k1, c);
    return i < 0 ? 0 : (i > VSB_SAH_BINS - 1 ? VSB_SAH_BINS - 1 : i);
}

VSB_HD float sah_center(float lo, float hi) { return rn_mul(rn_add(hi, lo), 0.5f); }     // aabb::center(): (max + min) * 0.5

// is a child with `size` references and centroid bounds cb split further?  (sah.h:677-690: size <= max_leaf_size or no
// centroid extent -> leaf)
VSB_HD bool sah_child_open(uint32_t size, const float* cb)
{
    if (size <= (uint32_t)VSB_SAH_MAX_LEAF) return false;
    float k0, k1; int axis;
    return sah_projection(cb, k0, k1, axis);
}

#if defined(__CUDACC__)

// ---------------------------------------------------------------------------------------------------------------------
// set-up: references, root bounds

// root[0..11] = prim min / max, centroid min / max as ordered uints (initialised by the host to the invalid box)
__global__ void __launch_bounds__(256) sah_init_kernel(int prim_kind, const void* __restrict__ prims, uint32_t n, SahRef* __restrict__ refs,
                                                       uint32_t* __restrict__ seg_of, uint32_t* root)
{
    __shared__ uint32_t acc[12];
    if (threadIdx.x < 12) acc[threadIdx.x] = sah_field_init((int)threadIdx.x);
    __syncthreads();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n)
    {
        const BNode b = prim_bounds(prim_kind, prims, i);
        SahRef r;
        r.lox = b.lox; r.loy = b.loy; r.loz = b.loz; r.index = i; r.hix = b.hix; r.hiy = b.hiy; r.hiz = b.hiz; r.pad = 0u;
        reinterpret_cast<float4*>(refs)[2 * (size_t)i] = make_float4(r.lox, r.loy, r.loz, __uint_as_float(i));
        reinterpret_cast<float4*>(refs)[2 * (size_t)i + 1] = make_float4(r.hix, r.hiy, r.hiz, 0.0f);
        seg_of[i] = 0u;
        const float c[3] = { sah_center(r.lox, r.hix), sah_center(r.loy, r.hiy), sah_center(r.loz, r.hiz) };
        const float lo[3] = { r.lox, r.loy, r.loz }, hi[3] = { r.hix, r.hiy, r.hiz };
        for (int k = 0; k < 3; ++k)
        {
            atomicMin(&acc[k], f2o(lo[k]));     atomicMax(&acc[3 + k], f2o(hi[k]));
            atomicMin(&acc[6 + k], f2o(c[k]));  atomicMax(&acc[9 + k], f2o(c[k]));
        }
    }
    __syncthreads();
    if (threadIdx.x < 12)
    {
        if ((threadIdx.x % 6) < 3) atomicMin(&root[threadIdx.x], acc[threadIdx.x]);
        else                       atomicMax(&root[threadIdx.x], acc[threadIdx.x]);
    }
}

// state[2 L] = open segments of level L, state[2 L + 1] = build nodes so far
__global__ void sah_root_kernel(const uint32_t* __restrict__ root, uint32_t n, SahSeg* segs, SahNode* nodes, uint32_t* state)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    SahSeg s;
    for (int k = 0; k < 6; ++k) { s.pb[k] = o2f(root[k]); s.cb[k] = o2f(root[6 + k]); }
    s.first = 0u; s.end = n; s.node = 0u; s.pad = 0u;
    segs[0] = s;
    SahNode nd;
    for (int k = 0; k < 6; ++k) nd.pb[k] = s.pb[k];
    nd.first = 0u; nd.end = n; nd.left = VSB_SAH_NONE; nd.parent = VSB_SAH_NONE;
    nodes[0] = nd;
    state[0] = sah_child_open(n, s.cb) ? 1u : 0u;
    state[1] = 1u;
}

__global__ void sah_fill_kernel(uint32_t* p, size_t n_words, int slot_layout)
{
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_words) return;
    p[i] = slot_layout ? sah_field_init((int)((i % VSB_SAH_SLOT_WORDS) / VSB_SAH_BINS)) : sah_field_init((int)(i % 12));
}

// ---------------------------------------------------------------------------------------------------------------------
// find_split (sah.h:305-368) by a group of 16 lanes, lane i = bin i = candidate boundary i.  bins[field * 16 + bin] holds
// ordered uints (shared or global memory).  EVERY lane of the warp calls this (shuffles); lanes of a group without a
// segment pass valid = false.  The sweep of the reference runs i = 15 .. 1 and keeps a candidate only if it is strictly
// cheaper, so among equal costs the LARGEST i wins; a cost that is not < FLT_MAX (NaN, inf) is never taken.
__device__ __forceinline__ uint32_t sah_merge(int f, uint32_t a, uint32_t b) { return f < 3 ? min(a, b) : (f < 6 ? max(a, b) : a + b); }

__device__ void sah_find_split(const uint32_t* bins, bool valid, const SahSeg* sg, SahSplit* out)
{
    const unsigned full = 0xffffffffu;
    const int l16 = (int)(threadIdx.x & 15u);
    uint32_t pre[7], suf[7];
#pragma unroll
    for (int f = 0; f < 7; ++f)
    {
        const int field = f < 6 ? f : 12;
        pre[f] = suf[f] = valid ? bins[field * VSB_SAH_BINS + l16] : sah_field_init(field);
    }
#pragma unroll
    for (int d = 1; d < 16; d <<= 1)
    {
#pragma unroll
        for (int f = 0; f < 7; ++f)
        {
            const uint32_t v = __shfl_up_sync(full, pre[f], d, 16);
            if (l16 >= d) pre[f] = sah_merge(f, pre[f], v);
            const uint32_t w = __shfl_down_sync(full, suf[f], d, 16);
            if (l16 + d < 16) suf[f] = sah_merge(f, suf[f], w);
        }
    }
    // candidate i: L = bins 0 .. i-1 (inclusive prefix of lane i-1), R = bins i .. 15 (inclusive suffix of lane i)
    float Lb[6], Rb[6];
    uint32_t Lc, Rc = suf[6];
#pragma unroll
    for (int f = 0; f < 6; ++f)
    {
        Lb[f] = o2f(__shfl_up_sync(full, pre[f], 1, 16));
        Rb[f] = o2f(suf[f]);
    }
    Lc = __shfl_up_sync(full, pre[6], 1, 16);
    float cost = VSB_SAH_FLT_MAX;
    int idx = -1;
    if (valid && l16 >= 1)
    {
        const float hsa = sah_half_area(sg->pb);
        const float hl = sah_half_area(Lb), hr = sah_half_area(Rb);
        const float c = rn_add(rn_add(1.0f, rn_mul(rn_div(hl, hsa), rn_mul(3.0f, (float)(int)Lc))),
                               rn_mul(rn_div(hr, hsa), rn_mul(3.0f, (float)(int)Rc)));
        if (c < VSB_SAH_FLT_MAX) { cost = c; idx = l16; }
    }
#pragma unroll
    for (int d = 8; d >= 1; d >>= 1)
    {
        const float oc = __shfl_xor_sync(full, cost, d, 16);
        const int oi = __shfl_xor_sync(full, idx, d, 16);
        if (oc < cost || (oc == cost && oi > idx)) { cost = oc; idx = oi; }
    }
    if (!valid) return;
    if (idx < 0)
    {
        if (l16 == 0) out->do_split = 0u;       // degenerate costs: the node stays a leaf
        return;
    }
    if (l16 != idx) return;
    SahSplit sp;
    uint32_t lc[6], rc[6];
#pragma unroll
    for (int f = 0; f < 6; ++f) { sp.lpb[f] = Lb[f]; sp.rpb[f] = Rb[f]; lc[f] = rc[f] = sah_field_init(6 + f); }
    for (int b = 0; b < VSB_SAH_BINS; ++b)
#pragma unroll
        for (int f = 0; f < 6; ++f)
        {
            const uint32_t v = bins[(6 + f) * VSB_SAH_BINS + b];
            if (b < idx) lc[f] = sah_merge(f, lc[f], v); else rc[f] = sah_merge(f, rc[f], v);
        }
#pragma unroll
    for (int f = 0; f < 6; ++f) { sp.lcb[f] = o2f(lc[f]); sp.rcb[f] = o2f(rc[f]); }
    sah_projection(sg->cb, sp.k0, sp.k1, sp.axis);
    sp.index = idx;
    const uint32_t size = sg->end - sg->first;
    sp.do_split = (cost > rn_mul(3.0f, (float)(int)size)) ? 0u : 1u;       // a leaf is cheaper (sah.h:738-745)
    sp.pad[0] = sp.pad[1] = sp.pad[2] = 0u;
    *out = sp;
}

// slots: one set of bins per chunk, for the segment that STARTS in that chunk and reaches beyond it (at most one);
// span_seg[chunk] names it.  Slots are in the cleared state between levels (sah_span_split_kernel clears what it used).
// prev_child != nullptr: seg_of[] still holds the segments of the level above; every position moves to the child
// segment it landed in (or to NONE when its segment stayed a leaf) first -- the reassignment pass of the level above,
// fused into this level's first read of seg_of[].
__global__ void __launch_bounds__(VSB_SAH_CHUNK) sah_bin_kernel(const SahRef* __restrict__ refs, uint32_t* __restrict__ seg_of, uint32_t n,
                                                                const SahSeg* __restrict__ segs, SahSplit* splits, uint32_t* slots,
                                                                uint32_t* __restrict__ span_seg, const SahChild* __restrict__ prev_child)
{
    __shared__ uint32_t sbins[VSB_SAH_MAX_LOCAL * VSB_SAH_SLOT_WORDS];
    __shared__ uint32_t s_lo, s_hi;
    const uint32_t chunk_start = blockIdx.x * VSB_SAH_CHUNK;
    const uint32_t chunk_end = chunk_start + VSB_SAH_CHUNK < n ? chunk_start + VSB_SAH_CHUNK : n;
    const uint32_t p = chunk_start + threadIdx.x;
    uint32_t seg = p < n ? seg_of[p] : VSB_SAH_NONE;
    if (prev_child && seg != VSB_SAH_NONE)
    {
        const SahChild c = prev_child[seg];
        seg = p < c.pivot ? c.lseg : c.rseg;          // a segment that stayed a leaf: lseg = rseg = NONE
        seg_of[p] = seg;
    }
    if (threadIdx.x == 0) { s_lo = VSB_SAH_NONE; s_hi = 0u; }
    __syncthreads();
    if (seg != VSB_SAH_NONE) { atomicMin(&s_lo, seg); atomicMax(&s_hi, seg); }
    __syncthreads();
    const uint32_t seg_base = s_lo;
    if (seg_base == VSB_SAH_NONE)
    {
        if (threadIdx.x == 0) span_seg[blockIdx.x] = VSB_SAH_NONE;
        return;
    }
    const uint32_t n_local = s_hi - seg_base + 1u;
    if (n_local > (uint32_t)VSB_SAH_MAX_LOCAL) __trap();       // cannot happen: open segments hold >= 5 references
    for (uint32_t w = threadIdx.x; w < n_local * VSB_SAH_SLOT_WORDS; w += VSB_SAH_CHUNK)
        sbins[w] = sah_field_init((int)((w % VSB_SAH_SLOT_WORDS) / VSB_SAH_BINS));
    __syncthreads();
    if (seg != VSB_SAH_NONE)
    {
        const SahSeg* sg = segs + seg;
        float cb[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) cb[k] = sg->cb[k];
        float k0 = 0.0f, k1 = 0.0f; int axis = 0;
        sah_projection(cb, k0, k1, axis);
        const float4 a = reinterpret_cast<const float4*>(refs)[2 * (size_t)p], b = reinterpret_cast<const float4*>(refs)[2 * (size_t)p + 1];
        const float lo[3] = { a.x, a.y, a.z }, hi[3] = { b.x, b.y, b.z };
        const float c[3] = { sah_center(lo[0], hi[0]), sah_center(lo[1], hi[1]), sah_center(lo[2], hi[2]) };
        const int bin = sah_project(k0, k1, axis == 0 ? c[0] : (axis == 1 ? c[1] : c[2]));
        uint32_t* sb = sbins + (seg - seg_base) * VSB_SAH_SLOT_WORDS + bin;
        // read before the atomic: once a bin holds a few references most of them extend none of its bounds
        volatile uint32_t* vb = sb;
#pragma unroll
        for (int k = 0; k < 3; ++k)
        {
            const uint32_t l = f2o(lo[k]), h = f2o(hi[k]), m = f2o(c[k]);
            if (l < vb[k * VSB_SAH_BINS]) atomicMin(sb + k * VSB_SAH_BINS, l);
            if (h > vb[(3 + k) * VSB_SAH_BINS]) atomicMax(sb + (3 + k) * VSB_SAH_BINS, h);
            if (m < vb[(6 + k) * VSB_SAH_BINS]) atomicMin(sb + (6 + k) * VSB_SAH_BINS, m);
            if (m > vb[(9 + k) * VSB_SAH_BINS]) atomicMax(sb + (9 + k) * VSB_SAH_BINS, m);
        }
        atomicAdd(sb + 12 * VSB_SAH_BINS, 1u);
    }
    __syncthreads();
    const uint32_t group = threadIdx.x >> 4, l16 = threadIdx.x & 15u;
    for (uint32_t base = 0; base < n_local; base += VSB_SAH_CHUNK / 16)
    {
        const uint32_t l = base + group;
        const bool valid = l < n_local;
        const SahSeg* sg = segs + seg_base + (valid ? l : 0u);
        const uint32_t first = sg->first, end = sg->end;
        const bool contained = valid && first >= chunk_start && end <= chunk_end;
        const uint32_t* mine = sbins + (valid ? l : 0u) * VSB_SAH_SLOT_WORDS;
        sah_find_split(mine, contained, sg, splits + seg_base + (valid ? l : 0u));
        if (valid && !contained)
        {
            uint32_t* slot = slots + (size_t)(first / VSB_SAH_CHUNK) * VSB_SAH_SLOT_WORDS;
#pragma unroll
            for (int f = 0; f < VSB_SAH_FIELDS; ++f)
            {
                const uint32_t v = mine[f * VSB_SAH_BINS + l16];
                if (v == sah_field_init(f)) continue;
                if (f == 12) atomicAdd(slot + f * VSB_SAH_BINS + l16, v);
                else if ((f % 6) < 3) atomicMin(slot + f * VSB_SAH_BINS + l16, v);
                else atomicMax(slot + f * VSB_SAH_BINS + l16, v);
            }
        }
    }
    if (threadIdx.x == 0)
    {
        const SahSeg* last = segs + seg_base + n_local - 1u;
        span_seg[blockIdx.x] = (last->first >= chunk_start && last->end > chunk_end) ? seg_base + n_local - 1u : VSB_SAH_NONE;
    }
}

__global__ void __launch_bounds__(128) sah_span_split_kernel(uint32_t n_chunks, const SahSeg* __restrict__ segs, SahSplit* splits, uint32_t* slots,
                                                             const uint32_t* __restrict__ span_seg)
{
    const uint32_t c = (blockIdx.x * 128u + threadIdx.x) >> 4, l16 = threadIdx.x & 15u;
    const uint32_t s = c < n_chunks ? span_seg[c] : VSB_SAH_NONE;
    const bool valid = s != VSB_SAH_NONE;
    uint32_t* slot = slots + (size_t)(valid ? c : 0u) * VSB_SAH_SLOT_WORDS;
    sah_find_split(slot, valid, segs + (valid ? s : 0u), splits + (valid ? s : 0u));
    __syncwarp();
    if (valid)
#pragma unroll
        for (int f = 0; f < VSB_SAH_FIELDS; ++f) slot[f * VSB_SAH_BINS + l16] = sah_field_init(f);
}

// ---------------------------------------------------------------------------------------------------------------------
// the partition: std::partition(first, last, [&](ref) { return project_unsafe(center) < split index; }) of libstdc++
// (bidirectional iterators: Hoare's scheme) leaves the k-th misplaced reference from the left and the k-th misplaced one
// from the right swapped, everything else where it was.  F = inclusive prefix sum of the predicate over ALL positions.

struct SahFlag
{
    const SahRef* refs; const uint32_t* seg_of; const SahSplit* splits;
    __device__ __forceinline__ uint32_t operator()(uint32_t p) const
    {
        const uint32_t seg = seg_of[p];
        if (seg == VSB_SAH_NONE) return 0u;
        const SahSplit* sp = splits + seg;
        if (!sp->do_split) return 0u;
        const int axis = sp->axis;
        const float* r = reinterpret_cast<const float*>(refs + p);
        return sah_project_unsafe(sp->k0, sp->k1, sah_center(r[axis], r[4 + axis])) < sp->index ? 1u : 0u;
    }
};

struct SahRange { uint32_t first, end, pivot, before; };      // before = F[first - 1]

__device__ __forceinline__ SahRange sah_range(const SahSeg* sg, const uint32_t* __restrict__ F)
{
    SahRange r;
    r.first = sg->first; r.end = sg->end;
    r.before = r.first ? F[r.first - 1u] : 0u;
    r.pivot = r.first + (F[r.end - 1u] - r.before);
    return r;
}

__global__ void sah_mislist_kernel(const uint32_t* __restrict__ seg_of, uint32_t n, const SahSeg* __restrict__ segs,
                                   const SahSplit* __restrict__ splits, const uint32_t* __restrict__ F,
                                   uint32_t* __restrict__ pos_l, uint32_t* __restrict__ pos_r)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t seg = seg_of[p];
    if (seg == VSB_SAH_NONE || !splits[seg].do_split) return;
    const SahRange r = sah_range(segs + seg, F);
    const uint32_t fp = F[p];
    const bool flag = fp != (p ? F[p - 1u] : 0u);
    if (p < r.pivot && !flag)
    {
        const uint32_t k = (p - r.first + 1u) - (fp - r.before);          // k-th "false" from the left, 1-based
        pos_l[r.first + k - 1u] = p;
    }
    else if (p >= r.pivot && flag)
    {
        const uint32_t k = (F[r.end - 1u] - fp) + 1u;                       // k-th "true" from the right, 1-based
        pos_r[r.first + k - 1u] = p;
    }
}

__global__ void sah_swap_kernel(SahRef* refs, const uint32_t* __restrict__ seg_of, uint32_t n, const SahSeg* __restrict__ segs,
                                const SahSplit* __restrict__ splits, const uint32_t* __restrict__ F,
                                const uint32_t* __restrict__ pos_l, const uint32_t* __restrict__ pos_r)
{
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const uint32_t seg = seg_of[p];
    if (seg == VSB_SAH_NONE || !splits[seg].do_split) return;
    const SahRange r = sah_range(segs + seg, F);
    if (r.pivot == r.first) return;
    const uint32_t misplaced = (r.pivot - r.first) - (F[r.pivot - 1u] - r.before);     // "false" left of the pivot
    const uint32_t j = p - r.first;
    if (j >= misplaced) return;
    const uint32_t a = pos_l[r.first + j], b = pos_r[r.first + j];
    float4* q = reinterpret_cast<float4*>(refs);
    const float4 a0 = q[2 * (size_t)a], a1 = q[2 * (size_t)a + 1], b0 = q[2 * (size_t)b], b1 = q[2 * (size_t)b + 1];
    q[2 * (size_t)a] = b0; q[2 * (size_t)a + 1] = b1;
    q[2 * (size_t)b] = a0; q[2 * (size_t)b + 1] = a1;
}

// ---------------------------------------------------------------------------------------------------------------------
// the next level.  packed[s] = (splits ? 1 : 0) << 32 | open children (0..2); its exclusive scan gives each splitting
// segment its two new build nodes and its open children their places in the next level's (position-ordered) list.

__global__ void sah_child_count_kernel(const uint32_t* __restrict__ state, uint32_t bound, const SahSeg* __restrict__ segs,
                                       const SahSplit* __restrict__ splits, const uint32_t* __restrict__ F, uint64_t* __restrict__ packed)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= bound) return;
    uint64_t v = 0ull;
    if (s < state[0] && splits[s].do_split)
    {
        const SahRange r = sah_range(segs + s, F);
        const uint32_t open = (sah_child_open(r.pivot - r.first, splits[s].lcb) ? 1u : 0u) + (sah_child_open(r.end - r.pivot, splits[s].rcb) ? 1u : 0u);
        v = (1ull << 32) | open;
    }
    packed[s] = v;
}

__global__ void sah_emit_children_kernel(uint32_t* state, uint32_t bound, const SahSeg* __restrict__ segs, const SahSplit* __restrict__ splits,
                                         const uint32_t* __restrict__ F, const uint64_t* __restrict__ packed, const uint64_t* __restrict__ scan,
                                         SahSeg* __restrict__ next, SahNode* nodes, SahChild* __restrict__ child)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= bound) return;
    const uint32_t n_open = state[0], n_nodes = state[1];
    if (s == bound - 1u)
    {
        const uint64_t t = scan[s] + packed[s];
        state[2] = (uint32_t)(t & 0xffffffffull);
        state[3] = n_nodes + 2u * (uint32_t)(t >> 32);
    }
    if (s >= n_open) return;
    const SahSplit* sp = splits + s;
    if (!sp->do_split)                              // stays a leaf: nodes[].left is NONE already
    {
        SahChild c; c.lseg = c.rseg = VSB_SAH_NONE; c.pivot = 0u; c.pad = 0u;
        child[s] = c;
        return;
    }
    const SahSeg* sg = segs + s;
    const SahRange r = sah_range(sg, F);
    uint32_t slot = (uint32_t)(scan[s] & 0xffffffffull);
    const uint32_t left = n_nodes + 2u * (uint32_t)(scan[s] >> 32);
    nodes[sg->node].left = left;
    SahNode nl, nr;
#pragma unroll
    for (int k = 0; k < 6; ++k) { nl.pb[k] = sp->lpb[k]; nr.pb[k] = sp->rpb[k]; }
    nl.first = r.first; nl.end = r.pivot; nl.left = VSB_SAH_NONE; nl.parent = sg->node;
    nr.first = r.pivot; nr.end = r.end;   nr.left = VSB_SAH_NONE; nr.parent = sg->node;
    nodes[left] = nl; nodes[left + 1u] = nr;
    SahChild c;
    c.lseg = c.rseg = VSB_SAH_NONE; c.pivot = r.pivot; c.pad = 0u;
    if (sah_child_open(r.pivot - r.first, sp->lcb))
    {
        SahSeg t;
#pragma unroll
        for (int k = 0; k < 6; ++k) { t.pb[k] = sp->lpb[k]; t.cb[k] = sp->lcb[k]; }
        t.first = r.first; t.end = r.pivot; t.node = left; t.pad = 0u;
        next[slot] = t; c.lseg = slot; ++slot;
    }
    if (sah_child_open(r.end - r.pivot, sp->rcb))
    {
        SahSeg t;
#pragma unroll
        for (int k = 0; k < 6; ++k) { t.pb[k] = sp->rpb[k]; t.cb[k] = sp->rcb[k]; }
        t.first = r.pivot; t.end = r.end; t.node = left + 1u; t.pad = 0u;
        next[slot] = t; c.rseg = slot;
    }
    child[s] = c;
}

// ---------------------------------------------------------------------------------------------------------------------
// layout in the reference's numbering (see the head of this file)

__global__ void sah_sort_key_kernel(const SahNode* __restrict__ nodes, uint32_t n_nodes, uint32_t n, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const SahNode* nd = nodes + i;
    keys[i] = nd->left != VSB_SAH_NONE ? (((uint64_t)(n - nd->end) << 32) | nd->first) : ~0ull;
    vals[i] = i;
}

__global__ void sah_rank_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, uint32_t n_nodes, uint32_t* __restrict__ rank_of)
{
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_nodes || keys[r] == ~0ull) return;
    rank_of[vals[r]] = r;
}

__global__ void sah_emit_kernel(const SahNode* __restrict__ nodes, uint32_t n_nodes, uint32_t n, const uint32_t* __restrict__ rank_of,
                                const SahRef* __restrict__ refs, Node32* __restrict__ out, uint32_t* __restrict__ indices)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const SahNode nd = nodes[i];
    uint32_t pos = 0u;
    if (nd.parent != VSB_SAH_NONE) pos = 1u + 2u * rank_of[nd.parent] + (i - nodes[nd.parent].left);
    Node32 o;
    o.min_x = nd.pb[0]; o.min_y = nd.pb[1]; o.min_z = nd.pb[2];
    o.max_x = nd.pb[3]; o.max_y = nd.pb[4]; o.max_z = nd.pb[5];
    if (nd.left != VSB_SAH_NONE) { o.first = 1u + 2u * rank_of[i]; o.num_prims = 0u; }
    else
    {
        o.first = n - nd.end; o.num_prims = nd.end - nd.first;
        for (uint32_t k = nd.first; k < nd.end; ++k) indices[n - nd.end + (k - nd.first)] = refs[k].index;
    }
    out[pos] = o;
}

#endif // __CUDACC__

} // namespace vsb

// build_ploc.cuh -- BVH2 construction on the GPU.
//
// Replaces the reference's host builder build<index_bvh<P>>(first, last) (bvh.h:455 ->
// detail/bvh/build.inl:28-140 -> binned_sah_builder, detail/bvh/sah.h:150-762: one thread,
// recursive, 1.2 s per million triangles) with a bottom-up build that is parallel at every step:
//
//   1. primitive bounds + centroid bounds                       (prim_bounds_kernel)
//   2. 63-bit Morton code of the centroid, radix sort           (morton_kernel + cub)
//   3. PLOC -- parallel locally-ordered clustering (Meister & Bittner, "Parallel Locally-Ordered
//      Clustering for Bounding Volume Hierarchy Construction", TVCG 2018): every cluster looks
//      for the neighbour within `radius` positions of the Morton-ordered cluster array whose
//      merged box has the smallest area; mutual nearest neighbours merge; the array is compacted
//      in order; repeat until one cluster is left.  The subtree statistics (primitive count,
//      SAH cost, "cheaper as a leaf of <= 4 primitives") are evaluated at merge time, because
//      both children are final by then.
//   4. top-down layout into the REFERENCE's format (bvh_node 32 B with adjacent children,
//      indices[]; bvh.h:52-119,189-236), level by level: the children of the inner node whose
//      left subtree ends at output leaf `s` go to nodes[1 + 2s], nodes[2 + 2s] -- a numbering
//      that needs no allocation counter, so the tree is the same on every run.
//
// What comes out is an ordinary index_bvh: it feeds the same BVH2 -> BVH8 re-encode as a tree from
// the reference builder, the exact traversal walks it in the reference's order, and it can be
// downloaded into the host adaptor's index_bvh<P>.  Any valid BVH gives the same closest hit up
// to t-ties, which is what the parity tests check against the oracle on the reference's tree.
//
// All arithmetic that DECIDES the tree (areas, costs, Morton normalisation) uses single-rounding
// operations (rn_*), so the CPU instantiation in tests/hostsim builds the same tree bit for bit.
#pragma once

#include "vsb_common.cuh"

namespace vsb
{

// build node / cluster record. Build node: leaf <=> right == BN_LEAF, left = primitive index.
// Cluster record (the PLOC working array): left = build node id.
struct BNode
{
    float lox, loy, loz; uint32_t left;
    float hix, hiy, hiz; uint32_t right;
};

// subtree statistics of a build node
struct BSub
{
    uint32_t cnt;     // primitives below | BS_LEAF when the node is emitted as a leaf
    uint32_t lcnt;    // emitted leaves below
    float    cost;    // SAH cost of the subtree as emitted
    uint32_t pad;
};

struct BTask { uint32_t node, out, prim_off, rank_off; };

#define BN_LEAF 0xffffffffu
#define BS_LEAF 0x80000000u
#define BN_NONE 0xffffffffu
#define VSB_BUILD_MAX_LEAF 4        // the BVH8 leaf children hold <= 4 primitives (Node80 meta byte)
#define VSB_BUILD_MAX_RADIUS 32

struct BuildParams
{
    int   radius;       // PLOC search radius (positions in the Morton-ordered cluster array)
    float cost_node;    // SAH: cost of visiting an inner node
    float cost_prim;    // SAH: cost of testing one primitive
};

VSB_HD float rn_min(float a, float b) { return a < b ? a : b; }     // no NaNs in contract
VSB_HD float rn_max(float a, float b) { return a < b ? b : a; }

VSB_HD float box_half_area(float lox, float loy, float loz, float hix, float hiy, float hiz)
{
    const float sx = rn_sub(hix, lox), sy = rn_sub(hiy, loy), sz = rn_sub(hiz, loz);
    return rn_add(rn_add(rn_mul(sx, sy), rn_mul(sy, sz)), rn_mul(sz, sx));
}

VSB_HD float merged_half_area(const BNode& a, const BNode& b)
{
    return box_half_area(rn_min(a.lox, b.lox), rn_min(a.loy, b.loy), rn_min(a.loz, b.loz),
                         rn_max(a.hix, b.hix), rn_max(a.hiy, b.hiy), rn_max(a.hiz, b.hiz));
}

// bounds of primitive i in the reference's layouts: triangle (v1, e1, e2) -> corners v1, v1+e1,
// v1+e2 (get_bounds, math/triangle.inl); sphere -> center +- radius (math/sphere.inl)
VSB_HD BNode prim_bounds(int prim_kind, const void* prims, uint32_t i)
{
    BNode b;
    const float* p = reinterpret_cast<const float*>(reinterpret_cast<const char*>(prims) + (size_t)i * prim_stride(prim_kind));
    if (prim_is_triangle(prim_kind, p))
    {
        const float ax = p[4], ay = p[5], az = p[6];
        const float bx = rn_add(ax, p[8]),  by = rn_add(ay, p[9]),  bz = rn_add(az, p[10]);
        const float cx = rn_add(ax, p[12]), cy = rn_add(ay, p[13]), cz = rn_add(az, p[14]);
        b.lox = rn_min(ax, rn_min(bx, cx)); b.loy = rn_min(ay, rn_min(by, cy)); b.loz = rn_min(az, rn_min(bz, cz));
        b.hix = rn_max(ax, rn_max(bx, cx)); b.hiy = rn_max(ay, rn_max(by, cy)); b.hiz = rn_max(az, rn_max(bz, cz));
    }
    else
    {
        const float r = p[8];
        b.lox = rn_sub(p[4], r); b.loy = rn_sub(p[5], r); b.loz = rn_sub(p[6], r);
        b.hix = rn_add(p[4], r); b.hiy = rn_add(p[5], r); b.hiz = rn_add(p[6], r);
    }
    b.left = i; b.right = BN_LEAF;
    return b;
}

VSB_HD uint64_t spread21(uint32_t v)      // 21 bits -> every third bit of 63
{
    uint64_t x = v & 0x1fffffu;
    x = (x | (x << 32)) & 0x1f00000000ffffull;
    x = (x | (x << 16)) & 0x1f0000ff0000ffull;
    x = (x | (x << 8))  & 0x100f00f00f00f00full;
    x = (x | (x << 4))  & 0x10c30c30c30c30c3ull;
    x = (x | (x << 2))  & 0x1249249249249249ull;
    return x;
}

VSB_HD uint32_t grid21(float c, float lo, float inv_extent)
{
    float f = rn_mul(rn_mul(rn_sub(c, lo), inv_extent), 2097152.0f);
    if (!(f > 0.0f)) return 0u;
    if (f > 2097151.0f) return 2097151u;
    return (uint32_t)f;
}

// cb[0..2] = min, cb[3..5] = max of all centroids
VSB_HD uint64_t morton_of(const BNode& b, const float* cb)
{
    const float cx = rn_mul(0.5f, rn_add(b.lox, b.hix)), cy = rn_mul(0.5f, rn_add(b.loy, b.hiy)), cz = rn_mul(0.5f, rn_add(b.loz, b.hiz));
    const float ex = rn_sub(cb[3], cb[0]), ey = rn_sub(cb[4], cb[1]), ez = rn_sub(cb[5], cb[2]);
    const float ix = ex > 0.0f ? rn_div(1.0f, ex) : 0.0f, iy = ey > 0.0f ? rn_div(1.0f, ey) : 0.0f, iz = ez > 0.0f ? rn_div(1.0f, ez) : 0.0f;
    return (spread21(grid21(cx, cb[0], ix)) << 2) | (spread21(grid21(cy, cb[1], iy)) << 1) | spread21(grid21(cz, cb[2], iz));
}

VSB_HD BSub leaf_sub(const BNode& b, const BuildParams& bp)
{
    BSub s;
    s.cnt = 1u | BS_LEAF; s.lcnt = 1u; s.pad = 0u;
    s.cost = rn_mul(box_half_area(b.lox, b.loy, b.loz, b.hix, b.hiy, b.hiz), bp.cost_prim);
    return s;
}

// nearest neighbour of cluster i among positions [i - radius, i + radius] of tile[], where
// tile[k] holds cluster (tile_first + k) and clusters [0, m) exist.  Smallest merged area wins.
// Equal areas are ordered by a key that is SYMMETRIC in the pair -- position distance, then pairs
// whose lower position is even, then the lower position -- so the order is a strict total order on
// unordered pairs: the globally smallest pair is each other's choice and every iteration merges at
// least one pair, and n identical primitives pair up as (0,1)(2,3).. instead of one merge per pass.
VSB_HD uint32_t ploc_pair_key(uint32_t i, uint32_t j)
{
    const uint32_t lo = i < j ? i : j, d = i < j ? j - i : i - j;
    return (d << 1) | (lo & 1u);
}

VSB_HD uint32_t ploc_nearest(const BNode* tile, long tile_first, uint32_t i, uint32_t m, int radius)
{
    const BNode self = tile[(long)i - tile_first];
    float best = FLT_MAX; uint32_t bj = i, bkey = 0xffffffffu;
    const long j0 = (long)i - radius < 0 ? 0 : (long)i - radius;
    const long j1 = (long)i + radius > (long)m - 1 ? (long)m - 1 : (long)i + radius;
    for (long j = j0; j <= j1; ++j)
    {
        if (j == (long)i) continue;
        const float a = merged_half_area(self, tile[j - tile_first]);
        const uint32_t key = ploc_pair_key(i, (uint32_t)j);
        // same distance and same key bits: j = i - d before j = i + d (the smaller lower position)
        if (a < best || (a == best && key < bkey)) { best = a; bj = (uint32_t)j; bkey = key; }
    }
    return bj;
}

// scan input of cluster i: low word 1 if the cluster survives the iteration (as itself or as the
// merged node), high word 1 if it leads a merge (the lower position of a mutual pair)
VSB_HD uint64_t ploc_flags(const uint32_t* nn, uint32_t i)
{
    const uint32_t j = nn[i];
    const bool mutual = j != i && nn[j] == i;
    const uint64_t survive = (mutual && i > j) ? 0ull : 1ull;
    const uint64_t leader = (mutual && i < j) ? 1ull : 0ull;
    return survive | (leader << 32);
}

// merge / carry cluster i into the compacted output array
VSB_HD void ploc_emit(const BNode* cl_in, const uint32_t* nn, const uint64_t* scan, uint32_t i, uint32_t n_leaves,
                      uint32_t merged_base, const BuildParams& bp, BNode* bn, BSub* sub, uint32_t* parent, BNode* cl_out)
{
    const uint32_t j = nn[i];
    const bool mutual = j != i && nn[j] == i;
    if (mutual && i > j) return;
    const uint32_t pos = (uint32_t)(scan[i] & 0xffffffffull);
    if (!mutual) { cl_out[pos] = cl_in[i]; return; }
    const BNode a = cl_in[i], b = cl_in[j];
    const uint32_t id = n_leaves + merged_base + (uint32_t)(scan[i] >> 32);
    BNode u;
    u.lox = rn_min(a.lox, b.lox); u.loy = rn_min(a.loy, b.loy); u.loz = rn_min(a.loz, b.loz);
    u.hix = rn_max(a.hix, b.hix); u.hiy = rn_max(a.hiy, b.hiy); u.hiz = rn_max(a.hiz, b.hiz);
    u.left = a.left; u.right = b.left;
    bn[id] = u;
    parent[a.left] = id; parent[b.left] = id;
    // subtree statistics: both children are final (created in earlier iterations)
    const BSub sa = sub[a.left], sb = sub[b.left];
    const uint32_t cnt = (sa.cnt & ~BS_LEAF) + (sb.cnt & ~BS_LEAF);
    const float area = box_half_area(u.lox, u.loy, u.loz, u.hix, u.hiy, u.hiz);
    const float c_inner = rn_add(rn_mul(area, bp.cost_node), rn_add(sa.cost, sb.cost));
    const float c_leaf = rn_mul(rn_mul(area, (float)cnt), bp.cost_prim);
    BSub s; s.pad = 0u;
    if (cnt <= VSB_BUILD_MAX_LEAF && c_leaf <= c_inner) { s.cnt = cnt | BS_LEAF; s.lcnt = 1u; s.cost = c_leaf; }
    else { s.cnt = cnt; s.lcnt = sa.lcnt + sb.lcnt; s.cost = c_inner; }
    sub[id] = s;
    u.left = id; u.right = 0u;
    cl_out[pos] = u;
}

// write build node b as a reference leaf: nodes[out] = (bounds, first = prim_off, num_prims) and
// the primitives of its subtree into indices[prim_off ...] in left-to-right order
VSB_HD void layout_leaf(const BNode* bn, const BSub* sub, uint32_t b, uint32_t out, uint32_t prim_off,
                        Node32* nodes, uint32_t* indices)
{
    const BNode n = bn[b];
    Node32 o;
    o.min_x = n.lox; o.min_y = n.loy; o.min_z = n.loz; o.first = prim_off;
    o.max_x = n.hix; o.max_y = n.hiy; o.max_z = n.hiz; o.num_prims = sub[b].cnt & ~BS_LEAF;
    nodes[out] = o;
    uint32_t st[VSB_BUILD_MAX_LEAF + 1]; int sp = 0;
    uint32_t cur = b;
    for (;;)
    {
        const BNode c = bn[cur];
        if (c.right == BN_LEAF)
        {
            indices[prim_off++] = c.left;
            if (sp == 0) break;
            cur = st[--sp];
        }
        else { st[sp++] = c.right; cur = c.left; }
    }
}

// one task of the level-synchronous layout; inner children are appended to `next` through `alloc`
template <typename AllocU32>
VSB_HD void layout_task(const BTask t, const BNode* bn, const BSub* sub, Node32* nodes, uint32_t* indices,
                        BTask* next, AllocU32& alloc)
{
    const BSub s = sub[t.node];
    if (s.cnt & BS_LEAF) { layout_leaf(bn, sub, t.node, t.out, t.prim_off, nodes, indices); return; }
    const BNode n = bn[t.node];
    const BSub sl = sub[n.left], sr = sub[n.right];
    const uint32_t c0 = 1u + 2u * (t.rank_off + sl.lcnt - 1u);
    Node32 o;
    o.min_x = n.lox; o.min_y = n.loy; o.min_z = n.loz; o.first = c0;
    o.max_x = n.hix; o.max_y = n.hiy; o.max_z = n.hiz; o.num_prims = 0u;
    nodes[t.out] = o;
    BTask l, r;
    l.node = n.left;  l.out = c0;      l.prim_off = t.prim_off;                        l.rank_off = t.rank_off;
    r.node = n.right; r.out = c0 + 1u; r.prim_off = t.prim_off + (sl.cnt & ~BS_LEAF);  r.rank_off = t.rank_off + sl.lcnt;
    if (sl.cnt & BS_LEAF) layout_leaf(bn, sub, l.node, l.out, l.prim_off, nodes, indices); else next[alloc(1u)] = l;
    if (sr.cnt & BS_LEAF) layout_leaf(bn, sub, r.node, r.out, r.prim_off, nodes, indices); else next[alloc(1u)] = r;
}

#if defined(__CUDACC__)

// ---- kernels ---------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t float_ordered(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ordered_float(uint32_t o)
{
    const uint32_t u = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

// boxes[i] = bounds of primitive i; cbounds[0..2] = min, [3..5] = max of the centroids (ordered uints)
__global__ void prim_bounds_kernel(int prim_kind, const void* __restrict__ prims, uint32_t n, BNode* __restrict__ boxes,
                                   uint32_t* cbounds)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
    if (i < n)
    {
        const BNode b = prim_bounds(prim_kind, prims, i);
        boxes[i] = b;
        lo[0] = hi[0] = rn_mul(0.5f, rn_add(b.lox, b.hix));
        lo[1] = hi[1] = rn_mul(0.5f, rn_add(b.loy, b.hiy));
        lo[2] = hi[2] = rn_mul(0.5f, rn_add(b.loz, b.hiz));
    }
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o > 0; o >>= 1)
        {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    __shared__ float s[8][6];
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) for (int k = 0; k < 3; ++k) { s[w][k] = lo[k]; s[w][3 + k] = hi[k]; }
    __syncthreads();
    if (threadIdx.x < 6)
    {
        float v = s[0][threadIdx.x];
        for (int k = 1; k < (int)(blockDim.x >> 5); ++k) v = threadIdx.x < 3 ? fminf(v, s[k][threadIdx.x]) : fmaxf(v, s[k][threadIdx.x]);
        if (threadIdx.x < 3) atomicMin(cbounds + threadIdx.x, float_ordered(v));
        else atomicMax(cbounds + threadIdx.x, float_ordered(v));
    }
}

__global__ void morton_kernel(const BNode* __restrict__ boxes, uint32_t n, const uint32_t* __restrict__ cbounds,
                              uint64_t* __restrict__ keys, uint32_t* __restrict__ vals)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float cb[6];
    for (int k = 0; k < 6; ++k) cb[k] = ordered_float(cbounds[k]);
    keys[i] = morton_of(boxes[i], cb);
    vals[i] = i;
}

// sorted position s -> leaf build node s and cluster s
__global__ void init_clusters_kernel(const BNode* __restrict__ boxes, const uint32_t* __restrict__ sorted_vals, uint32_t n,
                                     BuildParams bp, BNode* __restrict__ bn, BSub* __restrict__ sub,
                                     uint32_t* __restrict__ parent, BNode* __restrict__ cl)
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    BNode b = boxes[sorted_vals[s]];
    bn[s] = b;                      // left = primitive index, right = BN_LEAF
    sub[s] = leaf_sub(b, bp);
    parent[s] = BN_NONE;
    b.left = s; b.right = 0u;
    cl[s] = b;
}

#define VSB_PLOC_BLOCK 256

// The iteration state lives on the device: state[2k] = clusters before iteration k, state[2k + 1] = merges before it.
// The host launches iterations in chunks with grids sized for the last count it has read (counts only shrink) and
// syncs once per chunk instead of once per iteration; an iteration on m <= 1 clusters is a no-op that carries the
// state forward.  (The first version read the totals back after every iteration: 5.3 ms per 1 M triangles of which
// 2.6 ms were kernels.)

// nearest neighbours: a block stages its 256 clusters + `radius` on both sides in shared memory
__global__ void __launch_bounds__(VSB_PLOC_BLOCK) ploc_nn_kernel(const BNode* __restrict__ cl, const uint32_t* __restrict__ state, int radius,
                                                                  uint32_t* __restrict__ nn)
{
    __shared__ float4 tile4[2 * (VSB_PLOC_BLOCK + 2 * VSB_BUILD_MAX_RADIUS)];
    const uint32_t m = state[0];
    if (m <= 1u || (uint32_t)blockIdx.x * VSB_PLOC_BLOCK >= m) return;           // block-uniform
    BNode* tile = reinterpret_cast<BNode*>(tile4);
    const long first = (long)blockIdx.x * VSB_PLOC_BLOCK - radius;
    const int span = VSB_PLOC_BLOCK + 2 * radius;
    const float4* src = reinterpret_cast<const float4*>(cl);
    for (int k = threadIdx.x; k < 2 * span; k += VSB_PLOC_BLOCK)
    {
        const long c = first + (k >> 1);
        if (c >= 0 && c < (long)m) tile4[k] = src[2 * c + (k & 1)];
    }
    __syncthreads();
    const uint32_t i = blockIdx.x * VSB_PLOC_BLOCK + threadIdx.x;
    if (i < m) nn[i] = ploc_nearest(tile, first, i, m, radius);
}

__global__ void ploc_flags_kernel(const uint32_t* __restrict__ nn, const uint32_t* __restrict__ state, uint64_t* __restrict__ flags)
{
    const uint32_t m = state[0];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m > 1u && i < m) flags[i] = ploc_flags(nn, i);
}

// writes the next iteration's state: state[2] = clusters after this iteration, state[3] = merges so far
__global__ void ploc_emit_kernel(const BNode* __restrict__ cl_in, const uint32_t* __restrict__ nn,
                                 const uint64_t* __restrict__ scan, uint32_t* state, uint32_t n_leaves,
                                 BuildParams bp, BNode* bn, BSub* sub, uint32_t* parent, BNode* __restrict__ cl_out)
{
    const uint32_t m = state[0], merged_base = state[1];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (m <= 1u)
    {
        if (i == 0u) { state[2] = m; state[3] = merged_base; }       // finished: carry the state forward
        return;
    }
    if (i >= m) return;
    if (i == m - 1)
    {
        const uint64_t t = scan[i] + ploc_flags(nn, i);
        state[2] = (uint32_t)(t & 0xffffffffull);
        state[3] = merged_base + (uint32_t)(t >> 32);
    }
    ploc_emit(cl_in, nn, scan, i, n_leaves, merged_base, bp, bn, sub, parent, cl_out);
}

struct BuildAlloc
{
    uint32_t* counter;
    __device__ uint32_t operator()(uint32_t n) { return atomicAdd(counter, n); }
};

// counters[0] = tasks of this level (written by the level above), counters[1] = tasks emitted for the next level
// (zeroed by the host for the whole chunk of levels in advance)
__global__ void layout_level_kernel(const BTask* __restrict__ tasks, uint32_t* counters, const BNode* __restrict__ bn,
                                    const BSub* __restrict__ sub, Node32* nodes, uint32_t* indices, BTask* next)
{
    const uint32_t n_tasks = counters[0];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tasks) return;
    BuildAlloc a{ counters + 1 };
    layout_task(tasks[i], bn, sub, nodes, indices, next, a);
}

#endif // __CUDACC__

} // namespace vsb

// TEST-ONLY: host instantiation of the product's host/device headers (csrc/*.cuh) so that the
// BVH2 -> BVH8 collapse, the quantised slab test and both traversal loops can be checked on
// the CPU in the no-GPU test tier.  libvsnray_b200.so does not contain or export any of this.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <utility>
#include <vector>

#include "../../visionaray_b200/csrc/bvh8_traverse.cuh"
#include "../../visionaray_b200/csrc/bvh2_exact.cuh"
#include "../../visionaray_b200/csrc/build_ploc.cuh"
#include "../../visionaray_b200/csrc/multi_hit.cuh"
#include "../../visionaray_b200/csrc/watertight.cuh"
#include "../../visionaray_b200/csrc/shade.cuh"
#include <algorithm>
#include <numeric>

using namespace vsb;

namespace
{
struct Alloc
{
    uint32_t* counter;
    uint32_t operator()(uint32_t n) { uint32_t r = *counter; *counter += n; return r; }
};

struct HostWide
{
    std::vector<Node80> nodes;
    std::vector<uint8_t> prims;
    int prim_kind;
    int levels;
};
}

extern "C"
{

void* hs_build(int prim_kind, const void* prims, size_t n_prims, const void* nodes2, size_t n_nodes2,
               const uint32_t* indices, size_t n_idx)
{
    (void)n_prims;
    auto* w = new HostWide();
    w->prim_kind = prim_kind;
    const Node32* n2 = reinterpret_cast<const Node32*>(nodes2);
    size_t n_list = indices ? n_idx : n_prims;
    w->nodes.resize(n_nodes2 ? n_nodes2 : 1);
    std::vector<uint32_t> leaf_list(n_list + 1);
    std::vector<WideTask> cur(1), next(n_nodes2 + 1);
    cur[0].bvh2_node = 0; cur[0].wide_index = 0;
    uint32_t node_count = 1, prim_count = 0;
    w->levels = 0;
    // collapse DP in post-order (sequential form of collapse_dp_kernel)
    std::vector<CollapseCost> cost(n_nodes2);
    std::vector<CollapseDec> dec(n_nodes2);
    {
        std::vector<std::pair<uint32_t, int>> st;
        st.push_back({ 0u, 0 });
        while (!st.empty())
        {
            auto cur_ = st.back(); st.pop_back();
            const Node32& n = n2[cur_.first];
            if (n.num_prims != 0) { dp_leaf(n, cost[cur_.first], dec[cur_.first]); continue; }
            if (cur_.second == 0)
            {
                st.push_back({ cur_.first, 1 });
                st.push_back({ n.first, 0 });
                st.push_back({ n.first + 1, 0 });
            }
            else dp_inner(n, cost[n.first], cost[n.first + 1], cost[cur_.first], dec[cur_.first]);
        }
    }
    const CollapseDec* decp = getenv("VSB_HOSTSIM_GREEDY") ? nullptr : dec.data();
    while (!cur.empty())
    {
        uint32_t task_count = 0;
        Alloc an{ &node_count }, ap{ &prim_count }, at{ &task_count };
        for (auto const& t : cur) collapse_node(t, n2, w->nodes.data(), leaf_list.data(), next.data(), an, ap, at, decp);
        cur.assign(next.begin(), next.begin() + task_count);
        ++w->levels;
    }
    w->nodes.resize(node_count);
    size_t rec = (size_t)wide_rec_words16(prim_kind) * 16;
    w->prims.resize((size_t)prim_count * rec + 16);
    for (uint32_t i = 0; i < prim_count; ++i) make_prim_record(prim_kind, prims, indices, leaf_list[i], w->prims.data(), i);
    return w;
}

void hs_free(void* h) { delete reinterpret_cast<HostWide*>(h); }

void hs_stats(void* h, uint64_t* out)
{
    auto* w = reinterpret_cast<HostWide*>(h);
    out[0] = w->nodes.size();
    out[1] = w->prims.size() / ((size_t)wide_rec_words16(w->prim_kind) * 16);
    out[2] = (uint64_t)w->levels;
}

const void* hs_nodes(void* h) { return reinterpret_cast<HostWide*>(h)->nodes.data(); }
const void* hs_prims(void* h) { return reinterpret_cast<HostWide*>(h)->prims.data(); }

void hs_query_wide(void* h, int any, const void* rays, size_t n_rays, float max_t, const float* max_t_per_ray, void* hits)
{
    auto* w = reinterpret_cast<HostWide*>(h);
    WideScene s{ w->nodes.data(), w->prims.data(), w->prim_kind };
    const Ray32* r = reinterpret_cast<const Ray32*>(rays);
    Hit32* out = reinterpret_cast<Hit32*>(hits);
    for (size_t i = 0; i < n_rays; ++i)
    {
        RayCtx rc;
        ray_ctx_init(rc, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz);
        HitReg b;
        float mt = max_t_per_ray ? max_t_per_ray[i] : max_t;
        if (any) trace_wide<true>(s, rc, mt, b); else trace_wide<false>(s, rc, mt, b);
        out[i].hit = b.hit; out[i].prim_id = b.prim_id; out[i].geom_id = b.geom_id; out[i].t = b.t;
        out[i].u = b.u; out[i].v = b.v; out[i].prim_list_index = b.list_index; out[i].pad = 0;
    }
}

// Work counters of closest_hit on the BVH8 (tools/tree_stats.py): per ray, node visits, leaf groups entered
// (one ring entry of the group ring / one append batch), primitive tests, and the deepest stack
void hs_count_wide(void* h, const void* rays, size_t n_rays, uint64_t* out4)
{
    auto* w = reinterpret_cast<HostWide*>(h);
    WideScene s{ w->nodes.data(), w->prims.data(), w->prim_kind };
    const Ray32* r = reinterpret_cast<const Ray32*>(rays);
    uint64_t visits = 0, groups = 0, tests = 0, deepest = 0;
    for (size_t i = 0; i < n_rays; ++i)
    {
        RayCtx rc;
        ray_ctx_init(rc, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz);
        HitReg best; hit_reset(best);
        float limit = 3.4028234663852886e38f;
        TravSpill spill; TravState t; t.spill = &spill;
        trav_begin(t);
        for (;;)
        {
            const bool prim_mode = trav_prim_mode(t);
            const uint32_t before = t.pg_mask;
            if (prim_mode) tests += 1; else visits += 1;
            const bool more = trav_step(s, rc, t, best, limit, false);
            if (!prim_mode && t.pg_mask) groups += 1;
            (void)before;
            if ((uint64_t)t.sp > deepest) deepest = (uint64_t)t.sp;
            if (!more) break;
        }
    }
    out4[0] = visits; out4[1] = groups; out4[2] = tests; out4[3] = deepest;
}

// VSB_TRAVERSE_WATERTIGHT (csrc/watertight.cuh), host instantiation
void hs_query_watertight(void* h, int any, const void* rays, size_t n_rays, float max_t, void* hits)
{
    auto* w = reinterpret_cast<HostWide*>(h);
    WideScene s{ w->nodes.data(), w->prims.data(), w->prim_kind };
    const Ray32* r = reinterpret_cast<const Ray32*>(rays);
    Hit32* out = reinterpret_cast<Hit32*>(hits);
    for (size_t i = 0; i < n_rays; ++i)
    {
        RayCtx rc;
        ray_ctx_init(rc, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz);
        HitReg b; hit_reset(b);
        float limit = max_t;
        if (any) trace_wide_watertight<true>(s, rc, b, limit); else trace_wide_watertight<false>(s, rc, b, limit);
        out[i].hit = b.hit; out[i].prim_id = b.prim_id; out[i].geom_id = b.geom_id; out[i].t = b.t;
        out[i].u = b.u; out[i].v = b.v; out[i].prim_list_index = b.list_index; out[i].pad = 0;
    }
}

void hs_query_exact(int prim_kind, int any, const void* prims, const void* nodes2, const uint32_t* indices,
                    const void* rays, size_t n_rays, float max_t, const float* max_t_per_ray, void* hits)
{
    ExactScene s{ reinterpret_cast<const Node32*>(nodes2), prims, indices, prim_kind };
    const Ray32* r = reinterpret_cast<const Ray32*>(rays);
    Hit32* out = reinterpret_cast<Hit32*>(hits);
    for (size_t i = 0; i < n_rays; ++i)
    {
        HitReg b;
        float mt = max_t_per_ray ? max_t_per_ray[i] : max_t;
        if (any) trace_exact<true>(s, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz, mt, b);
        else     trace_exact<false>(s, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz, mt, b);
        out[i].hit = b.hit; out[i].prim_id = b.prim_id; out[i].geom_id = b.geom_id; out[i].t = b.t;
        out[i].u = b.u; out[i].v = b.v; out[i].prim_list_index = b.list_index; out[i].pad = 0;
    }
}

// multi_hit<N>: both traversals of csrc/multi_hit.cuh, one ray after the other
void hs_multi_hit_wide(void* h, const void* rays, size_t n_rays, int max_hits, void* hits)
{
    auto* w = reinterpret_cast<HostWide*>(h);
    WideScene s{ w->nodes.data(), w->prims.data(), w->prim_kind };
    const Ray32* r = reinterpret_cast<const Ray32*>(rays);
    for (size_t i = 0; i < n_rays; ++i)
    {
        RayCtx rc;
        ray_ctx_init(rc, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz);
        MultiHitSink sink; sink.rec = reinterpret_cast<Hit32*>(hits) + i * (size_t)max_hits; sink.n = max_hits;
        sink.reset();
        multi_hit_wide(s, rc, sink);
    }
}

void hs_multi_hit_exact(int prim_kind, const void* prims, const void* nodes2, const uint32_t* indices,
                        const void* rays, size_t n_rays, int max_hits, void* hits)
{
    ExactScene s{ reinterpret_cast<const Node32*>(nodes2), prims, indices, prim_kind };
    const Ray32* r = reinterpret_cast<const Ray32*>(rays);
    for (size_t i = 0; i < n_rays; ++i)
    {
        MultiHitSink sink; sink.rec = reinterpret_cast<Hit32*>(hits) + i * (size_t)max_hits; sink.n = max_hits;
        sink.reset();
        multi_hit_exact(s, r[i].ox, r[i].oy, r[i].oz, r[i].dx, r[i].dy, r[i].dz, sink);
    }
}

// Serial form of the GPU builder (csrc/build_ploc.cuh, driven like capi_build.inl drives the
// kernels): same element functions, same order of steps, so the tree is the GPU's bit for bit.
// Returns the number of nodes written (0 if the capacity is too small); info[0] = depth,
// info[1] = PLOC iterations, info[2] = emitted leaves.
size_t hs_build_ploc(int prim_kind, const void* prims, size_t n_prims, int radius, float cost_node, float cost_prim,
                     void* nodes_out, size_t cap_nodes, uint32_t* indices_out, uint32_t* info)
{
    const uint32_t n = (uint32_t)n_prims;
    BuildParams bp{ radius, cost_node, cost_prim };
    std::vector<BNode> boxes(n), bn(2 * (size_t)n - 1), cl[2];
    std::vector<BSub> sub(2 * (size_t)n - 1);
    std::vector<uint32_t> parent(2 * (size_t)n - 1, BN_NONE), nn(n), vals(n);
    std::vector<uint64_t> keys(n), flags(n), scan(n);
    float cb[6] = { FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX };
    for (uint32_t i = 0; i < n; ++i)
    {
        boxes[i] = prim_bounds(prim_kind, prims, i);
        const BNode& b = boxes[i];
        const float c[3] = { 0.5f * (b.lox + b.hix), 0.5f * (b.loy + b.hiy), 0.5f * (b.loz + b.hiz) };
        for (int k = 0; k < 3; ++k) { cb[k] = std::min(cb[k], c[k]); cb[3 + k] = std::max(cb[3 + k], c[k]); }
    }
    for (uint32_t i = 0; i < n; ++i) { keys[i] = morton_of(boxes[i], cb); vals[i] = i; }
    std::stable_sort(vals.begin(), vals.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    cl[0].resize(n); cl[1].resize(n);
    for (uint32_t s = 0; s < n; ++s)
    {
        BNode b = boxes[vals[s]];
        bn[s] = b; sub[s] = leaf_sub(b, bp);
        b.left = s; b.right = 0u; cl[0][s] = b;
    }
    uint32_t m = n, merged = 0, iters = 0; int cur = 0;
    while (m > 1)
    {
        for (uint32_t i = 0; i < m; ++i) nn[i] = ploc_nearest(cl[cur].data(), 0, i, m, radius);
        uint64_t acc = 0;
        for (uint32_t i = 0; i < m; ++i) { flags[i] = ploc_flags(nn.data(), i); scan[i] = acc; acc += flags[i]; }
        for (uint32_t i = 0; i < m; ++i)
            ploc_emit(cl[cur].data(), nn.data(), scan.data(), i, n, merged, bp, bn.data(), sub.data(), parent.data(), cl[cur ^ 1].data());
        m = (uint32_t)(acc & 0xffffffffull); merged += (uint32_t)(acc >> 32);
        cur ^= 1; ++iters;
    }
    const uint32_t root = cl[cur][0].left;
    const size_t n_out = 2 * (size_t)sub[root].lcnt - 1;
    info[1] = iters; info[2] = sub[root].lcnt;
    if (n_out > cap_nodes) return 0;
    Node32* nodes = reinterpret_cast<Node32*>(nodes_out);
    std::vector<BTask> tasks(1), next(n + 1);
    tasks[0].node = root; tasks[0].out = 0; tasks[0].prim_off = 0; tasks[0].rank_off = 0;
    uint32_t levels = 0;
    while (!tasks.empty())
    {
        uint32_t cnt = 0; Alloc a{ &cnt };
        for (auto const& t : tasks) layout_task(t, bn.data(), sub.data(), nodes, indices_out, next.data(), a);
        tasks.assign(next.begin(), next.begin() + cnt);
        ++levels;
    }
    info[0] = (sub[root].cnt & BS_LEAF) ? 1u : levels + 1u;
    return n_out;
}

// the product's ray_sincosf / ray_powf (csrc/shade.cuh), host instantiation: compared bit for bit with the oracle's
// restatement (oracle/vso_raymath.h) in tests/test_raymath.py
void hs_ray_sincosf(const float* x, size_t n, float* s, float* c)
{
    for (size_t i = 0; i < n; ++i) ray_sincosf(x[i], s[i], c[i]);
}
void hs_ray_powf(const float* x, const float* y, size_t n, float* out)
{
    for (size_t i = 0; i < n; ++i) out[i] = ray_powf(x[i], y[i]);
}

} // extern "C"

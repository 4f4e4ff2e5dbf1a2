// multi_hit.cuh -- multi-hit traversal: the N closest intersections along a ray, sorted by t.
//
// Replaces multi_hit<N>(r, begin, end) (detail/traverse_linear.inl:325-380) ->
// intersect<MultiHit, N>() (detail/bvh/intersect.inl:35-134 with RT = array<HR, N>) with the
// reference's record handling (detail/multi_hit.h):
//   * the result is N hit records, all misses (t = FLT_MAX) at first;
//   * a box is entered when it is closer than ANY record (detail/multi_hit.h:206-233), i.e. closer
//     than the last one, because the array stays sorted;
//   * a primitive hit is accepted under the same rule and inserted before the first record with a
//     LARGER t (algo::insert_sorted with is_closer_t, detail/algorithm.h:46-74: equal t goes behind
//     what is already there), the last record drops out;
//   * no early exit (exit_traversal is any-hit only).
// Two traversals, as for closest_hit: the reference-order BVH2 walk (bit-identical records, tie
// order included) and the BVH8 walk (same records except the order / survivor among exact t-ties;
// on trees with spatial splits it reports a primitive once where the reference may report the
// references it happens to visit, see MultiHitSink::contains).
// The records live in the caller's output array (32-byte vsb_hit each, N per ray): a thread only
// ever touches its own N records.
#pragma once

#include "bvh8_traverse.cuh"
#include "bvh2_exact.cuh"

namespace vsb
{

#define VSB_MULTI_HIT_MAX 16

// The N records of one ray in memory (Hit32 layout) + the cull bound in a register.
struct MultiHitSink
{
    Hit32* rec;
    int    n;
    float  limit;      // rec[n - 1].t: what a box / primitive must beat
    uint32_t scene_tag = 0u;   // index of the bvh being walked (rides in rec[].pad while a ray is in flight, see contains())

    VSB_HD void reset()
    {
        for (int k = 0; k < n; ++k)
        {
            Hit32 h; h.hit = 0; h.prim_id = 0u; h.geom_id = 0u; h.t = FLT_MAX; h.u = 0.0f; h.v = 0.0f; h.prim_list_index = 0u; h.pad = 0u;
            rec[k] = h;
        }
        limit = FLT_MAX;
    }

    // Is this primitive already recorded at this t?  A tree built with spatial splits (SBVH,
    // detail/bvh/sah.h:430-615) references a primitive from several leaves; the BVH8 walk may reach a
    // different subset of those references than the reference-order walk does, so it reports every
    // primitive once (the records are sorted: the scan stops at the first larger t).  Scoped to the bvh being
    // walked: the same geometry in two bvhs of a range is two hits, as in the reference.
    VSB_HD bool contains(float t, uint32_t prim_id, uint32_t geom_id) const
    {
        for (int k = 0; k < n; ++k)
        {
            const float tk = rec[k].t;
            if (tk > t) return false;
            if (tk == t && rec[k].prim_id == prim_id && rec[k].geom_id == geom_id && rec[k].pad == scene_tag) return true;
        }
        return false;
    }

    // caller has checked t >= 0 && t < limit
    VSB_HD void insert(float t, float u, float v, uint32_t prim_id, uint32_t geom_id, uint32_t list_index)
    {
        int pos = 0;
        while (pos < n - 1 && !(t < rec[pos].t)) ++pos;        // first record with a larger t (the last one qualifies: t < limit)
        for (int k = n - 1; k > pos; --k) rec[k] = rec[k - 1];
        Hit32 h; h.hit = 1; h.prim_id = prim_id; h.geom_id = geom_id; h.t = t; h.u = u; h.v = v; h.prim_list_index = list_index; h.pad = scene_tag;
        rec[pos] = h;
        limit = rec[n - 1].t;
    }
};

// Reference order: detail/bvh/intersect.inl:35-134 over the 32-byte nodes, near child first.
VSB_HD void multi_hit_exact(const ExactScene& s, float ox, float oy, float oz, float dx, float dy, float dz, MultiHitSink& out)
{
    uint32_t stack[VSB_EXACT_STACK];
    int sp = 0;
    const float ix = rn_div(1.0f, dx), iy = rn_div(1.0f, dy), iz = rn_div(1.0f, dz);
    stack[sp++] = 0u;
    while (sp > 0)
    {
        Node32 node = load_node32(s.nodes, stack[--sp]);
        bool culled = false;
        while (node.num_prims == 0u)
        {
            const Node32 c0 = load_node32(s.nodes, node.first);
            const Node32 c1 = load_node32(s.nodes, node.first + 1u);
            float tn0, tf0, tn1, tf1;
            const bool h0 = isect_aabb_exact(ox, oy, oz, ix, iy, iz, c0.min_x, c0.min_y, c0.min_z, c0.max_x, c0.max_y, c0.max_z, tn0, tf0);
            const bool h1 = isect_aabb_exact(ox, oy, oz, ix, iy, iz, c1.min_x, c1.min_y, c1.min_z, c1.max_x, c1.max_y, c1.max_z, tn1, tf1);
            const bool b0 = h0 && tn0 < out.limit && tf0 >= 0.0f;
            const bool b1 = h1 && tn1 < out.limit && tf1 >= 0.0f;
            if (b0 && b1)
            {
                const bool near0 = tn0 < tn1;
                if (sp >= VSB_EXACT_STACK) VSB_STACK_OVERFLOW();
                stack[sp++] = node.first + (near0 ? 1u : 0u);
                node = near0 ? c0 : c1;
            }
            else if (b0) node = c0;
            else if (b1) node = c1;
            else { culled = true; break; }
        }
        if (culled) continue;
        for (uint32_t i = node.first; i != node.first + node.num_prims; ++i)
        {
            const uint32_t p = s.indices ? s.indices[i] : i;
            float t, u = 0.0f, v = 0.0f;
            uint32_t prim_id, geom_id;
            bool hit;
            const float4* q = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(s.prims) + (size_t)p * prim_stride(s.prim_kind));
            if (prim_is_triangle(s.prim_kind, q))
            {
                const float4 h = q[0], a = q[1], b = q[2], c = q[3];
                geom_id = f2u(h.x); prim_id = f2u(h.y);
                hit = isect_triangle(ox, oy, oz, dx, dy, dz, a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z, t, u, v);
            }
            else
            {
                const float4 h = q[0], a = q[1], b = q[2];
                geom_id = f2u(h.x); prim_id = f2u(h.y);
                hit = isect_sphere(ox, oy, oz, dx, dy, dz, a.x, a.y, a.z, b.x, t);
            }
            if (hit && t >= 0.0f && t < out.limit) out.insert(t, u, v, prim_id, geom_id, i);
        }
    }
}

// BVH8: the quantised 8-child test of the closest-hit path with the N-th record as the cull bound.
VSB_HD void multi_hit_wide(const WideScene& s, const RayCtx& r, MultiHitSink& out)
{
    uint32_t sb[VSB_STACK_SIZE], sm[VSB_STACK_SIZE];
    int sp = 0;
    uint32_t ng_base = 0u, ng_mask = 0x80000000u, pg_base = 0u, pg_mask = 0u, pg_valid = 0u;     // the root as "child 31 of a virtual group"
    const int n16 = wide_rec_words16(s.prim_kind);
    NodeRegs n;
    for (;;)
    {
        if (pg_mask)
        {
            const int bit = find_msb(pg_mask);
            pg_mask &= ~(1u << bit);
            load_record(reinterpret_cast<const char*>(s.prims) + (size_t)prim_record(pg_base, pg_valid, bit) * (size_t)(n16 * 16), n16, n);
            if (s.prim_kind == 0 || (s.prim_kind == 2 && !(n.w[11] & VSB_GENERIC_SPHERE_TAG)))
            {
                float t = -1.0f, u = 0.0f, v = 0.0f;
                if (isect_triangle(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, u2f(n.w[0]), u2f(n.w[1]), u2f(n.w[2]),
                                   u2f(n.w[4]), u2f(n.w[5]), u2f(n.w[6]), u2f(n.w[8]), u2f(n.w[9]), u2f(n.w[10]), t, u, v)
                    && t >= 0.0f && t < out.limit && !out.contains(t, n.w[3], n.w[7]))
                    out.insert(t, u, v, n.w[3], n.w[7], n.w[11]);
            }
            else
            {
                float t;
                if (isect_sphere(r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, u2f(n.w[0]), u2f(n.w[1]), u2f(n.w[2]), u2f(n.w[3]), t)
                    && t >= 0.0f && t < out.limit && !out.contains(t, n.w[4], n.w[5]))
                    out.insert(t, 0.0f, 0.0f, n.w[4], n.w[5], n.w[6]);
            }
            if (pg_mask) continue;
        }
        else
        {
            const int bit = find_msb(ng_mask);
            ng_mask &= ~(1u << bit);
            const uint32_t imask = ng_mask & 0xffu;
            const uint32_t slot = ((uint32_t)(bit - 24) ^ (r.octinv4 & 0xffu)) & 7u;
            const uint32_t rel = (uint32_t)pop_count(imask & ~(0xffffffffu << slot));
            if (ng_mask & 0xff000000u)
            {
                if (sp >= VSB_STACK_SIZE) VSB_STACK_OVERFLOW();
                sb[sp] = ng_base; sm[sp] = ng_mask; ++sp;
            }
            load_record(reinterpret_cast<const char*>(s.nodes + (ng_base + rel)), 5, n);
            uint32_t ph;
            const uint32_t nh = test_node(n, r, out.limit, ph);
            ng_base = n.w[4]; ng_mask = nh | (n.w[3] >> 24);
            pg_base = n.w[5]; pg_mask = ph; pg_valid = n.w[6];
            if (pg_mask) continue;
        }
        if ((ng_mask & 0xff000000u) == 0u)
        {
            if (sp == 0) return;
            --sp; ng_base = sb[sp]; ng_mask = sm[sp];
        }
    }
}

#if defined(__CUDACC__)

// one thread per ray; rays / hits in the caller's order
// (SceneTable = DeviceScene of kernels.cuh: { WideScene wide; ExactScene exact; })
template <bool EXACT, typename SceneTable>
__global__ void __launch_bounds__(128) multi_hit_kernel(const SceneTable* __restrict__ scenes,
                                                        int n_scenes, const Ray32* __restrict__ rays, unsigned long long n_rays,
                                                        int max_hits, Hit32* hits)
{
    const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rays) return;
    const Ray32 ray = rays[i];
    MultiHitSink sink;
    sink.rec = hits + i * (unsigned long long)max_hits; sink.n = max_hits;
    sink.reset();
    // a range of BVHs: the records of every BVH merged by t (update_if(array, array), detail/multi_hit.h:181-202)
    for (int k = 0; k < n_scenes; ++k)
    {
        sink.scene_tag = (uint32_t)k;
        if (EXACT) multi_hit_exact(scenes[k].exact, ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz, sink);
        else
        {
            RayCtx rc;
            ray_ctx_init(rc, ray.ox, ray.oy, ray.oz, ray.dx, ray.dy, ray.dz);
            multi_hit_wide(scenes[k].wide, rc, sink);
        }
    }
    if (n_scenes > 1)
        for (int k = 0; k < max_hits; ++k) sink.rec[k].pad = 0u;      // the tag is scratch: records leave with pad 0
}

#endif

} // namespace vsb

// bvh2_exact.cuh -- VSB_TRAVERSE_EXACT: reference-order traversal of the reference's own BVH2.
//
// Same visit order, same slab arithmetic (no FMA), same tie rules as
// detail/bvh/intersect.inl:35-134 (near child first, ties -> child 1; strict '<' so the first
// visited primitive wins equal t; any-hit returns the first accepted primitive).  Results are
// bit-identical to the reference's scalar host traversal, including primitive_list_index and
// the any-hit winner.  It is the parity anchor of the fast BVH8 path and the drop-in for
// callers that depend on the reference's tie order.  The stack holds 64 entries (the
// reference: 31, unchecked, detail/stack.h:39-47); deeper trees are rejected at scene creation.
#pragma once

#include "vsb_common.cuh"

namespace vsb
{

struct ExactScene
{
    const Node32*   nodes;
    const void*     prims;      // reference layout: 64-B triangles or 48-B spheres
    const uint32_t* indices;    // NULL for a non-index bvh
    int             prim_kind;
};

#define VSB_EXACT_STACK 64

VSB_HD Node32 load_node32(const Node32* __restrict__ nodes, uint32_t i)
{
#if defined(__CUDA_ARCH__)
    const float4* p = reinterpret_cast<const float4*>(nodes + i);
    const float4 a = __ldg(p), b = __ldg(p + 1);
    Node32 n;
    n.min_x = a.x; n.min_y = a.y; n.min_z = a.z; n.first = __float_as_uint(a.w);
    n.max_x = b.x; n.max_y = b.y; n.max_z = b.z; n.num_prims = __float_as_uint(b.w);
    return n;
#else
    return nodes[i];
#endif
}

template <bool ANY>
VSB_HD void trace_exact(const ExactScene& s, float ox, float oy, float oz, float dx, float dy, float dz,
                        float max_t, HitReg& best)
{
    hit_reset(best);
    uint32_t stack[VSB_EXACT_STACK];
    int sp = 0;
    const float ix = rn_div(1.0f, dx), iy = rn_div(1.0f, dy), iz = rn_div(1.0f, dz);   // :63
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
            // is_closer(box, result, max_t): update_if.h:62-66,84-88
            const bool b0 = h0 && tn0 < best.t && tf0 >= 0.0f && tn0 < max_t;
            const bool b1 = h1 && tn1 < best.t && tf1 >= 0.0f && tn1 < max_t;
            if (b0 && b1)
            {
                const bool near0 = tn0 < tn1;                 // ties -> child 1 (:86)
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
            const uint32_t p = s.indices ? s.indices[i] : i;   // bvh.h:226-229 / :177-180
            float t, u = 0.0f, v = 0.0f;
            uint32_t prim_id, geom_id;
            bool hit;
            const float4* q = reinterpret_cast<const float4*>(reinterpret_cast<const char*>(s.prims) + (size_t)p * prim_stride(s.prim_kind));
            if (prim_is_triangle(s.prim_kind, q))
            {
#if defined(__CUDA_ARCH__)
                const float4 h = __ldg(q), a = __ldg(q + 1), b = __ldg(q + 2), c = __ldg(q + 3);
#else
                const float4 h = q[0], a = q[1], b = q[2], c = q[3];
#endif
                geom_id = f2u(h.x); prim_id = f2u(h.y);
                hit = isect_triangle(ox, oy, oz, dx, dy, dz, a.x, a.y, a.z, b.x, b.y, b.z, c.x, c.y, c.z, t, u, v);
            }
            else
            {
#if defined(__CUDA_ARCH__)
                const float4 h = __ldg(q), a = __ldg(q + 1), b = __ldg(q + 2);
#else
                const float4 h = q[0], a = q[1], b = q[2];
#endif
                geom_id = f2u(h.x); prim_id = f2u(h.y);
                hit = isect_sphere(ox, oy, oz, dx, dy, dz, a.x, a.y, a.z, b.x, t);
            }
            // is_closer(hr, result, max_t): update_if.h:50-56,75-79
            if (hit && t >= 0.0f && t < best.t && t < max_t)
            {
                best.t = t; best.u = u; best.v = v; best.hit = 1;
                best.prim_id = prim_id; best.geom_id = geom_id; best.list_index = i;
                if (ANY) return;
            }
        }
    }
}

} // namespace vsb

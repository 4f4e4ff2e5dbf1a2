// kernels.cuh -- the __global__ kernels of libvsnray_b200.so (sm_100a).
//
//  encode_level_kernel   BVH2 -> BVH8 collapse, one thread per wide node of the current level
//  fill_prims_kernel     leaf-order primitive records (removes the indices[] indirection)
//  trace_kernel<Program> persistent warps running ONE warp-synchronous traversal loop; lanes are
//                        refilled individually (ray compaction): a lane whose ray finishes runs
//                        its program's continuation (write the hit / shade / next bounce / next
//                        pixel) while the other lanes keep their rays.  Programs: RayListProgram
//                        (closest_hit / any_hit over a ray list) and the frame programs of
//                        shade.cuh (simple / whitted / ao / pathtracing: ray gen + kernel + store)
//  deinterleave_kernel   rank-major tile buffers -> row-major image (after the NCCL all-gather)
//
// Grids are sized as (resident CTAs per SM) x (SM count) and every warp pulls work from a
// global atomic counter (persistent-thread work queue), so load imbalance between rays that
// terminate early (sky, any-hit) and deep rays is absorbed by the queue, not by the tail of
// a static grid.
#pragma once

#include <cuda_runtime.h>

#include "bvh8_traverse.cuh"
#include "bvh2_exact.cuh"
#include "shade.cuh"

namespace vsb
{

struct DeviceScene
{
    WideScene  wide;
    ExactScene exact;
};

struct DeviceAlloc
{
    uint32_t* counter;
    __device__ uint32_t operator()(uint32_t n) { return atomicAdd(counter, n); }
};

// parent[] of the BVH2 (the reference layout has child links only)
__global__ void bvh2_parents_kernel(const Node32* __restrict__ nodes2, uint32_t n_nodes, uint32_t* __restrict__ parent)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    if (i == 0) parent[0] = 0xffffffffu;
    const Node32 n = nodes2[i];
    if (n.num_prims == 0) { parent[n.first] = i; parent[n.first + 1] = i; }
}

// Bottom-up collapse DP (bvh8.cuh): one thread per BVH2 leaf climbs towards the root; at every
// inner node the first arriving thread stops, the second (both children done) evaluates the node.
__global__ void collapse_dp_kernel(const Node32* __restrict__ nodes2, uint32_t n_nodes, const uint32_t* __restrict__ parent,
                                   uint32_t* __restrict__ arrived, CollapseCost* cost, CollapseDec* dec)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    Node32 n = nodes2[i];
    if (n.num_prims == 0) return;
    CollapseCost c; CollapseDec d;
    dp_leaf(n, c, d);
    cost[i] = c; dec[i] = d;
    for (;;)
    {
        const uint32_t p = parent[i];
        if (p == 0xffffffffu) return;
        __threadfence();
        if (atomicAdd(arrived + p, 1u) == 0u) return;
        __threadfence();
        n = nodes2[p];
        const volatile CollapseCost* vc = cost;
        CollapseCost l, r;
        for (int k = 0; k < 8; ++k) { l.c[k] = vc[n.first].c[k]; r.c[k] = vc[n.first + 1].c[k]; }
        dp_inner(n, l, r, c, d);
        cost[p] = c; dec[p] = d;
        i = p;
    }
}

// counters[0] = next free wide node, [1] = next free primitive slot, [2] = tasks emitted
__global__ void encode_level_kernel(const WideTask* __restrict__ tasks, uint32_t n_tasks,
                                    const Node32* __restrict__ nodes2, Node80* out_nodes,
                                    uint32_t* leaf_list, WideTask* next_tasks, uint32_t* counters,
                                    const CollapseDec* __restrict__ dec)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tasks) return;
    DeviceAlloc an{ counters + 0 }, ap{ counters + 1 }, at{ counters + 2 };
    collapse_node(tasks[i], nodes2, out_nodes, leaf_list, next_tasks, an, ap, at, dec);
}

__global__ void fill_prims_kernel(int prim_kind, const void* __restrict__ prims, const uint32_t* __restrict__ indices,
                                  const uint32_t* __restrict__ leaf_list, uint32_t n, void* out, uint32_t* __restrict__ rec_of_list)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    make_prim_record(prim_kind, prims, indices, leaf_list[i], out, i);
    rec_of_list[leaf_list[i]] = i;       // every list position sits in exactly one leaf (validate_tree): a permutation
}

// largest prim_id / geom_id of a primitive array (ids are caller data: per-primitive arrays are indexed by
// prim_id, materials and textures by geom_id -- checked once per scene instead of trusted)
__global__ void id_range_kernel(int prim_kind, const void* __restrict__ prims, size_t n, uint32_t* __restrict__ out2)
{
    uint32_t mp = 0u, mg = 0u;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    {
        const uint32_t* rec = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(prims) + i * prim_stride(prim_kind));
        mg = rec[0] > mg ? rec[0] : mg;
        mp = rec[1] > mp ? rec[1] : mp;
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        const uint32_t a = __shfl_xor_sync(0xffffffffu, mp, o), b = __shfl_xor_sync(0xffffffffu, mg, o);
        mp = a > mp ? a : mp; mg = b > mg ? b : mg;
    }
    if ((threadIdx.x & 31u) == 0u) { atomicMax(out2, mp); atomicMax(out2 + 1, mg); }
}

// ---- the persistent trace kernel ----------------------------------------------------------------
//
// Work items (rays of a list, pixels of a frame) are handed out in two levels: a warp reserves a
// CHUNK of consecutive items with one atomicAdd and then refills idle lanes from its chunk with
// ballot/popc arithmetic only.  Lane states:
//   TRAVERSING  the lane owns a ray and takes one trav_step() per loop iteration
//   WAITING     its ray finished; the program's done() continuation has not run yet
//   EMPTY       no work item; wants one from the chunk
//   DEAD        no work item and the queue is exhausted
// The continuations and the refill ("service") run only when at least service_min lanes are not
// traversing (or nobody is), so that shading / ray setup code is not executed for one lane at a
// time: this bounds both the idle lanes (< service_min / 32) and the per-ray service overhead.

enum { LANE_TRAVERSING = 0, LANE_WAITING = 1, LANE_EMPTY = 2, LANE_DEAD = 3 };

struct TraceLaunch
{
    const DeviceScene* scenes;
    int n_scenes;
    unsigned long long n_items;
    unsigned chunk;                       // items per warp reservation (multiple of 32)
    unsigned long long* work_counter;
    unsigned long long* ray_count;        // optional: += rays traced
    int service_min;                      // lanes that must be idle before continuations / refill run
    int phase_min;                        // phase vote threshold (phase_runs)
    int drain_min, pend_min;              // trace_kernel_q: pairs that trigger a drain; lanes waiting on their pairs that force one
    WideScene wide0;                      // trace_kernel_q (one bvh per launch): scenes[0].wide BY VALUE, i.e. in the constant bank --
                                          // as a copy read from the scene table it was spilled, and an LDL.64 of the node
                                          // array's base sat in front of every node load (profiles/r2_sass_trace_kernels.txt)
};

#define VSB_BLOCK 128
#ifndef VSB_MIN_BLOCKS
#define VSB_MIN_BLOCKS 6     // <= 80 registers: 24 warps per SM (measured best of 5..8)
#endif

// Phase vote among the traversing lanes of a warp (called by all of them): a lane is in
// primitive phase or node phase (trav_prim_mode).  Both phases run in the same iteration when each
// has at least `phase_min` lanes; otherwise only the majority phase runs and the minority lanes
// wait one iteration, so neither the node test nor the primitive test is executed for a handful of
// lanes.  Coherent packets stay in lockstep; incoherent warps gather lanes per phase.
__device__ __forceinline__ bool phase_runs(int phase_min, bool prim_mode)
{
    if (phase_min <= 1) return true;
    const unsigned m = __activemask();
    const unsigned mp = __ballot_sync(m, prim_mode);
    const int n_prim = __popc(mp), n_node = __popc(m) - n_prim;
    const bool run_prim = n_prim >= phase_min || n_prim >= n_node;
    const bool run_node = n_node >= phase_min || n_node > n_prim;
    return prim_mode ? run_prim : run_node;
}

template <typename Program, bool EXACT>
__global__ void __launch_bounds__(VSB_BLOCK, Program::MIN_BLOCKS)
trace_kernel(const TraceLaunch tl, const __grid_constant__ Program prog)
{
    extern __shared__ uint32_t lane_store_raw[];
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    // shared memory: [32 stack words per thread][Program::STATE_WORDS per thread], column per thread
    typename Program::Lane L(lane_store_raw + 2 * VSB_SHARED_STACK * VSB_BLOCK, threadIdx.x);

    unsigned long long chunk_next = 0, chunk_end = 0;     // warp-uniform
    bool exhausted = false;                               // warp-uniform
    int state = LANE_EMPTY;
    unsigned n_rays = 0;

    RayCtx rc; rc.ox = rc.oy = rc.oz = rc.dx = rc.dy = rc.dz = 0.0f; rc.ix = rc.iy = rc.iz = 0.0f; rc.octinv4 = 0u;
    HitReg best; hit_reset(best);
    float limit = 0.0f, ray_max_t = 0.0f;
    bool any = false;
    int scene_i = 0;
    WideScene sc = tl.scenes[0].wide;
    TravSpill spill;
    TravState t; trav_begin(t);
    t.spill = &spill;
    t.sstack = (uint32_t)__cvta_generic_to_shared(lane_store_raw + threadIdx.x);
    RayReq req;

    for (;;)
    {
        const unsigned not_trav = __ballot_sync(0xffffffffu, state == LANE_WAITING || state == LANE_EMPTY);
        const unsigned trav = __ballot_sync(0xffffffffu, state == LANE_TRAVERSING);
        if (not_trav != 0u && (__popc(not_trav) >= tl.service_min || trav == 0u))
        {
            // ---- continuations of finished rays ----
            bool start = false;
            if (state == LANE_WAITING)
            {
                if (prog.done(L, best, rc.ox, rc.oy, rc.oz, rc.dx, rc.dy, rc.dz, req)) start = true;
                else state = LANE_EMPTY;
            }
            // ---- refill empty lanes from the warp's chunk ----
            unsigned empty = __ballot_sync(0xffffffffu, state == LANE_EMPTY);
            while (empty != 0u && !exhausted)
            {
                if (chunk_next >= chunk_end)
                {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(tl.work_counter, (unsigned long long)tl.chunk);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base >= tl.n_items) { exhausted = true; break; }
                    chunk_next = base;
                    chunk_end = base + tl.chunk < tl.n_items ? base + tl.chunk : tl.n_items;
                }
                const unsigned long long item = chunk_next + (unsigned long long)__popc(empty & lt_mask);
                if (state == LANE_EMPTY && item < chunk_end)
                {
                    // begin() may reject the item (pixel outside the image / scissor): lane stays EMPTY
                    if (prog.begin(L, item, req)) { start = true; state = LANE_WAITING; }
                }
                const unsigned long long taken = (unsigned long long)__popc(empty);
                chunk_next = chunk_next + taken < chunk_end ? chunk_next + taken : chunk_end;
                empty = __ballot_sync(0xffffffffu, state == LANE_EMPTY);
            }
            if (exhausted && state == LANE_EMPTY) state = LANE_DEAD;
            if (start)
            {
                ++n_rays;
                any = req.any; ray_max_t = req.max_t; limit = req.max_t;
                hit_reset(best);
                if (EXACT) { rc.ox = req.ox; rc.oy = req.oy; rc.oz = req.oz; rc.dx = req.dx; rc.dy = req.dy; rc.dz = req.dz; }
                else ray_ctx_init(rc, req.ox, req.oy, req.oz, req.dx, req.dy, req.dz);
                scene_i = 0;
                if (tl.n_scenes > 1) sc = tl.scenes[0].wide;
                trav_begin(t);
                state = LANE_TRAVERSING;
            }
            if (__all_sync(0xffffffffu, state == LANE_DEAD)) break;
        }

        if (state == LANE_TRAVERSING)
        {
            if (EXACT)
            {
                // reference-order BVH2 walk (blocking); over the range of bvhs as
                // detail/traverse_linear.inl:106-127: each bvh from scratch, merged with strict '<'
                for (int k = 0; k < tl.n_scenes; ++k)
                {
                    HitReg h;
                    if (any) trace_exact<true>(tl.scenes[k].exact, rc.ox, rc.oy, rc.oz, rc.dx, rc.dy, rc.dz, ray_max_t, h);
                    else     trace_exact<false>(tl.scenes[k].exact, rc.ox, rc.oy, rc.oz, rc.dx, rc.dy, rc.dz, ray_max_t, h);
                    if (k == 0) best = h;
                    else if (h.hit && h.t >= 0.0f && h.t < best.t && h.t < ray_max_t) best = h;
                    if (any && best.hit) break;
                }
                state = LANE_WAITING;
            }
            else if (!phase_runs(tl.phase_min, trav_prim_mode(t))) { /* minority phase: wait for the lanes to gather */ }
            else if (!trav_step(sc, rc, t, best, limit, any))
            {
                // this bvh is finished: next bvh of the range (same best / limit == strict '<' merge)
                if (!(any && best.hit) && ++scene_i < tl.n_scenes)
                {
                    sc = tl.scenes[scene_i].wide;
                    trav_begin(t);
                }
                else state = LANE_WAITING;
            }
        }
    }

    if (tl.ray_count)
    {
        unsigned n = n_rays;
        for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
        if (lane == 0 && n) atomicAdd(tl.ray_count, (unsigned long long)n);
    }
}

// ---- the queue kernel: node visits per lane, primitive tests shared by the warp -------------------
//
// Same persistent loop, programs and work queue as trace_kernel, but a lane no longer tests its own
// primitives.  Every iteration all traversing lanes take ONE NODE VISIT; the primitives the node test
// selects are appended as (leaf-order record, owner) pairs to a per-warp ring in shared memory, and
// when the ring holds 32 pairs (or nothing else can run) ALL 32 LANES test one pair each: record loaded
// from global memory, the owner's ray from shared memory, result merged into the owner's 64-bit key
// (t bits << 32 | primitive_list_index) with atomicMin.  The primitive test thus runs at (nearly) full warp width
// however few lanes found primitives in a given step -- it ran at 9.4 of 32 lanes in trace_kernel
// (profiles/r1_cfg2_instruction_regions_v8.txt) -- and the node test is no longer skipped by lanes that
// are busy with primitives.  Cost: an owner's cull bound lags its pending pairs by an iteration or two.
// Replaces the leaf loop of detail/bvh/intersect.inl:104-130.
//
// Lane states: TRAVERSING (node work left), PENDING (node work done, pairs of this lane still queued:
// it is finished when the ring's head has passed the lane's last pair), WAITING / EMPTY / DEAD as above.
// Ties: the key orders (t, primitive_list_index), so among equal t the smallest position in the reference's
// primitive list wins -- a fixed rule, independent of scheduling AND of how the BVH8 happened to be laid out
// (its record order comes from atomic allocation and differs between builds; the first version keyed on it and
// two GPUs disagreed on tie pixels).  Tie candidates are never culled: the node test is conservative.
// Any-hit lanes leave as soon as their key holds a hit; pairs they leave behind are tested against
// whatever ray the lane holds then, which is a valid (ray, primitive) test of the same scene.
// One bvh per launch, BVH8 walk only: ranges of bvhs and the reference-order walk stay on trace_kernel.

#ifndef VSB_QCAP
#define VSB_QCAP 128              // pairs per warp ring (power of two, >= 64: a full ring always drains 32); 128 measured +2 % on cfg2 over 64
#endif
#define LANE_PENDING 4
#define VSB_KEY_NONE 0xffffffffu

// shared memory of trace_kernel_q in 32-bit words per CTA: stack | program state | ray (6) | key (2) | rings
template <typename Program>
__host__ __device__ constexpr size_t trace_q_smem_words()
{
    return (size_t)(2 * VSB_Q_SHARED_STACK + Program::STATE_WORDS + 8) * VSB_BLOCK + (size_t)(VSB_BLOCK / 32) * VSB_QCAP * 2;
}

template <typename Program>
__global__ void __launch_bounds__(VSB_BLOCK, Program::MIN_BLOCKS_Q)
trace_kernel_q(const TraceLaunch tl, const __grid_constant__ Program prog)
{
    extern __shared__ uint32_t lane_store_raw[];
    __shared__ uint8_t perm_lut[2048];                    // child-order permutation of a hit mask (test_node_lut)
    perm_lut_fill(perm_lut);
    __syncthreads();
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    uint32_t* const prog_state = lane_store_raw + 2 * VSB_Q_SHARED_STACK * VSB_BLOCK;
    typename Program::Lane L(prog_state, threadIdx.x);
    float* const ray_s = reinterpret_cast<float*>(prog_state + Program::STATE_WORDS * VSB_BLOCK);        // [6][VSB_BLOCK]
    unsigned long long* const key_s = reinterpret_cast<unsigned long long*>(ray_s + 6 * VSB_BLOCK);      // [VSB_BLOCK]
    uint2* const ring = reinterpret_cast<uint2*>(key_s + VSB_BLOCK) + (threadIdx.x >> 5) * VSB_QCAP;     // this warp's pairs

    unsigned long long chunk_next = 0, chunk_end = 0;     // warp-uniform
    bool exhausted = false;                               // warp-uniform
    unsigned q_head = 0u, q_tail = 0u;                    // warp-uniform pair sequence numbers (ring index = seq % VSB_QCAP)
    int state = LANE_EMPTY;
    unsigned n_rays = 0, my_seq = 0u;

    RayCtx rc; rc.ox = rc.oy = rc.oz = rc.dx = rc.dy = rc.dz = 0.0f; rc.ix = rc.iy = rc.iz = 0.0f; rc.octinv4 = 0u;
    float limit = 0.0f;
    bool any = false, live = false;                       // live: max_t > 0 (a ray with max_t <= 0 cannot hit and is not traced)
    const WideScene& sc = tl.wide0;
    const int n16 = wide_rec_words16(sc.prim_kind);
    TravSpill spill;
    TravState t; trav_begin(t);
    t.spill = &spill;
    t.sstack = (uint32_t)__cvta_generic_to_shared(lane_store_raw + threadIdx.x);
    RayReq req;

    // Program::GROUP_RING (ray lists, the AO and path-tracing frame programs; measured +4.5 % cfg3, +5.6 % cfg4, +5 % cfg5, -0.7 % cfg2: gpurun_out/r2_ab6.log):
    // one entry per (lane, node visit) = (first record, primitive mask, owner); a drain takes the leading
    // groups that hold <= 32 pairs together and expands them to one pair per lane cooperatively
    uint4* const gring = reinterpret_cast<uint4*>(ring);           // VSB_QCAP / 2 groups of 16 bytes in the same 1 KiB per warp
    unsigned q_pairs = 0u;                                         // warp-uniform: pairs in the ring
    auto drain = [&]()
    {
        if constexpr (Program::GROUP_RING)
        {
            const unsigned avail = q_tail - q_head;
            uint4 e = make_uint4(0u, 0u, 0u, 0u);
            if (lane < avail) e = gring[(q_head + lane) & (VSB_QCAP / 2 - 1u)];
            const int c = __popc(e.y);
            int incl = c;
    #pragma unroll
            for (int o = 1; o < 32; o <<= 1)
            {
                const int v = __shfl_up_sync(0xffffffffu, incl, o);
                if ((int)lane >= o) incl += v;
            }
            const int ntake = __popc(__ballot_sync(0xffffffffu, lane < avail && incl <= 32));     // leading groups (incl is monotone), >= 1
            const int n = __shfl_sync(0xffffffffu, incl, ntake - 1);
            const int excl = incl - c;
            const unsigned word = __reduce_or_sync(0xffffffffu, (int)lane < ntake ? 1u << excl : 0u);   // bit at the first pair of every taken group
            const int g = __popc(word & (0xffffffffu >> (31u - lane))) - 1;                           // the group lane's pair belongs to
            const uint32_t gb = __shfl_sync(0xffffffffu, e.x, g), go = __shfl_sync(0xffffffffu, e.z, g), gv = __shfl_sync(0xffffffffu, e.w, g);
            uint32_t gm = __shfl_sync(0xffffffffu, e.y, g);
            int k = (int)lane - __shfl_sync(0xffffffffu, excl, g);
            if ((int)lane < n)
            {
                while (k-- > 0) gm &= gm - 1u;                       // the k-th set bit of the group's mask
                const uint32_t rec = prim_record(gb, gv, __ffs((int)gm) - 1);
                NodeRegs r;
                load_record(reinterpret_cast<const char*>(sc.prims) + (size_t)rec * (size_t)(n16 * 16), n16, r);
                const float* o = ray_s + go;
                float tt;
                if (prim_candidate(sc.prim_kind, r, o[0], o[VSB_BLOCK], o[2 * VSB_BLOCK], o[3 * VSB_BLOCK], o[4 * VSB_BLOCK], o[5 * VSB_BLOCK], tt) && tt >= 0.0f)
                {
                    const unsigned long long cand = ((unsigned long long)(__float_as_uint(tt) & 0x7fffffffu) << 32) | record_list_index(sc.prim_kind, r);
                    if (cand < key_s[go]) atomicMin(key_s + go, cand);
                }
            }
            q_head += (unsigned)ntake;
            q_pairs -= (unsigned)n;
            __syncwarp();
        }
        else
        {
            // PAIR ring: one batch of up to 32 (record, owner) pairs, one per lane
            const unsigned n = q_tail - q_head < 32u ? q_tail - q_head : 32u;
            if (lane < n)
            {
                const uint2 e = ring[(q_head + lane) & (VSB_QCAP - 1u)];
                NodeRegs r;
                load_record(reinterpret_cast<const char*>(sc.prims) + (size_t)e.x * (size_t)(n16 * 16), n16, r);
                const float* o = ray_s + e.y;
                float tt;
                if (prim_candidate(sc.prim_kind, r, o[0], o[VSB_BLOCK], o[2 * VSB_BLOCK], o[3 * VSB_BLOCK], o[4 * VSB_BLOCK], o[5 * VSB_BLOCK], tt) && tt >= 0.0f)
                {
                    const unsigned long long cand = ((unsigned long long)(__float_as_uint(tt) & 0x7fffffffu) << 32) | record_list_index(sc.prim_kind, r);
                    if (cand < key_s[e.y]) atomicMin(key_s + e.y, cand);
                }
            }
            q_head += n;
            __syncwarp();
        }
    };

    for (;;)
    {
        const unsigned not_trav = __ballot_sync(0xffffffffu, state == LANE_WAITING || state == LANE_EMPTY);
        const unsigned trav = __ballot_sync(0xffffffffu, state == LANE_TRAVERSING || state == LANE_PENDING);
        if (not_trav != 0u && (__popc(not_trav) >= tl.service_min || trav == 0u))
        {
            // ---- continuations of finished rays ----
            bool start = false;
            if (state == LANE_WAITING)
            {
                const float* o = ray_s + threadIdx.x;
                const float ox = o[0], oy = o[VSB_BLOCK], oz = o[2 * VSB_BLOCK], dx = o[3 * VSB_BLOCK], dy = o[4 * VSB_BLOCK], dz = o[5 * VSB_BLOCK];
                const uint32_t li = (uint32_t)key_s[threadIdx.x];
                HitReg best;
                if (live && li != VSB_KEY_NONE) hit_from_record(sc, __ldg(sc.rec_of_list + li), ox, oy, oz, dx, dy, dz, best);
                else hit_reset(best);
                if (prog.done(L, best, ox, oy, oz, dx, dy, dz, req)) start = true;
                else state = LANE_EMPTY;
            }
            // ---- refill empty lanes from the warp's chunk ----
            unsigned empty = __ballot_sync(0xffffffffu, state == LANE_EMPTY);
            while (empty != 0u && !exhausted)
            {
                if (chunk_next >= chunk_end)
                {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(tl.work_counter, (unsigned long long)tl.chunk);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base >= tl.n_items) { exhausted = true; break; }
                    chunk_next = base;
                    chunk_end = base + tl.chunk < tl.n_items ? base + tl.chunk : tl.n_items;
                }
                const unsigned long long item = chunk_next + (unsigned long long)__popc(empty & lt_mask);
                if (state == LANE_EMPTY && item < chunk_end)
                {
                    if (prog.begin(L, item, req)) { start = true; state = LANE_WAITING; }
                }
                const unsigned long long taken = (unsigned long long)__popc(empty);
                chunk_next = chunk_next + taken < chunk_end ? chunk_next + taken : chunk_end;
                empty = __ballot_sync(0xffffffffu, state == LANE_EMPTY);
            }
            if (exhausted && state == LANE_EMPTY) state = LANE_DEAD;
            if (start)
            {
                ++n_rays;
                any = req.any;
                float* o = ray_s + threadIdx.x;
                o[0] = req.ox; o[VSB_BLOCK] = req.oy; o[2 * VSB_BLOCK] = req.oz;
                o[3 * VSB_BLOCK] = req.dx; o[4 * VSB_BLOCK] = req.dy; o[5 * VSB_BLOCK] = req.dz;
                // key: a candidate (t, rec) is accepted iff it is smaller; t < max_t strictly <=> bits(t) <= bits(max_t) - 1,
                // and t < FLT_MAX as against the reference's empty record (math/intersect.h:95-103)
                live = req.max_t > 0.0f;
                const float mt = req.max_t < FLT_MAX ? req.max_t : FLT_MAX;
                const uint32_t hi = live ? __float_as_uint(mt) - 1u : 0u;
                key_s[threadIdx.x] = ((unsigned long long)hi << 32) | (live ? VSB_KEY_NONE : 0u);
                limit = __uint_as_float(hi);
                ray_ctx_init(rc, req.ox, req.oy, req.oz, req.dx, req.dy, req.dz);
                trav_begin(t);
                state = live ? LANE_TRAVERSING : LANE_WAITING;
            }
            __syncwarp();
            if (__all_sync(0xffffffffu, state == LANE_DEAD)) break;
        }

        // ---- node phase: one node visit per traversing lane ----
        if (state == LANE_TRAVERSING && (t.ng_mask >> 24) != 0u)
        {
            const int bit = find_msb(t.ng_mask);
            t.ng_mask &= ~(1u << bit);
            const uint32_t imask = t.ng_mask & 0xffu;
            const uint32_t slot = ((uint32_t)(bit - 24) ^ (rc.octinv4 & 0xffu)) & 7u;
            const uint32_t rel = (uint32_t)pop_count(imask & ~(0xffffffffu << slot));
            const char* addr = reinterpret_cast<const char*>(sc.nodes + (t.ng_base + rel));
            if (t.ng_mask & 0xff000000u) stack_push<VSB_Q_SHARED_STACK>(t, t.ng_base, t.ng_mask);
            NodeRegs n;
            load_record(addr, 5, n);
            uint32_t ph;
            const uint32_t nh = test_node_lut(n, rc, limit, ph, perm_lut);
            t.ng_base = n.w[4]; t.ng_mask = nh | (n.w[3] >> 24);
            t.pg_base = n.w[5]; t.pg_mask = ph; t.pg_valid = n.w[6];
            if ((t.ng_mask >> 24) == 0u)
            {
                if (t.sp == 0) t.ng_mask = 0u;
                else stack_pop<VSB_Q_SHARED_STACK>(t, t.ng_base, t.ng_mask);
            }
        }

        bool drained = false;
        if constexpr (Program::GROUP_RING)
        {
            // ---- append one GROUP per lane that selected primitives (no per-primitive loop, no scan), drain while worthwhile ----
            {
                const bool mine = state == LANE_TRAVERSING && t.pg_mask != 0u;
                const unsigned has = __ballot_sync(0xffffffffu, mine);
                if (has != 0u)
                {
                    while (q_tail - q_head > VSB_QCAP / 2 - 32u) { drain(); drained = true; }       // room for 32 groups
                    if (mine)
                    {
                        gring[(q_tail + (unsigned)__popc(has & lt_mask)) & (VSB_QCAP / 2 - 1u)] = make_uint4(t.pg_base, t.pg_mask, threadIdx.x, t.pg_valid);
#if VSB_Q_PREFETCH
                        // the group's records are tested a drain later: start their way up the hierarchy now
                        const char* a = reinterpret_cast<const char*>(sc.prims) + (size_t)prim_record(t.pg_base, t.pg_valid, __ffs((int)t.pg_mask) - 1) * (size_t)(n16 * 16);
                        const char* e = reinterpret_cast<const char*>(sc.prims) + (size_t)(prim_record(t.pg_base, t.pg_valid, find_msb(t.pg_mask)) + 1u) * (size_t)(n16 * 16) - 1;
#if VSB_Q_PREFETCH == 1
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(a));
                        if (((uintptr_t)a >> 7) != ((uintptr_t)e >> 7)) asm volatile("prefetch.global.L1 [%0];" :: "l"(e));
#else
                        asm volatile("prefetch.global.L2 [%0];" :: "l"(a));
                        if (((uintptr_t)a >> 7) != ((uintptr_t)e >> 7)) asm volatile("prefetch.global.L2 [%0];" :: "l"(e));
#endif
#endif
                    }
                    q_pairs += __reduce_add_sync(0xffffffffu, mine ? (unsigned)__popc(t.pg_mask) : 0u);
                    t.pg_mask = 0u;
                    q_tail += (unsigned)__popc(has);
                    __syncwarp();
                }
                if (state == LANE_TRAVERSING && (t.ng_mask >> 24) == 0u) { state = LANE_PENDING; my_seq = q_tail; }
                while (q_pairs != 0u)
                {
                    const unsigned busy = __ballot_sync(0xffffffffu, state == LANE_TRAVERSING);
                    const int n_pend = __popc(__ballot_sync(0xffffffffu, state == LANE_PENDING && (int)(q_head - my_seq) < 0));
                    if (!(q_pairs >= (unsigned)tl.drain_min || busy == 0u || n_pend >= tl.pend_min)) break;
                    drain(); drained = true;
                }
            }
        }
        else
        {
            // ---- append the selected primitives to the warp's ring, drain while worthwhile (ONE drain site) ----
            for (;;)
            {
                const int c = state == LANE_TRAVERSING ? __popc(t.pg_mask) : 0;
                const unsigned has = __ballot_sync(0xffffffffu, c != 0);
                if (has != 0u)
                {
                    unsigned off, total;
                    if (__ballot_sync(0xffffffffu, c > 1) == 0u)
                    {
                        off = (unsigned)__popc(has & lt_mask); total = (unsigned)__popc(has);     // one pair per lane: no scan
                    }
                    else
                    {
                        int incl = c;
    #pragma unroll
                        for (int o = 1; o < 32; o <<= 1)
                        {
                            const int v = __shfl_up_sync(0xffffffffu, incl, o);
                            if ((int)lane >= o) incl += v;
                        }
                        total = (unsigned)__shfl_sync(0xffffffffu, incl, 31);
                        off = (unsigned)(incl - c);
                    }
                    const unsigned space = VSB_QCAP - (q_tail - q_head);
                    while (t.pg_mask != 0u && off < space)
                    {
                        const int bit = find_msb(t.pg_mask);
                        t.pg_mask &= ~(1u << bit);
                        ring[(q_tail + off) & (VSB_QCAP - 1u)] = make_uint2(prim_record(t.pg_base, t.pg_valid, bit), threadIdx.x);
                        ++off;
                    }
                    q_tail += total < space ? total : space;
                    __syncwarp();
                }
                // out of node work and everything appended: wait for the ring's head to pass this lane's last pair
                if (state == LANE_TRAVERSING && t.pg_mask == 0u && (t.ng_mask >> 24) == 0u) { state = LANE_PENDING; my_seq = q_tail; }
                const unsigned count = q_tail - q_head;
                if (count == 0u) break;
                const unsigned busy = __ballot_sync(0xffffffffu, state == LANE_TRAVERSING);                    // node work (or pairs) left
                const unsigned left = __ballot_sync(0xffffffffu, state == LANE_TRAVERSING && t.pg_mask != 0u); // the ring was full
                const int n_pend = __popc(__ballot_sync(0xffffffffu, state == LANE_PENDING && (int)(q_head - my_seq) < 0));
                if (!(count >= (unsigned)tl.drain_min || left != 0u || busy == 0u || n_pend >= tl.pend_min)) break;
                drain(); drained = true;
            }
        }
        if (drained && (state == LANE_TRAVERSING || state == LANE_PENDING))
        {
            const unsigned long long k = key_s[threadIdx.x];
            limit = __uint_as_float((uint32_t)(k >> 32));
            if (any && live && (uint32_t)k != VSB_KEY_NONE) state = LANE_WAITING;
        }
        if (state == LANE_PENDING && (int)(q_head - my_seq) >= 0) state = LANE_WAITING;
    }

    if (tl.ray_count)
    {
        unsigned n = n_rays;
        for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
        if (lane == 0 && n) atomicAdd(tl.ray_count, (unsigned long long)n);
    }
}

// ---- ray list program: closest_hit()/any_hit() over a ray list ----------------------------------
// Sort key of a ray for the coherence pre-pass of ray lists: a 21-bit Morton code of the origin
// inside the scene bounds (7 bits per axis, major) and the direction octant (3 bits, minor).  Rays
// that are neighbours after the sort start close together and head the same way, so the lanes of a
// warp walk the same part of the tree (L1/L2 hits instead of HBM sector misses).  Origin-major
// measured 10 % faster on the random-ray stress than octant-major (short mean free path: where a
// ray starts decides what it touches); 24 key bits = 3 radix passes.
__device__ __forceinline__ uint32_t morton_spread8(uint32_t v)
{
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

#ifndef VSB_SORT_AXIS_BITS
#define VSB_SORT_AXIS_BITS 7        // Morton bits per axis of the origin cell (21 + 3 octant bits = 3 radix passes)
#endif
#ifndef VSB_SORT_OCT_LOW
#define VSB_SORT_OCT_LOW 1          // 0: direction octant is the major key; 1: the origin cell is (measured +10 % on cfg5)
#endif
#ifndef VSB_SORT_NO_OCT
#define VSB_SORT_NO_OCT 0
#endif
#define VSB_SORT_KEY_BITS (3 * VSB_SORT_AXIS_BITS + (VSB_SORT_NO_OCT ? 0 : 3))

__global__ void ray_key_kernel(const float4* __restrict__ rays, uint32_t n, float3 lo, float3 inv_extent,
                               uint32_t* __restrict__ keys, uint32_t* __restrict__ index)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 o = __ldg(rays + 2 * (size_t)i), d = __ldg(rays + 2 * (size_t)i + 1);
    const float fx = fminf(fmaxf((o.x - lo.x) * inv_extent.x, 0.0f), 0.999f);
    const float fy = fminf(fmaxf((o.y - lo.y) * inv_extent.y, 0.0f), 0.999f);
    const float fz = fminf(fmaxf((o.z - lo.z) * inv_extent.z, 0.0f), 0.999f);
    const float cells = (float)(1 << VSB_SORT_AXIS_BITS);
    const uint32_t qx = (uint32_t)(fx * cells), qy = (uint32_t)(fy * cells), qz = (uint32_t)(fz * cells);
    const uint32_t oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    const uint32_t cell = morton_spread8(qx) | (morton_spread8(qy) << 1) | (morton_spread8(qz) << 2);
    keys[i] = VSB_SORT_NO_OCT ? cell : VSB_SORT_OCT_LOW ? ((cell << 3) | oct) : ((oct << (3 * VSB_SORT_AXIS_BITS)) | cell);
    index[i] = i;
}

struct RayListProgram
{
    const uint32_t* order;         // optional: item -> ray index (coherence sort); NULL = identity
    const float4* rays;
    const float*  max_t_per_ray;
    float         max_t;
    uint4*        hits;
    int           any;

    struct Lane
    {
        unsigned long long item;
        __device__ Lane(uint32_t*, unsigned) : item(0) {}
    };
    static constexpr int STATE_WORDS = 0;
#ifndef VSB_LIST_MIN_BLOCKS
#define VSB_LIST_MIN_BLOCKS VSB_MIN_BLOCKS
#endif
    static constexpr int MIN_BLOCKS = VSB_LIST_MIN_BLOCKS;
#ifndef VSB_LIST_MIN_BLOCKS_Q
#define VSB_LIST_MIN_BLOCKS_Q 8     // 64 registers, 32 warps per SM: measured +2 % on the incoherent ray lists (frames: 6 is better)
#endif
    static constexpr int MIN_BLOCKS_Q = VSB_LIST_MIN_BLOCKS_Q;
    static constexpr bool GROUP_RING = VSB_GROUP_RING_ALL >= 0;    // group ring: closest-hit +5 %, any-hit +1 % (records addressed by rank since the `valid` node word)

    __device__ bool begin(Lane& L, unsigned long long item, RayReq& req) const
    {
        if (order) item = __ldg(order + item);
        L.item = item;
        const float4 o = __ldg(rays + 2 * item), d = __ldg(rays + 2 * item + 1);
        req.ox = o.x; req.oy = o.y; req.oz = o.z; req.dx = d.x; req.dy = d.y; req.dz = d.z;
        req.max_t = max_t_per_ray ? __ldg(max_t_per_ray + item) : max_t;
        req.any = any != 0;
        return true;
    }
    __device__ bool done(Lane& L, const HitReg& h, float, float, float, float, float, float, RayReq&) const
    {
        uint4 a, b;
        a.x = (uint32_t)h.hit; a.y = h.prim_id; a.z = h.geom_id; a.w = __float_as_uint(h.t);
        b.x = __float_as_uint(h.u); b.y = __float_as_uint(h.v); b.z = h.list_index; b.w = 0u;
        hits[2 * L.item] = a;
        hits[2 * L.item + 1] = b;
        return false;
    }
};

// gathered: [rank][tiles_per_rank_padded][64*64] float4 -> image[y*width + x]
__global__ void deinterleave_kernel(const float4* __restrict__ gathered, int width, int height, int nranks,
                                    int tiles_per_rank_padded, float4* image)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= width || y >= height) return;
    const int tiles_x = (width + VSB_TILE - 1) / VSB_TILE;
    const int tile = (y / VSB_TILE) * tiles_x + (x / VSB_TILE);
    const int rank = tile % nranks, local_tile = tile / nranks;
    const size_t src = ((size_t)rank * tiles_per_rank_padded + local_tile) * (VSB_TILE * VSB_TILE)
                     + (size_t)(y % VSB_TILE) * VSB_TILE + (size_t)(x % VSB_TILE);
    image[(size_t)y * width + x] = gathered[src];
}

} // namespace vsb

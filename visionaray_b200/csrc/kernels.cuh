// kernels.cuh -- the __global__ kernels of libvsnray_b200.so (sm_100a).
//
//  encode_level_kernel   BVH2 -> BVH8 collapse, one thread per wide node of the current level
//  fill_prims_kernel     leaf-order primitive records (removes the indices[] indirection)
//  query_kernel          persistent warps over a ray list: closest_hit / any_hit
//  frame_kernel          persistent warps over 8x4-pixel packets: ray gen + kernel + store/blend
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

// counters[0] = next free wide node, [1] = next free primitive slot, [2] = tasks emitted
__global__ void encode_level_kernel(const WideTask* __restrict__ tasks, uint32_t n_tasks,
                                    const Node32* __restrict__ nodes2, Node80* out_nodes,
                                    uint32_t* leaf_list, WideTask* next_tasks, uint32_t* counters)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_tasks) return;
    DeviceAlloc an{ counters + 0 }, ap{ counters + 1 }, at{ counters + 2 };
    collapse_node(tasks[i], nodes2, out_nodes, leaf_list, next_tasks, an, ap, at);
}

__global__ void fill_prims_kernel(int prim_kind, const void* __restrict__ prims, const uint32_t* __restrict__ indices,
                                  const uint32_t* __restrict__ leaf_list, uint32_t n, void* out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    make_prim_record(prim_kind, prims, indices, leaf_list[i], out, i);
}

// ---- ray list ----------------------------------------------------------------------------------

template <bool ANY, bool EXACT>
__device__ __forceinline__ void trace_one(const DeviceScene& s, float ox, float oy, float oz,
                                          float dx, float dy, float dz, float max_t, HitReg& h)
{
    if (EXACT)
    {
        trace_exact<ANY>(s.exact, ox, oy, oz, dx, dy, dz, max_t, h);
    }
    else
    {
        RayCtx rc;
        ray_ctx_init(rc, ox, oy, oz, dx, dy, dz);
        trace_wide<ANY>(s.wide, rc, max_t, h);
    }
}

// is_closer + update_if merge over the range of bvhs (detail/traverse_linear.inl:106-127):
// each bvh is traversed from scratch, then merged with strict '<'.
template <bool ANY, bool EXACT>
__device__ __forceinline__ void trace_scenes(const DeviceScene* __restrict__ scenes, int n_scenes,
                                             float ox, float oy, float oz, float dx, float dy, float dz,
                                             float max_t, HitReg& best)
{
    trace_one<ANY, EXACT>(scenes[0], ox, oy, oz, dx, dy, dz, max_t, best);
    for (int k = 1; k < n_scenes; ++k)
    {
        if (ANY && best.hit) return;
        HitReg h;
        trace_one<ANY, EXACT>(scenes[k], ox, oy, oz, dx, dy, dz, max_t, h);
        if (h.hit && h.t >= 0.0f && h.t < best.t && h.t < max_t) best = h;
    }
}

template <bool ANY, bool EXACT>
__global__ void __launch_bounds__(128)
query_kernel(const DeviceScene* __restrict__ scenes, int n_scenes, const float4* __restrict__ rays,
             unsigned long long n_rays, float max_t, const float* __restrict__ max_t_per_ray,
             uint4* __restrict__ hits, unsigned long long* work_counter)
{
    const unsigned lane = threadIdx.x & 31u;
    for (;;)
    {
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(work_counter, 32ull);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n_rays) break;
        const unsigned long long i = base + lane;
        if (i < n_rays)
        {
            const float4 o = __ldg(rays + 2 * i), d = __ldg(rays + 2 * i + 1);
            const float mt = max_t_per_ray ? __ldg(max_t_per_ray + i) : max_t;
            HitReg h;
            trace_scenes<ANY, EXACT>(scenes, n_scenes, o.x, o.y, o.z, d.x, d.y, d.z, mt, h);
            uint4 a, b;
            a.x = (uint32_t)h.hit; a.y = h.prim_id; a.z = h.geom_id; a.w = __float_as_uint(h.t);
            b.x = __float_as_uint(h.u); b.y = __float_as_uint(h.v); b.z = h.list_index; b.w = 0u;
            hits[2 * i] = a;
            hits[2 * i + 1] = b;
        }
        __syncwarp();
    }
}

// ---- frame -------------------------------------------------------------------------------------

struct FrameLaunch
{
    FrameCtx ctx;
    int scissor[4];
    float4* color;
    int tile_rank, tile_nranks;
    int tiles_x, tiles_y;            // 64x64 tiles
    int n_owned_tiles;               // tiles this rank renders
    unsigned long long* ray_count;   // may be NULL
    const DeviceScene* scenes;
    int n_scenes;
};

template <bool EXACT>
struct Tracer
{
    const DeviceScene* scenes;
    int n_scenes;
    unsigned n_rays;
    __device__ void closest(V3 o, V3 d, HitReg& h)
    {
        ++n_rays;
        trace_scenes<false, EXACT>(scenes, n_scenes, o.x, o.y, o.z, d.x, d.y, d.z, FLT_MAX, h);
    }
    __device__ bool occluded(V3 o, V3 d, float max_t)
    {
        ++n_rays;
        HitReg h;
        trace_scenes<true, EXACT>(scenes, n_scenes, o.x, o.y, o.z, d.x, d.y, d.z, max_t, h);
        return h.hit != 0;
    }
};

#define VSB_TILE 64
#define VSB_PACKETS_PER_TILE ((VSB_TILE / 8) * (VSB_TILE / 4))

template <int KERNEL, bool EXACT>
__global__ void __launch_bounds__(128)
frame_kernel(const FrameLaunch fl, unsigned long long* work_counter)
{
    const unsigned lane = threadIdx.x & 31u;
    const FrameCtx& f = fl.ctx;
    Tracer<EXACT> tr{ fl.scenes, fl.n_scenes, 0u };
    const unsigned long long n_packets = (unsigned long long)fl.n_owned_tiles * VSB_PACKETS_PER_TILE;
    const int nr = fl.tile_nranks > 1 ? fl.tile_nranks : 1;
    const int rk = fl.tile_nranks > 1 ? fl.tile_rank : 0;
    for (;;)
    {
        unsigned long long pk = 0;
        if (lane == 0) pk = atomicAdd(work_counter, 1ull);
        pk = __shfl_sync(0xffffffffu, pk, 0);
        if (pk >= n_packets) break;
        const int local_tile = (int)(pk / VSB_PACKETS_PER_TILE);
        const int in_tile = (int)(pk % VSB_PACKETS_PER_TILE);
        const int tile = local_tile * nr + rk;                   // interleaved: tile k -> rank k % nranks
        const int tx = tile % fl.tiles_x, ty = tile / fl.tiles_x;
        const int px = in_tile % (VSB_TILE / 8), py = in_tile / (VSB_TILE / 8);
        const int lx = px * 8 + (int)(lane & 7u), ly = py * 4 + (int)(lane >> 3);
        const int x = tx * VSB_TILE + lx, y = ty * VSB_TILE + ly;
        const bool inside = x < f.p.width && y < f.p.height &&
                            x >= fl.scissor[0] && y >= fl.scissor[1] &&
                            x < fl.scissor[0] + fl.scissor[2] && y < fl.scissor[1] + fl.scissor[3];
        if (inside)
        {
            const uint32_t pixel = (uint32_t)y * (uint32_t)f.p.width + (uint32_t)x;
            uint32_t samp = sampler_init(pixel, f.p.frame_num, f.p.seed);
            float fx = (float)x, fy = (float)y;
            if (f.p.pixel_sampler != 0)
            {
                // jittered (detail/sched_common.h:197-217)
                const float jx = rn_sub(sampler_next(samp), 0.5f);
                const float jy = rn_sub(sampler_next(samp), 0.5f);
                fx = rn_add(fx, jx); fy = rn_add(fy, jy);
            }
            V3 o, d;
            primary_ray(f, fx, fy, o, d);
            float col[4];
            if (KERNEL == 0)      kernel_simple(f, tr, o, d, col);
            else if (KERNEL == 1) kernel_whitted(f, tr, o, d, col);
            else if (KERNEL == 2) kernel_ao(f, tr, o, d, samp, col);
            else                  kernel_pathtracing(f, tr, o, d, samp, col);

            const size_t dst = fl.tile_nranks > 1
                ? (size_t)local_tile * (VSB_TILE * VSB_TILE) + (size_t)ly * VSB_TILE + (size_t)lx
                : (size_t)pixel;
            float4 c = make_float4(col[0], col[1], col[2], col[3]);
            if (f.p.pixel_sampler == 2)
            {
                // jittered_blend (detail/sched_common.h:516-528)
                const float alpha = rn_div(1.0f, (float)f.p.frame_num);
                const float ia = rn_sub(1.0f, alpha);
                const float4 old = fl.color[dst];
                c.x = rn_add(rn_mul(c.x, alpha), rn_mul(old.x, ia));
                c.y = rn_add(rn_mul(c.y, alpha), rn_mul(old.y, ia));
                c.z = rn_add(rn_mul(c.z, alpha), rn_mul(old.z, ia));
                c.w = rn_add(rn_mul(c.w, alpha), rn_mul(old.w, ia));
            }
            fl.color[dst] = c;
        }
        __syncwarp();
    }
    if (fl.ray_count)
    {
        unsigned n = tr.n_rays;
        for (int o = 16; o > 0; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
        if (lane == 0 && n) atomicAdd(fl.ray_count, (unsigned long long)n);
    }
}

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

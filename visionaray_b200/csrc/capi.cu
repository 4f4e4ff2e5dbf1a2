// capi.cu -- the C ABI of libvsnray_b200.so (include/vsnray_b200.h) over the sm_100a kernels.
// No CPU fallback exists in this library: without a CUDA device vsb_init() fails and every
// compute entry point returns VSB_ERR_NOT_INIT.

#include <cstdio>
#include <cstring>
#include <cmath>
#include <utility>
#include <vector>

#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include "../../include/vsnray_b200.h"
#include "kernels.cuh"
#include "build_ploc.cuh"
#include "build_sah.cuh"
#include "multi_hit.cuh"
#include "watertight.cuh"

#include "capi_scene.inl"
#include "capi_build.inl"

namespace
{

static_assert(sizeof(vsb_ray) == 32 && sizeof(vsb_hit) == 32 && sizeof(vsb_bvh_node) == 32, "ABI layouts");
static_assert(sizeof(vsb_triangle) == 64 && sizeof(vsb_sphere) == 48, "ABI layouts");
static_assert(sizeof(vsb_material) == 64 && sizeof(vsb_point_light) == 64, "ABI layouts");
static_assert(sizeof(vsb_frame_params) == sizeof(vsb::FrameParams), "frame params layout");
static_assert(sizeof(vsb::Material) == 64 && sizeof(vsb::PointLight) == 64, "device layouts");
static_assert(sizeof(vsb::Node80) == 80 && sizeof(vsb::TriRec) == 48 && sizeof(vsb::SphRec) == 32, "wide layouts");

const int BLOCK = 128;

int upload_scene_table(const vsb_scene* const* scenes, int n_scenes, cudaStream_t st)
{
    if (!scenes || n_scenes < 1 || n_scenes > MAX_SCENES) return VSB_ERR_BAD_ARG;
    vsb::DeviceScene tab[MAX_SCENES];
    for (int i = 0; i < n_scenes; ++i)
    {
        if (!scenes[i]) return VSB_ERR_BAD_ARG;
        tab[i] = scenes[i]->dev;
    }
    VSB_CUDA(cudaMemcpyAsync(g.d_scenes, tab, sizeof(vsb::DeviceScene) * n_scenes, cudaMemcpyHostToDevice, st));
    return VSB_OK;
}

// the reference-order walk keeps its stack in VSB_EXACT_STACK entries; a GPU-built tree may be deeper
bool exact_supported(const vsb_scene* const* scenes, int n_scenes)
{
    for (int i = 0; i < n_scenes; ++i)
        if (scenes[i]->depth2 > VSB_EXACT_STACK - 1) return false;
    return true;
}

// grid = resident CTAs per SM x SM count (persistent warps pull work from the queue), capped by the
// number of warps the work can feed
template <typename K>
int grid_for(K kernel, size_t smem, unsigned long long items, unsigned chunk)
{
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, BLOCK, smem) != cudaSuccess || per_sm < 1) per_sm = 4;
    const unsigned long long chunks = (items + chunk - 1) / chunk;
    unsigned long long want = (chunks + (BLOCK / 32) - 1) / (BLOCK / 32);
    unsigned long long full = (unsigned long long)per_sm * (unsigned long long)g.sm_count;
    unsigned long long grid = want < full ? want : full;
    return (int)(grid < 1 ? 1 : grid);
}

// tuning knobs of the persistent loop (kernels.cuh), settable through vsb_tune():
//   query_service_min / frame_service_min  lanes that must be idle before continuations + refill run
//   phase_min                              phase vote threshold
// frame_service_min 0 = per kernel: 32 (warp-wide lockstep; measured best for the coherent simple /
// whitted frames) or 24 (AO and path tracing, whose secondary rays decorrelate the lanes; gpurun_out/r2_ab8.log)
int g_query_service_min = 8, g_frame_service_min = 0, g_phase_min = 1;     // ray lists: 8 (gpurun_out/r2_ab9.log: 2 / 4 / 8 / 16 = 1156 / 1217 / 1244 / 1196 Mrays/s)
// sort_rays: 1 = ray lists of >= 64 Ki rays are traced in coherence-sorted order (wide path), 0 = never
int g_sort_rays = 1;
// frame_bands: a host render target is traced in this many bands so that the D2H copy of one band
// overlaps the tracing of the next (1 = off)
int g_host_store = 0;      // 1: frames into PINNED host targets are stored by the trace kernel itself (mapped memory, no staging / D2H pass)
int g_list_piece_krays = 1024;   // host ray lists are staged, traced and read back in pieces of this many Ki rays
int g_band_min_mib = 16;   // a band carries at least this much of the target (per-launch ramps vs overlap)
int g_band_trace = 0;      // developer switch: print the band timeline of host-target frames to stderr
int g_band_streams = 2;    // 2: alternate bands over two streams (tail of band i overlaps the head of band i+1)
int g_frame_bands = 12;    // cfg2 into pinned host memory: 1 band 11.8 ms, 2: 9.6 (round 1); 8 / 12 / 16 / 24 interleaved bands on two streams: 5.90 / 5.72 / 5.82 / 6.02 ms

// trace_queue: 1 = single-bvh BVH8 launches run trace_kernel_q (primitive tests shared by the warp), 0 = trace_kernel
int g_trace_queue = 1, g_drain_min = 24, g_pend_min = 8;     // drain_min 16..24 measured best (cfg5 +3 % over 32)

template <typename Program, bool EXACT>
cudaError_t launch_trace(const Program& prog, vsb::TraceLaunch tl, cudaStream_t st)
{
    tl.drain_min = g_drain_min < 1 ? 1 : (g_drain_min > 32 ? 32 : g_drain_min);
    tl.pend_min = g_pend_min < 1 ? 1 : g_pend_min;
    if (!EXACT && g_trace_queue && tl.n_scenes == 1)
    {
        auto kernel = vsb::trace_kernel_q<Program>;
        const size_t smem = vsb::trace_q_smem_words<Program>() * sizeof(uint32_t);
        const int grid = grid_for(kernel, smem, tl.n_items, tl.chunk);
        kernel<<<grid, BLOCK, smem, st>>>(tl, prog);
        return cudaGetLastError();
    }
    auto kernel = vsb::trace_kernel<Program, EXACT>;
    const size_t smem = (size_t)(Program::STATE_WORDS + 2 * VSB_SHARED_STACK) * BLOCK * sizeof(uint32_t);
    const int grid = grid_for(kernel, smem, tl.n_items, tl.chunk);
    kernel<<<grid, BLOCK, smem, st>>>(tl, prog);
    return cudaGetLastError();
}

int run_query(bool any, const vsb_scene* const* scenes, int n_scenes, const vsb_ray* rays, size_t n_rays,
              float max_t, const float* max_t_per_ray, vsb_hit* hits, int mem, int traverse, void* cuda_stream)
{
    int rc = require_init(); if (rc) return rc;
    if (n_rays == 0) return VSB_OK;
    if (!rays || !hits) return VSB_ERR_BAD_ARG;
    if (traverse != VSB_TRAVERSE_WIDE && traverse != VSB_TRAVERSE_EXACT && traverse != VSB_TRAVERSE_WATERTIGHT) return VSB_ERR_BAD_ARG;
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : g.stream;
    CallOrder order(st);
    rc = upload_scene_table(scenes, n_scenes, st); if (rc) return rc;

    if (mem != VSB_MEM_HOST && mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    if (traverse == VSB_TRAVERSE_WATERTIGHT)
    {
        // opt-in mode (csrc/watertight.cuh): one thread per ray, triangle scenes only
        for (int i = 0; i < n_scenes; ++i) if (scenes[i]->prim_kind != VSB_PRIM_TRIANGLE) return VSB_ERR_UNSUPPORTED;
        const vsb::Ray32* d_rays = reinterpret_cast<const vsb::Ray32*>(rays);
        vsb::Hit32* d_hits = reinterpret_cast<vsb::Hit32*>(hits);
        const float* d_mt = max_t_per_ray;
        if (mem == VSB_MEM_HOST)
        {
            void *a, *b, *c = nullptr;
            rc = stage(0, n_rays * sizeof(vsb_ray), &a); if (rc) return rc;
            rc = stage(1, n_rays * sizeof(vsb_hit), &b); if (rc) return rc;
            VSB_CUDA(cudaMemcpyAsync(a, rays, n_rays * sizeof(vsb_ray), cudaMemcpyHostToDevice, st));
            if (max_t_per_ray)
            {
                rc = stage(2, n_rays * sizeof(float), &c); if (rc) return rc;
                VSB_CUDA(cudaMemcpyAsync(c, max_t_per_ray, n_rays * sizeof(float), cudaMemcpyHostToDevice, st));
            }
            d_rays = reinterpret_cast<const vsb::Ray32*>(a); d_hits = reinterpret_cast<vsb::Hit32*>(b); d_mt = reinterpret_cast<const float*>(c);
        }
        const unsigned blocks = (unsigned)((n_rays + 127) / 128);
        VSB_CUDA(cudaEventRecord(g.ev0, st));
        if (any) vsb::watertight_kernel<true><<<blocks, 128, 0, st>>>(g.d_scenes, n_scenes, d_rays, n_rays, max_t, d_mt, d_hits);
        else     vsb::watertight_kernel<false><<<blocks, 128, 0, st>>>(g.d_scenes, n_scenes, d_rays, n_rays, max_t, d_mt, d_hits);
        ++g.launches;
        VSB_CUDA(cudaGetLastError());
        VSB_CUDA(cudaEventRecord(g.ev1, st));
        if (mem == VSB_MEM_HOST)
        {
            VSB_CUDA(cudaMemcpyAsync(hits, d_hits, n_rays * sizeof(vsb_hit), cudaMemcpyDeviceToHost, st));
            VSB_CUDA(cudaStreamSynchronize(st));
        }
        return VSB_OK;
    }
    const bool exact = traverse == VSB_TRAVERSE_EXACT;
    if (exact && !exact_supported(scenes, n_scenes)) return VSB_ERR_UNSUPPORTED;

    // sort (optional) + trace of one contiguous range of the list, all on `st`
    auto trace_range = [&](const float4* d_rays, uint4* d_hits, const float* d_mt, size_t n_range, unsigned long long* counter,
                           cudaStream_t st, int slot_off) -> int
    {
        // coherence sort (order only: rays are gathered and hits scattered by index, so results land in
        // the caller's order)
        const uint32_t* d_order = nullptr;
        if (g_sort_rays && !exact && n_range >= (1u << 16) && n_range < 0x7fffffffull)
        {
            void *k0, *k1, *i0, *i1, *tmp;
            const uint32_t n = (uint32_t)n_range;
            int rc2;
            rc2 = stage(6 + slot_off, (size_t)n * 4, &k0); if (rc2) return rc2;
            rc2 = stage(7 + slot_off, (size_t)n * 4, &k1); if (rc2) return rc2;
            rc2 = stage(8 + slot_off, (size_t)n * 4, &i0); if (rc2) return rc2;
            rc2 = stage(9 + slot_off, (size_t)n * 4, &i1); if (rc2) return rc2;
            size_t tmp_bytes = 0;
            VSB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint32_t*)k0, (uint32_t*)k1, (const uint32_t*)i0, (uint32_t*)i1, (int)n, 0, VSB_SORT_KEY_BITS, st));
            rc2 = stage(10 + slot_off, tmp_bytes, &tmp); if (rc2) return rc2;
            const vsb_scene* s0 = scenes[0];
            float3 lo = make_float3(s0->lo[0], s0->lo[1], s0->lo[2]);
            float3 inv = make_float3(1.0f / fmaxf(s0->hi[0] - s0->lo[0], 1e-30f), 1.0f / fmaxf(s0->hi[1] - s0->lo[1], 1e-30f),
                                     1.0f / fmaxf(s0->hi[2] - s0->lo[2], 1e-30f));
            vsb::ray_key_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_rays, n, lo, inv, (uint32_t*)k0, (uint32_t*)i0);
            VSB_CUDA(cudaGetLastError());
            VSB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, (const uint32_t*)k0, (uint32_t*)k1, (const uint32_t*)i0, (uint32_t*)i1, (int)n, 0, VSB_SORT_KEY_BITS, st));
            g.launches += 2;
            d_order = (const uint32_t*)i1;
        }
        vsb::RayListProgram prog;
        prog.order = d_order;
        prog.rays = d_rays; prog.max_t_per_ray = d_mt; prog.max_t = max_t; prog.hits = d_hits; prog.any = any ? 1 : 0;
        vsb::TraceLaunch tl;
        tl.scenes = g.d_scenes; tl.n_scenes = n_scenes; tl.n_items = (unsigned long long)n_range;
        tl.wide0 = scenes[0]->dev.wide;
        tl.work_counter = counter; tl.ray_count = nullptr;
        tl.service_min = g_query_service_min; tl.phase_min = g_phase_min;
        {
            // chunk: ~16 reservations per resident warp, 32..256 rays
            unsigned long long per = (unsigned long long)n_range / ((unsigned long long)g.sm_count * 32ull * 16ull);
            per = (per / 32ull) * 32ull;
            tl.chunk = (unsigned)(per < 32ull ? 32ull : (per > 256ull ? 256ull : per));
        }
        if (exact) VSB_CUDA((launch_trace<vsb::RayListProgram, true>(prog, tl, st)));
        else       VSB_CUDA((launch_trace<vsb::RayListProgram, false>(prog, tl, st)));
        ++g.launches;
        VSB_CUDA(cudaGetLastError());
        return VSB_OK;
    };

    VSB_CUDA(cudaMemsetAsync(g.d_work, 0, 64 * sizeof(unsigned long long), st));
    if (mem == VSB_MEM_DEVICE)
    {
        VSB_CUDA(cudaEventRecord(g.ev0, st));
        rc = trace_range(reinterpret_cast<const float4*>(rays), reinterpret_cast<uint4*>(hits), max_t_per_ray, n_rays, g.d_work, st, 0); if (rc) return rc;
        VSB_CUDA(cudaEventRecord(g.ev1, st));
        return VSB_OK;
    }

    // Host lists: staged in pieces.  The rays of piece i+1 go up (H2D engine, upload stream) and the
    // hits of piece i-1 come down (D2H engine, copy stream) while piece i is traced.
    void *a, *b, *c = nullptr;
    rc = stage(0, n_rays * sizeof(vsb_ray), &a); if (rc) return rc;
    rc = stage(1, n_rays * sizeof(vsb_hit), &b); if (rc) return rc;
    if (max_t_per_ray) { rc = stage(2, n_rays * sizeof(float), &c); if (rc) return rc; }
    size_t piece = (size_t)(g_list_piece_krays < 64 ? 64 : g_list_piece_krays) << 10;
    if ((n_rays + piece - 1) / piece > 32) piece = (((n_rays + 31) / 32 + 65535) >> 16) << 16;
    const int n_pieces = (int)((n_rays + piece - 1) / piece);
    VSB_CUDA(cudaEventRecord(g.band_fork, st));
    VSB_CUDA(cudaStreamWaitEvent(g.upload_stream, g.band_fork, 0));
    for (int i = 0; i < n_pieces; ++i)
    {
        const size_t r0 = (size_t)i * piece, n = r0 + piece < n_rays ? piece : n_rays - r0;
        VSB_CUDA(cudaMemcpyAsync((char*)a + r0 * sizeof(vsb_ray), rays + r0, n * sizeof(vsb_ray), cudaMemcpyHostToDevice, g.upload_stream));
        if (max_t_per_ray) VSB_CUDA(cudaMemcpyAsync((char*)c + r0 * sizeof(float), max_t_per_ray + r0, n * sizeof(float), cudaMemcpyHostToDevice, g.upload_stream));
        VSB_CUDA(cudaEventRecord(g.band_up[i], g.upload_stream));
    }
    VSB_CUDA(cudaEventRecord(g.ev0, st));
    // pieces alternate between two streams (own sort buffers each): the persistent CTAs of piece i+1
    // are dispatched as those of piece i retire
    const bool two_streams = n_pieces > 1 && g_band_streams > 1;
    if (two_streams) VSB_CUDA(cudaStreamWaitEvent(g.band_stream, g.band_fork, 0));
    for (int i = 0; i < n_pieces; ++i)
    {
        const size_t r0 = (size_t)i * piece, n = r0 + piece < n_rays ? piece : n_rays - r0;
        cudaStream_t ps = (two_streams && (i & 1)) ? g.band_stream : st;
        VSB_CUDA(cudaStreamWaitEvent(ps, g.band_up[i], 0));
        rc = trace_range(reinterpret_cast<const float4*>(a) + 2 * r0, reinterpret_cast<uint4*>(b) + 2 * r0,
                         c ? reinterpret_cast<const float*>(c) + r0 : nullptr, n, g.d_work + 2 + i, ps, (two_streams && (i & 1)) ? 6 : 0);
        if (rc) return rc;
        VSB_CUDA(cudaEventRecord(g.band_done[i], ps));
        VSB_CUDA(cudaStreamWaitEvent(g.copy_stream, g.band_done[i], 0));
        VSB_CUDA(cudaMemcpyAsync(hits + r0, (char*)b + r0 * sizeof(vsb_hit), n * sizeof(vsb_hit), cudaMemcpyDeviceToHost, g.copy_stream));
    }
    if (two_streams)
    {
        VSB_CUDA(cudaEventRecord(g.band_join, g.band_stream));
        VSB_CUDA(cudaStreamWaitEvent(st, g.band_join, 0));
    }
    VSB_CUDA(cudaEventRecord(g.ev1, st));
    VSB_CUDA(cudaStreamSynchronize(st));
    VSB_CUDA(cudaStreamSynchronize(g.copy_stream));
    return VSB_OK;
}

} // namespace

#include "capi_noalloc.inl"

extern "C" int vsb_version(void) { return 100; }

extern "C" int vsb_tune(const char* key, int value)
{
    if (!key || value < 0 || value > (1 << 20)) return VSB_ERR_BAD_ARG;
    if (!std::strcmp(key, "query_service_min")) g_query_service_min = value;
    else if (!std::strcmp(key, "frame_service_min")) g_frame_service_min = value;
    else if (!std::strcmp(key, "phase_min")) g_phase_min = value;
    else if (!std::strcmp(key, "trace_queue")) g_trace_queue = value;
    else if (!std::strcmp(key, "drain_min")) g_drain_min = value;
    else if (!std::strcmp(key, "pend_min")) g_pend_min = value;
    else if (!std::strcmp(key, "sort_rays")) g_sort_rays = value;
    else if (!std::strcmp(key, "frame_bands")) g_frame_bands = value;
    else if (!std::strcmp(key, "host_store")) g_host_store = value;
    else if (!std::strcmp(key, "band_streams")) g_band_streams = value;
    else if (!std::strcmp(key, "band_trace")) g_band_trace = value;
    else if (!std::strcmp(key, "band_min_mib")) g_band_min_mib = value;
    else if (!std::strcmp(key, "list_piece_krays")) g_list_piece_krays = value;
    else if (!std::strcmp(key, "optimal_collapse")) g_optimal_collapse = value;
    else if (!std::strcmp(key, "build_algorithm")) { if (value != 0 && value != 1) return VSB_ERR_BAD_ARG; g_build_algorithm = value; }
    else if (!std::strcmp(key, "build_radius")) g_build_radius = value;
    else if (!std::strcmp(key, "build_cost_node_x100")) g_build_cost_node = 0.01f * (float)value;
    else if (!std::strcmp(key, "build_cost_prim_x100")) g_build_cost_prim = 0.01f * (float)value;
    else return VSB_ERR_BAD_ARG;
    return VSB_OK;
}

extern "C" const char* vsb_strerror(int code)
{
    switch (code)
    {
    case VSB_OK: return "ok";
    case VSB_ERR_NO_DEVICE: return "no CUDA device (this library has no CPU fallback)";
    case VSB_ERR_BAD_ARG: return "bad argument";
    case VSB_ERR_BAD_TREE: return "malformed or too deep BVH";
    case VSB_ERR_NOT_INIT: return "vsb_init() has not succeeded";
    case VSB_ERR_UNSUPPORTED: return "unsupported";
    case VSB_ERR_TOO_LARGE: return "size beyond the fixed maximum of this entry point (vsb_closest_hit_static: VSB_STATIC_MAX_RAYS rays)";
    default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
    }
}

extern "C" int vsb_closest_hit(const vsb_scene* const* scenes, int n_scenes, const vsb_ray* rays, size_t n_rays,
                               float max_t, const float* max_t_per_ray, vsb_hit* hits, int mem, int traverse, void* cuda_stream)
{
    return run_query(false, scenes, n_scenes, rays, n_rays, max_t, max_t_per_ray, hits, mem, traverse, cuda_stream);
}

extern "C" int vsb_any_hit(const vsb_scene* const* scenes, int n_scenes, const vsb_ray* rays, size_t n_rays,
                           float max_t, const float* max_t_per_ray, vsb_hit* hits, int mem, int traverse, void* cuda_stream)
{
    return run_query(true, scenes, n_scenes, rays, n_rays, max_t, max_t_per_ray, hits, mem, traverse, cuda_stream);
}

// multi_hit<N>() over a ray list: N records per ray, sorted by t (csrc/multi_hit.cuh)
extern "C" int vsb_multi_hit(const vsb_scene* const* scenes, int n_scenes, const vsb_ray* rays, size_t n_rays,
                             int max_hits, vsb_hit* hits, int mem, int traverse, void* cuda_stream)
{
    int rc = require_init(); if (rc) return rc;
    if (max_hits < 1 || max_hits > VSB_MULTI_HIT_MAX) return VSB_ERR_BAD_ARG;
    if (n_rays == 0) return VSB_OK;
    if (!rays || !hits) return VSB_ERR_BAD_ARG;
    if (traverse != VSB_TRAVERSE_WIDE && traverse != VSB_TRAVERSE_EXACT) return VSB_ERR_BAD_ARG;
    if (mem != VSB_MEM_HOST && mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : g.stream;
    CallOrder order(st);
    rc = upload_scene_table(scenes, n_scenes, st); if (rc) return rc;
    const bool exact = traverse == VSB_TRAVERSE_EXACT;
    if (exact && !exact_supported(scenes, n_scenes)) return VSB_ERR_UNSUPPORTED;
    const vsb::Ray32* d_rays = reinterpret_cast<const vsb::Ray32*>(rays);
    vsb::Hit32* d_hits = reinterpret_cast<vsb::Hit32*>(hits);
    const size_t hit_bytes = n_rays * (size_t)max_hits * sizeof(vsb_hit);
    if (mem == VSB_MEM_HOST)
    {
        void *a = nullptr, *b = nullptr;
        rc = stage(0, n_rays * sizeof(vsb_ray), &a); if (rc) return rc;
        rc = stage(1, hit_bytes, &b); if (rc) return rc;
        VSB_CUDA(cudaMemcpyAsync(a, rays, n_rays * sizeof(vsb_ray), cudaMemcpyHostToDevice, st));
        d_rays = reinterpret_cast<const vsb::Ray32*>(a); d_hits = reinterpret_cast<vsb::Hit32*>(b);
    }
    const unsigned blocks = (unsigned)((n_rays + 127) / 128);
    VSB_CUDA(cudaEventRecord(g.ev0, st));
    if (exact) vsb::multi_hit_kernel<true><<<blocks, 128, 0, st>>>(g.d_scenes, n_scenes, d_rays, n_rays, max_hits, d_hits);
    else       vsb::multi_hit_kernel<false><<<blocks, 128, 0, st>>>(g.d_scenes, n_scenes, d_rays, n_rays, max_hits, d_hits);
    ++g.launches;
    VSB_CUDA(cudaGetLastError());
    VSB_CUDA(cudaEventRecord(g.ev1, st));
    if (mem == VSB_MEM_HOST)
    {
        VSB_CUDA(cudaMemcpyAsync(hits, d_hits, hit_bytes, cudaMemcpyDeviceToHost, st));
        VSB_CUDA(cudaStreamSynchronize(st));
    }
    return VSB_OK;
}

// Host-side camera basis, exactly the arithmetic of cuda_sched_impl_frame
// (detail/cuda_sched.inl:259-266); this TU is built with -Xcompiler -ffp-contract=off.
extern "C" int vsb_camera_basis(vsb_frame_desc* d)
{
    if (!d) return VSB_ERR_BAD_ARG;
    const vsb_frame_params& p = d->p;
    auto norm = [](const float v[3], float o[3])
    {
        float s = 1.0f / std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        o[0] = v[0] * s; o[1] = v[1] * s; o[2] = v[2] * s;
    };
    auto cross = [](const float u[3], const float v[3], float o[3])
    {
        float x = u[1] * v[2] - u[2] * v[1], y = u[2] * v[0] - u[0] * v[2], z = u[0] * v[1] - u[1] * v[0];
        o[0] = x; o[1] = y; o[2] = z;
    };
    float dd[3] = { p.eye[0] - p.center[0], p.eye[1] - p.center[1], p.eye[2] - p.center[2] };
    float f[3], c[3], s[3], u[3];
    norm(dd, f);
    cross(p.up, f, c);
    norm(c, s);
    cross(f, s, u);
    const float th = std::tan(p.fovy / 2.0f);
    for (int i = 0; i < 3; ++i)
    {
        d->cam_eye[i] = p.eye[i];
        d->cam_u[i] = s[i] * th * p.aspect;
        d->cam_v[i] = u[i] * th;
        d->cam_w[i] = -f[i];
    }
    d->cam_eye[3] = d->cam_u[3] = d->cam_v[3] = d->cam_w[3] = 0.0f;
    return VSB_OK;
}

// inverse of a column-major 4x4 by cofactors of 2x2 sub-determinants, in float like the reference
// scheduler's inverse(sparams.view_matrix) (detail/cuda_sched.inl:231-234; -ffp-contract=off TU)
static bool mat4_inverse(const float* m, float* out)
{
    auto at = [&](int r, int c) { return m[c * 4 + r]; };
    auto d2 = [](float a, float b, float c, float d) { return a * d - c * b; };
    float s[6], c[6];
    const int pr[6][2] = { {0, 1}, {0, 2}, {0, 3}, {1, 2}, {1, 3}, {2, 3} };
    for (int k = 0; k < 6; ++k)
    {
        s[k] = d2(at(0, pr[k][0]), at(0, pr[k][1]), at(1, pr[k][0]), at(1, pr[k][1]));   // rows 0,1
        c[k] = d2(at(2, pr[k][0]), at(2, pr[k][1]), at(3, pr[k][0]), at(3, pr[k][1]));   // rows 2,3
    }
    const float det = s[0] * c[5] - s[1] * c[4] + s[2] * c[3] + s[3] * c[2] - s[4] * c[1] + s[5] * c[0];
    if (det == 0.0f || det != det) return false;
    // adjugate: out(col j of the inverse, stored contiguously) from rows of m
    out[0]  = (+at(1,1) * c[5] - at(1,2) * c[4] + at(1,3) * c[3]) / det;
    out[1]  = (-at(1,0) * c[5] + at(1,2) * c[2] + at(1,3) * c[1]) / det;
    out[2]  = (+at(1,0) * c[4] - at(1,1) * c[2] + at(1,3) * c[0]) / det;
    out[3]  = (-at(1,0) * c[3] + at(1,1) * c[1] + at(1,2) * c[0]) / det;
    out[4]  = (-at(0,1) * c[5] + at(0,2) * c[4] - at(0,3) * c[3]) / det;
    out[5]  = (+at(0,0) * c[5] - at(0,2) * c[2] + at(0,3) * c[1]) / det;
    out[6]  = (-at(0,0) * c[4] + at(0,1) * c[2] - at(0,3) * c[0]) / det;
    out[7]  = (+at(0,0) * c[3] - at(0,1) * c[1] + at(0,2) * c[0]) / det;
    out[8]  = (+at(3,1) * s[5] - at(3,2) * s[4] + at(3,3) * s[3]) / det;
    out[9]  = (-at(3,0) * s[5] + at(3,2) * s[2] - at(3,3) * s[1]) / det;
    out[10] = (+at(3,0) * s[4] - at(3,1) * s[2] + at(3,3) * s[0]) / det;
    out[11] = (-at(3,0) * s[3] + at(3,1) * s[1] - at(3,2) * s[0]) / det;
    out[12] = (-at(2,1) * s[5] + at(2,2) * s[4] - at(2,3) * s[3]) / det;
    out[13] = (+at(2,0) * s[5] - at(2,2) * s[2] + at(2,3) * s[1]) / det;
    out[14] = (-at(2,0) * s[4] + at(2,1) * s[2] - at(2,3) * s[0]) / det;
    out[15] = (+at(2,0) * s[3] - at(2,1) * s[1] + at(2,2) * s[0]) / det;
    return true;
}

extern "C" int vsb_tiles_of_rank(int width, int height, int rank, int nranks, int* n_tiles)
{
    if (!n_tiles || width < 1 || height < 1 || nranks < 1 || rank < 0 || rank >= nranks) return VSB_ERR_BAD_ARG;
    const int tiles = ((width + VSB_TILE - 1) / VSB_TILE) * ((height + VSB_TILE - 1) / VSB_TILE);
    *n_tiles = (tiles - rank + nranks - 1) / nranks;   // tiles k with k % nranks == rank
    return VSB_OK;
}

extern "C" int vsb_frame(const vsb_scene* const* scenes, int n_scenes, const vsb_frame_desc* d, void* cuda_stream)
{
    int rc = require_init(); if (rc) return rc;
    if (!d || !d->color) return VSB_ERR_BAD_ARG;
    const vsb_frame_params& p = d->p;
    if (p.width < 1 || p.height < 1 || p.kernel < 0 || p.kernel > 3 || p.pixel_sampler < 0 || p.pixel_sampler > 5) return VSB_ERR_BAD_ARG;
    if (p.pixel_sampler == VSB_PS_JITTERED_BLEND && p.frame_num == 0) return VSB_ERR_BAD_ARG;
    if (d->camera_kind != VSB_CAM_BASIS && d->camera_kind != VSB_CAM_MATRIX) return VSB_ERR_BAD_ARG;
    if (d->depth && d->camera_kind != VSB_CAM_MATRIX) return VSB_ERR_BAD_ARG;   // depth_transform needs the matrices
    if (p.kernel != VSB_KERNEL_AO && (!d->materials || p.n_materials < 1)) return VSB_ERR_BAD_ARG;
    if (p.n_lights > 0 && !d->lights && p.kernel <= VSB_KERNEL_WHITTED) return VSB_ERR_BAD_ARG;
    for (int i = 0; i < n_scenes; ++i)
    {
        if (!scenes || !scenes[i]) return VSB_ERR_BAD_ARG;
        // a sphere scene shades with get_normal(hr, sphere) (surface attached with VSB_NORMALS_FROM_PRIMITIVE); the
        // reference's path tracer reads isect_pos before setting it there (detail/pathtracing.inl:74,107), so
        // simple / whitted only
        const bool sphere_ok = scenes[i]->prim_kind == VSB_PRIM_SPHERE && scenes[i]->surf.enabled &&
                               scenes[i]->surf.normal_binding == VSB_NORMALS_FROM_PRIMITIVE && p.kernel <= VSB_KERNEL_WHITTED;
        const bool generic_ok = scenes[i]->prim_kind == VSB_PRIM_GENERIC && scenes[i]->surf.enabled && p.kernel <= VSB_KERNEL_WHITTED;
        if (scenes[i]->prim_kind != VSB_PRIM_TRIANGLE && !sphere_ok && !generic_ok) return VSB_ERR_UNSUPPORTED;
    }
    for (int i = 0; i < n_scenes; ++i)
    {
        rc = scene_id_range(scenes[i]); if (rc) return rc;
        if (p.kernel != VSB_KERNEL_AO && scenes[i]->max_geom_id >= (uint32_t)p.n_materials) return VSB_ERR_BAD_ARG;   // materials[geom_id]
    }
    // normals[prim_id] is ONE array over all bvhs of the range (get_normal.h:28-37): it must cover the largest prim_id
    size_t normals_needed = 0;
    for (int i = 0; i < n_scenes; ++i)
        if ((size_t)scenes[i]->max_prim_id + 1 > normals_needed) normals_needed = (size_t)scenes[i]->max_prim_id + 1;
    if (n_scenes == 1 && normals_needed > scenes[0]->n_prims) return VSB_ERR_BAD_ARG;
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : g.stream;
    CallOrder order(st);
    rc = upload_scene_table(scenes, n_scenes, st); if (rc) return rc;

    struct { vsb::FrameCtx ctx; int scissor[4]; float4* color; float* depth; int tile_rank, tile_nranks, tiles_x, tiles_y, n_owned_tiles, in_place;
             unsigned long long* ray_count; } fl;
    std::memset(&fl, 0, sizeof(fl));
    std::memcpy(&fl.ctx.p, &p, sizeof(p));
    for (int i = 0; i < 3; ++i) { fl.ctx.eye[i] = d->cam_eye[i]; fl.ctx.cu[i] = d->cam_u[i]; fl.ctx.cv[i] = d->cam_v[i]; fl.ctx.cw[i] = d->cam_w[i]; }
    if (d->camera_kind == VSB_CAM_MATRIX)
    {
        fl.ctx.camera_kind = 1;
        std::memcpy(fl.ctx.view, d->view_matrix, 64); std::memcpy(fl.ctx.proj, d->proj_matrix, 64);
        if (!mat4_inverse(d->view_matrix, fl.ctx.inv_view) || !mat4_inverse(d->proj_matrix, fl.ctx.inv_proj)) return VSB_ERR_BAD_ARG;
    }
    const int nranks = d->tile_nranks > 1 ? d->tile_nranks : 1;
    const int rank = d->tile_nranks > 1 ? d->tile_rank : 0;
    if (rank < 0 || rank >= nranks) return VSB_ERR_BAD_ARG;
    fl.tile_rank = rank; fl.tile_nranks = d->tile_nranks > 1 ? nranks : 0;
    fl.tiles_x = (p.width + VSB_TILE - 1) / VSB_TILE; fl.tiles_y = (p.height + VSB_TILE - 1) / VSB_TILE;
    vsb_tiles_of_rank(p.width, p.height, rank, nranks, &fl.n_owned_tiles);
    std::memcpy(fl.scissor, d->scissor, sizeof(fl.scissor));
    if (fl.scissor[2] <= 0 || fl.scissor[3] <= 0) { fl.scissor[0] = 0; fl.scissor[1] = 0; fl.scissor[2] = p.width; fl.scissor[3] = p.height; }

    const size_t n_normals = n_scenes == 1 ? scenes[0]->n_prims : normals_needed;      // what a host `normals` array must hold
    fl.in_place = (nranks > 1 && d->tile_target == VSB_TILES_IN_PLACE) ? 1 : 0;
    // in-place tiles go to device memory (a peer's frame over NVLink) or, for a host frame that several ranks share, to
    // PINNED + MAPPED host memory the kernel stores into directly (checked below); never with a blend, which re-reads the target
    if (fl.in_place && p.pixel_sampler == VSB_PS_JITTERED_BLEND) return VSB_ERR_UNSUPPORTED;
    if (d->color_format < VSB_PF_RGBA32F || d->color_format > VSB_PF_RGB8) return VSB_ERR_BAD_ARG;
    // the mirror frame: compact RGBA32F tile buffers in device memory only (the pixel's position is recovered from its compact index)
    if (d->mirror_color && (nranks < 2 || fl.in_place || d->color_format != VSB_PF_RGBA32F || d->mem != VSB_MEM_DEVICE)) return VSB_ERR_UNSUPPORTED;
    if (d->color_format != VSB_PF_RGBA32F && nranks > 1 && !fl.in_place) return VSB_ERR_UNSUPPORTED;   // compact tile buffers + de-interleave are RGBA32F
    const size_t bpp = d->color_format == VSB_PF_RGBA8 ? 4 : (d->color_format == VSB_PF_RGB8 ? 3 : 16);
    const size_t color_px = (nranks > 1 && !fl.in_place) ? (size_t)fl.n_owned_tiles * VSB_TILE * VSB_TILE : (size_t)p.width * p.height;
    unsigned long long* d_ray_count = nullptr;
    if (d->mem != VSB_MEM_HOST && d->mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    if (d->params_mem != VSB_MEM_HOST && d->params_mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    fl.ctx.surf = scenes[0]->surf;
    if (fl.ctx.surf.enabled && (n_scenes != 1 || p.kernel == VSB_KERNEL_AO)) return VSB_ERR_UNSUPPORTED;
    if (!d->normals && !scenes[0]->d_normals && !fl.ctx.surf.enabled) return VSB_ERR_BAD_ARG;
    if (!d->normals && !fl.ctx.surf.enabled && scenes[0]->n_normals < normals_needed) return VSB_ERR_BAD_ARG;   // attached array too small for the range
    fl.ctx.normals = d->normals ? reinterpret_cast<const float4*>(d->normals) : scenes[0]->d_normals;
    bool generic_materials = d->generic_materials != 0;
    if (d->params_mem == VSB_MEM_HOST && d->materials)
        for (int i = 0; i < p.n_materials; ++i) generic_materials |= d->materials[i].kind >= VSB_MAT_EMISSIVE;
    // spot lights ride on the same (SURF) instantiations as the generic material list
    bool spot_lights = d->spot_lights != 0;
    if (d->params_mem == VSB_MEM_HOST && d->lights)
        for (int i = 0; i < p.n_lights; ++i) spot_lights |= d->lights[i].kind == VSB_LIGHT_SPOT;
    if (d->params_mem == VSB_MEM_HOST)
    {
        // small per-frame parameter block: materials | lights (what cuda_sched copies as kernel
        // arguments, detail/cuda_sched.inl:60-99); explicit normals ride along when given
        void *a, *b = nullptr, *c = nullptr;
        if (d->normals)
        {
            rc = stage(3, n_normals * 16, &a); if (rc) return rc;
            VSB_CUDA(cudaMemcpyAsync(a, d->normals, n_normals * 16, cudaMemcpyHostToDevice, st));
            fl.ctx.normals = reinterpret_cast<const float4*>(a);
        }
        rc = stage(4, (size_t)(p.n_materials > 0 ? p.n_materials : 1) * 64 + (size_t)(p.n_lights > 0 ? p.n_lights : 1) * sizeof(vsb_point_light) + 64, &b); if (rc) return rc;
        if (d->materials) VSB_CUDA(cudaMemcpyAsync(b, d->materials, (size_t)p.n_materials * 64, cudaMemcpyHostToDevice, st));
        c = (char*)b + (size_t)(p.n_materials > 0 ? p.n_materials : 1) * 64;
        if (d->lights && p.n_lights > 0) VSB_CUDA(cudaMemcpyAsync(c, d->lights, (size_t)p.n_lights * sizeof(vsb_point_light), cudaMemcpyHostToDevice, st));
        fl.ctx.materials = reinterpret_cast<const vsb::Material*>(b);
        fl.ctx.lights = reinterpret_cast<const vsb::PointLight*>(c);
    }
    else
    {
        fl.ctx.materials = reinterpret_cast<const vsb::Material*>(d->materials);
        fl.ctx.lights = reinterpret_cast<const vsb::PointLight*>(d->lights);
    }
    // A pinned (page-locked, mapped) host target can be written by the kernel itself: every finished
    // pixel goes over PCIe while the other lanes keep traversing, so there is no D2H pass after the
    // trace.  Pageable targets and progressive blending (which re-reads the target) are staged.
    bool direct_host = false;
    if (d->mem == VSB_MEM_HOST && (g_host_store || fl.in_place) && p.pixel_sampler != VSB_PS_JITTERED_BLEND)
    {
        cudaPointerAttributes ac, az;
        void* dev_c = nullptr; void* dev_z = nullptr;
        bool ok = cudaPointerGetAttributes(&ac, d->color) == cudaSuccess && ac.type == cudaMemoryTypeHost && ac.devicePointer;
        if (ok) dev_c = ac.devicePointer;
        if (ok && d->depth)
        {
            ok = cudaPointerGetAttributes(&az, d->depth) == cudaSuccess && az.type == cudaMemoryTypeHost && az.devicePointer;
            if (ok) dev_z = az.devicePointer;
        }
        cudaGetLastError();
        if (ok)
        {
            direct_host = true;
            fl.color = reinterpret_cast<float4*>(dev_c);
            fl.depth = reinterpret_cast<float*>(dev_z);
            if (d->ray_count) d_ray_count = g.d_work + 1;
        }
    }
    if (fl.in_place && d->mem == VSB_MEM_HOST && !direct_host) return VSB_ERR_UNSUPPORTED;   // pageable / unmapped shared frame
    if (direct_host) {}
    else if (d->mem == VSB_MEM_HOST)
    {
        void* e;
        rc = stage(5, color_px * bpp + 16, &e); if (rc) return rc;
        fl.color = reinterpret_cast<float4*>(e);
        if (d->depth)
        {
            void* z;
            rc = stage(6, color_px * 4 + 16, &z); if (rc) return rc;
            fl.depth = reinterpret_cast<float*>(z);
        }
        if (d->ray_count) d_ray_count = g.d_work + 1;
    }
    else
    {
        fl.depth = reinterpret_cast<float*>(d->depth);
        fl.color = reinterpret_cast<float4*>(d->color);
        d_ray_count = reinterpret_cast<unsigned long long*>(d->ray_count);
    }
    fl.ray_count = d_ray_count;

    VSB_CUDA(cudaMemsetAsync(g.d_work, 0, 64 * sizeof(unsigned long long), st));
    const bool exact = d->traverse == VSB_TRAVERSE_EXACT;
    if (exact && !exact_supported(scenes, n_scenes)) return VSB_ERR_UNSUPPORTED;
    vsb::TraceLaunch tl;
    tl.scenes = g.d_scenes; tl.n_scenes = n_scenes;
    tl.wide0 = scenes[0]->dev.wide;
    tl.chunk = VSB_FRAME_CHUNK; tl.ray_count = fl.ray_count;
    tl.service_min = g_frame_service_min > 0 ? g_frame_service_min : (p.kernel >= VSB_KERNEL_AO ? 24 : 32);
    tl.phase_min = g_phase_min;

    // Host render targets are traced in bands so that the D2H copy of a finished band runs on the
    // copy stream while the next bands are traced.  A band is every n_bands-th UNIT (a tile row of a
    // row-major target, 64 consecutive tiles of a compact tile buffer; each contiguous in the
    // target): sky and geometry are spread over all bands, so bands finish at an even pace and
    // the copies never queue up behind a burst of cheap bands.
    const bool compact = nranks > 1 && !fl.in_place;
    int n_bands = 1;
    if (d->mem == VSB_MEM_HOST && !direct_host)
    {
        // a band should carry >= 16 MiB of pixels: below that the per-launch ramps cost more than the overlap buys
        const size_t by_size = (color_px * bpp >> 20) / (size_t)(g_band_min_mib < 1 ? 1 : g_band_min_mib);
        n_bands = g_frame_bands < 1 ? 1 : (g_frame_bands > 32 ? 32 : g_frame_bands);
        if ((size_t)n_bands > by_size) n_bands = by_size < 1 ? 1 : (int)by_size;
    }
    const int unit_tiles = compact ? 64 : fl.tiles_x;
    const int units = compact ? (fl.n_owned_tiles + unit_tiles - 1) / unit_tiles : fl.tiles_y;
    if (n_bands > units) n_bands = units;
    auto unit_pixels = [&](int u, size_t& px0, size_t& px1)      // contiguous pixel range of unit u in the target
    {
        if (n_bands == 1) { px0 = 0; px1 = color_px; }
        else if (compact)
        {
            px0 = (size_t)u * unit_tiles * VSB_TILE * VSB_TILE; px1 = px0 + (size_t)unit_tiles * VSB_TILE * VSB_TILE;
            if (px1 > color_px) px1 = color_px;
        }
        else
        {
            const size_t row0 = (size_t)u * VSB_TILE, row1 = (size_t)(u + 1) * VSB_TILE < (size_t)p.height ? (size_t)(u + 1) * VSB_TILE : (size_t)p.height;
            px0 = row0 * (size_t)p.width; px1 = row1 * (size_t)p.width;
        }
    };
    // progressive blending: the accumulated frame lives with the caller; it is uploaded band by band
    // on its own stream (the H2D engine runs ahead of the trace, the D2H engine behind it)
    // (a PF_RGB32F pixel has a 4th word the path never writes: it makes the same round trip)
    const bool upload = d->mem == VSB_MEM_HOST && !direct_host &&
                        (p.pixel_sampler == VSB_PS_JITTERED_BLEND || d->color_format == VSB_PF_RGB32F);
    if (upload)
    {
        VSB_CUDA(cudaEventRecord(g.band_fork, st));
        VSB_CUDA(cudaStreamWaitEvent(g.upload_stream, g.band_fork, 0));
        for (int band = 0; band < n_bands; ++band)
        {
            for (int u = band; u < units; u += n_bands)
            {
                size_t px0, px1;
                unit_pixels(u, px0, px1);
                VSB_CUDA(cudaMemcpyAsync(reinterpret_cast<char*>(fl.color) + px0 * bpp, reinterpret_cast<const char*>(d->color) + px0 * bpp,
                                         (px1 - px0) * bpp, cudaMemcpyHostToDevice, g.upload_stream));
                if (d->depth)
                    VSB_CUDA(cudaMemcpyAsync(reinterpret_cast<char*>(fl.depth) + px0 * 4, reinterpret_cast<const char*>(d->depth) + px0 * 4,
                                             (px1 - px0) * 4, cudaMemcpyHostToDevice, g.upload_stream));
                if (n_bands == 1) break;
            }
            VSB_CUDA(cudaEventRecord(g.band_up[band], g.upload_stream));
        }
    }

    cudaEvent_t tr_k[32], tr_c[32];
    VSB_CUDA(cudaEventRecord(g.ev0, st));
    // bands alternate between the caller's stream and a second one: the persistent CTAs of band
    // i+1 are dispatched as those of band i retire, so no SM idles in a band's tail
    const bool two_streams = n_bands > 1 && g_band_streams > 1;
    if (two_streams)
    {
        VSB_CUDA(cudaEventRecord(g.band_fork, st));
        VSB_CUDA(cudaStreamWaitEvent(g.band_stream, g.band_fork, 0));
    }
    for (int band = 0; band < n_bands; ++band)
    {
        cudaStream_t bs = (two_streams && (band & 1)) ? g.band_stream : st;
        if (upload) VSB_CUDA(cudaStreamWaitEvent(bs, g.band_up[band], 0));
        const int units_in_band = (units - band + n_bands - 1) / n_bands;
        const int tile_begin = n_bands == 1 ? 0 : band * unit_tiles;
        tl.n_items = (unsigned long long)(n_bands == 1 ? fl.n_owned_tiles : units_in_band * unit_tiles) * VSB_TILE * VSB_TILE;
        tl.work_counter = n_bands == 1 ? g.d_work : g.d_work + 2 + band;
#define VSB_LAUNCH_FRAME(K, SS, SF)                                                                                \
    do {                                                                                                     \
        vsb::FrameProgram<K, SS, SF> prog;                                                                       \
        prog.ctx = fl.ctx; std::memcpy(prog.scissor, fl.scissor, sizeof(prog.scissor)); prog.color = fl.color; prog.color_format = d->color_format; \
        prog.depth = fl.depth;                                                                               \
        prog.tile_rank = fl.tile_rank; prog.tile_nranks = fl.tile_nranks; prog.tiles_x = fl.tiles_x;         \
        prog.in_place = fl.in_place; prog.tile_begin = tile_begin;                                           \
        prog.unit_tiles = unit_tiles; prog.unit_stride = n_bands; prog.n_local_tiles = fl.n_owned_tiles;     \
        prog.mirror = reinterpret_cast<float4*>(d->mirror_color);                                            \
        if (exact) VSB_CUDA((launch_trace<vsb::FrameProgram<K, SS, SF>, true>(prog, tl, bs)));                   \
        else       VSB_CUDA((launch_trace<vsb::FrameProgram<K, SS, SF>, false>(prog, tl, bs)));                  \
    } while (0)
        // EXT instantiation: SSAA samplers, matrix camera, depth target, colour formats; SURF instantiation:
        // kernel_params with a surface description or a generic material list (see shade.cuh)
        {
            const bool ext = p.pixel_sampler >= VSB_PS_SSAA2 || d->camera_kind == VSB_CAM_MATRIX || d->color_format != VSB_PF_RGBA32F;
            const bool surf = p.kernel != VSB_KERNEL_AO && (fl.ctx.surf.enabled || generic_materials || spot_lights);
            switch (p.kernel * 4 + (ext ? 2 : 0) + (surf ? 1 : 0))
            {
            case 0:  VSB_LAUNCH_FRAME(0, false, false); break;
            case 1:  VSB_LAUNCH_FRAME(0, false, true); break;
            case 2:  VSB_LAUNCH_FRAME(0, true, false); break;
            case 3:  VSB_LAUNCH_FRAME(0, true, true); break;
            case 4:  VSB_LAUNCH_FRAME(1, false, false); break;
            case 5:  VSB_LAUNCH_FRAME(1, false, true); break;
            case 6:  VSB_LAUNCH_FRAME(1, true, false); break;
            case 7:  VSB_LAUNCH_FRAME(1, true, true); break;
            case 8:  VSB_LAUNCH_FRAME(2, false, false); break;
            case 10: VSB_LAUNCH_FRAME(2, true, false); break;
            case 12: VSB_LAUNCH_FRAME(3, false, false); break;
            case 13: VSB_LAUNCH_FRAME(3, false, true); break;
            case 14: VSB_LAUNCH_FRAME(3, true, false); break;
            default: VSB_LAUNCH_FRAME(3, true, true); break;
            }
        }
#undef VSB_LAUNCH_FRAME
        ++g.launches;
        VSB_CUDA(cudaGetLastError());
        if (direct_host) VSB_CUDA(cudaEventRecord(g.ev1, st));
        else if (d->mem == VSB_MEM_HOST)
        {
            VSB_CUDA(cudaEventRecord(g.band_done[band], bs));
            if (g_band_trace) { cudaEventCreate(&tr_k[band]); cudaEventRecord(tr_k[band], bs); }
            VSB_CUDA(cudaStreamWaitEvent(g.copy_stream, g.band_done[band], 0));
            for (int u = band; u < units; u += n_bands)
            {
                size_t px0, px1;
                unit_pixels(u, px0, px1);
                VSB_CUDA(cudaMemcpyAsync(reinterpret_cast<char*>(d->color) + px0 * bpp, reinterpret_cast<const char*>(fl.color) + px0 * bpp,
                                         (px1 - px0) * bpp, cudaMemcpyDeviceToHost, g.copy_stream));
                if (d->depth)
                    VSB_CUDA(cudaMemcpyAsync(reinterpret_cast<char*>(d->depth) + px0 * 4, reinterpret_cast<const char*>(fl.depth) + px0 * 4,
                                             (px1 - px0) * 4, cudaMemcpyDeviceToHost, g.copy_stream));
                if (n_bands == 1) break;
            }
            if (g_band_trace) { cudaEventCreate(&tr_c[band]); cudaEventRecord(tr_c[band], g.copy_stream); }
        }
    }
    if (two_streams)
    {
        VSB_CUDA(cudaEventRecord(g.band_join, g.band_stream));
        VSB_CUDA(cudaStreamWaitEvent(st, g.band_join, 0));
    }
    if (d->mem == VSB_MEM_HOST && !direct_host) VSB_CUDA(cudaEventRecord(g.ev1, st));
    if (d->mem == VSB_MEM_HOST)
    {
        if (d->ray_count) VSB_CUDA(cudaMemcpyAsync(d->ray_count, d_ray_count, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        VSB_CUDA(cudaStreamSynchronize(st));
        VSB_CUDA(cudaStreamSynchronize(g.copy_stream));   // host-visible target: frame() returns with pixels in place
        if (g_band_trace && !direct_host)
            for (int b = 0; b < n_bands; ++b)
            {
                float a = 0.0f, c = 0.0f;
                cudaEventElapsedTime(&a, g.ev0, tr_k[b]); cudaEventElapsedTime(&c, g.ev0, tr_c[b]);
                std::fprintf(stderr, "band %2d traced at %7.3f ms, on the host at %7.3f ms\n", b, a, c);
                cudaEventDestroy(tr_k[b]); cudaEventDestroy(tr_c[b]);
            }
    }
    else VSB_CUDA(cudaEventRecord(g.ev1, st));
    return VSB_OK;
}

extern "C" int vsb_deinterleave_tiles(const void* gathered, int width, int height, int nranks,
                                      int tiles_per_rank_padded, void* image, void* cuda_stream)
{
    int rc = require_init(); if (rc) return rc;
    if (!gathered || !image || width < 1 || height < 1 || nranks < 1) return VSB_ERR_BAD_ARG;
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : g.stream;
    dim3 block(32, 8), grid((width + 31) / 32, (height + 7) / 8);
    vsb::deinterleave_kernel<<<grid, block, 0, st>>>(reinterpret_cast<const float4*>(gathered), width, height, nranks,
                                                      tiles_per_rank_padded, reinterpret_cast<float4*>(image));
    ++g.launches;
    VSB_CUDA(cudaGetLastError());
    return VSB_OK;
}

extern "C" int vsb_device_alloc(size_t bytes, void** out)
{
    int rc = require_init(); if (rc) return rc;
    if (!out || bytes == 0) return VSB_ERR_BAD_ARG;
    VSB_CUDA(cudaMalloc(out, bytes));
    return VSB_OK;
}

extern "C" int vsb_device_free(void* p)
{
    if (p) VSB_CUDA(cudaFree(p));
    return VSB_OK;
}

extern "C" int vsb_device_copy(void* dst, const void* src, size_t bytes, int dst_mem, int src_mem)
{
    int rc = require_init(); if (rc) return rc;
    if (!dst || !src) return VSB_ERR_BAD_ARG;
    const cudaMemcpyKind kind = dst_mem == VSB_MEM_DEVICE ? (src_mem == VSB_MEM_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice)
                                                          : (src_mem == VSB_MEM_DEVICE ? cudaMemcpyDeviceToHost : cudaMemcpyHostToHost);
    if (g.call_done_valid) VSB_CUDA(cudaStreamWaitEvent(g.stream, g.call_done, 0));   // behind the last query / frame call of the context
    VSB_CUDA(cudaMemcpyAsync(dst, src, bytes, kind, g.stream));
    VSB_CUDA(cudaStreamSynchronize(g.stream));
    return VSB_OK;
}

extern "C" int vsb_ipc_export(const void* dev_ptr, void* handle64)
{
    int rc = require_init(); if (rc) return rc;
    if (!dev_ptr || !handle64) return VSB_ERR_BAD_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
    cudaIpcMemHandle_t h;
    VSB_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(dev_ptr)));
    std::memcpy(handle64, &h, 64);
    return VSB_OK;
}

extern "C" int vsb_ipc_open(const void* handle64, void** out)
{
    int rc = require_init(); if (rc) return rc;
    if (!handle64 || !out) return VSB_ERR_BAD_ARG;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle64, 64);
    VSB_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return VSB_OK;
}

extern "C" int vsb_ipc_close(void* p)
{
    if (p) VSB_CUDA(cudaIpcCloseMemHandle(p));
    return VSB_OK;
}

extern "C" int vsb_last_kernel_ms(float* ms)
{
    int rc = require_init(); if (rc) return rc;
    if (!ms) return VSB_ERR_BAD_ARG;
    VSB_CUDA(cudaEventSynchronize(g.ev1));
    VSB_CUDA(cudaEventElapsedTime(ms, g.ev0, g.ev1));
    return VSB_OK;
}

extern "C" int vsb_launch_count(uint64_t* n)
{
    if (!n) return VSB_ERR_BAD_ARG;
    *n = g.launches;
    return VSB_OK;
}

// capi_noalloc.inl -- vsb_closest_hit_static(): closest_hit over a device ray list with NO allocation of any kind
// (included by capi.cu: same translation unit as the kernels it launches).
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  vsb_closest_hit()
// keeps its scratch -- the scene table, the work-queue counter, the coherence sort's keys / order / radix-sort
// temporaries -- in buffers it cudaMalloc's on first use and grows on demand.  This variant keeps all of it in ONE
// file-scope __device__ array of fixed size, so it calls no cudaMalloc / cudaFree / allocating library routine
// (cub::DeviceRadixSort is given its temporary storage; it allocates nothing itself).
//
// The array is part of the loaded module: EVERY GPU that loads libvsnray_b200.so holds VSB_STATIC_SCRATCH_BYTES
// (24 MiB + 8 KiB) of device memory until the process exits, used or not.  All calls share it and it keeps the last
// call's contents: what must start at zero (the work counter) is cleared by the set-up kernel of every call, and two
// calls that overlap on the GPU would race on it -- calls of one context are ordered by the library (CallOrder), calls
// from several host threads or processes sharing a context are the caller's to serialise.

namespace
{

// layout of the scratch array
//   [0, 256)            work-queue counter (8 bytes used)
//   [256, 1024)         scene table entry (vsb::DeviceScene)
//   [8 KiB, +4 n) x 4   sort keys in / out, ray order in / out   (n <= VSB_STATIC_MAX_RAYS)
//   [.., +8 MiB)        radix-sort temporaries
constexpr size_t STATIC_HEADER_BYTES = 8192;
constexpr size_t STATIC_SORT_TEMP_BYTES = (size_t)8 << 20;
constexpr size_t STATIC_SCRATCH_BYTES = STATIC_HEADER_BYTES + 4 * 4 * (size_t)VSB_STATIC_MAX_RAYS + STATIC_SORT_TEMP_BYTES;
static_assert(sizeof(vsb::DeviceScene) <= 768, "scene table entry must fit the header");
static_assert(STATIC_SCRATCH_BYTES == VSB_STATIC_SCRATCH_BYTES, "the header states the size");

} // namespace

// In HOST code this name is a host-side stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress (static_scratch_address below).  Device code may name it directly.
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

namespace
{

unsigned char* g_static_scratch_ptr[64] = {};      // per device: the array's device address, fetched once with that device current

int static_scratch_address(unsigned char** out)
{
    int dev = -1;
    VSB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return VSB_ERR_UNSUPPORTED;
    if (!g_static_scratch_ptr[dev])
    {
        void* p = nullptr;
        VSB_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return VSB_ERR_UNSUPPORTED;
        g_static_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_static_scratch_ptr[dev];
    return VSB_OK;
}

// the scene table entry arrives BY VALUE as a kernel argument (no host-side staging copy), the counter is cleared
__global__ void static_setup_kernel(vsb::DeviceScene scene, vsb::DeviceScene* table, unsigned long long* counter)
{
    if (threadIdx.x == 0 && blockIdx.x == 0)
    {
        *table = scene;
        counter[0] = 0ull;
    }
}

} // namespace

extern "C" int vsb_closest_hit_static(const vsb_scene* scene, const vsb_ray* rays, size_t n_rays, float max_t,
                                      vsb_hit* hits, void* cuda_stream)
{
    int rc = require_init(); if (rc) return rc;
    if (n_rays == 0) return VSB_OK;
    if (!scene || !rays || !hits) return VSB_ERR_BAD_ARG;
    if (n_rays > (size_t)VSB_STATIC_MAX_RAYS) return VSB_ERR_TOO_LARGE;
    unsigned char* base = nullptr;
    rc = static_scratch_address(&base); if (rc) return rc;
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : g.stream;
    CallOrder order(st);

    unsigned long long* counter = reinterpret_cast<unsigned long long*>(base);
    vsb::DeviceScene* table = reinterpret_cast<vsb::DeviceScene*>(base + 256);
    static_setup_kernel<<<1, 32, 0, st>>>(scene->dev, table, counter);
    VSB_CUDA(cudaGetLastError());
    ++g.launches;

    const float4* d_rays = reinterpret_cast<const float4*>(rays);
    const uint32_t n = (uint32_t)n_rays;
    const uint32_t* d_order = nullptr;
    if (g_sort_rays && n_rays >= (1u << 16))
    {
        // coherence sort (order only), all of its buffers inside the static array
        uint32_t* k0 = reinterpret_cast<uint32_t*>(base + STATIC_HEADER_BYTES);
        uint32_t* k1 = k0 + VSB_STATIC_MAX_RAYS;
        uint32_t* i0 = k1 + VSB_STATIC_MAX_RAYS;
        uint32_t* i1 = i0 + VSB_STATIC_MAX_RAYS;
        void* tmp = i1 + VSB_STATIC_MAX_RAYS;
        size_t tmp_bytes = 0;
        VSB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint32_t*)k0, k1, (const uint32_t*)i0, i1, (int)n, 0, VSB_SORT_KEY_BITS, st));
        if (tmp_bytes <= STATIC_SORT_TEMP_BYTES)        // else: trace in the caller's order (same hit records, slower)
        {
            float3 lo = make_float3(scene->lo[0], scene->lo[1], scene->lo[2]);
            float3 inv = make_float3(1.0f / fmaxf(scene->hi[0] - scene->lo[0], 1e-30f), 1.0f / fmaxf(scene->hi[1] - scene->lo[1], 1e-30f),
                                     1.0f / fmaxf(scene->hi[2] - scene->lo[2], 1e-30f));
            vsb::ray_key_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_rays, n, lo, inv, k0, i0);
            VSB_CUDA(cudaGetLastError());
            VSB_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, (const uint32_t*)k0, k1, (const uint32_t*)i0, i1, (int)n, 0, VSB_SORT_KEY_BITS, st));
            g.launches += 2;
            d_order = i1;
        }
    }
    vsb::RayListProgram prog;
    prog.order = d_order;
    prog.rays = d_rays; prog.max_t_per_ray = nullptr; prog.max_t = max_t; prog.hits = reinterpret_cast<uint4*>(hits); prog.any = 0;
    vsb::TraceLaunch tl;
    tl.scenes = table; tl.n_scenes = 1; tl.n_items = (unsigned long long)n_rays;
    tl.wide0 = scene->dev.wide;
    tl.work_counter = counter; tl.ray_count = nullptr;
    tl.service_min = g_query_service_min; tl.phase_min = g_phase_min;
    {
        unsigned long long per = (unsigned long long)n_rays / ((unsigned long long)g.sm_count * 32ull * 16ull);
        per = (per / 32ull) * 32ull;
        tl.chunk = (unsigned)(per < 32ull ? 32ull : (per > 256ull ? 256ull : per));
    }
    VSB_CUDA((launch_trace<vsb::RayListProgram, false>(prog, tl, st)));
    ++g.launches;
    return VSB_OK;
}

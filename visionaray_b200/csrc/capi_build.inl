// capi_build.inl -- vsb_scene_build(): the BVH is built on the GPU instead of being handed in from the host builder:
// by default the reference's own binned-SAH tree, breadth-first (build_sah.cuh, capi_build_sah.inl); with
// vsb_tune("build_algorithm", 0) a Morton + PLOC tree (build_ploc.cuh).  Included by capi.cu after capi_scene.inl.

namespace
{

int   g_build_radius = 8;          // vsb_tune("build_radius", 1..32)
float g_build_cost_node = 1.0f;    // vsb_tune("build_cost_node_x100", ...)
float g_build_cost_prim = 1.0f;    // vsb_tune("build_cost_prim_x100", ...)
float g_last_build_ms[4] = { 0, 0, 0, 0 };   // total, sort + setup, PLOC, layout + wide re-encode
int   g_last_build_iters = 0;
int   g_build_algorithm = 1;       // vsb_tune("build_algorithm", 0 = Morton + PLOC, 1 = the reference's binned SAH, level by level)

// scratch from the stream-ordered pool (vsb_init keeps it cached between builds)
struct BuildScratch
{
    void* p[40] = {};
    int n = 0;
    cudaStream_t st = nullptr;
    ~BuildScratch() { for (int i = 0; i < n; ++i) if (p[i]) cudaFreeAsync(p[i], st); }
    template <typename T> cudaError_t alloc(T** out, size_t count)
    {
        void* q = nullptr;
        cudaError_t e = cudaMallocAsync(&q, count * sizeof(T) + 16, st);
        if (e == cudaSuccess) { p[n++] = q; *out = reinterpret_cast<T*>(q); }
        return e;
    }
};

} // namespace

#include "capi_build_sah.inl"

extern "C" int vsb_scene_build(int prim_kind, const void* prims, size_t n_prims, int mem, vsb_scene** out)
{
    int rc = require_init(); if (rc) return rc;
    if (!prims || !out || n_prims == 0) return VSB_ERR_BAD_ARG;
    if (prim_kind != VSB_PRIM_TRIANGLE && prim_kind != VSB_PRIM_SPHERE && prim_kind != VSB_PRIM_GENERIC) return VSB_ERR_BAD_ARG;
    if (mem != VSB_MEM_HOST && mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    if (n_prims >= 0x3fffffffull) return VSB_ERR_UNSUPPORTED;
    const uint32_t n = (uint32_t)n_prims;
    const size_t prim_bytes = vsb::prim_stride(prim_kind);
    cudaStream_t st = g.stream;

    vsb_scene* s = new vsb_scene();
    s->prim_kind = prim_kind; s->n_prims = n_prims; s->n_idx = n_prims;
    BuildScratch scratch;
    scratch.st = st;
    cudaEvent_t ev[4] = { nullptr, nullptr, nullptr, nullptr };
    auto fail = [&](int code) { for (auto e : ev) if (e) cudaEventDestroy(e); vsb_scene_destroy(s); return code; };
#define VSB_BUILD_CUDA(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return fail((int)e__); } while (0)
    for (auto& e : ev) VSB_BUILD_CUDA(cudaEventCreate(&e));

    VSB_BUILD_CUDA(cudaMalloc(&s->d_prims, n_prims * prim_bytes));
    VSB_BUILD_CUDA(cudaMemcpyAsync(s->d_prims, prims, n_prims * prim_bytes,
                                   mem == VSB_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
    VSB_BUILD_CUDA(cudaMalloc((void**)&s->d_indices, n_prims * sizeof(uint32_t)));

    vsb::BuildParams bp;
    bp.radius = g_build_radius < 1 ? 1 : (g_build_radius > VSB_BUILD_MAX_RADIUS ? VSB_BUILD_MAX_RADIUS : g_build_radius);
    bp.cost_node = g_build_cost_node; bp.cost_prim = g_build_cost_prim;

    int iters = 0;
    if (g_build_algorithm == 1)
    {
        // the reference's own tree (build_sah.cuh)
        VSB_BUILD_CUDA(cudaEventRecord(ev[0], st));
        int levels = 0;
        const int brc = build_tree_sah(s, prim_kind, n, st, scratch, ev, levels);
        if (brc != VSB_OK) return fail(brc);
        iters = levels;
        s->depth2 = levels + 1;
        vsb::Node32 r;
        VSB_BUILD_CUDA(cudaMemcpyAsync(&r, s->d_nodes2, sizeof(r), cudaMemcpyDeviceToHost, st));
        VSB_BUILD_CUDA(cudaStreamSynchronize(st));
        s->lo[0] = r.min_x; s->lo[1] = r.min_y; s->lo[2] = r.min_z; s->hi[0] = r.max_x; s->hi[1] = r.max_y; s->hi[2] = r.max_z;
    }
    else
    {
    const size_t n_bn = 2 * (size_t)n - 1;
    vsb::BNode *d_bn = nullptr, *d_cl[2] = { nullptr, nullptr };
    vsb::BSub* d_sub = nullptr;
    uint32_t *d_parent = nullptr, *d_nn = nullptr, *d_vals[2] = { nullptr, nullptr }, *d_small = nullptr;
    uint64_t *d_keys[2] = { nullptr, nullptr }, *d_flags = nullptr, *d_scan = nullptr;
    void* d_tmp = nullptr;
    VSB_BUILD_CUDA(scratch.alloc(&d_bn, n_bn));
    VSB_BUILD_CUDA(scratch.alloc(&d_sub, n_bn));
    VSB_BUILD_CUDA(scratch.alloc(&d_parent, n_bn));
    VSB_BUILD_CUDA(scratch.alloc(&d_cl[0], (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_cl[1], (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_nn, (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_keys[0], (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_keys[1], (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_vals[0], (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_vals[1], (size_t)n));
    VSB_BUILD_CUDA(scratch.alloc(&d_small, (size_t)16));        // [0..5] centroid bounds, [8..9] totals, [12] task counter
    d_flags = d_keys[0]; d_scan = d_keys[1];                     // the sort buffers are free once the clusters exist
    size_t tmp_sort = 0, tmp_scan = 0;
    VSB_BUILD_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_sort, d_keys[0], d_keys[1], d_vals[0], d_vals[1], (int)n, 0, 63, st));
    VSB_BUILD_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tmp_scan, d_flags, d_scan, (int)n, st));
    const size_t tmp_bytes = tmp_sort > tmp_scan ? tmp_sort : tmp_scan;
    {
        char* t = nullptr;
        VSB_BUILD_CUDA(scratch.alloc(&t, tmp_bytes + 256));
        d_tmp = t;
    }

    VSB_BUILD_CUDA(cudaEventRecord(ev[0], st));
    // 1. + 2. bounds, Morton codes, sort, leaf clusters.  d_cl[1] holds the unsorted boxes until the
    // first PLOC iteration overwrites it.
    {
        const uint32_t init[6] = { 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u };
        VSB_BUILD_CUDA(cudaMemcpyAsync(d_small, init, sizeof(init), cudaMemcpyHostToDevice, st));
    }
    const uint32_t blocks = (n + 255) / 256;
    vsb::prim_bounds_kernel<<<blocks, 256, 0, st>>>(prim_kind, s->d_prims, n, d_cl[1], d_small);
    vsb::morton_kernel<<<blocks, 256, 0, st>>>(d_cl[1], n, d_small, d_keys[0], d_vals[0]);
    size_t tb = tmp_bytes;
    VSB_BUILD_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tb, d_keys[0], d_keys[1], d_vals[0], d_vals[1], (int)n, 0, 63, st));
    vsb::init_clusters_kernel<<<blocks, 256, 0, st>>>(d_cl[1], d_vals[1], n, bp, d_bn, d_sub, d_parent, d_cl[0]);
    g.launches += 3 + 8;
    VSB_BUILD_CUDA(cudaGetLastError());
    VSB_BUILD_CUDA(cudaEventRecord(ev[1], st));

    // 3. PLOC iterations, state on the device (build_ploc.cuh), one host sync per chunk of iterations
    const int MAX_ITERS = 1024, CHUNK = 8;
    uint32_t* d_state = nullptr;
    VSB_BUILD_CUDA(scratch.alloc(&d_state, (size_t)2 * (MAX_ITERS + CHUNK + 1)));
    {
        const uint32_t init2[2] = { n, 0u };
        VSB_BUILD_CUDA(cudaMemcpyAsync(d_state, init2, sizeof(init2), cudaMemcpyHostToDevice, st));
    }
    uint32_t m = n, merged = 0; int cur = 0;
    while (m > 1)
    {
        if (iters + CHUNK > MAX_ITERS) return fail(VSB_ERR_BAD_TREE);
        const uint32_t b = (m + VSB_PLOC_BLOCK - 1) / VSB_PLOC_BLOCK;            // m only shrinks within the chunk
        for (int k = 0; k < CHUNK; ++k)
        {
            uint32_t* stt = d_state + 2 * (size_t)(iters + k);
            vsb::ploc_nn_kernel<<<b, VSB_PLOC_BLOCK, 0, st>>>(d_cl[cur], stt, bp.radius, d_nn);
            vsb::ploc_flags_kernel<<<b, VSB_PLOC_BLOCK, 0, st>>>(d_nn, stt, d_flags);
            tb = tmp_bytes;
            VSB_BUILD_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp, tb, d_flags, d_scan, (int)m, st));   // entries beyond the live count are never read
            vsb::ploc_emit_kernel<<<b, VSB_PLOC_BLOCK, 0, st>>>(d_cl[cur], d_nn, d_scan, stt, n, bp, d_bn, d_sub, d_parent, d_cl[cur ^ 1]);
            g.launches += 4;
            cur ^= 1;
        }
        std::vector<uint32_t> hist(2 * (CHUNK + 1));
        VSB_BUILD_CUDA(cudaMemcpyAsync(hist.data(), d_state + 2 * (size_t)iters, hist.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        VSB_BUILD_CUDA(cudaStreamSynchronize(st));
        for (int k = 0; k < CHUNK && m > 1; ++k)
        {
            const uint32_t m1 = hist[2 * (k + 1)], g1 = hist[2 * (k + 1) + 1];
            if (g1 == merged || m1 != m - (g1 - merged)) return fail(VSB_ERR_BAD_TREE);   // no progress: cannot happen (see ploc_nearest)
            m = m1; merged = g1; ++iters;
        }
    }
    if (merged != n - 1) return fail(VSB_ERR_BAD_TREE);
    VSB_BUILD_CUDA(cudaEventRecord(ev[2], st));

    // 4. layout into the reference's node / indices arrays, level by level
    const uint32_t root = n == 1 ? 0u : n + merged - 1u;
    vsb::BSub root_sub;
    VSB_BUILD_CUDA(cudaMemcpyAsync(&root_sub, d_sub + root, sizeof(root_sub), cudaMemcpyDeviceToHost, st));
    VSB_BUILD_CUDA(cudaStreamSynchronize(st));
    const size_t n_nodes = 2 * (size_t)root_sub.lcnt - 1;
    s->n_nodes2 = n_nodes;
    VSB_BUILD_CUDA(cudaMalloc((void**)&s->d_nodes2, n_nodes * sizeof(vsb::Node32)));
    vsb::BTask* d_tasks[2] = { reinterpret_cast<vsb::BTask*>(d_cl[0]), reinterpret_cast<vsb::BTask*>(d_cl[1]) };   // the cluster arrays are free now
    {
        vsb::BTask t; t.node = root; t.out = 0; t.prim_off = 0; t.rank_off = 0;
        VSB_BUILD_CUDA(cudaMemcpyAsync(d_tasks[0], &t, sizeof(t), cudaMemcpyHostToDevice, st));
    }
    // task counts per level live on the device; levels are launched in chunks with grids sized for the bound
    // "tasks double per level" and synced once per chunk
    const int MAX_LEVELS = 4096, LCHUNK = 8;
    uint32_t* d_lcount = nullptr;
    VSB_BUILD_CUDA(scratch.alloc(&d_lcount, (size_t)MAX_LEVELS + LCHUNK + 2));
    VSB_BUILD_CUDA(cudaMemsetAsync(d_lcount, 0, ((size_t)MAX_LEVELS + LCHUNK + 2) * sizeof(uint32_t), st));
    {
        const uint32_t one = 1u;
        VSB_BUILD_CUDA(cudaMemcpyAsync(d_lcount, &one, sizeof(one), cudaMemcpyHostToDevice, st));
    }
    uint32_t n_tasks = 1; int levels = 0; cur = 0;
    while (n_tasks > 0)
    {
        if (levels + LCHUNK > MAX_LEVELS) return fail(VSB_ERR_BAD_TREE);
        unsigned long long bound = n_tasks;
        for (int k = 0; k < LCHUNK; ++k)
        {
            const uint32_t nb = (uint32_t)(bound < (unsigned long long)n ? bound : (unsigned long long)n);
            vsb::layout_level_kernel<<<(nb + 127) / 128, 128, 0, st>>>(d_tasks[cur], d_lcount + levels + k, d_bn, d_sub, s->d_nodes2, s->d_indices,
                                                                    d_tasks[cur ^ 1]);
            ++g.launches;
            cur ^= 1;
            bound *= 2;
        }
        uint32_t cnt[LCHUNK + 1];
        VSB_BUILD_CUDA(cudaMemcpyAsync(cnt, d_lcount + levels, sizeof(cnt), cudaMemcpyDeviceToHost, st));
        VSB_BUILD_CUDA(cudaStreamSynchronize(st));
        for (int k = 0; k < LCHUNK && n_tasks > 0; ++k) { n_tasks = cnt[k + 1]; ++levels; }
    }
    s->depth2 = (root_sub.cnt & BS_LEAF) ? 1 : levels + 1;
    {
        vsb::Node32 r;
        VSB_BUILD_CUDA(cudaMemcpy(&r, s->d_nodes2, sizeof(r), cudaMemcpyDeviceToHost));
        s->lo[0] = r.min_x; s->lo[1] = r.min_y; s->lo[2] = r.min_z; s->hi[0] = r.max_x; s->hi[1] = r.max_y; s->hi[2] = r.max_z;
    }
    }
#undef VSB_BUILD_CUDA

    // 5. the same BVH2 -> BVH8 re-encode as for a host-built tree (leaves already hold <= 4 primitives)
    rc = encode_wide(s, s->d_nodes2, s->n_nodes2, n_prims);
    if (rc != VSB_OK) return fail(rc);
    cudaEventRecord(ev[3], st);
    cudaEventSynchronize(ev[3]);
    cudaEventElapsedTime(&g_last_build_ms[0], ev[0], ev[3]);
    cudaEventElapsedTime(&g_last_build_ms[1], ev[0], ev[1]);
    cudaEventElapsedTime(&g_last_build_ms[2], ev[1], ev[2]);
    cudaEventElapsedTime(&g_last_build_ms[3], ev[2], ev[3]);
    g_last_build_iters = iters;
    for (auto e : ev) cudaEventDestroy(e);
    *out = s;
    return VSB_OK;
}

extern "C" int vsb_scene_download_bvh(const vsb_scene* s, vsb_bvh_node* nodes_out, size_t nodes_cap,
                                      uint32_t* indices_out, size_t indices_cap)
{
    int rc = require_init(); if (rc) return rc;
    if (!s) return VSB_ERR_BAD_ARG;
    if (nodes_out)
    {
        if (nodes_cap < s->n_nodes2) return VSB_ERR_BAD_ARG;
        VSB_CUDA(cudaMemcpy(nodes_out, s->d_nodes2, s->n_nodes2 * sizeof(vsb::Node32), cudaMemcpyDeviceToHost));
    }
    if (indices_out)
    {
        if (!s->d_indices || indices_cap < s->n_idx) return VSB_ERR_BAD_ARG;
        VSB_CUDA(cudaMemcpy(indices_out, s->d_indices, s->n_idx * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    }
    return VSB_OK;
}

extern "C" int vsb_last_build_ms(float* ms4, int* iterations)
{
    if (!ms4) return VSB_ERR_BAD_ARG;
    for (int i = 0; i < 4; ++i) ms4[i] = g_last_build_ms[i];
    if (iterations) *iterations = g_last_build_iters;
    return VSB_OK;
}

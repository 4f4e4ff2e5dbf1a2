// capi_scene.inl -- context + scene half of capi.cu (included there).

namespace
{

struct Context
{
    bool         inited = false;
    int          device = -1;
    int          sm_count = 0;
    cudaStream_t stream = nullptr;          // internal stream for uploads / encode
    cudaStream_t copy_stream = nullptr;     // D2H of finished frame bands, overlapping the next band's kernel
    cudaStream_t band_stream = nullptr;     // odd bands are traced here, so a band's tail overlaps the next band's start
    cudaStream_t upload_stream = nullptr;   // H2D of the caller's progressive frame, band by band ahead of the trace
    cudaEvent_t  band_done[32] = {}, band_up[32] = {};
    cudaEvent_t  band_fork = nullptr, band_join = nullptr;
    cudaEvent_t  ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t  call_done = nullptr;       // recorded behind the device work of every query / frame call (CallOrder below)
    bool         call_done_valid = false;
    unsigned long long* d_work = nullptr;   // persistent-kernel work counter
    vsb::DeviceScene* d_scenes = nullptr;   // scene table handed to kernels (up to MAX_SCENES)
    void*        staging[24] = {};
    size_t       staging_cap[24] = {};
    float        last_ms = 0.0f;
    unsigned long long launches = 0;
};

const int MAX_SCENES = 16;
int g_optimal_collapse = 1;     // vsb_tune("optimal_collapse", 0|1), read at vsb_scene_create
Context g;

#define VSB_CUDA(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return (int)e__; } while (0)

int require_init()
{
    if (!g.inited) return VSB_ERR_NOT_INIT;
    return VSB_OK;
}

// The scene table, the work counters and the staging slots are per-context device state that every query / frame
// call rewrites.  Calls may come on different streams (the library's own, the caller's, torch's current one) and
// device-memory calls return asynchronously, so each call's stream first waits for the work of the previous call:
// calls on one context are ordered on the device however the caller mixes streams (the reference's frame() is
// not re-entrant either, detail/cuda_sched.inl:307-320).  Host THREADS must still serialise their calls.
struct CallOrder
{
    cudaStream_t st;
    explicit CallOrder(cudaStream_t s) : st(s)
    {
        if (g.call_done_valid) cudaStreamWaitEvent(st, g.call_done, 0);
    }
    ~CallOrder()
    {
        if (cudaEventRecord(g.call_done, st) == cudaSuccess) g.call_done_valid = true;
    }
};

// grow-only device staging buffer for host-memory callers
int stage(int slot, size_t bytes, void** out)
{
    if (bytes > g.staging_cap[slot])
    {
        if (g.staging[slot]) cudaFree(g.staging[slot]);
        g.staging[slot] = nullptr; g.staging_cap[slot] = 0;
        size_t cap = bytes + bytes / 4 + 256;
        VSB_CUDA(cudaMalloc(&g.staging[slot], cap));
        g.staging_cap[slot] = cap;
    }
    *out = g.staging[slot];
    return VSB_OK;
}

} // namespace

struct vsb_scene
{
    int      prim_kind = 0;
    size_t   n_prims = 0, n_nodes2 = 0, n_idx = 0;
    void*    d_prims = nullptr;          // reference layout (exact traversal)
    vsb::Node32* d_nodes2 = nullptr;
    uint32_t* d_indices = nullptr;       // NULL for a non-index bvh
    vsb::Node80* d_wide = nullptr;
    void*    d_wide_prims = nullptr;
    uint32_t* d_rec_of_list = nullptr;   // primitive_list_index -> leaf-order record
    float4*  d_normals = nullptr;        // optional, vsb_scene_set_normals()
    size_t   n_normals = 0;              // entries of d_normals (>= max_prim_id + 1; a frame over several bvhs indexes ONE array by prim_id)
    mutable int id_range_known = 0;      // largest prim_id / geom_id of the primitives, computed on first use
    mutable uint32_t max_prim_id = 0, max_geom_id = 0;
    vsb::SurfCtx surf = {};              // optional, vsb_scene_set_surface(); device copies listed in surf_allocs
    std::vector<void*> surf_allocs;
    uint32_t n_wide = 0, n_wide_prims = 0;
    int      levels = 0, depth2 = 0;
    float    lo[3] = { 0, 0, 0 }, hi[3] = { 0, 0, 0 };   // root bounds
    vsb::DeviceScene dev;
};

namespace
{

// Host validation of the incoming BVH2 (the reference only asserts, bvh.h:74,86,112): every
// node reachable exactly once, children adjacent and in range, leaf ranges inside indices[],
// indices inside prims[].  Returns max depth (root = 1) or a negative error.
int validate_tree(const vsb_bvh_node* nodes, size_t n_nodes, const uint32_t* indices, size_t n_idx,
                  size_t n_prims, bool* big_leaves)
{
    if (n_nodes == 0) return VSB_ERR_BAD_TREE;
    std::vector<uint8_t> seen(n_nodes, 0);
    std::vector<std::pair<uint32_t, int>> st;
    st.push_back({ 0u, 1 });
    int maxd = 0;
    size_t visited = 0, leaf_total = 0;
    *big_leaves = false;
    const size_t list_len = indices ? n_idx : n_prims;
    while (!st.empty())
    {
        auto cur = st.back(); st.pop_back();
        uint32_t i = cur.first; int d = cur.second;
        if (i >= n_nodes || seen[i]) return VSB_ERR_BAD_TREE;
        seen[i] = 1; ++visited;
        if (d > maxd) maxd = d;
        const vsb_bvh_node& n = nodes[i];
        if (n.num_prims == 0)
        {
            if ((size_t)n.first + 1 >= n_nodes || n.first == 0) return VSB_ERR_BAD_TREE;
            st.push_back({ n.first, d + 1 });
            st.push_back({ n.first + 1, d + 1 });
        }
        else
        {
            if ((size_t)n.first + n.num_prims > list_len) return VSB_ERR_BAD_TREE;
            leaf_total += n.num_prims;
            if (leaf_total > list_len) return VSB_ERR_BAD_TREE;   // overlapping leaf ranges: the wide encode's scratch holds list_len records
            if (n.num_prims > 4) *big_leaves = true;
            if (indices)
                for (uint32_t k = 0; k < n.num_prims; ++k)
                    if (indices[n.first + k] >= n_prims) return VSB_ERR_BAD_TREE;
        }
    }
    (void)visited;
    return maxd;
}

// Leaves with more than 4 primitives (degenerate centroids or SAH "leaf is cheaper",
// detail/bvh/sah.h:677-762) are split into chains of virtual BVH2 nodes that keep the parent
// bounds, so that the collapse can assume <= 4 primitives per leaf.  Only the copy used for
// the wide encode is changed; the exact traversal keeps the caller's tree.
void normalise_leaves(std::vector<vsb_bvh_node>& nodes)
{
    for (size_t i = 0; i < nodes.size(); ++i)
    {
        vsb_bvh_node n = nodes[i];
        if (n.num_prims <= 4) continue;
        uint32_t half = n.num_prims / 2;
        vsb_bvh_node a = n, b = n;
        a.num_prims = half;
        b.first = n.first + half; b.num_prims = n.num_prims - half;
        n.first = (uint32_t)nodes.size(); n.num_prims = 0;
        nodes[i] = n;
        nodes.push_back(a);
        nodes.push_back(b);   // revisited later by this same loop if still > 4
    }
}

} // namespace

extern "C" int vsb_init(int device)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return VSB_ERR_NO_DEVICE;
    if (device < 0 || device >= count) return VSB_ERR_BAD_ARG;
    if (g.inited && g.device == device) return VSB_OK;
    if (g.inited) vsb_shutdown();
    VSB_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    VSB_CUDA(cudaGetDeviceProperties(&prop, device));
    g.device = device;
    g.sm_count = prop.multiProcessorCount;
    {
        // scratch of scene builds comes from the stream-ordered pool (cudaMallocAsync); keep up to 8 GiB of it
        // cached between builds so that a rebuild does not pay cudaMalloc / cudaFree round trips
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess)
        {
            unsigned long long keep = 8ull << 30;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    VSB_CUDA(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    VSB_CUDA(cudaStreamCreateWithFlags(&g.copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 32; ++i) VSB_CUDA(cudaEventCreateWithFlags(&g.band_done[i], cudaEventDisableTiming));
    VSB_CUDA(cudaStreamCreateWithFlags(&g.band_stream, cudaStreamNonBlocking));
    VSB_CUDA(cudaStreamCreateWithFlags(&g.upload_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 32; ++i) VSB_CUDA(cudaEventCreateWithFlags(&g.band_up[i], cudaEventDisableTiming));
    VSB_CUDA(cudaEventCreateWithFlags(&g.band_fork, cudaEventDisableTiming));
    VSB_CUDA(cudaEventCreateWithFlags(&g.band_join, cudaEventDisableTiming));
    VSB_CUDA(cudaEventCreate(&g.ev0));
    VSB_CUDA(cudaEventCreate(&g.ev1));
    VSB_CUDA(cudaEventCreateWithFlags(&g.call_done, cudaEventDisableTiming));
    g.call_done_valid = false;
    VSB_CUDA(cudaMalloc(&g.d_work, 64 * sizeof(unsigned long long)));   // [0] queue, [1] ray count, [2..33] band queues
    VSB_CUDA(cudaMalloc(&g.d_scenes, sizeof(vsb::DeviceScene) * MAX_SCENES));
    {
        float table[256];
        for (int u = 0; u < 256; ++u) table[u] = (float)((double)(float)u / 255.0);      // unorm_to_float<8>, math/detail/unorm.inl:26-31
        VSB_CUDA(cudaMemcpyToSymbol(vsb::c_unorm8_to_float, table, sizeof(table)));
    }
    g.inited = true;
    return VSB_OK;
}

extern "C" int vsb_shutdown(void)
{
    if (!g.inited) return VSB_OK;
    cudaSetDevice(g.device);
    cudaDeviceSynchronize();
    for (int i = 0; i < 24; ++i) { if (g.staging[i]) cudaFree(g.staging[i]); g.staging[i] = nullptr; g.staging_cap[i] = 0; }
    if (g.d_work) cudaFree(g.d_work);
    if (g.d_scenes) cudaFree(g.d_scenes);
    if (g.ev0) cudaEventDestroy(g.ev0);
    if (g.ev1) cudaEventDestroy(g.ev1);
    if (g.call_done) cudaEventDestroy(g.call_done);
    g.call_done = nullptr; g.call_done_valid = false;
    if (g.stream) cudaStreamDestroy(g.stream);
    if (g.copy_stream) cudaStreamDestroy(g.copy_stream);
    if (g.band_stream) cudaStreamDestroy(g.band_stream);
    if (g.upload_stream) cudaStreamDestroy(g.upload_stream);
    for (int i = 0; i < 32; ++i) if (g.band_up[i]) cudaEventDestroy(g.band_up[i]);
    if (g.band_fork) cudaEventDestroy(g.band_fork);
    if (g.band_join) cudaEventDestroy(g.band_join);
    for (int i = 0; i < 32; ++i) if (g.band_done[i]) cudaEventDestroy(g.band_done[i]);
    g = Context();
    return VSB_OK;
}

extern "C" int vsb_sync(void)
{
    int rc = require_init(); if (rc) return rc;
    VSB_CUDA(cudaStreamSynchronize(g.stream));
    if (g.call_done_valid) VSB_CUDA(cudaEventSynchronize(g.call_done));   // the last query / frame call, whichever stream it ran on
    return VSB_OK;
}

extern "C" int vsb_device_sm_count(int* out)
{
    int rc = require_init(); if (rc) return rc;
    *out = g.sm_count;
    return VSB_OK;
}

extern "C" int vsb_scene_destroy(vsb_scene* s)
{
    if (!s) return VSB_OK;
    if (s->d_prims) cudaFree(s->d_prims);
    if (s->d_nodes2) cudaFree(s->d_nodes2);
    if (s->d_indices) cudaFree(s->d_indices);
    if (s->d_wide) cudaFree(s->d_wide);
    if (s->d_wide_prims) cudaFree(s->d_wide_prims);
    if (s->d_rec_of_list) cudaFree(s->d_rec_of_list);
    if (s->d_normals) cudaFree(s->d_normals);
    for (void* q : s->surf_allocs) cudaFree(q);
    delete s;
    return VSB_OK;
}

namespace
{

// GPU re-encode BVH2 -> BVH8 of a scene whose reference-layout arrays (d_prims, d_nodes2, d_indices)
// are on the device.  d_enc_nodes: the BVH2 to collapse (the scene's own, or a normalised copy with
// <= 4 primitives per leaf).  On failure the caller destroys the scene.
int encode_wide(vsb_scene* s, vsb::Node32* d_enc_nodes, size_t n_enc_nodes, size_t list_len)
{
    const size_t rec_bytes = (size_t)vsb::wide_rec_words16(s->prim_kind) * 16;
    uint32_t* d_leaf_list = nullptr; vsb::WideTask* d_tasks[2] = { nullptr, nullptr }; uint32_t* d_counters = nullptr;
    uint32_t* d_parent = nullptr; uint32_t* d_arrived = nullptr; vsb::CollapseCost* d_cost = nullptr; vsb::CollapseDec* d_dec = nullptr;
    vsb::Node80* d_wide_tmp = nullptr;     // encoded into scratch of the worst-case size, then copied to an exact-size array
    const size_t wide_cap = n_enc_nodes;   // a wide node consumes >= 1 BVH2 inner node
    cudaError_t err = cudaSuccess;
    cudaStream_t st = g.stream;
    auto cleanup = [&]()
    {
        void* all[] = { d_leaf_list, d_tasks[0], d_tasks[1], d_counters, d_parent, d_arrived, d_cost, d_dec, d_wide_tmp };
        for (void* q : all) if (q) cudaFreeAsync(q, st);
    };
#define VSB_ENC_CUDA(expr) do { err = (expr); if (err != cudaSuccess) { cleanup(); return (int)err; } } while (0)
    VSB_ENC_CUDA(cudaMallocAsync((void**)&d_wide_tmp, wide_cap * sizeof(vsb::Node80), st));
    VSB_ENC_CUDA(cudaMallocAsync((void**)&d_leaf_list, (list_len + 1) * sizeof(uint32_t), st));
    VSB_ENC_CUDA(cudaMallocAsync((void**)&d_tasks[0], (wide_cap + 1) * sizeof(vsb::WideTask), st));
    VSB_ENC_CUDA(cudaMallocAsync((void**)&d_tasks[1], (wide_cap + 1) * sizeof(vsb::WideTask), st));
    VSB_ENC_CUDA(cudaMallocAsync((void**)&d_counters, 4 * sizeof(uint32_t), st));
    {
        vsb::WideTask root; root.bvh2_node = 0; root.wide_index = 0;
        uint32_t init[4] = { 1u, 0u, 0u, 0u };
        VSB_ENC_CUDA(cudaMemcpyAsync(d_tasks[0], &root, sizeof(root), cudaMemcpyHostToDevice, g.stream));
        VSB_ENC_CUDA(cudaMemcpyAsync(d_counters, init, sizeof(init), cudaMemcpyHostToDevice, g.stream));
        VSB_ENC_CUDA(cudaStreamSynchronize(g.stream));      // `root` / `init` are stack variables
    }
    // optimal collapse: bottom-up DP over the BVH2 (g_optimal_collapse = 0: greedy largest-area-first)
    if (g_optimal_collapse)
    {
        const uint32_t nn = (uint32_t)n_enc_nodes;
        VSB_ENC_CUDA(cudaMallocAsync((void**)&d_parent, (size_t)nn * 4, st));
        VSB_ENC_CUDA(cudaMallocAsync((void**)&d_arrived, (size_t)nn * 4, st));
        VSB_ENC_CUDA(cudaMallocAsync((void**)&d_cost, (size_t)nn * sizeof(vsb::CollapseCost), st));
        VSB_ENC_CUDA(cudaMallocAsync((void**)&d_dec, (size_t)nn * sizeof(vsb::CollapseDec), st));
        VSB_ENC_CUDA(cudaMemsetAsync(d_arrived, 0, (size_t)nn * 4, g.stream));
        vsb::bvh2_parents_kernel<<<(nn + 255) / 256, 256, 0, g.stream>>>(d_enc_nodes, nn, d_parent);
        vsb::collapse_dp_kernel<<<(nn + 255) / 256, 256, 0, g.stream>>>(d_enc_nodes, nn, d_parent, d_arrived, d_cost, d_dec);
        g.launches += 2;
        VSB_ENC_CUDA(cudaGetLastError());
    }
    uint32_t n_tasks = 1; int cur = 0; s->levels = 0;
    uint32_t counters[4] = { 1u, 0u, 0u, 0u };
    while (n_tasks > 0)
    {
        VSB_ENC_CUDA(cudaMemsetAsync(d_counters + 2, 0, sizeof(uint32_t), g.stream));
        vsb::encode_level_kernel<<<(n_tasks + 127) / 128, 128, 0, g.stream>>>(
            d_tasks[cur], n_tasks, d_enc_nodes, d_wide_tmp, d_leaf_list, d_tasks[cur ^ 1], d_counters, d_dec);
        ++g.launches;
        VSB_ENC_CUDA(cudaGetLastError());
        VSB_ENC_CUDA(cudaMemcpyAsync(counters, d_counters, sizeof(counters), cudaMemcpyDeviceToHost, g.stream));
        VSB_ENC_CUDA(cudaStreamSynchronize(g.stream));
        n_tasks = counters[2];
        cur ^= 1;
        ++s->levels;
        if (s->levels > 256) { cleanup(); return VSB_ERR_BAD_TREE; }
    }
    s->n_wide = counters[0];
    s->n_wide_prims = counters[1];
    if (s->n_wide_prims != list_len || s->levels > VSB_STACK_SIZE - 2) { cleanup(); return VSB_ERR_BAD_TREE; }
    VSB_ENC_CUDA(cudaMalloc((void**)&s->d_wide, (size_t)s->n_wide * sizeof(vsb::Node80) + 64));
    VSB_ENC_CUDA(cudaMemcpyAsync(s->d_wide, d_wide_tmp, (size_t)s->n_wide * sizeof(vsb::Node80), cudaMemcpyDeviceToDevice, st));
    VSB_ENC_CUDA(cudaMalloc(&s->d_wide_prims, (size_t)s->n_wide_prims * rec_bytes + 64));
    VSB_ENC_CUDA(cudaMalloc((void**)&s->d_rec_of_list, (size_t)s->n_wide_prims * sizeof(uint32_t) + 64));
    vsb::fill_prims_kernel<<<(s->n_wide_prims + 255) / 256, 256, 0, g.stream>>>(
        s->prim_kind, s->d_prims, s->d_indices, d_leaf_list, s->n_wide_prims, s->d_wide_prims, s->d_rec_of_list);
    ++g.launches;
    VSB_ENC_CUDA(cudaGetLastError());
    VSB_ENC_CUDA(cudaStreamSynchronize(g.stream));
    cleanup();
#undef VSB_ENC_CUDA

    s->dev.wide.nodes = s->d_wide; s->dev.wide.prims = s->d_wide_prims; s->dev.wide.prim_kind = s->prim_kind;
    s->dev.wide.rec_of_list = s->d_rec_of_list;
    s->dev.exact.nodes = s->d_nodes2; s->dev.exact.prims = s->d_prims; s->dev.exact.indices = s->d_indices;
    s->dev.exact.prim_kind = s->prim_kind;
    return VSB_OK;
}

} // namespace

extern "C" int vsb_scene_create(int prim_kind, const void* prims, size_t n_prims,
                                const vsb_bvh_node* nodes, size_t n_nodes,
                                const uint32_t* indices, size_t n_indices, vsb_scene** out)
{
    int rc = require_init(); if (rc) return rc;
    if (!prims || !nodes || !out || n_prims == 0 || n_nodes == 0) return VSB_ERR_BAD_ARG;
    if (prim_kind != VSB_PRIM_TRIANGLE && prim_kind != VSB_PRIM_SPHERE && prim_kind != VSB_PRIM_GENERIC) return VSB_ERR_BAD_ARG;
    if (n_nodes >= 0x7fffffffull || n_prims >= 0x7fffffffull) return VSB_ERR_UNSUPPORTED;

    bool big = false;
    const int depth = validate_tree(nodes, n_nodes, indices, n_indices, n_prims, &big);
    if (depth < 0) return depth;
    if (depth > VSB_EXACT_STACK - 1 || depth > VSB_STACK_SIZE - 1) return VSB_ERR_BAD_TREE;

    vsb_scene* s = new vsb_scene();
    s->prim_kind = prim_kind; s->n_prims = n_prims; s->n_nodes2 = n_nodes; s->n_idx = indices ? n_indices : 0;
    s->depth2 = depth;
    for (int i = 0; i < 3; ++i) { s->lo[i] = nodes[0].bbox_min[i]; s->hi[i] = nodes[0].bbox_max[i]; }
    const size_t prim_bytes = vsb::prim_stride(prim_kind);
    const size_t list_len = indices ? n_indices : n_prims;

#define VSB_SCENE_CUDA(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { vsb_scene_destroy(s); return (int)e__; } } while (0)
    VSB_SCENE_CUDA(cudaMalloc(&s->d_prims, n_prims * prim_bytes));
    VSB_SCENE_CUDA(cudaMalloc((void**)&s->d_nodes2, n_nodes * sizeof(vsb::Node32)));
    if (indices) VSB_SCENE_CUDA(cudaMalloc((void**)&s->d_indices, n_indices * sizeof(uint32_t)));
    VSB_SCENE_CUDA(cudaMemcpyAsync(s->d_prims, prims, n_prims * prim_bytes, cudaMemcpyHostToDevice, g.stream));
    VSB_SCENE_CUDA(cudaMemcpyAsync(s->d_nodes2, nodes, n_nodes * sizeof(vsb::Node32), cudaMemcpyHostToDevice, g.stream));
    if (indices) VSB_SCENE_CUDA(cudaMemcpyAsync(s->d_indices, indices, n_indices * sizeof(uint32_t), cudaMemcpyHostToDevice, g.stream));

    // BVH2 used for the encode: the caller's tree, or a normalised copy when it has big leaves
    vsb::Node32* d_enc_nodes = s->d_nodes2;
    size_t n_enc_nodes = n_nodes;
    vsb::Node32* d_norm = nullptr;
    if (big)
    {
        std::vector<vsb_bvh_node> norm(nodes, nodes + n_nodes);
        normalise_leaves(norm);
        n_enc_nodes = norm.size();
        VSB_SCENE_CUDA(cudaMalloc((void**)&d_norm, n_enc_nodes * sizeof(vsb::Node32)));
        cudaError_t e = cudaMemcpy(d_norm, norm.data(), n_enc_nodes * sizeof(vsb::Node32), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { cudaFree(d_norm); vsb_scene_destroy(s); return (int)e; }
        d_enc_nodes = d_norm;
    }
#undef VSB_SCENE_CUDA
    rc = encode_wide(s, d_enc_nodes, n_enc_nodes, list_len);
    if (d_norm) cudaFree(d_norm);
    if (rc != VSB_OK) { vsb_scene_destroy(s); return rc; }
    *out = s;
    return VSB_OK;
}

namespace
{
// prim_id indexes the per-primitive arrays (normals, colours, tex coords), geom_id the materials and textures:
// the reference trusts them; here the scene's largest ids are computed once and every consumer checks its array sizes
int scene_id_range(const vsb_scene* s)
{
    if (s->id_range_known) return VSB_OK;
    uint32_t* d2 = nullptr;
    VSB_CUDA(cudaMalloc((void**)&d2, 8));
    cudaError_t e = cudaMemsetAsync(d2, 0, 8, g.stream);
    if (e == cudaSuccess)
    {
        vsb::id_range_kernel<<<(unsigned)(g.sm_count * 8), 256, 0, g.stream>>>(s->prim_kind, s->d_prims, s->n_prims, d2);
        ++g.launches;
        uint32_t h[2] = { 0u, 0u };
        e = cudaMemcpyAsync(h, d2, 8, cudaMemcpyDeviceToHost, g.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(g.stream);
        if (e == cudaSuccess) { s->max_prim_id = h[0]; s->max_geom_id = h[1]; s->id_range_known = 1; }
    }
    cudaFree(d2);
    return e == cudaSuccess ? VSB_OK : (int)e;
}
} // namespace

extern "C" int vsb_scene_set_normals(vsb_scene* s, const float* normals16, size_t n, int mem)
{
    int rc = require_init(); if (rc) return rc;
    // n is normally n_prims; a scene that is ONE of several bvhs of a frame carries the whole range's array
    // (the reference indexes normals[prim_id] over all bvhs of [begin, end), get_normal.h:28-37)
    if (!s || !normals16 || n < 1) return VSB_ERR_BAD_ARG;
    if (mem != VSB_MEM_HOST && mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    rc = scene_id_range(s); if (rc) return rc;
    if (s->max_prim_id >= n) return VSB_ERR_BAD_ARG;          // normals[prim_id] would leave the array
    if (s->d_normals && s->n_normals != n) { cudaFree(s->d_normals); s->d_normals = nullptr; }
    if (!s->d_normals) VSB_CUDA(cudaMalloc((void**)&s->d_normals, n * 16));
    s->n_normals = n;
    VSB_CUDA(cudaMemcpyAsync(s->d_normals, normals16, n * 16,
                             mem == VSB_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, g.stream));
    VSB_CUDA(cudaStreamSynchronize(g.stream));
    return VSB_OK;
}

extern "C" int vsb_scene_set_surface(vsb_scene* s, const vsb_surface_desc* desc, int mem)
{
    int rc = require_init(); if (rc) return rc;
    if (!s) return VSB_ERR_BAD_ARG;
    VSB_CUDA(cudaStreamSynchronize(g.stream));      // frames in flight may still read the old copies
    for (void* q : s->surf_allocs) cudaFree(q);
    s->surf_allocs.clear();
    s->surf = vsb::SurfCtx{};
    if (!desc) return VSB_OK;
    if (mem != VSB_MEM_HOST && mem != VSB_MEM_DEVICE) return VSB_ERR_BAD_ARG;
    if (desc->normal_binding < 0 || desc->normal_binding > VSB_NORMALS_FROM_PRIMITIVE || desc->color_binding < 0 || desc->color_binding > 2)
        return VSB_ERR_BAD_ARG;
    if (!desc->normals && desc->normal_binding != VSB_NORMALS_FROM_PRIMITIVE) return VSB_ERR_BAD_ARG;
    // spheres: only get_normal(hr, sphere) applies (bindings, lerped colours and tex coords are triangle rules); a
    // generic_primitive needs a normals array in the reference too (get_surface.h:218-245) and has no rule here
    // generic_primitive: per-face / per-vertex normals for its triangles, get_normal(hr, sphere) for its spheres
    // (detail/generic_primitive_get_normal.h); no colours / tex coords, and a normals array is required (get_surface.h:218-245)
    if (s->prim_kind == VSB_PRIM_GENERIC &&
        (desc->normal_binding == VSB_NORMALS_FROM_PRIMITIVE || desc->color_binding != VSB_COLORS_NONE || desc->tex_coords))
        return VSB_ERR_UNSUPPORTED;
    if (s->prim_kind == VSB_PRIM_SPHERE &&
        (desc->normal_binding != VSB_NORMALS_FROM_PRIMITIVE || desc->color_binding == VSB_COLORS_PER_VERTEX || desc->tex_coords))
        return VSB_ERR_UNSUPPORTED;
    if (desc->color_binding && !desc->colors) return VSB_ERR_BAD_ARG;
    if (desc->n_textures < 0 || (desc->n_textures > 0 && !desc->textures)) return VSB_ERR_BAD_ARG;
    rc = scene_id_range(s); if (rc) return rc;
    if (s->max_prim_id >= s->n_prims) return VSB_ERR_BAD_ARG;  // normals / colours / tex coords are indexed by prim_id
    const cudaMemcpyKind kind = mem == VSB_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    int err = VSB_OK;
    auto copy = [&](const void* src, size_t bytes) -> void*
    {
        void* q = nullptr;
        if (err || bytes == 0) return q;
        cudaError_t e = cudaMalloc(&q, bytes);
        if (e == cudaSuccess) { s->surf_allocs.push_back(q); e = cudaMemcpyAsync(q, src, bytes, kind, g.stream); }
        if (e != cudaSuccess) err = (int)e;
        return q;
    };
    vsb::SurfCtx c{};
    c.normal_binding = desc->normal_binding; c.color_binding = desc->color_binding;
    const size_t n = s->n_prims;
    if (desc->normal_binding != VSB_NORMALS_FROM_PRIMITIVE)
        c.normals = (const float4*)copy(desc->normals, (desc->normal_binding ? 3 * n : n) * 16);
    if (desc->color_binding) c.colors = (const float4*)copy(desc->colors, (desc->color_binding == 2 ? 3 * n : n) * 16);
    if (desc->tex_coords)
    {
        c.tex_coords = (const float2*)copy(desc->tex_coords, 3 * n * 8);
        // texture descriptors are host structures either way; their texels live where `mem` says
        std::vector<vsb::DevTexture> tab((size_t)(desc->n_textures > 0 ? desc->n_textures : 1));
        std::memset(tab.data(), 0, tab.size() * sizeof(vsb::DevTexture));
        for (int i = 0; i < desc->n_textures && !err; ++i)
        {
            const vsb_texture& t = desc->textures[i];
            if (t.width <= 0 || t.height <= 0 || !t.data) continue;
            if (t.filter < VSB_TEX_NEAREST || t.filter > VSB_TEX_CARDINAL_SPLINE || t.filter == 3 /* BSplineInterpol: needs the prefilter */ || t.address[0] < 0 || t.address[0] > VSB_TEX_BORDER ||
                t.address[1] < 0 || t.address[1] > VSB_TEX_BORDER) { err = VSB_ERR_UNSUPPORTED; break; }
            tab[i].data = (const uchar4*)copy(t.data, (size_t)t.width * (size_t)t.height * 4);
            tab[i].width = t.width; tab[i].height = t.height; tab[i].filter = t.filter;
            tab[i].addr_s = t.address[0]; tab[i].addr_t = t.address[1];
        }
        c.n_textures = desc->n_textures;
        if (!err)
        {
            void* q = nullptr;
            cudaError_t e = cudaMalloc(&q, tab.size() * sizeof(vsb::DevTexture));
            if (e == cudaSuccess) { s->surf_allocs.push_back(q); e = cudaMemcpyAsync(q, tab.data(), tab.size() * sizeof(vsb::DevTexture), cudaMemcpyHostToDevice, g.stream); }
            if (e != cudaSuccess) err = (int)e;
            c.textures = (const vsb::DevTexture*)q;
            if (!err) err = (int)cudaStreamSynchronize(g.stream);    // tab is a local
        }
    }
    if (!err) err = (int)cudaStreamSynchronize(g.stream);
    if (err)
    {
        for (void* q : s->surf_allocs) cudaFree(q);
        s->surf_allocs.clear();
        return err;
    }
    if (s->prim_kind == VSB_PRIM_SPHERE)
    {
        c.ref_spheres = reinterpret_cast<const float4*>(s->d_prims);
        c.indices = s->d_indices;
    }
    if (s->prim_kind == VSB_PRIM_GENERIC)
    {
        c.ref_generic = reinterpret_cast<const float4*>(s->d_prims);
        c.indices = s->d_indices;
    }
    if (s->prim_kind != VSB_PRIM_SPHERE && desc->normal_binding != VSB_NORMALS_PER_FACE)
    {
        const size_t n_list = s->d_indices ? s->n_idx : s->n_prims;
        float4* gn = nullptr;
        cudaError_t e = cudaMalloc((void**)&gn, n_list * sizeof(float4));
        if (e != cudaSuccess) { for (void* q : s->surf_allocs) cudaFree(q); s->surf_allocs.clear(); return (int)e; }
        s->surf_allocs.push_back(gn);
        vsb::surface_geo_normals_kernel<<<(unsigned)((n_list + 255) / 256), 256, 0, g.stream>>>(
            s->prim_kind, s->d_prims, s->d_indices, n_list, gn);
        ++g.launches;
        e = cudaStreamSynchronize(g.stream);
        if (e != cudaSuccess) { for (void* q : s->surf_allocs) cudaFree(q); s->surf_allocs.clear(); return (int)e; }
        c.geo_normals = gn;
    }
    c.enabled = 1;
    s->surf = c;
    return VSB_OK;
}

extern "C" int vsb_scene_stats(const vsb_scene* s, uint64_t* out, int n)
{
    if (!s || !out) return VSB_ERR_BAD_ARG;
    uint64_t v[6] = { s->n_wide, (uint64_t)s->n_wide * sizeof(vsb::Node80), s->n_wide_prims,
                      (uint64_t)s->n_wide_prims * ((size_t)vsb::wide_rec_words16(s->prim_kind) * 16),
                      (uint64_t)s->depth2, (uint64_t)s->n_nodes2 };
    for (int i = 0; i < n && i < 6; ++i) out[i] = v[i];
    if (n > 6) out[6] = (uint64_t)s->levels;
    return VSB_OK;
}

extern "C" int vsb_scene_download_wide(const vsb_scene* s, void* nodes_out, size_t node_bytes_cap,
                                       void* prims_out, size_t prim_bytes_cap)
{
    if (!s) return VSB_ERR_BAD_ARG;
    const size_t nb = (size_t)s->n_wide * sizeof(vsb::Node80);
    const size_t pb = (size_t)s->n_wide_prims * ((size_t)vsb::wide_rec_words16(s->prim_kind) * 16);
    if (nodes_out) { if (node_bytes_cap < nb) return VSB_ERR_BAD_ARG; VSB_CUDA(cudaMemcpy(nodes_out, s->d_wide, nb, cudaMemcpyDeviceToHost)); }
    if (prims_out) { if (prim_bytes_cap < pb) return VSB_ERR_BAD_ARG; VSB_CUDA(cudaMemcpy(prims_out, s->d_wide_prims, pb, cudaMemcpyDeviceToHost)); }
    return VSB_OK;
}

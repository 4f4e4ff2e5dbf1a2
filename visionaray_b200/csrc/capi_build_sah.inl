// capi_build_sah.inl -- host side of build_sah.cuh: the reference's binned-SAH tree built level by level on the GPU.
// Included by capi_build.inl; fills s->d_nodes2 / s->d_indices like the PLOC path and leaves the BVH8 re-encode to
// the caller.

namespace
{

#define VSB_SAH_CUDA(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return (int)e__; } while (0)

// ev[1]: set-up done, ev[2]: all levels done
int build_tree_sah(vsb_scene* s, int prim_kind, uint32_t n, cudaStream_t st, BuildScratch& scratch, cudaEvent_t* ev, int& levels_out)
{
    const uint32_t n_chunks = (n + VSB_SAH_CHUNK - 1) / VSB_SAH_CHUNK;
    const uint32_t seg_cap = n / (VSB_SAH_MAX_LEAF + 1) + 2;            // open segments hold > max_leaf_size references
    const size_t node_cap = 2 * (size_t)n;                               // every leaf holds >= 1 reference
    const int MAX_LEVELS = 4096, CHUNK = 4;

    vsb::SahRef* d_refs = nullptr;
    uint32_t *d_seg_of = nullptr, *d_F = nullptr, *d_pos_l = nullptr, *d_pos_r = nullptr;
    vsb::SahSeg* d_segs[2] = { nullptr, nullptr };
    vsb::SahSplit* d_splits = nullptr;
    vsb::SahChild* d_child[2] = { nullptr, nullptr };
    vsb::SahNode* d_nodes = nullptr;
    uint64_t *d_packed = nullptr, *d_scan = nullptr, *d_keys[2] = { nullptr, nullptr };
    uint32_t *d_vals[2] = { nullptr, nullptr }, *d_rank = nullptr, *d_slots = nullptr, *d_span = nullptr, *d_root = nullptr, *d_state = nullptr;
    VSB_SAH_CUDA(scratch.alloc(&d_refs, (size_t)n));
    VSB_SAH_CUDA(scratch.alloc(&d_seg_of, (size_t)n));
    VSB_SAH_CUDA(scratch.alloc(&d_F, (size_t)n));
    VSB_SAH_CUDA(scratch.alloc(&d_pos_l, (size_t)n));
    VSB_SAH_CUDA(scratch.alloc(&d_pos_r, (size_t)n));
    VSB_SAH_CUDA(scratch.alloc(&d_segs[0], (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_segs[1], (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_splits, (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_child[0], (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_child[1], (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_packed, (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_scan, (size_t)seg_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_nodes, node_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_keys[0], node_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_keys[1], node_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_vals[0], node_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_vals[1], node_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_rank, node_cap));
    VSB_SAH_CUDA(scratch.alloc(&d_slots, (size_t)n_chunks * VSB_SAH_SLOT_WORDS));
    VSB_SAH_CUDA(scratch.alloc(&d_span, (size_t)n_chunks));
    VSB_SAH_CUDA(scratch.alloc(&d_root, (size_t)16));
    VSB_SAH_CUDA(scratch.alloc(&d_state, (size_t)2 * (MAX_LEVELS + CHUNK + 2)));
    size_t t_flag = 0, t_seg = 0, t_sort = 0;
    using FlagIt = cub::TransformInputIterator<uint32_t, vsb::SahFlag, cub::CountingInputIterator<uint32_t>>;
    const vsb::SahFlag flag_op{ d_refs, d_seg_of, d_splits };
    const FlagIt flags(cub::CountingInputIterator<uint32_t>(0u), flag_op);
    VSB_SAH_CUDA(cub::DeviceScan::InclusiveSum(nullptr, t_flag, flags, d_F, (int)n, st));
    VSB_SAH_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, t_seg, d_packed, d_scan, (int)seg_cap, st));
    VSB_SAH_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, t_sort, d_keys[0], d_keys[1], d_vals[0], d_vals[1], (int)node_cap, 0, 64, st));
    size_t tmp_bytes = t_flag > t_seg ? t_flag : t_seg;
    if (t_sort > tmp_bytes) tmp_bytes = t_sort;
    char* d_tmp = nullptr;
    VSB_SAH_CUDA(scratch.alloc(&d_tmp, tmp_bytes + 256));

    // set-up: references, root bounds, cleared slots
    const uint32_t blocks = (n + 255) / 256;
    vsb::sah_fill_kernel<<<1, 32, 0, st>>>(d_root, 12, 0);
    {
        const size_t words = (size_t)n_chunks * VSB_SAH_SLOT_WORDS;
        vsb::sah_fill_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(d_slots, words, 1);
    }
    vsb::sah_init_kernel<<<blocks, 256, 0, st>>>(prim_kind, s->d_prims, n, d_refs, d_seg_of, d_root);
    vsb::sah_root_kernel<<<1, 32, 0, st>>>(d_root, n, d_segs[0], d_nodes, d_state);
    g.launches += 4;
    VSB_SAH_CUDA(cudaGetLastError());
    uint32_t st2[2] = { 0u, 0u };
    VSB_SAH_CUDA(cudaMemcpyAsync(st2, d_state, sizeof(st2), cudaMemcpyDeviceToHost, st));
    VSB_SAH_CUDA(cudaEventRecord(ev[1], st));
    VSB_SAH_CUDA(cudaStreamSynchronize(st));

    // levels
    uint32_t n_open = st2[0], n_nodes = st2[1];
    int level = 0, cur = 0;
    while (n_open > 0)
    {
        if (level + CHUNK > MAX_LEVELS) return VSB_ERR_UNSUPPORTED;
        unsigned long long bound = n_open;
        for (int k = 0; k < CHUNK; ++k)
        {
            const uint32_t sb = (uint32_t)(bound < (unsigned long long)seg_cap ? bound : (unsigned long long)seg_cap);
            uint32_t* stt = d_state + 2 * (size_t)(level + k);
            vsb::sah_bin_kernel<<<n_chunks, VSB_SAH_CHUNK, 0, st>>>(d_refs, d_seg_of, n, d_segs[cur], d_splits, d_slots, d_span,
                                                                    level + k > 0 ? d_child[cur ^ 1] : nullptr);
            vsb::sah_span_split_kernel<<<(n_chunks * 16 + 127) / 128, 128, 0, st>>>(n_chunks, d_segs[cur], d_splits, d_slots, d_span);
            size_t tb = tmp_bytes;
            VSB_SAH_CUDA(cub::DeviceScan::InclusiveSum(d_tmp, tb, flags, d_F, (int)n, st));
            vsb::sah_mislist_kernel<<<blocks, 256, 0, st>>>(d_seg_of, n, d_segs[cur], d_splits, d_F, d_pos_l, d_pos_r);
            vsb::sah_swap_kernel<<<blocks, 256, 0, st>>>(d_refs, d_seg_of, n, d_segs[cur], d_splits, d_F, d_pos_l, d_pos_r);
            vsb::sah_child_count_kernel<<<(sb + 127) / 128, 128, 0, st>>>(stt, sb, d_segs[cur], d_splits, d_F, d_packed);
            tb = tmp_bytes;
            VSB_SAH_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp, tb, d_packed, d_scan, (int)sb, st));
            vsb::sah_emit_children_kernel<<<(sb + 127) / 128, 128, 0, st>>>(stt, sb, d_segs[cur], d_splits, d_F, d_packed, d_scan, d_segs[cur ^ 1],
                                                                             d_nodes, d_child[cur]);
            g.launches += 8;
            cur ^= 1;
            bound *= 2;
        }
        VSB_SAH_CUDA(cudaGetLastError());
        uint32_t hist[2 * (4 + 1)];
        VSB_SAH_CUDA(cudaMemcpyAsync(hist, d_state + 2 * (size_t)level, sizeof(hist), cudaMemcpyDeviceToHost, st));
        VSB_SAH_CUDA(cudaStreamSynchronize(st));
        for (int k = 0; k < CHUNK && n_open > 0; ++k) { n_open = hist[2 * (k + 1)]; n_nodes = hist[2 * (k + 1) + 1]; ++level; }
    }
    VSB_SAH_CUDA(cudaEventRecord(ev[2], st));
    if (n_nodes == 0 || n_nodes > node_cap || (n_nodes & 1u) == 0u) return VSB_ERR_BAD_TREE;

    // layout in the reference's numbering
    s->n_nodes2 = n_nodes;
    VSB_SAH_CUDA(cudaMalloc((void**)&s->d_nodes2, (size_t)n_nodes * sizeof(vsb::Node32)));
    const uint32_t nb = (n_nodes + 255) / 256;
    vsb::sah_sort_key_kernel<<<nb, 256, 0, st>>>(d_nodes, n_nodes, n, d_keys[0], d_vals[0]);
    {
        size_t tb = tmp_bytes;
        VSB_SAH_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tb, d_keys[0], d_keys[1], d_vals[0], d_vals[1], (int)n_nodes, 0, 64, st));
    }
    vsb::sah_rank_kernel<<<nb, 256, 0, st>>>(d_keys[1], d_vals[1], n_nodes, d_rank);
    vsb::sah_emit_kernel<<<nb, 256, 0, st>>>(d_nodes, n_nodes, n, d_rank, d_refs, s->d_nodes2, s->d_indices);
    g.launches += 3 + 8;
    VSB_SAH_CUDA(cudaGetLastError());
    levels_out = level;
    return VSB_OK;
}

#undef VSB_SAH_CUDA

} // namespace

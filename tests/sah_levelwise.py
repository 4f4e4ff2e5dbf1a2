"""numpy restatement of csrc/build_sah.cuh -- the reference's binned-SAH builder (detail/bvh/sah.h:150-762, driven by
detail/bvh/build.inl:28-140) evaluated BREADTH-FIRST, the way the GPU builder does it -- for the CPU test tier: the three
claims the CUDA builder rests on can be checked without a GPU (tests/test_sah_levelwise.py compares the result with the
host builder's tree, which is byte-identical to the reference's):

  1. everything that decides the tree is order-independent (minima / maxima, integer counts, a fixed cost expression), so
     processing all open nodes of a level at once changes nothing;
  2. std::partition (libstdc++, bidirectional iterators = Hoare's scheme) has a closed form over a prefix sum of the
     predicate: the k-th misplaced element from the left swaps with the k-th misplaced element from the right;
  3. the recursion's node numbering is the order of the key (n - end, first) over the inner nodes (right-first pre-order),
     children at nodes[1 + 2 k], nodes[2 + 2 k]; a leaf over positions [first, end) owns indices[n - end : n - first].

All arithmetic is float32 in the reference's operation order (numpy does not contract)."""
import numpy as np

F = np.float32
FLT_MAX = F(3.4028234663852886e38)
BINS = 16
MAX_LEAF = 4
INT_MIN = -(2 ** 31)


def half_area(lo, hi):
    s = (hi - lo).astype(F)
    s = np.where(F(0) < s, s, F(0)).astype(F)
    return F(F(F(s[0] * s[1]) + F(s[1] * s[2])) + F(s[2] * s[0]))


def projection(cb_lo, cb_hi):
    s = (cb_hi - cb_lo).astype(F)
    a = 0 if s[1] < s[0] else 1
    axis = a if s[2] < s[a] else 2
    if not (s[axis] > 0):
        return None
    k0 = cb_lo[axis]
    k1 = F(F(BINS) / F(cb_hi[axis] - k0))
    return axis, k0, k1


def project_unsafe(k0, k1, c):
    x = (k1 * (c - k0).astype(F)).astype(F)
    with np.errstate(invalid="ignore"):
        ok = (x > F(-2147483648.0)) & (x < F(2147483648.0))
        return np.where(ok, np.trunc(np.where(ok, x, 0)).astype(np.int64), INT_MIN)


def child_open(size, cb_lo, cb_hi):
    return size > MAX_LEAF and projection(cb_lo, cb_hi) is not None


def hoare_partition(pred):
    """permutation p with new[i] = old[p[i]] that std::partition (Hoare) applies to a range with predicate `pred`"""
    m = pred.shape[0]
    t = int(pred.sum())
    perm = np.arange(m)
    left_false = np.flatnonzero(~pred[:t])                     # misplaced on the left, from the left
    right_true = t + np.flatnonzero(pred[t:])[::-1]            # misplaced on the right, from the right
    assert left_false.shape == right_true.shape
    perm[left_false], perm[right_true] = right_true, left_false
    return perm, t


def build(lo, hi):
    """lo, hi: (n, 3) float32 primitive bounds.  -> (nodes: structured array like NODE_DTYPE, indices uint32)"""
    n = lo.shape[0]
    lo = lo.astype(F); hi = hi.astype(F)
    cen = ((hi + lo).astype(F) * F(0.5)).astype(F)
    order = np.arange(n)
    nodes = [dict(lo=lo.min(axis=0), hi=hi.max(axis=0), first=0, end=n, left=None, parent=None)]
    root_cb = (cen.min(axis=0), cen.max(axis=0))
    open_segs = [(0, 0, n, root_cb[0], root_cb[1])] if child_open(n, *root_cb) else []
    levels = 0
    while open_segs:
        levels += 1
        nxt = []
        for node_id, first, end, cb_lo, cb_hi in open_segs:          # any order gives the same result (claim 1)
            nd = nodes[node_id]
            axis, k0, k1 = projection(cb_lo, cb_hi)
            ids = order[first:end]
            unsafe = project_unsafe(k0, k1, cen[ids, axis])
            bins = np.clip(unsafe, 0, BINS - 1)
            blo = np.full((BINS, 3), FLT_MAX, F); bhi = np.full((BINS, 3), -FLT_MAX, F)
            clo = np.full((BINS, 3), FLT_MAX, F); chi = np.full((BINS, 3), -FLT_MAX, F)
            cnt = np.zeros(BINS, np.int64)
            for b in range(BINS):
                sel = ids[bins == b]
                if sel.size:
                    blo[b], bhi[b] = lo[sel].min(axis=0), hi[sel].max(axis=0)
                    clo[b], chi[b] = cen[sel].min(axis=0), cen[sel].max(axis=0)
                    cnt[b] = sel.size
            hsa = half_area(nd["lo"], nd["hi"])
            best_cost, best = FLT_MAX, -1
            with np.errstate(all="ignore"):
                for i in range(BINS - 1, 0, -1):                         # the reference's sweep: strict <, so ties keep the larger i
                    hl = half_area(blo[:i].min(axis=0), bhi[:i].max(axis=0))
                    hr = half_area(blo[i:].min(axis=0), bhi[i:].max(axis=0))
                    cost = F(F(F(1.0) + F(F(hl / hsa) * F(F(3.0) * F(cnt[:i].sum())))) + F(F(hr / hsa) * F(F(3.0) * F(cnt[i:].sum()))))
                    if cost < best_cost:
                        best_cost, best = cost, i
            size = end - first
            if best < 0 or best_cost > F(F(3.0) * F(size)):
                continue                                                # stays a leaf
            perm, t = hoare_partition(unsafe < best)                    # claim 2
            order[first:end] = ids[perm]
            pivot = first + t
            left = len(nodes)
            nd["left"] = left
            nodes.append(dict(lo=blo[:best].min(axis=0), hi=bhi[:best].max(axis=0), first=first, end=pivot, left=None, parent=node_id))
            nodes.append(dict(lo=blo[best:].min(axis=0), hi=bhi[best:].max(axis=0), first=pivot, end=end, left=None, parent=node_id))
            lcb = (clo[:best].min(axis=0), chi[:best].max(axis=0)); rcb = (clo[best:].min(axis=0), chi[best:].max(axis=0))
            if child_open(pivot - first, *lcb):
                nxt.append((left, first, pivot, lcb[0], lcb[1]))
            if child_open(end - pivot, *rcb):
                nxt.append((left + 1, pivot, end, rcb[0], rcb[1]))
        open_segs = nxt
    # numbering (claim 3)
    inner = [i for i, nd in enumerate(nodes) if nd["left"] is not None]
    inner.sort(key=lambda i: (n - nodes[i]["end"], nodes[i]["first"]))
    rank = {i: r for r, i in enumerate(inner)}
    out = np.zeros(len(nodes), dtype=[("bbox_min", "<f4", (3,)), ("first", "<u4"), ("bbox_max", "<f4", (3,)), ("num_prims", "<u4")])
    indices = np.zeros(n, np.uint32)
    for i, nd in enumerate(nodes):
        pos = 0 if nd["parent"] is None else 1 + 2 * rank[nd["parent"]] + (i - nodes[nd["parent"]]["left"])
        out[pos]["bbox_min"], out[pos]["bbox_max"] = nd["lo"], nd["hi"]
        if nd["left"] is not None:
            out[pos]["first"], out[pos]["num_prims"] = 1 + 2 * rank[i], 0
        else:
            out[pos]["first"], out[pos]["num_prims"] = n - nd["end"], nd["end"] - nd["first"]
            indices[n - nd["end"]:n - nd["first"]] = order[nd["first"]:nd["end"]]
    return out, indices, levels

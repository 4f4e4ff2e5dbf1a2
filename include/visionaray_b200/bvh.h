// visionaray_b200/bvh.h -- BVH types of the drop-in boundary + host binned-SAH builder.
//
// Mirrors (names, accessors, memory layout) the reference's
//   bvh_node                         bvh.h:52-119   (32 B, inner <=> num_prims == 0)
//   index_bvh_t / index_bvh_ref_t    bvh.h:189-236, 298-403
//   bvh_t / bvh_ref_t                bvh.h:145-187, 242-292
//   build<Tree>(prims, n, splits)    bvh.h:455, detail/bvh/build.inl:134-178
//   cuda_index_bvh<P> / cuda_bvh<P>  bvh.h:443-448  -> here a handle to a vsb_scene (the
//                                    uploaded + GPU re-encoded tree), see cuda_bvh.h
// The builder restates the reference's binned-SAH algorithm (detail/bvh/sah.h:150-428,
// 615-762: 16 bins, longest centroid axis, leaf <= 4, leaf if split cost > 3*size, right
// subtree first) so that trees built here have the same shape and quality as the
// reference's, with and without spatial splits (SBVH, detail/bvh/sah.h:428-615).
#pragma once
#ifndef VISIONARAY_B200_BVH_H
#define VISIONARAY_B200_BVH_H 1

#include <algorithm>
#include <array>
#include <cassert>
#include <cstring>
#include <stdexcept>
#include <type_traits>
#include <vector>

#include "math.h"

namespace visionaray_b200
{

struct alignas(32) bvh_node
{
    float bbox_min[3];
    union { unsigned first_child; unsigned first_prim; };
    float bbox_max[3];
    unsigned num_prims;

    bool is_inner() const { return num_prims == 0; }
    bool is_leaf() const { return num_prims != 0; }
    aabb const& get_bounds() const { return *reinterpret_cast<aabb const*>(this); }
    unsigned get_child(unsigned i = 0) const { assert(is_inner()); return first_child + i; }
    struct index_range { unsigned first; unsigned last; };
    index_range get_indices() const { assert(is_leaf()); return { first_prim, first_prim + num_prims }; }
    unsigned get_first_primitive() const { return first_prim; }
    unsigned get_num_primitives() const { return num_prims; }

    void set_inner(aabb const& b, unsigned first_child_index)
    {
        std::memcpy(bbox_min, &b.min, sizeof(bbox_min));
        std::memcpy(bbox_max, &b.max, sizeof(bbox_max));
        first_child = first_child_index;
        num_prims = 0;
    }
    void set_leaf(aabb const& b, unsigned first_primitive_index, unsigned count)
    {
        assert(count > 0);
        std::memcpy(bbox_min, &b.min, sizeof(bbox_min));
        std::memcpy(bbox_max, &b.max, sizeof(bbox_max));
        first_prim = first_primitive_index;
        num_prims = count;
    }
};
static_assert(sizeof(bvh_node) == 32, "bvh_node layout");

inline bool is_inner(bvh_node const& n) { return n.is_inner(); }
inline bool is_leaf(bvh_node const& n) { return n.is_leaf(); }

// Non-owning views (PODs handed to kernels in the reference; here handed to the C ABI).
template <typename P>
struct index_bvh_ref_t
{
    using primitive_type = P;
    P const* first_prim; P const* last_prim;
    bvh_node const* first_node; bvh_node const* last_node;
    unsigned const* first_index; unsigned const* last_index;

    index_bvh_ref_t() = default;
    index_bvh_ref_t(P const* p0, P const* p1, bvh_node const* n0, bvh_node const* n1,
                    unsigned const* i0, unsigned const* i1)
        : first_prim(p0), last_prim(p1), first_node(n0), last_node(n1), first_index(i0), last_index(i1) {}

    size_t num_primitives() const { return size_t(last_prim - first_prim); }
    size_t num_nodes() const { return size_t(last_node - first_node); }
    size_t num_indices() const { return size_t(last_index - first_index); }
    bvh_node const& node(size_t i) const { return first_node[i]; }
    P const& primitive(size_t i) const { return first_prim[first_index[i]]; }
};

template <typename P>
struct bvh_ref_t
{
    using primitive_type = P;
    P const* first_prim; P const* last_prim;
    bvh_node const* first_node; bvh_node const* last_node;

    bvh_ref_t() = default;
    bvh_ref_t(P const* p0, P const* p1, bvh_node const* n0, bvh_node const* n1)
        : first_prim(p0), last_prim(p1), first_node(n0), last_node(n1) {}

    size_t num_primitives() const { return size_t(last_prim - first_prim); }
    size_t num_nodes() const { return size_t(last_node - first_node); }
    bvh_node const& node(size_t i) const { return first_node[i]; }
    P const& primitive(size_t i) const { return first_prim[i]; }
};

// Owning containers.
template <typename P>
class index_bvh
{
public:
    using primitive_type = P;
    using bvh_ref = index_bvh_ref_t<P>;
    using tag_type = std::true_type;   // is_index_bvh

    index_bvh() = default;
    index_bvh(P const* prims, size_t count)
        : primitives_(prims, prims + count), nodes_(count == 0 ? 0 : 2 * count - 1), indices_(count) {}

    bvh_ref ref() const
    {
        return bvh_ref(primitives_.data(), primitives_.data() + primitives_.size(),
                       nodes_.data(), nodes_.data() + nodes_.size(),
                       indices_.data(), indices_.data() + indices_.size());
    }

    P const& primitive(size_t i) const { return primitives_[indices_[i]]; }
    bvh_node const& node(size_t i) const { return nodes_[i]; }
    std::vector<P>&              primitives()       { return primitives_; }
    std::vector<P> const&        primitives() const { return primitives_; }
    std::vector<bvh_node>&       nodes()            { return nodes_; }
    std::vector<bvh_node> const& nodes() const      { return nodes_; }
    std::vector<unsigned>&       indices()          { return indices_; }
    std::vector<unsigned> const& indices() const    { return indices_; }
    size_t num_primitives() const { return primitives_.size(); }
    size_t num_nodes() const { return nodes_.size(); }
    size_t num_indices() const { return indices_.size(); }

    void clear(size_t capacity = 0)
    {
        nodes_.clear(); nodes_.reserve(capacity);
        indices_.clear(); indices_.reserve(capacity);
    }

private:
    std::vector<P> primitives_;
    std::vector<bvh_node> nodes_;
    std::vector<unsigned> indices_;
};

template <typename P>
class bvh
{
public:
    using primitive_type = P;
    using bvh_ref = bvh_ref_t<P>;
    using tag_type = std::false_type;

    bvh() = default;
    bvh(P const* prims, size_t count) : primitives_(prims, prims + count), nodes_(count == 0 ? 0 : 2 * count - 1) {}

    bvh_ref ref() const
    {
        return bvh_ref(primitives_.data(), primitives_.data() + primitives_.size(),
                       nodes_.data(), nodes_.data() + nodes_.size());
    }
    P const& primitive(size_t i) const { return primitives_[i]; }
    bvh_node const& node(size_t i) const { return nodes_[i]; }
    std::vector<P>&              primitives()       { return primitives_; }
    std::vector<P> const&        primitives() const { return primitives_; }
    std::vector<bvh_node>&       nodes()            { return nodes_; }
    std::vector<bvh_node> const& nodes() const      { return nodes_; }
    size_t num_primitives() const { return primitives_.size(); }
    size_t num_nodes() const { return nodes_.size(); }
    void clear(size_t capacity = 0) { nodes_.clear(); nodes_.reserve(capacity); }

private:
    std::vector<P> primitives_;
    std::vector<bvh_node> nodes_;
};

template <typename T> struct is_index_bvh : std::false_type {};
template <typename P> struct is_index_bvh<index_bvh<P>> : std::true_type {};
template <typename P> struct is_index_bvh<index_bvh_ref_t<P>> : std::true_type {};
template <typename T> struct is_any_bvh : std::false_type {};
template <typename P> struct is_any_bvh<index_bvh<P>> : std::true_type {};
template <typename P> struct is_any_bvh<index_bvh_ref_t<P>> : std::true_type {};
template <typename P> struct is_any_bvh<bvh<P>> : std::true_type {};
template <typename P> struct is_any_bvh<bvh_ref_t<P>> : std::true_type {};

namespace detail
{

// Where a primitive crosses an axis-aligned plane: bounds of its part on either side
// (split_primitive customisation point, detail/bvh/sah.h:23-128).  Triangles: every edge contributes
// its end points to the side(s) they lie on and, when it crosses, the crossing point (its coordinate
// on the axis set to the plane).
inline void split_edge(aabb& L, aabb& R, vec3 const& a, vec3 const& b, float plane, int axis)
{
    float ta = a[size_t(axis)], tb = b[size_t(axis)];
    if (ta <= plane) L.insert(a);
    if (ta >= plane) R.insert(a);
    if ((ta < plane && plane < tb) || (tb < plane && plane < ta))
    {
        float w = (plane - ta) / (tb - ta);
        vec3 x = a * (1.0f - w) + b * w;              // lerp(a, b, w), math/detail/math.h:459-464
        x[size_t(axis)] = plane;
        L.insert(x);
        R.insert(x);
    }
}

template <typename T>
inline void split_primitive(aabb& L, aabb& R, float plane, int axis, basic_triangle<3, T> const& tri)
{
    vec3 a = tri.v1, b = a + tri.e1, c = a + tri.e2;
    L.invalidate(); R.invalidate();
    split_edge(L, R, a, b, plane, axis);
    split_edge(L, R, b, c, plane, axis);
    split_edge(L, R, c, a, plane, axis);
}

// Spheres: the two caps' bounds (detail/bvh/sah.h:66-126)
template <typename T>
inline void split_primitive(aabb& L, aabb& R, float plane, int axis, basic_sphere<T> const& sph)
{
    vec3 cen = sph.center;
    float rad = sph.radius;
    size_t ax = size_t(axis);
    L.invalidate(); R.invalidate();
    if (plane < cen[ax] - rad)      { R.min = cen - vec3(rad); R.max = cen + vec3(rad); }
    else if (plane > cen[ax] + rad) { L.min = cen - vec3(rad); L.max = cen + vec3(rad); }
    else
    {
        float del = cen[ax] - plane;
        float da = 0.5f * (rad - del), sa = 0.5f * (rad + del);
        float r1 = std::sqrt(rad * rad - del * del), r2 = rad;
        if (del < 0) std::swap(r1, r2);
        vec3 c = cen, h = vec3(r1);
        c[ax] -= sa; h[ax] = da;
        L.insert(c - h); L.insert(c + h);
        c = cen; h = vec3(r2);
        c[ax] += da; h[ax] = sa;
        R.insert(c - h); R.insert(c + h);
    }
}

// generic_primitive: the visitor forwards to the alternative (detail/generic_primitive.inl:72-128)
inline void split_primitive(aabb& L, aabb& R, float plane, int axis, generic_triangle_sphere const& g)
{
    if (g.is_triangle()) split_primitive(L, R, plane, axis, g.as_triangle());
    else                 split_primitive(L, R, plane, axis, g.as_sphere());
}

inline aabb intersect(aabb const& a, aabb const& b)      // math/detail/aabb.inl:168-173
{
    aabb r; r.min = max(a.min, b.min); r.max = min(a.max, b.max);
    return r;
}

// Binned-SAH builder; restates detail/bvh/sah.h: object splits over the centroid bounds and,
// when enabled, spatial splits over the primitive bounds (SBVH, :428-615).
class binned_sah_builder
{
public:
    enum { NumBins = 16 };

    struct prim_ref { aabb bounds; int index; };
    struct leaf_info { aabb prim_bounds; aabb cent_bounds; int first; };

    void enable_spatial_splits(bool enable) { use_spatial_splits_ = enable; }
    void set_alpha(float value) { alpha_ = value; }

    template <typename P>
    leaf_info init(P const* prims, size_t n)
    {
        refs_.resize(n);
        leaf_info root;
        root.prim_bounds.invalidate();
        root.cent_bounds.invalidate();
        root.first = 0;
        for (size_t i = 0; i < n; ++i)
        {
            refs_[i].bounds = get_bounds(prims[i]);
            refs_[i].index = int(i);
            root.prim_bounds.insert(refs_[i].bounds);
            root.cent_bounds.insert(refs_[i].bounds.center());
        }
        sa_threshold_ = alpha_ * (2 * safe_half_surface_area(root.prim_bounds));     // sah.h:651
        return root;
    }

    // append the leaf's indices and drop its refs (refs act as a stack, sah.h:657-672)
    int insert_indices(std::vector<unsigned>& indices, leaf_info const& leaf)
    {
        int leaf_size = int(refs_.size()) - leaf.first;
        for (int i = leaf.first; i != int(refs_.size()); ++i) indices.push_back(unsigned(refs_[i].index));
        refs_.resize(size_t(leaf.first));
        return leaf_size;
    }

    template <typename P>
    bool split(leaf_info childs[2], leaf_info const& leaf, P const* prims, int max_leaf_size)
    {
        int leaf_size = int(refs_.size()) - leaf.first;
        if (leaf_size <= max_leaf_size) return false;

        vec3 size = leaf.cent_bounds.size();
        size_t axis = max_index(size);
        if (size[axis] <= 0.0f) return false;

        // object split: every reference goes to the bin of its centroid (sah.h:374-405)
        projection pr(leaf.cent_bounds, int(axis));
        bin_list bins;
        for (auto& b : bins) b.clear();
        for (size_t r = size_t(leaf.first); r < refs_.size(); ++r)
        {
            vec3 cen = refs_[r].bounds.center();
            bin& b = bins[size_t(pr.project(cen))];
            b.prim_bounds.insert(refs_[r].bounds);
            b.cent_bounds.insert(cen);
            b.enter++;
            b.leave++;
        }
        split_result sr;
        if (!find_split(sr, bins, leaf.prim_bounds)) return false;       // degenerate (NaN costs); the reference asserts

        // spatial split: tried when the children of the object split overlap (sah.h:708-736)
        bool spatial = false;
        if (use_spatial_splits_)
        {
            float sa = 2 * safe_half_surface_area(intersect(sr.prim_bounds[0], sr.prim_bounds[1]));
            if (sa > sa_threshold_)
            {
                size = leaf.prim_bounds.size();
                axis = max_index(size);
                if (size[axis] <= 0.0f) return false;
                projection pr2(leaf.prim_bounds, int(axis));
                for (auto& b : bins) b.clear();
                for (size_t r = size_t(leaf.first); r < refs_.size(); ++r) bin_chopped(bins, refs_[r], pr2, prims);
                split_result sr2;
                if (find_split(sr2, bins, leaf.prim_bounds) && sr2.cost < sr.cost) { spatial = true; pr = pr2; sr = sr2; }
            }
        }

        if (sr.cost > 3.0f * leaf_size) return false;       // a leaf is cheaper

        if (spatial) { partition_spatial(childs, sr, leaf, pr, prims); return true; }

        childs[0].prim_bounds = sr.prim_bounds[0]; childs[0].cent_bounds = sr.cent_bounds[0];
        childs[1].prim_bounds = sr.prim_bounds[1]; childs[1].cent_bounds = sr.cent_bounds[1];
        auto pivot = std::partition(refs_.begin() + leaf.first, refs_.end(),
                                    [&](prim_ref const& x) { return pr.project_unsafe(x.bounds.center()) < sr.index; });
        childs[0].first = leaf.first;
        childs[1].first = int(pivot - refs_.begin());
        return true;
    }

private:
    struct bin
    {
        aabb prim_bounds; aabb cent_bounds; int enter; int leave;
        void clear() { prim_bounds.invalidate(); cent_bounds.invalidate(); enter = 0; leave = 0; }
    };
    using bin_list = std::array<bin, NumBins>;

    // bin index of a point along the split axis, and the left plane of a bin (sah.h:230-275)
    struct projection
    {
        float k0, k1; int axis;
        projection(aabb const& bounds, int ax) : k0(bounds.min[size_t(ax)]), k1(float(NumBins) / (bounds.max[size_t(ax)] - k0)), axis(ax) {}
        int project_unsafe(vec3 const& v) const { return static_cast<int>(k1 * (v[size_t(axis)] - k0)); }
        int project(vec3 const& v) const { int i = project_unsafe(v); return i < 0 ? 0 : (i > NumBins - 1 ? NumBins - 1 : i); }
        float unproject(int i) const { return float(i) / k1 + k0; }
    };

    struct split_result { aabb prim_bounds[2]; aabb cent_bounds[2]; float cost; int index; };

    static size_t max_index(vec3 const& v)          // math/detail/vector.inl:811-821
    {
        size_t a = v[1] < v[0] ? 0 : 1;
        return v[2] < v[a] ? a : 2;
    }

    // sweep over the bin boundaries for the cheapest one (sah.h:305-368)
    static bool find_split(split_result& sr, bin_list const& bins, aabb const& bounds)
    {
        auto merge = [](bin l, bin const& r)
        {
            l.prim_bounds.insert(r.prim_bounds); l.cent_bounds.insert(r.cent_bounds);
            l.enter += r.enter; l.leave += r.leave;
            return l;
        };
        float hsa_parent = safe_half_surface_area(bounds);
        bin_list acc_l, acc_r;
        acc_l[0] = bins[0];
        for (int i = 1; i < NumBins; ++i) acc_l[size_t(i)] = merge(acc_l[size_t(i - 1)], bins[size_t(i)]);
        acc_r[NumBins - 1] = bins[NumBins - 1];
        float best_cost = FLT_MAX;
        int best_index = -1;
        for (int i = NumBins - 1; i > 0; --i)
        {
            acc_r[size_t(i - 1)] = merge(acc_r[size_t(i)], bins[size_t(i - 1)]);
            bin const& L = acc_l[size_t(i - 1)];
            bin const& R = acc_r[size_t(i)];
            float hl = safe_half_surface_area(L.prim_bounds);
            float hr = safe_half_surface_area(R.prim_bounds);
            float cost = 1.0f + (hl / hsa_parent) * (3.0f * L.enter) + (hr / hsa_parent) * (3.0f * R.leave);
            if (cost < best_cost) { best_cost = cost; best_index = i; }
        }
        if (best_index < 0) return false;
        bin const& L = acc_l[size_t(best_index - 1)];
        bin const& R = acc_r[size_t(best_index)];
        sr.prim_bounds[0] = L.prim_bounds; sr.cent_bounds[0] = L.cent_bounds;
        sr.prim_bounds[1] = R.prim_bounds; sr.cent_bounds[1] = R.cent_bounds;
        sr.cost = best_cost; sr.index = best_index;
        return true;
    }

    // spatial binning: the reference is chopped at every bin plane it spans; each piece's bounds
    // (clipped to what is left of the reference) go to the piece's bin; it "enters" in its first bin
    // and "leaves" in its last (sah.h:446-487)
    template <typename P>
    static void bin_chopped(bin_list& bins, prim_ref const& ref, projection const& pr, P const* prims)
    {
        int imin = pr.project(ref.bounds.min), imax = pr.project(ref.bounds.max);
        aabb rest = ref.bounds;
        for (int i = imin; i < imax; ++i)
        {
            aabb L, R;
            split_primitive(L, R, pr.unproject(i + 1), pr.axis, prims[ref.index]);
            L = intersect(L, rest);
            rest = intersect(R, rest);
            bins[size_t(i)].prim_bounds.insert(L);
            bins[size_t(i)].cent_bounds.insert(L.center());
        }
        bins[size_t(imax)].prim_bounds.insert(rest);
        bins[size_t(imax)].cent_bounds.insert(rest.center());
        bins[size_t(imin)].enter++;
        bins[size_t(imax)].leave++;
    }

    // references left of the plane move to the front, those right of it stay, a straddling one is
    // replaced by its right part and its left part is appended, then swapped to the front block
    // (sah.h:489-615; no reference unsplitting there either)
    template <typename P>
    void partition_spatial(leaf_info childs[2], split_result const& sr, leaf_info const& leaf, projection const& pr, P const* prims)
    {
        float plane = pr.unproject(sr.index);
        size_t ax = size_t(pr.axis);
        int pivot = leaf.first, i = leaf.first, last = int(refs_.size());
        for (int k = 0; k < 2; ++k) { childs[k].prim_bounds.invalidate(); childs[k].cent_bounds.invalidate(); }
        for (; i != last; ++i)
        {
            prim_ref& cur = refs_[size_t(i)];
            if (cur.bounds.max[ax] <= plane)
            {
                childs[0].prim_bounds.insert(cur.bounds);
                childs[0].cent_bounds.insert(cur.bounds.center());
                if (i != pivot) std::swap(refs_[size_t(pivot)], cur);
                ++pivot;
            }
            else if (cur.bounds.min[ax] >= plane)
            {
                childs[1].prim_bounds.insert(cur.bounds);
                childs[1].cent_bounds.insert(cur.bounds.center());
            }
            else
            {
                prim_ref L, R;
                split_primitive(L.bounds, R.bounds, plane, pr.axis, prims[cur.index]);
                L.bounds = intersect(L.bounds, cur.bounds);
                R.bounds = intersect(R.bounds, cur.bounds);
                L.index = R.index = cur.index;
                childs[0].prim_bounds.insert(L.bounds);
                childs[0].cent_bounds.insert(L.bounds.center());
                childs[1].prim_bounds.insert(R.bounds);
                childs[1].cent_bounds.insert(R.bounds.center());
                refs_[size_t(i)] = R;
                refs_.push_back(L);                      // may reallocate: `cur` is dead from here on
                std::swap(refs_[size_t(pivot)], refs_.back());
                ++pivot;
            }
        }
        childs[0].first = leaf.first;
        childs[1].first = pivot;
    }

    std::vector<prim_ref> refs_;
    float sa_threshold_ = 1.0e+38f, alpha_ = 1.0e-5f;
    bool use_spatial_splits_ = false;
};

// detail/bvh/build.inl:28-81, made iterative (explicit work stack) so that 20M-primitive
// trees cannot overflow the C++ call stack; the visiting order is the same (right subtree
// fully built before the left one).
template <typename P>
inline void build_tree_impl(std::vector<bvh_node>& nodes, std::vector<unsigned>& indices,
                            binned_sah_builder& b, binned_sah_builder::leaf_info const& root, P const* prims, int max_leaf_size)
{
    struct job { int index; binned_sah_builder::leaf_info leaf; };
    std::vector<job> work;
    work.push_back({ 0, root });
    while (!work.empty())
    {
        job j = work.back();
        work.pop_back();
        binned_sah_builder::leaf_info childs[2];
        if (b.split(childs, j.leaf, prims, max_leaf_size))
        {
            int first_child = int(nodes.size());
            nodes[size_t(j.index)].set_inner(j.leaf.prim_bounds, unsigned(first_child));
            nodes.emplace_back();
            nodes.emplace_back();
            work.push_back({ first_child + 0, childs[0] });   // popped second
            work.push_back({ first_child + 1, childs[1] });   // popped first: right subtree first
        }
        else
        {
            unsigned first = unsigned(indices.size());
            int count = b.insert_indices(indices, j.leaf);
            nodes[size_t(j.index)].set_leaf(j.leaf.prim_bounds, first, unsigned(count));
        }
    }
}

} // namespace detail

// build<index_bvh<P>>(prims, n[, spatial_splits]) -- bvh.h:455, detail/bvh/build.inl:165-178
// (index_bvh only: use reorder() for bvh<P>; build<b200_index_bvh<P>>() in scheduler.h builds on the GPU)
template <typename Tree, typename P>
typename std::enable_if<is_index_bvh<Tree>::value, Tree>::type build(P const* prims, size_t n, bool spatial_splits = false)
{
    Tree tree(prims, n);
    detail::binned_sah_builder builder;
    builder.enable_spatial_splits(spatial_splits);
    builder.set_alpha(1.0e-5f);
    auto root = builder.init(prims, n);
    tree.clear(2 * (n / 4));
    if (n == 0) return tree;
    tree.nodes().emplace_back();
    detail::build_tree_impl(tree.nodes(), tree.indices(), builder, root, prims, 4);
    return tree;
}

// index_bvh -> bvh: primitives re-ordered into leaf order (detail/bvh/build.inl:102-132)
template <typename P>
bvh<P> reorder(index_bvh<P> const& it)
{
    std::vector<P> prims(it.num_indices());
    for (size_t i = 0; i < it.num_indices(); ++i) prims[i] = it.primitives()[it.indices()[i]];
    bvh<P> out(prims.data(), prims.size());
    out.nodes() = it.nodes();
    return out;
}

} // namespace visionaray_b200

#endif

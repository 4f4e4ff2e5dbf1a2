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
// reference's; spatial splits (SBVH) are not implemented.
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

// Binned-SAH builder; restates detail/bvh/sah.h (object splits only).
class binned_sah_builder
{
public:
    enum { NumBins = 16 };

    struct prim_ref { aabb bounds; int index; };
    struct leaf_info { aabb prim_bounds; aabb cent_bounds; int first; };

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

    bool split(leaf_info childs[2], leaf_info const& leaf, int max_leaf_size)
    {
        int leaf_size = int(refs_.size()) - leaf.first;
        if (leaf_size <= max_leaf_size) return false;

        vec3 size = leaf.cent_bounds.size();
        size_t axis = size[1] < size[0] ? 0 : 1;          // max_index, math/detail/vector.inl:811-821
        axis = size[2] < size[axis] ? axis : 2;
        if (size[axis] <= 0.0f) return false;

        float k0 = leaf.cent_bounds.min[axis];
        float k1 = float(NumBins) / (leaf.cent_bounds.max[axis] - k0);
        auto project_unsafe = [&](vec3 const& v) { return static_cast<int>(k1 * (v[axis] - k0)); };

        struct bin { aabb prim_bounds; aabb cent_bounds; int enter; int leave; };
        std::array<bin, NumBins> bins;
        for (auto& b : bins) { b.prim_bounds.invalidate(); b.cent_bounds.invalidate(); b.enter = 0; b.leave = 0; }

        for (size_t r = size_t(leaf.first); r < refs_.size(); ++r)
        {
            vec3 cen = refs_[r].bounds.center();
            int i = project_unsafe(cen);
            i = i < 0 ? 0 : (i > NumBins - 1 ? NumBins - 1 : i);
            bins[size_t(i)].prim_bounds.insert(refs_[r].bounds);
            bins[size_t(i)].cent_bounds.insert(cen);
            bins[size_t(i)].enter++;
            bins[size_t(i)].leave++;
        }

        auto merge = [](bin l, bin const& r)
        {
            l.prim_bounds.insert(r.prim_bounds); l.cent_bounds.insert(r.cent_bounds);
            l.enter += r.enter; l.leave += r.leave;
            return l;
        };

        float hsa_parent = safe_half_surface_area(leaf.prim_bounds);
        std::array<bin, NumBins> acc_l, acc_r;
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
        if (best_index < 0) return false;            // degenerate (NaN costs); the reference asserts
        if (best_cost > 3.0f * leaf_size) return false;

        bin const& L = acc_l[size_t(best_index - 1)];
        bin const& R = acc_r[size_t(best_index)];
        childs[0].prim_bounds = L.prim_bounds; childs[0].cent_bounds = L.cent_bounds;
        childs[1].prim_bounds = R.prim_bounds; childs[1].cent_bounds = R.cent_bounds;

        auto pivot = std::partition(refs_.begin() + leaf.first, refs_.end(),
                                    [&](prim_ref const& x) { return project_unsafe(x.bounds.center()) < best_index; });
        childs[0].first = leaf.first;
        childs[1].first = int(pivot - refs_.begin());
        return true;
    }

private:
    std::vector<prim_ref> refs_;
};

// detail/bvh/build.inl:28-81, made iterative (explicit work stack) so that 20M-primitive
// trees cannot overflow the C++ call stack; the visiting order is the same (right subtree
// fully built before the left one).
inline void build_tree_impl(std::vector<bvh_node>& nodes, std::vector<unsigned>& indices,
                            binned_sah_builder& b, binned_sah_builder::leaf_info const& root, int max_leaf_size)
{
    struct job { int index; binned_sah_builder::leaf_info leaf; };
    std::vector<job> work;
    work.push_back({ 0, root });
    while (!work.empty())
    {
        job j = work.back();
        work.pop_back();
        binned_sah_builder::leaf_info childs[2];
        if (b.split(childs, j.leaf, max_leaf_size))
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
    if (spatial_splits) throw std::runtime_error("visionaray_b200::build: spatial splits are not supported");
    Tree tree(prims, n);
    detail::binned_sah_builder builder;
    auto root = builder.init(prims, n);
    tree.clear(2 * (n / 4));
    if (n == 0) return tree;
    tree.nodes().emplace_back();
    detail::build_tree_impl(tree.nodes(), tree.indices(), builder, root, 4);
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

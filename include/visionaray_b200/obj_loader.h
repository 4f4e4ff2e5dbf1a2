// visionaray_b200/obj_loader.h -- Wavefront OBJ + MTL ingest into the reference's `model` arrays, without Boost.
//
// Replaces load_obj(filename, model&) (src/common/obj_loader.cpp:299-533 of the reference; grammar
// src/common/obj_grammar.cpp:33-60; Boost.Spirit / Boost.Iostreams / Boost.Filesystem) and class model
// (src/common/model.h:20-48), so that a reference user's scene -- not only synthetic ones -- reaches
// build<index_bvh<P>>() / b200_index_bvh / b200_sched::frame().  The RULES are the reference's, statement for statement:
//   * faces are triangle fans over (f0, f[k-1], f[k]); negative indices count from the end (remap_index, :54-58)
//   * a triangle whose |cross(e1, e2)| is 0 is rejected (store_triangle, :65-88); prim_id = position in `primitives`,
//     geom_id = materials.size() - 1 at the time the face is read (0 while there is no material)
//   * tex coords / shading normals are appended (3 per triangle) only when all three corners carry the index (:115-141);
//     at the end tex_coords is padded with zeros up to 3 per triangle, shading_normals is NOT (reference quirk, kept:
//     a file whose faces carry normals only in part yields a misaligned array in the reference too)
//   * usemtl NAME appends a plastic material (ca = Ka, cd = Kd, cs = Ks, ka = kd = ks = 1, exponent = Ns; mtl entries start
//     from the default gray material) and one texture ref per material (map_Kd, else an empty 0 x 0 ref); an unknown
//     name adds nothing (warning)
//   * at the end: geometric normals normalize(cross(e1, e2)), default materials / empty textures up to geom_id, bbox
// Deviations, stated: the reference pads tex_coords with `for (i = tex_coords.size(); i < primitives.size(); ++i)` (:512-517),
// which leaves the array SHORTER than 3 x primitives when only some faces carry tex coords (out-of-bounds reads later);
// here it is always padded to 3 x primitives.  Face indices beyond the arrays are refused, not dereferenced.
// Numbers are converted with strtof (correctly rounded; Spirit's float_ may differ in the last
// bit on long mantissas); map_Kd images are read from binary PNM (P6 / P5) only -- the reference's image class needs
// libjpeg / libpng / libtiff -- every other file leaves the empty texture ref and a warning.  Parity of the arrays
// is pinned against a restatement of these rules in the tests (tests/test_obj_loader.py); the reference's own loader
// cannot be built here (Boost is absent).
#pragma once
#ifndef VISIONARAY_B200_OBJ_LOADER_H
#define VISIONARAY_B200_OBJ_LOADER_H 1

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <string>
#include <vector>

#include "scheduler.h"

namespace visionaray_b200
{

// texture<vector<4, unorm<8>>, 2>: owns its texels (texture/detail/texture2d.h)
struct texture2d
{
    std::vector<unorm8x4> texels;
    size_t width = 0, height = 0;
    texture_ref2d ref() const
    {
        texture_ref2d r(width, height);
        r.reset(texels.data());
        r.set_address_mode(Wrap);
        r.set_filter_mode(Linear);
        return r;
    }
};

// src/common/model.h:20-48
class model
{
public:
    using triangle_type  = basic_triangle<3, float>;
    using normal_type    = vec3;
    using tex_coord_type = vec2;
    using material_type  = plastic<float>;
    using texture_type   = texture2d;

    std::vector<triangle_type>       primitives;
    std::vector<normal_type>         shading_normals;
    std::vector<normal_type>         geometric_normals;
    std::vector<tex_coord_type>      tex_coords;
    std::vector<material_type>       materials;
    std::map<std::string, texture2d> texture_map;
    std::vector<texture_ref2d>       textures;
    aabb                             bbox;
    std::vector<std::string>         warnings;      // what the reference prints to std::cerr
};

namespace detail
{

inline plastic<float> make_default_material()       // obj_loader.cpp:36-47
{
    plastic<float> m;
    m.set_ca(vec3(0.2f, 0.2f, 0.2f)); m.set_cd(vec3(0.8f, 0.8f, 0.8f)); m.set_cs(vec3(0.1f, 0.1f, 0.1f));
    m.set_ka(1.0f); m.set_kd(1.0f); m.set_ks(1.0f);
    m.set_specular_exp(32.0f);
    return m;
}

struct obj_mtl                                      // obj_loader.cpp:167-186
{
    vec3 ka = vec3(0.2f, 0.2f, 0.2f), kd = vec3(0.8f, 0.8f, 0.8f), ke = vec3(0.0f), ks = vec3(0.1f, 0.1f, 0.1f);
    float ns = 32.0f;
    std::string map_kd;
};

struct obj_face_index { int v = 0; bool has_t = false, has_n = false; int t = 0, n = 0; };

inline std::string obj_dirname(std::string const& path)
{
    size_t p = path.find_last_of('/');
    return p == std::string::npos ? std::string() : path.substr(0, p);
}

inline bool obj_file_exists(std::string const& path) { std::ifstream f(path.c_str(), std::ios::binary); return f.good(); }

// a line split into its keyword and the rest, blanks (space, tab) trimmed on both sides, '\r' dropped
inline void obj_split(std::string const& line, std::string& key, std::string& rest)
{
    size_t a = 0, n = line.size();
    while (n > 0 && (line[n - 1] == '\r' || line[n - 1] == ' ' || line[n - 1] == '\t')) --n;
    while (a < n && (line[a] == ' ' || line[a] == '\t')) ++a;
    size_t b = a;
    while (b < n && line[b] != ' ' && line[b] != '\t') ++b;
    key = line.substr(a, b - a);
    while (b < n && (line[b] == ' ' || line[b] == '\t')) ++b;
    rest = line.substr(b, n - b);
}

// up to `max` floats from `s`; returns how many were read
inline int obj_floats(std::string const& s, float* out, int max)
{
    char const* p = s.c_str();
    int k = 0;
    while (k < max)
    {
        char* end = nullptr;
        float f = std::strtof(p, &end);
        if (end == p) break;
        out[k++] = f;
        p = end;
    }
    return k;
}

// "v", "v/t", "v//n", "v/t/n" (r_face_idx, obj_grammar.cpp:56)
inline bool obj_face_token(char const*& p, obj_face_index& out)
{
    while (*p == ' ' || *p == '\t') ++p;
    char* end = nullptr;
    long v = std::strtol(p, &end, 10);
    if (end == p) return false;
    out = obj_face_index();
    out.v = int(v);
    p = end;
    if (*p == '/')
    {
        ++p;
        long t = std::strtol(p, &end, 10);
        if (end != p) { out.has_t = true; out.t = int(t); p = end; }
        if (*p == '/')
        {
            ++p;
            long n = std::strtol(p, &end, 10);
            if (end != p) { out.has_n = true; out.n = int(n); p = end; }
        }
    }
    return true;
}

template <typename Int> inline Int obj_remap_index(Int idx, Int size) { return idx < 0 ? size + idx : idx - 1; }   // obj_loader.cpp:54-58

} // namespace detail
} // namespace visionaray_b200

#include "obj_loader.inl"

#endif // VISIONARAY_B200_OBJ_LOADER_H

// scenes.h -- seeded synthetic scene and ray generators.  Host C++.
//
// One generator feeds the oracle, the reference harness and the GPU path with identical
// arrays.  RNG: SplitMix64 -> float in [0,1) from the top 24 bits.  Triangles are stored the
// reference's way (v1, e1 = v2 - v1, e2 = v3 - v1; math/triangle.h), prim_id = index,
// geom_id = 0 unless stated.
#pragma once

#include <cmath>
#include <cstdint>
#include <vector>

#include "visionaray_b200/math.h"

namespace visionaray_b200
{
namespace scenes
{

using triangle = basic_triangle<3, float>;
using sphere   = basic_sphere<float>;

struct splitmix64
{
    uint64_t s;
    explicit splitmix64(uint64_t seed) : s(seed) {}
    uint64_t next_u64()
    {
        uint64_t z = (s += 0x9E3779B97F4A7C15ull);
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        return z ^ (z >> 31);
    }
    float next() { return float(next_u64() >> 40) * (1.0f / 16777216.0f); }
    float range(float a, float b) { return a + (b - a) * next(); }
};

inline triangle make_tri(vec3 const& a, vec3 const& b, vec3 const& c, unsigned prim_id, unsigned geom_id = 0)
{
    triangle t(a, b - a, c - a);
    t.prim_id = prim_id;
    t.geom_id = geom_id;
    return t;
}

// cfg1: UV sphere, `slices` x `stacks`; 2*slices*(stacks-1) triangles (224 x 224 -> 99,904)
inline void uv_sphere(std::vector<triangle>& out, int slices, int stacks, float radius)
{
    out.clear();
    const double pi = 3.14159265358979323846;
    auto vert = [&](int i, int j)
    {
        double th = pi * double(i) / double(stacks);
        double ph = 2.0 * pi * double(j % slices) / double(slices);
        return vec3(float(radius * std::sin(th) * std::cos(ph)), float(radius * std::cos(th)),
                    float(radius * std::sin(th) * std::sin(ph)));
    };
    for (int i = 0; i < stacks; ++i)
        for (int j = 0; j < slices; ++j)
        {
            vec3 a = vert(i, j), b = vert(i + 1, j), c = vert(i + 1, j + 1), d = vert(i, j + 1);
            if (i != stacks - 1) out.push_back(make_tri(a, b, c, unsigned(out.size())));
            if (i != 0)          out.push_back(make_tri(a, c, d, unsigned(out.size())));
        }
}

// seeded value-noise lattice + 4-octave fbm
struct value_noise
{
    enum { N = 64 };
    float lattice[N * N];
    explicit value_noise(uint64_t seed)
    {
        splitmix64 rng(seed);
        for (int i = 0; i < N * N; ++i) lattice[i] = rng.next() * 2.0f - 1.0f;
    }
    float at(int x, int y) const { return lattice[(y & (N - 1)) * N + (x & (N - 1))]; }
    float sample(float x, float y) const
    {
        float fx = std::floor(x), fy = std::floor(y);
        int ix = int(fx), iy = int(fy);
        float tx = x - fx, ty = y - fy;
        tx = tx * tx * (3.0f - 2.0f * tx);
        ty = ty * ty * (3.0f - 2.0f * ty);
        float a = at(ix, iy), b = at(ix + 1, iy), c = at(ix, iy + 1), d = at(ix + 1, iy + 1);
        return (a + (b - a) * tx) + ((c + (d - c) * tx) - (a + (b - a) * tx)) * ty;
    }
    float fbm(float x, float y) const
    {
        float sum = 0.0f, amp = 0.5f, f = 4.0f;
        for (int o = 0; o < 4; ++o) { sum += amp * sample(x * f, y * f); amp *= 0.5f; f *= 2.0f; }
        return sum;
    }
};

// cfg2 / cfg4: displaced grid, n x n quads over [-1,1]^2 in xz, y = amp * fbm(x,z);
// 2*n*n triangles (708 -> 1,002,528; 2236 -> 9,999,392)
inline void displaced_grid(std::vector<triangle>& out, int n, uint64_t seed, float amp = 0.15f)
{
    value_noise noise(seed);
    out.clear();
    out.reserve(size_t(n) * size_t(n) * 2);
    std::vector<float> h(size_t(n + 1) * size_t(n + 1));
    for (int j = 0; j <= n; ++j)
        for (int i = 0; i <= n; ++i)
        {
            float x = -1.0f + 2.0f * float(i) / float(n), z = -1.0f + 2.0f * float(j) / float(n);
            h[size_t(j) * size_t(n + 1) + size_t(i)] = amp * noise.fbm(x * 0.5f + 0.5f, z * 0.5f + 0.5f);
        }
    auto vert = [&](int i, int j)
    {
        float x = -1.0f + 2.0f * float(i) / float(n), z = -1.0f + 2.0f * float(j) / float(n);
        return vec3(x, h[size_t(j) * size_t(n + 1) + size_t(i)], z);
    };
    for (int j = 0; j < n; ++j)
        for (int i = 0; i < n; ++i)
        {
            vec3 a = vert(i, j), b = vert(i + 1, j), c = vert(i + 1, j + 1), d = vert(i, j + 1);
            out.push_back(make_tri(a, c, b, unsigned(out.size())));
            out.push_back(make_tri(a, d, c, unsigned(out.size())));
        }
}

// cfg3: axis-aligned cubes, flattened (no instancing in the reference): 12 triangles each;
// centres uniform in [-1,1]^3, edge in [0.01, 0.02] (416,667 cubes -> 5,000,004 triangles)
inline void cube_field(std::vector<triangle>& out, int n_cubes, uint64_t seed, float edge_lo = 0.01f, float edge_hi = 0.02f)
{
    splitmix64 rng(seed);
    out.clear();
    out.reserve(size_t(n_cubes) * 12);
    static const int faces[6][4] = { { 0, 2, 3, 1 }, { 4, 5, 7, 6 }, { 0, 1, 5, 4 }, { 2, 6, 7, 3 }, { 0, 4, 6, 2 }, { 1, 3, 7, 5 } };
    for (int k = 0; k < n_cubes; ++k)
    {
        vec3 c(rng.range(-1.0f, 1.0f), rng.range(-1.0f, 1.0f), rng.range(-1.0f, 1.0f));
        float hh = 0.5f * rng.range(edge_lo, edge_hi);
        vec3 v[8];
        for (int i = 0; i < 8; ++i)
            v[i] = vec3(c.x + ((i & 1) ? hh : -hh), c.y + ((i & 2) ? hh : -hh), c.z + ((i & 4) ? hh : -hh));
        for (auto const& f : faces)
        {
            out.push_back(make_tri(v[f[0]], v[f[1]], v[f[2]], unsigned(out.size())));
            out.push_back(make_tri(v[f[0]], v[f[2]], v[f[3]], unsigned(out.size())));
        }
    }
}

// cfg5: triangle soup; centre uniform in [0,1]^3, two edge vectors uniform in [-s,s]^3
inline void soup(std::vector<triangle>& out, size_t n, uint64_t seed, float s = 0.004f)
{
    splitmix64 rng(seed);
    out.resize(n);
    for (size_t i = 0; i < n; ++i)
    {
        vec3 c(rng.next(), rng.next(), rng.next());
        vec3 e1(rng.range(-s, s), rng.range(-s, s), rng.range(-s, s));
        vec3 e2(rng.range(-s, s), rng.range(-s, s), rng.range(-s, s));
        triangle t(c - (e1 + e2) * (1.0f / 3.0f), e1, e2);
        t.prim_id = unsigned(i);
        t.geom_id = 0;
        out[i] = t;
    }
}

// random spheres (basic_sphere leaves): centres uniform in [-1,1]^3, radius in [r_lo, r_hi]
inline void sphere_cloud(std::vector<sphere>& out, size_t n, uint64_t seed, float r_lo = 0.005f, float r_hi = 0.05f)
{
    splitmix64 rng(seed);
    out.resize(n);
    for (size_t i = 0; i < n; ++i)
    {
        sphere s(vec3(rng.range(-1.0f, 1.0f), rng.range(-1.0f, 1.0f), rng.range(-1.0f, 1.0f)), rng.range(r_lo, r_hi));
        s.prim_id = unsigned(i);
        s.geom_id = 0;
        out[i] = s;
    }
}

// cfg5 rays: origin uniform in [lo,hi]^3, direction uniform on the sphere (z = 1-2u, phi = 2 pi v)
inline void random_rays(ray* out, size_t n, uint64_t seed, float lo = 0.0f, float hi = 1.0f)
{
    splitmix64 rng(seed);
    for (size_t i = 0; i < n; ++i)
    {
        vec3 o(rng.range(lo, hi), rng.range(lo, hi), rng.range(lo, hi));
        float z = 1.0f - 2.0f * rng.next();
        float phi = 6.28318530717958647692f * rng.next();
        float r = std::sqrt(0.0f < 1.0f - z * z ? 1.0f - z * z : 0.0f);
        out[i].ori = o;
        out[i].dir = vec3(r * std::cos(phi), r * std::sin(phi), z);
    }
}

// per-face geometric normals normalize(cross(e1, e2)) (what the viewer's model carries as
// geometric_normals and binds with normals_per_face_binding, src/viewer/viewer.cpp:507-513)
inline void face_normals(std::vector<vec3>& out, std::vector<triangle> const& tris)
{
    out.resize(tris.size());
    for (size_t i = 0; i < tris.size(); ++i) out[i] = normalize(cross(tris[i].e1, tris[i].e2));
}

} // namespace scenes
} // namespace visionaray_b200

// visionaray_b200/math.h -- host-side POD math types of the drop-in boundary.
//
// Memory layouts are the ABI and match the reference's
//   vector<3,float>   16-byte aligned, 16 B   (math/detail/vector3f.inl:12)
//   basic_ray<float>  32 B                    (math/ray.h)
//   basic_aabb<float> 32 B                    (math/aabb.h)
//   basic_triangle<3,float,unsigned> 64 B     (math/triangle.h, math/primitive.h)
//   basic_sphere<float,unsigned>     48 B     (math/sphere.h)
// so arrays of the reference's types can be handed to this library unchanged
// (reinterpret_cast) and vice versa.  Only what the ray-query path needs is provided.
#pragma once
#ifndef VISIONARAY_B200_MATH_H
#define VISIONARAY_B200_MATH_H 1

#include <cfloat>
#include <cstring>
#include <cmath>
#include <cstddef>
#include <cstdint>

namespace visionaray_b200
{

struct vec2 { float x, y; };
struct vec2ui { unsigned x, y; };

struct alignas(16) vec3
{
    float x, y, z;

    vec3() = default;
    vec3(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    explicit vec3(float const* d) : x(d[0]), y(d[1]), z(d[2]) {}

    float&       operator[](size_t i)       { return (&x)[i]; }
    float const& operator[](size_t i) const { return (&x)[i]; }
    float*       data()       { return &x; }
    float const* data() const { return &x; }
};

struct alignas(16) vec4
{
    float x, y, z, w;
    vec4() = default;
    vec4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    vec4(vec3 const& v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    float*       data()       { return &x; }
    float const* data() const { return &x; }
};

// matrix<4,4,float> (math/matrix.h:74-130): four columns, 16 floats column-major; (row, col) access
struct alignas(16) mat4
{
    float m[16];
    mat4() = default;
    explicit mat4(float const* d) { for (int i = 0; i < 16; ++i) m[i] = d[i]; }
    mat4(float m00, float m10, float m20, float m30, float m01, float m11, float m21, float m31,
         float m02, float m12, float m22, float m32, float m03, float m13, float m23, float m33)
        : m{ m00, m10, m20, m30, m01, m11, m21, m31, m02, m12, m22, m32, m03, m13, m23, m33 } {}
    static mat4 identity() { return mat4(1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1); }
    float&       operator()(size_t row, size_t col)       { return m[col * 4 + row]; }
    float const& operator()(size_t row, size_t col) const { return m[col * 4 + row]; }
    float*       data()       { return m; }
    float const* data() const { return m; }
};

struct recti { int x, y, w, h; recti() = default; recti(int x_, int y_, int w_, int h_) : x(x_), y(y_), w(w_), h(h_) {} };

// Operation order follows the reference (math/detail/vector3.inl:293-336) so that host-side
// quantities derived here (camera basis, normals, bounds) are bit-identical to the
// reference's when built without FMA contraction.
inline vec3 operator+(vec3 const& a, vec3 const& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 const& a, vec3 const& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 const& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 const& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 const& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator*(vec3 const& a, vec3 const& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float dot(vec3 const& a, vec3 const& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 const& u, vec3 const& v)
{
    return vec3(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
inline float length(vec3 const& v) { return std::sqrt(dot(v, v)); }
inline vec3 normalize(vec3 const& v) { return v * (1.0f / std::sqrt(dot(v, v))); }
inline vec3 min(vec3 const& a, vec3 const& b)
{
    return vec3(a.x < b.x ? a.x : b.x, a.y < b.y ? a.y : b.y, a.z < b.z ? a.z : b.z);
}
inline vec3 max(vec3 const& a, vec3 const& b)
{
    return vec3(a.x < b.x ? b.x : a.x, a.y < b.y ? b.y : a.y, a.z < b.z ? b.z : a.z);
}

struct alignas(16) aabb
{
    vec3 min;
    vec3 max;

    void invalidate() { min = vec3(FLT_MAX); max = vec3(-FLT_MAX); }
    bool invalid() const { return min.x > max.x || min.y > max.y || min.z > max.z; }
    void insert(vec3 const& v) { min = visionaray_b200::min(min, v); max = visionaray_b200::max(max, v); }
    void insert(aabb const& b) { min = visionaray_b200::min(min, b.min); max = visionaray_b200::max(max, b.max); }
    vec3 center() const { return (max + min) * 0.5f; }
    vec3 size() const { return max - min; }
    vec3 safe_size() const
    {
        vec3 s = max - min;
        s.x = 0.0f < s.x ? s.x : 0.0f; s.y = 0.0f < s.y ? s.y : 0.0f; s.z = 0.0f < s.z ? s.z : 0.0f;
        return s;
    }
};

inline float safe_half_surface_area(aabb const& b)
{
    vec3 s = b.safe_size();
    return s.x * s.y + s.y * s.z + s.z * s.x;
}

template <typename T = float>
struct basic_ray
{
    using scalar_type = T;
    vec3 ori;
    vec3 dir;
    basic_ray() = default;
    basic_ray(vec3 const& o, vec3 const& d) : ori(o), dir(d) {}
};
using ray = basic_ray<float>;

// primitive<unsigned>: {geom_id, prim_id} (math/primitive.h)
struct primitive_ids { unsigned geom_id; unsigned prim_id; };

template <size_t Dim = 3, typename T = float, typename P = unsigned>
struct basic_triangle : primitive_ids
{
    static_assert(Dim == 3, "only 3-D triangles are on the path");
    vec3 v1;
    vec3 e1;
    vec3 e2;
    basic_triangle() = default;
    basic_triangle(vec3 const& v1_, vec3 const& e1_, vec3 const& e2_) : v1(v1_), e1(e1_), e2(e2_) {}
};

template <typename T = float, typename P = unsigned>
struct basic_sphere : primitive_ids
{
    vec3  center;
    float radius;
    basic_sphere() = default;
    basic_sphere(vec3 const& c, float r) : center(c), radius(r) {}
};

static_assert(sizeof(vec3) == 16, "vec3 layout");
static_assert(sizeof(ray) == 32, "ray layout");
static_assert(sizeof(aabb) == 32, "aabb layout");
static_assert(sizeof(basic_triangle<3, float>) == 64, "triangle layout");
static_assert(sizeof(basic_sphere<float>) == 48, "sphere layout");

// get_bounds: math/detail/triangle.inl:34-44, math/detail/sphere.inl:29-38
inline aabb get_bounds(basic_triangle<3, float> const& t)
{
    aabb b; b.invalidate();
    b.insert(t.v1); b.insert(t.v1 + t.e1); b.insert(t.v1 + t.e2);
    return b;
}
inline aabb get_bounds(basic_sphere<float> const& s)
{
    aabb b; b.invalidate();
    b.insert(s.center - vec3(s.radius)); b.insert(s.center + vec3(s.radius));
    return b;
}

// generic_primitive<basic_triangle<3,float>, basic_sphere<float>> (generic_primitive.h, variant.h): a
// variant holding either; the alternative sits in place, its 1-based index follows at +64 (80 bytes,
// 16-byte aligned -- the reference's layout, so arrays pass unchanged).  geom_id / prim_id are the first
// two words of either alternative.
template <typename ...Ts> class generic_primitive;

template <>
class alignas(16) generic_primitive<basic_triangle<3, float>, basic_sphere<float>>
{
public:
    using triangle_type = basic_triangle<3, float>;
    using sphere_type = basic_sphere<float>;

    generic_primitive() { std::memset(this, 0, sizeof(*this)); }
    generic_primitive(triangle_type const& t) { std::memset(this, 0, sizeof(*this)); std::memcpy(storage_, &t, sizeof(t)); which_ = 1; }
    generic_primitive(sphere_type const& s)   { std::memset(this, 0, sizeof(*this)); std::memcpy(storage_, &s, sizeof(s)); which_ = 2; }

    bool is_triangle() const { return which_ == 1; }
    bool is_sphere() const { return which_ == 2; }
    triangle_type const& as_triangle() const { return *reinterpret_cast<triangle_type const*>(storage_); }
    sphere_type const& as_sphere() const { return *reinterpret_cast<sphere_type const*>(storage_); }
    unsigned geom_id() const { return reinterpret_cast<unsigned const*>(storage_)[0]; }
    unsigned prim_id() const { return reinterpret_cast<unsigned const*>(storage_)[1]; }

private:
    alignas(16) unsigned char storage_[64];
    unsigned which_;
    unsigned pad_[3];
};
using generic_triangle_sphere = generic_primitive<basic_triangle<3, float>, basic_sphere<float>>;
static_assert(sizeof(generic_triangle_sphere) == 80, "generic_primitive layout");

// get_bounds through the visitor (detail/generic_primitive.inl:46-70)
inline aabb get_bounds(generic_triangle_sphere const& g)
{
    return g.is_triangle() ? get_bounds(g.as_triangle()) : get_bounds(g.as_sphere());
}

} // namespace visionaray_b200

#endif

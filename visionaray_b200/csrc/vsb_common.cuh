// vsb_common.cuh -- shared device/host definitions of the B200 ray-query path.
//
// Everything that decides a hit (ray/triangle, ray/sphere, the exact slab test) is written
// with explicit round-to-nearest single operations (rn_mul/rn_add/rn_sub/rn_rcp) so that nvcc
// can never contract them into FMAs: the arithmetic is then the same IEEE-754 sequence as the
// reference's scalar host path built without contraction (math/intersect.h), and hit
// decisions, t, u, v are bit-identical to it.
//
// The file also compiles as plain host C++ (VSB_HOSTSIM) for tests/hostsim, which runs the
// re-encode and the wide traversal logic on the CPU in the no-GPU test tier.  That host
// instantiation is test-only; libvsnray_b200.so exports no CPU path.
#pragma once

#include <stdint.h>
#include <float.h>

#if defined(__CUDACC__)
#define VSB_HD __host__ __device__ __forceinline__
#define VSB_D  __device__ __forceinline__
#else
#define VSB_HD inline
#define VSB_D  inline
#include <cmath>
#include <cstring>
struct float4 { float x, y, z, w; };
struct uint4 { uint32_t x, y, z, w; };
struct float2 { float x, y; };
struct uchar4 { unsigned char x, y, z, w; };
#endif

// A traversal stack that would overflow is a broken invariant (tree depths are checked when a scene is created /
// before the reference-order walk is chosen), never a case to drop a subtree silently: the kernel traps (the call
// returns a CUDA error), the host instantiation aborts.
#if defined(__CUDA_ARCH__)
#define VSB_STACK_OVERFLOW() __trap()
#else
#include <cstdlib>
#define VSB_STACK_OVERFLOW() std::abort()
#endif

namespace vsb
{

// ---- reference-layout records (include/vsnray_b200.h) --------------------------------------

struct Ray32 { float ox, oy, oz, pad0, dx, dy, dz, pad1; };

struct Node32            // bvh_node, bvh.h:52-119
{
    float    min_x, min_y, min_z;
    uint32_t first;
    float    max_x, max_y, max_z;
    uint32_t num_prims;
};

struct Hit32             // vsb_hit
{
    int32_t  hit;
    uint32_t prim_id, geom_id;
    float    t, u, v;
    uint32_t prim_list_index, pad;
};

// ---- leaf-order primitive records of the wide BVH ------------------------------------------
// Triangles: 48 B = 3 x 128-bit loads; the index indirection of index_bvh (bvh.h:226-229) is
// gone, ids and the list index ride in the w lanes.
struct TriRec
{
    float v1x, v1y, v1z; uint32_t prim_id;
    float e1x, e1y, e1z; uint32_t geom_id;
    float e2x, e2y, e2z; uint32_t list_index;
};
// Spheres: 32 B = 2 x 128-bit loads.
struct SphRec
{
    float cx, cy, cz, radius;
    uint32_t prim_id, geom_id, list_index, pad;
};

// ---- BVH8 node: 80 B = 5 x 128-bit records ---------------------------------------------------
//  q0: origin p (3 x f32) | e_x e_y e_z imask (4 x u8; child box = p + q * 2^(e-127... see below))
//  q1: child_base (u32) | prim_base (u32) | valid (u32) | 0
//  q2: qlo_x[8] | qlo_y[8]      q3: qlo_z[8] | qhi_x[8]      q4: qhi_y[8] | qhi_z[8]
// valid: nibble i = (1 << count) - 1 for a leaf child in slot i with count = 1..4 primitives, 0 for an inner child (bit i
// of imask set) or an empty slot.  The node's leaf-order records lie in slot order from prim_base: the record of bit b
// of `valid` is prim_base + popc(valid & ((1 << b) - 1)).  A node test turns its 8 child hit bits into a primitive mask
// with one nibble expansion and one AND (round 1 kept count << 5 | offset per child and built the mask child by child:
// ~40 of the ~180 ALU-pipe instructions of a node visit).  The exponents are
// biased float exponents: scale = bits_to_float(e << 23).
struct Node80
{
    uint32_t w[20];
};

#define VSB_WIDE 8
#ifndef VSB_Q_PREFETCH
#define VSB_Q_PREFETCH 0          // group ring: prefetch a group's records when it is appended (1: into L1, 2: into L2)
#endif
#ifndef VSB_GROUP_RING_ALL
#define VSB_GROUP_RING_ALL 0      // development switch: 1 / -1 = group ring / pair ring for every program (0: per program)
#endif
#define VSB_MAX_LEAF_BITS 32

// ---- exact single-rounding arithmetic ---------------------------------------------------------
#if defined(__CUDA_ARCH__)
VSB_D float rn_mul(float a, float b) { return __fmul_rn(a, b); }
VSB_D float rn_add(float a, float b) { return __fadd_rn(a, b); }
VSB_D float rn_sub(float a, float b) { return __fsub_rn(a, b); }
VSB_D float rn_rcp(float a) { return __frcp_rn(a); }          // correctly rounded == 1.0f / a
VSB_D float rn_div(float a, float b) { return __fdiv_rn(a, b); }
VSB_D float rn_sqrt(float a) { return __fsqrt_rn(a); }
#else
// host build: compiled with -ffp-contract=off, so plain operators are single roundings
VSB_HD float rn_mul(float a, float b) { return a * b; }
VSB_HD float rn_add(float a, float b) { return a + b; }
VSB_HD float rn_sub(float a, float b) { return a - b; }
VSB_HD float rn_rcp(float a) { return 1.0f / a; }
VSB_HD float rn_div(float a, float b) { return a / b; }
VSB_HD float rn_sqrt(float a) { return sqrtf(a); }
#endif

// reference min/max: x < y ? x : y  /  x < y ? y : x  (math/detail/math.h:48-60)
VSB_HD float ref_min(float x, float y) { return x < y ? x : y; }
VSB_HD float ref_max(float x, float y) { return x < y ? y : x; }

// dot = (x*x' + y*y') + z*z'   (math/detail/vector3.inl:305-308)
VSB_HD float rn_dot(float ax, float ay, float az, float bx, float by, float bz)
{
    return rn_add(rn_add(rn_mul(ax, bx), rn_mul(ay, by)), rn_mul(az, bz));
}

VSB_HD uint32_t f2u(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
VSB_HD float u2f(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

// ---- hit record in registers -------------------------------------------------------------------
// Primitive kinds (include/vsnray_b200.h VSB_PRIM_*): 0 basic_triangle (64 B), 1 basic_sphere (48 B),
// 2 generic_primitive<basic_triangle, basic_sphere> (80 B: the alternative in place, then its 1-based
// index at +64 -- generic_primitive.h, variant.h).  Every reader of the reference-layout array goes
// through these two.
VSB_HD uint32_t prim_stride(int kind) { return kind == 0 ? 64u : (kind == 1 ? 48u : 80u); }
VSB_HD bool prim_is_triangle(int kind, const void* rec)
{
    return kind == 0 || (kind == 2 && reinterpret_cast<const uint32_t*>(rec)[16] == 1u);
}
// leaf-order records of the wide BVH: 48 B (3 x 128 bit) for triangles and for the generic kind, 32 B for
// spheres.  In a generic scene a sphere keeps the 32-byte sphere layout in the first two words and
// marks itself with bit 31 of word 11 (a triangle's list_index lives there, < 2^31).
#define VSB_GENERIC_SPHERE_TAG 0x80000000u
VSB_HD int wide_rec_words16(int kind) { return kind == 1 ? 2 : 3; }

struct HitReg
{
    float    t;          // FLT_MAX while no hit (math/intersect.h:95-103)
    float    u, v;
    uint32_t prim_id, geom_id, list_index;
    int      hit;
};

VSB_HD void hit_reset(HitReg& h)
{
    h.t = FLT_MAX; h.u = 0.0f; h.v = 0.0f; h.prim_id = 0; h.geom_id = 0; h.list_index = 0; h.hit = 0;
}

// Ray / triangle: Moeller-Trumbore on (v1, e1, e2), operation for operation as
// math/intersect.h:122-179.  Returns true and fills t/u/v when the primitive test hits (no
// t-range test here: is_closer() applies t >= 0, t < best, t < max_t; update_if.h:50-79).
VSB_HD bool isect_triangle(float ox, float oy, float oz, float dx, float dy, float dz,
                           float v1x, float v1y, float v1z, float e1x, float e1y, float e1z,
                           float e2x, float e2y, float e2z, float& t, float& u, float& v)
{
    // s1 = cross(dir, e2)
    float s1x = rn_sub(rn_mul(dy, e2z), rn_mul(dz, e2y));
    float s1y = rn_sub(rn_mul(dz, e2x), rn_mul(dx, e2z));
    float s1z = rn_sub(rn_mul(dx, e2y), rn_mul(dy, e2x));
    float div = rn_dot(s1x, s1y, s1z, e1x, e1y, e1z);
    if (!(div != 0.0f)) return false;
    float inv_div = rn_rcp(div);
    float ddx = rn_sub(ox, v1x), ddy = rn_sub(oy, v1y), ddz = rn_sub(oz, v1z);
    float b1 = rn_mul(rn_dot(ddx, ddy, ddz, s1x, s1y, s1z), inv_div);
    if (!(b1 >= 0.0f && b1 <= 1.0f)) return false;
    // s2 = cross(d, e1)
    float s2x = rn_sub(rn_mul(ddy, e1z), rn_mul(ddz, e1y));
    float s2y = rn_sub(rn_mul(ddz, e1x), rn_mul(ddx, e1z));
    float s2z = rn_sub(rn_mul(ddx, e1y), rn_mul(ddy, e1x));
    float b2 = rn_mul(rn_dot(dx, dy, dz, s2x, s2y, s2z), inv_div);
    if (!(b2 >= 0.0f && rn_add(b1, b2) <= 1.0f)) return false;
    t = rn_mul(rn_dot(e2x, e2y, e2z, s2x, s2y, s2z), inv_div);
    u = b1;
    v = b2;
    return true;
}

// Ray / sphere: math/intersect.h:186-221.  `valid` is the reference's hit flag; t is the
// smaller root even when negative (so rays starting inside miss), u = v = 0.
VSB_HD bool isect_sphere(float ox, float oy, float oz, float dx, float dy, float dz,
                         float cx, float cy, float cz, float radius, float& t)
{
    float rx = rn_sub(ox, cx), ry = rn_sub(oy, cy), rz = rn_sub(oz, cz);
    float A = rn_dot(dx, dy, dz, dx, dy, dz);
    float B = rn_mul(rn_dot(dx, dy, dz, rx, ry, rz), 2.0f);
    float C = rn_sub(rn_dot(rx, ry, rz, rx, ry, rz), rn_mul(radius, radius));
    float disc = rn_sub(rn_mul(B, B), rn_mul(rn_mul(4.0f, A), C));
    bool valid = disc >= 0.0f;
    float root_disc = valid ? rn_sqrt(disc) : disc;
    float q = B < 0.0f ? rn_mul(-0.5f, rn_sub(B, root_disc)) : rn_mul(-0.5f, rn_add(B, root_disc));
    float t1 = rn_div(q, A);
    float t2 = rn_div(C, q);
    t = valid ? (t1 > t2 ? t2 : t1) : -1.0f;
    return valid;
}

// Exact slab test: math/intersect.h:52-70 (same operation order and min/max semantics).
VSB_HD bool isect_aabb_exact(float ox, float oy, float oz, float ix, float iy, float iz,
                             float minx, float miny, float minz, float maxx, float maxy, float maxz,
                             float& tnear, float& tfar)
{
    float t1x = rn_mul(rn_sub(minx, ox), ix), t1y = rn_mul(rn_sub(miny, oy), iy), t1z = rn_mul(rn_sub(minz, oz), iz);
    float t2x = rn_mul(rn_sub(maxx, ox), ix), t2y = rn_mul(rn_sub(maxy, oy), iy), t2z = rn_mul(rn_sub(maxz, oz), iz);
    tnear = ref_max(ref_min(t1x, t2x), ref_max(ref_min(t1y, t2y), ref_min(t1z, t2z)));
    tfar  = ref_min(ref_max(t1x, t2x), ref_min(ref_max(t1y, t2y), ref_max(t1z, t2z)));
    return tfar >= tnear;
}

// Deterministic counter-based sampler shared with the oracle (oracle/vsnray_oracle.c:
// vso_sampler_init / vso_sampler_next): PCG-RXS-M-XS-32 seeded by a hash of (pixel, frame, seed).
VSB_HD uint32_t sampler_init(uint32_t pixel, uint32_t frame, uint32_t seed)
{
    uint32_t h = seed * 0x9E3779B9u + 0x7F4A7C15u;
    h ^= pixel + 0x85EBCA6Bu + (h << 6) + (h >> 2);
    h ^= frame + 0xC2B2AE35u + (h << 6) + (h >> 2);
    h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
    return h;
}
VSB_HD float sampler_next(uint32_t& s)
{
    s = s * 747796405u + 2891336453u;
    uint32_t w = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
    w = (w >> 22u) ^ w;
    return (float)(w >> 8) * (1.0f / 16777216.0f);
}

} // namespace vsb

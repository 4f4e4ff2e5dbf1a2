/*
 * vsnray_oracle.c -- TEST INFRASTRUCTURE ONLY (see vsnray_oracle.h).
 *
 * Plain-C scalar restatement of the reference's ray-query hot path.  It restates the
 * *algorithm* (operation order included, so that results are bit-identical to the scalar
 * float instantiation of the reference headers when both are built with -ffp-contract=off);
 * no reference code is copied.  Paths cited are relative to
 * /root/reference/include/visionaray/.
 *
 * Parity status: PINNED against the reference itself (tests/test_oracle_vs_ref.py) and
 * against the reference's known-answer test (test/unittests/get_normal.cpp:17-75).
 */
#define _GNU_SOURCE 1
#include "vsnray_oracle.h"

#include <float.h>
#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* math/detail/math.h:48-60: min/max are `x < y ? x : y` / `x < y ? y : x` (NaN-propagation
 * order matters for 0*inf in the slab test) */
static inline float vmin(float x, float y) { return x < y ? x : y; }
static inline float vmax(float x, float y) { return x < y ? y : x; }

/* math/detail/vector3.inl:305-308: dot = (x*x' + y*y') + z*z' */
static inline float dot3(const float a[3], const float b[3])
{
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

/* math/detail/vector3.inl:293-301 */
static inline void cross3(const float u[3], const float v[3], float out[3])
{
    float x = u[1] * v[2] - u[2] * v[1];
    float y = u[2] * v[0] - u[0] * v[2];
    float z = u[0] * v[1] - u[1] * v[0];
    out[0] = x; out[1] = y; out[2] = z;
}

/* math/detail/vector3.inl:333-336 with rsqrt = 1/sqrt (math/detail/math.h:479-482) */
static inline void normalize3(const float v[3], float out[3])
{
    float s = 1.0f / sqrtf(dot3(v, v));
    out[0] = v[0] * s; out[1] = v[1] * s; out[2] = v[2] * s;
}

static inline float length3(const float v[3]) { return sqrtf(dot3(v, v)); }

static inline void hit_init(vso_hit* h)
{
    /* math/intersect.h:95-103 + detail/bvh/hit_record.h:46 */
    h->hit = 0; h->prim_id = 0; h->geom_id = 0; h->t = FLT_MAX; h->u = 0.0f; h->v = 0.0f;
    h->prim_list_index = 0; h->pad = 0;
}

/* ray / aabb slab test -- math/intersect.h:52-70 */
int vso_intersect_aabb(const vso_ray* r, const float bmin[3], const float bmax[3],
                       const float inv_dir[3], float* tnear, float* tfar)
{
    float t1x = (bmin[0] - r->ori[0]) * inv_dir[0];
    float t1y = (bmin[1] - r->ori[1]) * inv_dir[1];
    float t1z = (bmin[2] - r->ori[2]) * inv_dir[2];
    float t2x = (bmax[0] - r->ori[0]) * inv_dir[0];
    float t2y = (bmax[1] - r->ori[1]) * inv_dir[1];
    float t2z = (bmax[2] - r->ori[2]) * inv_dir[2];
    /* min_max(x,y,z) = max(min(x,y), z); max_min(x,y,z) = min(max(x,y), z)
     * (math/detail/math.h:126-140) */
    float tn = vmax(vmin(t1x, t2x), vmax(vmin(t1y, t2y), vmin(t1z, t2z)));
    float tf = vmin(vmax(t1x, t2x), vmin(vmax(t1y, t2y), vmax(t1z, t2z)));
    *tnear = tn;
    *tfar = tf;
    return tf >= tn;
}

/* ray / triangle, Moeller-Trumbore on (v1,e1,e2) -- math/intersect.h:122-179 */
void vso_intersect_triangle(const vso_ray* r, const vso_triangle* tri, vso_hit* out)
{
    float s1[3], s2[3], d[3];
    hit_init(out);
    out->t = -1.0f;

    cross3(r->dir, tri->e2, s1);
    float div = dot3(s1, tri->e1);
    if (!(div != 0.0f)) return;

    float inv_div = 1.0f / div;
    d[0] = r->ori[0] - tri->v1[0];
    d[1] = r->ori[1] - tri->v1[1];
    d[2] = r->ori[2] - tri->v1[2];
    float b1 = dot3(d, s1) * inv_div;
    if (!(b1 >= 0.0f && b1 <= 1.0f)) return;

    cross3(d, tri->e1, s2);
    float b2 = dot3(r->dir, s2) * inv_div;
    if (!(b2 >= 0.0f && b1 + b2 <= 1.0f)) return;

    out->hit = 1;
    out->prim_id = tri->prim_id;
    out->geom_id = tri->geom_id;
    out->t = dot3(tri->e2, s2) * inv_div;
    out->u = b1;
    out->v = b2;
}

/* ray / sphere -- math/intersect.h:186-221.  Returns the smaller root even when negative,
 * so a ray starting inside a sphere misses it (is_closer needs t >= 0, update_if.h:50-56). */
void vso_intersect_sphere(const vso_ray* r, const vso_sphere* s, vso_hit* out)
{
    float o[3] = { r->ori[0] - s->center[0], r->ori[1] - s->center[1], r->ori[2] - s->center[2] };
    float A = dot3(r->dir, r->dir);
    float B = dot3(r->dir, o) * 2.0f;
    float C = dot3(o, o) - s->radius * s->radius;
    float disc = B * B - 4.0f * A * C;
    int valid = disc >= 0.0f;
    float root_disc = valid ? sqrtf(disc) : disc;
    float q = B < 0.0f ? -0.5f * (B - root_disc) : -0.5f * (B + root_disc);
    float t1 = q / A;
    float t2 = C / q;
    hit_init(out);
    out->hit = valid;
    out->prim_id = s->prim_id;
    out->geom_id = s->geom_id;
    out->t = valid ? (t1 > t2 ? t2 : t1) : -1.0f;
}

/* intersect(ray, generic_primitive<triangle, sphere>) -- detail/generic_primitive.inl:20-44: the visitor
 * calls intersect() of the alternative the variant holds (which = 1-based index, variant.h) */
void vso_intersect_generic(const vso_ray* r, const vso_generic_primitive* g, vso_hit* out)
{
    if (g->which == 1) vso_intersect_triangle(r, (const vso_triangle*)g->storage, out);
    else               vso_intersect_sphere(r, (const vso_sphere*)g->storage, out);
}

static inline void isect_prim(const vso_ray* r, const vso_bvh* b, uint32_t i, vso_hit* hr)
{
    /* index_bvh: primitive(i) = prims[indices[i]] (bvh.h:226-229); bvh: prims[i] (:177-180) */
    uint32_t p = b->indices ? b->indices[i] : i;
    if (b->prim_kind == 0)
        vso_intersect_triangle(r, (const vso_triangle*)b->prims + p, hr);
    else if (b->prim_kind == 1)
        vso_intersect_sphere(r, (const vso_sphere*)b->prims + p, hr);
    else
        vso_intersect_generic(r, (const vso_generic_primitive*)b->prims + p, hr);
    hr->prim_list_index = i;   /* HR(isect(ray, prim), i), detail/bvh/intersect.inl:112 */
}

/* is_closer(box, result, max_t) -- update_if.h:62-66,84-88 */
static inline int box_closer(int hit, float tnear, float tfar, float best_t, float max_t)
{
    return hit && tnear < best_t && tfar >= 0.0f && tnear < max_t;
}

/* The traversal loop -- detail/bvh/intersect.inl:35-134.
 * The reference's stack<32> holds 31 entries and is unchecked (detail/stack.h:39-47); this
 * restatement uses 128 entries, which differs only where the reference has undefined
 * behaviour (tree depth > 31, reported by vso_bvh_max_depth()). */
void vso_intersect_bvh(int any, const vso_ray* r, const vso_bvh* b, float max_t, vso_hit* out,
                       vso_counters* c)
{
    vso_hit result;
    uint32_t stack[128];
    int sp = 0;
    uint64_t n_inner = 0, n_prim = 0;
    float inv_dir[3] = { 1.0f / r->dir[0], 1.0f / r->dir[1], 1.0f / r->dir[2] };   /* :63 */

    hit_init(&result);
    stack[sp++] = 0;   /* root */

    while (sp > 0)
    {
        const vso_bvh_node* node = &b->nodes[stack[--sp]];
        int culled = 0;

        while (node->num_prims == 0)   /* inner, bvh.h:63 */
        {
            const vso_bvh_node* ch = &b->nodes[node->first];   /* siblings adjacent, :76 */
            float tn1, tf1, tn2, tf2;
            int h1 = vso_intersect_aabb(r, ch[0].bbox_min, ch[0].bbox_max, inv_dir, &tn1, &tf1);
            int h2 = vso_intersect_aabb(r, ch[1].bbox_min, ch[1].bbox_max, inv_dir, &tn2, &tf2);
            int b1 = box_closer(h1, tn1, tf1, result.t, max_t);
            int b2 = box_closer(h2, tn2, tf2, result.t, max_t);
            ++n_inner;

            if (b1 && b2)
            {
                unsigned near_addr = (tn1 < tn2) ? 0u : 1u;   /* ties -> child 1, :86 */
                stack[sp++] = node->first + !near_addr;
                node = &ch[near_addr];
            }
            else if (b1) node = &ch[0];
            else if (b2) node = &ch[1];
            else { culled = 1; break; }   /* goto next, :100 */
        }
        if (culled) continue;

        for (uint32_t i = node->first; i != node->first + node->num_prims; ++i)
        {
            vso_hit hr;
            isect_prim(r, b, i, &hr);
            ++n_prim;
            /* is_closer(hr, result, max_t): update_if.h:50-56,75-79 (strict <: first wins ties) */
            int closer = hr.hit && (hr.t >= 0.0f && hr.t < result.t) && hr.t < max_t;
            if (!closer) continue;
            result = hr;   /* update_if with a true scalar mask, update_if.h:29-37 */
            result.hit = 1;
            if (any) goto done;   /* exit_traversal<AnyHit>, detail/exit_traversal.h:49-57 */
        }
    }
done:
    if (!result.hit)
    {
        /* a miss keeps the default record (t = FLT_MAX) */
        hit_init(&result);
    }
    *out = result;
    if (c)
    {
        c->n_inner += n_inner; c->n_prim += n_prim; c->n_rays += 1; c->n_hits += (uint64_t)result.hit;
    }
}

int vso_bvh_max_depth(const vso_bvh* b)
{
    /* iterative DFS with explicit (node, depth) stack sized by node count */
    size_t cap = 256, sp = 0;
    uint32_t* st = (uint32_t*)malloc(cap * 2 * sizeof(uint32_t));
    int maxd = 0;
    if (b->n_nodes == 0) { free(st); return 0; }
    st[0] = 0; st[1] = 1; sp = 1;
    while (sp > 0)
    {
        --sp;
        uint32_t n = st[sp * 2], d = st[sp * 2 + 1];
        if ((int)d > maxd) maxd = (int)d;
        if (b->nodes[n].num_prims == 0)
        {
            if (sp + 2 > cap) { cap *= 2; st = (uint32_t*)realloc(st, cap * 2 * sizeof(uint32_t)); }
            st[sp * 2] = b->nodes[n].first;     st[sp * 2 + 1] = d + 1; ++sp;
            st[sp * 2] = b->nodes[n].first + 1; st[sp * 2 + 1] = d + 1; ++sp;
        }
    }
    free(st);
    return maxd;
}

/* ------------------------------------------------------------------------------------------
 * parallel driver (pthread; dynamic chunks)
 * ---------------------------------------------------------------------------------------- */

typedef void (*range_fn)(size_t b, size_t e, void* ctx, vso_counters* c);

typedef struct
{
    size_t n, chunk;
    size_t* next;
    pthread_mutex_t* mu;
    range_fn fn;
    void* ctx;
    vso_counters c;
} worker_t;

static void* worker_main(void* p)
{
    worker_t* w = (worker_t*)p;
    for (;;)
    {
        size_t b;
        pthread_mutex_lock(w->mu);
        b = *w->next;
        *w->next += w->chunk;
        pthread_mutex_unlock(w->mu);
        if (b >= w->n) break;
        size_t e = b + w->chunk < w->n ? b + w->chunk : w->n;
        w->fn(b, e, w->ctx, &w->c);
    }
    return NULL;
}

static void parallel_ranges(size_t n, int nthreads, range_fn fn, void* ctx, vso_counters* total)
{
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    size_t next = 0;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    worker_t* ws = (worker_t*)calloc((size_t)nthreads, sizeof(worker_t));
    pthread_t* ts = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
    for (int i = 0; i < nthreads; ++i)
    {
        ws[i].n = n; ws[i].chunk = 2048; ws[i].next = &next; ws[i].mu = &mu; ws[i].fn = fn; ws[i].ctx = ctx;
    }
    if (nthreads == 1) worker_main(&ws[0]);
    else
    {
        for (int i = 0; i < nthreads; ++i) pthread_create(&ts[i], NULL, worker_main, &ws[i]);
        for (int i = 0; i < nthreads; ++i) pthread_join(ts[i], NULL);
    }
    if (total)
    {
        for (int i = 0; i < nthreads; ++i)
        {
            total->n_inner += ws[i].c.n_inner; total->n_prim += ws[i].c.n_prim;
            total->n_rays += ws[i].c.n_rays;   total->n_hits += ws[i].c.n_hits;
        }
    }
    free(ws); free(ts);
}

/* ------------------------------------------------------------------------------------------
 * closest_hit() / any_hit() over a ray list -- detail/traverse_linear.inl:186-318 with a
 * one-element range of bvh refs (the merge step :120 is then the identity)
 * ---------------------------------------------------------------------------------------- */

typedef struct
{
    int any; const vso_bvh* b; const vso_ray* rays; float max_t; const float* max_t_per_ray; vso_hit* hits;
} query_ctx;

static void query_range(size_t b, size_t e, void* p, vso_counters* c)
{
    query_ctx* q = (query_ctx*)p;
    for (size_t i = b; i < e; ++i)
    {
        float mt = q->max_t_per_ray ? q->max_t_per_ray[i] : q->max_t;
        vso_intersect_bvh(q->any, &q->rays[i], q->b, mt, &q->hits[i], c);
    }
}

void vso_query(int any, const vso_bvh* b, const vso_ray* rays, size_t n_rays, float max_t,
               const float* max_t_per_ray, vso_hit* hits, vso_counters* counters, int nthreads)
{
    query_ctx q = { any, b, rays, max_t, max_t_per_ray, hits };
    parallel_ranges(n_rays, nthreads, query_range, &q, counters);
}

/* ------------------------------------------------------------------------------------------
 * multi_hit<N>() -- detail/traverse_linear.inl:325-380 -> intersect<MultiHit, N>()
 * (detail/bvh/intersect.inl:35-134 with RT = array<HR, N>, detail/multi_hit.h:29-41).
 *   box:   is_closer(box, result_array, max_t) = OR over the records (multi_hit.h:206-233);
 *   prim:  same OR, then update_if(array, hr) = algo::insert_sorted(hr, begin, end, is_closer_t())
 *          (multi_hit.h:163-176, detail/algorithm.h:46-74): before the first record with a larger t,
 *          records behind it shift by one, the last drops out;
 *   no early exit (exit_traversal<MultiHit> is the generic "never").
 * max_t is numeric_limits::max() (multi_hit() has no max_t overload).
 * ---------------------------------------------------------------------------------------- */
static int multi_closer_box(int hit, float tnear, float tfar, const vso_hit* res, int n, float max_t)
{
    for (int i = 0; i < n; ++i)
        if (hit && tnear < res[i].t && tfar >= 0.0f && tnear < max_t) return 1;
    return 0;
}

static int multi_closer_prim(const vso_hit* hr, const vso_hit* res, int n, float max_t)
{
    for (int i = 0; i < n; ++i)
        if (hr->hit && (hr->t >= 0.0f && hr->t < res[i].t) && hr->t < max_t) return 1;
    return 0;
}

static void insert_sorted(const vso_hit* item, vso_hit* first, int n)
{
    int pos = n;
    for (int it = 0; it < n; ++it)
        if (item->hit && (item->t >= 0.0f && item->t < first[it].t)) { pos = it; break; }
    for (int it = (pos != n) ? n - 1 : n; it > pos; --it) first[it] = first[it - 1];
    if (pos != n) first[pos] = *item;
}

void vso_multi_hit_bvh(const vso_ray* r, const vso_bvh* b, int n, vso_hit* result)
{
    const float max_t = FLT_MAX;
    uint32_t stack[128];
    int sp = 0;
    float inv_dir[3] = { 1.0f / r->dir[0], 1.0f / r->dir[1], 1.0f / r->dir[2] };
    for (int i = 0; i < n; ++i) hit_init(&result[i]);
    stack[sp++] = 0;
    while (sp > 0)
    {
        const vso_bvh_node* node = &b->nodes[stack[--sp]];
        int culled = 0;
        while (node->num_prims == 0)
        {
            const vso_bvh_node* ch = &b->nodes[node->first];
            float tn1, tf1, tn2, tf2;
            int h1 = vso_intersect_aabb(r, ch[0].bbox_min, ch[0].bbox_max, inv_dir, &tn1, &tf1);
            int h2 = vso_intersect_aabb(r, ch[1].bbox_min, ch[1].bbox_max, inv_dir, &tn2, &tf2);
            int b1 = multi_closer_box(h1, tn1, tf1, result, n, max_t);
            int b2 = multi_closer_box(h2, tn2, tf2, result, n, max_t);
            if (b1 && b2)
            {
                unsigned near_addr = (tn1 < tn2) ? 0u : 1u;
                stack[sp++] = node->first + !near_addr;
                node = &ch[near_addr];
            }
            else if (b1) node = &ch[0];
            else if (b2) node = &ch[1];
            else { culled = 1; break; }
        }
        if (culled) continue;
        for (uint32_t i = node->first; i != node->first + node->num_prims; ++i)
        {
            vso_hit hr;
            isect_prim(r, b, i, &hr);
            if (!multi_closer_prim(&hr, result, n, max_t)) continue;
            hr.hit = 1;
            insert_sorted(&hr, result, n);
        }
    }
}

/* a range of BVHs: traverse() (detail/traverse_linear.inl:71-130) merges the per-BVH arrays with
 * update_if(array, array) (detail/multi_hit.h:181-202) */
typedef struct { const vso_bvh* bvhs; int n_bvhs; const vso_ray* rays; int n; vso_hit* hits; } multi_ctx;

static void multi_range(size_t b, size_t e, void* p, vso_counters* c)
{
    multi_ctx* q = (multi_ctx*)p;
    vso_hit tmp[64];
    (void)c;
    for (size_t i = b; i < e; ++i)
    {
        vso_hit* dst = q->hits + i * (size_t)q->n;
        for (int k = 0; k < q->n; ++k) hit_init(&dst[k]);
        for (int s = 0; s < q->n_bvhs; ++s)
        {
            vso_multi_hit_bvh(&q->rays[i], &q->bvhs[s], q->n, tmp);
            if (!dst[0].hit) { memcpy(dst, tmp, sizeof(vso_hit) * (size_t)q->n); continue; }
            for (int k = 0; k < q->n && tmp[k].hit; ++k) insert_sorted(&tmp[k], dst, q->n);
        }
        for (int k = 0; k < q->n; ++k) if (!dst[k].hit) hit_init(&dst[k]);
    }
}

void vso_multi_hit(const vso_bvh* bvhs, int n_bvhs, const vso_ray* rays, size_t n_rays, int n, vso_hit* hits, int nthreads)
{
    multi_ctx q = { bvhs, n_bvhs, rays, n, hits };
    if (n < 1 || n > 64) return;
    parallel_ranges(n_rays, nthreads, multi_range, &q, NULL);
}

/* brute force in list order -- detail/traverse_linear.inl:26-63 */
typedef struct
{
    int any, prim_kind; const void* prims; size_t n_prims; const vso_ray* rays; float max_t; vso_hit* hits;
} brute_ctx;

static void brute_range(size_t b, size_t e, void* p, vso_counters* c)
{
    brute_ctx* q = (brute_ctx*)p;
    (void)c;
    for (size_t i = b; i < e; ++i)
    {
        vso_hit result, hr;
        hit_init(&result);
        for (size_t k = 0; k < q->n_prims; ++k)
        {
            if (q->prim_kind == 0)      vso_intersect_triangle(&q->rays[i], (const vso_triangle*)q->prims + k, &hr);
            else if (q->prim_kind == 1) vso_intersect_sphere(&q->rays[i], (const vso_sphere*)q->prims + k, &hr);
            else                        vso_intersect_generic(&q->rays[i], (const vso_generic_primitive*)q->prims + k, &hr);
            hr.prim_list_index = (uint32_t)k;
            if (hr.hit && hr.t >= 0.0f && hr.t < result.t && hr.t < q->max_t)
            {
                result = hr; result.hit = 1;
                if (q->any) break;
            }
        }
        if (!result.hit) hit_init(&result);
        q->hits[i] = result;
    }
}

void vso_query_brute(int any, int prim_kind, const void* prims, size_t n_prims,
                     const vso_ray* rays, size_t n_rays, float max_t, vso_hit* hits, int nthreads)
{
    brute_ctx q = { any, prim_kind, prims, n_prims, rays, max_t, hits };
    parallel_ranges(n_rays, nthreads, brute_range, &q, NULL);
}

/* ------------------------------------------------------------------------------------------
 * camera basis, primary rays, sampler
 * ---------------------------------------------------------------------------------------- */

/* detail/cuda_sched.inl:259-266 (identical in detail/tiled_sched.inl:296-304) */
void vso_camera_basis(const vso_frame_params* p, float eye[3], float cu[3], float cv[3], float cw[3])
{
    float d[3] = { p->eye[0] - p->center[0], p->eye[1] - p->center[1], p->eye[2] - p->center[2] };
    float f[3], s[3], u[3], c[3];
    float up[3] = { p->up[0], p->up[1], p->up[2] };
    normalize3(d, f);
    cross3(up, f, c);
    normalize3(c, s);
    cross3(f, s, u);
    float th = tanf(p->fovy / 2.0f);
    for (int i = 0; i < 3; ++i)
    {
        eye[i] = p->eye[i];
        cu[i] = s[i] * th * p->aspect;
        cv[i] = u[i] * th;
        cw[i] = -f[i];
    }
}

/* detail/sched_common.h:130-150 */
void vso_primary_ray(float x, float y, int width, int height, const float eye[3],
                     const float cu[3], const float cv[3], const float cw[3], vso_ray* out)
{
    float u = 2.0f * (x + 0.5f) / (float)width - 1.0f;
    float v = 2.0f * (y + 0.5f) / (float)height - 1.0f;
    float d[3];
    for (int i = 0; i < 3; ++i) d[i] = cu[i] * u + cv[i] * v + cw[i];
    out->ori[0] = eye[0]; out->ori[1] = eye[1]; out->ori[2] = eye[2]; out->pad0 = 0.0f;
    normalize3(d, out->dir);
    out->pad1 = 0.0f;
}

/* PCG-RXS-M-XS-32 seeded by a hash of (pixel, frame, seed).  Not from the reference (its
 * schedulers use a wall-clock-seeded std::default_random_engine, detail/tiled_sched.inl:191,
 * random_sampler.h:34): shared by the oracle, the reference harness and the CUDA kernels so
 * that RNG-consuming kernels can be compared at all. */
uint32_t vso_sampler_init(uint32_t pixel, uint32_t frame, uint32_t seed)
{
    uint32_t h = seed * 0x9E3779B9u + 0x7F4A7C15u;
    h ^= pixel + 0x85EBCA6Bu + (h << 6) + (h >> 2);
    h ^= frame + 0xC2B2AE35u + (h << 6) + (h >> 2);
    h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
    return h;
}

float vso_sampler_next(uint32_t* state)
{
    uint32_t s = *state * 747796405u + 2891336453u;
    *state = s;
    uint32_t w = ((s >> ((s >> 28u) + 4u)) ^ s) * 277803737u;
    w = (w >> 22u) ^ w;
    return (float)(w >> 8) * (1.0f / 16777216.0f);
}

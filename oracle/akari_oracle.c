/*
 * oracle/akari_oracle.c — TEST INFRASTRUCTURE ONLY. Not part of the product path.
 *
 * Plain-C CPU restatement of AkariRender's unidirectional path-tracing hot path, written from the reference's
 * algorithm (every function cites the /root/reference/src/akari file:line it follows). It exists so that the
 * CUDA path can be checked on boxes where /root/reference is absent, and so the extensions the reference does
 * not have (Russian roulette, dielectric) have a CPU statement at all.
 *
 * PARITY PIN: this file is checked bit-for-bit against the reference's own code (oracle/_ref/libakari_ref.so,
 * built from the reference sources by oracle/Makefile) in tests/test_oracle_vs_ref.py, and against the golden
 * vectors generated from that build (tests/golden/). The reference's own tests hold no vector for this path
 * (SURVEY.md §4), so those generated vectors are the pin. The extensions are "parity unpinned by the reference".
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * Arithmetic contract: IEEE fp32, no FMA contraction (the reference host build is -mavx2 without -mfma,
 * CMakeLists.txt:28), libm sinf/cosf/atanf/sqrtf, every expression in the reference's operator order.
 * Intersection acceleration is NOT a restatement of the reference's SBVH: a median-split binary BVH with a
 * conservative (padded) slab test returns the nearest hit over ALL triangles under the reference's triangle
 * test. It differs from the reference BVH only (i) between candidates with bit-identical t (visit-order ties)
 * and (ii) where the reference's unpadded slab test (kernel/bvh-accelerator.h:89-103) culls a box whose
 * triangle the ray does hit (a reference false-miss).
 */
#define _POSIX_C_SOURCE 200809L
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../include/akari_b200.h"

typedef struct { float x, y, z; } v3;

/* common/array.h:203-210 dot: s = a0*b0; s += a1*b1; s += a2*b2 */
static inline float dot3(v3 a, v3 b) { float s = a.x * b.x; s += a.y * b.y; s += a.z * b.z; return s; }
/* common/math.h:177-181 */
static inline v3 cross3(v3 a, v3 b) {
    v3 r = {(a.y * b.z) - (a.z * b.y), (a.z * b.x) - (a.x * b.z), (a.x * b.y) - (a.y * b.x)};
    return r;
}
static inline v3 add3(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static inline v3 sub3(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline v3 mul3(v3 a, v3 b) { v3 r = {a.x * b.x, a.y * b.y, a.z * b.z}; return r; }
static inline v3 scale3(float s, v3 a) { v3 r = {s * a.x, s * a.y, s * a.z}; return r; }
static inline v3 div3s(v3 a, float s) { v3 r = {a.x / s, a.y / s, a.z / s}; return r; }
static inline v3 neg3(v3 a) { v3 r = {-a.x, -a.y, -a.z}; return r; }
static inline v3 mk3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
/* common/array.h:283-286 normalize: a / sqrt(dot(a,a)) (true divisions) */
static inline v3 normalize3(v3 a) { return div3s(a, sqrtf(dot3(a, a))); }
/* std::max / std::min semantics: max(a,b) = (a<b)?b:a ; min(a,b) = (b<a)?b:a  (common/array.h:41-46) */
static inline float maxf_(float a, float b) { return (a < b) ? b : a; }
static inline float minf_(float a, float b) { return (b < a) ? b : a; }
/* common/math.h:49-51 lerp3 */
static inline v3 lerp3v(v3 a, v3 b, v3 c, float u, float v) {
    float w = 1.0f - u - v;
    return add3(add3(scale3(w, a), scale3(u, b)), scale3(v, c));
}

/* common/math.h:34-42 */
#define K_PI ((float)3.1415926535897932384f)
#define K_PI2 (K_PI / 2.0f)
#define K_PI4 (K_PI / 4.0f)
#define K_INVPI (1.0f / K_PI)
#define K_EPS 0.001f
#define K_SHADOW_EPS 0.0001f

/* ---- samplers: kernel/sampler.h:28-66 ------------------------------------------------------------------- */
typedef struct { int kind; uint32_t lcg; uint64_t pcg; } sampler_t;

static uint32_t pcg32(sampler_t *s) { /* kernel/sampler.h:33-41 */
    uint64_t x = s->pcg;
    unsigned count = (unsigned)(x >> 59);
    s->pcg = x * 6364136223846793005ull + 1442695040888963407ull;
    x ^= x >> 18;
    uint32_t v = (uint32_t)(x >> 27);
    return v >> count | v << (-count & 31);
}
static void sampler_set_index(sampler_t *s, uint64_t idx) {
    if (s->kind == AKB_SAMPLER_PCG) { /* :42-45 */
        s->pcg = idx + 1442695040888963407ull;
        (void)pcg32(s);
    } else { /* :59 */
        s->lcg = (uint32_t)(idx & 0xffffffffu);
    }
}
static float next1d(sampler_t *s) {
    if (s->kind == AKB_SAMPLER_PCG) /* :50: Float(pcg32()) / (float)0xffffffff ; (float)0xffffffff == 2^32 */
        return (float)pcg32(s) / (float)0xffffffffu;
    s->lcg = 1103515245u * s->lcg + 12345u; /* :60-63 */
    return (float)s->lcg / (float)0xFFFFFFFFu;
}
/* next2d() = float2(next1d(), next1d()) (:51,:64). Argument evaluation order is compiler-defined; the oracle
 * build (g++ 13, x86-64) evaluates right to left: y gets the FIRST draw, x the SECOND. Pinned by the KAT in
 * tests/golden. */
static void next2d(sampler_t *s, float *x, float *y) { *y = next1d(s); *x = next1d(s); }

/* ---- camera: kernel/camera.h:36-86 ------------------------------------------------------------------------ */
typedef struct { float m[4][4]; } m44;
static m44 m44_mul(m44 a, m44 b) { /* common/math.h:97-105: m(i,j) = dot(row(i), rhs.col(j)) */
    m44 r;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float s = a.m[i][0] * b.m[0][j];
            for (int k = 1; k < 4; k++) s += a.m[i][k] * b.m[k][j];
            r.m[i][j] = s;
        }
    return r;
}
static m44 m44_scale(float x, float y, float z) { m44 r; memset(&r, 0, sizeof r); r.m[0][0] = x; r.m[1][1] = y; r.m[2][2] = z; r.m[3][3] = 1; return r; }
static m44 m44_translate(float x, float y, float z) { m44 r = m44_scale(1, 1, 1); r.m[0][3] = x; r.m[1][3] = y; r.m[2][3] = z; return r; }

typedef struct { m44 c2w, r2c; int w, h; } camera_t;

static void camera_init(camera_t *c, const akb_camera *d) {
    memcpy(c->c2w.m, d->c2w, sizeof(float) * 16);
    c->w = d->width; c->h = d->height;
    /* preprocess(), kernel/camera.h:45-59 */
    m44 m = m44_scale(1, 1, 1);
    m = m44_mul(m44_scale(1.0f / (float)c->w, 1.0f / (float)c->h, 1), m);
    m = m44_mul(m44_scale(2, 2, 1), m);
    m = m44_mul(m44_translate(-1, -1, 0), m);
    m = m44_mul(m44_scale(1, -1, 1), m);
    float s = atanf(d->fov / 2);
    if (c->w > c->h) m = m44_mul(m44_scale(s, s * (float)c->h / (float)c->w, 1), m);
    else m = m44_mul(m44_scale(s * (float)c->w / (float)c->h, s, 1), m);
    c->r2c = m;
}
typedef struct { v3 o, d; float tmin, tmax; } ray_t;

/* Transform::apply_point (common/math.h:251-259) */
static v3 apply_point(const m44 *m, v3 p) {
    float v[4];
    for (int i = 0; i < 4; i++) { float s = m->m[i][0] * p.x; s += m->m[i][1] * p.y; s += m->m[i][2] * p.z; s += m->m[i][3] * 1.0f; v[i] = s; }
    v3 q = {v[0], v[1], v[2]};
    if (v[3] != 1.0) q = div3s(q, v[3]);
    return q;
}
/* PerspectiveCamera::generate_ray, pinhole branch (kernel/camera.h:67-86) */
static ray_t camera_generate_ray(const camera_t *c, float u2x, float u2y, int px, int py) {
    float fx = (float)px + u2x, fy = (float)py + u2y;
    v3 p = apply_point(&c->r2c, mk3(fx, fy, 0.0f));
    v3 d = normalize3(sub3(mk3(p.x, p.y, 0), mk3(0, 0, 1)));
    ray_t r;
    r.o = apply_point(&c->c2w, mk3(0, 0, 0));
    /* apply_vector: m3 * v (common/math.h:261) */
    r.d.x = dot3(mk3(c->c2w.m[0][0], c->c2w.m[0][1], c->c2w.m[0][2]), d);
    r.d.y = dot3(mk3(c->c2w.m[1][0], c->c2w.m[1][1], c->c2w.m[1][2]), d);
    r.d.z = dot3(mk3(c->c2w.m[2][0], c->c2w.m[2][1], c->c2w.m[2][2]), d);
    r.tmin = K_EPS; r.tmax = INFINITY; /* Ray ctor defaults, common/math.h:192-194 */
    return r;
}

/* ---- scene: flattened Scene<C> (kernel/scene.h:50-91, kernel/instance.h:30-97) -------------------------- */
typedef struct {
    v3 v[3], n[3];
    float uv[3][2];
    int32_t mat; /* index into the scene material table, -1 = none */
    int32_t geom, prim;
} tri_t;

typedef struct { float lo[3], hi[3]; int32_t left, right, first, count; } bnode_t;

typedef struct {
    v3 v[3];
    v3 color;
    float area_dummy;
    float uv[3][2];
    int32_t tex; /* image texture of the emissive colour, or -1 */
} light_t;

typedef struct oracle_scene {
    uint64_t n_tris;
    tri_t *tris;
    uint64_t *geom_offset; /* first flattened triangle of each mesh */
    uint32_t n_geoms;
    uint32_t n_mats;
    akb_material *mats;
    uint32_t n_tex;
    akb_texture *tex; /* deep copies */
    camera_t cam;
    /* lights: core/nodes/scene.cpp:51-92 ; Distribution1D: common/distribution.h:47-81 */
    uint32_t n_lights;
    light_t *lights;
    float *func, *cdf, func_int;
    /* acceleration (not a restatement, see header) */
    bnode_t *nodes; int32_t n_nodes; int32_t *order;
    int brute_force;
} oracle_scene;

/* MeshInstance::intersect (kernel/instance.h:42-80). best_t plays RayHit::t. Returns 1 when the hit record was
 * updated (closest) — or, with best_t == NULL, when any t in (tmin,tmax) passes (occlude path). */
static int tri_intersect(const ray_t *ray, const tri_t *tr, float *best_t, float *hu, float *hv) {
    v3 v0 = tr->v[0];
    v3 e1 = sub3(tr->v[1], v0), e2 = sub3(tr->v[2], v0);
    v3 h = cross3(ray->d, e2);
    float a = dot3(e1, h);
    if (a > -1e-6f && a < 1e-6f) return 0;
    float f = 1.0f / a;
    v3 s = sub3(ray->o, v0);
    float u = f * dot3(s, h);
    if (u < 0.0 || u > 1.0) return 0;
    v3 q = cross3(s, e1);
    float v = f * dot3(ray->d, q);
    if (v < 0.0 || u + v > 1.0) return 0;
    float t = f * dot3(e2, q);
    if (t > ray->tmin && t < ray->tmax) {
        if (best_t) {
            if (t < *best_t) { *hu = u; *hv = v; *best_t = t; return 1; }
            return 0;
        }
        return 1;
    }
    return 0;
}

/* Conservative slab test: entry distance or -1. Padded so that it never culls a box the exact ray enters. */
static int box_hit(const bnode_t *n, const ray_t *ray, const float *invd, float best) {
    float t0 = 0.0f, t1 = ray->tmax < best ? ray->tmax : best;
    const float o[3] = {ray->o.x, ray->o.y, ray->o.z};
    for (int a = 0; a < 3; a++) {
        float tn = (n->lo[a] - o[a]) * invd[a], tf = (n->hi[a] - o[a]) * invd[a];
        if (tn != tn || tf != tf) continue; /* 0*inf: ray lies in the slab plane -> no constraint */
        if (tn > tf) { float x = tn; tn = tf; tf = x; }
        tn = tn - fabsf(tn) * 4e-7f; tf = tf + fabsf(tf) * 4e-7f;
        if (tn > t0) t0 = tn;
        if (tf < t1) t1 = tf;
        if (t0 > t1) return 0;
    }
    return 1;
}

static void tri_bounds(const tri_t *t, float *lo, float *hi) {
    for (int a = 0; a < 3; a++) {
        float x0 = ((const float *)&t->v[0])[a], x1 = ((const float *)&t->v[1])[a], x2 = ((const float *)&t->v[2])[a];
        lo[a] = fminf(x0, fminf(x1, x2)); hi[a] = fmaxf(x0, fmaxf(x1, x2));
    }
}
typedef struct { float c[3]; int32_t id; } cent_t;
static int g_sort_axis;
static int cent_cmp(const void *a, const void *b) {
    float x = ((const cent_t *)a)->c[g_sort_axis], y = ((const cent_t *)b)->c[g_sort_axis];
    return (x > y) - (x < y);
}
static int32_t build_rec(oracle_scene *s, cent_t *c, int32_t first, int32_t count) {
    int32_t id = s->n_nodes++;
    bnode_t *n = &s->nodes[id];
    float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int a = 0; a < 3; a++) { n->lo[a] = INFINITY; n->hi[a] = -INFINITY; }
    for (int32_t i = first; i < first + count; i++) {
        float lo[3], hi[3];
        tri_bounds(&s->tris[c[i].id], lo, hi);
        for (int a = 0; a < 3; a++) {
            n->lo[a] = fminf(n->lo[a], lo[a]); n->hi[a] = fmaxf(n->hi[a], hi[a]);
            clo[a] = fminf(clo[a], c[i].c[a]); chi[a] = fmaxf(chi[a], c[i].c[a]);
        }
    }
    n->first = first; n->count = count; n->left = n->right = -1;
    if (count <= 4) return id;
    int ax = 0;
    if (chi[1] - clo[1] > chi[ax] - clo[ax]) ax = 1;
    if (chi[2] - clo[2] > chi[ax] - clo[ax]) ax = 2;
    if (!(chi[ax] - clo[ax] > 0)) return id;
    g_sort_axis = ax;
    qsort(c + first, count, sizeof(cent_t), cent_cmp);
    int32_t half = count / 2;
    int32_t l = build_rec(s, c, first, half);
    int32_t r = build_rec(s, c, first + half, count - half);
    n = &s->nodes[id];
    n->left = l; n->right = r; n->count = 0;
    return id;
}
static void build_accel(oracle_scene *s) {
    int64_t n = (int64_t)s->n_tris;
    s->nodes = (bnode_t *)malloc(sizeof(bnode_t) * (size_t)(2 * n + 1));
    s->order = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n + 1));
    cent_t *c = (cent_t *)malloc(sizeof(cent_t) * (size_t)(n + 1));
    for (int64_t i = 0; i < n; i++) {
        float lo[3], hi[3];
        tri_bounds(&s->tris[i], lo, hi);
        for (int a = 0; a < 3; a++) c[i].c[a] = 0.5f * (lo[a] + hi[a]);
        c[i].id = (int32_t)i;
    }
    s->n_nodes = 0;
    if (n > 0) build_rec(s, c, 0, (int32_t)n);
    for (int64_t i = 0; i < n; i++) s->order[i] = c[i].id;
    free(c);
}

/* Scene::intersect semantics (kernel/scene.cpp:26-45 -> kernel/bvh-accelerator.h:488-518,651-664): nearest t under
 * tri_intersect over all triangles; strict t < best keeps the first-visited among equal t. Returns tri index or -1. */
static int64_t scene_intersect(const oracle_scene *s, const ray_t *ray, float *t_out, float *u_out, float *v_out) {
    float best = INFINITY, hu = 0, hv = 0;
    int64_t hit = -1;
    if (s->brute_force || s->n_nodes == 0) {
        for (uint64_t i = 0; i < s->n_tris; i++)
            if (tri_intersect(ray, &s->tris[i], &best, &hu, &hv)) hit = (int64_t)i;
    } else {
        float invd[3] = {1.0f / ray->d.x, 1.0f / ray->d.y, 1.0f / ray->d.z};
        int32_t stack[128]; int sp = 0; stack[sp++] = 0;
        while (sp) {
            const bnode_t *n = &s->nodes[stack[--sp]];
            if (!box_hit(n, ray, invd, best)) continue;
            if (n->left < 0) {
                for (int32_t i = n->first; i < n->first + n->count; i++)
                    if (tri_intersect(ray, &s->tris[s->order[i]], &best, &hu, &hv)) hit = s->order[i];
            } else { stack[sp++] = n->right; stack[sp++] = n->left; }
        }
    }
    *t_out = best; *u_out = hu; *v_out = hv;
    return hit;
}
/* Scene::occlude (kernel/bvh-accelerator.h:519-547): any triangle with tmin < t < tmax */
static int scene_occlude(const oracle_scene *s, const ray_t *ray) {
    if (s->brute_force || s->n_nodes == 0) {
        for (uint64_t i = 0; i < s->n_tris; i++)
            if (tri_intersect(ray, &s->tris[i], NULL, NULL, NULL)) return 1;
        return 0;
    }
    float invd[3] = {1.0f / ray->d.x, 1.0f / ray->d.y, 1.0f / ray->d.z};
    int32_t stack[128]; int sp = 0; stack[sp++] = 0;
    while (sp) {
        const bnode_t *n = &s->nodes[stack[--sp]];
        if (!box_hit(n, ray, invd, INFINITY)) continue;
        if (n->left < 0) {
            for (int32_t i = n->first; i < n->first + n->count; i++)
                if (tri_intersect(ray, &s->tris[s->order[i]], NULL, NULL, NULL)) return 1;
        } else { stack[sp++] = n->right; stack[sp++] = n->left; }
    }
    return 0;
}

/* ---- BSDF helpers: kernel/bsdf-funcs.h:26-89 (local frame is y-up) --------------------------------------- */
static inline int same_hemisphere(v3 wo, v3 wi) { return wo.y * wi.y >= 0; }
static inline float cos2_theta(v3 w) { return w.y * w.y; }
static inline float sin2_theta(v3 w) { return 1 - cos2_theta(w); }
static inline float tan2_theta(v3 w) { return sin2_theta(w) / cos2_theta(w); }
/* bsdf::reflect (:58-60): -1.0f * w + 2.0f * dot(w, n) * n */
static inline v3 reflect3(v3 w, v3 n) { return add3(scale3(-1.0f, w), scale3(2.0f * dot3(w, n), n)); }
/* bsdf::refract (:62-73) */
static int refract3(v3 wi, v3 n, float eta, v3 *wt) {
    float cosThetaI = dot3(n, wi);
    float sin2ThetaI = fmaxf(0.0f, 1.0f - cosThetaI * cosThetaI);
    float sin2ThetaT = eta * eta * sin2ThetaI;
    if (sin2ThetaT >= 1) return 0;
    float cosThetaT = sqrtf(1 - sin2ThetaT);
    *wt = add3(scale3(eta, neg3(wi)), scale3(eta * cosThetaI - cosThetaT, n));
    return 1;
}
/* bsdf::fr_dielectric (:74-89) */
static float fr_dielectric(float cosThetaI, float etaI, float etaT) {
    int entering = cosThetaI > 0.f;
    if (!entering) { float x = etaI; etaI = etaT; etaT = x; cosThetaI = fabsf(cosThetaI); }
    float sinThetaI = sqrtf(maxf_(0.0f, 1 - cosThetaI * cosThetaI));
    float sinThetaT = etaI / etaT * sinThetaI;
    if (sinThetaT >= 1) return 1;
    float cosThetaT = sqrtf(maxf_(0.0f, 1 - sinThetaT * sinThetaT));
    float Rpar = ((etaT * cosThetaI) - (etaI * cosThetaT)) / ((etaT * cosThetaI) + (etaI * cosThetaT));
    float Rper = ((etaI * cosThetaI) - (etaT * cosThetaT)) / ((etaI * cosThetaI) + (etaT * cosThetaT));
    return (float)(0.5 * (double)(Rpar * Rpar + Rper * Rper));
}

/* ---- sampling: kernel/sampling.h:32-69 ---------------------------------------------------------------------- */
static void concentric_disk(float ux, float uy, float *ox, float *oy) {
    float x = 2.f * ux - 1, y = 2.f * uy - 1;
    if (x == 0 && y == 0) { *ox = 0; *oy = 0; return; }
    float theta, r;
    if (fabsf(x) > fabsf(y)) { r = x; theta = K_PI4 * (y / x); }
    else { r = y; theta = K_PI2 - K_PI4 * (x / y); }
    *ox = r * cosf(theta); *oy = r * sinf(theta);
}
static v3 cosine_hemisphere(float ux, float uy) {
    float x, y; concentric_disk(ux, uy, &x, &y);
    float r = x * x; r += y * y;
    float h = sqrtf(maxf_(0.0f, 1.0f - r));
    return mk3(x, h, y);
}

/* ---- Frame: common/math.h:201-225 ---------------------------------------------------------------------------- */
typedef struct { v3 n, T, B; } frame_t;
static frame_t frame_make(v3 v1) {
    frame_t f; f.n = v1;
    if (fabsf(v1.x) > fabsf(v1.y)) f.T = div3s(mk3(-v1.z, 0, v1.x), sqrtf(v1.x * v1.x + v1.z * v1.z));
    else f.T = div3s(mk3(0, v1.z, -v1.y), sqrtf(v1.y * v1.y + v1.z * v1.z));
    f.B = normalize3(cross3(v1, f.T));
    return f;
}
static v3 w2l(const frame_t *f, v3 v) { return mk3(dot3(f->T, v), dot3(f->n, v), dot3(f->B, v)); }
static v3 l2w(const frame_t *f, v3 v) { return add3(add3(scale3(v.x, f->T), scale3(v.y, f->n)), scale3(v.z, f->B)); }

/* ---- closures: kernel/material.h:57-138, kernel/microfacet.h:74-157 ------------------------------------------- */
enum { CL_DIFFUSE = 0, CL_GGX = 1, CL_DIELECTRIC = 2, CL_NULL = 3 };
typedef struct { int kind; v3 R; float alpha; float ior; frame_t frame; float choice_pdf; } bsdf_t;

static float ggx_D(float alpha, v3 m) { /* microfacet.h:74-82 */
    if (m.y <= 0.0f) return 0.0f;
    float a2 = alpha * alpha, c2 = cos2_theta(m), t2 = tan2_theta(m), at = (a2 + t2);
    return a2 / (K_PI * c2 * c2 * at * at);
}
static float ggx_G1(float alpha, v3 v, v3 m) { /* microfacet.h:84-89 — theta of m, evaluated in double */
    if (dot3(v, m) * v.y <= 0) return 0.0f;
    return (float)(2.0 / (1.0 + sqrt(1.0 + (double)(alpha * alpha * tan2_theta(m)))));
}
static v3 closure_evaluate(const bsdf_t *b, v3 wo, v3 wi) {
    v3 zero = {0, 0, 0};
    if (b->kind == CL_DIFFUSE) { /* material.h:69-74 */
        if (same_hemisphere(wo, wi)) return scale3(K_INVPI, b->R);
        return zero;
    }
    if (b->kind == CL_GGX) { /* material.h:103-121 */
        if (!same_hemisphere(wo, wi)) return zero;
        float cosThetaO = fabsf(wo.y), cosThetaI = fabsf(wi.y);
        v3 wh = add3(wo, wi);
        if (cosThetaI == 0 || cosThetaO == 0) return zero;
        if (wh.x == 0 && wh.y == 0 && wh.z == 0) return zero;
        wh = normalize3(wh);
        if (wh.y < 0) wh = neg3(wh);
        float F = 1.0f;
        float s = ggx_D(b->alpha, wh) * (ggx_G1(b->alpha, wo, wh) * ggx_G1(b->alpha, wi, wh)) * F /
                  (4.0f * cosThetaI * cosThetaO);
        return mk3(b->R.x * s, b->R.y * s, b->R.z * s);
    }
    return zero; /* specular dielectric (extension) and null closure evaluate to 0 */
}
/* returns f; writes wi (local), pdf */
static v3 closure_sample(const bsdf_t *b, float ux, float uy, v3 wo, v3 *wi, float *pdf) {
    v3 zero = {0, 0, 0};
    if (b->kind == CL_DIFFUSE) { /* material.h:76-85 */
        *wi = cosine_hemisphere(ux, uy);
        if (!same_hemisphere(wo, *wi)) wi->y = -wi->y;
        *pdf = fabsf(wi->y) * K_INVPI;
        return scale3(K_INVPI, b->R);
    }
    if (b->kind == CL_GGX) { /* material.h:122-137 + microfacet.h:123-148 */
        float phi = 2 * K_PI * uy;
        float t2 = b->alpha * b->alpha * ux / (1 - ux);
        float cosTheta = 1.0f / sqrtf(1 + t2);
        float sinTheta = sqrtf(maxf_(0.0f, 1 - cosTheta * cosTheta));
        v3 wh = mk3(cosf(phi) * sinTheta, cosTheta, sinf(phi) * sinTheta);
        if (!same_hemisphere(wo, wh)) wh = neg3(wh);
        *wi = reflect3(wo, wh);
        if (!same_hemisphere(wo, *wi)) { *pdf = 0; return zero; }
        if (wh.y < 0) wh = neg3(wh);
        *pdf = (ggx_D(b->alpha, wh) * fabsf(wh.y)) / (4.0f * fabsf(dot3(wo, wh)));
        return closure_evaluate(b, wo, *wi);
    }
    if (b->kind == CL_DIELECTRIC) {
        /* EXTENSION (not in the reference): smooth dielectric built from the reference's own
         * fr_dielectric/refract helpers. Reflect with probability F, else refract; no radiance eta^2 scaling. */
        float F = fr_dielectric(wo.y, 1.0f, b->ior);
        if (ux < F) {
            *wi = mk3(-wo.x, wo.y, -wo.z);
            *pdf = F;
            float s = F / fabsf(wi->y);
            return mk3(b->R.x * s, b->R.y * s, b->R.z * s);
        }
        int entering = wo.y > 0;
        float eta = entering ? 1.0f / b->ior : b->ior;
        v3 n = mk3(0, entering ? 1.0f : -1.0f, 0);
        if (!refract3(wo, n, eta, wi)) { *pdf = 0; return zero; }
        *pdf = 1.0f - F;
        float s = (1.0f - F) / fabsf(wi->y);
        return mk3(b->R.x * s, b->R.y * s, b->R.z * s);
    }
    *pdf = 0;
    return zero;
}

/* Material::select_material + get_bsdf (kernel/material.h:255-297). u is ctx.u1[0] (drawn from a COPY of the
 * sampler, material.h:198-202). Emissive/null leaves give a null closure (get_bsdf0 returns BSDF()). */
/* ImageTexture::evaluate (kernel/texture.h:45-49) through TImage::operator()(float2) (core/image.hpp:46-68) */
static v3 tex_eval(const oracle_scene *s, int32_t tex, float tu, float tv) {
    const akb_texture *t = &s->tex[tex];
    float x = fmodf(tu, 1.0f), y = fmodf(tv, 1.0f);
    y = 1.0f - y;
    int ix = (int)(x * (float)t->width), iy = (int)(y * (float)t->height);
    if (ix < 0) ix = 0; if (ix > t->width - 1) ix = t->width - 1;
    if (iy < 0) iy = 0; if (iy > t->height - 1) iy = t->height - 1;
    const float *px = t->rgba + 4 * ((size_t)ix + (size_t)iy * (size_t)t->width);
    return mk3(px[0], px[1], px[2]);
}
/* Triangle::texcoord(uv), kernel/shape.h:40-42 */
static void tri_texcoord(const float uv[3][2], float u, float v, float *tu, float *tv) {
    v3 t = lerp3v(mk3(uv[0][0], uv[0][1], 0.0f), mk3(uv[1][0], uv[1][1], 0.0f), mk3(uv[2][0], uv[2][1], 0.0f), u, v);
    *tu = t.x; *tv = t.y;
}

static bsdf_t material_get_bsdf(const oracle_scene *s, int32_t mat, float u, v3 ns, float tu, float tv) {
    bsdf_t b; memset(&b, 0, sizeof b);
    float choice_pdf = 1.0f;
    const akb_material *m = &s->mats[mat];
    int guard = 0;
    while (m->type == AKB_MAT_MIX && guard++ < 64) {
        float frac = m->fraction_tex >= 0 ? tex_eval(s, m->fraction_tex, tu, tv).x : m->fraction;
        if (u < frac) { u = u / frac; m = &s->mats[m->child_b]; choice_pdf *= 1.0f / frac; }
        else { u = (u - frac) / (1.0f - frac); m = &s->mats[m->child_a]; choice_pdf *= 1.0f / (1.0f - frac); }
    }
    b.frame = frame_make(ns);
    b.choice_pdf = choice_pdf;
    b.R = mk3(m->color[0], m->color[1], m->color[2]);
    if (m->color_tex >= 0 && (m->type == AKB_MAT_DIFFUSE || m->type == AKB_MAT_GLOSSY)) b.R = tex_eval(s, m->color_tex, tu, tv);
    if (m->type == AKB_MAT_DIFFUSE) b.kind = CL_DIFFUSE;
    else if (m->type == AKB_MAT_GLOSSY) {
        b.kind = CL_GGX;
        float r = m->roughness_tex >= 0 ? tex_eval(s, m->roughness_tex, tu, tv).x : m->roughness;
        r *= r; b.alpha = r;
    }
    else if (m->type == AKB_MAT_DIELECTRIC) { b.kind = CL_DIELECTRIC; b.ior = m->ior; }
    else b.kind = CL_NULL;
    return b;
}

/* ---- lights: kernel/light.h:47-72, common/distribution.h:32-81, kernel/scene.h:79-90 -------------------- */
static int select_light(const oracle_scene *s, float u, float *pdf) {
    if (s->n_lights == 0) { *pdf = 0.0f; return -1; }
    /* upper_bound(0, cdf.size(), cdf[idx] <= u) */
    int lo = 0, hi = (int)s->n_lights + 1, last = hi;
    while (lo < hi) { int mid = (lo + hi) / 2; if (s->cdf[mid] <= u) lo = mid + 1; else hi = mid; }
    int i = hi - 1; if (i < 0) i = 0; if (i > last - 2) i = last - 2;
    *pdf = s->func[i] / (s->func_int * (float)s->n_lights);
    return i;
}

typedef struct { v3 L, beta; int depth, max_depth; sampler_t smp; uint64_t n_closest, n_shadow; } path_t;

static inline int is_black(v3 c) { return !((c.x > 0.0f) || (c.y > 0.0f) || (c.z > 0.0f)); }

/* Optional per-thread log of the rays a path traces (akor_pixel_rays: diagnosis of single-pixel differences).
 * 9 floats per ray: kind (0 closest, 1 any hit), o, d, tmin, tmax. */
static __thread float *g_raylog = NULL;
static __thread int g_raylog_n = 0, g_raylog_cap = 0;
static void log_ray(int kind, const ray_t *r) {
    if (!g_raylog || g_raylog_n >= g_raylog_cap) return;
    float *o = g_raylog + 9 * g_raylog_n++;
    o[0] = (float)kind; o[1] = r->o.x; o[2] = r->o.y; o[3] = r->o.z; o[4] = r->d.x; o[5] = r->d.y; o[6] = r->d.z;
    o[7] = r->tmin; o[8] = r->tmax;
}

/* GenericPathTracer::run_megakernel (kernel/pathtracer.h:61-163) + extensions (rr_start_depth, dielectric). */
static void run_path(const oracle_scene *s, path_t *pt, int px, int py, int rr_start) {
    float u1x, u1y, u2x, u2y;
    next2d(&pt->smp, &u2x, &u2y); /* g++ evaluates generate_ray's arguments right to left: u2 first */
    next2d(&pt->smp, &u1x, &u1y);
    (void)u1x; (void)u1y;
    ray_t ray = camera_generate_ray(&s->cam, u2x, u2y, px, py);
    for (;;) {
        float t, hu, hv;
        pt->n_closest++;
        log_ray(0, &ray);
        int64_t ti = scene_intersect(s, &ray, &t, &hu, &hv);
        if (ti < 0) break; /* on_miss is empty */
        const tri_t *tr = &s->tris[ti];
        v3 wo = neg3(ray.d);
        /* SurfaceInteraction(uv, triangle), kernel/interaction.h:40-42 */
        v3 p = lerp3v(tr->v[0], tr->v[1], tr->v[2], hu, hv);
        v3 ng = normalize3(cross3(sub3(tr->v[1], tr->v[0]), sub3(tr->v[2], tr->v[0])));
        v3 ns = lerp3v(tr->n[0], tr->n[1], tr->n[2], hu, hv);
        if (tr->mat < 0) break; /* reference dereferences null here; defined as "terminate" */
        const akb_material *mat = &s->mats[tr->mat];
        float tu, tv; /* si.texcoords, kernel/interaction.h:40-42 */
        tri_texcoord(tr->uv, hu, hv, &tu, &tv);
        /* on_surface_scatter, kernel/pathtracer.h:96-132 */
        if (mat->type == AKB_MAT_EMISSIVE) {
            if (pt->depth == 0) {
                int face_front = dot3(neg3(wo), ng) < 0.0f;
                if (mat->double_sided || face_front) {
                    v3 Le = mat->color_tex >= 0 ? tex_eval(s, mat->color_tex, tu, tv) : mk3(mat->color[0], mat->color[1], mat->color[2]);
                    pt->L = add3(pt->L, mul3(pt->beta, Le));
                }
            }
            break;
        }
        if (!(pt->depth < pt->max_depth)) break;
        /* MaterialEvalContext draws from a copy: ctx.u1[0] == the x of the next2d() the BSDF sample uses */
        float bux, buy;
        next2d(&pt->smp, &bux, &buy);
        bsdf_t bsdf = material_get_bsdf(s, tr->mat, bux, ns, tu, tv);
        v3 wo_l = w2l(&bsdf.frame, wo), wi_l;
        float pdf;
        v3 f = closure_sample(&bsdf, bux, buy, wo_l, &wi_l, &pdf);
        v3 wi = l2w(&bsdf.frame, wi_l);
        pdf *= bsdf.choice_pdf;
        if (pdf == 0.0f) break;
        float angi = fabsf(dot3(ng, wi));
        ray_t next; next.o = p; next.d = wi; next.tmin = K_EPS / angi; next.tmax = INFINITY;
        v3 ev_beta = div3s(scale3(angi, f), pdf); /* sample.f * |ng.wi| / sample.pdf */
        /* compute_direct_lighting(si, hit, select_light(scene)), kernel/pathtracer.h:65-91 */
        float slx, sly; next2d(&pt->smp, &slx, &sly);
        float light_pdf; int li = select_light(s, slx, &light_pdf);
        if (li >= 0) {
            float lux, luy; next2d(&pt->smp, &lux, &luy);
            const light_t *lt = &s->lights[li];
            /* AreaLight::sample, kernel/light.h:58-71 */
            float su0 = sqrtf(lux); float b0 = 1 - su0, b1 = luy * su0;
            v3 lp = lerp3v(lt->v[0], lt->v[1], lt->v[2], b0, b1);
            v3 lcr = cross3(sub3(lt->v[1], lt->v[0]), sub3(lt->v[2], lt->v[0]));
            v3 lng = normalize3(lcr);
            v3 lwi = sub3(lp, p);
            float dist_sqr = dot3(lwi, lwi);
            lwi = div3s(lwi, sqrtf(dist_sqr));
            float area = sqrtf(dot3(lcr, lcr)) * 0.5f;
            float lpdf = dist_sqr / maxf_(0.0f, -dot3(lwi, lng)) / area;
            ray_t sh; sh.o = lp; sh.d = neg3(lwi); sh.tmin = K_EPS / fabsf(dot3(lwi, lng));
            sh.tmax = sqrtf(dist_sqr) * (1.0f - K_SHADOW_EPS);
            if (!(lpdf <= 0.0)) {
                light_pdf *= lpdf;
                v3 fe = closure_evaluate(&bsdf, w2l(&bsdf.frame, wo), w2l(&bsdf.frame, lwi));
                v3 Le = lt->color;
                if (lt->tex >= 0) { /* sample.L = color->evaluate(triangle.texcoord(coords)), kernel/light.h:65 */
                    float lu, lv;
                    tri_texcoord(lt->uv, b0, b1, &lu, &lv);
                    Le = tex_eval(s, lt->tex, lu, lv);
                }
                v3 ff = scale3(fabsf(dot3(ns, lwi)), mul3(Le, fe));
                v3 color = div3s(mul3(pt->beta, ff), light_pdf);
                if (!is_black(color)) {
                    pt->n_shadow++;
                    log_ray(1, &sh);
                    if (!scene_occlude(s, &sh)) pt->L = add3(pt->L, color);
                }
            }
        }
        pt->beta = mul3(pt->beta, ev_beta);
        pt->depth++;
        ray = next;
        if (rr_start > 0 && pt->depth >= rr_start) {
            /* EXTENSION: Russian roulette. q = min(max(beta), 0.95); one next1d() draw. */
            float q = minf_(maxf_(maxf_(pt->beta.x, pt->beta.y), pt->beta.z), 0.95f);
            float u = next1d(&pt->smp);
            if (!(u < q)) break;
            pt->beta = div3s(pt->beta, q);
        }
    }
}

/* AmbientOcclusion::render's Li (kernel/integrators/cpu/integrator.cpp:42-60): primary hit -> one cosine-weighted ray
 * about the geometric normal from trig.p(uv) (Ray defaults tmin = Eps, tmax = inf); 0 if its closest hit is nearer
 * than `occlude`, else 1; a primary miss gives 0. Returns the grey value; *n_rays counts Scene::intersect calls. */
static float run_ao(const oracle_scene *s, sampler_t *smp, int px, int py, float occlude, uint64_t *n_rays) {
    float u1x, u1y, u2x, u2y;
    next2d(smp, &u2x, &u2y); /* right-to-left argument evaluation, as in run_path */
    next2d(smp, &u1x, &u1y);
    (void)u1x; (void)u1y;
    ray_t ray = camera_generate_ray(&s->cam, u2x, u2y, px, py);
    float t, hu, hv;
    (*n_rays)++;
    int64_t ti = scene_intersect(s, &ray, &t, &hu, &hv);
    if (ti < 0) return 0.0f;
    const tri_t *tr = &s->tris[ti];
    v3 ng = normalize3(cross3(sub3(tr->v[1], tr->v[0]), sub3(tr->v[2], tr->v[0])));
    frame_t fr = frame_make(ng);
    float ux, uy;
    next2d(smp, &ux, &uy);
    v3 w = l2w(&fr, cosine_hemisphere(ux, uy));
    ray_t r2; r2.o = lerp3v(tr->v[0], tr->v[1], tr->v[2], hu, hv); r2.d = w; r2.tmin = K_EPS; r2.tmax = INFINITY;
    (*n_rays)++;
    if (scene_intersect(s, &r2, &t, &hu, &hv) >= 0 && t < occlude) return 0.0f;
    return 1.0f;
}

/* ---- scene construction: SceneNode::compile (core/nodes/scene.cpp:43-95) --------------------------------- */
static v3 ld3(const float *p) { return mk3(p[0], p[1], p[2]); }

void akor_scene_destroy(oracle_scene *s) {
    if (!s) return;
    free(s->tris); free(s->geom_offset); free(s->mats); free(s->lights); free(s->func); free(s->cdf);
    for (uint32_t i = 0; i < s->n_tex; i++) free((void *)s->tex[i].rgba);
    free(s->tex);
    free(s->nodes); free(s->order); free(s);
}

oracle_scene *akor_scene_create(const akb_scene_desc *d, int brute_force) {
    if (!d || d->abi_version != AKB_ABI_VERSION) return NULL;
    oracle_scene *s = (oracle_scene *)calloc(1, sizeof *s);
    s->brute_force = brute_force;
    s->n_mats = d->n_materials;
    s->mats = (akb_material *)malloc(sizeof(akb_material) * (d->n_materials + 1));
    memcpy(s->mats, d->materials, sizeof(akb_material) * d->n_materials);
    s->n_tex = d->n_textures;
    s->tex = (akb_texture *)calloc(d->n_textures + 1, sizeof(akb_texture));
    float *tex_integral = (float *)calloc(d->n_textures + 1, sizeof(float));
    for (uint32_t i = 0; i < d->n_textures; i++) {
        size_t n = (size_t)d->textures[i].width * d->textures[i].height;
        float *px = (float *)malloc(16 * n);
        memcpy(px, d->textures[i].rgba, 16 * n);
        s->tex[i] = d->textures[i]; s->tex[i].rgba = px;
        float I = 0; /* ImageTexture::integral, kernel/texture.h:50-56 */
        for (size_t q = 0; q < n; q++) I += dot3(mk3(px[4 * q], px[4 * q + 1], px[4 * q + 2]), mk3((float)0.2126, (float)0.7152, (float)0.0722));
        tex_integral[i] = I / (float)(d->textures[i].width * d->textures[i].height);
    }
    s->n_geoms = d->n_meshes;
    s->geom_offset = (uint64_t *)malloc(sizeof(uint64_t) * (d->n_meshes + 1));
    uint64_t nt = 0;
    for (uint32_t g = 0; g < d->n_meshes; g++) { s->geom_offset[g] = nt; nt += d->meshes[g].n_triangles; }
    s->geom_offset[d->n_meshes] = nt;
    s->n_tris = nt;
    s->tris = (tri_t *)malloc(sizeof(tri_t) * (size_t)(nt + 1));
    uint32_t n_lights = 0;
    for (uint32_t g = 0; g < d->n_meshes; g++) {
        const akb_mesh *m = &d->meshes[g];
        for (uint64_t p = 0; p < m->n_triangles; p++) {
            tri_t *t = &s->tris[s->geom_offset[g] + p];
            for (int i = 0; i < 3; i++) { /* get_triangle, kernel/instance.h:83-97 */
                int idx = m->indices[3 * p + i];
                t->v[i] = ld3(m->vertices + 3 * (size_t)idx);
                size_t vidx = 3 * p + i;
                t->n[i] = ld3(m->normals + 3 * vidx);
                t->uv[i][0] = m->texcoords[2 * vidx]; t->uv[i][1] = m->texcoords[2 * vidx + 1];
            }
            int32_t mi = m->material_indices[p];
            t->mat = (mi >= 0 && mi < m->n_mesh_materials) ? m->materials[mi] : -1;
            t->geom = (int32_t)g; t->prim = (int32_t)p;
            if (t->mat >= 0 && s->mats[t->mat].type == AKB_MAT_EMISSIVE) n_lights++;
        }
    }
    /* area lights in (mesh, prim) order + power distribution, core/nodes/scene.cpp:58-92 */
    s->n_lights = n_lights;
    s->lights = (light_t *)malloc(sizeof(light_t) * (n_lights + 1));
    s->func = (float *)malloc(sizeof(float) * (n_lights + 1));
    s->cdf = (float *)malloc(sizeof(float) * (n_lights + 2));
    uint32_t k = 0;
    for (uint64_t i = 0; i < nt; i++) {
        const tri_t *t = &s->tris[i];
        if (t->mat < 0 || s->mats[t->mat].type != AKB_MAT_EMISSIVE) continue;
        light_t *l = &s->lights[k];
        l->v[0] = t->v[0]; l->v[1] = t->v[1]; l->v[2] = t->v[2];
        const float *c = s->mats[t->mat].color;
        l->color = mk3(c[0], c[1], c[2]);
        l->tex = s->mats[t->mat].color_tex;
        memcpy(l->uv, t->uv, sizeof l->uv);
        /* ConstantTexture::integral = luminance (kernel/texture.h:36, common/color.h:52-56) */
        float I = l->tex >= 0 ? tex_integral[l->tex] : dot3(l->color, mk3((float)0.2126, (float)0.7152, (float)0.0722));
        v3 tc[3];
        for (int j = 0; j < 3; j++) tc[j] = mk3(t->uv[j][0], t->uv[j][1], 0.0f);
        v3 tcc = cross3(sub3(tc[1], tc[0]), sub3(tc[2], tc[0]));
        double tc_area = (double)sqrtf(dot3(tcc, tcc)) * 0.5;
        v3 cr = cross3(sub3(t->v[1], t->v[0]), sub3(t->v[2], t->v[0]));
        float area = sqrtf(dot3(cr, cr));
        s->func[k] = (float)((double)area * tc_area * (double)I);
        k++;
    }
    /* Distribution1D ctor, common/distribution.h:52-67 */
    s->cdf[0] = 0;
    for (uint32_t i = 0; i < n_lights; i++) s->cdf[i + 1] = s->cdf[i] + s->func[i] / (float)n_lights;
    s->func_int = s->cdf[n_lights];
    if (s->func_int == 0) { for (uint32_t i = 1; i < n_lights + 1; i++) s->cdf[i] = (float)i / (float)n_lights; }
    else { for (uint32_t i = 1; i < n_lights + 1; i++) s->cdf[i] /= s->func_int; }
    free(tex_integral);
    camera_init(&s->cam, &d->camera);
    if (!brute_force) build_accel(s);
    return s;
}

uint64_t akor_n_lights(const oracle_scene *s) { return s->n_lights; }

/* ---- threaded drivers ---------------------------------------------------------------------------------------- */
typedef struct job {
    const oracle_scene *s;
    /* trace */
    const float *org, *dir; size_t n; int any_hit; akb_hit *hits; int32_t *occ;
    /* render */
    const akb_render_params *p; const int32_t *tiles; long n_tiles; float *out; uint64_t counts[3];
    long long cpu_ns; /* CPU time summed over the workers (CLOCK_THREAD_CPUTIME_ID) */
    volatile long next; pthread_mutex_t mu;
} job_t;

static long long thread_cpu_ns(void) {
    struct timespec ts; clock_gettime(CLOCK_THREAD_CPUTIME_ID, &ts);
    return (long long)ts.tv_sec * 1000000000ll + ts.tv_nsec;
}

static long job_fetch(job_t *j, long step) { return __sync_fetch_and_add(&j->next, step); }

static void *trace_worker(void *arg) {
    job_t *j = (job_t *)arg;
    for (;;) {
        long b = job_fetch(j, 4096);
        if ((size_t)b >= j->n) break;
        size_t e = (size_t)b + 4096; if (e > j->n) e = j->n;
        for (size_t i = (size_t)b; i < e; i++) {
            ray_t r; r.o = ld3(j->org + 4 * i); r.tmin = j->org[4 * i + 3];
            r.d = ld3(j->dir + 4 * i); r.tmax = j->dir[4 * i + 3];
            if (j->any_hit) j->occ[i] = scene_occlude(j->s, &r);
            else {
                akb_hit h; int64_t ti = scene_intersect(j->s, &r, &h.t, &h.u, &h.v);
                if (ti >= 0) { h.prim_id = j->s->tris[ti].prim; h.geom_id = j->s->tris[ti].geom; }
                else { h.t = INFINITY; h.u = h.v = 0; h.prim_id = h.geom_id = -1; }
                j->hits[i] = h;
            }
        }
    }
    return NULL;
}

/* cpu::PathTracer::render's tile + per-pixel loop (kernel/integrators/cpu/integrator.cpp:115-135) + Tile::add_sample /
 * Film::merge_tile (core/film.h:66-95). Work items are 16x16-pixel tiles (the reference's tile_size) drained by all
 * workers through one counter (core/parallel.cpp:119-129). ray_clamp > 0 applies the GPU integrator's per-sample clamp
 * (gpu/cuda/integrator.cpp:196-198: clamp_zero then min(L, ray_clamp)). */
static void *render_worker(void *arg) {
    job_t *j = (job_t *)arg;
    const akb_render_params *p = j->p;
    const int W = j->s->cam.w;
    uint64_t c0 = 0, c1 = 0, c2 = 0;
    const long long t_start = thread_cpu_ns();
    for (;;) {
        const long k = job_fetch(j, 1);
        if (k >= j->n_tiles) break;
        const int32_t *tj = j->tiles + 4 * k;
        for (int y = tj[1]; y < tj[3]; y++)
        for (int x = tj[0]; x < tj[2]; x++) {
            sampler_t smp; memset(&smp, 0, sizeof smp); smp.kind = p->sampler;
            /* AKB_SPLIT_SAMPLE (include/akari_b200.h): this rank's share of the samples, per-rank seed offset */
            const int world = p->world_size > 0 ? p->world_size : 1;
            const int sample_split = p->split == AKB_SPLIT_SAMPLE;
            const int spp = sample_split ? p->spp / world + (p->rank < p->spp % world ? 1 : 0) : p->spp;
            const uint64_t seed_off = sample_split ? (uint64_t)p->rank * (uint64_t)W * (uint64_t)j->s->cam.h : 0;
            sampler_set_index(&smp, (uint64_t)(x + (int)y * W) + seed_off);
            v3 sum = {0, 0, 0}; float weight = 0;
            for (int sidx = 0; sidx < spp; sidx++) {
                if (p->integrator == AKB_INTEGRATOR_AO) { /* cpu/integrator.cpp:72-80 */
                    float g = run_ao(j->s, &smp, x, (int)y, p->ao_occlude, &c1);
                    weight += 1.0f; sum = add3(sum, mk3(g, g, g));
                    c0++;
                    continue;
                }
                path_t pt; memset(&pt, 0, sizeof pt);
                pt.beta = mk3(1, 1, 1); pt.max_depth = p->max_depth; pt.smp = smp;
                run_path(j->s, &pt, x, (int)y, p->rr_start_depth);
                smp = pt.smp;
                v3 L = pt.L;
                if (p->ray_clamp > 0.0f) {
                    float *c = (float *)&L;
                    for (int k = 0; k < 3; k++) { float v = c[k]; v = (v != v) ? 0.0f : maxf_(0.0f, v); c[k] = minf_(v, p->ray_clamp); }
                }
                weight += 1.0f; sum = add3(sum, L);
                c0++; c1 += pt.n_closest; c2 += pt.n_shadow;
            }
            float *o = j->out + 4 * ((size_t)y * W + x);
            o[0] = sum.x; o[1] = sum.y; o[2] = sum.z; o[3] = weight;
        }
    }
    pthread_mutex_lock(&j->mu);
    j->counts[0] += c0; j->counts[1] += c1; j->counts[2] += c2;
    j->cpu_ns += thread_cpu_ns() - t_start;
    pthread_mutex_unlock(&j->mu);
    return NULL;
}

static void run_threads(void *(*fn)(void *), job_t *j, int n_threads) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256];
    pthread_mutex_init(&j->mu, NULL);
    for (int i = 1; i < n_threads; i++) pthread_create(&th[i], NULL, fn, j);
    fn(j);
    for (int i = 1; i < n_threads; i++) pthread_join(th[i], NULL);
    pthread_mutex_destroy(&j->mu);
}

int akor_trace(const oracle_scene *s, const float *org_tmin, const float *dir_tmax, size_t n, int any_hit,
               akb_hit *out_hits, int32_t *out_occ, int n_threads) {
    job_t j; memset(&j, 0, sizeof j);
    j.s = s; j.org = org_tmin; j.dir = dir_tmax; j.n = n; j.any_hit = any_hit; j.hits = out_hits; j.occ = out_occ;
    run_threads(trace_worker, &j, n_threads);
    return 0;
}

/* Renders the pixel rectangles rects[4*r..] = (x0, y0, x1, y1). counts (may be NULL): [0] samples, [1] closest-hit
 * rays, [2] shadow rays. busy (may be NULL): [0] CPU seconds summed over the workers, [1] wall seconds. */
int akor_render_rects(const oracle_scene *s, const akb_render_params *p, const int32_t *rects, int n_rects, float *out_rgbw,
                      int n_threads, uint64_t *counts, double *busy) {
    if (p->integrator != AKB_INTEGRATOR_PATH && p->integrator != AKB_INTEGRATOR_AO) return 1;
    const int T = 16;
    long n_tiles = 0;
    for (int r = 0; r < n_rects; r++) {
        const long w = rects[4 * r + 2] - rects[4 * r], h = rects[4 * r + 3] - rects[4 * r + 1];
        if (w > 0 && h > 0) n_tiles += ((w + T - 1) / T) * ((h + T - 1) / T);
    }
    int32_t *tiles = (int32_t *)malloc(sizeof(int32_t) * 4 * (size_t)(n_tiles > 0 ? n_tiles : 1));
    long k = 0;
    for (int r = 0; r < n_rects; r++)
        for (int ty = rects[4 * r + 1]; ty < rects[4 * r + 3]; ty += T)
            for (int tx = rects[4 * r]; tx < rects[4 * r + 2]; tx += T) {
                tiles[4 * k] = tx; tiles[4 * k + 1] = ty;
                tiles[4 * k + 2] = tx + T < rects[4 * r + 2] ? tx + T : rects[4 * r + 2];
                tiles[4 * k + 3] = ty + T < rects[4 * r + 3] ? ty + T : rects[4 * r + 3];
                k++;
            }
    job_t j; memset(&j, 0, sizeof j);
    j.s = s; j.p = p; j.tiles = tiles; j.n_tiles = n_tiles; j.out = out_rgbw; j.next = 0;
    struct timespec w0, w1; clock_gettime(CLOCK_MONOTONIC, &w0);
    run_threads(render_worker, &j, n_threads);
    clock_gettime(CLOCK_MONOTONIC, &w1);
    free(tiles);
    if (counts) { counts[0] = j.counts[0]; counts[1] = j.counts[1]; counts[2] = j.counts[2]; }
    if (busy) { busy[0] = (double)j.cpu_ns * 1e-9; busy[1] = (double)(w1.tv_sec - w0.tv_sec) + 1e-9 * (double)(w1.tv_nsec - w0.tv_nsec); }
    return 0;
}

int akor_render(const oracle_scene *s, const akb_render_params *p, int x0, int y0, int x1, int y1, float *out_rgbw,
                int n_threads, uint64_t *counts) {
    const int32_t rect[4] = {x0, y0, x1, y1};
    return akor_render_rects(s, p, rect, 1, out_rgbw, n_threads, counts, NULL);
}

/* Diagnosis aid: every ray the Path integrator traces for pixel (x, y) over its p->spp samples, in order
 * (9 floats each, see log_ray). Returns the number of rays (at most cap are stored). */
int akor_pixel_rays(const oracle_scene *s, const akb_render_params *p, int x, int y, float *out, int cap) {
    sampler_t smp; memset(&smp, 0, sizeof smp); smp.kind = p->sampler;
    sampler_set_index(&smp, (uint64_t)(x + y * s->cam.w));
    g_raylog = out; g_raylog_n = 0; g_raylog_cap = cap;
    for (int k = 0; k < p->spp; k++) {
        path_t pt; memset(&pt, 0, sizeof pt);
        pt.beta = mk3(1, 1, 1); pt.max_depth = p->max_depth; pt.smp = smp;
        run_path(s, &pt, x, y, p->rr_start_depth);
        smp = pt.smp;
    }
    g_raylog = NULL;
    return g_raylog_n;
}

int akor_generate_rays(const oracle_scene *s, const int32_t *xy, const float *u2, size_t n, float *org_tmin,
                       float *dir_tmax) {
    for (size_t i = 0; i < n; i++) {
        ray_t r = camera_generate_ray(&s->cam, u2[2 * i], u2[2 * i + 1], xy[2 * i], xy[2 * i + 1]);
        org_tmin[4 * i] = r.o.x; org_tmin[4 * i + 1] = r.o.y; org_tmin[4 * i + 2] = r.o.z; org_tmin[4 * i + 3] = r.tmin;
        dir_tmax[4 * i] = r.d.x; dir_tmax[4 * i + 1] = r.d.y; dir_tmax[4 * i + 2] = r.d.z; dir_tmax[4 * i + 3] = r.tmax;
    }
    return 0;
}

int akor_sampler_kat(int kind, uint64_t seed, int n, float *out, float *out_next2d) {
    sampler_t s; memset(&s, 0, sizeof s); s.kind = kind;
    sampler_set_index(&s, seed);
    sampler_t s2 = s;
    for (int i = 0; i < n; i++) out[i] = next1d(&s);
    next2d(&s2, &out_next2d[0], &out_next2d[1]);
    return 0;
}

/* Film::write_image post-process (core/film.h:97-113): radiance/weight; GammaCorrection -> linear_to_srgb in fp32
 * (common/color.h:57-61: L*12.92 or 1.055*pow(L,1/2.4)-0.055, scalars converted to float by Array's operators);
 * DefaultImageWriter (core/image.cpp:48-53): clamp to [0,1], (int)std::round(x * 255.5) in double, clamp to [0,255]. */
void akor_film_to_srgb8(const float *rgbw, int w, int h, uint8_t *out) {
    for (size_t i = 0; i < (size_t)w * h; i++) {
        float wt = rgbw[4 * i + 3];
        for (int c = 0; c < 3; c++) {
            float L = rgbw[4 * i + c];
            if (wt != 0) L = L / wt;
            float v = (L < 0.0031308f) ? (L * 12.92f) : (1.055f * powf(L, 1.0f / 2.4f) - 0.055f);
            v = minf_(maxf_(v, 0.0f), 1.0f);
            int q = (int)round((double)v * 255.5);
            out[3 * i + c] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
        }
    }
}

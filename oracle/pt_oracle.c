/* oracle/pt_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.  See pt_oracle.h.
 *
 * Plain-C restatement of the reference hot path.  Every function cites the reference lines it
 * follows (paths relative to /root/reference).  Arithmetic is written operation-by-operation in the
 * order the reference's C++ evaluates it (including its float->double promotions, e.g. `T k = 1.0/t`
 * in core/vector.h:93-94 and `2.0f * M_PI * v` in rendering/integrator.h:67) so that, compiled with
 * -ffp-contract=off on the same host, it reproduces oracle/_ref/ref_host bit for bit.
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC pt_oracle.c -lm   (oracle/Makefile)
 */
#define _GNU_SOURCE   /* sincosf */
#include "pt_oracle.h"
#include <ctype.h>
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846 /* double, as glibc defines it (sensor.h:7-9 never triggers on Linux) */
#endif

/* ============================================================================================ */
/* core/vector.h                                                                                 */
/* ============================================================================================ */
typedef struct { float e[3]; } v3;

static inline v3 V(float x, float y, float z) { v3 r = {{x, y, z}}; return r; }
static inline v3 vadd(v3 a, v3 b) { return V(a.e[0] + b.e[0], a.e[1] + b.e[1], a.e[2] + b.e[2]); }   /* vector.h:144-149 */
static inline v3 vsub(v3 a, v3 b) { return V(a.e[0] - b.e[0], a.e[1] - b.e[1], a.e[2] - b.e[2]); }   /* :151-156 */
static inline v3 vmul(v3 a, v3 b) { return V(a.e[0] * b.e[0], a.e[1] * b.e[1], a.e[2] * b.e[2]); }   /* :158-163 */
static inline v3 vdivv(v3 a, v3 b) { return V(a.e[0] / b.e[0], a.e[1] / b.e[1], a.e[2] / b.e[2]); }  /* :165-170 */
static inline v3 vscale(float t, v3 a) { return V(t * a.e[0], t * a.e[1], t * a.e[2]); }              /* :173-183 */
static inline v3 vneg(v3 a) { return V(-a.e[0], -a.e[1], -a.e[2]); }                                  /* :56-60 */
/* operator/(Vector, T): `T k = 1.0 / t` is a DOUBLE division rounded to float, then multiplies (:185-191) */
static inline v3 vdiv(v3 a, float t) { float k = (float)(1.0 / (double)t); return V(a.e[0] * k, a.e[1] * k, a.e[2] * k); }
static inline float vdot(v3 a, v3 b) { float s = 0; s += a.e[0] * b.e[0]; s += a.e[1] * b.e[1]; s += a.e[2] * b.e[2]; return s; } /* :194-199 */
static inline float vlen2(v3 a) { float s = 0; s += a.e[0] * a.e[0]; s += a.e[1] * a.e[1]; s += a.e[2] * a.e[2]; return s; }     /* :99-103 */
static inline float vlen(v3 a) { return sqrtf(vlen2(a)); }                                                                      /* :105-107 */
static inline v3 vunit(v3 a) { return vdiv(a, vlen(a)); }                                                                       /* :201-204 */
static inline v3 vcross(v3 a, v3 b) {                                                                                           /* :207-214 */
    return V(a.e[1] * b.e[2] - a.e[2] * b.e[1], (-(a.e[0] * b.e[2] - a.e[2] * b.e[0])), a.e[0] * b.e[1] - a.e[1] * b.e[0]);
}
static inline v3 vload(const float* p) { return V(p[0], p[1], p[2]); }
static inline void vstore(float* p, v3 a) { p[0] = a.e[0]; p[1] = a.e[1]; p[2] = a.e[2]; }

/* ============================================================================================ */
/* cuRAND XORWOW (CUDA toolkit 12.9, curand_kernel.h: curand() :863-874, _curand_init_scratch      */
/* :758-783, _skipahead_sequence_scratch :645-699; curand_uniform.h:69-72).  The toolkit ships     */
/* precomputed jump matrices; this oracle derives them itself by repeated squaring of the one-step */
/* GF(2) transition matrix -- the result M^(2^67 * 4^k) is unique, and is pinned by the ref_host    */
/* KATs in tests/golden/xorwow_kat.json.                                                           */
/* ============================================================================================ */
#define XW_WORDS 5
#define XW_ROWS 160
#define XW_NMAT 32
static uint32_t g_jump[XW_NMAT][XW_ROWS * XW_WORDS];
static int g_jump_ready = 0;

static void xw_step_linear(uint32_t v[5]) {
    uint32_t t = (v[0] ^ (v[0] >> 2));
    v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
    v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
}
/* row-vector times matrix, the convention of __curand_matvec (curand_kernel.h:336-352) */
static void xw_matvec(const uint32_t* vec, const uint32_t* m, uint32_t* out) {
    uint32_t r[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < 5; i++)
        for (int j = 0; j < 32; j++)
            if (vec[i] & (1u << j)) {
                const uint32_t* row = m + 5 * (i * 32 + j);
                r[0] ^= row[0]; r[1] ^= row[1]; r[2] ^= row[2]; r[3] ^= row[3]; r[4] ^= row[4];
            }
    memcpy(out, r, sizeof r);
}
static void xw_square(uint32_t* m) {
    static uint32_t tmp[XW_ROWS * XW_WORDS];
    for (int r = 0; r < XW_ROWS; r++) xw_matvec(m + 5 * r, m, tmp + 5 * r);
    memcpy(m, tmp, sizeof tmp);
}
static void xw_prepare(void) {
    #pragma omp critical(pto_xw_prepare)
    {
        if (!g_jump_ready) {
            uint32_t* m = g_jump[0];
            for (int r = 0; r < XW_ROWS; r++) {
                uint32_t e[5] = {0, 0, 0, 0, 0};
                e[r / 32] = 1u << (r % 32);
                xw_step_linear(e);
                memcpy(m + 5 * r, e, sizeof e);
            }
            for (int s = 0; s < 67; s++) xw_square(m);               /* one subsequence = 2^67 draws */
            for (int k = 1; k < XW_NMAT; k++) {                       /* PRECALC_BLOCK_SIZE = 2 -> powers of 4 */
                memcpy(g_jump[k], g_jump[k - 1], sizeof g_jump[0]);
                xw_square(g_jump[k]); xw_square(g_jump[k]);
            }
            g_jump_ready = 1;
        }
    }
}
void pto_xorwow_jump_tables(uint32_t* out) { xw_prepare(); memcpy(out, g_jump, sizeof g_jump); }

void pto_xorwow_init(uint64_t seed, uint64_t subsequence, pto_xorwow* st) {
    if (!g_jump_ready) xw_prepare();
    uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
    uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
    uint32_t t0 = 1099087573u * s0;
    uint32_t t1 = 2591861531u * s1;
    st->d = 6615241u + t1 + t0;
    st->v[0] = 123456789u + t0;
    st->v[1] = 362436069u ^ t0;
    st->v[2] = 521288629u + t1;
    st->v[3] = 88675123u ^ t1;
    st->v[4] = 5783321u + t0;
    uint64_t p = subsequence;
    for (int k = 0; p && k < XW_NMAT; k++, p >>= 2)
        for (unsigned t = 0; t < (unsigned)(p & 3); t++) xw_matvec(st->v, g_jump[k], st->v);
    /* d unchanged: 2^67 * n * 362437 == 0 mod 2^32 (curand_kernel.h:698); offset is always 0 (integrator.h:279) */
}
uint32_t pto_xorwow_next(pto_xorwow* st) {
    xw_step_linear(st->v);
    st->d += 362437u;
    return st->v[4] + st->d;
}
float pto_uniform(pto_xorwow* st) {
    uint32_t x = pto_xorwow_next(st);
    return x * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}

/* ============================================================================================ */
/* geometry: triangle.h:64-96, quad.h:49-132, primitive.h:83-98                                    */
/* ============================================================================================ */
typedef struct { v3 o, d; } ray_t;
static inline ray_t make_ray(v3 o, v3 dir) { ray_t r; r.o = o; r.d = vunit(dir); return r; }   /* ray.h:9-12 */

typedef struct { float t, u, v; v3 p; } hit_t;

/* one Moeller-Trumbore test; `strict_upper` selects `t <= t_max` (Triangle) or `t < closest` (Quad halves) */
static inline int mt_test(v3 v0, v3 v1, v3 v2, const ray_t* r, float t_min, float t_max, int strict_upper, hit_t* h) {
    const float EPSILON = 1e-8f;
    v3 edge1 = vsub(v1, v0), edge2 = vsub(v2, v0);
    v3 hh = vcross(r->d, edge2);
    float a = vdot(edge1, hh);
    if (strict_upper ? !(fabsf(a) > EPSILON) : (fabsf(a) < EPSILON)) return 0;   /* quad.h:66 vs triangle.h:71 */
    float f = 1.0f / a;
    v3 s = vsub(r->o, v0);
    float u = f * vdot(s, hh);
    if (strict_upper ? !(u >= 0.0f && u <= 1.0f) : (u < 0.0f || u > 1.0f)) return 0;
    v3 q = vcross(s, edge1);
    float v = f * vdot(r->d, q);
    if (strict_upper ? !(v >= 0.0f && u + v <= 1.0f) : (v < 0.0f || u + v > 1.0f)) return 0;
    float t = f * vdot(edge2, q);
    if (t > EPSILON && t >= t_min && (strict_upper ? (t < t_max) : (t <= t_max))) { h->t = t; h->u = u; h->v = v; return 1; }
    return 0;
}
static inline int prim_intersect(const pto_prim* p, const ray_t* r, float t_min, float t_max, hit_t* h) {
    if (p->type == 0) {
        if (!mt_test(vload(p->v[0]), vload(p->v[1]), vload(p->v[2]), r, t_min, t_max, 0, h)) return 0;
        h->p = vadd(r->o, vscale(h->t, r->d));                                   /* triangle.h:90 */
        return 1;
    }
    int hit = 0; float closest = t_max; hit_t tmp;
    if (mt_test(vload(p->v[0]), vload(p->v[1]), vload(p->v[2]), r, t_min, closest, 1, &tmp)) { closest = tmp.t; *h = tmp; hit = 1; }
    if (mt_test(vload(p->v[0]), vload(p->v[2]), vload(p->v[3]), r, t_min, closest, 1, &tmp)) { closest = tmp.t; *h = tmp; hit = 1; }
    if (hit) h->p = vadd(r->o, vscale(closest, r->d));                           /* quad.h:124 */
    return hit;
}
static inline v3 prim_centroid(const pto_prim* p) {                              /* primitive.h:92-98 */
    if (p->type == 0) return vdiv(vadd(vadd(vload(p->v[0]), vload(p->v[1])), vload(p->v[2])), 3.0f);
    return vscale(0.25f, vadd(vadd(vadd(vload(p->v[0]), vload(p->v[1])), vload(p->v[2])), vload(p->v[3])));
}

/* ============================================================================================ */
/* BVHBuilder, rendering/bvh.h:76-219 (host recursive, longest-centroid-axis midpoint, leaf <= 4)    */
/* ============================================================================================ */
static void bounds_of(const pto_scene* s, int start, int end, float bmin[3], float bmax[3]) {       /* bvh.h:110-150 */
    const float eps = 1e-6f;
    for (int a = 0; a < 3; a++) { bmin[a] = 1e30f; bmax[a] = -1e30f; }
    for (int i = start; i < end; i++) {
        const pto_prim* p = &s->prims[s->idx[i]];
        for (int a = 0; a < 3; a++) {
            float lo, hi;
            if (p->type == 0) {
                lo = fminf(fminf(p->v[0][a], p->v[1][a]), p->v[2][a]);
                hi = fmaxf(fmaxf(p->v[0][a], p->v[1][a]), p->v[2][a]);
            } else {
                lo = fminf(fminf(p->v[0][a], p->v[1][a]), fminf(p->v[2][a], p->v[3][a]));
                hi = fmaxf(fmaxf(p->v[0][a], p->v[1][a]), fmaxf(p->v[2][a], p->v[3][a]));
            }
            bmin[a] = fminf(bmin[a], lo - eps);
            bmax[a] = fmaxf(bmax[a], hi + eps);
        }
    }
}
static int build_rec(pto_scene* s, int start, int end) {                                            /* bvh.h:156-218 */
    if (s->n_nodes == s->cap_nodes) {
        s->cap_nodes = s->cap_nodes ? s->cap_nodes * 2 : 64;
        s->nodes = (pto_node*)realloc(s->nodes, sizeof(pto_node) * (size_t)s->cap_nodes);
    }
    int me = s->n_nodes++;
    float bmin[3], bmax[3];
    bounds_of(s, start, end, bmin, bmax);
    int count = end - start;
    pto_node nd; memcpy(nd.bmin, bmin, 12); memcpy(nd.bmax, bmax, 12); nd.left = -1; nd.right = -1; nd.count = 0;
    if (count <= 4) { nd.left = start; nd.count = count; s->nodes[me] = nd; return me; }
    float cmin[3] = {1e30f, 1e30f, 1e30f}, cmax[3] = {-1e30f, -1e30f, -1e30f};
    for (int i = start; i < end; i++) {
        v3 c = prim_centroid(&s->prims[s->idx[i]]);
        for (int a = 0; a < 3; a++) { cmin[a] = fminf(cmin[a], c.e[a]); cmax[a] = fmaxf(cmax[a], c.e[a]); }
    }
    float ext[3] = {cmax[0] - cmin[0], cmax[1] - cmin[1], cmax[2] - cmin[2]};
    int axis = 0;
    if (ext[1] > ext[0]) axis = 1;
    if (ext[2] > ext[axis]) axis = 2;
    if (ext[axis] < 1e-6f) { nd.left = start; nd.count = count; s->nodes[me] = nd; return me; }
    float split = (cmin[axis] + cmax[axis]) * 0.5f;
    int mid = start;
    for (int i = start; i < end; i++)
        if (prim_centroid(&s->prims[s->idx[i]]).e[axis] < split) { int t = s->idx[i]; s->idx[i] = s->idx[mid]; s->idx[mid] = t; mid++; }
    if (mid == start || mid == end) mid = start + count / 2;
    int l = build_rec(s, start, mid);
    int r = build_rec(s, mid, end);
    nd.left = l; nd.right = r; nd.count = 0;
    s->nodes[me] = nd;
    return me;
}

pto_scene* pto_scene_new(const pto_prim* prims, int n) {
    pto_scene* s = (pto_scene*)calloc(1, sizeof(pto_scene));
    s->prims = (pto_prim*)malloc(sizeof(pto_prim) * (size_t)(n > 0 ? n : 1));
    memcpy(s->prims, prims, sizeof(pto_prim) * (size_t)n);
    s->n_prims = n;
    s->idx = (int*)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) s->idx[i] = i;
    if (n > 0) build_rec(s, 0, n);
    return s;
}
void pto_scene_free(pto_scene* s) { if (!s) return; free(s->prims); free(s->nodes); free(s->idx); free(s->guide); free(s->mirror); free(s); }
int pto_scene_num_prims(const pto_scene* s) { return s->n_prims; }
int pto_scene_num_nodes(const pto_scene* s) { return s->n_nodes; }
const pto_prim* pto_scene_prims(const pto_scene* s) { return s->prims; }
const pto_node* pto_scene_nodes(const pto_scene* s) { return s->nodes; }
const int* pto_scene_indices(const pto_scene* s) { return s->idx; }

/* ============================================================================================ */
/* loadMTL / loadOBJ, utils/file_manager.h:39-79, 93-273 (iostream token semantics restated)         */
/* ============================================================================================ */
typedef struct { char name[128]; float bsdf[3], Le[3]; } mtl_t;

static const char* next_token(const char* p, char* tok, size_t cap) {   /* `iss >> std::string` */
    while (*p && isspace((unsigned char)*p)) p++;
    if (!*p) return NULL;
    size_t n = 0;
    while (*p && !isspace((unsigned char)*p)) { if (n + 1 < cap) tok[n++] = *p; p++; }
    tok[n] = 0;
    return p;
}
static int parse_floats(const char* p, float* out, int n) {             /* `iss >> x >> y >> z` */
    for (int i = 0; i < n; i++) {
        char* e; float f = strtof(p, &e);
        if (e == p) return 0;
        out[i] = f; p = e;
    }
    return 1;
}
static int load_mtl(const char* path, mtl_t** out) {
    FILE* f = fopen(path, "r");
    *out = NULL;
    if (!f) return 0;
    int n = 0, cap = 0, have = 0; mtl_t cur; memset(&cur, 0, sizeof cur);
    mtl_t* v = NULL;
    char* line = NULL; size_t lcap = 0;
    #define MTL_PUSH() do { int found = -1; for (int i_ = 0; i_ < n; i_++) if (!strcmp(v[i_].name, cur.name)) found = i_; \
        if (found >= 0) v[found] = cur; else { if (n == cap) { cap = cap ? cap * 2 : 8; v = (mtl_t*)realloc(v, sizeof(mtl_t) * cap); } v[n++] = cur; } } while (0)
    while (getline(&line, &lcap, f) >= 0) {
        char tok[128]; const char* p = next_token(line, tok, sizeof tok);
        if (!p) continue;
        if (!strcmp(tok, "newmtl")) {
            if (have && cur.name[0]) MTL_PUSH();
            memset(&cur, 0, sizeof cur);
            cur.bsdf[0] = cur.bsdf[1] = cur.bsdf[2] = 0.8f;              /* Material defaults, file_manager.h:24-27 */
            char nm[128]; if (next_token(p, nm, sizeof nm)) strcpy(cur.name, nm);
            have = 1;
        } else if (!strcmp(tok, "Kd")) { float c[3]; if (parse_floats(p, c, 3)) memcpy(cur.bsdf, c, 12); }
        else if (!strcmp(tok, "Ke")) { float c[3]; if (parse_floats(p, c, 3)) memcpy(cur.Le, c, 12); }
    }
    if (have && cur.name[0]) MTL_PUSH();
    free(line); fclose(f);
    *out = v;
    return n;
}
/* `std::stringstream(token) >> size_t` : optional sign, digits; a leading '-' wraps like strtoull */
static int parse_index(const char** pp, unsigned long long* out) {
    const char* p = *pp; int neg = 0;
    if (*p == '+' || *p == '-') { neg = (*p == '-'); p++; }
    if (!isdigit((unsigned char)*p)) return 0;
    unsigned long long v = 0;
    while (isdigit((unsigned char)*p)) { v = v * 10 + (unsigned)(*p - '0'); p++; }
    *out = neg ? (0ull - v) : v; *pp = p;
    return 1;
}

pto_scene* pto_load_obj(const char* path, int convert_quads) {
    FILE* f = fopen(path, "r");
    if (!f) return NULL;
    char base[1024] = "";
    const char* sl = strrchr(path, '/'); const char* bs = strrchr(path, '\\');
    if (bs && (!sl || bs > sl)) sl = bs;
    if (sl) { size_t n = (size_t)(sl - path) + 1; if (n >= sizeof base) n = sizeof base - 1; memcpy(base, path, n); base[n] = 0; }
    float* verts = NULL; size_t nv = 0, capv = 0;
    float* norms = NULL; size_t nn = 0, capn = 0;
    pto_prim* prims = NULL; size_t np = 0, capp = 0;
    mtl_t* mtls = NULL; int nm = 0;
    mtl_t cur; memset(&cur, 0, sizeof cur); cur.bsdf[0] = cur.bsdf[1] = cur.bsdf[2] = 0.8f;
    char* line = NULL; size_t lcap = 0;
    while (getline(&line, &lcap, f) >= 0) {
        size_t L = strlen(line);
        while (L && (line[L - 1] == '\n')) line[--L] = 0;
        if (L == 0 || line[0] == '#' || line[0] == 'o' || line[0] == 's') continue;           /* file_manager.h:120 */
        char tok[256]; const char* p = next_token(line, tok, sizeof tok);
        if (!p) continue;
        if (!strcmp(tok, "v")) {
            float c[3]; if (!parse_floats(p, c, 3)) continue;
            if (nv == capv) { capv = capv ? capv * 2 : 64; verts = (float*)realloc(verts, capv * 12); }
            memcpy(verts + 3 * nv++, c, 12);
        } else if (!strcmp(tok, "vn")) {
            float c[3]; if (!parse_floats(p, c, 3)) continue;
            v3 u = vunit(V(c[0], c[1], c[2]));                                                 /* :141 */
            if (nn == capn) { capn = capn ? capn * 2 : 64; norms = (float*)realloc(norms, capn * 12); }
            vstore(norms + 3 * nn++, u);
        } else if (!strcmp(tok, "mtllib")) {
            char nm_[512]; if (!next_token(p, nm_, sizeof nm_)) nm_[0] = 0;
            char full[2048]; snprintf(full, sizeof full, "%s%s", base, nm_);
            free(mtls); nm = load_mtl(full, &mtls);
        } else if (!strcmp(tok, "usemtl")) {
            char nm_[128]; if (!next_token(p, nm_, sizeof nm_)) nm_[0] = 0;
            int found = -1; for (int i = 0; i < nm; i++) if (!strcmp(mtls[i].name, nm_)) found = i;
            if (found >= 0) cur = mtls[found];
            else { memset(&cur, 0, sizeof cur); cur.bsdf[0] = cur.bsdf[1] = cur.bsdf[2] = 0.8f; }
        } else if (!strcmp(tok, "f")) {
            unsigned long long vi[64], ni[64]; int cnt = 0;
            char t[256];
            while ((p = next_token(p, t, sizeof t)) != NULL) {
                const char* q = t; unsigned long long v = 0, vt = 0, vn = 0;
                if (!parse_index(&q, &v)) continue;                                             /* :166-170 */
                if (*q == '/') {
                    q++;
                    if (*q == '/') { q++; if (!parse_index(&q, &vn)) vn = 0; }
                    else if (parse_index(&q, &vt)) { if (*q == '/') { q++; if (!parse_index(&q, &vn)) vn = 0; } }
                }
                if (cnt < 64) { vi[cnt] = v; ni[cnt] = vn; cnt++; }
            }
            if (cnt != 3 && cnt != 4) continue;
            int bad = 0; for (int k = 0; k < cnt; k++) if (vi[k] == 0 || vi[k] > nv) bad = 1;
            if (bad) continue;
            pto_prim pr; memset(&pr, 0, sizeof pr);
            for (int k = 0; k < cnt; k++) memcpy(pr.v[k], verts + 3 * (vi[k] - 1), 12);
            int has_n = (ni[0] != 0 && ni[0] <= nn);
            if (cnt == 3) {
                pr.type = 0;
                v3 n = has_n ? vload(norms + 3 * (ni[0] - 1))
                             : vunit(vcross(vsub(vload(pr.v[1]), vload(pr.v[0])), vsub(vload(pr.v[2]), vload(pr.v[0]))));   /* :206-210 */
                vstore(pr.n, n);
            } else {
                pr.type = 1;
                v3 n = has_n ? vload(norms + 3 * (ni[0] - 1))
                             : vunit(vcross(vsub(vload(pr.v[1]), vload(pr.v[0])), vsub(vload(pr.v[3]), vload(pr.v[0]))));   /* quad.h:27-29 */
                vstore(pr.n, n);
            }
            memcpy(pr.bsdf, cur.bsdf, 12); memcpy(pr.Le, cur.Le, 12);
            if (np == capp) { capp = capp ? capp * 2 : 64; prims = (pto_prim*)realloc(prims, capp * sizeof(pto_prim)); }
            prims[np++] = pr;
        }
    }
    free(line); fclose(f); free(verts); free(norms); free(mtls);
    if (np == 0) { free(prims); return NULL; }
    if (convert_quads) {                                     /* convertQuadsToTriangles, application_state.h:323-365 */
        pto_prim* out = (pto_prim*)malloc(sizeof(pto_prim) * np * 2); size_t m = 0;
        for (size_t i = 0; i < np; i++) {
            if (prims[i].type == 0) { out[m++] = prims[i]; continue; }
            for (int h = 0; h < 2; h++) {
                pto_prim t = prims[i]; t.type = 0;
                memcpy(t.v[1], prims[i].v[1 + h], 12); memcpy(t.v[2], prims[i].v[2 + h], 12); memset(t.v[3], 0, 12);
                vstore(t.n, vunit(vcross(vsub(vload(t.v[1]), vload(t.v[0])), vsub(vload(t.v[2]), vload(t.v[0])))));  /* triangle.h:27-29 */
                out[m++] = t;
            }
        }
        free(prims); prims = out; np = m;
    }
    pto_scene* s = pto_scene_new(prims, (int)np);
    free(prims);
    return s;
}

/* ============================================================================================ */
/* Sensor, rendering/sensor.h:16-67; RenderState, application_state.h:88, 106-109; application.h:162 */
/* ============================================================================================ */
void pto_camera_lookat(const float eye[3], const float at[3], const float up[3], float vfov, int w, int h,
                       int orbit, pto_camera* cam) {
    v3 origin = vload(eye), lookat = vload(at), vup = vload(up);
    float radius = vlen(vsub(origin, lookat));
    float aspect = (float)w / (float)h;
    if (orbit) {                                                           /* updateCameraOrbit, yaw = 90, pitch = 0 */
        float yaw = 90.0f, pitch = 0.0f;
        float yawRad = (float)(((yaw) * M_PI / 180.0f));
        float pitchRad = (float)(((pitch) * M_PI / 180.0f));
        origin.e[0] = lookat.e[0] + radius * cosf(pitchRad) * cosf(yawRad);
        origin.e[1] = lookat.e[1] + radius * sinf(pitchRad);
        origin.e[2] = lookat.e[2] + radius * cosf(pitchRad) * sinf(yawRad);
    }
    float theta = vfov * M_PI / 180.0f;                                    /* double arithmetic, rounded to float */
    float half_height = tanf(theta / 2.0f);
    float half_width = aspect * half_height;
    v3 ww = vunit(vsub(origin, lookat));
    v3 uu = vunit(vcross(vup, ww));
    v3 vv = vcross(ww, uu);
    v3 llc = vsub(vsub(vsub(origin, vscale(half_width, uu)), vscale(half_height, vv)), ww);
    vstore(cam->origin, origin); vstore(cam->llc, llc);
    vstore(cam->horizontal, vscale(2 * half_width, uu));
    vstore(cam->vertical, vscale(2 * half_height, vv));
    cam->width = w; cam->height = h;
}
void pto_camera_default(int w, int h, pto_camera* cam) {
    const float eye[3] = {0.5f, 3.0f, 8.5f}, at[3] = {0.0f, 2.5f, 0.0f}, up[3] = {0.0f, 1.0f, 0.0f};
    pto_camera_lookat(eye, at, up, 40.0f, w, h, 1, cam);
}
static inline ray_t camera_ray(const pto_camera* c, float u, float v) {   /* sensor.h:31-33 */
    v3 o = vload(c->origin);
    v3 dir = vsub(vadd(vadd(vload(c->llc), vscale(u, vload(c->horizontal))), vscale(v, vload(c->vertical))), o);
    return make_ray(o, dir);
}

/* ============================================================================================ */
/* Scene::intersect_bvh_optimized (scene.h:50-110) and intersect_linear (scene.h:113-129)           */
/* ============================================================================================ */
static int scene_intersect(const pto_scene* s, const ray_t* r, float t_min, float t_max, int mode, hit_t* best, uint64_t* ctr) {
    int best_prim = -1;
    if (mode == 1 || s->n_nodes == 0) {
        float cur = t_max; hit_t tmp;
        for (int i = 0; i < s->n_prims; i++) {
            if (ctr) ctr[1]++;
            if (prim_intersect(&s->prims[i], r, t_min, cur, &tmp) && tmp.t < cur) { *best = tmp; cur = tmp.t; best_prim = i; }
        }
        return best_prim;
    }
    float closest = t_max;
    int stack[64], sp = 0;
    stack[sp++] = 0;
    float inv[3] = {1.0f / r->d.e[0], 1.0f / r->d.e[1], 1.0f / r->d.e[2]};
    while (sp > 0) {
        const pto_node* nd = &s->nodes[stack[--sp]];
        if (ctr) ctr[0]++;
        float tmin_box = t_min, tmax_box = closest;
        for (int a = 0; a < 3; a++) {
            float t0 = (nd->bmin[a] - r->o.e[a]) * inv[a];
            float t1 = (nd->bmax[a] - r->o.e[a]) * inv[a];
            if (inv[a] < 0.0f) { float tmp = t0; t0 = t1; t1 = tmp; }
            tmin_box = t0 > tmin_box ? t0 : tmin_box;
            tmax_box = t1 < tmax_box ? t1 : tmax_box;
        }
        if (tmax_box < tmin_box) continue;
        if (nd->count > 0) {
            for (int i = 0; i < nd->count; i++) {
                int pi = s->idx[nd->left + i]; hit_t tmp;
                if (ctr) ctr[1]++;
                if (prim_intersect(&s->prims[pi], r, t_min, closest, &tmp) && tmp.t < closest) { *best = tmp; closest = tmp.t; best_prim = pi; }
            }
        } else if (sp < 62) { stack[sp++] = nd->right; stack[sp++] = nd->left; }
    }
    return best_prim;
}
int pto_intersect(const pto_scene* s, const float o[3], const float d[3], float t_min, float t_max, int mode,
                  float* t_out, uint64_t* counters) {
    ray_t r = make_ray(vload(o), vload(d));
    hit_t h; h.t = -1.0f;
    int id = scene_intersect(s, &r, t_min, t_max, mode, &h, counters);
    if (t_out) *t_out = id >= 0 ? h.t : -1.0f;
    return id;
}
void pto_primary_ids(const pto_scene* s, const pto_camera* cam, int32_t* ids, float* ts) {
    int w = cam->width, hgt = cam->height;
    #pragma omp parallel for schedule(dynamic, 4)
    for (int y = 0; y < hgt; y++)
        for (int x = 0; x < w; x++) {
            float u = (x + 0.5f) / (float)w, v = (y + 0.5f) / (float)hgt;
            ray_t r = camera_ray(cam, u, v);
            hit_t h; int id = scene_intersect(s, &r, 1e-4f, FLT_MAX, 0, &h, NULL);
            ids[y * w + x] = id; ts[y * w + x] = id >= 0 ? h.t : -1.0f;
        }
}

/* ============================================================================================ */
/* Scene::occluded (scene.h:132-149) with occluded_bvh (scene.h:152-190); mode 1 = its linear branch  */
/* ============================================================================================ */
int pto_occluded(const pto_scene* s, const float from[3], const float to[3], int mode) {
    v3 f = vload(from);
    v3 dir = vsub(vload(to), f);
    float dist = vlen(dir);
    dir = vdiv(dir, dist);
    ray_t r = make_ray(vadd(f, vscale(1e-4f, dir)), dir);                 /* Ray ctor normalises again */
    const float t_min = 1e-4f, t_max = dist - 1e-4f;
    hit_t tmp;
    if (mode == 1 || s->n_nodes == 0) {
        for (int i = 0; i < s->n_prims; i++)
            if (prim_intersect(&s->prims[i], &r, t_min, t_max, &tmp)) return 1;
        return 0;
    }
    int stack[64], sp = 0;
    stack[sp++] = 0;
    float inv[3] = {1.0f / r.d.e[0], 1.0f / r.d.e[1], 1.0f / r.d.e[2]};
    while (sp > 0) {
        const pto_node* nd = &s->nodes[stack[--sp]];
        float tmin_box = t_min, tmax_box = t_max;
        for (int a = 0; a < 3; a++) {
            float t0 = (nd->bmin[a] - r.o.e[a]) * inv[a];
            float t1 = (nd->bmax[a] - r.o.e[a]) * inv[a];
            if (inv[a] < 0.0f) { float t = t0; t0 = t1; t1 = t; }
            tmin_box = t0 > tmin_box ? t0 : tmin_box;
            tmax_box = t1 < tmax_box ? t1 : tmax_box;
        }
        if (tmax_box < tmin_box) continue;
        if (nd->count > 0) {
            for (int i = 0; i < nd->count; i++)
                if (prim_intersect(&s->prims[s->idx[nd->left + i]], &r, t_min, t_max, &tmp)) return 1;
        } else if (sp < 62) { stack[sp++] = nd->right; stack[sp++] = nd->left; }
    }
    return 0;
}

/* ============================================================================================ */
/* render_radiosity (integrator.h:460-504): primary visibility, Le + per-primitive radiosity, sqrt    */
/* `radiosity` = 3 floats per primitive (what the out-of-scope solver stores in the primitives) or NULL */
/* ============================================================================================ */
void pto_visibility(const pto_scene* s, const pto_camera* cam, pto_xorwow* rng, int spp, const float* radiosity, uint8_t* rgb8) {
    int w = cam->width, hgt = cam->height;
    #pragma omp parallel for schedule(dynamic, 1)
    for (int y = 0; y < hgt; y++)
        for (int x = 0; x < w; x++) {
            int pix = y * w + x;
            pto_xorwow st = rng[pix];
            v3 color = V(0, 0, 0);
            for (int k = 0; k < spp; k++) {
                float u = (x + pto_uniform(&st)) / (float)w;
                float v = (y + pto_uniform(&st)) / (float)hgt;
                ray_t r = camera_ray(cam, u, v);
                hit_t h;
                int id = scene_intersect(s, &r, 1e-4f, FLT_MAX, 0, &h, NULL);
                if (id >= 0) {
                    color = vadd(color, vload(s->prims[id].Le));
                    if (radiosity) color = vadd(color, vload(radiosity + 3 * id));
                    else color = vadd(color, V(0, 0, 0));
                }
            }
            color = vdiv(color, (float)spp);
            for (int c = 0; c < 3; c++) rgb8[3 * pix + c] = (unsigned char)(255.99f * sqrtf(fminf(color.e[c], 1.0f)));
            rng[pix] = st;
        }
}

/* ============================================================================================ */
/* sampleCosineHemisphere (integrator.h:62-85), integrator() BSDF branch (integrator.h:189-268)       */
/* ============================================================================================ */
static v3 sample_cosine_hemisphere(v3 n, pto_xorwow* rng) {
    float u = pto_uniform(rng);
    float v = pto_uniform(rng);
    float r = sqrtf(u);
    float phi = 2.0f * M_PI * v;                     /* double product, rounded to float on assignment */
    float x = r * cosf(phi);
    float y = r * sinf(phi);
    float z = sqrtf(fmaxf(0.0f, 1.0f - u));
    v3 tangent, bitangent;
    if (n.e[2] < -0.9999999f) { tangent = V(0.0f, -1.0f, 0.0f); bitangent = V(-1.0f, 0.0f, 0.0f); }
    else {
        float a = 1.0f / (1.0f + n.e[2]);
        float b = -n.e[0] * n.e[1] * a;
        tangent = V(1.0f - n.e[0] * n.e[0] * a, b, -n.e[0]);
        bitangent = V(b, 1.0f - n.e[1] * n.e[1] * a, -n.e[1]);
    }
    return vunit(vadd(vadd(vscale(x, tangent), vscale(y, bitangent)), vscale(z, n)));
}
/* ============================================================================================ */
/* Grid (rendering/grid.h), PrecomputedCDF (render_config.h:27-34), sampleMIS (integrator.h:112-166)   */
/* M_PI is the DOUBLE constant of <math.h> here: grid.h:22-24's float fallback never triggers on Linux, */
/* so every product with M_PI below is evaluated in double and rounded on assignment, as in the C++.    */
/* ============================================================================================ */
#define G_RES 16
#define G_SIZE 256
#define G_HALF 8
#define G_INV_RES (1.0f / G_RES)
#define G_INV_HALF (1.0f / G_HALF)
#define G_D_THETA ((M_PI * 0.5f) / G_HALF)
#define G_D_PHI (2.0f * M_PI / G_RES)
typedef struct {
    float pdf[G_SIZE], row_sums[G_HALF], marginal_cdf[G_HALF], row_cdfs[G_SIZE];
    float total_weight;
    int is_valid;
} cdf_t;

static void build_cdf(const float* rgb, cdf_t* c) {       /* initFromRadiosity + buildCDFs, grid.h:62-137 */
    for (int i = 0; i < G_SIZE; i++)
        c->pdf[i] = 0.2126f * rgb[3 * i] + 0.7152f * rgb[3 * i + 1] + 0.0722f * rgb[3 * i + 2];
    c->total_weight = 0.0f;
    for (int v = 0; v < G_HALF; v++) {
        float row_sum = 0.0f;
        for (int u = 0; u < G_RES; u++) row_sum += c->pdf[v * G_RES + u];
        c->row_sums[v] = row_sum;
        c->total_weight += row_sum;
    }
    float running = 0.0f;
    const float inv_total = (c->total_weight > 1e-6f) ? (1.0f / c->total_weight) : 0.0f;
    for (int v = 0; v < G_HALF; v++) { running += c->row_sums[v]; c->marginal_cdf[v] = running * inv_total; }
    c->marginal_cdf[G_HALF - 1] = 1.0f;
    for (int i = 0; i < G_SIZE; i++) c->row_cdfs[i] = ((i % G_RES) + 1) * G_INV_RES;   /* lower rows: precomputeCDFs :563-568; unused by sampling */
    for (int v = 0; v < G_HALF; v++) {
        float row_sum = c->row_sums[v];
        if (row_sum < 1e-6f) {
            for (int u = 0; u < G_RES; u++) c->row_cdfs[v * G_RES + u] = (u + 1) * G_INV_RES;
        } else {
            float running_row = 0.0f;
            const float inv_row_sum = 1.0f / row_sum;
            for (int u = 0; u < G_RES; u++) { running_row += c->pdf[v * G_RES + u]; c->row_cdfs[v * G_RES + u] = running_row * inv_row_sum; }
            c->row_cdfs[v * G_RES + G_RES - 1] = 1.0f;
        }
    }
    c->is_valid = c->total_weight > 1e-6f;
}
void pto_scene_set_grids(pto_scene* s, const float* rgb, float mis_bsdf_fraction) {
    free(s->guide); s->guide = NULL;
    s->mis_bsdf_fraction = mis_bsdf_fraction;
    if (!rgb) return;
    cdf_t* c = (cdf_t*)malloc(sizeof(cdf_t) * (size_t)(s->n_prims > 0 ? s->n_prims : 1));
    for (int p = 0; p < s->n_prims; p++) build_cdf(rgb + (size_t)p * G_SIZE * 3, &c[p]);
    s->guide = c;
}
void pto_scene_get_cdf(const pto_scene* s, int prim, float* out) {
    const cdf_t* c = &((const cdf_t*)s->guide)[prim];
    memcpy(out, c->pdf, sizeof c->pdf); memcpy(out + 256, c->row_sums, sizeof c->row_sums);
    memcpy(out + 264, c->marginal_cdf, sizeof c->marginal_cdf); memcpy(out + 272, c->row_cdfs, sizeof c->row_cdfs);
    out[528] = c->total_weight; out[529] = (float)c->is_valid;
}
static void build_frame(v3 n, v3* t, v3* b) {             /* grid.h:300-309 */
    if (n.e[2] < -0.9999999f) { *t = V(0.0f, -1.0f, 0.0f); *b = V(-1.0f, 0.0f, 0.0f); return; }
    const float a = 1.0f / (1.0f + n.e[2]);
    const float c = -n.e[0] * n.e[1] * a;
    *t = V(1.0f - n.e[0] * n.e[0] * a, c, -n.e[0]);
    *b = V(c, 1.0f - n.e[1] * n.e[1] * a, -n.e[1]);
}
static int linear_search_cdf(const float* cdf, int size, float xi) {   /* grid.h:248-255 */
    xi = fminf(fmaxf(xi, 0.0f), 0.999999f);
    for (int i = 0; i < size; i++) if (xi < cdf[i]) return i;
    return size - 1;
}
static float pdf_for_cell(const cdf_t* c, int theta_idx, int phi_idx) {   /* grid.h:258-274 */
    float cell_value = c->pdf[theta_idx * G_RES + phi_idx];
    if (cell_value < 1e-8f) return 1e-6f;
    float cell_prob = cell_value / fmaxf(c->total_weight, 1e-6f);
    float theta_center = (theta_idx + 0.5f) * G_INV_HALF * (M_PI * 0.5f);
    float sin_theta = fmaxf(sinf(theta_center), 0.01f);
    float solid_angle = sin_theta * G_D_THETA * G_D_PHI;
    return cell_prob / fmaxf(solid_angle, 1e-6f);
}
static v3 grid_sample(const cdf_t* c, v3 normal, pto_xorwow* rng, float* out_pdf) {   /* Grid::sample, grid.h:143-190 (valid grids only) */
    float xi1 = pto_uniform(rng);
    float xi2 = pto_uniform(rng);
    int theta_idx = linear_search_cdf(c->marginal_cdf, G_HALF, xi1);
    int phi_idx = linear_search_cdf(&c->row_cdfs[theta_idx * G_RES], G_RES, xi2);
    float jitter_theta = pto_uniform(rng);
    float jitter_phi = pto_uniform(rng);
    float theta = (theta_idx + jitter_theta) * G_INV_HALF * (M_PI * 0.5f);
    theta = fminf(theta, (float)(M_PI * 0.5f - 0.01f));
    float phi = (phi_idx + jitter_phi) * G_INV_RES * 2.0f * M_PI;
    float sin_t, cos_t, sin_p, cos_p;
    sincosf(theta, &sin_t, &cos_t);
    sincosf(phi, &sin_p, &cos_p);
    v3 local = V(sin_t * cos_p, sin_t * sin_p, cos_t);
    v3 tangent, bitangent;
    build_frame(normal, &tangent, &bitangent);
    v3 world = vunit(vadd(vadd(vscale(local.e[0], tangent), vscale(local.e[1], bitangent)), vscale(local.e[2], normal)));
    *out_pdf = pdf_for_cell(c, theta_idx, phi_idx);
    return world;
}
static float grid_compute_pdf(const cdf_t* c, v3 dir, v3 normal) {       /* Grid::computePDF, grid.h:200-216 + worldToSpherical :312-325 */
    v3 tangent, bitangent;
    build_frame(normal, &tangent, &bitangent);
    float lx = vdot(dir, tangent), ly = vdot(dir, bitangent), lz = vdot(dir, normal);
    float theta = acosf(fminf(fmaxf(lz, -1.0f), 1.0f));
    float phi = atan2f(ly, lx);
    if (phi < 0.0f) phi += 2.0f * M_PI;
    if (theta > M_PI * 0.5f) return 0.0f;
    int theta_idx = (int)(theta * (2.0f / M_PI) * G_HALF);
    int phi_idx = (int)(phi * (0.5f / M_PI) * G_RES);
    theta_idx = theta_idx < 0 ? 0 : (theta_idx > G_HALF - 1 ? G_HALF - 1 : theta_idx);
    phi_idx = phi_idx < 0 ? 0 : (phi_idx > G_RES - 1 ? G_RES - 1 : phi_idx);
    return pdf_for_cell(c, theta_idx, phi_idx);
}
static float mis_power(float a, float b) {                                /* integrator.h:90-95 */
    if (a <= 0.0f) return 0.0f;
    float a2 = a * a, b2 = b * b;
    return a2 / (a2 + b2);
}
static v3 sample_cosine_hemisphere(v3 n, pto_xorwow* rng);
static v3 sample_mis(const cdf_t* c, v3 normal, pto_xorwow* rng, float* weight, float bsdf_prob) {   /* integrator.h:112-166 */
    const float BSDF_PROB = fmaxf(fminf(bsdf_prob, 0.99f), 0.01f);
    const float GRID_PROB = 1.0f - BSDF_PROB;
    float xi = pto_uniform(rng);
    v3 dir;
    float pdf_grid, pdf_bsdf, mis_w;
    if (xi < BSDF_PROB) {
        dir = sample_cosine_hemisphere(normal, rng);
        float cos_theta = fmaxf(vdot(dir, normal), 0.0f);
        pdf_bsdf = cos_theta / M_PI;
        pdf_grid = grid_compute_pdf(c, dir, normal);
        mis_w = mis_power(pdf_bsdf, pdf_grid);
        *weight = pdf_bsdf > 1e-6f ? mis_w / BSDF_PROB : 0.0f;
    } else {
        dir = grid_sample(c, normal, rng, &pdf_grid);
        float cos_theta = fmaxf(vdot(dir, normal), 0.0f);
        pdf_bsdf = cos_theta / M_PI;
        mis_w = mis_power(pdf_grid, pdf_bsdf);
        if (pdf_grid > 1e-6f && cos_theta > 0.0f) {
            float w = mis_w * cos_theta / (M_PI * pdf_grid * GRID_PROB);
            *weight = fminf(w, 10.0f);
        } else *weight = 0.0f;
    }
    return dir;
}

/* mode: SamplingMode (render_config.h:41-47).  0 = BSDF, 3 = MIS, anything else = pure grid sampling (integrator.h:226-262) */
static v3 integrate_mode(const pto_scene* s, ray_t r, int max_depth, pto_xorwow* rng, int mode, uint64_t* stats);
static v3 integrate(const pto_scene* s, ray_t r, int max_depth, pto_xorwow* rng, uint64_t* stats) {
    return integrate_mode(s, r, max_depth, rng, 0, stats);
}
static v3 integrate_mode(const pto_scene* s, ray_t r, int max_depth, pto_xorwow* rng, int mode, uint64_t* stats) {
    v3 L = V(0, 0, 0), beta = V(1.0f, 1.0f, 1.0f);
    for (int depth = 0; depth < max_depth; depth++) {
        hit_t h;
        if (stats) stats[0]++;
        int id = scene_intersect(s, &r, 1e-4f, FLT_MAX, 0, &h, stats ? stats + 2 : NULL);
        if (id < 0) break;
        const pto_prim* p = &s->prims[id];
        L = vadd(L, vmul(beta, vload(p->Le)));
        if (depth > 2) {
            float mx = fmaxf(beta.e[0], fmaxf(beta.e[1], beta.e[2]));
            float rr = fminf(mx, 0.95f);
            if (pto_uniform(rng) > rr) break;
            float k = (float)(1.0 / (double)rr);                                  /* operator/=(T), vector.h:92-96 */
            beta = V(beta.e[0] * k, beta.e[1] * k, beta.e[2] * k);
        }
        beta = vmul(beta, vload(p->bsdf));
        if (vlen(beta) < 1e-5f) break;
        v3 n = vload(p->n);
        v3 ns = vdot(r.d, n) < 0 ? n : vneg(n);
        v3 next;
        const cdf_t* c = (mode != 0 && s->guide) ? &((const cdf_t*)s->guide)[id] : NULL;
        if (c && c->is_valid) {                                                   /* has_grid && grid.isValid() */
            float weight;
            if (mode == 3) next = sample_mis(c, ns, rng, &weight, s->mis_bsdf_fraction);
            else {
                float grid_pdf;
                next = grid_sample(c, ns, rng, &grid_pdf);
                float cos_theta = fmaxf(vdot(next, ns), 0.0f);
                weight = cos_theta / (M_PI * fmaxf(grid_pdf, 1e-6f));
                weight = fminf(fmaxf(weight, 0.0f), 10.0f);
            }
            beta = V(beta.e[0] * weight, beta.e[1] * weight, beta.e[2] * weight);   /* operator*=(T), vector.h */
        } else if (s->mirror && s->mirror[id]) next = vsub(r.d, vscale(2.0f * vdot(r.d, ns), ns));   /* extension: mirror, no draws */
        else next = sample_cosine_hemisphere(ns, rng);
        r = make_ray(vadd(h.p, vscale(1e-4f, ns)), next);
    }
    return L;
}

void pto_scene_set_mirrors(pto_scene* s, const unsigned char* flags) {
    free(s->mirror); s->mirror = NULL;
    if (!flags || s->n_prims <= 0) return;
    s->mirror = (unsigned char*)malloc((size_t)s->n_prims);
    memcpy(s->mirror, flags, (size_t)s->n_prims);
}

void pto_render_init(int w, int h, pto_xorwow* rng, int threads) {
    xw_prepare();
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads); else omp_set_num_threads(omp_get_num_procs());
#endif
    #pragma omp parallel for schedule(dynamic, 64)
    for (int p = 0; p < w * h; p++) pto_xorwow_init((uint64_t)(2023 + p), (uint64_t)p, &rng[p]);   /* integrator.h:278-279 */
}
void pto_tonemap(const float* linear, int n_pixels, uint8_t* rgb8) {            /* integrator.h:393-407 */
    const float gamma = 1.0f / 2.2f;
    for (int i = 0; i < n_pixels * 3; i++) {
        float c = linear[i];
        c = c / (c + 1.0f);
        c = powf(c, gamma);
        rgb8[i] = (unsigned char)(255.99f * fminf(c, 1.0f));
    }
}
void pto_render(const pto_scene* s, const pto_camera* cam, pto_xorwow* rng, int spp, int max_depth,
                float* linear, uint8_t* rgb8, uint64_t* stats, int threads) {
    int w = cam->width, hgt = cam->height;
    uint64_t tot[4] = {0, 0, 0, 0};
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads); else omp_set_num_threads(omp_get_num_procs());
#endif
    #pragma omp parallel
    {
        uint64_t loc[4] = {0, 0, 0, 0};
        #pragma omp for schedule(dynamic, 1)
        for (int y = 0; y < hgt; y++)
            for (int x = 0; x < w; x++) {
                int pix = y * w + x;
                pto_xorwow* st = &rng[pix];
                v3 color = V(0, 0, 0);
                for (int sidx = 0; sidx < spp; sidx++) {
                    float u = (x + pto_uniform(st)) / (float)w;
                    float v = (y + pto_uniform(st)) / (float)hgt;
                    ray_t r = camera_ray(cam, u, v);
                    color = vadd(color, integrate(s, r, max_depth, st, stats ? loc : NULL));
                    loc[1]++;
                }
                color = vdiv(color, (float)spp);
                if (linear) vstore(linear + 3 * pix, color);
                if (rgb8) pto_tonemap(color.e, 1, rgb8 + 3 * pix);
            }
        #pragma omp critical(pto_stats)
        for (int k = 0; k < 4; k++) tot[k] += loc[k];
    }
    if (stats) for (int k = 0; k < 4; k++) stats[k] += tot[k];
}
void pto_render_guided(const pto_scene* s, const pto_camera* cam, pto_xorwow* rng, int spp, int max_depth, int mode,
                       float* linear, uint64_t* stats, int threads) {
    int w = cam->width, hgt = cam->height;
    uint64_t tot[4] = {0, 0, 0, 0};
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads); else omp_set_num_threads(omp_get_num_procs());
#endif
    #pragma omp parallel
    {
        uint64_t loc[4] = {0, 0, 0, 0};
        #pragma omp for schedule(dynamic, 1)
        for (int y = 0; y < hgt; y++)
            for (int x = 0; x < w; x++) {
                int pix = y * w + x;
                pto_xorwow* st = &rng[pix];
                v3 color = V(0, 0, 0);
                for (int sidx = 0; sidx < spp; sidx++) {
                    float u = (x + pto_uniform(st)) / (float)w;
                    float v = (y + pto_uniform(st)) / (float)hgt;
                    ray_t r = camera_ray(cam, u, v);
                    color = vadd(color, integrate_mode(s, r, max_depth, st, mode, loc));
                    loc[1]++;
                }
                color = vdiv(color, (float)spp);
                if (linear) vstore(linear + 3 * pix, color);
            }
        #pragma omp critical(pto_stats)
        for (int k = 0; k < 4; k++) tot[k] += loc[k];
    }
    if (stats) for (int k = 0; k < 4; k++) stats[k] += tot[k];
}
void pto_render_moments(const pto_scene* s, const pto_camera* cam, pto_xorwow* rng, int spp, int max_depth,
                        double* sum, double* sumsq, int threads) {
    int w = cam->width, hgt = cam->height;
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads); else omp_set_num_threads(omp_get_num_procs());
#endif
    #pragma omp parallel for schedule(dynamic, 1)
    for (int y = 0; y < hgt; y++)
        for (int x = 0; x < w; x++) {
            int pix = y * w + x;
            pto_xorwow* st = &rng[pix];
            for (int sidx = 0; sidx < spp; sidx++) {
                float u = (x + pto_uniform(st)) / (float)w;
                float v = (y + pto_uniform(st)) / (float)hgt;
                ray_t r = camera_ray(cam, u, v);
                v3 c = integrate(s, r, max_depth, st, NULL);
                for (int k = 0; k < 3; k++) { sum[3 * pix + k] += c.e[k]; sumsq[3 * pix + k] += (double)c.e[k] * c.e[k]; }
            }
        }
}

/* mcpt_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement, in plain C and double precision, of the hot path of Arieys/MonteCarloPathTracing
 * (citations into /root/reference).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it, and only as the checker.
 *
 * Pinning: the reference has no golden vectors of its own.  This restatement is pinned
 * against outputs of the reference itself, built unmodified by oracle/build_ref.sh and run in the
 * build container by tests/golden/make_golden.py: Morton codes + sorted order (exact), closest-hit
 * records on exported ray batches (exact ids, t/p/pn to 1e-12), and images (statistical: the
 * reference's RNG is time-seeded and racy, MTPC/pathTracing.cpp:5,32,68,169).
 *
 * Deliberate differences from the reference, all documented in DESIGN.md:
 *   - RNG: Philox4x32-10 keyed (pixel, sample), counter (bounce, stream, seed) — the same mapping
 *     the CUDA integrator uses, so GPU and oracle images can be compared pixel by pixel.
 *   - sort: stable (ties by OBJ face index) instead of std::sort (MTPC/MTPC.cpp:44).
 *   - a light whose CDF selects no triangle (rnd >= its area, quirk Q6) contributes nothing
 *     (the reference evaluates NaN there; cannot happen in the shipped scenes).
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../include/mcpt.h"

typedef struct { double x, y, z; } V3;
static inline V3 v3(double x, double y, double z) { V3 r = {x, y, z}; return r; }
static inline V3 vadd(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline V3 vsub(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline V3 vmul(V3 a, double t) { return v3(a.x * t, a.y * t, a.z * t); }
static inline V3 vdiv(V3 a, double t) { return v3(a.x / t, a.y / t, a.z / t); }
static inline V3 vneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
static inline double vdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
/* Vertex::cross, MTPC/sceneManagement.h:69-75 (same operand order) */
static inline V3 vcross(V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, b.x * a.z - a.x * b.z, a.x * b.y - b.x * a.y); }
static inline double vnorm(V3 a) { return sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
static inline V3 vnormalize(V3 a) { double d = vnorm(a); return v3(a.x / d, a.y / d, a.z / d); }

/* ---- Morton code: MTPC/morton code.cpp:3-32, MINP/MAXP MTPC/morton code.h:6-7 ---------------- */
static uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static uint32_t morton3d(float x, float y, float z) {
    x = fminf(fmaxf(x * 1024.0f, 0.0f), 1023.0f);
    y = fminf(fmaxf(y * 1024.0f, 0.0f), 1023.0f);
    z = fminf(fmaxf(z * 1024.0f, 0.0f), 1023.0f);
    return expand_bits((uint32_t)x) * 4 + expand_bits((uint32_t)y) * 2 + expand_bits((uint32_t)z);
}
static uint32_t get_morton_code(float x, float y, float z) {
    float mx = (x - (-1)) / (float)(4 - (-1));
    float my = (y - (-1)) / (float)(4 - (-1));
    float mz = (z - (-1)) / (float)(4 - (-1));
    return morton3d(mx, my, mz);
}
/* centroid in double then narrowed at the call: MTPC/sceneManagement.cpp:176-179 */
MCPT_API void orc_morton(const double* pos, int n, uint32_t* codes) {
    for (int i = 0; i < n; i++) {
        const double* p = pos + 9 * (size_t)i;
        double cx = (p[0] + p[3] + p[6]) / 3, cy = (p[1] + p[4] + p[7]) / 3, cz = (p[2] + p[5] + p[8]) / 3;
        codes[i] = get_morton_code((float)cx, (float)cy, (float)cz);
    }
}
/* stable counting-free merge sort of indices by code (canonical tie rule = face index) */
MCPT_API void orc_sort(const uint32_t* codes, int n, int32_t* perm) {
    int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) perm[i] = i;
    for (int w = 1; w < n; w *= 2) {
        for (int lo = 0; lo < n; lo += 2 * w) {
            int mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
            int a = lo, b = mid, k = lo;
            while (a < mid && b < hi) tmp[k++] = (codes[perm[b]] < codes[perm[a]]) ? perm[b++] : perm[a++];
            while (a < mid) tmp[k++] = perm[a++];
            while (b < hi) tmp[k++] = perm[b++];
        }
        memcpy(perm, tmp, sizeof(int32_t) * (size_t)n);
    }
    free(tmp);
}

/* ---- scene + the reference's balanced BVH (MTPC/BVH.cpp:44-132) -------------------------------- */
typedef struct { double mn[3], mx[3]; int empty; } Box;
typedef struct {
    V3 v1, v2, v3, n1, n2, n3, norm; /* norm = Face::calNorm, MTPC/sceneManagement.cpp:408-412 */
    double uv[6];
    int mat, obj_idx;
} Tri;
typedef struct orc_scene {
    int n, Lc, Level;       /* faces, leaf slots 2^ceil(log2 n), tree depth */
    Tri* tri;               /* Morton-sorted */
    Box* box;               /* complete tree, 2*Lc-1 slots, heap order; virtual nodes empty */
    uint32_t* codes;        /* OBJ order */
    int32_t* perm;          /* sorted k -> OBJ index */
    int n_mat, n_tex, n_light;
    mcpt_material* mat;
    int* mat_light;         /* material -> light index or -1 (light_map, MTPC/sceneManagement.cpp:256) */
    int32_t* tex_h; int32_t* tex_w; uint8_t** tex;
    mcpt_light* light;
    int** light_tris; int* light_ntri; double** light_cdf; double* light_area; /* OBJ order CDFs */
    Tri* obj_tri;           /* OBJ order copy for light sampling */
} orc_scene;

static double dmin3(double a, double b, double c) { if (a <= b && a <= c) return a; else if (b <= a && b <= c) return b; else return c; }
static double dmax3(double a, double b, double c) { if (a >= b && a >= c) return a; else if (b >= a && b >= c) return b; else return c; }

/* Face::calAera (law of cosines), MTPC/sceneManagement.cpp:399-406 */
static double tri_area(const Tri* f) {
    double a = vnorm(vsub(f->v2, f->v1)), b = vnorm(vsub(f->v3, f->v1)), c = vnorm(vsub(f->v3, f->v2));
    double cos_c = (a * a + b * b - c * c) / (2 * a * b);
    double sin_c = sqrt(1 - pow(cos_c, 2));
    return a * b * sin_c / 2;
}

MCPT_API void orc_free(orc_scene* S) {
    if (!S) return;
    free(S->tri); free(S->box); free(S->codes); free(S->perm); free(S->mat); free(S->mat_light);
    for (int i = 0; i < S->n_tex; i++) free(S->tex[i]);
    free(S->tex); free(S->tex_h); free(S->tex_w); free(S->light);
    for (int i = 0; i < S->n_light; i++) { free(S->light_tris[i]); free(S->light_cdf[i]); }
    free(S->light_tris); free(S->light_ntri); free(S->light_cdf); free(S->light_area); free(S->obj_tri);
    free(S);
}

MCPT_API orc_scene* orc_build(const mcpt_scene_desc* d) {
    orc_scene* S = (orc_scene*)calloc(1, sizeof(orc_scene));
    int n = d->n_faces;
    S->n = n;
    S->codes = (uint32_t*)malloc(4 * (size_t)n); S->perm = (int32_t*)malloc(4 * (size_t)n);
    orc_morton(d->pos, n, S->codes);
    orc_sort(S->codes, n, S->perm);
    S->obj_tri = (Tri*)malloc(sizeof(Tri) * (size_t)n); S->tri = (Tri*)malloc(sizeof(Tri) * (size_t)n);
    for (int i = 0; i < n; i++) {
        Tri* t = &S->obj_tri[i];
        const double* p = d->pos + 9 * (size_t)i; const double* q = d->nrm + 9 * (size_t)i;
        t->v1 = v3(p[0], p[1], p[2]); t->v2 = v3(p[3], p[4], p[5]); t->v3 = v3(p[6], p[7], p[8]);
        t->n1 = v3(q[0], q[1], q[2]); t->n2 = v3(q[3], q[4], q[5]); t->n3 = v3(q[6], q[7], q[8]);
        memcpy(t->uv, d->uv + 6 * (size_t)i, 48);
        t->norm = vnormalize(vcross(vsub(t->v1, t->v2), vsub(t->v3, t->v1)));
        t->mat = d->mat_id[i]; t->obj_idx = i;
    }
    for (int k = 0; k < n; k++) S->tri[k] = S->obj_tri[S->perm[k]];
    /* complete tree: Lc = 2^ceil(log2 t), Level = log2(Lc) (BVH.cpp:46-52); leaf k = sorted face k (:62-67);
       inner = union of children or copy of the left child when the right one is virtual (:70-79) */
    int Lc = 1, Level = 0;
    while (Lc < n) { Lc *= 2; Level++; }
    S->Lc = Lc; S->Level = Level;
    S->box = (Box*)malloc(sizeof(Box) * (size_t)(2 * Lc - 1));
    for (int k = 0; k < Lc; k++) {
        Box* b = &S->box[Lc - 1 + k];
        if (k >= n) { b->empty = 1; continue; }
        const Tri* t = &S->tri[k];
        b->empty = 0; /* BVHNode::findBondingBox(Face&), BVH.cpp:87-97 */
        b->mx[0] = dmax3(t->v1.x, t->v2.x, t->v3.x); b->mx[1] = dmax3(t->v1.y, t->v2.y, t->v3.y); b->mx[2] = dmax3(t->v1.z, t->v2.z, t->v3.z);
        b->mn[0] = dmin3(t->v1.x, t->v2.x, t->v3.x); b->mn[1] = dmin3(t->v1.y, t->v2.y, t->v3.y); b->mn[2] = dmin3(t->v1.z, t->v2.z, t->v3.z);
    }
    for (int i = Lc - 2; i >= 0; i--) {
        Box* b = &S->box[i]; const Box* l = &S->box[2 * i + 1]; const Box* r = &S->box[2 * i + 2];
        if (l->empty) { b->empty = 1; continue; }
        *b = *l;
        if (!r->empty) for (int a = 0; a < 3; a++) { b->mx[a] = fmax(l->mx[a], r->mx[a]); b->mn[a] = fmin(l->mn[a], r->mn[a]); }
    }
    /* materials / textures / lights */
    S->n_mat = d->n_materials; S->n_tex = d->n_textures; S->n_light = d->n_lights;
    S->mat = (mcpt_material*)malloc(sizeof(mcpt_material) * (size_t)(S->n_mat + 1));
    memcpy(S->mat, d->materials, sizeof(mcpt_material) * (size_t)S->n_mat);
    S->mat_light = (int*)malloc(sizeof(int) * (size_t)(S->n_mat + 1));
    for (int m = 0; m < S->n_mat; m++) S->mat_light[m] = -1;
    S->tex = (uint8_t**)calloc((size_t)S->n_tex + 1, sizeof(uint8_t*));
    S->tex_h = (int32_t*)calloc((size_t)S->n_tex + 1, 4); S->tex_w = (int32_t*)calloc((size_t)S->n_tex + 1, 4);
    for (int i = 0; i < S->n_tex; i++) {
        size_t sz = (size_t)d->textures[i].height * d->textures[i].width * 3;
        S->tex_h[i] = d->textures[i].height; S->tex_w[i] = d->textures[i].width;
        S->tex[i] = (uint8_t*)malloc(sz ? sz : 1); memcpy(S->tex[i], d->textures[i].bgr, sz);
    }
    S->light = (mcpt_light*)malloc(sizeof(mcpt_light) * (size_t)(S->n_light + 1));
    memcpy(S->light, d->lights, sizeof(mcpt_light) * (size_t)S->n_light);
    S->light_tris = (int**)calloc((size_t)S->n_light + 1, sizeof(int*)); S->light_cdf = (double**)calloc((size_t)S->n_light + 1, sizeof(double*));
    S->light_ntri = (int*)calloc((size_t)S->n_light + 1, sizeof(int)); S->light_area = (double*)calloc((size_t)S->n_light + 1, sizeof(double));
    for (int l = 0; l < S->n_light; l++) {
        int m = S->light[l].mat_id, cnt = 0;
        S->mat_light[m] = l; /* later mtlname lines overwrite light_map[name] */
        for (int i = 0; i < n; i++) cnt += (d->mat_id[i] == m);
        S->light_tris[l] = (int*)malloc(sizeof(int) * (size_t)(cnt + 1)); S->light_cdf[l] = (double*)malloc(8 * (size_t)(cnt + 1));
        S->light_ntri[l] = cnt;
        double total = 0; int k = 0;
        for (int i = 0; i < n; i++) if (d->mat_id[i] == m) { /* CDF in OBJ order, pathTracing.cpp:177-184 */
            total += tri_area(&S->obj_tri[i]); S->light_tris[l][k] = i; S->light_cdf[l][k] = total; k++;
        }
        S->light_area[l] = total;
    }
    return S;
}
MCPT_API void orc_get_order(const orc_scene* S, uint32_t* codes_obj, int32_t* perm) {
    if (codes_obj) memcpy(codes_obj, S->codes, 4 * (size_t)S->n);
    if (perm) memcpy(perm, S->perm, 4 * (size_t)S->n);
}
MCPT_API double orc_light_area(const orc_scene* S, int l) { return S->light_area[l]; }

/* ---- intersection core ------------------------------------------------------------------------ */
typedef struct { V3 o, d; } Ray;
typedef struct { int hit, k; double t; V3 p, pn; } Isect; /* k = sorted leaf index */

/* intersect(Ray&, boundingBox&), MTPC/sceneManagement.cpp:340-391 */
static int hit_box(const Ray* r, const Box* b) {
    double txmin = (b->mn[0] - r->o.x) / r->d.x, txmax = (b->mx[0] - r->o.x) / r->d.x;
    double tymin = (b->mn[1] - r->o.y) / r->d.y, tymax = (b->mx[1] - r->o.y) / r->d.y;
    double tzmin = (b->mn[2] - r->o.z) / r->d.z, tzmax = (b->mx[2] - r->o.z) / r->d.z;
    double tmp;
    if (txmin > txmax) { tmp = txmin; txmin = txmax; txmax = tmp; }
    if (tymin > tymax) { tmp = tymin; tymin = tymax; tymax = tmp; }
    if (tzmin > tzmax) { tmp = tzmin; tzmin = tzmax; tzmax = tmp; }
    if (txmax < 0 || tymax < 0 || tzmax < 0) return 0;
    if (txmin <= 0 && tymin <= 0 && tzmin <= 0) return 1;
    if (dmax3(txmin, tymin, tzmin) <= dmin3(txmax, tymax, tzmax)) return 1;
    return 0;
}
/* intersect(Ray&, Face&, Vertex&), MTPC/sceneManagement.cpp:316-338 */
static int hit_tri(const Ray* r, const Tri* f, V3* ret) {
    V3 norm = f->norm;
    double t = vdot(vsub(f->v1, r->o), norm) / vdot(norm, r->d);
    V3 p = vadd(r->o, vmul(r->d, t));
    V3 ap = vsub(p, f->v1), bp = vsub(p, f->v2), cp = vsub(p, f->v3);
    V3 ab = vsub(f->v2, f->v1), bc = vsub(f->v3, f->v2), ca = vsub(f->v1, f->v3);
    double dir1 = vdot(vcross(ab, ap), norm), dir2 = vdot(vcross(bc, bp), norm), dir3 = vdot(vcross(ca, cp), norm);
    double j1 = dir1 * dir2, j2 = dir1 * dir3, j3 = dir2 * dir3;
    *ret = p;
    return (j1 >= 0 && j2 >= 0 && j3 >= 0);
}
/* findGarCor, MTPC/pathTracing.cpp:394-432 */
static V3 bary_of(const Tri* f, V3 p) {
    V3 e1 = vsub(f->v3, f->v2), e2 = vsub(f->v1, f->v3), e3 = vsub(f->v2, f->v1);
    V3 d1 = vsub(p, f->v1), d2 = vsub(p, f->v2), d3 = vsub(p, f->v3);
    V3 n = vcross(e1, e2);
    double an = vdot(n, n);
    return v3(vdot(vcross(e1, d3), n) / an, vdot(vcross(e2, d1), n) / an, vdot(vcross(e3, d2), n) / an);
}
/* bvh_intersect, MTPC/pathTracing.cpp:334-374: un-pruned DFS, left child first; heap node i, level l */
static void bvh_walk(const orc_scene* S, const Ray* r, Isect* v, int node, int level, long* nbox, long* ntri) {
    const Box* b = &S->box[node];
    if (b->empty) return;
    if (nbox) (*nbox)++;
    if (!hit_box(r, b)) return;
    if (level == S->Level) {
        int k = node - (S->Lc - 1);
        V3 p;
        if (ntri) (*ntri)++;
        if (hit_tri(r, &S->tri[k], &p)) {
            double t = (p.x - r->o.x) / r->d.x;               /* :347 — t from the x coordinate only */
            if (t > 0 && (!v->hit || t < v->t)) {            /* :355-361 strict <, first found wins ties */
                V3 g = bary_of(&S->tri[k], p);
                const Tri* f = &S->tri[k];
                v->hit = 1; v->k = k; v->t = t; v->p = p;
                v->pn = vadd(vadd(vmul(f->n1, g.x), vmul(f->n2, g.y)), vmul(f->n3, g.z)); /* :351, not normalised */
            }
        }
        return;
    }
    bvh_walk(S, r, v, 2 * node + 1, level + 1, nbox, ntri);
    bvh_walk(S, r, v, 2 * node + 2, level + 1, nbox, ntri);
}
/* ray_intersect, MTPC/pathTracing.cpp:382-390 */
static int ray_intersect(const orc_scene* S, const Ray* r, Isect* ret) {
    ret->hit = 0; ret->k = -1; ret->t = 0;
    bvh_walk(S, r, ret, 0, 0, 0, 0);
    return ret->hit;
}

/* rays: n x 6 doubles.  Per ray out: hit, face (OBJ order), leaf (sorted order), rec7 = t, p[3], pn[3]. */
MCPT_API void orc_trace(const orc_scene* S, const double* rays, int n, int32_t* hit, int32_t* face, int32_t* leaf,
                        double* rec7, double* counts2) {
    long nb = 0, nt = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : nb, nt)
    for (int i = 0; i < n; i++) {
        Ray r = {v3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]), v3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5])};
        Isect v; v.hit = 0; v.k = -1; v.t = 0; v.p = v3(0, 0, 0); v.pn = v3(0, 0, 0);
        long b = 0, t = 0;
        bvh_walk(S, &r, &v, 0, 0, &b, &t);
        nb += b; nt += t;
        hit[i] = v.hit; face[i] = v.hit ? S->tri[v.k].obj_idx : -1; leaf[i] = v.k;
        double* o = rec7 + 7 * (size_t)i;
        o[0] = v.t; o[1] = v.p.x; o[2] = v.p.y; o[3] = v.p.z; o[4] = v.pn.x; o[5] = v.pn.y; o[6] = v.pn.z;
    }
    if (counts2) { counts2[0] = (double)nb; counts2[1] = (double)nt; }
}

/* ---- RNG: Philox4x32-10 (Salmon et al. 2011), the mapping shared with the CUDA integrator ------ */
static void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
MCPT_API void orc_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1], out);
}
typedef struct { uint32_t pixel, sample, seed_lo, seed_hi; long n_shadow, n_extend; int flags; } Rng;
enum { STREAM_PATH = 0, STREAM_FRESNEL = 1, STREAM_LIGHT0 = 2, STREAM_JITTER = 0x7FFFFFFF };
static void draw4(const Rng* g, int bounce, int stream, double u[4]) {
    uint32_t o[4];
    philox4x32_10((uint32_t)bounce, (uint32_t)stream, g->seed_lo, g->seed_hi, g->pixel, g->sample, o);
    for (int i = 0; i < 4; i++) u[i] = (double)(o[i] >> 8) * (1.0 / 16777216.0);
}

#define ORC_PI 3.1415926 /* MTPC/pathTracing.h:11 */
#define ORC_P_RR 0.6     /* MTPC/pathTracing.cpp:237 */
#define ORC_MAX_DEPTH 64 /* the reference recursion is unbounded; 0.6^64 ~ 6e-15 */
enum { RT_DIFFUSE = 0, RT_SPECULAR = 1, RT_TRANSMISSION = 2 };

/* Refract, MTPC/pathTracing.cpp:13-27 (float intermediates kept) */
static int refract_dir(V3 i, V3 n, double eta, V3* out) {
    float cosi = (float)vdot(i, n);
    float cost2 = (float)(1.0f - eta * eta * (1.0f - cosi * cosi));
    if (cost2 >= 0.0f) { *out = vsub(vmul(i, eta), vmul(n, eta * cosi + sqrtf(cost2))); return 1; }
    return 0;
}
/* BRDFImportanceSampling, MTPC/pathTracing.cpp:30-64; phi's uniform is drawn before theta's */
static V3 brdf_sample(V3 axis, int type, double Ns, double u_phi, double u_theta) {
    double phi = u_phi * 2 * ORC_PI, theta;
    if (type == RT_DIFFUSE) theta = asin(sqrt(u_theta));
    else theta = acos(pow(u_theta, (double)1 / (Ns + 1)));
    V3 s = v3(sin(theta) * cos(phi), cos(theta), sin(theta) * sin(phi));
    V3 front;
    if (fabs(axis.x) > fabs(axis.y)) front = vnormalize(v3(axis.z, 0, -axis.x));
    else front = vnormalize(v3(0, -axis.z, axis.y));
    V3 right = vcross(axis, front);
    return vnormalize(vadd(vadd(vmul(right, s.x), vmul(axis, s.y)), vmul(front, s.z)));
}
/* nextRay, MTPC/pathTracing.cpp:66-134 */
static Ray next_ray(const orc_scene* S, const Isect* p, V3 wo, V3 kd, const mcpt_material* m, const Rng* g, int bounce, int* type) {
    double u[4], uf[4];
    draw4(g, bounce, STREAM_PATH, u); /* u[0] was the Russian-roulette coin; u[1] lobe, u[2] phi, u[3] theta */
    Ray r; V3 direction;
    if (m->Ni > 1) {
        double n1, n2, cos_in = vdot(vneg(wo), p->pn);
        V3 normal;
        if (cos_in > 0) { normal = vneg(p->pn); n1 = m->Ni; n2 = 1.0; } else { normal = p->pn; n1 = 1.0; n2 = m->Ni; }
        double rf0 = pow((n1 - n2) / (n1 + n2), 2);
        double fresnel = rf0 + (1.0f - rf0) * pow(1.0f - fabs(cos_in), 5);
        draw4(g, bounce, STREAM_FRESNEL, uf);
        if (fresnel < uf[0]) {
            if (refract_dir(vneg(wo), normal, n1 / n2, &direction)) { r.o = p->p; r.d = direction; *type = RT_TRANSMISSION; return r; }
            V3 inc = vneg(wo);
            r.o = p->p; r.d = vsub(inc, vmul(vmul(normal, vdot(inc, normal)), 2)); *type = RT_SPECULAR; return r;
        }
    }
    double kd_norm = vnorm(kd), ks_norm = vnorm(v3(m->ks[0], m->ks[1], m->ks[2]));
    if (ks_norm != 0 && kd_norm / ks_norm < u[1]) {
        V3 inc = vneg(wo);
        V3 reflect = vsub(inc, vmul(vmul(p->pn, vdot(inc, p->pn)), 2));
        direction = brdf_sample(reflect, RT_SPECULAR, m->Ns, u[2], u[3]);
        *type = RT_SPECULAR;
    } else {
        direction = brdf_sample(p->pn, RT_DIFFUSE, m->Ns, u[2], u[3]);
        *type = RT_DIFFUSE;
    }
    r.o = vadd(p->p, vmul(direction, 0.01)); r.d = direction;
    return r;
}

/* shade, MTPC/pathTracing.cpp:137-266 (recursive, as in the reference) */
static V3 shade(const orc_scene* S, const Isect* p, V3 wo, Rng* g, int bounce) {
    const Tri* f = &S->tri[p->k];
    if (S->mat_light[f->mat] >= 0) {                              /* :141-144 */
        const double* le = S->light[S->mat_light[f->mat]].radiance;
        return v3(le[0], le[1], le[2]);
    }
    const mcpt_material* m = &S->mat[f->mat];
    V3 kd;
    if (m->tex_id >= 0) {                                          /* :149-160, quirk Q5: u -> row, v -> column */
        V3 gc = bary_of(f, p->p);
        double row = f->uv[0] * gc.x + f->uv[2] * gc.y + f->uv[4] * gc.z;
        double col = f->uv[1] * gc.x + f->uv[3] * gc.y + f->uv[5] * gc.z;
        double irow = row - floor(row), icol = col - floor(col);
        int H = S->tex_h[m->tex_id], W = S->tex_w[m->tex_id];
        int r = (int)(irow * H), c = (int)(icol * W);
        if (r > H - 1) r = H - 1;
        if (c > W - 1) c = W - 1;
        const uint8_t* px = S->tex[m->tex_id] + ((size_t)r * W + c) * 3;
        kd = v3((double)px[2] / 255, (double)px[1] / 255, (double)px[0] / 255);
    } else kd = v3(m->kd[0], m->kd[1], m->kd[2]);

    V3 L_dir = v3(0, 0, 0);
    for (int i = 0; i < S->n_light; i++) {                        /* :171-232 */
        double u[4];
        draw4(g, bounce, STREAM_LIGHT0 + i, u);
        double total_area = S->light_area[i];
        double rnd = u[0] * S->light_area[0];                     /* quirk Q6: range frozen to light 0's area (:185) */
        int sel = -1;
        for (int j = 0; j < S->light_ntri[i]; j++) if (rnd < S->light_cdf[i][j]) { sel = j; break; }
        if (sel < 0) continue;                                     /* documented deviation (reference: NaN) */
        const Tri* sf = &S->obj_tri[S->light_tris[i][sel]];
        double sum = u[1] + u[2] + u[3];
        double p1 = u[1] / sum, p2 = u[2] / sum, p3 = u[3] / sum; /* quirk Q7 (:194-197) */
        V3 xl = vadd(vadd(vmul(sf->v1, p1), vmul(sf->v2, p2)), vmul(sf->v3, p3));
        V3 vn = vadd(vadd(vmul(sf->n1, p1), vmul(sf->n2, p2)), vmul(sf->n3, p3));
        V3 direction = vnormalize(vsub(xl, p->p));
        Ray rl; rl.o = vadd(p->p, vmul(direction, 0.01)); rl.d = direction;
        Isect inter;
        double visibility = 1;
        g->n_shadow++;
        if (g->flags & MCPT_FLAG_ANYHIT_SHADOW) {   /* quality extension (twin of the CUDA flag): visible iff nothing lies in between */
            double tmax = (vnorm(vsub(xl, p->p)) - 0.01) * 0.9999;
            if (ray_intersect(S, &rl, &inter) && inter.t < tmax) visibility = 0;
        } else
        if (!ray_intersect(S, &rl, &inter) || S->tri[inter.k].mat != sf->mat) visibility = 0; /* :212-215 closest hit, by material */
        if (vdot(direction, p->pn) > 0) {
            double pdf_light = (double)1 / total_area;
            double cos_theta = fabs(vdot(direction, vn) / vnorm(direction) / vnorm(vn));
            double cos_theta_hat = fabs(vdot(direction, p->pn) / vnorm(direction) / vnorm(p->pn));
            double dist = fmax(1.0, vnorm(vsub(xl, p->p)));
            const double* le = S->light[i].radiance;
            V3 inten = vmul(vdiv(vdiv(vmul(vmul(v3(le[0], le[1], le[2]), cos_theta), cos_theta_hat), pow(dist, 2)), pdf_light), visibility);
            double kd_dots = vdot(direction, p->pn);
            if (kd_dots > 0) {                                     /* quirk Q8: diffuse only, cosine twice */
                L_dir.x += kd.x * inten.x * kd_dots / ORC_PI;
                L_dir.y += kd.y * inten.y * kd_dots / ORC_PI;
                L_dir.z += kd.z * inten.z * kd_dots / ORC_PI;
            }
        }
    }
    V3 L_indir = v3(0, 0, 0);
    double u[4];
    draw4(g, bounce, STREAM_PATH, u);
    if (u[0] < ORC_P_RR && bounce + 1 < ORC_MAX_DEPTH) {         /* russian_Roulette, :3-11, :240 */
        int type;
        Ray r = next_ray(S, p, wo, kd, m, g, bounce, &type);
        Isect ret;
        g->n_extend++;
        if (ray_intersect(S, &r, &ret)) {
            V3 inten = vdiv(shade(S, &ret, vneg(r.d), g, bounce + 1), ORC_P_RR);
            if (type == RT_DIFFUSE) {
                if (S->mat_light[S->tri[ret.k].mat] < 0) { L_indir.x += kd.x * inten.x; L_indir.y += kd.y * inten.y; L_indir.z += kd.z * inten.z; }
            } else if (type == RT_SPECULAR) {
                L_indir.x += m->ks[0] * inten.x; L_indir.y += m->ks[1] * inten.y; L_indir.z += m->ks[2] * inten.z;
            } else L_indir = vadd(L_indir, inten);
        }
    }
    return vadd(L_dir, L_indir);
}


/* ---- MCPT_FLAG_MIS_BALANCE: FP64 twin of the CUDA integrator's physically based mode (csrc/shade.cuh) -------------------------
 * NOT the reference's estimator: Lambert + normalised Phong on the normalised shading normal turned toward the viewer, uniform
 * area sampling of every light + BSDF sampling combined with the balance heuristic, any-hit visibility.  Same Philox streams. */
enum { ORC_MIS_BSDF_ONLY = 64, ORC_MIS_NEE_ONLY = 128 };   /* test-only variants: each strategy alone (weights 1 / 0) */
typedef struct { V3 n, r, kd, ks; double pd, Ns; } MisFrame;
static double mis_lobe(const MisFrame* F, V3 wi) { double c = vdot(F->r, wi); return pow(c > 0 ? c : 0, F->Ns); }
static V3 mis_brdf(const MisFrame* F, double lobe) {
    double sp = (F->Ns + 2) * (0.5 / ORC_PI) * lobe;
    return v3(F->kd.x / ORC_PI + F->ks.x * sp, F->kd.y / ORC_PI + F->ks.y * sp, F->kd.z / ORC_PI + F->ks.z * sp);
}
static double mis_pdf_bsdf(const MisFrame* F, double cosi, double lobe) {
    return F->pd * cosi / ORC_PI + (1 - F->pd) * (F->Ns + 1) * (0.5 / ORC_PI) * lobe;
}
static V3 texel_or_kd(const orc_scene* S, const Tri* f, const mcpt_material* m, V3 p) {
    if (m->tex_id < 0) return v3(m->kd[0], m->kd[1], m->kd[2]);
    V3 gc = bary_of(f, p);
    double row = f->uv[0] * gc.x + f->uv[2] * gc.y + f->uv[4] * gc.z;
    double col = f->uv[1] * gc.x + f->uv[3] * gc.y + f->uv[5] * gc.z;
    double irow = row - floor(row), icol = col - floor(col);
    int H = S->tex_h[m->tex_id], W = S->tex_w[m->tex_id];
    int r = (int)(irow * H), c = (int)(icol * W);
    if (r > H - 1) r = H - 1;
    if (c > W - 1) c = W - 1;
    const uint8_t* px = S->tex[m->tex_id] + ((size_t)r * W + c) * 3;
    return v3((double)px[2] / 255, (double)px[1] / 255, (double)px[0] / 255);
}
/* radiance arriving along -wo at hit p; pdf_b = pdf of the BSDF sample that found p (< 0: camera ray or delta branch) */
static V3 shade_mis(const orc_scene* S, const Isect* p, Ray in, double pdf_b, Rng* g, int bounce) {
    const Tri* f = &S->tri[p->k];
    V3 wo = vneg(in.d);
    if (S->mat_light[f->mat] >= 0) {
        int li = S->mat_light[f->mat];
        const double* le = S->light[li].radiance;
        double w = 1;
        if (pdf_b >= 0) {
            V3 ng = vnormalize(vcross(vsub(f->v2, f->v1), vsub(f->v3, f->v1)));
            double cosl = fabs(vdot(ng, in.d));
            if (cosl > 1e-6) w = pdf_b / (pdf_b + p->t * p->t / (S->light_area[li] * cosl));
            if (g->flags & ORC_MIS_BSDF_ONLY) w = 1;   /* validation variants of the tests: one strategy alone must agree */
            if (g->flags & ORC_MIS_NEE_ONLY) w = 0;
        }
        return v3(le[0] * w, le[1] * w, le[2] * w);
    }
    const mcpt_material* m = &S->mat[f->mat];
    MisFrame F;
    F.kd = texel_or_kd(S, f, m, p->p);
    F.ks = v3(m->ks[0], m->ks[1], m->ks[2]); F.Ns = m->Ns;
    F.n = vnormalize(p->pn);
    if (vdot(F.n, wo) < 0) F.n = vneg(F.n);
    F.r = vsub(vmul(F.n, 2 * vdot(F.n, wo)), wo);
    double kdn = vnorm(F.kd), ksn = vnorm(F.ks);
    F.pd = (kdn + ksn) > 0 ? kdn / (kdn + ksn) : 1;
    V3 L = v3(0, 0, 0);
    for (int i = 0; i < S->n_light; i++) {
        double u[4];
        draw4(g, bounce, STREAM_LIGHT0 + i, u);
        double A = S->light_area[i], rnd = u[0] * A;
        int sel = S->light_ntri[i] - 1;
        for (int j = 0; j < S->light_ntri[i]; j++) if (rnd < S->light_cdf[i][j]) { sel = j; break; }
        const Tri* sf = &S->obj_tri[S->light_tris[i][sel]];
        double su = sqrt(u[1]);
        V3 xl = vadd(vadd(vmul(sf->v1, 1 - su), vmul(sf->v2, su * (1 - u[2]))), vmul(sf->v3, su * u[2]));
        V3 ng = vnormalize(vcross(vsub(sf->v2, sf->v1), vsub(sf->v3, sf->v1)));
        V3 dv = vsub(xl, p->p);
        double dist = vnorm(dv);
        V3 w = vdiv(dv, dist);
        double cosi = vdot(F.n, w), cosl = fabs(vdot(ng, w));
        if (!(cosi > 0) || !(cosl > 1e-6) || !(dist > 0)) continue;
        double lobe = mis_lobe(&F, w);
        double gg = cosi / (dist * dist / (A * cosl) + mis_pdf_bsdf(&F, cosi, lobe));
        if (g->flags & ORC_MIS_NEE_ONLY) gg = cosi / (dist * dist / (A * cosl));
        if (g->flags & ORC_MIS_BSDF_ONLY) continue;
        V3 fr = mis_brdf(&F, lobe);
        const double* le = S->light[i].radiance;
        V3 c = v3(fr.x * le[0] * gg, fr.y * le[1] * gg, fr.z * le[2] * gg);
        if (c.x == 0 && c.y == 0 && c.z == 0) continue;
        Ray rl; rl.o = vadd(p->p, vmul(w, 0.01)); rl.d = w;
        Isect inter;
        g->n_shadow++;
        if (ray_intersect(S, &rl, &inter) && inter.t < (dist - 0.01) * 0.9999) continue;   /* any-hit visibility */
        L = vadd(L, c);
    }
    double u[4];
    draw4(g, bounce, STREAM_PATH, u);
    if (u[0] < ORC_P_RR && bounce + 1 < ORC_MAX_DEPTH) {
        Ray r; V3 weight; double pb;
        if (m->Ni > 1) {   /* the reference's delta refraction branch, emission counted in full */
            int type;
            r = next_ray(S, p, wo, F.kd, m, g, bounce, &type);
            weight = type == RT_DIFFUSE ? F.kd : type == RT_SPECULAR ? F.ks : v3(1, 1, 1);
            pb = -1;
        } else {
            V3 d = u[1] < F.pd ? brdf_sample(F.n, RT_DIFFUSE, F.Ns, u[2], u[3]) : brdf_sample(F.r, RT_SPECULAR, F.Ns, u[2], u[3]);
            double cosi = vdot(F.n, d);
            if (!(cosi > 0)) return L;
            double lobe = mis_lobe(&F, d);
            pb = mis_pdf_bsdf(&F, cosi, lobe);
            if (!(pb > 0)) return L;
            V3 fr = mis_brdf(&F, lobe);
            weight = vmul(fr, cosi / pb);
            r.o = vadd(p->p, vmul(d, 0.01)); r.d = d;
        }
        Isect ret;
        g->n_extend++;
        if (ray_intersect(S, &r, &ret)) {
            V3 Li = vdiv(shade_mis(S, &ret, r, pb, g, bounce + 1), ORC_P_RR);
            L.x += weight.x * Li.x; L.y += weight.y * Li.y; L.z += weight.z * Li.z;
        }
    }
    return L;
}

/* generateImg, MTPC/pathTracing.cpp:274-331.  Samples [s0,s1) of spp are integrated; img gets their SUM / spp.
 * counts3 = {primary, shadow, extend} closest-hit queries. */
MCPT_API void orc_render_flags(const orc_scene* S, const mcpt_camera* cam, int width, int height, int spp, int s0, int s1,
                               uint64_t seed, int flags, double* img, double* counts3);
MCPT_API void orc_render(const orc_scene* S, const mcpt_camera* cam, int width, int height, int spp, int s0, int s1,
                         uint64_t seed, double* img, double* counts3) {
    orc_render_flags(S, cam, width, height, spp, s0, s1, seed, 0, img, counts3);
}
/* flags: MCPT_FLAG_JITTER / MCPT_FLAG_ANYHIT_SHADOW, the CPU twins of the CUDA integrator's quality extensions (include/mcpt.h). */
MCPT_API void orc_render_flags(const orc_scene* S, const mcpt_camera* cam, int width, int height, int spp, int s0, int s1,
                               uint64_t seed, int flags, double* img, double* counts3) {
    V3 up = vnormalize(v3(cam->up[0], cam->up[1], cam->up[2]));
    V3 eye = v3(cam->eye[0], cam->eye[1], cam->eye[2]), look = v3(cam->lookat[0], cam->lookat[1], cam->lookat[2]);
    if (width <= 0) width = cam->width;
    if (height <= 0) height = cam->height;
    V3 dir = vsub(look, eye);
    double l = vnorm(dir);
    double dy = tan(cam->fovy / 2 / 180 * ORC_PI) * l;
    double dx = dy / height * width;
    double pdx = 2 * dx / width, pdy = 2 * dy / height;
    V3 sx = vnormalize(vcross(dir, up)), sy = up;
    V3 screen_pdy = vmul(sy, pdy), screen_pdx = vmul(sx, pdx);
    V3 start = vadd(vsub(look, vmul(sx, dx)), vmul(up, dy));
    long np = 0, ns = 0, ne = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : np, ns, ne)
    for (int i = 0; i < height; i++) {
        V3 pos = vsub(start, vmul(screen_pdy, i));
        for (int j = 0; j < width; j++) {
            double acc[3] = {0, 0, 0};
            for (int k = s0; k < s1; k++) {
                Ray ray; ray.o = eye; ray.d = vnormalize(vsub(pos, eye));   /* quirk Q4: same ray for every sample */
                Rng g = {(uint32_t)(i * width + j), (uint32_t)k, (uint32_t)seed, (uint32_t)(seed >> 32), 0, 0, flags};
                if (flags & MCPT_FLAG_JITTER) {   /* through (j + u, i + v) of the pixel grid instead of its corner */
                    double u[4];
                    draw4(&g, 0, STREAM_JITTER, u);
                    V3 pj = vadd(vsub(start, vmul(screen_pdy, (double)i + u[1])), vmul(screen_pdx, (double)j + u[0]));
                    ray.d = vnormalize(vsub(pj, eye));
                }
                Isect ret;
                np++;
                if (ray_intersect(S, &ray, &ret)) {
                    V3 rad = (flags & MCPT_FLAG_MIS_BALANCE) ? shade_mis(S, &ret, ray, -1.0, &g, 0) : shade(S, &ret, vneg(ray.d), &g, 0);
                    acc[0] += rad.x / spp; acc[1] += rad.y / spp; acc[2] += rad.z / spp;
                    ns += g.n_shadow; ne += g.n_extend;
                }
            }
            double* o = img + ((size_t)i * width + j) * 3;
            o[0] = acc[0]; o[1] = acc[1]; o[2] = acc[2];
            pos = vadd(pos, screen_pdx);
        }
    }
    if (counts3) { counts3[0] = (double)np; counts3[1] = (double)ns; counts3[2] = (double)ne; }
}

/* TEST INFRASTRUCTURE ONLY -- CPU restatements (plain C) that the GPU path is checked against.
 * Never linked into or called from the product path (cuda_raytrace_b200/).
 *
 * Pinned against the reference itself: tests/test_oracle.py runs these functions and the reference's
 * own pbrt-v2 code (oracle/_ref/libpbrt_oracle.so, built from /root/reference) on the same inputs in
 * the build container and requires bit-identical (t, prim) -- and the resulting vectors are committed
 * under tests/golden/ so the check also runs where /root/reference does not exist.
 *
 * Must be compiled with -ffp-contract=off (no FMA contraction), x86-64 SSE float semantics.       */
#define _GNU_SOURCE
#include "oracle_port.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>

/* ---- (1) primitive tests ---------------------------------------------------------------------- */

/* pbrt Cross(): products and difference in double, result rounded to float
 * (pbrt-v2/src/core/geometry.h:477-484) */
static inline void cross_d(const float a[3], const float b[3], float out[3]) {
    double ax = a[0], ay = a[1], az = a[2], bx = b[0], by = b[1], bz = b[2];
    out[0] = (float)((ay * bz) - (az * by));
    out[1] = (float)((az * bx) - (ax * bz));
    out[2] = (float)((ax * by) - (ay * bx));
}
static inline float dot_f(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

/* Triangle::Intersect, pbrt-v2/src/shapes/trianglemesh.cpp:146-179 (the accept/reject part) */
static int tri_intersect(const float o[3], const float d[3], float tmin, float tmax, const float *p1,
                         const float *p2, const float *p3, float *t, float *b1o, float *b2o) {
    float e1[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    float e2[3] = {p3[0] - p1[0], p3[1] - p1[1], p3[2] - p1[2]};
    float s1[3];
    cross_d(d, e2, s1);
    float divisor = dot_f(s1, e1);
    if (divisor == 0.f) return 0;
    float invDivisor = 1.f / divisor;
    float dv[3] = {o[0] - p1[0], o[1] - p1[1], o[2] - p1[2]};
    float b1 = dot_f(dv, s1) * invDivisor;
    if (b1 < 0.f || b1 > 1.f) return 0;
    float s2[3];
    cross_d(dv, e1, s2);
    float b2 = dot_f(d, s2) * invDivisor;
    if (b2 < 0.f || b1 + b2 > 1.f) return 0;
    float tt = dot_f(e2, s2) * invDivisor;
    if (tt < tmin || tt > tmax) return 0;
    *t = tt; *b1o = b1; *b2o = b2;
    return 1;
}

/* Transform::operator()(Point) / (Vector), pbrt-v2/src/core/transform.h:192-220 (affine, w == 1) */
static inline void xf_point(const float *m, const float p[3], float out[3]) {
    out[0] = m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3];
    out[1] = m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7];
    out[2] = m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11];
}
static inline void xf_vec(const float *m, const float v[3], float out[3]) {
    out[0] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
    out[1] = m[4] * v[0] + m[5] * v[1] + m[6] * v[2];
    out[2] = m[8] * v[0] + m[9] * v[1] + m[10] * v[2];
}

/* Sphere::Intersect for a full sphere, pbrt-v2/src/shapes/sphere.cpp:58-83; Quadratic core/pbrt.h:308-322 */
/* phi > phimax without atan2f: csrc/b2rt_device.cuh phi_clipped, same operations */
void port_sincos_quarter(float q, float *s, float *c);
static int phi_clipped(float x, float y, float phimax) {
    float s, c;
    port_sincos_quarter(phimax * 0.63661977236758134308f, &s, &c);
    const float cr = c * y - s * x;
    return phimax <= 3.14159265358979323846f ? !(y >= 0.f && cr <= 0.f) : (y < 0.f && cr > 0.f);
}
static int sphere_intersect(const float o_[3], const float d_[3], float tmin, float tmax, const float *A28, float phimax,
                            float *t) {
    float o[3], d[3];
    xf_point(A28 + 12, o_, o);
    xf_vec(A28 + 12, d_, d);
    float radius = A28[24];
    float A = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
    float B = 2 * (d[0] * o[0] + d[1] * o[1] + d[2] * o[2]);
    float C = o[0] * o[0] + o[1] * o[1] + o[2] * o[2] - radius * radius;
    float discrim = B * B - 4.f * A * C;
    if (discrim < 0.) return 0;
    float rootDiscrim = sqrtf(discrim);
    float q;
    if (B < 0) q = -.5f * (B - rootDiscrim);
    else q = -.5f * (B + rootDiscrim);
    float t0 = q / A, t1 = C / q;
    if (t0 > t1) { float s = t0; t0 = t1; t1 = s; }
    if (t0 > tmax || t1 < tmin) return 0;
    float thit = t0;
    if (t0 < tmin) {
        thit = t1;
        if (thit > tmax) return 0;
    }
    /* partial sphere, shapes/sphere.cpp:84-106: record slots 25 / 26 hold zmin / zmax for a sphere (already clamped to
     * [-radius, radius]; zmin >= zmax = no z clipping).  phimax (radians, 0 = full turn) sits at the end of the record
     * when the caller passes the long form; the 28-float record has none. */
    float zmin = A28[25], zmax = A28[26];
    if (zmin < zmax || phimax > 0.f) {
        for (int pass = 0; pass < 2; ++pass) {
            float px = o[0] + d[0] * thit, py = o[1] + d[1] * thit, pz = o[2] + d[2] * thit;
            if (px == 0.f && py == 0.f) px = 1e-5f * radius;
            int c = (zmin < zmax) && ((zmin > -radius && pz < zmin) || (zmax < radius && pz > zmax));
            if (phimax > 0.f) c = c || phi_clipped(px, py, phimax);
            if (!c) break;
            if (pass == 1) return 0;
            if (thit == t1) return 0;
            if (t1 > tmax) return 0;
            thit = t1;
        }
    }
    *t = thit;
    return 1;
}

/* Disk::Intersect, pbrt-v2/src/shapes/disk.cpp:56-72 (phiMax = 2pi) */
static int disk_intersect(const float o_[3], const float d_[3], float tmin, float tmax, const float *A28, float phimax,
                          float *t) {
    float o[3], d[3];
    xf_point(A28 + 12, o_, o);
    xf_vec(A28 + 12, d_, d);
    float radius = A28[24], height = A28[25], inner = A28[26];
    if (fabsf(d[2]) < 1e-7) return 0;
    float thit = (height - o[2]) / d[2];
    if (thit < tmin || thit > tmax) return 0;
    float px = o[0] + d[0] * thit, py = o[1] + d[1] * thit;
    float dist2 = px * px + py * py;
    if (dist2 > radius * radius || dist2 < inner * inner) return 0;
    if (phimax > 0.f && phi_clipped(px, py, phimax)) return 0;
    *t = thit;
    return 1;
}

typedef struct {
    int nprims; const int32_t *kind; const float *verts; const int32_t *idx; const float *analytic;
    const float *rays; float *t_out; int32_t *prim_out; float *b1_out, *b2_out; uint8_t *occ;
    int64_t lo, hi;
} BruteJob;

static int prim_test(const BruteJob *J, int p, const float *r, float *t, float *b1, float *b2) {
    int k = J->kind[p];
    *b1 = 0.f; *b2 = 0.f;
    if (k >= 0) {
        const int32_t *v = J->idx + 3 * (int64_t)k;
        return tri_intersect(r, r + 4, r[3], r[7], J->verts + 3 * (int64_t)v[0], J->verts + 3 * (int64_t)v[1],
                             J->verts + 3 * (int64_t)v[2], t, b1, b2);
    }
    const float *A = J->analytic + PORT_ANALYTIC_STRIDE * (int64_t)(~k);
    if ((int)A[27] == 1) return sphere_intersect(r, r + 4, r[3], r[7], A, A[28], t);
    return disk_intersect(r, r + 4, r[3], r[7], A, A[28], t);
}

static void *brute_worker(void *arg) {
    BruteJob *J = (BruteJob *)arg;
    for (int64_t i = J->lo; i < J->hi; ++i) {
        const float *r = J->rays + 8 * i;
        float best = INFINITY, bb1 = 0, bb2 = 0; int bp = -1;
        for (int p = 0; p < J->nprims; ++p) {
            float t, b1, b2;
            if (prim_test(J, p, r, &t, &b1, &b2)) {
                if (J->occ) { bp = p; break; }
                if (t < best) { best = t; bp = p; bb1 = b1; bb2 = b2; }   /* ascending p: ties keep min id */
            }
        }
        if (J->occ) J->occ[i] = bp >= 0;
        else { J->t_out[i] = best; J->prim_out[i] = bp; J->b1_out[i] = bb1; J->b2_out[i] = bb2; }
    }
    return NULL;
}

static void run_brute(BruteJob proto, int64_t nrays, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 64) nthreads = 64;
    pthread_t th[64]; BruteJob jobs[64];
    int64_t chunk = (nrays + nthreads - 1) / nthreads;
    int started = 0;
    for (int t = 0; t < nthreads; ++t) {
        jobs[t] = proto; jobs[t].lo = t * chunk; jobs[t].hi = jobs[t].lo + chunk;
        if (jobs[t].hi > nrays) jobs[t].hi = nrays;
        if (jobs[t].lo >= jobs[t].hi) break;
        if (nthreads == 1) { brute_worker(&jobs[t]); return; }
        pthread_create(&th[t], NULL, brute_worker, &jobs[t]);
        ++started;
    }
    for (int t = 0; t < started; ++t) pthread_join(th[t], NULL);
}

void port_intersect_brute(int nprims, const int32_t *kind, const float *verts, const int32_t *idx,
                          const float *analytic, int64_t nrays, const float *rays, float *t_out,
                          int32_t *prim_out, float *b1_out, float *b2_out, int nthreads) {
    BruteJob J = {nprims, kind, verts, idx, analytic, rays, t_out, prim_out, b1_out, b2_out, NULL, 0, 0};
    run_brute(J, nrays, nthreads);
}

void port_occluded_brute(int nprims, const int32_t *kind, const float *verts, const int32_t *idx,
                         const float *analytic, int64_t nrays, const float *rays, uint8_t *occ, int nthreads) {
    BruteJob J = {nprims, kind, verts, idx, analytic, rays, NULL, NULL, NULL, NULL, occ, 0, 0};
    run_brute(J, nrays, nthreads);
}

/* ---- (3) stable LSD radix sort (restates csrc/sort.cu's contract: stable on the low key_bits) ----- */
void port_sort_pairs(int64_t n, uint32_t *keys, uint32_t *vals, int key_bits) {
    uint32_t *k2 = (uint32_t *)malloc((size_t)(n ? n : 1) * 4), *v2 = (uint32_t *)malloc((size_t)(n ? n : 1) * 4);
    int passes = (key_bits + 7) / 8;
    uint32_t *ki = keys, *vi = vals, *ko = k2, *vo = v2;
    for (int p = 0; p < passes; ++p) {
        int64_t cnt[257];
        memset(cnt, 0, sizeof(cnt));
        for (int64_t i = 0; i < n; ++i) cnt[((ki[i] >> (8 * p)) & 0xff) + 1]++;
        for (int d = 0; d < 256; ++d) cnt[d + 1] += cnt[d];
        for (int64_t i = 0; i < n; ++i) {
            int64_t pos = cnt[(ki[i] >> (8 * p)) & 0xff]++;
            ko[pos] = ki[i]; vo[pos] = vi[i];
        }
        uint32_t *t;
        t = ki; ki = ko; ko = t;
        t = vi; vi = vo; vo = t;
    }
    if (ki != keys) { memcpy(keys, ki, (size_t)n * 4); memcpy(vals, vi, (size_t)n * 4); }
    free(k2); free(v2);
}

/* ---- (2) LBVH builder restatement (mirrors cuda_raytrace_b200/csrc/build.cu step by step) ---------- */
void port_prim_boxes(int nprims, const int32_t *kind, const float *verts, const int32_t *idx, float *box) {
    for (int p = 0; p < nprims; ++p) {
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        int k = kind[p];
        if (k >= 0) {
            const int32_t *t = idx + 3 * (int64_t)k;
            for (int j = 0; j < 3; ++j) {
                const float *v = verts + 3 * (int64_t)t[j];
                for (int a = 0; a < 3; ++a) { lo[a] = fminf(lo[a], v[a]); hi[a] = fmaxf(hi[a], v[a]); }
            }
        }
        for (int a = 0; a < 3; ++a) { box[6 * (int64_t)p + a] = lo[a]; box[6 * (int64_t)p + 3 + a] = hi[a]; }
    }
}

static uint32_t expand_bits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
static int delta_k(const uint32_t *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + clz32((uint32_t)i ^ (uint32_t)j);
    return clz32(a ^ b);
}
static float box_area(const float *b) {
    float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
    return 2.f * (dx * dy + dy * dz + dz * dx);
}
#define LEAF_MAX 4
#define CHILD_EMPTY 0x7fffffff

typedef struct { int n; const int32_t *child; const float *bin_box; const uint8_t *collapse; } TreeView;
static int openable(const TreeView *T, int node) { return node < T->n - 1 && !T->collapse[node]; }

static void refit_rec(int node, int n, const int32_t *child, const int32_t (*range)[2], const float *leaf_box,
                      const int32_t *order, float *bin_box, float *cost, uint8_t *collapse) {
    /* iterative post-order to survive deep (degenerate) trees */
    int *stack = (int *)malloc(sizeof(int) * (size_t)(2 * n + 2));
    uint8_t *state = (uint8_t *)calloc((size_t)(2 * n), 1);
    int sp = 0;
    stack[sp++] = node;
    while (sp) {
        int cur = stack[sp - 1];
        if (cur >= n - 1) {
            const float *s = leaf_box + 6 * (int64_t)order[cur - (n - 1)];
            memcpy(bin_box + 6 * (int64_t)cur, s, 24);
            cost[cur] = box_area(s);
            --sp;
            continue;
        }
        if (!state[cur]) {
            state[cur] = 1;
            stack[sp++] = child[2 * cur + 1];
            stack[sp++] = child[2 * cur];
            continue;
        }
        --sp;
        int l = child[2 * cur], r = child[2 * cur + 1];
        const float *bl = bin_box + 6 * (int64_t)l, *br = bin_box + 6 * (int64_t)r;
        float *b = bin_box + 6 * (int64_t)cur;
        for (int a = 0; a < 3; ++a) { b[a] = fminf(bl[a], br[a]); b[3 + a] = fmaxf(bl[3 + a], br[3 + a]); }
        float area = box_area(b);
        float c_int = area + (cost[l] + cost[r]);
        int cnt = range[cur][1] - range[cur][0] + 1;
        float c_leaf = area * (float)cnt;
        int col = (cnt <= LEAF_MAX) && (c_leaf <= c_int);
        collapse[cur] = (uint8_t)col;
        cost[cur] = col ? c_leaf : c_int;
    }
    free(stack); free(state);
}

int port_lbvh_build(int n, const float *box, uint32_t *morton, int32_t *order, int32_t *parent,
                    int32_t *child, float *bin_box, uint8_t *collapse, uint32_t *wide_nodes) {
    if (n <= 0) return 0;
    /* K1: centroid bounds, Morton codes */
    float cmin[3] = {INFINITY, INFINITY, INFINITY}, cmax[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int p = 0; p < n; ++p)
        for (int a = 0; a < 3; ++a) {
            float c = 0.5f * (box[6 * (int64_t)p + a] + box[6 * (int64_t)p + 3 + a]);
            cmin[a] = fminf(cmin[a], c); cmax[a] = fmaxf(cmax[a], c);
        }
    uint32_t *vals = (uint32_t *)malloc((size_t)n * 4);
    for (int p = 0; p < n; ++p) {
        uint32_t q[3];
        for (int a = 0; a < 3; ++a) {
            float ext = cmax[a] - cmin[a];
            float scale = ext > 0.f ? 1024.f / ext : 0.f;
            float c = 0.5f * (box[6 * (int64_t)p + a] + box[6 * (int64_t)p + 3 + a]);
            float f = (c - cmin[a]) * scale;
            int qi = (int)f;
            q[a] = (uint32_t)(qi < 0 ? 0 : (qi > 1023 ? 1023 : qi));
        }
        morton[p] = (expand_bits10(q[0]) << 2) | (expand_bits10(q[1]) << 1) | expand_bits10(q[2]);
        vals[p] = (uint32_t)p;
    }
    /* K2 */
    port_sort_pairs(n, morton, vals, 30);
    for (int p = 0; p < n; ++p) order[p] = (int32_t)vals[p];
    free(vals);
    if (n == 1) { parent[0] = -1; memcpy(bin_box, box + 6 * (int64_t)order[0], 24); }
    int32_t (*range)[2] = (int32_t(*)[2])malloc(sizeof(int32_t) * 2 * (size_t)n);
    /* K3: Karras topology */
    for (int i = 0; i < n - 1; ++i) {
        int d = (delta_k(morton, n, i, i + 1) - delta_k(morton, n, i, i - 1)) >= 0 ? 1 : -1;
        int dmin = delta_k(morton, n, i, i - d);
        int lmax = 2;
        while (delta_k(morton, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int t = lmax >> 1; t >= 1; t >>= 1)
            if (delta_k(morton, n, i, i + (l + t) * d) > dmin) l += t;
        int j = i + l * d;
        int dnode = delta_k(morton, n, i, j);
        int s = 0;
        for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
            if (delta_k(morton, n, i, i + (s + t) * d) > dnode) s += t;
            if (t == 1) break;
        }
        int gamma = i + s * d + (d < 0 ? d : 0);
        int first = i < j ? i : j, last = i < j ? j : i;
        int left = (first == gamma) ? (n - 1) + gamma : gamma;
        int right = (last == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
        child[2 * i] = left; child[2 * i + 1] = right;
        parent[left] = i; parent[right] = i;
        range[i][0] = first; range[i][1] = last;
        if (i == 0) parent[0] = -1;
    }
    int nwide = 1;
    uint32_t *W = wide_nodes;
    float inf = INFINITY, ninf = -INFINITY;
#define WF(node, field, k) (*(float *)&W[32 * (size_t)(node) + 4 * (field) + (k)])
#define WI(node, field, k) (*(int32_t *)&W[32 * (size_t)(node) + 4 * (field) + (k)])
#define W_EMPTY(node, k) do { WF(node,0,k)=inf; WF(node,2,k)=inf; WF(node,4,k)=inf; WF(node,1,k)=ninf; \
        WF(node,3,k)=ninf; WF(node,5,k)=ninf; WI(node,6,k)=CHILD_EMPTY; } while (0)
    /* absolute margin on every stored child box: 2^-21 * largest |root-box coordinate| (build.cu node_box_pad) */
    float pad = 0.f;        /* set by W_PAD() once the root box bin_box[0..5] exists */
#define W_PAD() do { float m_ = 0.f; for (int a_ = 0; a_ < 6; ++a_) m_ = fmaxf(m_, fabsf(bin_box[a_])); \
        pad = m_ * 4.76837158203125e-07f; } while (0)
#define W_BOX(node, k, b) do { WF(node,0,k)=(b)[0]-pad; WF(node,2,k)=(b)[1]-pad; WF(node,4,k)=(b)[2]-pad; \
        WF(node,1,k)=(b)[3]+pad; WF(node,3,k)=(b)[4]+pad; WF(node,5,k)=(b)[5]+pad; } while (0)
    if (n == 1) {
        for (int k = 0; k < 4; ++k) { W_EMPTY(0, k); WI(0, 7, k) = 0; }
        W_PAD();
        W_BOX(0, 0, bin_box); WI(0, 6, 0) = (int32_t)(0x80000000u | 0u); WI(0, 7, 0) = 1;
        free(range);
        return 1;
    }
    float *cost = (float *)malloc(sizeof(float) * (size_t)(2 * n));
    memset(collapse, 0, (size_t)(n - 1));
    refit_rec(0, n, child, (const int32_t(*)[2])range, box, order, bin_box, cost, collapse);
    free(cost);
    W_PAD();
    if (collapse[0]) {
        for (int k = 0; k < 4; ++k) { W_EMPTY(0, k); WI(0, 7, k) = 0; }
        W_BOX(0, 0, bin_box); WI(0, 6, 0) = (int32_t)(0x80000000u | (uint32_t)(n - 1)); WI(0, 7, 0) = 1;
        free(range);
        return 1;
    }
    /* K4: greedy wide collapse; wide index = rank of the binary node id among all wide roots */
    TreeView T = {n, child, bin_box, collapse};
    int32_t (*tmp)[4] = (int32_t(*)[4])malloc(sizeof(int32_t) * 4 * (size_t)n);
    uint8_t *flag = (uint8_t *)calloc((size_t)n, 1);
    int *queue = (int *)malloc(sizeof(int) * (size_t)n);
    int qh = 0, qt = 0;
    queue[qt++] = 0;
    while (qh < qt) {
        int root = queue[qh++];
        int c[4] = {child[2 * root], child[2 * root + 1], -1, -1};
        int m = 2;
        while (m < 4) {
            int best = -1; float best_area = -1.f;
            for (int k = 0; k < m; ++k) {
                if (!openable(&T, c[k])) continue;
                float a = box_area(bin_box + 6 * (int64_t)c[k]);
                if (a > best_area) { best_area = a; best = k; }
            }
            if (best < 0) break;
            int node = c[best];
            for (int k = m; k > best + 1; --k) c[k] = c[k - 1];
            c[best] = child[2 * node]; c[best + 1] = child[2 * node + 1];
            ++m;
        }
        memcpy(tmp[root], c, sizeof(c));
        flag[root] = 1;
        for (int k = 0; k < m; ++k)
            if (openable(&T, c[k])) queue[qt++] = c[k];
    }
    int32_t *windex = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    nwide = 0;
    for (int i = 0; i < n - 1; ++i) { windex[i] = nwide; nwide += flag[i]; }
    for (int i = 0; i < n - 1; ++i) {
        if (!flag[i]) continue;
        int w = windex[i], used = 0;
        for (int k = 0; k < 4; ++k) {
            int c = tmp[i][k];
            if (c < 0) { W_EMPTY(w, k); continue; }
            ++used;
            W_BOX(w, k, bin_box + 6 * (int64_t)c);
            if (c >= n - 1) WI(w, 6, k) = (int32_t)(0x80000000u | ((uint32_t)(c - (n - 1)) << 3));
            else if (collapse[c])
                WI(w, 6, k) = (int32_t)(0x80000000u | ((uint32_t)range[c][0] << 3) | (uint32_t)(range[c][1] - range[c][0]));
            else WI(w, 6, k) = windex[c];
        }
        WI(w, 7, 0) = used; WI(w, 7, 1) = i; WI(w, 7, 2) = 0; WI(w, 7, 3) = 0;
    }
    free(tmp); free(flag); free(queue); free(windex); free(range);
    return nwide;
}

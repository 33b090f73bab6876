/* TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatements used as the parity oracle for the GPU path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this.
 * Build: make -C oracle port   (gcc -O2 -std=c11 -ffp-contract=off)                              */
#ifndef ORACLE_PORT_H
#define ORACLE_PORT_H
#include <stdint.h>

/* (1) pbrt primitive tests + the exhaustive loop of renderers/aggregatetest.cpp:93-102.
 * rays: 8 floats {o,tmin,d,tmax}.  prim kinds: kind[p] >= 0 -> triangle index, < 0 -> ~analytic index.
 * analytic record = 28 floats {o2w[12], w2o[12], radius, height, inner, type(1 sphere,2 disk)}.
 * Winner = (min t, then min prim id).  Outputs t (INF on miss), prim (-1), b1, b2.           */
void port_intersect_brute(int nprims, const int32_t *kind, const float *verts, const int32_t *idx,
                          const float *analytic, int64_t nrays, const float *rays, float *t_out,
                          int32_t *prim_out, float *b1_out, float *b2_out, int nthreads);
void port_occluded_brute(int nprims, const int32_t *kind, const float *verts, const int32_t *idx,
                         const float *analytic, int64_t nrays, const float *rays, uint8_t *occ, int nthreads);

/* (2) restatement of OUR deterministic LBVH builder (csrc/build.cu) for bit-exact topology checks.
 * Inputs: per-prim AABBs box[6n].  Outputs (caller-allocated):
 *   morton[n], order[n], parent[2n-1], child[2(n-1)], bin_box[6(2n-1)], collapse[n-1],
 *   wide nodes as raw 32-uint32 records (capacity n), returns number of wide nodes.        */
int port_lbvh_build(int n, const float *box, uint32_t *morton, int32_t *order, int32_t *parent,
                    int32_t *child, float *bin_box, uint8_t *collapse, uint32_t *wide_nodes);
void port_prim_boxes(int nprims, const int32_t *kind, const float *verts, const int32_t *idx, float *box);

/* (3) stable LSD radix sort restatement (std stable sort semantics) */
void port_sort_pairs(int64_t n, uint32_t *keys, uint32_t *vals, int key_bits);
#endif

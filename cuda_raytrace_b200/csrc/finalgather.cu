// Final gathering: pbrt's PhotonIntegrator with finalgather true, the integrator's default.
// The reference's GPU code has no equivalent (its finalGatheringCamera, photon_mapping/gathering.cu:129-146, averages
// the k-nearest estimate); what is followed here is pbrt-v2/src/integrators/photonmap.cpp:
//   K15 fg_photon_kernel      :440-490  direct photons + radiance photons (u < 0.125 per deposit) of a path range
//   K16 irradiance_kernel     :263-283  EPhoton from the k-nearest sets the gather kernel (K12) selects
//       radiance_lo_kernel    :585-621  ComputeRadianceTask: Lo = INV_PI * rho_r * (E_direct + E_indirect + E_caustic)
//   K17 radiance_lookup_kernel:156-172  nearest radiance photon with Dot(rp.n, n) > 0, unbounded radius (hash grid
//                                       shells instead of pbrt's kd-tree; exact, b2rt_fg.cuh)
//   K18 final_gather_kernel   :645-758  one warp per shading point: lane = gather ray (BSDF-sampled and photon-cone-sampled,
//                                       power heuristic), traversal + radiance-photon lookup inline, fixed-order butterfly sum
// Randomness is Philox(seed; pixel, sample | tree node, ray): an image tile is bit-identical on any GPU.
#include "b2rt_shade.cuh"
#include "b2rt_fg.cuh"
#include <algorithm>

namespace {

constexpr int FG_THREADS = 128;
constexpr int FG_NDIR = 50;              // nIndirSamplePhotons (photonmap.cpp:652)

// ---- K15 ---------------------------------------------------------------------------------------------------------------
// slots[(path - first) * (max_depth + 1) + j]: the path's direct photon (if any) first, then its radiance photons in
// bounce order.  Same emission / bounce / roulette stream as photon_trace_kernel (photon.cu), so the paths are the very
// paths whose caustic / indirect photons sit in maps 0 and 1; the radiance-photon draw is block 2 of the bounce counter,
// which the path itself never reads.
// 8 resident blocks per SM (64 registers) like photon_trace_regen_kernel: 2.45 -> 2.40 ms for 4 M paths (r08s)
#ifndef B2RT_FGPHOTON_BLOCKS
#define B2RT_FGPHOTON_BLOCKS 8
#endif
template <bool INST>
__global__ void __launch_bounds__(FG_THREADS, B2RT_FGPHOTON_BLOCKS)
fg_photon_kernel(DScene sc, uint32_t seed_lo, uint32_t seed_hi, int64_t first_path, int64_t n_paths, int max_depth,
                 b2rt_photon *__restrict__ slots, uint32_t *__restrict__ cnt_d, uint32_t *__restrict__ cnt_r, uint32_t *status,
                 int64_t lim_c, int64_t lim_i) {
    int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_paths) return;
    const uint64_t path = (uint64_t)(first_path + li);
    const uint32_t p_lo = (uint32_t)path, p_hi = (uint32_t)(path >> 32);
    uint32_t nd = 0, nr = 0;
    b2rt_photon *my = slots + li * (max_depth + 1);
    u4 r0 = philox4x32_10(p_lo, p_hi, 0u, 0u, seed_lo, seed_hi);
    u4 r1 = philox4x32_10(p_lo, p_hi, 0u, 1u, seed_lo, seed_hi);
    f3 o, d, alpha;
    float tmin;
    bool alive = emit_photon(sc, u01(r0.x), u01(r0.y), u01(r0.z), u01(r1.x), u01(r1.y), o, d, tmin, alpha);
    int n_isect = 0;
    bool overflow = false, specular_path = true;
    while (alive) {
        HitD h;
        if (!traverse<false, INST>(sc.nodes, sc.recs, sc.analytic, o.x, o.y, o.z, d.x, d.y, d.z, tmin, INFINITY, h, overflow))
            break;
        ++n_isect;
        Surface s = make_surface<INST>(sc, h, o, d);
        Mat m = load_material(sc, s.material);
        f3 wo = -d;
        // pbrt's depositedPhoton (photonmap.cpp:455-490): a caustic-type hit is a deposit only while the caustic map still
        // collects (path < lim_c), a direct / indirect one while the indirect map does (path < lim_i); b2rt_fg_deposit_limits
        const bool caustic_hit = specular_path && n_isect > 1;
        const bool deposited = caustic_hit ? ((int64_t)path < lim_c) : ((int64_t)path < lim_i);
        if (m.type == B2RT_MAT_MATTE && deposited) {     // hasNonSpecular: a photon is deposited here (direct, indirect or caustic)
            if (n_isect == 1) {
                float4 *dst = reinterpret_cast<float4 *>(my);
                dst[0] = make_float4(s.p.x, s.p.y, s.p.z, __uint_as_float(fg_oct_encode(wo)));
                dst[1] = make_float4(alpha.x, alpha.y, alpha.z, __uint_as_float(p_lo));
                nd = 1;
            }
            u4 rr = philox4x32_10(p_lo, p_hi, (uint32_t)n_isect, 2u, seed_lo, seed_hi);
            if (u01(rr.x) < .125f) {
                f3 n = dot3(s.ng, wo) < 0.f ? -s.ng : s.ng;                     // Faceforward(dg.nn, -photonRay.d)
                float4 *dst = reinterpret_cast<float4 *>(my + nd + nr);
                dst[0] = make_float4(s.p.x, s.p.y, s.p.z, __uint_as_float(fg_oct_encode(n)));
                dst[1] = make_float4(m.kd.x, m.kd.y, m.kd.z, __uint_as_float(p_lo));   // rho_r of the Lambertian lobe
                ++nr;
            }
        }
        if (n_isect >= max_depth) break;
        u4 rb = philox4x32_10(p_lo, p_hi, (uint32_t)n_isect, 0u, seed_lo, seed_hi);
        BsdfSample bs;
        if (!sample_bsdf(s, m, wo, u01(rb.x), u01(rb.y), u01(rb.z), bs)) break;
        if (is_black(bs.f) || bs.pdf == 0.f) break;
        f3 anew = mk3(alpha.x * bs.f.x * bs.absdot / bs.pdf, alpha.y * bs.f.y * bs.absdot / bs.pdf,
                      alpha.z * bs.f.z * bs.absdot / bs.pdf);
        float cont = fminf(1.f, lum(anew) / lum(alpha));
        if (u01(rb.w) > cont) break;
        alpha = mk3(anew.x / cont, anew.y / cont, anew.z / cont);
        specular_path = specular_path && bs.specular;
        o = s.p; d = bs.wi; tmin = s.ray_eps;
    }
    cnt_d[li] = nd;
    cnt_r[li] = nr;
    if (overflow) atomicAdd(status, 1u);
}

__global__ void __launch_bounds__(256)
fg_compact_kernel(const b2rt_photon *__restrict__ slots, const uint32_t *__restrict__ off_d, const uint32_t *__restrict__ off_r,
                  const uint32_t *__restrict__ tot, int64_t n_paths, int stride, const PhotonDst dst) {
    int64_t li = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n_paths) return;
    uint32_t od = off_d[li], orr = off_r[li];
    uint32_t nd = ((li + 1 < n_paths) ? off_d[li + 1] : tot[0]) - od;
    uint32_t nr = ((li + 1 < n_paths) ? off_r[li + 1] : tot[1]) - orr;
    const float4 *src = reinterpret_cast<const float4 *>(slots + li * stride);
    // like photon_compact_kernel: the slice goes to its place in the arrays of every GPU (peer stores over NVLink)
    for (uint32_t j = 0; j < nd; ++j) {
        const float4 a = src[2 * j], b = src[2 * j + 1];
        for (int g = 0; g < dst.n; ++g) {
            float4 *d = reinterpret_cast<float4 *>(dst.a[g] + od + j);
            d[0] = a; d[1] = b;
        }
    }
    for (uint32_t j = 0; j < nr; ++j) {
        const float4 a = src[2 * (nd + j)], b = src[2 * (nd + j) + 1];
        for (int g = 0; g < dst.n; ++g) {
            float4 *d = reinterpret_cast<float4 *>(dst.b[g] + orr + j);
            d[0] = a; d[1] = b;
        }
    }
}

// ---- K16 ---------------------------------------------------------------------------------------------------------------
// EPhoton over the set K12 selected (idx: k entries per query, ascending (d2, index), -1 padded): E = sum of alpha over
// the photons with Dot(n, wi) > 0, in that order, / (nPaths * md2 * M_PI) with pbrt's float * float * double product.
__global__ void __launch_bounds__(256)
irradiance_kernel(const b2rt_photon *__restrict__ sorted, const b2rt_query *__restrict__ queries, const int32_t *__restrict__ idx,
                  const float *__restrict__ r2, const int32_t *__restrict__ found, int64_t nq, int k, float n_paths,
                  float *__restrict__ E, int accumulate) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float4 *Q = reinterpret_cast<const float4 *>(queries + q);
    const float4 q0 = Q[0], q1 = Q[1];
    const f3 n = mk3(q0.w, q1.x, q1.y);
    f3 acc = mk3(0.f, 0.f, 0.f);
    const int nf = found[q];
    for (int j = 0; j < nf; ++j) {
        const int32_t i = idx[q * k + j];
        const float4 *pp = reinterpret_cast<const float4 *>(sorted + i);
        const float4 A = __ldg(pp);
        if (dot3(n, fg_oct_decode(__float_as_uint(A.w))) > 0.f) {
            const float4 B = __ldg(pp + 1);
            acc = acc + mk3(B.x, B.y, B.z);
        }
    }
    f3 e = mk3(0.f, 0.f, 0.f);
    if (nf > 0) {
        const float den = (float)((double)(n_paths * r2[q]) * 3.14159265358979323846);
        e = mk3(acc.x / den, acc.y / den, acc.z / den);
    }
    if (accumulate) { e.x = E[3 * q] + e.x; e.y = E[3 * q + 1] + e.y; e.z = E[3 * q + 2] + e.z; }
    E[3 * q] = e.x; E[3 * q + 1] = e.y; E[3 * q + 2] = e.z;
}

// radiance photon i -> the query (p, n) its irradiance is estimated at
__global__ void __launch_bounds__(256)
radiance_queries_kernel(const b2rt_photon *__restrict__ rp, int64_t n, b2rt_query *__restrict__ q) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 A = reinterpret_cast<const float4 *>(rp + i)[0];
    const f3 nn = fg_oct_decode(__float_as_uint(A.w));
    float4 *dst = reinterpret_cast<float4 *>(q + i);
    dst[0] = make_float4(A.x, A.y, A.z, nn.x);
    dst[1] = make_float4(nn.y, nn.z, 0.f, 0.f);
    dst[2] = make_float4(0.f, 0.f, 0.f, 0.f);
}
// rp.Lo += INV_PI * rho_r * E   (photonmap.cpp:608); rho_t is black for the Lambertian lobe
__global__ void __launch_bounds__(256)
radiance_lo_kernel(b2rt_photon *__restrict__ rp, const float *__restrict__ E, int64_t n) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float *a = rp[i].alpha;
    const bool blk = a[0] == 0.f && a[1] == 0.f && a[2] == 0.f;
    for (int c = 0; c < 3; ++c) a[c] = blk ? 0.f : (B2_INV_PI * a[c]) * E[3 * i + c];
}

// ---- K17 ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
radiance_lookup_kernel(FgGrid g, const float4 *__restrict__ pos, const float4 *__restrict__ photons, const uint32_t *__restrict__ cell_start,
                       const uint32_t *__restrict__ coarse1, const uint32_t *__restrict__ coarse2,
                       const float *__restrict__ nbox1, const float *__restrict__ nbox2,
                       const b2rt_query *__restrict__ queries, int64_t nq, float *__restrict__ out_rgb, int32_t *__restrict__ out_idx) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const float4 *Q = reinterpret_cast<const float4 *>(queries + q);
    const float4 q0 = Q[0], q1 = Q[1];
    float d2;
    const int i = fg_radiance_nearest(g, pos, cell_start, coarse1, coarse2, nbox1, nbox2, mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), d2);
    if (out_idx) out_idx[q] = i;
    if (out_rgb) {
        float4 B = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i >= 0) B = photons[2 * (size_t)i + 1];
        out_rgb[3 * q] = B.x; out_rgb[3 * q + 1] = B.y; out_rgb[3 * q + 2] = B.z;
    }
}

// ---- photon directions of the 50 nearest indirect photons: radius doubling for the queries that found fewer ------------
__global__ void __launch_bounds__(256)
fg_flag_kernel(const int32_t *__restrict__ found, int64_t nq, int need, uint32_t *__restrict__ flag) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q < nq) flag[q] = found[q] < need ? 1u : 0u;
}
__global__ void __launch_bounds__(256)
fg_pending_kernel(const b2rt_query *__restrict__ queries, const int32_t *__restrict__ found, const uint32_t *__restrict__ off,
                  int64_t nq, int need, b2rt_query *__restrict__ pq, uint32_t *__restrict__ pidx) {
    int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq || found[q] >= need) return;
    const uint32_t j = off[q];
    const float4 *src = reinterpret_cast<const float4 *>(queries + q);
    float4 *dst = reinterpret_cast<float4 *>(pq + j);
    dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
    pidx[j] = (uint32_t)q;
}
__global__ void __launch_bounds__(256)
fg_scatter_kernel(const uint32_t *__restrict__ pidx, const int32_t *__restrict__ tidx, const int32_t *__restrict__ tfound,
                  int64_t np, int k, int32_t *__restrict__ idx, int32_t *__restrict__ found) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= np * k) return;
    const int64_t j = t / k;
    const int e = (int)(t % k);
    const uint32_t q = pidx[j];
    idx[(int64_t)q * k + e] = tidx[t];
    if (e == 0) found[q] = tfound[j];
}

// ---- K18 ---------------------------------------------------------------------------------------------------------------
struct FgArgs {
    DScene sc;
    FgGrid g;
    const float4 *rad;                 // sorted radiance photons, float4 pairs (the winner's Lo is read from here)
    const float4 *rad_pos;             // what the lookup scans: rad itself (g.pstride 2) or the packed copy (g.pstride 1)
    const uint32_t *rad_cells, *rad_coarse1, *rad_coarse2;   // cell ranges and the two occupancy levels (4^3, 16^3 cells)
    const float *rad_nbox1, *rad_nbox2;                      // normal bounds of the two levels
    const b2rt_photon *ind;            // sorted indirect photons (directions for the cone sampling)
    const int32_t *idx;                // [nq * FG_NDIR] nearest indirect photons per query
    const int32_t *found;
    const b2rt_query *queries;
    int64_t nq;
    int gs;                            // gather rays per strategy (pbrt: gatherSamples / 2)
    float cos_angle, cone_pdf;
    uint32_t k0, k1;
    float *out_rgb;
    uint32_t *status;
    // split pipeline (default): per gather ray of the chunk
    float4 *ray_recs;                  // [2 * nrays] {hit point, -} {facing normal of the hit, -}
    uint32_t *ray_keys, *ray_vals;     // radiance-grid cell of the hit (n_cells = no hit) / ray index
    int32_t *ray_ri;                   // nearest facing radiance photon of the ray's hit, -1 = none / no hit
    uint32_t n_cells;
};

// Gather ray r of a shading point (photonmap.cpp:662-687 BSDF-sampled, :708-733 cone-sampled): lanes / rays [0, gs) sample
// the Lambertian lobe, [gs, 2 gs) pick one of the nd photon directions and sample its cone.  Pure function of the query's
// Philox words and r, so the trace and the combine launches of the split pipeline regenerate the very same direction.
struct FgRay { f3 wi; float cosw; bool act, cone; };
template <class DirOf>
__device__ __forceinline__ FgRay fg_make_ray(uint32_t gpix, uint32_t kc, int r, int gs, int nd, f3 nf, f3 sn, f3 tn,
                                             float cos_angle, uint32_t k0, uint32_t k1, DirOf dir_of) {
    FgRay R;
    R.act = r < 2 * gs;
    R.cone = r >= gs;
    R.wi = mk3(0.f, 0.f, 1.f);
    R.cosw = 0.f;
    if (R.act) {
        const u4 u = philox4x32_10(gpix, kc, 0x200u + (uint32_t)r, 0u, k0, k1);
        if (!R.cone) {
            // Lambertian::Sample_f: cosine hemisphere on wo's side (= nf's side), pdf = |cos| / pi
            const f3 l = cosine_sample_hemisphere(u01(u.x), u01(u.y));
            R.wi = mk3(sn.x * l.x + tn.x * l.y + nf.x * l.z, sn.y * l.x + tn.y * l.y + nf.y * l.z,
                       sn.z * l.x + tn.z * l.y + nf.z * l.z);
            R.cosw = dot3(R.wi, nf);
            R.act = (l.z > 0.f) && (R.cosw > 0.f);
        } else if (nd == 0) {
            R.act = false;
        } else {
            const int pn = min(nd - 1, (int)floorf(u01(u.z) * (float)nd));
            const f3 z = dir_of(pn);
            f3 vx, vy;
            fg_coordinate_system(z, vx, vy);
            R.wi = fg_sample_cone(u01(u.x), u01(u.y), cos_angle, vx, vy, z);
            R.cosw = dot3(R.wi, nf);
            R.act = R.cosw > 0.f;                                        // else bsdf->f(wo, wi) is black
        }
    }
    return R;
}
// MIS-weighted contribution of one gather ray whose hit found the radiance photon with outgoing radiance Lo
// (photonmap.cpp:689-706 / :735-754)
__device__ __forceinline__ void fg_accumulate(const FgRay &R, f3 fr, float4 Lo, const f3 *dirs, int nd, int gs, float cos_angle,
                                              float cone_pdf, f3 &accB, f3 &accC) {
    const float bsdf_pdf = fabsf(R.cosw) * B2_INV_PI;
    const float photon_pdf = fg_photon_pdf(dirs, nd, R.wi, cos_angle, cone_pdf);
    const float absdot = fabsf(R.cosw);
    if (!R.cone) {
        const float wt = fg_power_heuristic(gs, bsdf_pdf, gs, photon_pdf);
        const float w = absdot * wt / bsdf_pdf;
        accB = accB + mk3(fr.x * Lo.x * w, fr.y * Lo.y * w, fr.z * Lo.z * w);
    } else {
        const float wt = fg_power_heuristic(gs, photon_pdf, gs, bsdf_pdf);
        accC = accC + mk3(fr.x * Lo.x * absdot * wt / photon_pdf, fr.y * Lo.y * absdot * wt / photon_pdf,
                          fr.z * Lo.z * absdot * wt / photon_pdf);
    }
}
__device__ __forceinline__ void fg_warp_sum_store(f3 accB, f3 accC, int lane, int gs, float *out) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        accB.x += __shfl_xor_sync(0xffffffffu, accB.x, o); accB.y += __shfl_xor_sync(0xffffffffu, accB.y, o);
        accB.z += __shfl_xor_sync(0xffffffffu, accB.z, o);
        accC.x += __shfl_xor_sync(0xffffffffu, accC.x, o); accC.y += __shfl_xor_sync(0xffffffffu, accC.y, o);
        accC.z += __shfl_xor_sync(0xffffffffu, accC.z, o);
    }
    if (lane == 0) {
        const float g = (float)gs;
        out[0] = accB.x / g + accC.x / g;
        out[1] = accB.y / g + accC.y / g;
        out[2] = accB.z / g + accC.z / g;
    }
}

// B2RT_FG_VARIANT=0, the first version, kept as the A/B reference of the split pipeline below (identical bytes): one warp
// per shading point, lane = gather ray, traversal and the radiance-photon lookup inline.  A warp's 32 lanes walk 32
// unrelated lookups with their own trip counts: 2.9 of 32 threads active per instruction (ncu r03g), 344 ms for 960 x 540.
template <bool INST>
__global__ void __launch_bounds__(FG_THREADS)
final_gather_kernel(const FgArgs A) {
    __shared__ f3 s_dirs[FG_THREADS / 32][FG_NDIR];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t wstride = (int64_t)gridDim.x * (FG_THREADS / 32);
    bool overflow = false;
    for (int64_t q = (int64_t)blockIdx.x * (FG_THREADS / 32) + warp; q < A.nq; q += wstride) {
        const float4 *Q = reinterpret_cast<const float4 *>(A.queries + q);
        const float4 q0 = __ldg(Q), q1 = __ldg(Q + 1), q2 = __ldg(Q + 2);
        const f3 p = mk3(q0.x, q0.y, q0.z), nf = mk3(q0.w, q1.x, q1.y), kd = mk3(q1.z, q1.w, q2.x);
        const float ray_eps = q2.y;
        const uint32_t gpix = __float_as_uint(q2.z), kc = __float_as_uint(q2.w);
        const int nd = min(A.found[q], FG_NDIR);
        __syncwarp();
        for (int j = lane; j < nd; j += 32) {
            const int32_t i = A.idx[q * FG_NDIR + j];
            s_dirs[warp][j] = fg_oct_decode(__float_as_uint(__ldg(reinterpret_cast<const float4 *>(A.ind + i)).w));
        }
        __syncwarp();
        f3 sn, tn;
        fg_coordinate_system(nf, sn, tn);
        const f3 fr = kd * B2_INV_PI;                                    // Lambertian f, throughput folded into kd
        f3 accB = mk3(0.f, 0.f, 0.f), accC = mk3(0.f, 0.f, 0.f);
        const f3 *dirs = s_dirs[warp];
        for (int base = 0; base < 2 * A.gs; base += 32) {
            const FgRay R = fg_make_ray(__float_as_uint(q2.z), kc, base + lane, A.gs, nd, nf, sn, tn, A.cos_angle, A.k0, A.k1,
                                        [dirs](int pn) { return dirs[pn]; });
            (void)gpix;
            if (!R.act) continue;
            HitD h;
            if (!traverse<false, INST>(A.sc.nodes, A.sc.recs, A.sc.analytic, p.x, p.y, p.z, R.wi.x, R.wi.y, R.wi.z, ray_eps, INFINITY, h,
                                 overflow))
                continue;
            const Surface s = make_surface<INST>(A.sc, h, p, R.wi);
            const f3 ng = dot3(s.ng, -R.wi) < 0.f ? -s.ng : s.ng;          // Faceforward(nGather, -bounceRay.d)
            float d2;
            const int ri = fg_radiance_nearest(A.g, A.rad_pos, A.rad_cells, A.rad_coarse1, A.rad_coarse2, A.rad_nbox1, A.rad_nbox2,
                                               s.p, ng, d2);
            if (ri >= 0)
                fg_accumulate(R, fr, __ldg(A.rad + 2 * (size_t)ri + 1), dirs, nd, A.gs, A.cos_angle, A.cone_pdf, accB, accC);
        }
        __syncwarp();
        fg_warp_sum_store(accB, accC, lane, A.gs, A.out_rgb + 3 * q);
    }
    if (overflow) atomicAdd(A.status, 1u);
}

// ---- the split pipeline (default): trace -> sort the hits by radiance-grid cell -> lookup -> combine -------------------------
// What bounded the fused kernel was not memory but divergent issue: the lanes of a warp were the gather rays of one shading
// point, their hits land all over the scene, and every lane ran its own shell walk.  Here the lookups are a launch of their
// own over the hits SORTED BY THE CELL THEY FALL IN (K2): the lanes of a warp start in the same or neighbouring cells, walk
// the same rows with the same trip counts and read the same photons (L1).  Every ray's arithmetic is unchanged, and the
// combine launch sums in the same butterfly order, so the output is bit-identical to the fused kernel (tested).
//
// launch 1: one thread per gather ray (ray i of the chunk = query i / R, ray i % R): direction, traversal, hit record
// (ptxas' own 64 registers = 8 blocks per SM; 10 blocks measured slower, r08s)
template <bool INST>
__global__ void __launch_bounds__(FG_THREADS)
fg_trace_kernel(const FgArgs A, int64_t nrays, int R) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrays) return;
    const int64_t q = i / R;
    const int r = (int)(i - q * R);
    const float4 *Q = reinterpret_cast<const float4 *>(A.queries + q);
    const float4 q0 = __ldg(Q), q1 = __ldg(Q + 1), q2 = __ldg(Q + 2);
    const f3 p = mk3(q0.x, q0.y, q0.z), nf = mk3(q0.w, q1.x, q1.y);
    const int nd = min(A.found[q], FG_NDIR);
    f3 sn, tn;
    fg_coordinate_system(nf, sn, tn);
    const int32_t *qidx = A.idx + q * FG_NDIR;
    const b2rt_photon *ind = A.ind;
    const FgRay Ry = fg_make_ray(__float_as_uint(q2.z), __float_as_uint(q2.w), r, A.gs, nd, nf, sn, tn, A.cos_angle, A.k0, A.k1,
                                 [qidx, ind](int pn) {
                                     return fg_oct_decode(__float_as_uint(__ldg(reinterpret_cast<const float4 *>(ind + __ldg(qidx + pn))).w));
                                 });
    uint32_t key = A.n_cells;
    bool overflow = false;
    if (Ry.act) {
        HitD h;
        if (traverse<false, INST>(A.sc.nodes, A.sc.recs, A.sc.analytic, p.x, p.y, p.z, Ry.wi.x, Ry.wi.y, Ry.wi.z, q2.y, INFINITY, h, overflow)) {
            const Surface s = make_surface<INST>(A.sc, h, p, Ry.wi);
            const f3 ng = dot3(s.ng, -Ry.wi) < 0.f ? -s.ng : s.ng;          // Faceforward(nGather, -bounceRay.d)
            A.ray_recs[2 * i] = make_float4(s.p.x, s.p.y, s.p.z, 0.f);
            A.ray_recs[2 * i + 1] = make_float4(ng.x, ng.y, ng.z, 0.f);
            // the fine cell the lookup starts from (fg_radiance_phase0 clamps the same way)
            const float pa[3] = {s.p.x, s.p.y, s.p.z};
            int c[3];
            for (int a = 0; a < 3; ++a) {
                float f = floorf((pa[a] - A.g.lo[a]) * A.g.inv);
                f = fminf(fmaxf(f, 0.f), (float)(A.g.dims[a] - 1));
                c[a] = (int)f;
            }
            key = ((uint32_t)c[2] * (uint32_t)A.g.dims[1] + (uint32_t)c[1]) * (uint32_t)A.g.dims[0] + (uint32_t)c[0];
        }
    }
    A.ray_keys[i] = key;
    A.ray_vals[i] = (uint32_t)i;
    A.ray_ri[i] = -1;
    if (overflow) atomicAdd(A.status, 1u);
}
// launch 2: thread j = the j-th hit in cell order
// resident blocks per SM: 12 (40 registers) instead of ptxas' own 64 registers -- 960 x 540 final-gather frame 23.5 -> 22.9 ms;
// 6 / 10 blocks: 24.3 / 23.0 ms (r07v)
#ifndef B2RT_FGLOOKUP_BLOCKS
#define B2RT_FGLOOKUP_BLOCKS 12
#endif
__global__ void __launch_bounds__(FG_THREADS, B2RT_FGLOOKUP_BLOCKS)
fg_lookup_kernel(const FgArgs A, const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, int64_t nrays) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nrays || keys[j] >= A.n_cells) return;       // misses sort to the end
    const uint32_t i = vals[j];
    const float4 P = A.ray_recs[2 * (size_t)i], N = A.ray_recs[2 * (size_t)i + 1];
    float d2;
    A.ray_ri[i] = fg_radiance_nearest(A.g, A.rad_pos, A.rad_cells, A.rad_coarse1, A.rad_coarse2, A.rad_nbox1, A.rad_nbox2,
                                      mk3(P.x, P.y, P.z), mk3(N.x, N.y, N.z), d2);
}
// launch 3: one warp per shading point, lane = gather ray, in the fused kernel's order
__global__ void __launch_bounds__(FG_THREADS)
fg_combine_kernel(const FgArgs A) {
    __shared__ f3 s_dirs[FG_THREADS / 32][FG_NDIR];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t wstride = (int64_t)gridDim.x * (FG_THREADS / 32);
    const int R = 2 * A.gs;
    for (int64_t q = (int64_t)blockIdx.x * (FG_THREADS / 32) + warp; q < A.nq; q += wstride) {
        const float4 *Q = reinterpret_cast<const float4 *>(A.queries + q);
        const float4 q0 = __ldg(Q), q1 = __ldg(Q + 1), q2 = __ldg(Q + 2);
        const f3 nf = mk3(q0.w, q1.x, q1.y), kd = mk3(q1.z, q1.w, q2.x);
        const int nd = min(A.found[q], FG_NDIR);
        __syncwarp();
        for (int j = lane; j < nd; j += 32) {
            const int32_t i = A.idx[q * FG_NDIR + j];
            s_dirs[warp][j] = fg_oct_decode(__float_as_uint(__ldg(reinterpret_cast<const float4 *>(A.ind + i)).w));
        }
        __syncwarp();
        f3 sn, tn;
        fg_coordinate_system(nf, sn, tn);
        const f3 fr = kd * B2_INV_PI;
        f3 accB = mk3(0.f, 0.f, 0.f), accC = mk3(0.f, 0.f, 0.f);
        const f3 *dirs = s_dirs[warp];
        for (int base = 0; base < R; base += 32) {
            const int r = base + lane;
            const int ri = r < R ? A.ray_ri[q * R + r] : -1;
            if (ri < 0) continue;
            const FgRay Ry = fg_make_ray(__float_as_uint(q2.z), __float_as_uint(q2.w), r, A.gs, nd, nf, sn, tn, A.cos_angle, A.k0,
                                         A.k1, [dirs](int pn) { return dirs[pn]; });
            fg_accumulate(Ry, fr, __ldg(A.rad + 2 * (size_t)ri + 1), dirs, nd, A.gs, A.cos_angle, A.cone_pdf, accB, accC);
        }
        __syncwarp();
        fg_warp_sum_store(accB, accC, lane, A.gs, A.out_rgb + 3 * q);
    }
}
// packed copies of the sorted radiance photons for the lookups: {p, oct normal} and the decoded normal
__global__ void __launch_bounds__(256)
fg_pack_kernel(const float4 *__restrict__ sorted, int64_t n, float4 *__restrict__ pos, float4 *__restrict__ nrm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 A = sorted[2 * i];
    const f3 m = fg_oct_decode(__float_as_uint(A.w));
    pos[i] = A;
    nrm[i] = make_float4(m.x, m.y, m.z, 0.f);
}

// search tables of the radiance-photon grid: normal and position bounds per level-1 cell (4^3 fine cells), then count + normal bounds
// per level-2 cell (16^3 fine cells); arithmetic in b2rt_fg.cuh
__global__ void __launch_bounds__(128)
fg_nbox1_kernel(FgGrid g, const float4 *__restrict__ photons, const uint32_t *__restrict__ cell_start, float *__restrict__ nbox1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.cdims[0] * g.cdims[1] * g.cdims[2]) return;
    fg_nbox_level1(g, photons, cell_start, i % g.cdims[0], (i / g.cdims[0]) % g.cdims[1], i / (g.cdims[0] * g.cdims[1]),
                   nbox1 + 12 * (size_t)i);
}
__global__ void __launch_bounds__(128)
fg_level2_kernel(FgGrid g, const float *__restrict__ nbox1, const uint32_t *__restrict__ coarse1, float *__restrict__ nbox2,
                 uint32_t *__restrict__ coarse2) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= g.c2dims[0] * g.c2dims[1] * g.c2dims[2]) return;
    fg_nbox_level2(g, nbox1, coarse1, i % g.c2dims[0], (i / g.c2dims[0]) % g.c2dims[1], i / (g.c2dims[0] * g.c2dims[1]),
                   nbox2 + 12 * (size_t)i, coarse2 + i);
}

FgGrid fg_grid_of(const PhotonMap &m) {
    FgGrid g;
    for (int a = 0; a < 3; ++a) { g.cdims[a] = m.cdims[a]; g.c2dims[a] = (m.cdims[a] + 3) / 4; }
    float mx = 0.f;
    for (int a = 0; a < 3; ++a) {
        g.lo[a] = m.grid.origin[a]; g.dims[a] = m.grid.dims[a];
        float hi = m.grid.origin[a] + (float)m.grid.dims[a] * m.grid.cell_size;
        mx = fmaxf(mx, fmaxf(fabsf(m.grid.origin[a]), fabsf(hi)));
    }
    g.cell = m.grid.cell_size; g.inv = m.grid.inv_cell;
    g.slack = 1e-5f * mx + 1e-4f * m.grid.cell_size;
    g.pstride = 2; g.nrm = nullptr;
    return g;
}

}  // namespace

// The lookup's view of the radiance-photon grid: packed positions + decoded normals (fg_pack_kernel) and the search tables
// of the two coarse levels.  Built once per grid build (grid_epoch), not per call: a frame runs the gather chunk by chunk.
// On return g describes the packed layout and *pos points at the array the lookups scan.
static int fg_prepare_radiance(b2rt_ctx *ctx, FgGrid &g, const float4 **pos) {
    PhotonMap &mr = ctx->maps[B2RT_MAP_RADIANCE];
    const int n1 = g.cdims[0] * g.cdims[1] * g.cdims[2], n2 = g.c2dims[0] * g.c2dims[1] * g.c2dims[2];
    const bool fresh = ctx->fg_tables_epoch != mr.grid_epoch || mr.grid_epoch == 0;
    if (fresh) {
        B2RT_CUDA(ctx, ctx->d_fg_rpos.reserve((size_t)std::max<int64_t>(mr.n, 1) * 16));
        B2RT_CUDA(ctx, ctx->d_fg_rnrm.reserve((size_t)std::max<int64_t>(mr.n, 1) * 16));
        B2RT_CUDA(ctx, ctx->d_fg_nbox1.reserve((size_t)n1 * 48));
        B2RT_CUDA(ctx, ctx->d_fg_nbox2.reserve((size_t)n2 * 48));
        B2RT_CUDA(ctx, ctx->d_fg_coarse2.reserve((size_t)n2 * 4));
    }
    g.pstride = 1; g.nrm = ctx->d_fg_rnrm.as<float4>();
    *pos = ctx->d_fg_rpos.as<float4>();
    if (!fresh) return B2RT_OK;
    if (mr.n > 0) {
        fg_pack_kernel<<<(unsigned)((mr.n + 255) / 256), 256, 0, ctx->stream>>>(mr.sorted.as<float4>(), mr.n, ctx->d_fg_rpos.as<float4>(),
                                                                                ctx->d_fg_rnrm.as<float4>());
        B2RT_LAUNCH_CHECK(ctx);
    }
    fg_nbox1_kernel<<<(n1 + 127) / 128, 128, 0, ctx->stream>>>(g, *pos, mr.cell_start.as<uint32_t>(), ctx->d_fg_nbox1.as<float>());
    B2RT_LAUNCH_CHECK(ctx);
    fg_level2_kernel<<<(n2 + 127) / 128, 128, 0, ctx->stream>>>(g, ctx->d_fg_nbox1.as<float>(), mr.coarse.as<uint32_t>(),
                                                                ctx->d_fg_nbox2.as<float>(), ctx->d_fg_coarse2.as<uint32_t>());
    B2RT_LAUNCH_CHECK(ctx);
    ctx->fg_tables_epoch = mr.grid_epoch;
    return B2RT_OK;
}

int b2rt_photons_trace_fg_begin(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, uint32_t tot[2]) {
    cudaStream_t st = ctx->stream;
    tot[0] = tot[1] = 0;
    if (n_paths == 0) return B2RT_OK;
    if (n_paths > (1ll << 28)) { ctx->err = "photons_trace_fg: at most 2^28 paths per call"; return B2RT_E_ARG; }
    if (ctx->radiance_computed) { ctx->err = "photons_trace_fg: radiance already computed; clear the photon maps first"; return B2RT_E_STATE; }
    DScene sc;
    b2rt_fill_dscene(ctx, &sc);
    const int stride = max_depth + 1;
    B2RT_CUDA(ctx, ctx->d_fg_slots.reserve((size_t)n_paths * stride * sizeof(b2rt_photon)));
    B2RT_CUDA(ctx, ctx->d_fg_cnt.reserve((size_t)n_paths * 4));
    B2RT_CUDA(ctx, ctx->d_fg_cnt2.reserve((size_t)n_paths * 4));
    uint32_t *cnt_d = ctx->d_fg_cnt.as<uint32_t>(), *cnt_r = ctx->d_fg_cnt2.as<uint32_t>();
    uint32_t *counters = ctx->d_counters.as<uint32_t>();
    (sc.instances ? fg_photon_kernel<true> : fg_photon_kernel<false>)<<<(unsigned)((n_paths + FG_THREADS - 1) / FG_THREADS), FG_THREADS, 0, st>>>(
        sc, (uint32_t)ctx->seed, (uint32_t)(ctx->seed >> 32), first_path, n_paths, max_depth,
        ctx->d_fg_slots.as<b2rt_photon>(), cnt_d, cnt_r, counters + 8, ctx->fg_lim_c, ctx->fg_lim_i);
    B2RT_LAUNCH_CHECK(ctx);
    int rc = b2rt_exclusive_scan_u32(ctx, cnt_d, n_paths, counters + 32);
    if (rc) return rc;
    rc = b2rt_exclusive_scan_u32(ctx, cnt_r, n_paths, counters + 33);
    if (rc) return rc;
    B2RT_CUDA(ctx, cudaMemcpyAsync(tot, counters + 32, 8, cudaMemcpyDeviceToHost, st));
    B2RT_CUDA(ctx, cudaStreamSynchronize(st));
    return B2RT_OK;
}
int b2rt_photons_trace_fg_commit(b2rt_ctx *ctx, int64_t n_paths, int max_depth, const PhotonDst &dst) {
    if (n_paths == 0) return B2RT_OK;
    fg_compact_kernel<<<(unsigned)((n_paths + 255) / 256), 256, 0, ctx->stream>>>(
        ctx->d_fg_slots.as<b2rt_photon>(), ctx->d_fg_cnt.as<uint32_t>(), ctx->d_fg_cnt2.as<uint32_t>(),
        ctx->d_counters.as<uint32_t>() + 32, n_paths, max_depth + 1, dst);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

int b2rt_photons_trace_fg_impl(b2rt_ctx *ctx, int64_t first_path, int64_t n_paths, int max_depth, int64_t n_stored[2]) {
    if (n_stored) n_stored[0] = n_stored[1] = 0;
    if (n_paths == 0) return B2RT_OK;
    uint32_t tot[2];
    int rc = b2rt_photons_trace_fg_begin(ctx, first_path, n_paths, max_depth, tot);
    if (rc) return rc;
    PhotonMap &md = ctx->maps[B2RT_MAP_DIRECT], &mr = ctx->maps[B2RT_MAP_RADIANCE];
    rc = b2rt_photons_reserve(ctx, md.photons, md.n, md.n + tot[0]);
    if (rc) return rc;
    rc = b2rt_photons_reserve(ctx, mr.photons, mr.n, mr.n + tot[1]);
    if (rc) return rc;
    PhotonDst dst;
    dst.n = 1; dst.a[0] = md.photons.as<b2rt_photon>() + md.n; dst.b[0] = mr.photons.as<b2rt_photon>() + mr.n;
    rc = b2rt_photons_trace_fg_commit(ctx, n_paths, max_depth, dst);
    if (rc) return rc;
    md.n += tot[0]; mr.n += tot[1];
    // nDirectPaths only counts the paths shot while the indirect map was collecting (photonmap.cpp:546-553)
    md.n_paths += std::max<int64_t>(0, std::min<int64_t>(first_path + n_paths, ctx->fg_lim_i) - first_path);
    mr.n_paths += n_paths;
    md.grid_built = mr.grid_built = false;
    if (n_stored) { n_stored[0] = tot[0]; n_stored[1] = tot[1]; }
    return B2RT_OK;
}

// E (or E += with accumulate) for nq device queries against `map`; chunks keep the k-nearest index buffer bounded
int b2rt_irradiance_impl(b2rt_ctx *ctx, int map, int64_t nq, const b2rt_query *d_queries, int k, float max_dist2,
                         float *d_E, bool accumulate) {
    PhotonMap &m = ctx->maps[map];
    const int64_t chunk = 1 << 20;
    const int64_t cq = std::min<int64_t>(nq, chunk);
    B2RT_CUDA(ctx, ctx->d_fg_idx.reserve((size_t)cq * k * 4));
    B2RT_CUDA(ctx, ctx->d_fg_r2.reserve((size_t)cq * 4));
    B2RT_CUDA(ctx, ctx->d_fg_found.reserve((size_t)cq * 4));
    for (int64_t q0 = 0; q0 < nq; q0 += chunk) {
        const int64_t n = std::min<int64_t>(chunk, nq - q0);
        int rc = b2rt_gather_impl(ctx, map, n, d_queries + q0, k, max_dist2, nullptr, ctx->d_fg_r2.as<float>(),
                                  ctx->d_fg_found.as<int32_t>(), ctx->d_fg_idx.as<int32_t>());
        if (rc) return rc;
        irradiance_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(
            m.sorted.as<b2rt_photon>(), d_queries + q0, ctx->d_fg_idx.as<int32_t>(), ctx->d_fg_r2.as<float>(),
            ctx->d_fg_found.as<int32_t>(), n, k, (float)m.n_paths, d_E + 3 * q0, accumulate ? 1 : 0);
        B2RT_LAUNCH_CHECK(ctx);
    }
    return B2RT_OK;
}

int b2rt_radiance_compute_impl(b2rt_ctx *ctx, int nused, float max_dist2) {
    PhotonMap &mr = ctx->maps[B2RT_MAP_RADIANCE];
    if (ctx->radiance_computed) { ctx->err = "radiance_compute: already computed for these radiance photons"; return B2RT_E_STATE; }
    const int64_t n = mr.n;
    if (n > 0) {
        B2RT_CUDA(ctx, ctx->d_fg_queries.reserve((size_t)n * sizeof(b2rt_query)));
        B2RT_CUDA(ctx, ctx->d_fg_E.reserve((size_t)n * 12));
        B2RT_CUDA(ctx, cudaMemsetAsync(ctx->d_fg_E.p, 0, (size_t)n * 12, ctx->stream));
        radiance_queries_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(mr.photons.as<b2rt_photon>(), n,
                                                                                        ctx->d_fg_queries.as<b2rt_query>());
        B2RT_LAUNCH_CHECK(ctx);
        // E = EPhoton(direct) + EPhoton(indirect) + EPhoton(caustic), in pbrt's order (photonmap.cpp:600-605)
        const int order[3] = {B2RT_MAP_DIRECT, B2RT_MAP_INDIRECT, B2RT_MAP_CAUSTIC};
        for (int t = 0; t < 3; ++t) {
            PhotonMap &m = ctx->maps[order[t]];
            if (m.n == 0) continue;
            if (!m.grid_built) { ctx->err = "radiance_compute: a non-empty photon map has no grid (b2rt_photon_grid_build)"; return B2RT_E_STATE; }
            int rc = b2rt_irradiance_impl(ctx, order[t], n, ctx->d_fg_queries.as<b2rt_query>(), nused, max_dist2,
                                          ctx->d_fg_E.as<float>(), true);
            if (rc) return rc;
        }
        radiance_lo_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(mr.photons.as<b2rt_photon>(),
                                                                                   ctx->d_fg_E.as<float>(), n);
        B2RT_LAUNCH_CHECK(ctx);
    }
    ctx->radiance_computed = true;
    mr.grid_built = false;
    if (n == 0) return B2RT_OK;          // no radiance photons: the gather has nothing to look up, the indirect term is zero
    return b2rt_photon_grid_build_impl(ctx, B2RT_MAP_RADIANCE, 0.f, 0.f);
}

int b2rt_radiance_lookup_impl(b2rt_ctx *ctx, int64_t nq, const b2rt_query *d_queries, float *d_rgb, int32_t *d_idx) {
    PhotonMap &mr = ctx->maps[B2RT_MAP_RADIANCE];
    if (nq <= 0) return B2RT_OK;
    FgGrid g = fg_grid_of(mr);
    const float4 *pos = nullptr;
    int rc = fg_prepare_radiance(ctx, g, &pos);
    if (rc) return rc;
    radiance_lookup_kernel<<<(unsigned)((nq + 127) / 128), 128, 0, ctx->stream>>>(
        g, pos, mr.sorted.as<float4>(), mr.cell_start.as<uint32_t>(), mr.coarse.as<uint32_t>(), ctx->d_fg_coarse2.as<uint32_t>(),
        ctx->d_fg_nbox1.as<float>(), ctx->d_fg_nbox2.as<float>(), d_queries, nq, d_rgb, d_idx);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

// the indirect term of the camera pass with final gathering: d_rgb[3 * nq]
int b2rt_final_gather_impl(b2rt_ctx *ctx, int64_t nq, const b2rt_query *d_queries, float max_dist2, float *d_rgb) {
    if (nq <= 0) return B2RT_OK;
    cudaStream_t st = ctx->stream;
    PhotonMap &mi = ctx->maps[B2RT_MAP_INDIRECT], &mr = ctx->maps[B2RT_MAP_RADIANCE];
    const int k = FG_NDIR;
    const int need = (int)std::min<int64_t>(k, mi.n);
    B2RT_CUDA(ctx, ctx->d_fg_dirs.reserve((size_t)nq * k * 4));
    B2RT_CUDA(ctx, ctx->d_fg_found.reserve((size_t)nq * 4));
    B2RT_CUDA(ctx, ctx->d_fg_r2.reserve((size_t)nq * 4));
    int32_t *idx = ctx->d_fg_dirs.as<int32_t>(), *found = ctx->d_fg_found.as<int32_t>();
    int rc = b2rt_gather_impl(ctx, B2RT_MAP_INDIRECT, nq, d_queries, k, max_dist2, nullptr, ctx->d_fg_r2.as<float>(), found, idx);
    if (rc) return rc;
    // pbrt doubles the squared search radius until 50 photons are found (photonmap.cpp:655-660); only the queries that
    // are still short are searched again
    uint32_t *counters = ctx->d_counters.as<uint32_t>();
    float md2 = max_dist2;
    for (int it = 0; it < 40 && need > 0; ++it) {
        B2RT_CUDA(ctx, ctx->d_fg_flag.reserve((size_t)nq * 4));
        uint32_t *flag = ctx->d_fg_flag.as<uint32_t>();
        fg_flag_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(found, nq, need, flag);
        B2RT_LAUNCH_CHECK(ctx);
        rc = b2rt_exclusive_scan_u32(ctx, flag, nq, counters + 34);
        if (rc) return rc;
        uint32_t np = 0;
        B2RT_CUDA(ctx, cudaMemcpyAsync(&np, counters + 34, 4, cudaMemcpyDeviceToHost, st));
        B2RT_CUDA(ctx, cudaStreamSynchronize(st));
        if (np == 0) break;
        md2 *= 2.f;
        B2RT_CUDA(ctx, ctx->d_fg_pq.reserve((size_t)np * sizeof(b2rt_query)));
        B2RT_CUDA(ctx, ctx->d_fg_pidx.reserve((size_t)np * 4));
        B2RT_CUDA(ctx, ctx->d_fg_idx.reserve((size_t)np * k * 4));
        B2RT_CUDA(ctx, ctx->d_fg_pend.reserve((size_t)np * 4));
        fg_pending_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, st>>>(d_queries, found, flag, nq, need,
                                                                         ctx->d_fg_pq.as<b2rt_query>(), ctx->d_fg_pidx.as<uint32_t>());
        B2RT_LAUNCH_CHECK(ctx);
        rc = b2rt_gather_impl(ctx, B2RT_MAP_INDIRECT, np, ctx->d_fg_pq.as<b2rt_query>(), k, md2, nullptr, ctx->d_fg_r2.as<float>(),
                              ctx->d_fg_pend.as<int32_t>(), ctx->d_fg_idx.as<int32_t>());
        if (rc) return rc;
        fg_scatter_kernel<<<(unsigned)(((int64_t)np * k + 255) / 256), 256, 0, st>>>(
            ctx->d_fg_pidx.as<uint32_t>(), ctx->d_fg_idx.as<int32_t>(), ctx->d_fg_pend.as<int32_t>(), np, k, idx, found);
        B2RT_LAUNCH_CHECK(ctx);
    }
    FgArgs A;
    b2rt_fill_dscene(ctx, &A.sc);
    A.g = fg_grid_of(mr);
    rc = fg_prepare_radiance(ctx, A.g, &A.rad_pos);
    if (rc) return rc;
    A.rad = mr.sorted.as<float4>(); A.rad_cells = mr.cell_start.as<uint32_t>();
    A.rad_coarse1 = mr.coarse.as<uint32_t>(); A.rad_coarse2 = ctx->d_fg_coarse2.as<uint32_t>();
    A.rad_nbox1 = ctx->d_fg_nbox1.as<float>(); A.rad_nbox2 = ctx->d_fg_nbox2.as<float>();
    A.ind = mi.sorted.as<b2rt_photon>();
    A.gs = std::max(1, ctx->fg_samples / 2);
    A.cos_angle = ctx->fg_cos_angle; A.cone_pdf = fg_cone_pdf(ctx->fg_cos_angle);
    A.k0 = (uint32_t)ctx->seed ^ 0x46475448u; A.k1 = (uint32_t)(ctx->seed >> 32);      // 'FGTH' stream
    A.status = counters + 8;
    A.ray_recs = nullptr; A.ray_keys = A.ray_vals = nullptr; A.ray_ri = nullptr;
    A.n_cells = (uint32_t)mr.grid.n_cells;
    const int64_t cap = (int64_t)ctx->sm_count * 8;
    if (ctx->fg_variant == 0) {
        A.idx = idx; A.found = found; A.queries = d_queries; A.nq = nq; A.out_rgb = d_rgb;
        int64_t blocks = std::min(cap, (nq + FG_THREADS / 32 - 1) / (FG_THREADS / 32));
        (A.sc.instances ? final_gather_kernel<true> : final_gather_kernel<false>)<<<(unsigned)blocks, FG_THREADS, 0, st>>>(A);
        B2RT_LAUNCH_CHECK(ctx);
        return B2RT_OK;
    }
    // split pipeline, in chunks of at most 2^26 gather rays (2 GB of hit records + 1.3 GB of sort buffers and results)
    const int R = 2 * A.gs;
    const int64_t cq = std::max<int64_t>(1, std::min<int64_t>(nq, (1ll << 26) / R));
    const int64_t cr = cq * R;
    B2RT_CUDA(ctx, ctx->d_fg_rayrec.reserve((size_t)cr * 32));
    B2RT_CUDA(ctx, ctx->d_fg_rkeys.reserve((size_t)cr * 4));
    B2RT_CUDA(ctx, ctx->d_fg_rvals.reserve((size_t)cr * 4));
    B2RT_CUDA(ctx, ctx->d_fg_rkeys2.reserve((size_t)cr * 4));
    B2RT_CUDA(ctx, ctx->d_fg_rvals2.reserve((size_t)cr * 4));
    B2RT_CUDA(ctx, ctx->d_fg_ri.reserve((size_t)cr * 4));
    A.ray_recs = ctx->d_fg_rayrec.as<float4>(); A.ray_keys = ctx->d_fg_rkeys.as<uint32_t>();
    A.ray_vals = ctx->d_fg_rvals.as<uint32_t>(); A.ray_ri = ctx->d_fg_ri.as<int32_t>();
    int key_bits = 1;
    while (key_bits < 32 && (1ull << key_bits) <= (uint64_t)A.n_cells) ++key_bits;     // keys 0 .. n_cells inclusive
    for (int64_t q0 = 0; q0 < nq; q0 += cq) {
        const int64_t n = std::min<int64_t>(cq, nq - q0), nrays = n * R;
        A.idx = idx + q0 * k; A.found = found + q0; A.queries = d_queries + q0; A.nq = n; A.out_rgb = d_rgb + 3 * q0;
        (A.sc.instances ? fg_trace_kernel<true> : fg_trace_kernel<false>)<<<(unsigned)((nrays + FG_THREADS - 1) / FG_THREADS), FG_THREADS, 0, st>>>(A, nrays, R);
        B2RT_LAUNCH_CHECK(ctx);
        bool in_alt = false;
        rc = b2rt_sort_pairs_impl(ctx, nrays, A.ray_keys, A.ray_vals, ctx->d_fg_rkeys2.as<uint32_t>(), ctx->d_fg_rvals2.as<uint32_t>(),
                                  key_bits, &in_alt);
        if (rc) return rc;
        const uint32_t *sk = in_alt ? ctx->d_fg_rkeys2.as<uint32_t>() : A.ray_keys;
        const uint32_t *sv = in_alt ? ctx->d_fg_rvals2.as<uint32_t>() : A.ray_vals;
        fg_lookup_kernel<<<(unsigned)((nrays + FG_THREADS - 1) / FG_THREADS), FG_THREADS, 0, st>>>(A, sk, sv, nrays);
        B2RT_LAUNCH_CHECK(ctx);
        int64_t blocks = std::min(cap, (n + FG_THREADS / 32 - 1) / (FG_THREADS / 32));
        fg_combine_kernel<<<(unsigned)blocks, FG_THREADS, 0, st>>>(A);
        B2RT_LAUNCH_CHECK(ctx);
    }
    return B2RT_OK;
}

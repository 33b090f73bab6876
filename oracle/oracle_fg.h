/* TEST INFRASTRUCTURE ONLY (checker for tests/, smoke() and bench.py's cpu_baseline leg; never on the product path).
 *
 * CPU restatement of pbrt-v2's final gathering for the GPU kernels of csrc/finalgather.cu.  Textually included at the
 * end of oracle_photon.c (it uses that file's static helpers: exhaustive closest hit, BSDF sampling, Philox, the
 * deterministic sin/cos).  Each function cites the pbrt code it follows (pbrt-v2/src/integrators/photonmap.cpp).
 *
 * Pin: port_radiance_lookup and port_irradiance are checked against the reference's own KdTree<RadiancePhoton>::Lookup +
 * RadiancePhotonProcess and EPhoton through oracle/_ref (ref_photon.cpp: ref_radiance_lookup / ref_ephoton).  The
 * gather itself (port_final_gather) draws its rays from Philox instead of pbrt's per-pixel low-discrepancy samples, so
 * it is pinned statistically: images against oracle/_ref/pbrt with finalgather true (tests/test_gpu_zfinalgather.py).
 */

/* photonmap.cpp:440-490 with finalgather true: the direct photon (first diffuse hit) of every path and a radiance photon
 * (p, Faceforward(nn, wo), rho_r) at every deposit that draws u < .125f.  The draw is Philox block 2 of the bounce
 * counter: the path's own stream (blocks 0, 1) is untouched, so the paths are those of port_photon_trace. */
/* deposit limits (pbrt's depositedPhoton, photonmap.cpp:455-490): a caustic-type hit only counts as a deposit while the
 * caustic map is still collecting (path < lim_c), a direct / indirect one while the indirect map is (path < lim_i) */
static int64_t g_fg_lim_c = INT64_MAX, g_fg_lim_i = INT64_MAX;
void port_fg_deposit_limits(int64_t lim_c, int64_t lim_i) { g_fg_lim_c = lim_c; g_fg_lim_i = lim_i; }
int port_photon_trace_fg(const port_scene *sc, uint64_t seed, int64_t first_path, int64_t n_paths, int max_depth,
                         port_photon *out_d, int64_t cap_d, int64_t *n_d, port_photon *out_r, int64_t cap_r,
                         int64_t *n_r) {
    olight L[64]; float cdf[130];
    if (sc->nlights < 1 || sc->nlights > 64) return -2;
    derive_lights(sc, L, cdf);
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    int64_t nd = *n_d, nr = *n_r;
    for (int64_t li = 0; li < n_paths; ++li) {
        uint64_t path = (uint64_t)(first_path + li);
        uint32_t p_lo = (uint32_t)path, p_hi = (uint32_t)(path >> 32);
        uint32_t r0[4], r1[4];
        philox(p_lo, p_hi, 0u, 0u, k0, k1, r0);
        philox(p_lo, p_hi, 0u, 1u, k0, k1, r1);
        float u0 = u01(r0[0]);
        int lsel = 0;
        while (lsel + 1 < sc->nlights && u0 >= cdf[lsel + 1]) ++lsel;
        float light_pdf = cdf[sc->nlights + 1 + lsel];
        const olight *Ls = &L[lsel];
        v3 o, d, Nl, Le; float tmin, pdf;
        o_emit(sc, Ls, r0, r1, &o, &d, &Nl, &Le, &tmin, &pdf);
        if (pdf == 0.f || black(Le)) continue;
        float ad = fabsf(vdot(Nl, d)), den = pdf * light_pdf;
        v3 alpha = V(ad * Le.x / den, ad * Le.y / den, ad * Le.z / den);
        if (black(alpha)) continue;
        int n_isect = 0, specular_path = 1;
        for (;;) {
            osurf s; int prim;
            if (!closest(sc, o, d, tmin, &s, &prim)) break;
            ++n_isect;
            omat m = get_material(sc, prim);
            v3 wo = vneg(d);
            const int caustic_hit = specular_path && n_isect > 1;
            const int deposited = caustic_hit ? ((int64_t)path < g_fg_lim_c) : ((int64_t)path < g_fg_lim_i);
            if (m.type == 0 && deposited) {
                if (n_isect == 1) {
                    port_photon ph;
                    ph.p[0] = s.p.x; ph.p[1] = s.p.y; ph.p[2] = s.p.z; ph.wi_oct = oct_encode(wo);
                    ph.alpha[0] = alpha.x; ph.alpha[1] = alpha.y; ph.alpha[2] = alpha.z; ph.path = p_lo;
                    if (nd >= cap_d) return -1;
                    out_d[nd++] = ph;
                }
                uint32_t rr[4];
                philox(p_lo, p_hi, (uint32_t)n_isect, 2u, k0, k1, rr);
                if (u01(rr[0]) < .125f) {
                    v3 n = vdot(s.ng, wo) < 0.f ? vneg(s.ng) : s.ng;
                    port_photon ph;
                    ph.p[0] = s.p.x; ph.p[1] = s.p.y; ph.p[2] = s.p.z; ph.wi_oct = oct_encode(n);
                    ph.alpha[0] = m.kd.x; ph.alpha[1] = m.kd.y; ph.alpha[2] = m.kd.z; ph.path = p_lo;
                    if (nr >= cap_r) return -1;
                    out_r[nr++] = ph;
                }
            }
            if (n_isect >= max_depth) break;
            uint32_t rb[4];
            philox(p_lo, p_hi, (uint32_t)n_isect, 0u, k0, k1, rb);
            osample bs;
            if (!sample_bsdf(&s, &m, wo, u01(rb[0]), u01(rb[1]), u01(rb[2]), &bs)) break;
            if (black(bs.f) || bs.pdf == 0.f) break;
            v3 anew = V(alpha.x * bs.f.x * bs.absdot / bs.pdf, alpha.y * bs.f.y * bs.absdot / bs.pdf,
                        alpha.z * bs.f.z * bs.absdot / bs.pdf);
            float cont = fminf(1.f, lumin(anew) / lumin(alpha));
            if (u01(rb[3]) > cont) break;
            alpha = V(anew.x / cont, anew.y / cont, anew.z / cont);
            specular_path = specular_path && bs.specular;
            o = s.p; d = bs.wi; tmin = s.ray_eps;
        }
    }
    *n_d = nd; *n_r = nr;
    return 0;
}

/* EPhoton (photonmap.cpp:263-283): brute-force k nearest by (d2, index) within max_dist2; the sum runs over them in that
 * order; E / (count * md2 * M_PI) with pbrt's float * float -> double product rounded to float */
void port_irradiance(int64_t n, const port_photon *sorted, int64_t n_paths, int64_t nq, const float *queries, int k,
                     float max_dist2, float *E) {
    uint64_t *keys = (uint64_t *)malloc((size_t)(n ? n : 1) * 8);
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        v3 nn = V(Q[3], Q[4], Q[5]);
        int64_t m = 0;
        for (int64_t i = 0; i < n; ++i) {
            float dx = sorted[i].p[0] - Q[0], dy = sorted[i].p[1] - Q[1], dz = sorted[i].p[2] - Q[2];
            float d2 = dx * dx + dy * dy + dz * dz;
            if (d2 < max_dist2) { uint32_t b; memcpy(&b, &d2, 4); keys[m++] = ((uint64_t)b << 32) | (uint32_t)i; }
        }
        qsort(keys, (size_t)m, 8, cmp_u64);
        int64_t cnt = m > k ? k : m;
        float md2 = max_dist2;
        if (cnt == k) { uint32_t b = (uint32_t)(keys[k - 1] >> 32); memcpy(&md2, &b, 4); }
        v3 acc = V(0, 0, 0);
        for (int64_t j = 0; j < cnt; ++j) {
            uint32_t i = (uint32_t)keys[j];
            if (vdot(nn, oct_decode(sorted[i].wi_oct)) > 0.f)
                acc = vadd(acc, V(sorted[i].alpha[0], sorted[i].alpha[1], sorted[i].alpha[2]));
        }
        float *o = E + 3 * q;
        o[0] = o[1] = o[2] = 0.f;
        if (cnt > 0) {
            float den = (float)((double)((float)n_paths * md2) * 3.14159265358979323846);
            o[0] = acc.x / den; o[1] = acc.y / den; o[2] = acc.z / den;
        }
    }
    free(keys);
}

/* RadiancePhotonProcess (photonmap.cpp:156-172) with md2 = INFINITY: nearest photon with Dot(rp.n, n) > 0, brute force;
 * ties to the lower index.  The photon's normal is the decoded octahedral word, as on the device. */
static int64_t radiance_nearest(int64_t n, const port_photon *rp, v3 p, v3 nn) {
    int64_t best = -1;
    float bd2 = INFINITY;
    for (int64_t i = 0; i < n; ++i) {
        float dx = rp[i].p[0] - p.x, dy = rp[i].p[1] - p.y, dz = rp[i].p[2] - p.z;
        float d2 = dx * dx + dy * dy + dz * dz;
        if (d2 < bd2 && vdot(oct_decode(rp[i].wi_oct), nn) > 0.f) { best = i; bd2 = d2; }
    }
    return best;
}
void port_radiance_lookup(int64_t n, const port_photon *rp, int64_t nq, const float *queries, float *rgb, int32_t *idx) {
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        int64_t b = radiance_nearest(n, rp, V(Q[0], Q[1], Q[2]), V(Q[3], Q[4], Q[5]));
        if (idx) idx[q] = (int32_t)b;
        if (rgb) for (int c = 0; c < 3; ++c) rgb[3 * q + c] = b >= 0 ? rp[b].alpha[c] : 0.f;
    }
}

/* The gather (photonmap.cpp:645-758) for a Lambertian shading point, one query = 12 floats {p, Nf, rho * throughput,
 * ray epsilon, bits(pixel id), bits(sample | node << 22)}: directions of the 50 nearest indirect photons (pbrt's radius
 * doubling ends with exactly that set), gs BSDF-sampled + gs cone-sampled gather rays, exhaustive closest hit, nearest
 * radiance photon facing the hit, power heuristic.  Ray r of a query accumulates into partial sum r % 32; the 32 partial
 * sums are combined by an xor butterfly -- the summation order of the device kernel (one warp per query). */
void port_final_gather(const port_scene *sc, int64_t n_ind, const port_photon *ind, int64_t n_rad, const port_photon *rad,
                       int64_t nq, const float *queries, int gs, float cos_angle, uint64_t seed, float *rgb) {
    const uint32_t k0 = (uint32_t)seed ^ 0x46475448u, k1 = (uint32_t)(seed >> 32);
    const float cone_pdf = (float)(1.0 / ((double)(2.f) * 3.14159265358979323846 * (double)(1.f - cos_angle)));
    const float thr = .999f * cos_angle;
    uint64_t *keys = (uint64_t *)malloc((size_t)(n_ind ? n_ind : 1) * 8);
    for (int64_t q = 0; q < nq; ++q) {
        const float *Q = queries + 12 * q;
        const v3 p = V(Q[0], Q[1], Q[2]), nf = V(Q[3], Q[4], Q[5]), kd = V(Q[6], Q[7], Q[8]);
        const float ray_eps = Q[9];
        uint32_t gpix, kc;
        memcpy(&gpix, Q + 10, 4); memcpy(&kc, Q + 11, 4);
        for (int64_t i = 0; i < n_ind; ++i) {
            float dx = ind[i].p[0] - p.x, dy = ind[i].p[1] - p.y, dz = ind[i].p[2] - p.z;
            float d2 = dx * dx + dy * dy + dz * dz;
            uint32_t b; memcpy(&b, &d2, 4);
            keys[i] = ((uint64_t)b << 32) | (uint32_t)i;
        }
        qsort(keys, (size_t)n_ind, 8, cmp_u64);
        const int nd = (int)(n_ind < 50 ? n_ind : 50);
        v3 dirs[50];
        for (int j = 0; j < nd; ++j) dirs[j] = oct_decode(ind[(uint32_t)keys[j]].wi_oct);
        v3 sn, tn;
        coord_system(nf, &sn, &tn);
        const v3 fr = V(kd.x * O_INV_PI, kd.y * O_INV_PI, kd.z * O_INV_PI);
        v3 pB[32], pC[32];
        for (int l = 0; l < 32; ++l) pB[l] = pC[l] = V(0, 0, 0);
        for (int r = 0; r < 2 * gs; ++r) {
            const int lane = r & 31, cone = r >= gs;
            uint32_t u[4];
            philox(gpix, kc, 0x200u + (uint32_t)r, 0u, k0, k1, u);
            v3 wi; float cosw;
            if (!cone) {
                v3 l = cosine_hemisphere(u01(u[0]), u01(u[1]));
                wi = V(sn.x * l.x + tn.x * l.y + nf.x * l.z, sn.y * l.x + tn.y * l.y + nf.y * l.z,
                       sn.z * l.x + tn.z * l.y + nf.z * l.z);
                cosw = vdot(wi, nf);
                if (!(l.z > 0.f) || !(cosw > 0.f)) continue;
            } else {
                if (nd == 0) continue;
                int pn = (int)floorf(u01(u[2]) * (float)nd);
                if (pn > nd - 1) pn = nd - 1;
                v3 z = dirs[pn], vx, vy;
                coord_system(z, &vx, &vy);
                float u1 = u01(u[0]), u2 = u01(u[1]);
                float costheta = (1.f - u1) * cos_angle + u1 * 1.f;
                float sintheta = sqrtf(fmaxf(0.f, 1.f - costheta * costheta));
                float s, c;
                port_sincos_quarter(4.f * u2, &s, &c);
                float a = c * sintheta, b = s * sintheta;
                wi = V(a * vx.x + b * vy.x + costheta * z.x, a * vx.y + b * vy.y + costheta * z.y,
                       a * vx.z + b * vy.z + costheta * z.z);
                cosw = vdot(wi, nf);
                if (!(cosw > 0.f)) continue;
            }
            osurf s; int prim;
            if (!closest(sc, p, wi, ray_eps, &s, &prim)) continue;
            v3 ng = vdot(s.ng, vneg(wi)) < 0.f ? vneg(s.ng) : s.ng;
            int64_t ri = radiance_nearest(n_rad, rad, s.p, ng);
            if (ri < 0) continue;
            v3 Lo = V(rad[ri].alpha[0], rad[ri].alpha[1], rad[ri].alpha[2]);
            float bsdf_pdf = fabsf(cosw) * O_INV_PI;
            float photon_pdf = 0.f;
            for (int j = 0; j < nd; ++j) if (vdot(dirs[j], wi) > thr) photon_pdf += cone_pdf;
            photon_pdf = nd > 0 ? photon_pdf / (float)nd : 0.f;
            float absdot = fabsf(cosw);
            if (!cone) {
                float f = (float)gs * bsdf_pdf, g = (float)gs * photon_pdf;
                float wt = (f * f) / (f * f + g * g);
                float w = absdot * wt / bsdf_pdf;
                pB[lane] = vadd(pB[lane], V(fr.x * Lo.x * w, fr.y * Lo.y * w, fr.z * Lo.z * w));
            } else {
                float f = (float)gs * photon_pdf, g = (float)gs * bsdf_pdf;
                float wt = (f * f) / (f * f + g * g);
                pC[lane] = vadd(pC[lane], V(fr.x * Lo.x * absdot * wt / photon_pdf, fr.y * Lo.y * absdot * wt / photon_pdf,
                                            fr.z * Lo.z * absdot * wt / photon_pdf));
            }
        }
        for (int o = 16; o > 0; o >>= 1) {
            v3 nb[32], nc[32];
            for (int l = 0; l < 32; ++l) { nb[l] = vadd(pB[l], pB[l ^ o]); nc[l] = vadd(pC[l], pC[l ^ o]); }
            memcpy(pB, nb, sizeof(pB)); memcpy(pC, nc, sizeof(pC));
        }
        const float g = (float)gs;
        rgb[3 * q] = pB[0].x / g + pC[0].x / g;
        rgb[3 * q + 1] = pB[0].y / g + pC[0].y / g;
        rgb[3 * q + 2] = pB[0].z / g + pC[0].z / g;
    }
    free(keys);
}

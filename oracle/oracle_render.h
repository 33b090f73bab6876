/* TEST INFRASTRUCTURE ONLY (included at the end of oracle_photon.c).  Per-sample oracle of the camera pass: a plain-C
 * restatement of csrc/render.cu render_li_kernel + direct_lighting at the same Philox counters, with the exhaustive
 * closest-hit / any-hit loops instead of the BVH.  What it restates follows pbrt-v2: PhotonIntegrator::Li head
 * (integrators/photonmap.cpp:624-638), UniformSampleAllLights / EstimateDirect with MIS (core/integrator.cpp:47-174),
 * SpecularReflect / SpecularTransmit (:177-258), DiffuseAreaLight / PointLight Sample_L (lights/diffuse.cpp:69-81,
 * lights/point.cpp:50-57), Shape::Pdf (core/shape.cpp:95-108).  Pin: images against pbrt (tests/golden/images_*.npz and the
 * converged goldens) -- per-sample values cannot be compared with pbrt (different random streams). */
float port_vdc(uint32_t n, uint32_t scramble);          /* oracle_film.c */
float port_sobol2(uint32_t n, uint32_t scramble);

static int r_disk_hit(const float *D, v3 o_, v3 d_, float tmin, float tmax, float *t) {   /* Disk::Intersect, full phi */
    v3 o, d;
    xfp(D + 12, o_, &o);
    d = V(D[12] * d_.x + D[13] * d_.y + D[14] * d_.z, D[16] * d_.x + D[17] * d_.y + D[18] * d_.z,
          D[20] * d_.x + D[21] * d_.y + D[22] * d_.z);
    if ((double)fabsf(d.z) < 1e-7) return 0;
    float thit = (D[25] - o.z) / d.z;
    if (thit < tmin || thit > tmax) return 0;
    float px = o.x + d.x * thit, py = o.y + d.y * thit;
    float dist2 = px * px + py * py;
    if (dist2 > D[24] * D[24] || dist2 < D[26] * D[26]) return 0;
    *t = thit;
    return 1;
}
static float r_disk_light_pdf(const float *D, const olight *L, v3 p, v3 wi) {
    float t;
    if (!r_disk_hit(D, p, wi, 1e-3f, INFINITY, &t)) return 0.f;
    v3 dv = vscale(wi, t);
    float d2 = dv.x * dv.x + dv.y * dv.y + dv.z * dv.z;
    float pdf = d2 / (fabsf(vdot(L->n, vneg(wi))) * L->area);
    return isinf(pdf) ? 0.f : pdf;
}
static int r_occluded(const port_scene *sc, v3 o, v3 d, float tmin, float tmax) {
    float ray[8] = {o.x, o.y, o.z, tmin, d.x, d.y, d.z, tmax};
    uint8_t occ = 0;
    port_occluded_brute(sc->nprims, sc->kind, sc->verts, sc->idx, sc->analytic, 1, ray, &occ, 1);
    return occ;
}
static float r_power(float f, float g) { return (f * f) / (f * f + g * g); }

static v3 r_direct(const port_scene *sc, const olight *Ls, const osurf *s, const omat *m, v3 wo, uint32_t c0, uint32_t c1,
                   uint32_t node, uint32_t k0, uint32_t k1, int sampler_type) {
    v3 L = V(0, 0, 0);
    const v3 f_lambert = vscale(m->kd, O_INV_PI);
    const int pattern = sampler_type != 0 && node == 1u;
    for (int li = 0; li < sc->nlights; ++li) {
        const olight *Lt = &Ls[li];
        const int light_prim = (int)sc->lights[8 * li + 7];
        const int delta = Lt->type != 1;                  /* point, spot, distant */
        const int ns = delta ? 1 : (sc->light_nsamples && sc->light_nsamples[li] > 1 ? sc->light_nsamples[li] : 1);
        v3 Ld = V(0, 0, 0);
        uint32_t scr[4] = {0, 0, 0, 0};
        if (pattern && !delta) philox(c0, 0xfffffffeu, (uint32_t)li, 0u, k0, k1, scr);
        for (int j = 0; j < ns; ++j) {
            uint32_t r[4];
            philox(c0, c1, node, 0x100u + (uint32_t)(li * 64 + j), k0, k1, r);
            float u_l1 = u01(r[0]), u_l2 = u01(r[1]), u_b1 = u01(r[2]), u_b2 = u01(r[3]);
            if (pattern) {
                const uint32_t n = c1 * (uint32_t)ns + (uint32_t)j;
                u_l1 = port_vdc(n, scr[0]); u_l2 = port_sobol2(n, scr[1]); u_b1 = port_vdc(n, scr[2]); u_b2 = port_sobol2(n, scr[3]);
            }
            if (Lt->type == 3) {                  /* DistantLight::Sample_L (lights/distant.cpp:48-55) */
                v3 wi = Lt->n;
                if (!(vdot(wi, s->ng) * vdot(wo, s->ng) > 0.f)) continue;
                if (r_occluded(sc, s->p, wi, s->ray_eps, INFINITY)) continue;
                float w = fabsf(vdot(wi, s->n));
                Ld = vadd(Ld, V(f_lambert.x * Lt->rgb.x * w, f_lambert.y * Lt->rgb.y * w, f_lambert.z * Lt->rgb.z * w));
                continue;
            }
            if (delta) {                          /* PointLight / SpotLight::Sample_L */
                v3 dl = vsub(Lt->p, s->p);
                float d2 = dl.x * dl.x + dl.y * dl.y + dl.z * dl.z;
                v3 wi = vnormalize(dl);
                if (!(vdot(wi, s->ng) * vdot(wo, s->ng) > 0.f)) continue;
                v3 I = Lt->rgb;
                if (Lt->type == 2) {
                    float fo = spot_falloff(Lt, vneg(wi));
                    if (fo == 0.f) continue;
                    I = vscale(I, fo);
                }
                if (r_occluded(sc, s->p, dl, s->ray_eps, 1.f)) continue;
                float w = fabsf(vdot(wi, s->n));
                Ld = vadd(Ld, V(f_lambert.x * (I.x / d2) * w, f_lambert.y * (I.y / d2) * w, f_lambert.z * (I.z / d2) * w));
                continue;
            }
            const float *D = sc->analytic + PORT_ANALYTIC_STRIDE * (int64_t)Lt->analytic;
            {   /* light sampling */
                float x, y;
                concentric_disk(u_l1, u_l2, &x, &y);
                x *= D[24]; y *= D[24];
                v3 ps;
                xfp(D, V(x, y, D[25]), &ps);
                v3 dl = vsub(ps, s->p);
                v3 wi = vnormalize(dl);
                float lpdf = r_disk_light_pdf(D, Lt, s->p, wi);
                int emits = vdot(Lt->n, vneg(wi)) > 0.f;
                int same = vdot(wi, s->ng) * vdot(wo, s->ng) > 0.f;
                if (lpdf > 0.f && emits && same) {
                    if (!r_occluded(sc, s->p, dl, s->ray_eps, 1.f - 1e-3f)) {
                        v3 wil = to_local(s, wi), wol = to_local(s, wo);
                        float bpdf = (wol.z * wil.z > 0.f) ? fabsf(wil.z) * O_INV_PI : 0.f;
                        float w = fabsf(vdot(wi, s->n)) * r_power(lpdf, bpdf) / lpdf;
                        Ld = vadd(Ld, V(f_lambert.x * Lt->rgb.x * w, f_lambert.y * Lt->rgb.y * w, f_lambert.z * Lt->rgb.z * w));
                    }
                }
            }
            {   /* BSDF sampling */
                osample bs;
                if (!sample_bsdf(s, m, wo, u_b1, u_b2, 0.f, &bs) || black(bs.f) || !(bs.pdf > 0.f)) continue;
                float lpdf = r_disk_light_pdf(D, Lt, s->p, bs.wi);
                if (lpdf == 0.f) continue;
                float weight = r_power(bs.pdf, lpdf);
                osurf hs; int hp;
                if (closest(sc, s->p, bs.wi, s->ray_eps, &hs, &hp) && hp == light_prim && vdot(Lt->n, vneg(bs.wi)) > 0.f) {
                    float w = bs.absdot * weight / bs.pdf;
                    Ld = vadd(Ld, V(bs.f.x * Lt->rgb.x * w, bs.f.y * Lt->rgb.y * w, bs.f.z * Lt->rgb.z * w));
                }
            }
        }
        float inv = 1.f / (float)ns;
        L = vadd(L, vscale(Ld, inv));
    }
    return L;
}

static int r_refract(v3 wo, float ior, v3 *wi) {           /* SpecularTransmission::Sample_f (reflection.cpp:147-169) */
    int entering = wo.z > 0.f;
    float ei = 1.f, et = ior;
    if (!entering) { ei = ior; et = 1.f; }
    float sini2 = fmaxf(0.f, 1.f - wo.z * wo.z);
    float eta = ei / et;
    float sint2 = eta * eta * sini2;
    if (sint2 >= 1.f) return 0;
    float cost = sqrtf(fmaxf(0.f, 1.f - sint2));
    if (entering) cost = -cost;
    *wi = V(eta * -wo.x, eta * -wo.y, cost);
    return 1;
}

/* cam: o[3], A[3], B[3], C[3] (b2rt_camera without the resolution).  xy: raster positions of the n samples
 * (port_sample_positions), gpix / ks: their global pixel id and sample number.  out_rgb[3n]: Le + direct lighting + both
 * specular branches (the camera-pass terms); out_nq[n]: number of diffuse leaves; out_q (may be NULL): the leaves as 12-float
 * queries, `slots` per sample. */
void port_camera_samples(const port_scene *sc, uint64_t seed, const float cam[12], int64_t n, const float *xy,
                         const uint32_t *gpix, const uint32_t *ks, int sampler_type, int max_specular_depth,
                         int direct_lighting, int want_queries, int slots, float *out_rgb, uint32_t *out_nq, float *out_q) {
    olight Ls[64]; float cdf[130];
    if (sc->nlights > 64) return;
    if (sc->nlights > 0) derive_lights(sc, Ls, cdf);
    const uint32_t k0 = (uint32_t)seed ^ 0x43414d52u, k1 = (uint32_t)(seed >> 32);
    typedef struct { v3 o, d; float tmin; v3 T; int depth; uint32_t id; } node_t;
    for (int64_t ls = 0; ls < n; ++ls) {
        const float fx = xy[2 * ls], fy = xy[2 * ls + 1];
        v3 d0 = V(cam[3] + fx * cam[6] + fy * cam[9], cam[4] + fx * cam[7] + fy * cam[10], cam[5] + fx * cam[8] + fy * cam[11]);
        node_t stack[9];
        int sp = 0;
        stack[sp].o = V(cam[0], cam[1], cam[2]); stack[sp].d = vnormalize(d0); stack[sp].tmin = 0.f;
        stack[sp].T = V(1.f, 1.f, 1.f); stack[sp].depth = 0; stack[sp].id = 1u; ++sp;
        v3 L = V(0, 0, 0);
        uint32_t nq = 0;
        while (sp > 0) {
            node_t nd = stack[--sp];
            osurf sf; int prim;
            if (!closest(sc, nd.o, nd.d, nd.tmin, &sf, &prim)) continue;
            omat m = get_material(sc, prim);
            v3 wo = vneg(nd.d);
            const int light = sc->prim_light ? sc->prim_light[prim] : -1;
            if (light >= 0 && light < sc->nlights && vdot(sf.ng, wo) > 0.f) {
                const float *lg = sc->lights + 8 * light;
                L = vadd(L, V(nd.T.x * lg[4], nd.T.y * lg[5], nd.T.z * lg[6]));
            }
            if (m.type == 0) {
                if (direct_lighting) {
                    v3 Ld = r_direct(sc, Ls, &sf, &m, wo, gpix[ls], ks[ls], nd.id, k0, k1, sampler_type);
                    L = vadd(L, V(nd.T.x * Ld.x, nd.T.y * Ld.y, nd.T.z * Ld.z));
                }
                if (want_queries) {
                    if ((int)nq < slots) {
                        v3 nf = vdot(sf.n, wo) < 0.f ? vneg(sf.n) : sf.n;
                        v3 kd = V(m.kd.x * nd.T.x, m.kd.y * nd.T.y, m.kd.z * nd.T.z);
                        if (out_q) {
                            float *q = out_q + 12 * ((int64_t)ls * slots + nq);
                            uint32_t w0 = gpix[ls], w1 = ks[ls] | (nd.id << 22);
                            q[0] = sf.p.x; q[1] = sf.p.y; q[2] = sf.p.z; q[3] = nf.x; q[4] = nf.y; q[5] = nf.z;
                            q[6] = kd.x; q[7] = kd.y; q[8] = kd.z; q[9] = sf.ray_eps;
                            memcpy(q + 10, &w0, 4); memcpy(q + 11, &w1, 4);
                        }
                        ++nq;
                    }
                }
                continue;
            }
            if (nd.depth + 1 >= max_specular_depth) continue;
            v3 wol = to_local(&sf, wo);
            float F = (m.type == 2) ? fresnel_dielectric(wol.z, 1.f, m.ior) : 1.f;
            if (m.type == 2 && !black(m.kt)) {
                v3 wil;
                if (r_refract(wol, m.ior, &wil) && wil.z != 0.f && sp < 8) {
                    float ac = fabsf(wil.z);
                    v3 w = V((1.f - F) * m.kt.x / ac * ac, (1.f - F) * m.kt.y / ac * ac, (1.f - F) * m.kt.z / ac * ac);
                    if (!black(w)) {
                        stack[sp].o = sf.p; stack[sp].d = to_world(&sf, wil); stack[sp].tmin = sf.ray_eps;
                        stack[sp].T = V(nd.T.x * w.x, nd.T.y * w.y, nd.T.z * w.z); stack[sp].depth = nd.depth + 1;
                        stack[sp].id = nd.id * 2u + 1u; ++sp;
                    }
                }
            }
            if (!black(m.kd) && wol.z != 0.f && sp < 8) {
                v3 wil = V(-wol.x, -wol.y, wol.z);
                float ac = fabsf(wil.z);
                v3 w = V(F * m.kd.x / ac * ac, F * m.kd.y / ac * ac, F * m.kd.z / ac * ac);
                if (!black(w)) {
                    stack[sp].o = sf.p; stack[sp].d = to_world(&sf, wil); stack[sp].tmin = sf.ray_eps;
                    stack[sp].T = V(nd.T.x * w.x, nd.T.y * w.y, nd.T.z * w.z); stack[sp].depth = nd.depth + 1;
                    stack[sp].id = nd.id * 2u; ++sp;
                }
            }
        }
        out_rgb[3 * ls] = L.x; out_rgb[3 * ls + 1] = L.y; out_rgb[3 * ls + 2] = L.z;
        out_nq[ls] = nq;
    }
}

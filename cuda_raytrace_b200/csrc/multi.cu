// Multi-GPU layer of the C ABI (include/b2rt.h, "b2rt_multi_*"): one process drives every GPU of the box, the process
// model SURVEY.md 8(e) asks for -- the caller stays single-threaded (the reference's contract, cuda_render/cudaapi.cpp:3),
// the library keeps one worker thread per device so the per-device host work (launches, the few host reads of a pass)
// overlaps.  The reference has no multi-GPU path at all (one implicit device, cuda_render/cudarender.cpp:17-36).
//
//   photon pass   path range [first, first + n) cut into one contiguous slice per GPU (a path's photons depend only on
//                 (seed, path index)).  Every GPU traces its slice (b2rt_photons_trace_begin), the host adds up the
//                 per-GPU photon counts, and then the COMPACTION kernel of GPU g writes g's photons to their final
//                 position in the photon arrays of ALL GPUs: peer stores over NVLink / NVSwitch, compute and exchange in
//                 one kernel, no send buffer, no staging copy, no collective afterwards.  Rank-order concatenation of
//                 path-ordered slices = the single-GPU array byte for byte.  That single-kernel form is B2RT_MGPU_EXCHANGE=fused;
//                 the DEFAULT is `push`: the compaction writes the local array and a second kernel stores the finished
//                 slice into the peers with whole 512-byte lines per warp (r04r on 8 x B200: 0.66 against 1.33 ms, the
//                 compaction's own peer stores being 32-byte pieces at path-dependent offsets).  Two more modes for A/B and
//                 for boxes without peer access: B2RT_MGPU_EXCHANGE=nccl (local compaction in place, then one grouped
//                 ncclBroadcast per slice straight between the final arrays; libnccl is dlopen'ed, never linked) and
//                 =copy (cudaMemcpyPeerAsync per slice).
//   grids         replicated: every GPU builds the same deterministic grid over the same bytes.  (A partition of the
//                 build was costed and dropped: shipping the sorted records back, 56 B / photon over NVLink, takes longer
//                 than sorting them locally, 188 B / photon of HBM traffic -- DESIGN.md section 6.)
//   camera pass   the image is cut into 64 x 64 tiles, tile t belongs to GPU t mod G (interleaved for load balance);
//                 each GPU renders ITS pixel list in one b2rt_render_pixels_impl call and its resolve kernel stores every
//                 pixel at its place in the whole image on GPU 0 (peer stores again: the composite is part of the
//                 kernel).  Pixel randomness is a function of the global pixel id, so the image is bit-identical to the
//                 single-GPU frame.  Pixel filters wider than a pixel fall back to one row band per GPU rendered through
//                 the crop window (every band re-traces its halo cells) and copied to GPU 0.
#include "b2rt_internal.cuh"
#include <dlfcn.h>
#include <string.h>
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>

namespace {

// ---- the slice of NCCL this file uses, resolved at run time (torch ships 2.28, the system 2.27: whichever is loaded) ----
typedef struct ncclComm *ncclComm_t;
typedef int ncclResult_t;
struct Nccl {
    void *lib = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load() {
        if (lib) return true;
        const char *names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char *n : names)
            if ((lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
        if (!lib) return false;
        CommInitAll = (decltype(CommInitAll))dlsym(lib, "ncclCommInitAll");
        CommDestroy = (decltype(CommDestroy))dlsym(lib, "ncclCommDestroy");
        GroupStart = (decltype(GroupStart))dlsym(lib, "ncclGroupStart");
        GroupEnd = (decltype(GroupEnd))dlsym(lib, "ncclGroupEnd");
        Broadcast = (decltype(Broadcast))dlsym(lib, "ncclBroadcast");
        GetErrorString = (decltype(GetErrorString))dlsym(lib, "ncclGetErrorString");
        return CommInitAll && CommDestroy && GroupStart && GroupEnd && Broadcast && GetErrorString;
    }
};
constexpr int NCCL_UINT8 = 1;      // ncclUint8 (nccl.h: ncclInt8 = 0, ncclUint8 = 1)

// one worker per device: runs the job it is handed with its device current
struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, done = false, quit = false;
    int rc = 0;
};

enum { EXCH_FUSED = 0, EXCH_NCCL = 1, EXCH_COPY = 2, EXCH_PUSH = 3 };

}  // namespace

struct b2rt_multi {
    int n = 0;
    std::vector<b2rt_ctx *> ctx;
    std::vector<int> dev;
    std::vector<Worker *> workers;
    bool peer = false;              // every pair of devices has peer access enabled
    int exchange = EXCH_PUSH;       // r04r, 8 x B200, 400 MB of photons: push 0.66 ms (528 GB/s bus), peer copies 0.63, NCCL 1.11,
                                    // stores from inside the compaction kernel 1.33 (32-byte pieces at path-dependent offsets)
    Nccl nccl;
    std::vector<ncclComm_t> comms;
    std::string err;
    b2rt_multi_stats stats{};
    // camera pass: per device pixel lists of the last (xres, yres, window) and the frame on GPU 0
    std::vector<DevBuf *> pixmaps;
    std::vector<int64_t> pixmap_n;
    int pm_xres = 0, pm_yres = 0, pm_win[4] = {0, 0, 0, 0}, pm_tile = 0;
    DevBuf *frame0 = nullptr;       // whole image on device 0
    std::vector<DevBuf *> bands;    // per-device band buffers (filtered frames)
};

namespace {

void worker_main(b2rt_multi *m, int i) {
    cudaSetDevice(m->dev[i]);
    Worker &w = *m->workers[i];
    for (;;) {
        std::function<int()> job;
        {
            std::unique_lock<std::mutex> lk(w.mu);
            w.cv.wait(lk, [&] { return w.has_job || w.quit; });
            if (w.quit) return;
            job = w.job;
            w.has_job = false;
        }
        int rc = job();
        {
            std::lock_guard<std::mutex> lk(w.mu);
            w.rc = rc; w.done = true;
        }
        w.cv.notify_all();
    }
}

// run f(i) on every device's worker at once; returns the first failing code (message copied from that context)
int parallel(b2rt_multi *m, const std::function<int(int)> &f) {
    for (int i = 0; i < m->n; ++i) {
        Worker &w = *m->workers[i];
        std::lock_guard<std::mutex> lk(w.mu);
        w.job = [&f, i] { return f(i); };
        w.has_job = true; w.done = false;
        w.cv.notify_all();
    }
    int rc = B2RT_OK;
    for (int i = 0; i < m->n; ++i) {
        Worker &w = *m->workers[i];
        std::unique_lock<std::mutex> lk(w.mu);
        w.cv.wait(lk, [&] { return w.done; });
        if (w.rc && !rc) { rc = w.rc; m->err = "gpu " + std::to_string(m->dev[i]) + ": " + m->ctx[i]->err; }
    }
    return rc;
}

void slice_of(int64_t first, int64_t n, int i, int G, int64_t *lo, int64_t *cnt) {
    const int64_t a = n * i / G, b = n * (i + 1) / G;
    *lo = first + a; *cnt = b - a;
}

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// The exchange of one sharded photon pass.  begin(i, lo, cnt, tot) traces device i's slice and returns its two photon
// counts; commit(i, cnt, dst) compacts it to the destinations.  mapA / mapB: the two maps the pass fills.
int sharded_pass(b2rt_multi *m, int64_t first, int64_t n_paths, int mapA, int mapB, bool countA, bool countB,
                 const std::function<int(int, int64_t, int64_t, uint32_t *)> &begin,
                 const std::function<int(int, int64_t, const PhotonDst &)> &commit, int64_t n_stored[2]) {
    const int G = m->n;
    std::vector<uint32_t> tot(2 * G, 0);
    const double t0 = now_ms();
    int rc = parallel(m, [&](int i) -> int {
        int64_t lo, cnt;
        slice_of(first, n_paths, i, G, &lo, &cnt);
        return begin(i, lo, cnt, &tot[2 * i]);
    });
    if (rc) return rc;
    const double t1 = now_ms();
    int64_t offA[B2RT_MAX_GPUS + 1] = {0}, offB[B2RT_MAX_GPUS + 1] = {0};
    for (int i = 0; i < G; ++i) { offA[i + 1] = offA[i] + tot[2 * i]; offB[i + 1] = offB[i] + tot[2 * i + 1]; }
    const int64_t baseA = m->ctx[0]->maps[mapA].n, baseB = m->ctx[0]->maps[mapB].n;
    if (baseA + offA[G] >= (1ll << 32) || baseB + offB[G] >= (1ll << 32)) { m->err = "photon pass: a map would exceed 2^32 photons"; return B2RT_E_ARG; }
    // every device makes room for the whole pass (a phase of its own only when some array has to grow)
    bool grow = false;
    for (int i = 0; i < G; ++i)
        grow = grow || m->ctx[i]->maps[mapA].photons.cap < (size_t)(baseA + offA[G]) * sizeof(b2rt_photon) ||
               m->ctx[i]->maps[mapB].photons.cap < (size_t)(baseB + offB[G]) * sizeof(b2rt_photon);
    if (grow) rc = parallel(m, [&](int i) -> int {
        b2rt_ctx *c = m->ctx[i];
        int r = b2rt_photons_reserve(c, c->maps[mapA].photons, baseA, baseA + offA[G]);
        if (r) return r;
        return b2rt_photons_reserve(c, c->maps[mapB].photons, baseB, baseB + offB[G]);
    });
    if (rc) return rc;
    const double t2 = now_ms();
    const int mode = ((m->exchange == EXCH_FUSED || m->exchange == EXCH_PUSH) && !m->peer && G > 1) ? EXCH_COPY : m->exchange;
    rc = parallel(m, [&](int i) -> int {
        b2rt_ctx *c = m->ctx[i];
        int64_t lo, cnt;
        slice_of(first, n_paths, i, G, &lo, &cnt);
        PhotonDst dst;
        if (mode == EXCH_FUSED) {
            dst.n = G;
            for (int g = 0; g < G; ++g) {
                dst.a[g] = m->ctx[g]->maps[mapA].photons.as<b2rt_photon>() + baseA + offA[i];
                dst.b[g] = m->ctx[g]->maps[mapB].photons.as<b2rt_photon>() + baseB + offB[i];
            }
        } else {
            dst.n = 1;
            dst.a[0] = c->maps[mapA].photons.as<b2rt_photon>() + baseA + offA[i];
            dst.b[0] = c->maps[mapB].photons.as<b2rt_photon>() + baseB + offB[i];
        }
        int r = commit(i, cnt, dst);
        if (r) return r;
        if (mode == EXCH_PUSH) {
            // local slice -> the same place in every peer's array, one coalesced copy kernel per map
            const int maps[2] = {mapA, mapB};
            const int64_t *offs[2] = {offA, offB};
            const int64_t bases[2] = {baseA, baseB};
            for (int t = 0; t < 2; ++t) {
                const int64_t np = offs[t][i + 1] - offs[t][i];
                PhotonDst peers;
                peers.n = 0;
                for (int k = 1; k < G; ++k) {
                    const int g = (i + k) % G;
                    peers.a[peers.n++] = m->ctx[g]->maps[maps[t]].photons.as<b2rt_photon>() + bases[t] + offs[t][i];
                }
                r = b2rt_photons_push(c, c->maps[maps[t]].photons.as<b2rt_photon>() + bases[t] + offs[t][i], np, peers);
                if (r) return r;
            }
        }
        if (mode == EXCH_COPY) {
            for (int k = 1; k < G; ++k) {           // staggered so that not everybody targets the same peer first
                const int g = (i + k) % G;
                const int maps[2] = {mapA, mapB};
                const int64_t *offs[2] = {offA, offB};
                const int64_t bases[2] = {baseA, baseB};
                for (int t = 0; t < 2; ++t) {
                    const size_t bytes = (size_t)(offs[t][i + 1] - offs[t][i]) * sizeof(b2rt_photon);
                    if (!bytes) continue;
                    const size_t at = (size_t)(bases[t] + offs[t][i]) * sizeof(b2rt_photon);
                    B2RT_CUDA(c, cudaMemcpyPeerAsync((char *)m->ctx[g]->maps[maps[t]].photons.p + at, m->dev[g],
                                                     (char *)c->maps[maps[t]].photons.p + at, m->dev[i], bytes, c->stream));
                }
            }
        }
        return b2rt_sync(c);            // stream sync + the traversal-overflow status word
    });
    if (rc) return rc;
    if (mode == EXCH_NCCL && G > 1) {
        // slices are in place on their owners; one grouped broadcast per (slice, map) between the final arrays
        ncclResult_t nr = m->nccl.GroupStart();
        for (int t = 0; t < 2 && nr == 0; ++t) {
            const int map = t ? mapB : mapA;
            const int64_t *off = t ? offB : offA;
            const int64_t base = t ? baseB : baseA;
            for (int r = 0; r < G && nr == 0; ++r) {
                const size_t bytes = (size_t)(off[r + 1] - off[r]) * sizeof(b2rt_photon);
                if (!bytes) continue;
                for (int g = 0; g < G && nr == 0; ++g) {
                    char *p = (char *)m->ctx[g]->maps[map].photons.p + (size_t)(base + off[r]) * sizeof(b2rt_photon);
                    nr = m->nccl.Broadcast(p, p, bytes, NCCL_UINT8, r, m->comms[g], m->ctx[g]->stream);
                }
            }
        }
        ncclResult_t ne = m->nccl.GroupEnd();
        if (nr || ne) { m->err = std::string("nccl: ") + m->nccl.GetErrorString(nr ? nr : ne); return B2RT_E_CUDA; }
        rc = parallel(m, [&](int i) -> int { B2RT_CUDA(m->ctx[i], cudaStreamSynchronize(m->ctx[i]->stream)); return B2RT_OK; });
        if (rc) return rc;
    }
    const double t3 = now_ms();
    for (int i = 0; i < G; ++i) {
        b2rt_ctx *c = m->ctx[i];
        PhotonMap &A = c->maps[mapA], &B = c->maps[mapB];
        A.n = baseA + offA[G]; B.n = baseB + offB[G];
        if (countA) A.n_paths += n_paths;
        if (countB) B.n_paths += n_paths;
        A.grid_built = B.grid_built = false;
    }
    if (n_stored) { n_stored[0] = offA[G]; n_stored[1] = offB[G]; }
    b2rt_multi_stats &s = m->stats;
    s.trace_ms += (float)(t1 - t0); s.reserve_ms += (float)(t2 - t1); s.exchange_ms += (float)(t3 - t2);
    s.exchange_bytes += (int64_t)(offA[G] + offB[G]) * (int64_t)sizeof(b2rt_photon);      // what every GPU ends up holding
    s.passes += 1;
    return B2RT_OK;
}

int check_state(b2rt_multi *m, const char *who) {
    for (int i = 0; i < m->n; ++i)
        if (!m->ctx[i]->built) { m->err = std::string(who) + ": b2rt_multi_build_accel has not been called"; return B2RT_E_STATE; }
    return B2RT_OK;
}

}  // namespace

extern "C" {

int b2rt_multi_photons_clear(b2rt_multi *m);

int b2rt_multi_create(int ngpus, const int *devices, uint64_t seed, b2rt_multi **out) {
    if (!out || ngpus < 1 || ngpus > B2RT_MAX_GPUS) return B2RT_E_ARG;
    *out = nullptr;
    b2rt_multi *m = new b2rt_multi();
    m->n = ngpus;
    for (int i = 0; i < ngpus; ++i) {
        const int d = devices ? devices[i] : i;
        b2rt_ctx *c = nullptr;
        int rc = b2rt_create(d, seed, &c);
        if (rc) {
            for (b2rt_ctx *p : m->ctx) b2rt_destroy(p);
            delete m;
            return rc;                                  // message: b2rt_last_error(NULL)
        }
        m->ctx.push_back(c); m->dev.push_back(d);
    }
    // peer access between every pair (NVSwitch box: all of them)
    m->peer = true;
    for (int i = 0; i < ngpus; ++i)
        for (int j = 0; j < ngpus; ++j) {
            if (i == j || m->dev[i] == m->dev[j]) continue;       // several contexts on one device share its address space
            int can = 0;
            cudaDeviceCanAccessPeer(&can, m->dev[i], m->dev[j]);
            if (!can) { m->peer = false; continue; }
            cudaSetDevice(m->dev[i]);
            cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) m->peer = false;
            cudaGetLastError();
        }
    if (const char *v = getenv("B2RT_MGPU_EXCHANGE")) {
        if (!strcmp(v, "nccl")) m->exchange = EXCH_NCCL;
        else if (!strcmp(v, "copy")) m->exchange = EXCH_COPY;
        else if (!strcmp(v, "push")) m->exchange = EXCH_PUSH;
        else if (!strcmp(v, "fused")) m->exchange = EXCH_FUSED;
    }
    if (m->exchange == EXCH_NCCL && ngpus > 1) {
        bool ok = m->nccl.load();
        if (ok) {
            m->comms.resize(ngpus);
            ok = m->nccl.CommInitAll(m->comms.data(), ngpus, m->dev.data()) == 0;
            if (!ok) m->comms.clear();
        }
        if (!ok) m->exchange = m->peer ? EXCH_PUSH : EXCH_COPY;        // reported by b2rt_multi_info
    }
    for (int i = 0; i < ngpus; ++i) {
        m->workers.push_back(new Worker());
        m->pixmaps.push_back(new DevBuf()); m->pixmap_n.push_back(0);
        m->bands.push_back(new DevBuf());
    }
    m->frame0 = new DevBuf();
    for (int i = 0; i < ngpus; ++i) m->workers[i]->th = std::thread(worker_main, m, i);
    *out = m;
    return B2RT_OK;
}

// Grow the group to ngpus devices: the new contexts get a copy of context 0's scene (everything b2rt_add_* staged on the
// host side), of its materials, lights, sampler, filter and final-gather settings -- so a caller that learns the GPU
// count late (the pbrt drop-in reads "integer gpus" from the Renderer line after the scene has been parsed) ingests
// the scene once.  Call b2rt_multi_build_accel afterwards.  Photon maps are not copied (cleared on all).
int b2rt_multi_resize(b2rt_multi *m, int ngpus, const int *devices) {
    if (!m || ngpus < m->n || ngpus > B2RT_MAX_GPUS) { if (m) m->err = "multi_resize: can only grow, up to 16 devices"; return B2RT_E_ARG; }
    if (ngpus == m->n) return B2RT_OK;
    if (m->exchange == EXCH_NCCL) { m->err = "multi_resize: not with the NCCL exchange (communicators are created for a fixed group)"; return B2RT_E_STATE; }
    const b2rt_ctx *c0 = m->ctx[0];
    const int old = m->n;
    for (int i = old; i < ngpus; ++i) {
        const int d = devices ? devices[i] : i;
        b2rt_ctx *c = nullptr;
        int rc = b2rt_create(d, c0->seed, &c);
        if (rc) { m->err = b2rt_last_error(nullptr); return rc; }
        rc = b2rt_copy_staged_scene(c0, c);      // meshes, analytic shapes, objects and instances, materials, lights
        if (rc) { m->err = c->err; b2rt_destroy(c); return rc; }
        c->film_filter = c0->film_filter; c->sampler_type = c0->sampler_type; c->sampler_nx = c0->sampler_nx; c->sampler_ny = c0->sampler_ny;
        c->fg_enable = c0->fg_enable; c->fg_samples = c0->fg_samples; c->fg_cos_angle = c0->fg_cos_angle;
        m->ctx.push_back(c); m->dev.push_back(d);
    }
    m->n = ngpus;
    for (int i = 0; i < ngpus; ++i)
        for (int j = 0; j < ngpus; ++j) {
            if (i == j || m->dev[i] == m->dev[j] || (i < old && j < old)) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, m->dev[i], m->dev[j]);
            if (!can) { m->peer = false; continue; }
            cudaSetDevice(m->dev[i]);
            cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) m->peer = false;
            cudaGetLastError();
        }
    for (int i = old; i < ngpus; ++i) {
        m->workers.push_back(new Worker());
        m->pixmaps.push_back(new DevBuf()); m->pixmap_n.push_back(0);
        m->bands.push_back(new DevBuf());
        m->workers[i]->th = std::thread(worker_main, m, i);
    }
    m->pm_xres = m->pm_yres = 0;           // tile ownership changes with the GPU count
    b2rt_multi_photons_clear(m);
    return B2RT_OK;
}

void b2rt_multi_destroy(b2rt_multi *m) {
    if (!m) return;
    for (Worker *w : m->workers) {
        { std::lock_guard<std::mutex> lk(w->mu); w->quit = true; }
        w->cv.notify_all();
        w->th.join();
        delete w;
    }
    for (ncclComm_t c : m->comms) m->nccl.CommDestroy(c);
    for (int i = 0; i < m->n; ++i) {
        cudaSetDevice(m->dev[i]);
        cudaDeviceSynchronize();
        delete m->pixmaps[i]; delete m->bands[i];
        if (i == 0) delete m->frame0;
    }
    for (b2rt_ctx *c : m->ctx) b2rt_destroy(c);
    delete m;
}

int b2rt_multi_photons_clear(b2rt_multi *m);
const char *b2rt_multi_last_error(b2rt_multi *m) { return m ? m->err.c_str() : b2rt_last_error(nullptr); }
int b2rt_multi_count(b2rt_multi *m) { return m ? m->n : 0; }
b2rt_ctx *b2rt_multi_ctx(b2rt_multi *m, int i) { return (m && i >= 0 && i < m->n) ? m->ctx[i] : nullptr; }

int b2rt_multi_info(b2rt_multi *m, int32_t *peer_access, int32_t *exchange_mode) {
    if (!m) return B2RT_E_ARG;
    if (peer_access) *peer_access = m->peer ? 1 : 0;
    if (exchange_mode) *exchange_mode = ((m->exchange == EXCH_FUSED || m->exchange == EXCH_PUSH) && !m->peer && m->n > 1) ? EXCH_COPY : m->exchange;
    return B2RT_OK;
}

int b2rt_multi_stats_get(b2rt_multi *m, b2rt_multi_stats *out, int reset) {
    if (!m || !out) return B2RT_E_ARG;
    *out = m->stats;
    if (reset) m->stats = b2rt_multi_stats{};
    return B2RT_OK;
}

int b2rt_multi_build_accel(b2rt_multi *m) {
    if (!m) return B2RT_E_ARG;
    return parallel(m, [&](int i) -> int { return b2rt_build_accel(m->ctx[i]); });
}

int b2rt_multi_photons_clear(b2rt_multi *m) {
    if (!m) return B2RT_E_ARG;
    for (b2rt_ctx *c : m->ctx) b2rt_photons_clear(c);
    return B2RT_OK;
}

int b2rt_multi_photons_trace(b2rt_multi *m, int64_t first_path, int64_t n_paths, int32_t max_depth, int32_t flags,
                             int64_t n_stored[2]) {
    if (!m) return B2RT_E_ARG;
    if (first_path < 0 || n_paths < 0 || max_depth < 1 || max_depth > 64 || !(flags & 3)) { m->err = "multi_photons_trace: bad arguments"; return B2RT_E_ARG; }
    int rc = check_state(m, "multi_photons_trace");
    if (rc) return rc;
    if (m->ctx[0]->h_lights.empty()) { m->err = "multi_photons_trace: the scene has no lights"; return B2RT_E_STATE; }
    rc = sharded_pass(m, first_path, n_paths, B2RT_MAP_CAUSTIC, B2RT_MAP_INDIRECT, (flags & 1) != 0, (flags & 2) != 0,
                      [&](int i, int64_t lo, int64_t cnt, uint32_t *tot) { return b2rt_photons_trace_begin(m->ctx[i], lo, cnt, max_depth, flags, tot); },
                      [&](int i, int64_t cnt, const PhotonDst &dst) { return b2rt_photons_trace_commit(m->ctx[i], cnt, max_depth, dst); },
                      n_stored);
    return rc;
}

int b2rt_multi_photons_trace_fg(b2rt_multi *m, int64_t first_path, int64_t n_paths, int32_t max_depth, int64_t n_stored[2]) {
    if (!m) return B2RT_E_ARG;
    if (first_path < 0 || n_paths < 0 || max_depth < 1 || max_depth > 64) { m->err = "multi_photons_trace_fg: bad arguments"; return B2RT_E_ARG; }
    int rc = check_state(m, "multi_photons_trace_fg");
    if (rc) return rc;
    rc = sharded_pass(m, first_path, n_paths, B2RT_MAP_DIRECT, B2RT_MAP_RADIANCE, true, true,
                      [&](int i, int64_t lo, int64_t cnt, uint32_t *tot) { return b2rt_photons_trace_fg_begin(m->ctx[i], lo, cnt, max_depth, tot); },
                      [&](int i, int64_t cnt, const PhotonDst &dst) { return b2rt_photons_trace_fg_commit(m->ctx[i], cnt, max_depth, dst); },
                      n_stored);
    if (rc) return rc;
    // nDirectPaths only counts the paths shot while the indirect map was collecting (sharded_pass added n_paths)
    const int64_t counted = std::max<int64_t>(0, std::min<int64_t>(first_path + n_paths, m->ctx[0]->fg_lim_i) - first_path);
    for (b2rt_ctx *c : m->ctx) c->maps[B2RT_MAP_DIRECT].n_paths += counted - n_paths;
    return B2RT_OK;
}

// pbrt's stopping rule over sharded batches: the same loop as b2rt_photons_shoot, every batch cut over the GPUs
int b2rt_multi_photons_shoot(b2rt_multi *m, int64_t n_caustic_wanted, int64_t n_indirect_wanted, int32_t max_depth,
                             int64_t batch_paths, int64_t n_paths_out[2], int64_t n_stored_out[2]) {
    if (!m) return B2RT_E_ARG;
    if (n_caustic_wanted < 0 || n_indirect_wanted < 0 || batch_paths <= 0) { m->err = "multi_photons_shoot: bad arguments"; return B2RT_E_ARG; }
    int rc = b2rt_multi_photons_clear(m);
    if (rc) return rc;
    int64_t first = 0;
    bool c_done = n_caustic_wanted == 0, i_done = n_indirect_wanted == 0;
    while (!(c_done && i_done)) {
        const int flags = (c_done ? 0 : 1) | (i_done ? 0 : 2);
        rc = b2rt_multi_photons_trace(m, first, batch_paths, max_depth, flags, nullptr);
        if (rc) return rc;
        first += batch_paths;
        if (m->ctx[0]->maps[0].n >= n_caustic_wanted) c_done = true;
        if (m->ctx[0]->maps[1].n >= n_indirect_wanted) i_done = true;
        if (first > 500000 && ((!c_done && m->ctx[0]->maps[0].n < 4) || (!i_done && m->ctx[0]->maps[1].n < 4))) {
            b2rt_multi_photons_clear(m);          // pbrt's give-up: maps erased, not an error (see b2rt_photons_shoot)
            m->err = "multi_photons_shoot: unable to store enough photons, giving up (photon maps erased, as pbrt does)";
            break;
        }
    }
    for (int k = 0; k < 2; ++k) {
        if (n_paths_out) n_paths_out[k] = m->ctx[0]->maps[k].n_paths;
        if (n_stored_out) n_stored_out[k] = m->ctx[0]->maps[k].n;
    }
    return B2RT_OK;
}

int b2rt_multi_photon_grid_build(b2rt_multi *m, int32_t map, float cell_size, float target_per_cell) {
    if (!m) return B2RT_E_ARG;
    const double t0 = now_ms();
    int rc = parallel(m, [&](int i) -> int {
        int r = b2rt_photon_grid_build(m->ctx[i], map, cell_size, target_per_cell);
        if (r) return r;
        B2RT_CUDA(m->ctx[i], cudaStreamSynchronize(m->ctx[i]->stream));
        return B2RT_OK;
    });
    m->stats.grid_ms += (float)(now_ms() - t0);
    return rc;
}

int b2rt_multi_radiance_compute(b2rt_multi *m, int32_t nused, float max_dist2) {
    if (!m) return B2RT_E_ARG;
    return parallel(m, [&](int i) -> int {
        int r = b2rt_radiance_compute(m->ctx[i], nused, max_dist2);
        if (r) return r;
        B2RT_CUDA(m->ctx[i], cudaStreamSynchronize(m->ctx[i]->stream));
        return B2RT_OK;
    });
}

int b2rt_multi_fg_deposit_limits(b2rt_multi *m, int64_t caustic_paths, int64_t indirect_paths) {
    if (!m) return B2RT_E_ARG;
    for (b2rt_ctx *c : m->ctx) b2rt_fg_deposit_limits(c, caustic_paths, indirect_paths);
    return B2RT_OK;
}
int b2rt_multi_final_gather_set(b2rt_multi *m, int32_t enable, int32_t gather_samples, float gather_angle_deg) {
    if (!m) return B2RT_E_ARG;
    for (b2rt_ctx *c : m->ctx) {
        int rc = b2rt_final_gather_set(c, enable, gather_samples, gather_angle_deg);
        if (rc) { m->err = c->err; return rc; }
    }
    return B2RT_OK;
}
int b2rt_multi_sampler_set(b2rt_multi *m, int32_t type, int32_t xsamples, int32_t ysamples) {
    if (!m) return B2RT_E_ARG;
    for (b2rt_ctx *c : m->ctx) {
        int rc = b2rt_sampler_set(c, type, xsamples, ysamples);
        if (rc) { m->err = c->err; return rc; }
    }
    return B2RT_OK;
}
int b2rt_multi_film_set_filter(b2rt_multi *m, const b2rt_filter *filter) {
    if (!m) return B2RT_E_ARG;
    for (b2rt_ctx *c : m->ctx) {
        int rc = b2rt_film_set_filter(c, filter);
        if (rc) { m->err = c->err; return rc; }
    }
    return B2RT_OK;
}

// 64 x 64 tiles of the crop window, tile t -> GPU t mod G; per GPU the pixels tile after tile, row-major inside a tile
// tile edge in pixels: 64; B2RT_MGPU_TILE = 16 / 32 / 64 / 128 for A/B runs of the load balance
static int tile_edge() {
    int T = 64;
    if (const char *v = getenv("B2RT_MGPU_TILE")) { const int t = atoi(v); if (t == 16 || t == 32 || t == 64 || t == 128) T = t; }
    return T;
}
static void tile_pixel_lists(int G, const b2rt_frame_params *fp, std::vector<std::vector<uint32_t>> &lists) {
    const int T = tile_edge();
    lists.assign(G, {});
    const int ntx = (fp->x1 - fp->x0 + T - 1) / T, nty = (fp->y1 - fp->y0 + T - 1) / T;
    for (int ty = 0; ty < nty; ++ty)
        for (int tx = 0; tx < ntx; ++tx) {
            std::vector<uint32_t> &L = lists[(ty * ntx + tx) % G];
            const int xa = fp->x0 + tx * T, xb = std::min(xa + T, (int)fp->x1), ya = fp->y0 + ty * T, yb = std::min(ya + T, (int)fp->y1);
            for (int y = ya; y < yb; ++y)
                for (int x = xa; x < xb; ++x) L.push_back((uint32_t)x | ((uint32_t)y << 16));
        }
}

int b2rt_multi_render_frame(b2rt_multi *m, const b2rt_camera *cam, const b2rt_frame_params *fp, float *out_rgb, int where) {
    if (!m) return B2RT_E_ARG;
    if (!cam || !fp || !out_rgb || fp->spp < 1 || fp->x0 < 0 || fp->y0 < 0 || fp->x1 > cam->xres || fp->y1 > cam->yres ||
        fp->x0 >= fp->x1 || fp->y0 >= fp->y1 || fp->nused > 224 || !(fp->max_dist2 > 0.f) || cam->xres > 65535 || cam->yres > 65535) {
        m->err = "multi_render_frame: bad arguments"; return B2RT_E_ARG;
    }
    if (where != B2RT_HOST && where != B2RT_DEVICE) { m->err = "multi_render_frame: where must be B2RT_HOST or B2RT_DEVICE (on GPU 0)"; return B2RT_E_ARG; }
    int rc = check_state(m, "multi_render_frame");
    if (rc) return rc;
    const int G = m->n;
    b2rt_ctx *c0 = m->ctx[0];
    const int cw = fp->x1 - fp->x0, chh = fp->y1 - fp->y0;
    const size_t full_bytes = (size_t)cam->xres * cam->yres * 12, crop_bytes = (size_t)cw * chh * 12;
    const b2rt_filter &flt = c0->film_filter;
    const bool filtered = !(flt.type == B2RT_FILTER_BOX && flt.xwidth <= 0.5f && flt.ywidth <= 0.5f);
    const double t0 = now_ms();
    cudaSetDevice(m->dev[0]);
    if (!filtered && (m->peer || G == 1)) {
        // ---- tiles: every GPU renders its pixel list, pixels land in the whole image on GPU 0 -----------------------
        if (m->frame0->reserve(full_bytes) != cudaSuccess) { m->err = "multi_render_frame: out of memory on GPU 0"; return B2RT_E_CUDA; }
        const bool fresh = m->pm_tile != tile_edge() || m->pm_xres != cam->xres || m->pm_yres != cam->yres || m->pm_win[0] != fp->x0 || m->pm_win[1] != fp->x1 ||
                           m->pm_win[2] != fp->y0 || m->pm_win[3] != fp->y1;
        std::vector<std::vector<uint32_t>> lists;
        if (fresh) tile_pixel_lists(G, fp, lists);
        float *full = m->frame0->as<float>();
        // Whole image into page-locked host memory (b2rt_host_alloc, cudaHostAlloc, cudaHostRegister: portable and mapped
        // under unified addressing): every GPU stores its pixels straight into the caller's buffer over its own PCIe link
        // instead of GPU 0 collecting the frame and downloading 12 B / pixel in one serial copy at the end.  Same bytes.
        // Measured on the Cfg3 camera pass (r08i / r08j): 8 GPUs 3.39 -> 3.00 ms, 4 GPUs 5.18 -> 4.97, 2 GPUs 9.08 -> 9.18,
        // 1 GPU 16.73 -> 17.38 (one GPU's 12-byte stores over one PCIe link lose against one DMA copy): on from 4 GPUs.
        // B2RT_MGPU_HOSTFRAME=0 / 1 forces it off / on (A/B).
        bool direct = false;
        if (where == B2RT_HOST && cw == cam->xres && chh == cam->yres) {
            const char *e = getenv("B2RT_MGPU_HOSTFRAME");
            const bool want = e ? atoi(e) != 0 : G >= 4;
            cudaPointerAttributes at;
            if (want && cudaPointerGetAttributes(&at, out_rgb) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
                full = static_cast<float *>(at.devicePointer);
                direct = true;
            }
            cudaGetLastError();
        }
        rc = parallel(m, [&](int i) -> int {
            b2rt_ctx *c = m->ctx[i];
            if (fresh) {
                m->pixmap_n[i] = (int64_t)lists[i].size();
                B2RT_CUDA(c, m->pixmaps[i]->reserve(std::max<size_t>(lists[i].size() * 4, 16)));
                if (!lists[i].empty())
                    B2RT_CUDA(c, cudaMemcpyAsync(m->pixmaps[i]->p, lists[i].data(), lists[i].size() * 4, cudaMemcpyHostToDevice, c->stream));
            }
            int r = b2rt_render_pixels_impl(c, cam, fp, m->pixmaps[i]->as<uint32_t>(), m->pixmap_n[i], full);
            if (r) return r;
            return b2rt_sync(c);                 // also reports a traversal-stack overflow
        });
        if (rc) return rc;
        m->pm_xres = cam->xres; m->pm_yres = cam->yres; m->pm_tile = tile_edge();
        m->pm_win[0] = fp->x0; m->pm_win[1] = fp->x1; m->pm_win[2] = fp->y0; m->pm_win[3] = fp->y1;
        // the crop window out of the whole image (row pitch xres -> cw)
        if (!direct) {
            const float *src = full + 3 * ((size_t)fp->y0 * cam->xres + fp->x0);
            B2RT_CUDA(c0, cudaMemcpy2DAsync(out_rgb, (size_t)cw * 12, src, (size_t)cam->xres * 12, (size_t)cw * 12, chh,
                                            where == B2RT_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, c0->stream));
            B2RT_CUDA(c0, cudaStreamSynchronize(c0->stream));
        }
    } else {
        // ---- bands: one row band of the crop window per GPU through the crop-window path (halo cells re-traced) --------
        if (m->frame0->reserve(crop_bytes) != cudaSuccess) { m->err = "multi_render_frame: out of memory on GPU 0"; return B2RT_E_CUDA; }
        rc = parallel(m, [&](int i) -> int {
            b2rt_ctx *c = m->ctx[i];
            b2rt_frame_params p = *fp;
            p.y0 = fp->y0 + (int)((int64_t)chh * i / G); p.y1 = fp->y0 + (int)((int64_t)chh * (i + 1) / G);
            if (p.y1 <= p.y0) return (int)B2RT_OK;
            const size_t bytes = (size_t)cw * (p.y1 - p.y0) * 12;
            float *dst0 = m->frame0->as<float>() + 3 * (size_t)cw * (p.y0 - fp->y0);
            float *mine = dst0;
            if (i != 0) { B2RT_CUDA(c, m->bands[i]->reserve(bytes)); mine = m->bands[i]->as<float>(); }
            int r = b2rt_render_frame(c, cam, &p, mine, B2RT_DEVICE);
            if (r) return r;
            if (i != 0) B2RT_CUDA(c, cudaMemcpyPeerAsync(dst0, m->dev[0], mine, m->dev[i], bytes, c->stream));
            B2RT_CUDA(c, cudaStreamSynchronize(c->stream));
            return (int)B2RT_OK;
        });
        if (rc) return rc;
        B2RT_CUDA(c0, cudaMemcpyAsync(out_rgb, m->frame0->p, crop_bytes, where == B2RT_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice,
                                      c0->stream));
        B2RT_CUDA(c0, cudaStreamSynchronize(c0->stream));
    }
    m->stats.render_ms += (float)(now_ms() - t0);
    return B2RT_OK;
}

}  // extern "C"

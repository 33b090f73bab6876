// K2: hand-written, CUB-free, stable LSD radix sort of (key32, value32) pairs + a generic
// exclusive scan.  Replaces nothing in the reference directly: the reference sorts nothing on the
// device -- its spatial structures are built by closed OptiX ("Sbvh", cuda_render/cudarender.cpp:44)
// and by pbrt's CPU KdTree (cuda_render/photon_mapping/photonmappingrenderer.cpp:150-180).
//
// Determinism: every step is integer arithmetic whose result does not depend on scheduling --
// per-block digit histograms, an exclusive scan, and a stable in-block ranking (warp w owns a
// contiguous sub-tile, lanes rank by __match_any_sync + popc of lower lanes).  Equal keys keep
// their input order, so with value = original index ties break by index.
#include "b2rt_internal.cuh"
#include <algorithm>
#include <mutex>
#include <utility>
#include <type_traits>

namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = 16;                       // keys per thread
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS; // 4096 keys per block
constexpr int RADIX = 256;
static_assert(SORT_THREADS == RADIX, "the scatter kernel maps one thread to one digit");

// ---- per-block digit histogram --------------------------------------------------------------
// One 16-byte load per thread per round (the tile is 4096 keys = 4 rounds of 256 threads x 4 keys); the cudaMalloc'ed
// key arrays are 256-byte aligned and a tile starts at a multiple of 4096 keys, so the vector loads are aligned.
// The histogram does not depend on the order in which a block's keys are counted.
__global__ void __launch_bounds__(SORT_THREADS)
radix_hist_kernel(const uint32_t *__restrict__ keys, int64_t n, int shift, uint32_t *__restrict__ hist,
                  int nblocks) {
    __shared__ uint32_t sh[SORT_WARPS][RADIX];
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&sh[0][0])[i] = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    const int64_t tile0 = (int64_t)blockIdx.x * SORT_TILE;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(keys) & 15) == 0;
    if (vec_ok && tile0 + SORT_TILE <= n) {
        const uint4 *kp = reinterpret_cast<const uint4 *>(keys + tile0);
        uint4 v[SORT_ITEMS / 4];
#pragma unroll
        for (int r = 0; r < SORT_ITEMS / 4; ++r) v[r] = __ldg(kp + r * SORT_THREADS + threadIdx.x);
#pragma unroll
        for (int r = 0; r < SORT_ITEMS / 4; ++r) {
            atomicAdd(&sh[warp][(v[r].x >> shift) & 0xff], 1u);
            atomicAdd(&sh[warp][(v[r].y >> shift) & 0xff], 1u);
            atomicAdd(&sh[warp][(v[r].z >> shift) & 0xff], 1u);
            atomicAdd(&sh[warp][(v[r].w >> shift) & 0xff], 1u);
        }
    } else {
        for (int64_t i = tile0 + threadIdx.x; i < n && i < tile0 + SORT_TILE; i += SORT_THREADS)
            atomicAdd(&sh[warp][(keys[i] >> shift) & 0xff], 1u);
    }
    __syncthreads();
    for (int d = threadIdx.x; d < RADIX; d += SORT_THREADS) {
        uint32_t s = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) s += sh[w][d];
        hist[(int64_t)d * nblocks + blockIdx.x] = s;
    }
}

// ---- stable scatter ---------------------------------------------------------------------------
// Ranks are computed as before (warp w owns a contiguous sub-tile; lanes rank by __match_any_sync + popc of lower
// lanes), but the tile is first permuted into digit order in shared memory and then written out by consecutive
// threads: all elements of one digit of one block are contiguous in the output, so the global stores are coalesced
// runs instead of 4-byte scatters (ncu r01b: the direct scatter reached ~0.8 TB/s of the 6.5 TB/s copy peak).
// RANK: how a lane finds the lanes of its warp that hold the same digit: 0 = match.any, 1 = eight ballots (one per digit
// bit) combined with and / not (default), 5 = a per-warp table of lane masks in shared memory (atomicOr, sync, read).
// FULL: the tile is complete (no bounds checks; the tail tile takes the other instantiation).
// Measured on B200, 2^26 pairs (profiles/r06h..r06k): match.any 2.50-2.80 ms, ballots 1.92-1.97 ms, shared-memory table 2.03 ms,
// a mix of match.any for some items and ballots for the rest 1.95-2.06 ms (worse the more match.any): MATCH is the slow
// instruction here, and once it is gone two very different rankings cost the same -- what is left (ncu r06i: issue-active
// 47 %, 4 blocks of 64 registers per SM, five block barriers per tile, the 16 ordered counter rounds) is not the peer search.
template <int RANK, bool FULL = false>
__device__ __forceinline__ uint32_t digit_peers(uint32_t d, bool ok) {
    if (RANK == 0) return __match_any_sync(0xffffffffu, (FULL || ok) ? d : 0x100u);
    uint32_t peers = 0xffffffffu;
    if (!FULL) {
        peers = __ballot_sync(0xffffffffu, ok);
        if (!ok) peers = ~peers;
    }
    // per bit: one LOP3 with a predicate result, one VOTE, one select-and (the sign-extended bit picks m or ~m)
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const uint32_t m = __ballot_sync(0xffffffffu, (d & (1u << b)) != 0u);
        const uint32_t s = (uint32_t)((int32_t)(d << (31 - b)) >> 31);     // all ones where the bit is set
        peers &= ~(m ^ s);
    }
    return peers;
}

#ifndef B2RT_SCATTER_BLOCKS
#define B2RT_SCATTER_BLOCKS 4
#endif
template <int RANK, bool FULL>
__global__ void __launch_bounds__(SORT_THREADS, B2RT_SCATTER_BLOCKS)
radix_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                     uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, int64_t n, int shift,
                     const uint32_t *__restrict__ scanned_hist, int nblocks, int tile_first) {
    __shared__ uint32_t sh[SORT_WARPS][RADIX];  // first: running warp histogram; then: exclusive count of earlier warps
    __shared__ uint32_t s_off[RADIX];           // first tile slot of digit d
    __shared__ uint32_t s_gadj[RADIX];          // global position of digit d's first element minus s_off[d]
    __shared__ uint32_t s_wsum[SORT_WARPS];
    __shared__ uint32_t s_key[SORT_TILE], s_val[SORT_TILE];
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&sh[0][0])[i] = 0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const int tile = tile_first + blockIdx.x;
    const int64_t tile0 = (int64_t)tile * SORT_TILE;
    const int64_t base = tile0 + warp * (32 * SORT_ITEMS);
    const int tile_n = FULL ? SORT_TILE : (int)(n - tile0 < SORT_TILE ? n - tile0 : SORT_TILE);
    const int wn = FULL ? 32 * SORT_ITEMS : max(0, min(32 * SORT_ITEMS, tile_n - warp * (32 * SORT_ITEMS)));   // valid keys of this warp
    const uint32_t *kp = keys_in + base + lane, *vp = vals_in + base + lane;

    // Three loops instead of one: the 16 loads and the 16 peer searches are independent of each other and pipeline; only
    // the counter updates are ordered (item r before item r+1 of the same digit: that order is the sort's stability).
    uint32_t k[SORT_ITEMS], rank[SORT_ITEMS], peers[SORT_ITEMS];
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) k[r] = (FULL || r * 32 + lane < wn) ? kp[r * 32] : 0u;
    if (RANK == 5) {
        // Digit peers through shared memory instead of 8 ballots (~40 ALU instructions) per key: a per-warp table of lane
        // masks indexed by digit -- every lane ORs its bit into T[d], the warp syncs, T[d] IS the peer mask.  Two tables
        // used in turn; a lane takes its own bit of round r-1 out again (atomicAnd) after the sync of round r, when every
        // lane has read it, so nothing has to be cleared wholesale.  The tables live where s_key is built later (zeroed
        // here; only complete tiles take this path).  Results are the same integers as with match.any / ballots.
        static_assert(2 * SORT_WARPS * RADIX <= SORT_TILE, "the mask tables alias s_key");
        for (int i = threadIdx.x; i < 2 * SORT_WARPS * RADIX; i += SORT_THREADS) s_key[i] = 0;
        __syncthreads();
        uint32_t *T0 = s_key + warp * RADIX, *T1 = s_key + (SORT_WARPS + warp) * RADIX;
        const uint32_t me = 1u << lane;
        uint32_t dprev = 0;
#pragma unroll
        for (int r = 0; r < SORT_ITEMS; ++r) {
            const uint32_t d = (k[r] >> shift) & 0xff;
            uint32_t *T = (r & 1) ? T1 : T0, *Tp = (r & 1) ? T0 : T1;
            atomicOr(&T[d], me);
            __syncwarp();
            const uint32_t pr = T[d];
            if (r > 0) atomicAnd(&Tp[dprev], ~me);
            dprev = d;
            const int leader = __ffs(pr) - 1;
            uint32_t before = 0;
            if (lane == leader) {
                before = sh[warp][d];
                sh[warp][d] = before + __popc(pr);
            }
            before = __shfl_sync(0xffffffffu, before, leader);
            rank[r] = before + __popc(pr & lt);
        }
        __syncthreads();
    } else {
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r)
        peers[r] = digit_peers<(RANK == 0 ? 0 : 1), FULL>((k[r] >> shift) & 0xff, FULL || r * 32 + lane < wn);
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        const bool ok = FULL || r * 32 + lane < wn;
        const uint32_t d = (k[r] >> shift) & 0xff;
        const int leader = __ffs(peers[r]) - 1;
        uint32_t before = 0;
        if (lane == leader && ok) {
            before = sh[warp][d];
            sh[warp][d] = before + __popc(peers[r]);
        }
        before = __shfl_sync(0xffffffffu, before, leader);
        rank[r] = before + __popc(peers[r] & lt);
        __syncwarp();
    }
    __syncthreads();
    }
    // thread d: per-warp counts of digit d -> exclusive over warps; block total -> exclusive scan over the digits
    {
        const int d = threadIdx.x;          // SORT_THREADS == RADIX
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) {
            uint32_t c = sh[w][d];
            sh[w][d] = run;
            run += c;
        }
        uint32_t inc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_wsum[warp] = inc;
        __syncthreads();
        uint32_t wbase = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w)
            if (w < warp) wbase += s_wsum[w];
        const uint32_t off = wbase + inc - run;
        s_off[d] = off;
        s_gadj[d] = scanned_hist[(int64_t)d * nblocks + tile] - off;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        if (FULL || r * 32 + lane < wn) {
            uint32_t d = (k[r] >> shift) & 0xff;
            uint32_t lp = s_off[d] + sh[warp][d] + rank[r];
            s_key[lp] = k[r];
            s_val[lp] = vp[r * 32];
        }
    }
    __syncthreads();
    if (FULL) {
#pragma unroll 4
        for (int i = threadIdx.x; i < SORT_TILE; i += SORT_THREADS) {
            const uint32_t key = s_key[i];
            const uint32_t pos = s_gadj[(key >> shift) & 0xff] + (uint32_t)i;
            keys_out[pos] = key;
            vals_out[pos] = s_val[i];
        }
    } else {
        for (int i = threadIdx.x; i < tile_n; i += SORT_THREADS) {
            uint32_t key = s_key[i];
            uint32_t pos = s_gadj[(key >> shift) & 0xff] + (uint32_t)i;
            keys_out[pos] = key;
            vals_out[pos] = s_val[i];
        }
    }
}

// ---- one-sweep variant (default): ONE histogram pass for all digits + one scatter launch per digit --------------------
// After Adinets & Merrill, "Onesweep" (2022), restated for this tile / ranking scheme.  The three-launch pass above reads
// the keys twice per pass and round-trips a per-block histogram through memory and a multi-launch scan: 20 P bytes per
// pair and ~5 launches per pass where 4 + 16 P bytes and one launch are needed (68 B/pair for P = 4).  Here
//   * onesweep_hist_kernel reads the keys once and counts every digit of every pass (P x 256 global counters);
//   * onesweep_scatter_kernel does one pass per launch.  A block takes the next tile from a global ticket counter (so a
//     block only ever waits for tiles that were handed out before its own: they are running or done, no deadlock whatever
//     the launch order of blocks), ranks its 4096 keys exactly as radix_scatter_kernel does, publishes its 256 digit counts
//     and finds its own start per digit by DECOUPLED LOOK-BACK over the tiles before it: state word = flag << 30 | value,
//     flag 1 = the tile's own count, flag 2 = inclusive prefix over all tiles up to it.  Thread d walks digit d.
// (Tried and dropped, r04d: ranking by shared-memory atomic tickets + REDUX min over the peer mask instead of leader +
// shuffle -- same bytes, 2x slower at every size: 5.9 against 2.9 ms at 2^26.)
// Determinism: the output position of a key is (keys with a smaller digit) + (same digit in earlier TILES) + (its stable
// rank inside the tile) -- integers that do not depend on which block ran the tile or when.  Stability as before.
constexpr uint32_t OS_FLAG_AGG = 1u << 30, OS_FLAG_INC = 2u << 30, OS_VALUE = (1u << 30) - 1u;
constexpr int OS_MAX_PASSES = 4;

__global__ void __launch_bounds__(SORT_THREADS)
onesweep_hist_kernel(const uint32_t *__restrict__ keys, int64_t n, int passes, uint32_t *__restrict__ ghist) {
    __shared__ uint32_t sh[OS_MAX_PASSES][RADIX];
    for (int i = threadIdx.x; i < OS_MAX_PASSES * RADIX; i += SORT_THREADS) (&sh[0][0])[i] = 0;
    __syncthreads();
    // The top digit of a short key takes few values (a 26-bit cell key: 4): plain shared atomics would serialise 32-way
    // inside every warp.  Its count is aggregated per warp first (match.any, the leader adds the group size); the lower
    // digits are spread out and take the atomic directly.  Whole warps iterate together (`ok` masks the tail).
    const bool vec_ok = (reinterpret_cast<uintptr_t>(keys) & 15) == 0;
    const int64_t n4 = vec_ok ? n / 4 : 0;
    const uint4 *kp = reinterpret_cast<const uint4 *>(keys);
    const int lane = threadIdx.x & 31, top = passes - 1;
    const int64_t stride = (int64_t)gridDim.x * SORT_THREADS;
    auto count = [&](uint32_t k, bool ok) {
        for (int p = 0; p < top; ++p)
            if (ok) atomicAdd(&sh[p][(k >> (8 * p)) & 0xff], 1u);
        const uint32_t d = ok ? ((k >> (8 * top)) & 0xff) : 0x100u;
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        if (ok && lane == __ffs(peers) - 1) atomicAdd(&sh[top][d], (uint32_t)__popc(peers));
    };
    const int64_t start = (int64_t)blockIdx.x * SORT_THREADS + (threadIdx.x & ~31);      // the warp's first index
    for (int64_t w = start; w < n4; w += stride) {
        const int64_t i = w + lane;
        const bool ok = i < n4;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (ok) v = __ldg(kp + i);
        count(v.x, ok); count(v.y, ok); count(v.z, ok); count(v.w, ok);
    }
    for (int64_t w = 4 * n4 + start; w < n; w += stride) {
        const int64_t i = w + lane;
        const bool ok = i < n;
        count(ok ? keys[i] : 0u, ok);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RADIX; i += SORT_THREADS) {
        const uint32_t c = (&sh[0][0])[i];
        if (c) atomicAdd(&ghist[i], c);
    }
}

__global__ void __launch_bounds__(SORT_THREADS, 4)
onesweep_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ vals_in,
                        uint32_t *__restrict__ keys_out, uint32_t *__restrict__ vals_out, int64_t n, int shift,
                        const uint32_t *__restrict__ ghist /* [256] of this pass */, uint32_t *__restrict__ state /* [tiles][256] */,
                        uint32_t *__restrict__ ticket) {
    __shared__ uint32_t sh[SORT_WARPS][RADIX];
    __shared__ uint32_t s_off[RADIX];
    __shared__ uint32_t s_gadj[RADIX];
    __shared__ uint32_t s_wsum[SORT_WARPS];
    __shared__ uint32_t s_key[SORT_TILE], s_val[SORT_TILE];
    __shared__ uint32_t s_tile;
    for (int i = threadIdx.x; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&sh[0][0])[i] = 0;
    if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const int64_t tile0 = (int64_t)tile * SORT_TILE;
    const int64_t base = tile0 + warp * (32 * SORT_ITEMS);
    const int tile_n = (int)(n - tile0 < SORT_TILE ? n - tile0 : SORT_TILE);

    uint32_t k[SORT_ITEMS], rank[SORT_ITEMS], peers[SORT_ITEMS];
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        int64_t i = base + r * 32 + lane;
        k[r] = i < n ? keys_in[i] : 0u;
    }
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) peers[r] = digit_peers<1>((k[r] >> shift) & 0xff, base + r * 32 + lane < n);
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        const bool ok = base + r * 32 + lane < n;
        const uint32_t d = (k[r] >> shift) & 0xff;
        const int leader = __ffs(peers[r]) - 1;
        uint32_t before = 0;
        if (lane == leader && ok) {
            before = sh[warp][d];
            sh[warp][d] = before + __popc(peers[r]);
        }
        before = __shfl_sync(0xffffffffu, before, leader);
        rank[r] = before + __popc(peers[r] & lt);
        __syncwarp();
    }
    __syncthreads();
    {
        const int d = threadIdx.x;          // SORT_THREADS == RADIX: thread d owns digit d
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) {
            uint32_t c = sh[w][d];
            sh[w][d] = run;
            run += c;
        }
        // publish this tile's count of digit d at once, then look back
        volatile uint32_t *st = state;
        const size_t me = (size_t)tile * RADIX + d;
        if (tile == 0) st[me] = OS_FLAG_INC | run;
        else st[me] = OS_FLAG_AGG | run;
        // block-wide exclusive scans over the digits: of the tile's counts (slot of digit d in the shared tile) and of
        // the global histogram (first output position of digit d)
        uint32_t inc = run, ginc = ghist[d];
        const uint32_t gcount = ginc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, o), g = __shfl_up_sync(0xffffffffu, ginc, o);
            if (lane >= o) { inc += t; ginc += g; }
        }
        __shared__ uint32_t s_gsum[SORT_WARPS];
        if (lane == 31) { s_wsum[warp] = inc; s_gsum[warp] = ginc; }
        __syncthreads();
        uint32_t wbase = 0, gbase = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w)
            if (w < warp) { wbase += s_wsum[w]; gbase += s_gsum[w]; }
        const uint32_t off = wbase + inc - run;
        const uint32_t gfirst = gbase + ginc - gcount;
        uint32_t excl = 0;                                  // digit d in the tiles before this one
        if (tile > 0) {
            size_t j = me - RADIX;
            for (;;) {
                const uint32_t v = st[j];
                const uint32_t f = v & ~OS_VALUE;
                if (f == 0u) continue;                      // not published yet: that tile is running, poll again
                excl += v & OS_VALUE;
                if (f == OS_FLAG_INC) break;
                j -= RADIX;
            }
            st[me] = OS_FLAG_INC | (excl + run);
        }
        s_off[d] = off;
        s_gadj[d] = gfirst + excl - off;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        int64_t i = base + r * 32 + lane;
        if (i < n) {
            uint32_t d = (k[r] >> shift) & 0xff;
            uint32_t lp = s_off[d] + sh[warp][d] + rank[r];
            s_key[lp] = k[r];
            s_val[lp] = vals_in[i];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < tile_n; i += SORT_THREADS) {
        uint32_t key = s_key[i];
        uint32_t pos = s_gadj[(key >> shift) & 0xff] + (uint32_t)i;
        keys_out[pos] = key;
        vals_out[pos] = s_val[i];
    }
}

// ---- exclusive scan (reduce-then-scan, recursive) ------------------------------------------------
constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS)
scan_tile_kernel(uint32_t *__restrict__ data, int64_t n, uint32_t *__restrict__ block_sums) {
    __shared__ uint32_t warp_sums[SCAN_THREADS / 32];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    uint32_t x[SCAN_ITEMS];
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        x[j] = (base + j < n) ? data[base + j] : 0u;
        sum += x[j];
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < SCAN_THREADS / 32) ? warp_sums[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += t;
        }
        if (lane < SCAN_THREADS / 32) warp_sums[lane] = winc - w;
        if (lane == SCAN_THREADS / 32 - 1 && block_sums) block_sums[blockIdx.x] = winc;
    }
    __syncthreads();
    uint32_t run = warp_sums[warp] + inc - sum;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j) {
        if (base + j < n) data[base + j] = run;
        run += x[j];
    }
}

__global__ void __launch_bounds__(SCAN_THREADS)
scan_add_kernel(uint32_t *__restrict__ data, int64_t n, const uint32_t *__restrict__ block_offsets) {
    uint32_t off = block_offsets[blockIdx.x];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int j = 0; j < SCAN_ITEMS; ++j)
        if (base + j < n) data[base + j] += off;
}

// the complete tiles in one launch, the ragged last tile (if any) in a second one-block launch of the checked instantiation
int launch_scatter(b2rt_ctx *ctx, const uint32_t *kin, const uint32_t *vin, uint32_t *kout, uint32_t *vout, int64_t n, int shift,
                   const uint32_t *hist, int nblocks) {
    const int full = (int)(n / SORT_TILE);
    auto launch = [&](auto rank) {
        constexpr int R = decltype(rank)::value;
        if (full) radix_scatter_kernel<R, true><<<full, SORT_THREADS, 0, ctx->stream>>>(kin, vin, kout, vout, n, shift, hist, nblocks, 0);
        if (full < nblocks) radix_scatter_kernel<(R == 0 ? 0 : 1), false><<<1, SORT_THREADS, 0, ctx->stream>>>(kin, vin, kout, vout, n, shift, hist, nblocks, full);
    };
    switch (ctx->sort_rank) {
        case 0: launch(std::integral_constant<int, 0>()); break;
        case 5: launch(std::integral_constant<int, 5>()); break;
        default: launch(std::integral_constant<int, 1>()); break;
    }
    B2RT_LAUNCH_CHECK(ctx);
    if (full && full < nblocks) ctx->launches++;
    return B2RT_OK;
}

int scan_rec(b2rt_ctx *ctx, uint32_t *d, int64_t n, uint32_t *tmp, uint32_t *d_total) {
    int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (nb <= 1) {
        scan_tile_kernel<<<1, SCAN_THREADS, 0, ctx->stream>>>(d, n, d_total);
        B2RT_LAUNCH_CHECK(ctx);
        return B2RT_OK;
    }
    scan_tile_kernel<<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(d, n, tmp);
    B2RT_LAUNCH_CHECK(ctx);
    int rc = scan_rec(ctx, tmp, nb, tmp + ((nb + 3) & ~3ll), d_total);
    if (rc) return rc;
    scan_add_kernel<<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(d, n, tmp);
    B2RT_LAUNCH_CHECK(ctx);
    return B2RT_OK;
}

}  // namespace

int b2rt_exclusive_scan_u32(b2rt_ctx *ctx, uint32_t *d, int64_t n, uint32_t *d_total) {
    if (n <= 0) return B2RT_OK;
    // temp: sum over levels of ceil(n/TILE^k) (+ padding)
    size_t need = 0;
    for (int64_t m = (n + SCAN_TILE - 1) / SCAN_TILE; m > 1; m = (m + SCAN_TILE - 1) / SCAN_TILE)
        need += ((m + 3) & ~3ll);
    need += 8;
    B2RT_CUDA(ctx, ctx->d_scan_tmp.reserve(need * sizeof(uint32_t)));
    return scan_rec(ctx, d, n, ctx->d_scan_tmp.as<uint32_t>(), d_total);
}

int b2rt_sort_pairs_impl(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, uint32_t *keys_alt,
                         uint32_t *vals_alt, int key_bits, bool *result_in_alt) {
    *result_in_alt = false;
    if (n <= 1) return B2RT_OK;
    if (n >= (1ll << 32)) { ctx->err = "sort: n must be < 2^32"; return B2RT_E_ARG; }
    int passes = (key_bits + 7) / 8;
    int nblocks = (int)((n + SORT_TILE - 1) / SORT_TILE);
    uint32_t *kin = keys, *vin = vals, *kout = keys_alt, *vout = vals_alt;
    // Measured on B200 (profiles/r04c_stage_rooflines*.json): one-sweep wins while the sort is launch-bound (2^20 pairs: 0.115
    // against 0.138 ms) and loses for large inputs (2^26: 3.2 against 2.5 ms) -- at ~40 tiles per microsecond the per-digit
    // look-back (one thread walks one digit's chain, one L2 round trip per predecessor) cannot keep up and the tiles
    // convoy.  sort_variant 1 = one-sweep below 2^22 pairs, 2 = always, 0 = never.
    const bool onesweep = ctx->sort_variant == 2 || (ctx->sort_variant == 1 && n < (1ll << 22));
    if (onesweep && n < (1ll << 30) && passes <= OS_MAX_PASSES) {
        // one-sweep: [passes * 256 digit counts][passes tickets, padded to 64][passes * tiles * 256 look-back states]
        const size_t head = (size_t)OS_MAX_PASSES * RADIX + 64;
        const size_t words = head + (size_t)passes * nblocks * RADIX;
        B2RT_CUDA(ctx, ctx->d_hist.reserve(words * sizeof(uint32_t)));
        uint32_t *ghist = ctx->d_hist.as<uint32_t>(), *tickets = ghist + OS_MAX_PASSES * RADIX, *states = ghist + head;
        B2RT_CUDA(ctx, cudaMemsetAsync(ghist, 0, words * sizeof(uint32_t), ctx->stream));
        const int hblocks = (int)std::min<int64_t>((n + 4 * SORT_THREADS - 1) / (4 * SORT_THREADS), (int64_t)ctx->sm_count * 8);
        onesweep_hist_kernel<<<hblocks, SORT_THREADS, 0, ctx->stream>>>(kin, n, passes, ghist);
        B2RT_LAUNCH_CHECK(ctx);
        for (int p = 0; p < passes; ++p) {
            onesweep_scatter_kernel<<<nblocks, SORT_THREADS, 0, ctx->stream>>>(kin, vin, kout, vout, n, 8 * p, ghist + p * RADIX,
                                                                              states + (size_t)p * nblocks * RADIX, tickets + p);
            B2RT_LAUNCH_CHECK(ctx);
            uint32_t *t;
            t = kin; kin = kout; kout = t;
            t = vin; vin = vout; vout = t;
        }
        *result_in_alt = (passes & 1) != 0;
        return B2RT_OK;
    }
    B2RT_CUDA(ctx, ctx->d_hist.reserve((size_t)RADIX * nblocks * sizeof(uint32_t)));
    uint32_t *hist = ctx->d_hist.as<uint32_t>();
    for (int p = 0; p < passes; ++p) {
        int shift = 8 * p;
        radix_hist_kernel<<<nblocks, SORT_THREADS, 0, ctx->stream>>>(kin, n, shift, hist, nblocks);
        B2RT_LAUNCH_CHECK(ctx);
        int rc = b2rt_exclusive_scan_u32(ctx, hist, (int64_t)RADIX * nblocks, nullptr);
        if (rc) return rc;
        rc = launch_scatter(ctx, kin, vin, kout, vout, n, shift, hist, nblocks);
        if (rc) return rc;
        uint32_t *t;
        t = kin; kin = kout; kout = t;
        t = vin; vin = vout; vout = t;
    }
    *result_in_alt = (passes & 1) != 0;
    return B2RT_OK;
}

// ---- b2rt_sort_pairs_noalloc: the same sort with its scratch in a static device array ---------------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart
// in the reference.  Same kernels, same schedule choice (ctx->sort_variant) and
// therefore the same bytes as b2rt_sort_pairs; only where the ping-pong arrays, histograms / look-back states and scan
// temporaries live differs: g_noalloc_scratch below, a file-scope array that every GPU that loads the library holds until
// the process exits (9.1 MB).  This path calls no cudaMalloc / cudaFree / DevBuf::reserve.
//
// The array's name is only a host-side stand-in in host code, not a device address: the address the kernels receive comes
// from cudaGetSymbolAddress, taken once per device with that device current (noalloc_scratch_ptr).
// Every call on a device shares the array: two calls that overlap on the GPU (different streams / contexts of one device)
// race.  The array keeps the previous call's contents; what must start at zero (the one-sweep digit counts, tickets and
// look-back states) is cleared by a cudaMemsetAsync in every call, everything else is written before it is read.
namespace {

constexpr int64_t NOALLOC_MAX_PAIRS = B2RT_SORT_NOALLOC_MAX;                      // 2^20 pairs
constexpr int64_t NOALLOC_TILES = (NOALLOC_MAX_PAIRS + SORT_TILE - 1) / SORT_TILE;
constexpr size_t NOALLOC_OS_HEAD = (size_t)OS_MAX_PASSES * RADIX + 64;           // digit counts + tickets (one-sweep)
// hist region: max(one-sweep: head + passes * tiles * 256 states, three-launch: 256 * tiles block histograms)
constexpr size_t NOALLOC_HIST_WORDS = NOALLOC_OS_HEAD + (size_t)OS_MAX_PASSES * NOALLOC_TILES * RADIX;
constexpr size_t NOALLOC_SCAN_WORDS = 256;      // scan of 256 * tiles <= 65536 words: 16 block sums + one level above, padded
constexpr size_t NOALLOC_OFF_KEYS = 0, NOALLOC_OFF_VALS = (size_t)NOALLOC_MAX_PAIRS,
                 NOALLOC_OFF_HIST = 2 * (size_t)NOALLOC_MAX_PAIRS, NOALLOC_OFF_SCAN = NOALLOC_OFF_HIST + NOALLOC_HIST_WORDS,
                 NOALLOC_WORDS = NOALLOC_OFF_SCAN + NOALLOC_SCAN_WORDS;
static_assert((int64_t)RADIX * NOALLOC_TILES <= (int64_t)SCAN_TILE * 64, "scan temporaries sized for two levels");

static __device__ __align__(256) uint32_t g_noalloc_scratch[NOALLOC_WORDS];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
uint32_t *g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};      // device address of g_noalloc_scratch on device i, nullptr = not resolved

// once per device, with that device current
int noalloc_scratch_ptr(b2rt_ctx *ctx, uint32_t **out) {
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES) { ctx->err = "sort_pairs_noalloc: device index out of range"; return B2RT_E_ARG; }
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[ctx->device]) {
        B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            ctx->err = std::string("sort_pairs_noalloc: cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e);
            return B2RT_E_CUDA;
        }
        g_noalloc_ptr[ctx->device] = static_cast<uint32_t *>(p);
    }
    *out = g_noalloc_ptr[ctx->device];
    return B2RT_OK;
}

}  // namespace

int b2rt_sort_pairs_noalloc(b2rt_ctx *ctx, int64_t n, uint32_t *keys, uint32_t *vals, int key_bits) {
    if (!ctx) return B2RT_E_ARG;
    if (n < 0 || key_bits < 1 || key_bits > 32 || (n && (!keys || !vals))) { ctx->err = "sort_pairs_noalloc: bad arguments"; return B2RT_E_ARG; }
    if (n > NOALLOC_MAX_PAIRS) {
        ctx->err = "sort_pairs_noalloc: n = " + std::to_string((long long)n) + " pairs exceeds the static scratch (B2RT_SORT_NOALLOC_MAX = " +
                   std::to_string((long long)NOALLOC_MAX_PAIRS) + " pairs); use b2rt_sort_pairs";
        return B2RT_E_ARG;
    }
    uint32_t *scratch = nullptr;
    int rc = noalloc_scratch_ptr(ctx, &scratch);       // n = 0: the per-device set-up only
    if (rc) return rc;
    if (n <= 1) return B2RT_OK;
    B2RT_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *keys_alt = scratch + NOALLOC_OFF_KEYS, *vals_alt = scratch + NOALLOC_OFF_VALS, *hist = scratch + NOALLOC_OFF_HIST,
             *scan_tmp = scratch + NOALLOC_OFF_SCAN;
    const int passes = (key_bits + 7) / 8;
    const int nblocks = (int)((n + SORT_TILE - 1) / SORT_TILE);
    uint32_t *kin = keys, *vin = vals, *kout = keys_alt, *vout = vals_alt;
    const bool onesweep = ctx->sort_variant == 2 || (ctx->sort_variant == 1 && n < (1ll << 22));     // as b2rt_sort_pairs_impl
    if (onesweep) {
        const size_t words = NOALLOC_OS_HEAD + (size_t)passes * nblocks * RADIX;
        uint32_t *ghist = hist, *tickets = hist + OS_MAX_PASSES * RADIX, *states = hist + NOALLOC_OS_HEAD;
        B2RT_CUDA(ctx, cudaMemsetAsync(ghist, 0, words * sizeof(uint32_t), ctx->stream));      // the last call's counts and flags
        const int hblocks = (int)std::min<int64_t>((n + 4 * SORT_THREADS - 1) / (4 * SORT_THREADS), (int64_t)ctx->sm_count * 8);
        onesweep_hist_kernel<<<hblocks, SORT_THREADS, 0, ctx->stream>>>(kin, n, passes, ghist);
        B2RT_LAUNCH_CHECK(ctx);
        for (int p = 0; p < passes; ++p) {
            onesweep_scatter_kernel<<<nblocks, SORT_THREADS, 0, ctx->stream>>>(kin, vin, kout, vout, n, 8 * p, ghist + p * RADIX,
                                                                              states + (size_t)p * nblocks * RADIX, tickets + p);
            B2RT_LAUNCH_CHECK(ctx);
            std::swap(kin, kout);
            std::swap(vin, vout);
        }
    } else {
        for (int p = 0; p < passes; ++p) {
            radix_hist_kernel<<<nblocks, SORT_THREADS, 0, ctx->stream>>>(kin, n, 8 * p, hist, nblocks);      // writes all 256 * nblocks words
            B2RT_LAUNCH_CHECK(ctx);
            rc = scan_rec(ctx, hist, (int64_t)RADIX * nblocks, scan_tmp, nullptr);
            if (rc) return rc;
            rc = launch_scatter(ctx, kin, vin, kout, vout, n, 8 * p, hist, nblocks);
            if (rc) return rc;
            std::swap(kin, kout);
            std::swap(vin, vout);
        }
    }
    if (passes & 1) {       // an odd number of passes leaves the result in the scratch: copy it home (stream-ordered, no allocation)
        B2RT_CUDA(ctx, cudaMemcpyAsync(keys, keys_alt, (size_t)n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        B2RT_CUDA(ctx, cudaMemcpyAsync(vals, vals_alt, (size_t)n * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    }
    return B2RT_OK;
}

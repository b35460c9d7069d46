// sm_100a kernels for the uPow nonce search (replaces the loop at miner.py:83-98).
//
// Work decomposition (contiguous search):  nonce = (H << 16) | L.
//   * a WARP owns an item = 32 consecutive L values x `hc` consecutive H values;
//     lane <-> L, the inner loop walks H.  Items are numbered H-chunk-major, so a higher
//     item index never holds a lower nonce than every nonce of a lower H-chunk: that is
//     what makes the grid-wide early exit exact.
//   * v1 (138-byte message): L is the low half of tail word W1 and H the high half of W2
//     (SURVEY Appendix B.1), so everything that depends on W0/W1 only -- round 1, sigma0(W1),
//     W16, sigma1(W16), the W17/W31 partial sums -- is loop invariant for a lane and is
//     hoisted out of the H loop; rounds 2..63 remain per nonce.
//   * v2 (108-byte message): the nonce is tail word W10, rounds 0..9 are done on the host
//     (Tail::pre), rounds 10..63 remain per nonce; schedule terms that depend only on the
//     template are uniform values the compiler keeps on the uniform datapath.
//   * only digest word 0 (= mid[0] + a64) is produced in the hot loop; the difficulty rule
//     is tested as one unsigned range compare on it, a warp ballot decides whether anybody
//     leaves the loop, and pre-filter hits are re-checked on a rolled slow path that
//     computes the full digest and applies the general nibble rule.
// HBM traffic: none in the loop (all operands are registers, immediates or the constant
// bank holding the kernel parameters); a hit costs one 8-byte atomic.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "sha256_core.h"
#include "template_prep.h"
#include "hot_loop.h"

namespace upow {

enum : int { MODE_MIN = 0, MODE_ENUM = 1 };

// Device-side scratch of one search (allocated by the library, one per stream slot).
struct DevResult {
    unsigned long long best;        // MODE_MIN: lowest valid nonce so far, ~0ull = none
    unsigned long long items_done;  // completed warp items (hashes_done = items_done * 32 * hc)
    unsigned int count;             // MODE_ENUM: valid nonces seen (may exceed the capacity)
    unsigned int abort_flag;        // written by the host to stop the kernel early
    unsigned long long next_item;   // dynamic work queue: next unclaimed item index
    unsigned long long clk_cycles;  // clock64() span of warp 0 (SM clock = cycles / ns)
    unsigned long long clk_ns;      // %globaltimer span of the same region
};

struct SearchArgs {
    Tail tail;
    Rule rule;
    unsigned long long nonce_lo, nonce_hi;  // exact bounds; hits outside are dropped on the slow path
    uint32_t h_lo;        // first H of this launch
    uint32_t h_count;     // number of H values
    uint32_t hc;          // H values per item (multiple of the kernel's NPT)
    uint32_t n_items;     // ceil(h_count / hc) * 2048
    DevResult* res;
    uint32_t* out;        // MODE_ENUM: valid nonces
    uint32_t cap;
    uint32_t pad;
    Opaque q;             // opaque 1 / 2^k operands of the pipe-controlled primitives (hot_loop.h)
    V2Consts v2;          // host-summed schedule constants of the 108-byte layout
};

struct StridedArgs {
    Tail tail;
    Rule rule;
    unsigned long long start, stride, n_steps;  // nonce_i = start + i*stride, i < n_steps, all < 2^32
    uint32_t ic;        // steps per lane per item
    uint32_t n_items;   // ceil(n_steps / (32*ic))
    DevResult* res;
    uint32_t* out;
    uint32_t cap;
    uint32_t pad;
    Opaque q;           // used by the TUNED build only
    V2Consts v2;
};

__constant__ uint32_t K_DEV[64];

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Claims the next item of the launch-wide queue for the calling warp (lane 0 does the atomic).
__device__ __forceinline__ unsigned long long claim_item(unsigned long long* counter, uint32_t lane) {
    unsigned long long item = 0;
    if (lane == 0) item = atomicAdd(counter, 1ull);
    return __shfl_sync(0xffffffffu, item, 0);
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    return *reinterpret_cast<const volatile unsigned long long*>(p);
}
__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p) {
    return *reinterpret_cast<const volatile unsigned int*>(p);
}

// Rounds FIRST..63 of the tail compression, fully unrolled, schedule expanded in place.
// `s` is the state BEFORE round FIRST; returns a after round 63 (digest word 0 minus mid[0]).
// Words of `w` that are literal zeros at the call site fold away at compile time.
template <int FIRST>
__device__ __forceinline__ uint32_t tail_rounds_a64(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e,
                                                    uint32_t f, uint32_t g, uint32_t h, uint32_t (&w)[16]) {
#pragma unroll
    for (int t = FIRST; t < 64; t++) {
        if (t >= 16) {
            w[t & 15] = w[t & 15] + small_sigma0(w[(t + 1) & 15]) + w[(t + 9) & 15] + small_sigma1(w[(t + 14) & 15]);
        }
        uint32_t t1 = h + big_sigma1(e) + ch(e, f, g) + (Kc(t) + w[t & 15]);
        uint32_t t2 = big_sigma0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    return a;
}

// Slow path: full digest of one nonce from the midstate (rolled code, rarely executed).
__device__ __noinline__ void full_digest(const Tail& tail, uint32_t nonce, uint32_t out[8]) {
    uint32_t w[16];
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = tail.w[i];
    place_nonce(w, tail.layout, nonce);
#pragma unroll
    for (int i = 0; i < 8; i++) out[i] = tail.mid[i];
    compress_rolled(out, w, K_DEV);
}

template <int MODE>
__device__ __noinline__ void report_hit(const Tail& tail, const Rule& rule, unsigned long long lo,
                                        unsigned long long hi, uint32_t nonce, DevResult* res, uint32_t* out,
                                        uint32_t cap) {
    if (nonce < lo || nonce >= hi) return;
    uint32_t d[8];
    full_digest(tail, nonce, d);
    if (!rule_accepts(rule, d)) return;
    if (MODE == MODE_MIN) {
        atomicMin(&res->best, (unsigned long long)nonce);
    } else {
        unsigned int idx = atomicAdd(&res->count, 1u);
        if (idx < cap) out[idx] = nonce;
    }
}

// Contiguous search.  LAYOUT: LAYOUT_V1 / LAYOUT_V2.  NPT: nonces per lane per loop trip.
template <int LAYOUT, int MODE, int NPT>
__global__ void __launch_bounds__(256) search_kernel(const __grid_constant__ SearchArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t done = 0;
    unsigned long long clk0 = 0, ns0 = 0;
    if (warp == 0) { clk0 = clock64(); ns0 = global_timer_ns(); }

    for (;;) {
        const unsigned long long claimed = claim_item(&A.res->next_item, lane);  // increasing order
        if (claimed >= A.n_items) break;
        const uint32_t item = (uint32_t)claimed;
        const uint32_t hchunk = item >> 11, lgroup = item & 2047;
        const uint32_t h0 = A.h_lo + hchunk * A.hc;
        uint32_t h_end = h0 + A.hc;
        if (h_end > A.h_lo + A.h_count) h_end = A.h_lo + A.h_count;
        // grid-wide early exit: nothing at or above this item can beat the recorded minimum
        bool stop = ld_volatile_u32(&A.res->abort_flag) != 0;
        if (MODE == MODE_MIN) stop = stop || ld_volatile_u64(&A.res->best) < ((unsigned long long)h0 << 16);
        if (__any_sync(0xffffffffu, stop)) break;

        const uint32_t L = (lgroup << 5) | lane;

        if (LAYOUT == LAYOUT_V1) {
            // ---- per-lane invariants: W1 and round 1 ----
            const uint32_t w0 = A.tail.w[0];
            const uint32_t w1 = A.tail.w[1] | v1_w1_bits(L);
            uint32_t a = A.tail.pre[0], b = A.tail.pre[1], c = A.tail.pre[2], d = A.tail.pre[3];
            uint32_t e = A.tail.pre[4], f = A.tail.pre[5], g = A.tail.pre[6], h = A.tail.pre[7];
            {
                uint32_t t1 = h + big_sigma1(e) + ch(e, f, g) + (Kc(1) + w1);
                uint32_t t2 = big_sigma0(a) + maj(a, b, c);
                h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
            }
            for (uint32_t H = h0; H < h_end; H += NPT) {
                uint32_t a64[NPT];
#pragma unroll
                for (int j = 0; j < NPT; j++) {
                    uint32_t w[16];
                    w[0] = w0; w[1] = w1;
                    w[2] = v1_w2(H + j);
#pragma unroll
                    for (int i = 3; i < 15; i++) w[i] = 0u;
                    w[15] = 138u * 8u;
                    a64[j] = tail_rounds_a64<2>(a, b, c, d, e, f, g, h, w);
                }
                bool hit = false;
#pragma unroll
                for (int j = 0; j < NPT; j++) hit = hit || (a64[j] + A.rule.sub0 < A.rule.range0);
                if (__ballot_sync(0xffffffffu, hit)) {
#pragma unroll
                    for (int j = 0; j < NPT; j++)
                        if (a64[j] + A.rule.sub0 < A.rule.range0 && H + j < h_end)
                            report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.res,
                                             A.out, A.cap);
                }
            }
        } else {
            const uint32_t a = A.tail.pre[0], b = A.tail.pre[1], c = A.tail.pre[2], d = A.tail.pre[3];
            const uint32_t e = A.tail.pre[4], f = A.tail.pre[5], g = A.tail.pre[6], h = A.tail.pre[7];
            const uint32_t lsw = v2_w10_from_L(L);
            for (uint32_t H = h0; H < h_end; H += NPT) {
                uint32_t a64[NPT];
#pragma unroll
                for (int j = 0; j < NPT; j++) {
                    uint32_t w[16];
#pragma unroll
                    for (int i = 0; i < 10; i++) w[i] = A.tail.w[i];
                    w[10] = lsw | v2_w10_from_H(H + j);
                    w[11] = 0x80000000u; w[12] = 0u; w[13] = 0u; w[14] = 0u;
                    w[15] = 108u * 8u;
                    a64[j] = tail_rounds_a64<10>(a, b, c, d, e, f, g, h, w);
                }
                bool hit = false;
#pragma unroll
                for (int j = 0; j < NPT; j++) hit = hit || (a64[j] + A.rule.sub0 < A.rule.range0);
                if (__ballot_sync(0xffffffffu, hit)) {
#pragma unroll
                    for (int j = 0; j < NPT; j++)
                        if (a64[j] + A.rule.sub0 < A.rule.range0 && H + j < h_end)
                            report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.res,
                                             A.out, A.cap);
                }
            }
        }
        done++;
    }
    if (lane == 0 && done) atomicAdd(&A.res->items_done, (unsigned long long)done);
    if (warp == 0 && lane == 0) {
        A.res->clk_cycles = clock64() - clk0;
        A.res->clk_ns = global_timer_ns() - ns0;
    }
}

// Contiguous search, hand-specialised per layout with explicit pipe assignment (hot_loop.h).
// Same work decomposition and hit protocol as search_kernel above; Tn selects how much of the
// integer work is issued on the fma pipe.
template <int LAYOUT, int MODE, class Tn, int MINB, int BLOCK = 256, int NPT = 1>
__global__ void __launch_bounds__(BLOCK, MINB) search_kernel_tuned(const __grid_constant__ SearchArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint32_t done = 0;
    unsigned long long clk0 = 0, ns0 = 0;
    if (warp == 0) { clk0 = clock64(); ns0 = global_timer_ns(); }

    // Items are claimed from a device-wide counter, in increasing order.  (A static split leaves
    // the SM half empty at the end: the warp scheduler favours some warps, those finish their
    // share early, and the stragglers then run without enough parallelism -- measured with ncu:
    // 17 of 24 warps active on average.)
    for (;;) {
        const unsigned long long claimed = claim_item(&A.res->next_item, lane);
        if (claimed >= A.n_items) break;
        const uint32_t item = (uint32_t)claimed;
        const uint32_t hchunk = item >> 11, lgroup = item & 2047;
        const uint32_t h0 = A.h_lo + hchunk * A.hc;
        uint32_t h_end = h0 + A.hc;
        if (h_end > A.h_lo + A.h_count) h_end = A.h_lo + A.h_count;
        bool stop = ld_volatile_u32(&A.res->abort_flag) != 0;
        if (MODE == MODE_MIN) stop = stop || ld_volatile_u64(&A.res->best) < ((unsigned long long)h0 << 16);
        if (__any_sync(0xffffffffu, stop)) break;
        const uint32_t L = (lgroup << 5) | lane;

        if (LAYOUT == LAYOUT_V1) {
            const V1Lane s = v1_lane_setup(A.tail, L);
            for (uint32_t H = h0; H < h_end; H += NPT) {
                uint32_t a64[NPT];
#pragma unroll
                for (int j = 0; j < NPT; j++) a64[j] = v1_nonce_a64<Tn>(s, v1_w2(H + j), A.q);   // independent chains
#pragma unroll
                for (int j = 0; j < NPT; j++) {
                    const bool hit = addu<Tn::ADD != 0>(a64[j], A.rule.sub0, A.q) < A.rule.range0 && (NPT == 1 || H + j < h_end);
                    if (__ballot_sync(0xffffffffu, hit)) {
                        if (hit) report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.res, A.out, A.cap);
                    }
                }
            }
        } else {
            const uint32_t lsw = v2_w10_from_L(L);
            for (uint32_t H = h0; H < h_end; H += NPT) {
                uint32_t a64[NPT];
#pragma unroll
                for (int j = 0; j < NPT; j++) a64[j] = v2_nonce_a64<Tn>(A.v2, lsw | v2_w10_from_H(H + j), A.q);
#pragma unroll
                for (int j = 0; j < NPT; j++) {
                    const bool hit = addu<Tn::ADD != 0>(a64[j], A.rule.sub0, A.q) < A.rule.range0 && (NPT == 1 || H + j < h_end);
                    if (__ballot_sync(0xffffffffu, hit)) {
                        if (hit) report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.res, A.out, A.cap);
                    }
                }
            }
        }
        done++;
    }
    if (lane == 0 && done) atomicAdd(&A.res->items_done, (unsigned long long)done);
    if (warp == 0 && lane == 0) {
        A.res->clk_cycles = clock64() - clk0;
        A.res->clk_ns = global_timer_ns() - ns0;
    }
}

// Strided search (the reference's residue-class worker layout, miner.py:139-148):
// nonce_i = start + i*stride.  Nothing is hoisted: all 64 rounds run per nonce from the
// midstate, which also makes this kernel an independent cross-check of search_kernel.
// TUNED = true runs the same pipe-assigned rounds as search_kernel_tuned instead (the per-lane set-up of the
// 138-byte layout is then redone for every nonce, because W1 changes with every step).
using StridedTune = Tune<1, 0, 0, 0, 0, 0, 0>;
template <int LAYOUT, int MODE, bool TUNED = false>
__global__ void __launch_bounds__(256, TUNED ? 3 : 1) search_strided_kernel(const __grid_constant__ StridedArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t done = 0;
    for (;;) {
        const unsigned long long item = claim_item(&A.res->next_item, lane);
        if (item >= A.n_items) break;
        const unsigned long long i0 = item * 32u * A.ic;
        bool stop = ld_volatile_u32(&A.res->abort_flag) != 0;
        if (MODE == MODE_MIN) stop = stop || ld_volatile_u64(&A.res->best) < A.start + i0 * A.stride;
        if (__any_sync(0xffffffffu, stop)) break;
        for (uint32_t k = 0; k < A.ic; k++) {
            const unsigned long long i = i0 + (unsigned long long)k * 32u + lane;
            const bool live = i < A.n_steps;
            const uint32_t nonce = (uint32_t)(A.start + i * A.stride);
            if (TUNED) {
                uint32_t a64t;
                if (LAYOUT == LAYOUT_V1) {
                    const V1Lane s = v1_lane_setup(A.tail, nonce & 0xffffu);
                    a64t = v1_nonce_a64<StridedTune>(s, v1_w2(nonce >> 16), A.q);
                } else {
                    a64t = v2_nonce_a64<StridedTune>(A.v2, bswap32(nonce), A.q);
                }
                const bool hit_t = live && (addu<true>(a64t, A.rule.sub0, A.q) < A.rule.range0);
                if (__ballot_sync(0xffffffffu, hit_t)) {
                    if (hit_t) report_hit<MODE>(A.tail, A.rule, A.start, 1ull << 32, nonce, A.res, A.out, A.cap);
                }
                continue;
            }
            uint32_t w[16];
            if (LAYOUT == LAYOUT_V1) {
                w[0] = A.tail.w[0];
                w[1] = A.tail.w[1] | v1_w1_bits(nonce);
                w[2] = v1_w2_from_nonce(nonce);
#pragma unroll
                for (int j = 3; j < 15; j++) w[j] = 0u;
                w[15] = 138u * 8u;
            } else {
#pragma unroll
                for (int j = 0; j < 10; j++) w[j] = A.tail.w[j];
                w[10] = bswap32(nonce);
                w[11] = 0x80000000u; w[12] = 0u; w[13] = 0u; w[14] = 0u;
                w[15] = 108u * 8u;
            }
            const uint32_t a64 = tail_rounds_a64<0>(A.tail.mid[0], A.tail.mid[1], A.tail.mid[2], A.tail.mid[3],
                                                    A.tail.mid[4], A.tail.mid[5], A.tail.mid[6], A.tail.mid[7], w);
            const bool hit = live && (a64 + A.rule.sub0 < A.rule.range0);
            if (__ballot_sync(0xffffffffu, hit)) {
                if (hit) report_hit<MODE>(A.tail, A.rule, A.start, 1ull << 32, nonce, A.res, A.out, A.cap);
            }
        }
        done++;
    }
    if (lane == 0 && done) atomicAdd(&A.res->items_done, (unsigned long long)done);
}

// ---------------------------------------------------------------------------------------
// Batched search (BASELINE config 5): many templates, each with its own midstate, mined
// concurrently.  The job table lives in global memory (a few hundred bytes per template, so
// thousands of templates stay L2 resident); a warp loads its template's words once per item
// and keeps them in registers across the H loop.
struct BatchJob {
    Tail tail;
    Rule rule;
};

struct BatchArgs {
    const BatchJob* jobs;         // [n_jobs]
    const uint32_t* active;       // [n_active] indices into jobs / best
    unsigned long long* best;     // [n_jobs] lowest valid nonce per template, ~0ull = none
    DevResult* res;               // items_done / abort_flag shared by the launch
    unsigned long long nonce_hi;  // exclusive bound (max_nonce)
    uint32_t n_active;
    uint32_t h_lo, h_count, hc;   // this round covers H in [h_lo, h_lo + h_count)
    uint32_t items_per_job;       // ceil(h_count / hc) * 2048
    uint32_t pad;
    Opaque q;                     // see hot_loop.h
};

using BatchTune = Tune<1, 0, 0, 0, 0, 0, 0>;   // same pipe assignment as the default search variant

template <int LAYOUT>
__global__ void __launch_bounds__(256, 3) batch_kernel(const __grid_constant__ BatchArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    const unsigned long long n_items = (unsigned long long)A.n_active * A.items_per_job;
    uint32_t done = 0;
    for (;;) {
        const unsigned long long item = claim_item(&A.res->next_item, lane);
        if (item >= n_items) break;
        const uint32_t slot = (uint32_t)(item / A.items_per_job);
        const uint32_t rem = (uint32_t)(item - (unsigned long long)slot * A.items_per_job);
        const uint32_t hchunk = rem >> 11, lgroup = rem & 2047;
        const uint32_t h0 = A.h_lo + hchunk * A.hc;
        uint32_t h_end = h0 + A.hc;
        if (h_end > A.h_lo + A.h_count) h_end = A.h_lo + A.h_count;
        if (__any_sync(0xffffffffu, ld_volatile_u32(&A.res->abort_flag) != 0)) break;
        const uint32_t j = A.active[slot];
        if (__any_sync(0xffffffffu, ld_volatile_u64(&A.best[j]) < ((unsigned long long)h0 << 16))) continue;
        const BatchJob& job = A.jobs[j];
        if (job.tail.layout != LAYOUT) continue;  // the other layout's launch handles it
        const uint32_t L = (lgroup << 5) | lane;
        const uint32_t sub0 = job.rule.sub0, range0 = job.rule.range0;
        uint32_t a = job.tail.pre[0], b = job.tail.pre[1], c = job.tail.pre[2], d = job.tail.pre[3];
        uint32_t e = job.tail.pre[4], f = job.tail.pre[5], g = job.tail.pre[6], h = job.tail.pre[7];
        if (LAYOUT == LAYOUT_V1) {
            // the 138-byte layout runs the hand-specialised loop of hot_loop.h; the per-template words
            // come from the job table instead of the constant bank, otherwise identical to search_kernel_tuned
            const V1Lane s = v1_lane_setup(job.tail, L);
            for (uint32_t H = h0; H < h_end; H++) {
                const uint32_t a64 = v1_nonce_a64<BatchTune>(s, v1_w2(H), A.q);
                const bool hit = add2<true>(a64, sub0, A.q) < range0;
                if (__ballot_sync(0xffffffffu, hit)) {
                    if (hit) {
                        const uint32_t nonce = (H << 16) | L;
                        uint32_t dg[8];
                        full_digest(job.tail, nonce, dg);
                        if (nonce < A.nonce_hi && rule_accepts(job.rule, dg))
                            atomicMin(&A.best[j], (unsigned long long)nonce);
                    }
                }
            }
        } else {
            uint32_t wt[10];
#pragma unroll
            for (int i = 0; i < 10; i++) wt[i] = job.tail.w[i];
            const uint32_t lsw = v2_w10_from_L(L);
            for (uint32_t H = h0; H < h_end; H++) {
                uint32_t w[16];
#pragma unroll
                for (int i = 0; i < 10; i++) w[i] = wt[i];
                w[10] = lsw | v2_w10_from_H(H);
                w[11] = 0x80000000u; w[12] = 0u; w[13] = 0u; w[14] = 0u;
                w[15] = 108u * 8u;
                const uint32_t a64 = tail_rounds_a64<10>(a, b, c, d, e, f, g, h, w);
                const bool hit = a64 + sub0 < range0;
                if (__ballot_sync(0xffffffffu, hit)) {
                    if (hit) {
                        const uint32_t nonce = (H << 16) | L;
                        uint32_t dg[8];
                        full_digest(job.tail, nonce, dg);
                        if (nonce < A.nonce_hi && rule_accepts(job.rule, dg))
                            atomicMin(&A.best[j], (unsigned long long)nonce);
                    }
                }
            }
        }
        done++;
    }
    if (lane == 0 && done) atomicAdd(&A.res->items_done, (unsigned long long)done);
}

// Fully unrolled generic compression (no folded words): verifier and batched hashing.
__device__ __forceinline__ void compress_unrolled(uint32_t st[8], uint32_t (&w)[16]) {
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t++) {
        if (t >= 16)
            w[t & 15] = w[t & 15] + small_sigma0(w[(t + 1) & 15]) + w[(t + 9) & 15] + small_sigma1(w[(t + 14) & 15]);
        uint32_t t1 = h + big_sigma1(e) + ch(e, f, g) + (Kc(t) + w[t & 15]);
        uint32_t t2 = big_sigma0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// ---------------------------------------------------------------------------------------
// Batch block verifier: check_block_is_valid (upow/manager.py:130-151) for many
// finished blocks at once -- what chain sync does one block at a time (upow/node/main.py:97-150).
// One thread hashes one block (2 or 3 compressions).  The records (192 B each, = upow_block) are
// staged through shared memory with coalesced 16-byte loads; this kernel is HBM-bound:
// 192 B in + 33 B out per block.
struct VerifyRecord {          // must match upow_block in include/upow_b200.h
    uint8_t content[140];
    uint32_t content_len;
    uint8_t prev_hash[32];
    int32_t n_int;
    int32_t charset_count;
    uint32_t pad[2];
};
static_assert(sizeof(VerifyRecord) == 192, "VerifyRecord must be 192 bytes");

constexpr int VERIFY_THREADS = 128;

__global__ void __launch_bounds__(VERIFY_THREADS) verify_kernel(const VerifyRecord* recs, uint32_t n, uint8_t* out_valid,
                                                               uint32_t* out_digests) {
    __shared__ uint4 stage[VERIFY_THREADS * 12];   // 128 records x 192 B
    const uint32_t base = blockIdx.x * VERIFY_THREADS;
    const uint32_t n_here = min((uint32_t)VERIFY_THREADS, n - base);
    const uint4* src = reinterpret_cast<const uint4*>(recs + base);
    for (uint32_t i = threadIdx.x; i < n_here * 12; i += VERIFY_THREADS) stage[i] = src[i];
    __syncthreads();
    if (threadIdx.x >= n_here) return;
    const uint32_t* rec = reinterpret_cast<const uint32_t*>(stage) + threadIdx.x * 48;
    const uint32_t len = rec[35];
    uint8_t verdict = 0;
    uint32_t st[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    if (len == 138 || len == 108) {   // the only two lengths the node accepts (upow/manager.py:401-419)
        const uint32_t n_blocks = (len + 9 + 63) / 64;
        for (uint32_t b = 0; b < n_blocks; b++) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; i++) {
                const uint32_t off = 64 * b + 4 * i;           // byte offset of this word in the padded message
                uint32_t v = 0;
                if (off + 4 <= len) v = bswap32(rec[off >> 2]);
                else if (off <= len) {                          // word holding the tail of the content and the 0x80
                    const uint32_t keep = len - off;            // 0..3 content bytes
                    const uint32_t raw = off < 140 ? bswap32(rec[off >> 2]) : 0u;
                    v = keep ? (raw & (0xffffffffu << (32 - 8 * keep))) : 0u;
                    v |= 0x80u << (24 - 8 * keep);
                }
                if (b == n_blocks - 1 && i == 15) v = len * 8;
                w[i] = v;
            }
            compress_unrolled(st, w);
        }
        uint32_t prev[8];
#pragma unroll
        for (int i = 0; i < 8; i++) prev[i] = bswap32(rec[36 + i]);
        verdict = digest_accepts_prev(prev, (int)rec[44], (int)rec[45], st) ? 1 : 0;
    } else {
        verdict = 0xff;   // malformed length: the reference's split_block_content asserts
#pragma unroll
        for (int i = 0; i < 8; i++) st[i] = 0;
    }
    out_valid[base + threadIdx.x] = verdict;
    if (out_digests) {
#pragma unroll
        for (int i = 0; i < 8; i++) out_digests[(size_t)(base + threadIdx.x) * 8 + i] = st[i];
    }
}

// ---------------------------------------------------------------------------------------
// Generic batched SHA-256 of variable-length messages: the per-transaction hashes
// of upow/node/main.py:692 / upow/manager.py:365-378 / Transaction.hash.  One thread per message,
// fully unrolled compression; messages sit back to back in one byte buffer (any alignment) and are read
// with aligned 32-bit loads re-aligned and byte-swapped by one PRMT per word.  The buffer must be
// readable 8 bytes past its end (the library pads its device copy).  Bound by the alu pipe like the
// search kernels (about 1300 instructions per 64 bytes), not by HBM.
__global__ void __launch_bounds__(128) sha256_batch_kernel(const uint8_t* data, const unsigned long long* offsets,
                                                          uint32_t n, uint32_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned long long beg = offsets[i];
    const unsigned long long len = offsets[i + 1] - beg;
    const uint32_t* words = reinterpret_cast<const uint32_t*>(data) + (beg >> 2);   // aligned word holding byte `beg`
    const uint32_t mis = (uint32_t)(beg & 3);
    // big-endian word from two consecutive little-endian aligned words lo, hi at byte misalignment `mis`
    const uint32_t sel = mis == 0 ? 0x0123u : mis == 1 ? 0x1234u : mis == 2 ? 0x2345u : 0x3456u;
    uint32_t st[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    const unsigned long long n_blocks = (len + 9 + 63) / 64;
    for (unsigned long long b = 0; b < n_blocks; b++) {
        uint32_t w[16];
        const unsigned long long base = 64 * b;
        uint32_t lo = (base < len) ? __ldg(words + base / 4) : 0u;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const unsigned long long pos = base + 4 * j;
            uint32_t v = 0;
            if (pos < len) {
                const uint32_t hi = __ldg(words + pos / 4 + 1);     // may lie past the message: masked below / padded buffer
                v = __byte_perm(lo, hi, sel);
                lo = hi;
                if (pos + 4 > len) {                                 // last, partial word: keep len-pos bytes, then 0x80
                    const uint32_t keep = (uint32_t)(len - pos);    // 1..3
                    v = (v & (0xffffffffu << (32 - 8 * keep))) | (0x80u << (24 - 8 * keep));
                }
            } else if (pos == len) {
                v = 0x80000000u;
            }
            w[j] = v;
        }
        if (b == n_blocks - 1) {
            w[14] = (uint32_t)((len * 8) >> 32);
            w[15] = (uint32_t)(len * 8);
        }
        compress_unrolled(st, w);
    }
#pragma unroll
    for (int k = 0; k < 8; k++) out[(size_t)i * 8 + k] = st[k];
}

// Digests of the batch winners: thread i handles job i (skipped when nothing was found).
__global__ void batch_digest_kernel(const BatchJob* jobs, const unsigned long long* best, uint32_t n_jobs,
                                    uint32_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_jobs || best[i] == ~0ull) return;
    uint32_t d[8];
    full_digest(jobs[i].tail, (uint32_t)best[i], d);
    for (int k = 0; k < 8; k++) out[i * 8 + k] = d[k];
}

// Full digest of one nonce (the winner), written to out[0..8).
__global__ void digest_kernel(const __grid_constant__ Tail tail, uint32_t nonce, uint32_t* out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        uint32_t d[8];
        full_digest(tail, nonce, d);
        for (int i = 0; i < 8; i++) out[i] = d[i];
    }
}

// Selftest: thread i hashes message block set i (already padded, n_blocks[i] blocks of 16
// words) with the rolled compression, and thread 0 additionally runs the unrolled hot-path
// rounds on block 0 of message 0; results go to out[i*8..] and unrolled_a64.
__global__ void selftest_kernel(const uint32_t* blocks, const uint32_t* n_blocks, const uint32_t* offsets,
                                uint32_t n_msgs, uint32_t* out, uint32_t* unrolled_a64) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_msgs) return;
    uint32_t st[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    for (uint32_t b = 0; b < n_blocks[i]; b++) compress_rolled(st, blocks + (offsets[i] + b) * 16, K_DEV);
    for (int j = 0; j < 8; j++) out[i * 8 + j] = st[j];
    if (i == 0) {
        uint32_t w[16];
        for (int j = 0; j < 16; j++) w[j] = blocks[offsets[0] * 16 + j];
        *unrolled_a64 = tail_rounds_a64<0>(0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c,
                                           0x1f83d9ab, 0x5be0cd19, w) + 0x6a09e667u;
    }
}

}  // namespace upow

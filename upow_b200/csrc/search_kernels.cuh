// sm_100a search kernels for the uPow nonce search (replaces the loop at miner.py:83-98).
// Everything in this header is a template, a `static` device function or a `static` table, so
// several translation units (upow_lib.cu, noalloc.cu) can include it side by side.
//
// Work decomposition (contiguous search):  nonce = (H << 16) | L.
//   * a WARP owns an item = 32 consecutive L values (one "lgroup", 2048 per H) x `hc` consecutive
//     H values; lane <-> L, the inner loop walks H.  Items are numbered H-chunk-major, so a
//     higher item index never holds a lower nonce than every nonce of a lower H-chunk: that is
//     what makes the grid-wide early exit exact.
//   * several GPUs share one nonce space by LGROUP INTERLEAVE: a launch handles the lgroups
//     lg_first, lg_first + lg_step, ... of every H-chunk.  All GPUs then walk H together, every
//     GPU does the same amount of work per H-chunk, and the lowest valid nonce is the minimum of
//     the per-GPU minima.  A find is published to the other GPUs' `limit` words (peer memory)
//     so that they stop at the same H.
//   * v1 (138-byte message): L is the low half of tail word W1 and H the high half of W2
//     (SURVEY Appendix B.1), so everything that depends on W0/W1 only -- round 1, sigma0(W1),
//     W16, sigma1(W16), the W17/W31 partial sums -- is loop invariant for a lane and is
//     hoisted out of the H loop; rounds 2..63 remain per nonce.
//   * v2 (108-byte message): the nonce is tail word W10, rounds 0..9 are done on the host
//     (Tail::pre), rounds 10..63 remain per nonce; schedule terms that depend only on the
//     template are summed on the host (V2Consts).
//   * only digest word 0 (= mid[0] + a64) is produced in the hot loop; the difficulty rule
//     is tested as one unsigned range compare on it, a warp ballot decides whether anybody
//     leaves the loop, and pre-filter hits are re-checked on a rolled slow path that
//     computes the full digest and applies the general nibble rule.
// HBM traffic: none in the loop (all operands are registers, immediates or the constant
// bank holding the kernel parameters); a hit costs a few 8-byte atomics.
#pragma once
#include <assert.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "sha256_core.h"
#include "template_prep.h"
#include "hot_loop.h"

// -DUPOW_CHECKED (python -m upow_b200.build --checked -> libupow_b200_checked.so): every computed index of the
// kernels is asserted on the device.  compute-sanitizer is not available on the GPU pool this was developed on;
// tools/sanitizer_target.py runs a small pass through every kernel against this build instead.
#ifdef UPOW_CHECKED
#define UPOW_ASSERT(cond) assert(cond)
#else
#define UPOW_ASSERT(cond) ((void)0)
#endif

namespace upow {

enum : int { MODE_MIN = 0, MODE_ENUM = 1 };
constexpr int MAX_PEERS = 8;
constexpr unsigned long long NO_NONCE = ~0ull;

// Device-side scratch of one launch (allocated by the library; a ring of these per GPU).
struct DevResult {
    unsigned long long best;        // MODE_MIN: lowest valid nonce found by THIS launch, NO_NONCE = none
    unsigned long long limit;       // items whose first nonce is >= limit are not searched: min over every known
                                    // find (own and the other GPUs') of nonce + 1; NO_NONCE while nothing is known
    unsigned long long next_item;   // dynamic work queue: next unclaimed item index
    unsigned long long items_done;  // completed warp items
    unsigned long long h_steps;     // sum over warps of H values actually hashed (hashes = h_steps * 32)
    unsigned int count;             // MODE_ENUM: valid nonces seen (may exceed the capacity)
    unsigned int warps_exited;      // the last warp to leave writes `digest`
    unsigned long long clk_cycles;  // clock64() span of warp 0 (SM clock = cycles / ns)
    unsigned long long clk_ns;      // %globaltimer span of the same region
    uint32_t digest[8];             // MODE_MIN: SHA-256 state words of `best`'s candidate
};

// Where a valid nonce is recorded.  peer_limit[i] points at the `limit` word of the launch that searches the
// same nonce space on another GPU (device memory reached over NVLink peer access; 0 entries on one GPU).
struct HitSink {
    DevResult* res;
    uint32_t* out;        // MODE_ENUM: valid nonces
    uint32_t cap;
    uint32_t n_peers;
    unsigned long long* peer_limit[MAX_PEERS];
};

// Host-to-kernel stop signal that cannot be lost or hit the wrong launch: every launch on a device carries
// a serial number, and the device-wide kill word holds a serial range [lo, hi] (lo in the high half).  A
// launch whose serial lies in the range leaves at its next item.  Serial numbers only grow, so a word
// written for an earlier call never matches a later launch.
struct KillSwitch {
    const unsigned long long* word;
    uint32_t serial;
    uint32_t pad;
};
constexpr unsigned long long KILL_NONE = 0xffffffff00000000ull;
__host__ __device__ inline unsigned long long kill_range(uint32_t lo, uint32_t hi) {
    return ((unsigned long long)lo << 32) | hi;
}

struct SearchArgs {
    Tail tail;
    Rule rule;
    unsigned long long nonce_lo, nonce_hi;  // exact bounds; hits outside are dropped on the slow path
    uint32_t h_lo;        // first H of this launch
    uint32_t h_count;     // number of H values
    uint32_t hc;          // H values per item
    uint32_t lg_first, lg_step, lg_count;   // this launch's lgroups: lg_first + i * lg_step, i < lg_count
    uint32_t n_items;     // ceil(h_count / hc) * lg_count
    uint32_t pad;
    HitSink sink;
    KillSwitch kill;
    Opaque q;             // opaque 1 / 2^k operands of the pipe-controlled primitives (hot_loop.h)
    V2Consts v2;          // host-summed schedule constants of the 108-byte layout
};

struct StridedArgs {
    Tail tail;
    Rule rule;
    unsigned long long start, stride, n_steps;  // nonce_i = start + i*stride, i < n_steps, all < 2^32
    uint32_t ic;        // steps per lane per item
    uint32_t n_items;   // ceil(n_steps / (32*ic))
    HitSink sink;
    KillSwitch kill;
    Opaque q;           // used by the TUNED build only
    V2Consts v2;
};

// Round constants for the rolled slow path (the unrolled loops take them as immediates from Kc()).
static __constant__ uint32_t K_DEV[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
    0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
    0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
    0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
    0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
    0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
    0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
    0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};

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
__device__ __forceinline__ bool killed(const KillSwitch& k) {
    const unsigned long long w = ld_volatile_u64(k.word);
    return k.serial >= (uint32_t)(w >> 32) && k.serial <= (uint32_t)w;
}
// How often the H loop of a MODE_MIN search re-reads `limit` (H values, power of two).  The read is software
// pipelined (issued at one poll, tested at the next), so a find anywhere stops every warp within twice this
// many hashes per lane instead of at its next item, and no warp ever waits for the load.
#ifndef UPOW_LIMIT_POLL
#define UPOW_LIMIT_POLL 8
#endif
constexpr uint32_t LIMIT_POLL = UPOW_LIMIT_POLL;

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
static __device__ __noinline__ void full_digest(const Tail& tail, uint32_t nonce, uint32_t out[8]) {
    uint32_t w[16];
#pragma unroll
    for (int i = 0; i < 16; i++) w[i] = tail.w[i];
    place_nonce(w, tail.layout, nonce);
#pragma unroll
    for (int i = 0; i < 8; i++) out[i] = tail.mid[i];
    compress_rolled(out, w, K_DEV);
}

// A lane's pre-filter hit: full digest, general rule, then the result slot.  MODE_MIN also lowers `limit`
// here and on the other GPUs that share the nonce space.
template <int MODE>
static __device__ __noinline__ void report_hit(const Tail& tail, const Rule& rule, unsigned long long lo,
                                               unsigned long long hi, uint32_t nonce, const HitSink& sink) {
    if (nonce < lo || nonce >= hi) return;
    uint32_t d[8];
    full_digest(tail, nonce, d);
    if (!rule_accepts(rule, d)) return;
    if (MODE == MODE_MIN) {
        atomicMin(&sink.res->best, (unsigned long long)nonce);
        atomicMin(&sink.res->limit, (unsigned long long)nonce + 1);
        for (uint32_t p = 0; p < sink.n_peers; p++) atomicMin_system(sink.peer_limit[p], (unsigned long long)nonce + 1);
    } else {
        unsigned int idx = atomicAdd(&sink.res->count, 1u);
        UPOW_ASSERT(sink.out != nullptr || sink.cap == 0);
        if (idx < sink.cap) sink.out[idx] = nonce;
    }
}

// Common kernel epilogue: work counters, and (MODE_MIN) the digest of the launch's best nonce by whichever
// warp leaves last.
template <int MODE>
__device__ __forceinline__ void leave_kernel(const Tail& tail, DevResult* res, uint32_t lane, uint32_t done,
                                             unsigned long long h_steps) {
    bool last = false;
    if (lane == 0) {
        if (done) {
            atomicAdd(&res->items_done, (unsigned long long)done);
            atomicAdd(&res->h_steps, h_steps);
        }
        if (MODE == MODE_MIN) {
            __threadfence();
            const unsigned int total = gridDim.x * (blockDim.x >> 5);
            last = atomicAdd(&res->warps_exited, 1u) == total - 1;
        }
    }
    if (MODE == MODE_MIN) {
        last = __shfl_sync(0xffffffffu, last, 0);
        if (last && lane == 0) {
            const unsigned long long best = ld_volatile_u64(&res->best);
            if (best != NO_NONCE) {
                uint32_t d[8];
                full_digest(tail, (uint32_t)best, d);
#pragma unroll
                for (int i = 0; i < 8; i++) res->digest[i] = d[i];
            }
        }
    }
}

// Item decode shared by the contiguous kernels.
struct ItemRange {
    uint32_t h0, h_end, L;
};
__device__ __forceinline__ ItemRange decode_item(const SearchArgs& A, uint32_t item, uint32_t lane) {
    const uint32_t hchunk = item / A.lg_count, lgi = item - hchunk * A.lg_count;
    UPOW_ASSERT(item < A.n_items && A.lg_first + lgi * A.lg_step < 2048 && A.h_lo + hchunk * A.hc < A.h_lo + A.h_count);
    ItemRange r;
    r.h0 = A.h_lo + hchunk * A.hc;
    r.h_end = r.h0 + A.hc;
    if (r.h_end > A.h_lo + A.h_count) r.h_end = A.h_lo + A.h_count;
    r.L = ((A.lg_first + lgi * A.lg_step) << 5) | lane;
    return r;
}

// Contiguous search, plain CUDA (compiler-chosen pipes).  LAYOUT: LAYOUT_V1 / LAYOUT_V2.  Kept as an
// independent implementation of the same search (variant 0).
template <int LAYOUT, int MODE, int NPT>
__global__ void __launch_bounds__(256) search_kernel(const __grid_constant__ SearchArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    DevResult* const res = A.sink.res;
    uint32_t done = 0;
    unsigned long long h_steps = 0;
    unsigned long long clk0 = 0, ns0 = 0;
    if (warp == 0) { clk0 = clock64(); ns0 = global_timer_ns(); }

    for (;;) {
        const unsigned long long claimed = claim_item(&res->next_item, lane);  // increasing order
        if (claimed >= A.n_items) break;
        const ItemRange it = decode_item(A, (uint32_t)claimed, lane);
        // grid-wide early exit: nothing at or above this item can beat the recorded minimum
        bool stop = killed(A.kill);
        unsigned long long lim = NO_NONCE;   // MODE_MIN: the limit as last read; refreshed every LIMIT_POLL H values
        if (MODE == MODE_MIN) {
            lim = ld_volatile_u64(&res->limit);
            stop = stop || ((unsigned long long)it.h0 << 16) >= lim;
        }
        if (__any_sync(0xffffffffu, stop)) break;
        const uint32_t L = it.L;
        uint32_t H = it.h0;

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
            for (; H < it.h_end; H += NPT) {
                if (MODE == MODE_MIN && (H & (LIMIT_POLL - 1)) == 0) {
                    // test the value read LIMIT_POLL iterations ago, then start the next read: its latency hides
                    // behind the hashes in between instead of stalling the warp
                    if (((unsigned long long)H << 16) >= lim) break;
                    lim = ld_volatile_u64(&res->limit);
                }
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
                        if (a64[j] + A.rule.sub0 < A.rule.range0 && H + j < it.h_end)
                            report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.sink);
                }
            }
        } else {
            const uint32_t a = A.tail.pre[0], b = A.tail.pre[1], c = A.tail.pre[2], d = A.tail.pre[3];
            const uint32_t e = A.tail.pre[4], f = A.tail.pre[5], g = A.tail.pre[6], h = A.tail.pre[7];
            const uint32_t lsw = v2_w10_from_L(L);
            for (; H < it.h_end; H += NPT) {
                if (MODE == MODE_MIN && (H & (LIMIT_POLL - 1)) == 0) {
                    // test the value read LIMIT_POLL iterations ago, then start the next read: its latency hides
                    // behind the hashes in between instead of stalling the warp
                    if (((unsigned long long)H << 16) >= lim) break;
                    lim = ld_volatile_u64(&res->limit);
                }
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
                        if (a64[j] + A.rule.sub0 < A.rule.range0 && H + j < it.h_end)
                            report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.sink);
                }
            }
        }
        h_steps += H - it.h0;
        done++;
    }
    leave_kernel<MODE>(A.tail, res, lane, done, h_steps);
    if (warp == 0 && lane == 0) {
        res->clk_cycles = clock64() - clk0;
        res->clk_ns = global_timer_ns() - ns0;
    }
}

// Contiguous search, hand-specialised per layout with explicit pipe assignment (hot_loop.h).
// Same work decomposition and hit protocol as search_kernel above; Tn selects how much of the
// integer work is issued on the fma pipe.
template <int LAYOUT, int MODE, class Tn, int MINB, int BLOCK = 256, int NPT = 1>
__global__ void __launch_bounds__(BLOCK, MINB) search_kernel_tuned(const __grid_constant__ SearchArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    DevResult* const res = A.sink.res;
    uint32_t done = 0;
    unsigned long long h_steps = 0;
    unsigned long long clk0 = 0, ns0 = 0;
    if (warp == 0) { clk0 = clock64(); ns0 = global_timer_ns(); }

    // Items are claimed from a device-wide counter, in increasing order.  (A static split leaves
    // the SM half empty at the end: the warp scheduler favours some warps, those finish their
    // share early, and the stragglers then run without enough parallelism -- measured with ncu:
    // 17 of 24 warps active on average.)
    for (;;) {
        const unsigned long long claimed = claim_item(&res->next_item, lane);
        if (claimed >= A.n_items) break;
        const ItemRange it = decode_item(A, (uint32_t)claimed, lane);
        bool stop = killed(A.kill);
        unsigned long long lim = NO_NONCE;   // MODE_MIN: the limit as last read; refreshed every LIMIT_POLL H values
        if (MODE == MODE_MIN) {
            lim = ld_volatile_u64(&res->limit);
            stop = stop || ((unsigned long long)it.h0 << 16) >= lim;
        }
        if (__any_sync(0xffffffffu, stop)) break;
        const uint32_t L = it.L;
        uint32_t H = it.h0;

        if (LAYOUT == LAYOUT_V1) {
            const V1Lane s = v1_lane_setup(A.tail, L);
            // MODE_MIN walks H in blocks of LIMIT_POLL: at every block boundary the limit read one block ago is
            // tested and the next read is started (its latency hides behind the block's hashes); the inner loop is
            // the same instruction stream as the enumerating kernel's.
            while (H < it.h_end) {
                uint32_t h_stop = it.h_end;
                if (MODE == MODE_MIN) {
                    if (((unsigned long long)H << 16) >= lim) break;
                    lim = ld_volatile_u64(&res->limit);
                    h_stop = min(h_stop, (H | (LIMIT_POLL - 1)) + 1);
                }
                for (; H < h_stop; H += NPT) {
                    uint32_t a64[NPT];
#pragma unroll
                    for (int j = 0; j < NPT; j++) a64[j] = v1_nonce_a64<Tn>(s, v1_w2(H + j), A.q);   // independent chains
#pragma unroll
                    for (int j = 0; j < NPT; j++) {
                        const bool hit = addu<Tn::ADD != 0>(a64[j], A.rule.sub0, A.q) < A.rule.range0 && (NPT == 1 || H + j < it.h_end);
                        if (__ballot_sync(0xffffffffu, hit)) {
                            if (hit) report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.sink);
                        }
                    }
                }
            }
        } else {
            const uint32_t lsw = v2_w10_from_L(L);
            while (H < it.h_end) {
                uint32_t h_stop = it.h_end;
                if (MODE == MODE_MIN) {
                    if (((unsigned long long)H << 16) >= lim) break;
                    lim = ld_volatile_u64(&res->limit);
                    h_stop = min(h_stop, (H | (LIMIT_POLL - 1)) + 1);
                }
                for (; H < h_stop; H += NPT) {
                    uint32_t a64[NPT];
#pragma unroll
                    for (int j = 0; j < NPT; j++) a64[j] = v2_nonce_a64<Tn>(A.v2, lsw | v2_w10_from_H(H + j), A.q);
#pragma unroll
                    for (int j = 0; j < NPT; j++) {
                        const bool hit = addu<Tn::ADD != 0>(a64[j], A.rule.sub0, A.q) < A.rule.range0 && (NPT == 1 || H + j < it.h_end);
                        if (__ballot_sync(0xffffffffu, hit)) {
                            if (hit) report_hit<MODE>(A.tail, A.rule, A.nonce_lo, A.nonce_hi, ((H + j) << 16) | L, A.sink);
                        }
                    }
                }
            }
        }
        h_steps += H - it.h0;
        done++;
    }
    leave_kernel<MODE>(A.tail, res, lane, done, h_steps);
    if (warp == 0 && lane == 0) {
        res->clk_cycles = clock64() - clk0;
        res->clk_ns = global_timer_ns() - ns0;
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
    DevResult* const res = A.sink.res;
    uint32_t done = 0;
    unsigned long long steps = 0;
    for (;;) {
        const unsigned long long item = claim_item(&res->next_item, lane);
        if (item >= A.n_items) break;
        const unsigned long long i0 = item * 32u * A.ic;
        bool stop = killed(A.kill);
        if (MODE == MODE_MIN) stop = stop || A.start + i0 * A.stride >= ld_volatile_u64(&res->limit);
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
                    if (hit_t) report_hit<MODE>(A.tail, A.rule, A.start, 1ull << 32, nonce, A.sink);
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
                if (hit) report_hit<MODE>(A.tail, A.rule, A.start, 1ull << 32, nonce, A.sink);
            }
        }
        steps += A.ic;
        done++;
    }
    leave_kernel<MODE>(A.tail, res, lane, done, steps);
}

// ---------------------------------------------------------------------------------------
// Batched search (BASELINE config 5): many templates, each with its own midstate, mined
// concurrently.  The job table lives in global memory (a few hundred bytes per template, so
// thousands of templates stay L2 resident); a warp loads its template's words once per item
// and keeps them in registers across the H loop.
//
// The search advances in ROUNDS of h_count H values, one launch per round, queued back to back on the
// stream with no host round trip in between.  A round's items are ordered H-chunk-major over the templates
// that were still unsolved when the round began (`active`, rebuilt on the device by batch_compact_kernel
// after every round).  Several GPUs split every template's lgroups (same interleave as the single-template
// search) and publish each find to the others' limit[] arrays, so no template is ever re-dealt and every
// GPU does the same work per round.
struct BatchJob {
    Tail tail;
    Rule rule;
    V2Consts v2;
};

struct BatchCtl {
    unsigned long long next_item;    // item queue of the current round (reset by batch_compact_kernel)
    unsigned long long items_done;   // totals over all rounds
    unsigned long long h_steps;
    unsigned int n_active[2];        // entries of active[parity]
    unsigned int first_find_round;   // lowest round in which anything was found (own or peer finds), ~0u = none
    unsigned int rounds_run;         // rounds that had at least one item
};

struct BatchArgs {
    const BatchJob* jobs;            // [n_jobs]
    const uint32_t* active;          // this round's list: indices into jobs / limit
    unsigned long long* limit;       // [n_jobs] min over known finds of nonce + 1, NO_NONCE = unsolved
    BatchCtl* ctl;
    unsigned long long* peer_limit[MAX_PEERS];        // the other GPUs' limit[] arrays
    unsigned int* peer_first_find[MAX_PEERS];         // the other GPUs' ctl->first_find_round
    unsigned long long nonce_hi;     // exclusive bound (max_nonce)
    uint32_t n_peers;
    uint32_t n_jobs;
    uint32_t parity;                 // which n_active entry describes `active`
    uint32_t round;
    uint32_t stop_at_first;          // 1: rounds after the first one with a find do nothing (upow_search_any)
    uint32_t h_lo, h_count, hc;      // this round covers H in [h_lo, h_lo + h_count)
    uint32_t lg_first, lg_step, lg_count;
    uint32_t poll_mask;              // limit[] is re-read every poll_mask + 1 H values (a power of two minus one)
    KillSwitch kill;
    Opaque q;                        // see hot_loop.h
};

using BatchTune = Tune<1, 0, 0, 0, 0, 0, 0>;   // same pipe assignment as the default search variant

static __device__ __noinline__ void batch_report_hit(const BatchJob& job, uint32_t j, uint32_t nonce, const BatchArgs& A) {
    if (nonce >= A.nonce_hi) return;
    uint32_t dg[8];
    full_digest(job.tail, nonce, dg);
    if (!rule_accepts(job.rule, dg)) return;
    atomicMin(&A.limit[j], (unsigned long long)nonce + 1);
    atomicMin(&A.ctl->first_find_round, A.round);
    for (uint32_t p = 0; p < A.n_peers; p++) {
        atomicMin_system(A.peer_limit[p] + j, (unsigned long long)nonce + 1);
        atomicMin_system(A.peer_first_find[p], A.round);
    }
}

template <int LAYOUT>
__global__ void __launch_bounds__(256, 3) batch_kernel(const __grid_constant__ BatchArgs A) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t n_active = A.ctl->n_active[A.parity];
    if (A.stop_at_first && ld_volatile_u32(&A.ctl->first_find_round) < A.round) return;
    const uint32_t n_hchunks = (A.h_count + A.hc - 1) / A.hc;
    const uint32_t per_chunk = n_active * A.lg_count;
    const unsigned long long n_items = (unsigned long long)n_hchunks * per_chunk;
    uint32_t done = 0;
    unsigned long long h_steps = 0;
    for (;;) {
        const unsigned long long item = claim_item(&A.ctl->next_item, lane);
        if (item >= n_items) break;
        if (__any_sync(0xffffffffu, killed(A.kill))) break;
        // (a 64-bit division is ~100 instructions; rounds almost always have fewer than 2^32 items)
        const uint32_t hchunk = n_items <= 0xffffffffull ? (uint32_t)item / per_chunk : (uint32_t)(item / per_chunk);
        const uint32_t rem = (uint32_t)(item - (unsigned long long)hchunk * per_chunk);
        const uint32_t slot = rem / A.lg_count, lgi = rem - slot * A.lg_count;
        const uint32_t h0 = A.h_lo + hchunk * A.hc;
        uint32_t h_end = h0 + A.hc;
        if (h_end > A.h_lo + A.h_count) h_end = A.h_lo + A.h_count;
        UPOW_ASSERT(slot < n_active && n_active <= A.n_jobs && h0 < A.h_lo + A.h_count);
        const uint32_t j = A.active[slot];
        UPOW_ASSERT(j < A.n_jobs);
        const unsigned long long* lim_p = &A.limit[j];
        unsigned long long lim = ld_volatile_u64(lim_p);
        if (__any_sync(0xffffffffu, ((unsigned long long)h0 << 16) >= lim)) continue;
        const BatchJob& job = A.jobs[j];
        if (job.tail.layout != LAYOUT) continue;  // the other layout's launch handles it
        const uint32_t L = ((A.lg_first + lgi * A.lg_step) << 5) | lane;
        const uint32_t sub0 = job.rule.sub0, range0 = job.rule.range0;
        uint32_t H = h0;
        if (LAYOUT == LAYOUT_V1) {
            // the hand-specialised loop of hot_loop.h; the per-template words come from the job table instead of
            // the constant bank, otherwise identical to search_kernel_tuned
            const V1Lane s = v1_lane_setup(job.tail, L);
            while (H < h_end) {
                if (((unsigned long long)H << 16) >= lim) break;
                lim = ld_volatile_u64(lim_p);
                const uint32_t h_stop = min(h_end, (H | A.poll_mask) + 1);
                for (; H < h_stop; H++) {
                    const uint32_t a64 = v1_nonce_a64<BatchTune>(s, v1_w2(H), A.q);
                    const bool hit = add2<true>(a64, sub0, A.q) < range0;
                    if (__ballot_sync(0xffffffffu, hit)) {
                        if (hit) batch_report_hit(job, j, (H << 16) | L, A);
                    }
                }
            }
        } else {
            const V2Consts u = job.v2;   // per-template schedule constants, in registers for the H loop
            const uint32_t lsw = v2_w10_from_L(L);
            while (H < h_end) {
                if (((unsigned long long)H << 16) >= lim) break;
                lim = ld_volatile_u64(lim_p);
                const uint32_t h_stop = min(h_end, (H | A.poll_mask) + 1);
                for (; H < h_stop; H++) {
                    const uint32_t a64 = v2_nonce_a64<BatchTune>(u, lsw | v2_w10_from_H(H), A.q);
                    const bool hit = add2<true>(a64, sub0, A.q) < range0;
                    if (__ballot_sync(0xffffffffu, hit)) {
                        if (hit) batch_report_hit(job, j, (H << 16) | L, A);
                    }
                }
            }
        }
        h_steps += H - h0;
        done++;
    }
    if (lane == 0 && done) {
        atomicAdd(&A.ctl->items_done, (unsigned long long)done);
        atomicAdd(&A.ctl->h_steps, h_steps);
    }
}

// After a round: the templates that still have nonces worth searching above h_next form the next round's
// list (order is irrelevant), and the item queue is rewound.  One thread per entry of the old list; the host
// zeroes n_active[parity_in ^ 1] on the stream beforehand.
static __global__ void batch_compact_kernel(const uint32_t* active_in, uint32_t* active_out, const unsigned long long* limit,
                                            BatchCtl* ctl, uint32_t parity_in, uint32_t h_next, uint32_t n_jobs) {
    const uint32_t n_in = ctl->n_active[parity_in];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        ctl->next_item = 0;
        if (n_in) ctl->rounds_run++;
    }
    bool keep = false;
    uint32_t j = 0;
    if (i < n_in) {
        j = active_in[i];
        UPOW_ASSERT(j < n_jobs && n_in <= n_jobs);
        keep = ((unsigned long long)h_next << 16) < limit[j];
    }
    const uint32_t mask = __ballot_sync(0xffffffffu, keep);
    if (!mask) return;
    uint32_t base = 0;
    if ((threadIdx.x & 31) == (uint32_t)(__ffs(mask) - 1)) base = atomicAdd(&ctl->n_active[parity_in ^ 1], (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
    UPOW_ASSERT(!keep || base + __popc(mask & ((1u << (threadIdx.x & 31)) - 1)) < n_jobs);
    if (keep) active_out[base + __popc(mask & ((1u << (threadIdx.x & 31)) - 1))] = j;
}
// Digests of the batch winners: thread i handles job i (skipped when nothing was found).
static __global__ void batch_digest_kernel(const BatchJob* jobs, const unsigned long long* limit, uint32_t n_jobs,
                                           uint32_t* out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_jobs || limit[i] == NO_NONCE) return;
    uint32_t d[8];
    full_digest(jobs[i].tail, (uint32_t)(limit[i] - 1), d);
    for (int k = 0; k < 8; k++) out[i * 8 + k] = d[k];
}

}  // namespace upow

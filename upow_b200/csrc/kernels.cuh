// Non-template kernels of libupow_b200.so that only upow_lib.cu needs: the batch block verifier, the batched
// variable-length SHA-256, the single-digest and selftest kernels.  The search kernels proper live in
// search_kernels.cuh (shared with noalloc.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "search_kernels.cuh"

namespace upow {

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

using BulkTune = Tune<1, 0, 0, 0, 0, 0, 0>;   // additions on the fma pipe, as in the search kernels
// TUNED: the compression with explicit pipe assignment (hot_loop.h compress_tuned); otherwise plain CUDA.
template <bool TUNED>
__global__ void __launch_bounds__(VERIFY_THREADS) verify_kernel(const VerifyRecord* recs, uint32_t n, uint8_t* out_valid,
                                                               uint32_t* out_digests, const __grid_constant__ Opaque q) {
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
                UPOW_ASSERT(off + 4 > len || (off >> 2) < 35);   // content words only
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
            if (TUNED) compress_tuned<BulkTune>(st, w, q);
            else compress_unrolled(st, w);
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
        for (int i = 0; i < 8; i++) out_digests[(size_t)(base + threadIdx.x) * 8 + i] = bswap32(st[i]);   // digest bytes in order
    }
}

// ---------------------------------------------------------------------------------------
// Generic batched SHA-256 of variable-length messages: the per-transaction hashes
// of upow/node/main.py:692 / upow/manager.py:365-378 / Transaction.hash.  One thread per message,
// fully unrolled compression; messages sit back to back in one byte buffer (any alignment) and are read
// with aligned 32-bit loads re-aligned and byte-swapped by one PRMT per word.  The buffer must be
// readable 8 bytes past its end (the library pads its device copy).  Bound by the alu pipe like the
// search kernels (about 1300 instructions per 64 bytes), not by HBM.
template <bool TUNED>
__global__ void __launch_bounds__(128) sha256_batch_kernel(const uint8_t* data, const unsigned long long* offsets,
                                                          uint32_t n, uint32_t* out, unsigned long long readable_bytes,
                                                          const __grid_constant__ Opaque q) {
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
        UPOW_ASSERT(base >= len || (beg >> 2) * 4 + base + 4 <= readable_bytes);
        uint32_t lo = (base < len) ? __ldg(words + base / 4) : 0u;
#pragma unroll
        for (int j = 0; j < 16; j++) {
            const unsigned long long pos = base + 4 * j;
            uint32_t v = 0;
            if (pos < len) {
                UPOW_ASSERT((beg >> 2) * 4 + pos + 8 <= readable_bytes);   // the one-word overread stays inside the padding
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
        if (TUNED) compress_tuned<BulkTune>(st, w, q);
        else compress_unrolled(st, w);
    }
#pragma unroll
    for (int k = 0; k < 8; k++) out[(size_t)i * 8 + k] = bswap32(st[k]);   // digest bytes in order
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

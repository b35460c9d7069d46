// SHA-256 building blocks shared by the CUDA kernels and the host-side template
// preparation.  Everything here is 32-bit unsigned integer arithmetic (FIPS 180-4);
// the reference reaches the same function through hashlib (miner.py:52,61).
//
// The header compiles both under nvcc (device + host) and under plain g++ (the host
// logic tests build a harness from it), hence the UPOW_HD / UPOW_DEVICE_CODE macros.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define UPOW_HD __host__ __device__ __forceinline__
#else
#define UPOW_HD inline
#endif
#if defined(__CUDA_ARCH__)
#define UPOW_DEVICE_CODE 1
#else
#define UPOW_DEVICE_CODE 0
#endif

namespace upow {

// Round constants.  Indexed with compile-time constants in the unrolled kernels, so
// they become immediates in SASS.
UPOW_HD constexpr uint32_t Kc(int i) {
    constexpr uint32_t k[64] = {
        0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5,
        0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174,
        0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
        0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967,
        0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85,
        0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
        0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3,
        0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
    return k[i];
}

UPOW_HD uint32_t rotr(uint32_t x, int n) {
#if UPOW_DEVICE_CODE
    return __funnelshift_r(x, x, n);  // SHF.R.W
#else
    return (x >> n) | (x << (32 - n));
#endif
}

// Three-input logic ops.  Written as plain expressions (not inline PTX) so that literal-zero
// schedule words still constant-fold; nvcc turns each into a single LOP3.LUT.
UPOW_HD uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) { return a ^ b ^ c; }
UPOW_HD uint32_t ch(uint32_t e, uint32_t f, uint32_t g) { return g ^ (e & (f ^ g)); }
UPOW_HD uint32_t maj(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (c & (a | b)); }

UPOW_HD uint32_t big_sigma0(uint32_t x) { return xor3(rotr(x, 2), rotr(x, 13), rotr(x, 22)); }
UPOW_HD uint32_t big_sigma1(uint32_t x) { return xor3(rotr(x, 6), rotr(x, 11), rotr(x, 25)); }
UPOW_HD uint32_t small_sigma0(uint32_t x) { return xor3(rotr(x, 7), rotr(x, 18), x >> 3); }
UPOW_HD uint32_t small_sigma1(uint32_t x) { return xor3(rotr(x, 17), rotr(x, 19), x >> 10); }

// PRMT: result byte i = byte number (nibble i of sel) of the 8-byte pool {x bytes 0-3, y bytes 4-7}.
UPOW_HD uint32_t byte_perm(uint32_t x, uint32_t y, uint32_t sel) {
#if UPOW_DEVICE_CODE
    return __byte_perm(x, y, sel);
#else
    const uint64_t pool = ((uint64_t)y << 32) | x;
    uint32_t r = 0;
    for (int i = 0; i < 4; i++) r |= (uint32_t)((pool >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
#endif
}

UPOW_HD uint32_t bswap32(uint32_t x) {
#if UPOW_DEVICE_CODE
    return __byte_perm(x, 0, 0x0123);  // PRMT
#else
    return (x >> 24) | ((x >> 8) & 0xff00) | ((x << 8) & 0xff0000) | (x << 24);
#endif
}

// Rolled, fully general compression.  Used by the host-side midstate computation, by the
// device slow path (re-check of a pre-filter hit, full digest of the winner) and by the
// selftest kernel.  `k` is a pointer so device code can hand in its __constant__ copy.
UPOW_HD void compress_rolled(uint32_t st[8], const uint32_t w_in[16], const uint32_t* k) {
    uint32_t w[16];
    for (int i = 0; i < 16; i++) w[i] = w_in[i];
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#if UPOW_DEVICE_CODE
#pragma unroll 1
#endif
    for (int t = 0; t < 64; t++) {
        if (t >= 16) {
            w[t & 15] += small_sigma0(w[(t + 1) & 15]) + w[(t + 9) & 15] + small_sigma1(w[(t + 14) & 15]);
        }
        uint32_t t1 = h + big_sigma1(e) + ch(e, f, g) + k[t] + w[t & 15];
        uint32_t t2 = big_sigma0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

}  // namespace upow

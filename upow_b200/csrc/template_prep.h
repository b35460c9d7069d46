// Host-side preparation of one block template for the device: SHA-256 midstate of the
// constant leading blocks, the tail-block words, the nonce-independent leading rounds and
// the nibble form of the difficulty rule.  Plain C++ (no CUDA) so the host-logic tests can
// compile it with g++.
//
// Reference behaviour restated here:
//   message layout            miner.py:74-82, 87  (prefix || nonce u32-LE; 138 or 108 bytes,
//                             upow/manager.py:401-419 asserts exactly these two lengths)
//   target_prefix             miner.py:43         prev_hash_hex[-int(difficulty):]  (N==0 -> whole string)
//   charset limit             miner.py:45-49,51-56 digest_hex[N] in "0123456789abcdef"[:count]
#pragma once
#include <stdint.h>
#include <string.h>

#include "sha256_core.h"

namespace upow {

enum : uint32_t { LAYOUT_V1 = 1 /* 138-byte message, 64-byte address */, LAYOUT_V2 = 2 /* 108-byte, 33-byte address */ };

// Difficulty rule on raw digest words (word 0 = first 8 hex chars of the hex digest).
struct Rule {
    uint32_t want[8];        // target nibbles, packed exactly like the digest words; unused nibbles 0
    uint32_t n_target;       // leading nibbles that must equal `want` (0..64)
    uint32_t frac_index;     // index of the charset-limited nibble (only read when charset_count < 16)
    uint32_t charset_count;  // nibble at frac_index must be < charset_count; 16 = unrestricted
    uint32_t sub0, range0;   // word-0 pre-filter: pass  <=>  (a64 + sub0) < range0   (unsigned)
};

// One (template, timestamp): everything the kernels need, all host-computed.
struct Tail {
    uint32_t mid[8];   // SHA-256 state after the constant leading blocks (the midstate)
    uint32_t w[16];    // tail block as big-endian words, nonce bytes zero
    uint32_t pre[8];   // state after the leading tail rounds that do not depend on the nonce
                       // (v1: round 0; v2: rounds 0..9)
    uint32_t layout;   // LAYOUT_V1 / LAYOUT_V2
    uint32_t pad;
};

static const uint32_t H_INIT[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a,
                                   0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};

struct KTable {
    uint32_t k[64];
    KTable() {
        for (int i = 0; i < 64; i++) k[i] = Kc(i);
    }
};
inline const uint32_t* k_table() {
    static const KTable t;
    return t.k;
}

inline uint32_t load_be32(const uint8_t* p) {
    return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3];
}
inline void store_be32(uint8_t* p, uint32_t v) {
    p[0] = v >> 24; p[1] = v >> 16; p[2] = v >> 8; p[3] = v;
}

// Rounds [first, last) of one compression on (st, w) without the final feed-forward.
inline void host_rounds(uint32_t st[8], const uint32_t w_in[16], int first, int last) {
    uint32_t w[64];
    for (int i = 0; i < 16; i++) w[i] = w_in[i];
    for (int t = 16; t < 64; t++) w[t] = w[t - 16] + small_sigma0(w[t - 15]) + w[t - 7] + small_sigma1(w[t - 2]);
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    for (int t = first; t < last; t++) {
        uint32_t t1 = h + big_sigma1(e) + ch(e, f, g) + Kc(t) + w[t];
        uint32_t t2 = big_sigma0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] = a; st[1] = b; st[2] = c; st[3] = d; st[4] = e; st[5] = f; st[6] = g; st[7] = h;
}

inline void host_sha256(const uint8_t* msg, size_t len, uint8_t out[32]) {
    uint32_t st[8];
    memcpy(st, H_INIT, sizeof st);
    uint32_t w[16];
    size_t off = 0;
    for (; off + 64 <= len; off += 64) {
        for (int i = 0; i < 16; i++) w[i] = load_be32(msg + off + 4 * i);
        compress_rolled(st, w, k_table());
    }
    uint8_t last[128];
    memset(last, 0, sizeof last);
    size_t rem = len - off;
    memcpy(last, msg + off, rem);
    last[rem] = 0x80;
    size_t total = rem + 9 <= 64 ? 64 : 128;
    uint64_t bits = (uint64_t)len * 8;
    for (int i = 0; i < 8; i++) last[total - 1 - i] = (uint8_t)(bits >> (8 * i));
    for (size_t o = 0; o < total; o += 64) {
        for (int i = 0; i < 16; i++) w[i] = load_be32(last + o + 4 * i);
        compress_rolled(st, w, k_table());
    }
    for (int i = 0; i < 8; i++) store_be32(out + 4 * i, st[i]);
}

// Byte offset of the nonce inside the tail block: v1 -> 6 (straddles W1/W2), v2 -> 40 (= W10).
inline int nonce_tail_offset(uint32_t layout) { return layout == LAYOUT_V1 ? 6 : 40; }

// Insert the nonce (u32 little-endian, miner.py:87) into tail words.
UPOW_HD void place_nonce(uint32_t w[16], uint32_t layout, uint32_t nonce) {
    if (layout == LAYOUT_V1) {
        // bytes 6..9 of the tail block: W1 low half = b0 b1, W2 high half = b2 b3
        w[1] = (w[1] & 0xffff0000u) | ((nonce & 0xff) << 8) | ((nonce >> 8) & 0xff);
        w[2] = (w[2] & 0x0000ffffu) | (((nonce >> 16) & 0xff) << 24) | ((nonce >> 24) << 16);
    } else {
        w[10] = bswap32(nonce);
    }
}

// Nonce-carrying tail words for nonce = (H << 16) | L, one PRMT each:
//   v1: W1 = w1_template | b0<<8 | b1          (b0,b1 = bytes of L)
//       W2 = b2<<24 | b3<<16 | 0x8000          (b2,b3 = bytes of H; 0x80 is the SHA pad byte)
//   v2: W10 = bswap32(nonce) = v2_w10_from_L(L) | v2_w10_from_H(H)
UPOW_HD uint32_t v1_w1_bits(uint32_t L) { return byte_perm(L, 0, 0x4401); }
UPOW_HD uint32_t v1_w2(uint32_t H) { return byte_perm(H, 0x00008000u, 0x0154); }
UPOW_HD uint32_t v1_w2_from_nonce(uint32_t nonce) { return byte_perm(nonce, 0x00008000u, 0x2354); }
UPOW_HD uint32_t v2_w10_from_L(uint32_t L) { return byte_perm(L, 0, 0x0144); }
UPOW_HD uint32_t v2_w10_from_H(uint32_t H) { return byte_perm(H, 0, 0x4401); }

// Builds Tail from the prefix bytes.  Returns false when prefix_len is neither 134 nor 104.
inline bool prepare_tail(const uint8_t* prefix, int prefix_len, Tail* t) {
    if (prefix_len != 134 && prefix_len != 104) return false;
    const int msg_len = prefix_len + 4;
    const int n_const = msg_len / 64;  // 2 or 1 constant leading blocks
    memset(t, 0, sizeof *t);
    t->layout = prefix_len == 134 ? LAYOUT_V1 : LAYOUT_V2;
    memcpy(t->mid, H_INIT, sizeof t->mid);
    uint32_t w[16];
    for (int b = 0; b < n_const; b++) {
        for (int i = 0; i < 16; i++) w[i] = load_be32(prefix + 64 * b + 4 * i);
        compress_rolled(t->mid, w, k_table());
    }
    uint8_t tail[64];
    memset(tail, 0, sizeof tail);
    const int rem = prefix_len - 64 * n_const;  // 6 or 40
    memcpy(tail, prefix + 64 * n_const, rem);
    tail[rem + 4] = 0x80;  // pad byte right after the 4 nonce bytes
    tail[62] = (uint8_t)((msg_len * 8) >> 8);
    tail[63] = (uint8_t)((msg_len * 8) & 0xff);
    for (int i = 0; i < 16; i++) t->w[i] = load_be32(tail + 4 * i);
    memcpy(t->pre, t->mid, sizeof t->pre);
    host_rounds(t->pre, t->w, 0, t->layout == LAYOUT_V1 ? 1 : 10);
    return true;
}

// Same template, another timestamp (the "ts <- ts+1, nonce <- 0" rollover): the timestamp
// is the u32-LE at prefix offset len-6 (miner.py:79).
inline void set_prefix_timestamp(uint8_t* prefix, int prefix_len, uint32_t ts) {
    uint8_t* p = prefix + prefix_len - 6;
    p[0] = ts; p[1] = ts >> 8; p[2] = ts >> 16; p[3] = ts >> 24;
}
inline uint32_t get_prefix_timestamp(const uint8_t* prefix, int prefix_len) {
    const uint8_t* p = prefix + prefix_len - 6;
    return (uint32_t)p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24;
}

// Rule from (prev_hash bytes, int(difficulty), charset count).  n_int == 0 reproduces Python's
// s[-0:] == s: the target is the WHOLE previous hash (miner.py:43, upow/manager.py:148).
inline bool prepare_rule(const uint8_t prev_hash[32], int n_int, int charset_count, const uint32_t mid[8], Rule* r) {
    if (n_int < 0 || charset_count < 1 || charset_count > 16) return false;
    if (charset_count < 16 && n_int > 63) return false;  // digest_hex[N] would raise IndexError in the reference
    memset(r, 0, sizeof *r);
    const int n_target = (n_int == 0 || n_int > 64) ? 64 : n_int;
    r->n_target = n_target;
    r->frac_index = n_int;
    r->charset_count = charset_count;
    for (int i = 0; i < n_target; i++) {
        int src = 64 - n_target + i;  // nibble index inside prev_hash hex
        uint32_t nib = (src & 1) ? (prev_hash[src >> 1] & 0xf) : (prev_hash[src >> 1] >> 4);
        r->want[i >> 3] |= nib << (28 - 4 * (i & 7));
    }
    // word-0 pre-filter, exact when every constrained nibble lies in word 0
    r->sub0 = mid[0] - r->want[0];
    if (n_target >= 8) {
        r->range0 = 1;  // word 0 must match exactly; the rest is checked on the slow path
    } else {
        uint32_t free_bits = 32 - 4 * n_target;  // >= 4
        if (charset_count < 16 && (int)r->frac_index == n_target)
            r->range0 = (uint32_t)charset_count << (free_bits - 4);
        else
            r->range0 = 1u << free_bits;  // n_target >= 1 here, so free_bits <= 28
    }
    return true;
}

// The node-side rule, upow/manager.py:130-151, straight from the previous block's hash (as digest-style
// big-endian words) without a prepared Rule: block_hash.startswith(last_hash[-N:]) and, with a fractional
// difficulty, block_hash[N] in charset[:count].  N == 0 and N > 64 select the whole hash (Python slices);
// N >= 64 with a charset limit would raise IndexError in the reference and is reported as not valid.
UPOW_HD uint32_t nibble_of(const uint32_t w[8], int i) { return (w[i >> 3] >> (28 - 4 * (i & 7))) & 0xf; }
UPOW_HD bool digest_accepts_prev(const uint32_t prev[8], int n_int, int charset_count, const uint32_t d[8]) {
    if (n_int < 0 || charset_count < 1 || charset_count > 16) return false;
    const int n_target = (n_int == 0 || n_int > 64) ? 64 : n_int;
    for (int i = 0; i < n_target; i++)
        if (nibble_of(d, i) != nibble_of(prev, 64 - n_target + i)) return false;
    if (charset_count < 16) {
        if (n_int > 63) return false;
        if (nibble_of(d, n_int) >= (uint32_t)charset_count) return false;
    }
    return true;
}

// Generic check of a full digest against the rule (device slow path + host cross-check).
UPOW_HD bool rule_accepts(const Rule& r, const uint32_t d[8]) {
    const uint32_t full = r.n_target >> 3, part = r.n_target & 7;
    for (uint32_t i = 0; i < 8; i++) {
        if (i < full) {
            if (d[i] != r.want[i]) return false;
        } else if (i == full && part) {
            uint32_t mask = 0xffffffffu << (32 - 4 * part);
            if ((d[i] ^ r.want[i]) & mask) return false;
        }
    }
    if (r.charset_count < 16) {
        uint32_t nib = (d[(r.frac_index >> 3) & 7] >> (28 - 4 * (r.frac_index & 7))) & 0xf;
        if (nib >= r.charset_count) return false;
    }
    return true;
}

}  // namespace upow

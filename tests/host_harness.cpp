// Host-only harness around upow_b200/csrc/template_prep.h for the CPU tests (g++, no CUDA).
// It re-plays on the CPU exactly the arithmetic the kernels do per nonce -- tail words built
// with the same PRMT selectors, rounds started from Tail::pre, word-0 pre-filter -- so layout
// and rule bugs are caught before any GPU time is spent.  Test code only; not shipped.
#include <stdint.h>
#include <string.h>

#include "../upow_b200/csrc/hot_loop.h"

using namespace upow;

extern "C" {

int hh_sizes(int* tail, int* rule) { *tail = sizeof(Tail); *rule = sizeof(Rule); return 0; }

int hh_prepare(const uint8_t* prefix, int prefix_len, const uint8_t* prev_hash, int n_int, int cc, Tail* t, Rule* r) {
    if (!prepare_tail(prefix, prefix_len, t)) return -1;
    if (!prepare_rule(prev_hash, n_int, cc, t->mid, r)) return -1;
    return 0;
}

void hh_sha256(const uint8_t* msg, uint64_t len, uint8_t out[32]) { host_sha256(msg, (size_t)len, out); }

// slow path: midstate + rolled compression with place_nonce
void hh_full_digest(const Tail* t, uint32_t nonce, uint32_t out[8]) {
    uint32_t w[16];
    memcpy(w, t->w, sizeof w);
    place_nonce(w, t->layout, nonce);
    memcpy(out, t->mid, 32);
    compress_rolled(out, w, k_table());
}

// hot path as the search kernel does it: nonce = (H<<16)|L, start from Tail::pre
uint32_t hh_hot_a64(const Tail* t, uint32_t nonce) {
    const uint32_t H = nonce >> 16, L = nonce & 0xffff;
    uint32_t w[16], st[8];
    memcpy(w, t->w, sizeof w);
    memcpy(st, t->pre, sizeof st);
    if (t->layout == LAYOUT_V1) {
        w[1] = t->w[1] | v1_w1_bits(L);
        w[2] = v1_w2(H);
        for (int i = 3; i < 15; i++) w[i] = 0;
        w[15] = 138 * 8;
        host_rounds(st, w, 1, 64);
    } else {
        w[10] = v2_w10_from_L(L) | v2_w10_from_H(H);
        w[11] = 0x80000000u; w[12] = w[13] = w[14] = 0; w[15] = 108 * 8;
        host_rounds(st, w, 10, 64);
    }
    return st[0];
}

// hot path of the strided kernel: words from the full nonce, all 64 rounds from the midstate
uint32_t hh_strided_a64(const Tail* t, uint32_t nonce) {
    uint32_t w[16], st[8];
    memcpy(w, t->w, sizeof w);
    memcpy(st, t->mid, sizeof st);
    if (t->layout == LAYOUT_V1) {
        w[1] = t->w[1] | v1_w1_bits(nonce);
        w[2] = v1_w2_from_nonce(nonce);
    } else {
        w[10] = bswap32(nonce);
    }
    host_rounds(st, w, 0, 64);
    return st[0];
}

// the hand-specialised hot loops of hot_loop.h (portable fallbacks of the pipe-controlled primitives)
uint32_t hh_tuned_a64(const Tail* t, uint32_t nonce) {
    using Tn = Tune<1, 1, 0x24, 2, 0xa, 3>;
    Opaque q;
    fill_opaque(&q);
    const uint32_t H = nonce >> 16, L = nonce & 0xffff;
    if (t->layout == LAYOUT_V1) {
        V1Lane lane = v1_lane_setup(*t, L);
        return v1_nonce_a64<Tn>(lane, v1_w2(H), q);
    }
    V2Consts u;
    prepare_v2_consts(*t, &u);
    return v2_nonce_a64<Tn>(u, v2_w10_from_L(L) | v2_w10_from_H(H), q);
}

// whole-message SHA-256 through compress_tuned (the compression of the block verifier / batched hashing kernels)
void hh_sha256_tuned(const uint8_t* msg, uint64_t len, uint8_t out[32]) {
    using Tn = Tune<1, 0, 0, 0, 0, 0, 0>;
    Opaque q;
    fill_opaque(&q);
    uint32_t st[8];
    memcpy(st, H_INIT, sizeof st);
    uint8_t buf[128];
    uint64_t off = 0;
    auto block = [&](const uint8_t* p) {
        uint32_t w[16];
        for (int i = 0; i < 16; i++) w[i] = load_be32(p + 4 * i);
        compress_tuned<Tn>(st, w, q);
    };
    for (; off + 64 <= len; off += 64) block(msg + off);
    memset(buf, 0, sizeof buf);
    const uint64_t rem = len - off;
    memcpy(buf, msg + off, rem);
    buf[rem] = 0x80;
    const uint64_t total = rem + 9 <= 64 ? 64 : 128, bits = len * 8;
    for (int i = 0; i < 8; i++) buf[total - 1 - i] = (uint8_t)(bits >> (8 * i));
    for (uint64_t o = 0; o < total; o += 64) block(buf + o);
    for (int i = 0; i < 8; i++) store_be32(out + 4 * i, st[i]);
}

int hh_prefilter(const Rule* r, uint32_t a64) { return a64 + r->sub0 < r->range0; }
int hh_rule_accepts(const Rule* r, const uint32_t d[8]) { return rule_accepts(*r, d); }

int hh_digest_accepts_prev(const uint32_t prev[8], int n_int, int cc, const uint32_t d[8]) {
    return digest_accepts_prev(prev, n_int, cc, d);
}

void hh_set_timestamp(uint8_t* prefix, int len, uint32_t ts) { set_prefix_timestamp(prefix, len, ts); }
uint32_t hh_get_timestamp(const uint8_t* prefix, int len) { return get_prefix_timestamp(prefix, len); }
}

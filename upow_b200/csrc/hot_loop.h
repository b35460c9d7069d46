// The per-nonce arithmetic of the search kernels, hand-specialised for the two uPow message
// layouts and with explicit control over WHICH ISSUE PIPE every operation uses.
//
// Why: on sm_100 the integer work of SHA-256 splits over two pipes that issue concurrently --
// LOP3 / SHF / PRMT / IADD3 on the "alu" pipe and IMAD* on the "fma" pipe, each 16 lanes per
// SM sub-partition (measured with upow_microbench on B200: 2.0 warp-instructions/clk/SM each,
// 4.0 when interleaved).  nvcc on its own leaves ~1180 of ~1290 instructions per hash on the
// alu pipe, so the kernel is alu-bound at 50 % issue utilisation.  Here every addition can be
// issued as IMAD (x*1+y), every logical shift as the high half of IMAD.WIDE (x * 2^(32-n)),
// and a chosen share of the rotates as IMAD.WIDE + IMAD (lo + hi of x * 2^(32-n)); the "1" and
// the powers of two come from kernel parameters so ptxas cannot fold them back into IADD3/SHF.
//
// Everything is UPOW_HD: tests/host_harness.cpp runs the very same code on the CPU (with the
// portable fallbacks below) against hashlib before any GPU time is spent.
#pragma once
#include <stdint.h>

#include "sha256_core.h"
#include "template_prep.h"

namespace upow {

// Values the device compiler must not see through.  m[n] = 2^(32-n) (n = 1..31).
// `one` feeds the register+register additions: ptxas keeps it in a UNIFORM register, so those
// IMADs read only two vector registers.  `one_k` feeds the additions of a round constant
// (IMAD Rd, Ra, Rone, imm): an immediate and a uniform operand cannot share one instruction, so
// this copy lives in a vector register.  Keeping the two apart is what stops ptxas from moving
// the shared value into a vector register for every IMAD (measured: +25 % register-file reads).
// (The two copies are deliberately NOT adjacent: ptxas otherwise fetches them with one 64-bit
// load into a vector register pair and the uniform form is lost.)
struct Opaque {
    uint32_t one_k;
    uint32_t m[32];
    uint32_t one;
};
inline void fill_opaque(Opaque* q) {
    q->one = 1;
    q->one_k = 1;
    q->m[0] = 0;
    for (int n = 1; n < 32; n++) q->m[n] = 1u << (32 - n);
}

// Tuning knobs (compile time).
//   ADD  1: every addition in the loop is an IMAD on the fma pipe; 0: plain C (compiler's choice)
//   SHR  1: x>>3 / x>>10 of the schedule via IMAD.WIDE
//   RR   bit mask of per-round rotates moved to the fma pipe: bits 0-2 = Sigma1's (6,11,25),
//        bits 3-5 = Sigma0's (2,13,22); applied on rounds with (t & 3) < RQ
//   SR   same for the schedule: bits 0-1 = sigma0's (7,18), bits 2-3 = sigma1's (17,19);
//        applied on words with (t & 3) < SQ
//   CH   1: Sigma1(e) is added separately into the new e and the new a (8 additions per round,
//        one-add dependency on Sigma1); 0: T1 is formed first (7 additions, longer chain)
//   SHR  0: SHF; 1: IMAD.WIDE (high half); 2: IMAD.HI -- on schedule words with (t & 3) < SHRQ
template <int ADD_, int SHR_, int RR_, int RQ_, int SR_, int SQ_, int CH_ = 1, int SHRQ_ = 4>
struct Tune {
    static constexpr int ADD = ADD_, SHR = SHR_, RR = RR_, RQ = RQ_, SR = SR_, SQ = SQ_, CH = CH_, SHRQ = SHRQ_;
};

// ---- pipe-controlled primitives ---------------------------------------------------------
template <bool FMA>
UPOW_HD uint32_t add2(uint32_t a, uint32_t b, const Opaque& q) {
#if UPOW_DEVICE_CODE
    if (FMA) {
        uint32_t r;
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(q.one), "r"(b));
        return r;
    }
#endif
    (void)q;
    return a + b;
}
// a + u where u is a per-launch (warp-uniform) value: the uniform operand occupies the IMAD's
// uniform/immediate slot, so the multiplier "1" has to be the vector-register copy.
template <bool FMA>
UPOW_HD uint32_t addu(uint32_t a, uint32_t u, const Opaque& q) {
#if UPOW_DEVICE_CODE
    if (FMA) {
        uint32_t r;
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(q.one_k), "r"(u));
        return r;
    }
#endif
    (void)q;
    return a + u;
}
template <bool FMA, uint32_t K>
UPOW_HD uint32_t addk(uint32_t a, const Opaque& q) {
#if UPOW_DEVICE_CODE
    if (FMA) {
        uint32_t r;
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(q.one_k), "n"(K));
        return r;
    }
#endif
    (void)q;
    return a + K;
}
template <int FMA, int N>
UPOW_HD uint32_t shr_n(uint32_t x, const Opaque& q) {
#if UPOW_DEVICE_CODE
    if (FMA == 2) {  // IMAD.HI: half issue rate on the fma pipe, but one destination register
        uint32_t hi;
        asm("mul.hi.u32 %0, %1, %2;" : "=r"(hi) : "r"(x), "r"(q.m[N]));
        return hi;
    }
    if (FMA == 1) {
        uint32_t hi;
        asm("{ .reg .b64 t; .reg .b32 lo; mul.wide.u32 t, %1, %2; mov.b64 {lo, %0}, t; }"
            : "=r"(hi) : "r"(x), "r"(q.m[N]));
        return hi;
    }
#endif
    (void)q;
    return x >> N;
}
template <bool FMA, int N>
UPOW_HD uint32_t rotr_n(uint32_t x, const Opaque& q) {
#if UPOW_DEVICE_CODE
    if (FMA) {
        uint32_t lo, hi, r;
        asm("{ .reg .b64 t; mul.wide.u32 t, %2, %3; mov.b64 {%0, %1}, t; }" : "=r"(lo), "=r"(hi) : "r"(x), "r"(q.m[N]));
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(q.one), "r"(hi));
        return r;
    }
#endif
    (void)q;
    return rotr(x, N);
}

template <class Tn, int T>
UPOW_HD uint32_t Sig1(uint32_t e, const Opaque& q) {
    constexpr bool on = (T & 3) < Tn::RQ;
    return xor3(rotr_n<on && (Tn::RR & 1), 6>(e, q), rotr_n<on && (Tn::RR & 2), 11>(e, q),
                rotr_n<on && (Tn::RR & 4), 25>(e, q));
}
template <class Tn, int T>
UPOW_HD uint32_t Sig0(uint32_t a, const Opaque& q) {
    constexpr bool on = (T & 3) < Tn::RQ;
    return xor3(rotr_n<on && (Tn::RR & 8), 2>(a, q), rotr_n<on && (Tn::RR & 16), 13>(a, q),
                rotr_n<on && (Tn::RR & 32), 22>(a, q));
}
template <class Tn, int T>
UPOW_HD uint32_t sig0(uint32_t x, const Opaque& q) {
    constexpr bool on = (T & 3) < Tn::SQ;
    return xor3(rotr_n<on && (Tn::SR & 1), 7>(x, q), rotr_n<on && (Tn::SR & 2), 18>(x, q),
                shr_n<(((T & 3) < Tn::SHRQ) ? Tn::SHR : 0), 3>(x, q));
}
template <class Tn, int T>
UPOW_HD uint32_t sig1(uint32_t x, const Opaque& q) {
    constexpr bool on = (T & 3) < Tn::SQ;
    return xor3(rotr_n<on && (Tn::SR & 4), 17>(x, q), rotr_n<on && (Tn::SR & 8), 19>(x, q),
                shr_n<(((T & 3) < Tn::SHRQ) ? Tn::SHR : 0), 10>(x, q));
}

// One round.  `t_hk` = h + K[t] + W[t] already summed (the caller knows whether that is a
// literal, a per-launch constant, a per-lane invariant or per-nonce).  Both outputs depend on
// Sigma1(e) through a single addition, which keeps the dependent chain per round short.
// UNI_HK / UNI_D: t_hk / d is a warp-uniform value in this round (early rounds only).
template <class Tn, int T, bool NEED_E = true, bool UNI_HK = false, bool UNI_D = false>
UPOW_HD void round_core(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d, uint32_t& e, uint32_t& f, uint32_t& g,
                        uint32_t& h, uint32_t t_hk, const Opaque& q) {
    constexpr bool F = Tn::ADD != 0;
    const uint32_t t = UNI_HK ? addu<F>(ch(e, f, g), t_hk, q) : add2<F>(t_hk, ch(e, f, g), q);
    const uint32_t t2 = add2<F>(Sig0<Tn, T>(a, q), maj(a, b, c), q);
    const uint32_t s1 = Sig1<Tn, T>(e, q);
    uint32_t e_new = 0, a_new;
    if (Tn::CH) {
        const uint32_t tt = add2<F>(t, t2, q);
        if (NEED_E) e_new = add2<F>(UNI_D ? addu<F>(t, d, q) : add2<F>(d, t, q), s1, q);
        a_new = add2<F>(tt, s1, q);
    } else {
        const uint32_t t1 = add2<F>(t, s1, q);
        if (NEED_E) e_new = UNI_D ? addu<F>(t1, d, q) : add2<F>(d, t1, q);
        a_new = add2<F>(t1, t2, q);
    }
    h = g; g = f; f = e; e = e_new; d = c; c = b; b = a; a = a_new;
}

// Generic rounds T..63 on a fully variable schedule held in w[64]; words from SCHED_FROM on are
// produced here with the full four-term recurrence, earlier ones by the layout-specific code.
template <class Tn, int T, int SCHED_FROM>
struct GenericTail {
    static UPOW_HD uint32_t run(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e, uint32_t f, uint32_t g,
                                uint32_t h, uint32_t (&w)[64], const Opaque& q) {
        constexpr bool F = Tn::ADD != 0;
        if (T >= SCHED_FROM) {
            w[T] = add2<F>(add2<F>(add2<F>(w[T - 16], w[T - 7], q), sig0<Tn, T>(w[T - 15], q), q),
                           sig1<Tn, T>(w[T - 2], q), q);
        }
        const uint32_t hk = add2<F>(h, addk<F, Kc(T)>(w[T], q), q);
        round_core<Tn, T, (T < 63)>(a, b, c, d, e, f, g, h, hk, q);
        return GenericTail<Tn, T + 1, SCHED_FROM>::run(a, b, c, d, e, f, g, h, w, q);
    }
};
template <class Tn, int SCHED_FROM>
struct GenericTail<Tn, 64, SCHED_FROM> {
    static UPOW_HD uint32_t run(uint32_t a, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t,
                                uint32_t (&)[64], const Opaque&) {
        return a;
    }
};

// A whole compression on a fully variable block, all eight state words kept (block verifier, batched
// hashing): the same pipe assignment as the search loops, nothing folded.
template <class Tn, int T>
struct FullRounds {
    static UPOW_HD void run(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d, uint32_t& e, uint32_t& f, uint32_t& g,
                            uint32_t& h, uint32_t (&w)[64], const Opaque& q) {
        constexpr bool F = Tn::ADD != 0;
        if (T >= 16) {
            w[T] = add2<F>(add2<F>(add2<F>(w[T - 16], w[T - 7], q), sig0<Tn, T>(w[T - 15], q), q),
                           sig1<Tn, T>(w[T - 2], q), q);
        }
        const uint32_t hk = add2<F>(h, addk<F, Kc(T)>(w[T], q), q);
        round_core<Tn, T>(a, b, c, d, e, f, g, h, hk, q);
        FullRounds<Tn, T + 1>::run(a, b, c, d, e, f, g, h, w, q);
    }
};
template <class Tn>
struct FullRounds<Tn, 64> {
    static UPOW_HD void run(uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&,
                            uint32_t (&)[64], const Opaque&) {}
};
template <class Tn>
UPOW_HD void compress_tuned(uint32_t st[8], const uint32_t (&w16)[16], const Opaque& q) {
    constexpr bool F = Tn::ADD != 0;
    uint32_t w[64];
#if UPOW_DEVICE_CODE
#pragma unroll
#endif
    for (int i = 0; i < 16; i++) w[i] = w16[i];
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    FullRounds<Tn, 0>::run(a, b, c, d, e, f, g, h, w, q);
    st[0] = add2<F>(st[0], a, q); st[1] = add2<F>(st[1], b, q); st[2] = add2<F>(st[2], c, q); st[3] = add2<F>(st[3], d, q);
    st[4] = add2<F>(st[4], e, q); st[5] = add2<F>(st[5], f, q); st[6] = add2<F>(st[6], g, q); st[7] = add2<F>(st[7], h, q);
}

// =========================================================================================
// v1: 138-byte message.  Tail words: W0 = bswap(ts) (per launch), W1 = difficulty|b0 b1 (per
// lane), W2 = b2 b3|0x8000 (per nonce), W3..W14 = 0, W15 = 1104.
// =========================================================================================
struct V1Lane {            // everything that is invariant while a lane walks H (per item)
    uint32_t a2, e2;       // state after round 1 (b2,c2,d2,f2,g2,h2 are Tail::pre values)
    uint32_t b2, c2, d2, f2, g2, h2;
    uint32_t pe3, pa3;     // round 2: e3 = pe3 + W2, a3 = pa3 + W2
    uint32_t hk3, hk4, hk5;   // h + K for rounds 3..5 (h is still invariant there)
    uint32_t w16, p17, p18, p31, kw16;
};

UPOW_HD V1Lane v1_lane_setup(const Tail& t, uint32_t L) {
    V1Lane s;
    const uint32_t w0 = t.w[0], w1 = t.w[1] | v1_w1_bits(L), w15 = 138u * 8u;
    uint32_t a = t.pre[0], b = t.pre[1], c = t.pre[2], d = t.pre[3], e = t.pre[4], f = t.pre[5], g = t.pre[6],
             h = t.pre[7];
    {   // round 1 (W1)
        uint32_t t1 = h + big_sigma1(e) + ch(e, f, g) + Kc(1) + w1;
        uint32_t t2 = big_sigma0(a) + maj(a, b, c);
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    s.a2 = a; s.b2 = b; s.c2 = c; s.d2 = d; s.e2 = e; s.f2 = f; s.g2 = g; s.h2 = h;
    {   // round 2 without its W2 term
        const uint32_t t1p = h + big_sigma1(e) + ch(e, f, g) + Kc(2);
        s.pe3 = d + t1p;
        s.pa3 = t1p + big_sigma0(a) + maj(a, b, c);
    }
    // after round 2: (a3, a2, b2, c2 | e3, e2, f2, g2): h of round 3 = g2, round 4 = f2, round 5 = e2
    s.hk3 = g + Kc(3);
    s.hk4 = f + Kc(4);
    s.hk5 = e + Kc(5);
    s.w16 = small_sigma0(w1) + w0;
    s.p17 = small_sigma1(w15) + w1;
    s.p18 = small_sigma1(s.w16);
    s.p31 = small_sigma0(s.w16) + w15;
    s.kw16 = Kc(16) + s.w16;
    return s;
}

// constant-K rounds 6..15 of v1 (W3..W14 = 0, W15 = 1104): h is variable, K+W is a literal
template <class Tn, int T>
struct V1ConstRounds {
    static UPOW_HD void run(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d, uint32_t& e, uint32_t& f,
                            uint32_t& g, uint32_t& h, const Opaque& q) {
        constexpr uint32_t KW = Kc(T) + (T == 15 ? 138u * 8u : 0u);
        round_core<Tn, T>(a, b, c, d, e, f, g, h, addk<Tn::ADD != 0, KW>(h, q), q);
        V1ConstRounds<Tn, T + 1>::run(a, b, c, d, e, f, g, h, q);
    }
};
template <class Tn>
struct V1ConstRounds<Tn, 16> {
    static UPOW_HD void run(uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&,
                            const Opaque&) {}
};

// rounds 17..31 of v1: K + W[t] per nonce, schedule words already in w[]
template <class Tn, int T, int END>
struct VarRounds {
    static UPOW_HD void run(uint32_t& a, uint32_t& b, uint32_t& c, uint32_t& d, uint32_t& e, uint32_t& f,
                            uint32_t& g, uint32_t& h, const uint32_t (&w)[64], const Opaque& q) {
        constexpr bool F = Tn::ADD != 0;
        round_core<Tn, T>(a, b, c, d, e, f, g, h, add2<F>(h, addk<F, Kc(T)>(w[T], q), q), q);
        VarRounds<Tn, T + 1, END>::run(a, b, c, d, e, f, g, h, w, q);
    }
};
template <class Tn, int END>
struct VarRounds<Tn, END, END> {
    static UPOW_HD void run(uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&, uint32_t&,
                            const uint32_t (&)[64], const Opaque&) {}
};

// a after round 63 for one nonce of lane `s` whose W2 is `w2`.
template <class Tn>
UPOW_HD uint32_t v1_nonce_a64(const V1Lane& s, uint32_t w2, const Opaque& q) {
    constexpr bool F = Tn::ADD != 0;
    constexpr uint32_t W15 = 138u * 8u;
    uint32_t w[64];
    // ---- schedule words 17..31 (everything that is zero or invariant is already folded) ----
    w[16] = s.w16;
    w[17] = add2<F>(s.p17, sig0<Tn, 17>(w2, q), q);
    w[18] = add2<F>(s.p18, w2, q);
    w[19] = sig1<Tn, 19>(w[17], q);
    w[20] = sig1<Tn, 20>(w[18], q);
    w[21] = sig1<Tn, 21>(w[19], q);
    w[22] = addk<F, W15>(sig1<Tn, 22>(w[20], q), q);
    w[23] = add2<F>(sig1<Tn, 23>(w[21], q), s.w16, q);
    w[24] = add2<F>(sig1<Tn, 24>(w[22], q), w[17], q);
    w[25] = add2<F>(sig1<Tn, 25>(w[23], q), w[18], q);
    w[26] = add2<F>(sig1<Tn, 26>(w[24], q), w[19], q);
    w[27] = add2<F>(sig1<Tn, 27>(w[25], q), w[20], q);
    w[28] = add2<F>(sig1<Tn, 28>(w[26], q), w[21], q);
    w[29] = add2<F>(sig1<Tn, 29>(w[27], q), w[22], q);
    constexpr uint32_t S0_W15 = ((W15 >> 7) | (W15 << 25)) ^ ((W15 >> 18) | (W15 << 14)) ^ (W15 >> 3);
    w[30] = add2<F>(addk<F, S0_W15>(w[23], q), sig1<Tn, 30>(w[28], q), q);
    w[31] = add2<F>(add2<F>(w[24], s.p31, q), sig1<Tn, 31>(w[29], q), q);

    // ---- rounds ----
    uint32_t a = add2<F>(s.pa3, w2, q), e = add2<F>(s.pe3, w2, q);   // round 2
    uint32_t b = s.a2, c = s.b2, d = s.c2, f = s.e2, g = s.f2, h = s.g2;
    round_core<Tn, 3, true, true, true>(a, b, c, d, e, f, g, h, s.hk3, q);   // h = g2, d = c2: uniform
    round_core<Tn, 4, true, true, true>(a, b, c, d, e, f, g, h, s.hk4, q);   // h = f2, d = b2: uniform
    round_core<Tn, 5>(a, b, c, d, e, f, g, h, s.hk5, q);                     // h = e2, d = a2: per lane
    V1ConstRounds<Tn, 6>::run(a, b, c, d, e, f, g, h, q);
    round_core<Tn, 16>(a, b, c, d, e, f, g, h, add2<F>(h, s.kw16, q), q);
    VarRounds<Tn, 17, 32>::run(a, b, c, d, e, f, g, h, w, q);
    return GenericTail<Tn, 32, 32>::run(a, b, c, d, e, f, g, h, w, q);
}

// =========================================================================================
// v2: 108-byte message.  W0..W9 per launch, W10 = bswap(nonce), W11 = 0x80000000,
// W12..W14 = 0, W15 = 864.  Rounds 0..9 are done on the host (Tail::pre); every schedule
// term that depends on the template only is summed on the host into V2Consts.
// =========================================================================================
struct V2Consts {
    uint32_t ue11, ua11;          // round 10: e11 = ue11 + W10, a11 = ua11 + W10
    uint32_t b11, c11, d11, f11, g11, h11;   // the rest of the state after round 10
    uint32_t hk11, hk12, hk13;    // h + K + W for rounds 11..13
    uint32_t kw16, kw18, kw20, kw22;   // K + W for the constant schedule words
    uint32_t w16, w18, w20, w22;
    uint32_t u17, u19, u21, u23, u24, u25, u26, u27, u29, u30, u31, u33, u35, u37;
};

inline void prepare_v2_consts(const Tail& t, V2Consts* u) {
    const uint32_t* w = t.w;  // W10 is zero in the template
    const uint32_t W11 = 0x80000000u, W15 = 108u * 8u;
    uint32_t a = t.pre[0], b = t.pre[1], c = t.pre[2], d = t.pre[3], e = t.pre[4], f = t.pre[5], g = t.pre[6],
             h = t.pre[7];
    const uint32_t t1p = h + big_sigma1(e) + ch(e, f, g) + Kc(10);
    u->ue11 = d + t1p;
    u->ua11 = t1p + big_sigma0(a) + maj(a, b, c);
    u->b11 = a; u->c11 = b; u->d11 = c; u->f11 = e; u->g11 = f; u->h11 = g;
    // after round 10: (a11, a, b, c | e11, e, f, g): h of round 11 = g, round 12 = f, round 13 = e
    u->hk11 = g + Kc(11) + W11;
    u->hk12 = f + Kc(12);
    u->hk13 = e + Kc(13);
    u->w16 = w[9] + small_sigma0(w[1]) + w[0];                              // sigma1(W14 = 0) = 0
    u->w18 = small_sigma1(u->w16) + W11 + small_sigma0(w[3]) + w[2];
    u->w20 = small_sigma1(u->w18) + small_sigma0(w[5]) + w[4];              // W13 = 0
    u->w22 = small_sigma1(u->w20) + W15 + small_sigma0(w[7]) + w[6];
    u->kw16 = Kc(16) + u->w16; u->kw18 = Kc(18) + u->w18; u->kw20 = Kc(20) + u->w20; u->kw22 = Kc(22) + u->w22;
    u->u17 = small_sigma1(W15) + small_sigma0(w[2]) + w[1];                 // W17 = u17 + W10
    u->u19 = small_sigma0(w[4]) + w[3];                                     // W19 = sigma1(W17) + u19   (W12 = 0)
    u->u21 = small_sigma0(w[6]) + w[5];                                     // W21 = sigma1(W19) + u21   (W14 = 0)
    u->u23 = u->w16 + small_sigma0(w[8]) + w[7];                            // W23 = sigma1(W21) + u23
    u->u24 = small_sigma1(u->w22) + small_sigma0(w[9]) + w[8];              // W24 = u24 + W17
    u->u25 = u->w18 + w[9];                                                 // W25 = sigma1(W23) + sigma0(W10) + u25
    u->u26 = small_sigma0(W11);                                             // W26 = sigma1(W24) + W19 + W10 + u26
    u->u27 = u->w20 + W11;                                                  // W27 = sigma1(W25) + u27   (sigma0(W12)=0)
    u->u29 = u->w22;                                                        // W29 = sigma1(W27) + u29
    u->u30 = small_sigma0(W15);                                             // W30 = sigma1(W28) + W23 + u30
    u->u31 = small_sigma0(u->w16) + W15;                                    // W31 = sigma1(W29) + W24 + u31
    u->u33 = small_sigma0(u->w18);                                          // W33 = sigma1(W31) + W26 + W17 + u33
    u->u35 = small_sigma0(u->w20);                                          // W35 = sigma1(W33) + W28 + W19 + u35
    u->u37 = small_sigma0(u->w22);                                          // W37 = sigma1(W35) + W30 + W21 + u37
}

template <class Tn>
UPOW_HD uint32_t v2_nonce_a64(const V2Consts& u, uint32_t w10, const Opaque& q) {
    constexpr bool F = Tn::ADD != 0;
    uint32_t w[64];
    w[16] = u.w16; w[18] = u.w18; w[20] = u.w20; w[22] = u.w22;
    w[17] = addu<F>(w10, u.u17, q);
    w[19] = addu<F>(sig1<Tn, 19>(w[17], q), u.u19, q);
    w[21] = addu<F>(sig1<Tn, 21>(w[19], q), u.u21, q);
    w[23] = addu<F>(sig1<Tn, 23>(w[21], q), u.u23, q);
    w[24] = addu<F>(w[17], u.u24, q);
    w[25] = add2<F>(addu<F>(sig0<Tn, 25>(w10, q), u.u25, q), sig1<Tn, 25>(w[23], q), q);
    w[26] = add2<F>(add2<F>(addu<F>(w10, u.u26, q), w[19], q), sig1<Tn, 26>(w[24], q), q);
    w[27] = addu<F>(sig1<Tn, 27>(w[25], q), u.u27, q);
    w[28] = add2<F>(sig1<Tn, 28>(w[26], q), w[21], q);
    w[29] = addu<F>(sig1<Tn, 29>(w[27], q), u.u29, q);
    w[30] = add2<F>(addu<F>(w[23], u.u30, q), sig1<Tn, 30>(w[28], q), q);
    w[31] = add2<F>(addu<F>(w[24], u.u31, q), sig1<Tn, 31>(w[29], q), q);
    w[32] = add2<F>(add2<F>(addu<F>(w[25], u.w16, q), sig0<Tn, 32>(w[17], q), q), sig1<Tn, 32>(w[30], q), q);
    w[33] = add2<F>(addu<F>(add2<F>(w[26], w[17], q), u.u33, q), sig1<Tn, 33>(w[31], q), q);
    w[34] = add2<F>(add2<F>(addu<F>(w[27], u.w18, q), sig0<Tn, 34>(w[19], q), q), sig1<Tn, 34>(w[32], q), q);
    w[35] = add2<F>(addu<F>(add2<F>(w[28], w[19], q), u.u35, q), sig1<Tn, 35>(w[33], q), q);
    w[36] = add2<F>(add2<F>(addu<F>(w[29], u.w20, q), sig0<Tn, 36>(w[21], q), q), sig1<Tn, 36>(w[34], q), q);
    w[37] = add2<F>(addu<F>(add2<F>(w[30], w[21], q), u.u37, q), sig1<Tn, 37>(w[35], q), q);
    w[38] = add2<F>(add2<F>(addu<F>(w[31], u.w22, q), sig0<Tn, 38>(w[23], q), q), sig1<Tn, 38>(w[36], q), q);
    uint32_t a = addu<F>(w10, u.ua11, q), e = addu<F>(w10, u.ue11, q);   // round 10
    uint32_t b = u.b11, c = u.c11, d = u.d11, f = u.f11, g = u.g11, h = u.h11;
    round_core<Tn, 11, true, true, true>(a, b, c, d, e, f, g, h, u.hk11, q);   // h, d still uniform
    round_core<Tn, 12, true, true, true>(a, b, c, d, e, f, g, h, u.hk12, q);
    round_core<Tn, 13, true, true, true>(a, b, c, d, e, f, g, h, u.hk13, q);
    round_core<Tn, 14>(a, b, c, d, e, f, g, h, addk<F, Kc(14)>(h, q), q);
    round_core<Tn, 15>(a, b, c, d, e, f, g, h, addk<F, Kc(15) + 108u * 8u>(h, q), q);
    round_core<Tn, 16>(a, b, c, d, e, f, g, h, addu<F>(h, u.kw16, q), q);
    round_core<Tn, 17>(a, b, c, d, e, f, g, h, add2<F>(h, addk<F, Kc(17)>(w[17], q), q), q);
    round_core<Tn, 18>(a, b, c, d, e, f, g, h, addu<F>(h, u.kw18, q), q);
    round_core<Tn, 19>(a, b, c, d, e, f, g, h, add2<F>(h, addk<F, Kc(19)>(w[19], q), q), q);
    round_core<Tn, 20>(a, b, c, d, e, f, g, h, addu<F>(h, u.kw20, q), q);
    round_core<Tn, 21>(a, b, c, d, e, f, g, h, add2<F>(h, addk<F, Kc(21)>(w[21], q), q), q);
    round_core<Tn, 22>(a, b, c, d, e, f, g, h, addu<F>(h, u.kw22, q), q);
    VarRounds<Tn, 23, 39>::run(a, b, c, d, e, f, g, h, w, q);
    return GenericTail<Tn, 39, 39>::run(a, b, c, d, e, f, g, h, w, q);
}

}  // namespace upow

// Opcode issue-rate microbenchmarks: the denominators of the INT-pipe roofline
// (H_peak = SMs * f_SM / max(N_alu/R_alu, N_fma/R_fma, N_total/R_issue)).
//
// Every SM is filled with 2048 threads; each thread runs ITERS x 256 inline-PTX ops on 8
// independent accumulator chains.  Thread 0 of every block samples clock64() around the
// loop; rate = warp-instructions issued per SM / max block cycles.  %globaltimer around the
// same region gives the SM clock the benchmark really ran at.
// The SASS opcode each case produces is checked offline with cuobjdump (profiles/).
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/upow_b200.h"

namespace {

enum Op : int {
    OP_IADD3 = 0,    // add.u32            -> IADD3              (alu pipe)
    OP_LOP3,         // lop3.b32           -> LOP3.LUT           (alu)
    OP_SHF_ROT,      // shf.r.wrap.b32     -> SHF.R.W.U32        (alu)
    OP_SHR,          // shr.u32            -> SHF.R.U32.HI       (alu)
    OP_PRMT,         // prmt.b32           -> PRMT               (alu)
    OP_IMAD,         // mad.lo.u32 a*m+b   -> IMAD               (fma pipe)
    OP_IMAD_ADD,     // mad.lo.u32 b*1+a, "1" opaque -> IMAD     (fma): a 2-input add on the fma pipe
    OP_IMAD_HI,      // mad.hi.u32         -> IMAD.HI.U32        (fma): x>>k as hi(x * 2^(32-k))
    OP_IMAD_WIDE,    // mad.wide.u32       -> IMAD.WIDE.U32      (fma)
    OP_MIX_SHF_IMAD,      // 1 SHF : 1 IMAD_ADD interleaved      (both pipes)
    OP_MIX_LOP3_IMAD,     // 1 LOP3 : 1 IMAD_ADD
    OP_MIX_2ALU_1IMAD,    // SHF, LOP3, IMAD_ADD
    OP_MIX_3ALU_1IMAD,    // SHF, SHF, LOP3, IMAD_ADD
    OP_MIX_SHF_IMADHI,    // 1 SHF : 1 IMAD.HI
    OP_MIX_SHF_LOP3,      // 1 SHF : 1 LOP3 (same pipe; should equal the single-op rate)
    // ---- register-file read bandwidth: operands rotate over 8 live registers so that the
    //      operand-reuse cache cannot help; "Nr" = distinct register reads per instruction ----
    OP_LOP3_3R,           // lop3 a_i, a_i, a_j, a_k
    OP_IMAD_3R,           // mad.lo a_i = a_j * a_k + a_i
    OP_IMAD_2R_IMM,       // mad.lo a_i = a_j * imm + a_i
    OP_IADD3_3R,          // a_i = a_i + a_j + a_k   (one IADD3 after ptxas fusion)
    OP_MIX_LOP3_3R_IMAD_3R,     // 6 reads per pair
    OP_MIX_LOP3_3R_IMAD_2R,     // 5 reads per pair
    OP_MIX_SHF_1R_IMAD_3R,      // 4 reads per pair
    OP_MIX_SHF_1R_IMAD_2R,      // 3 reads per pair
    OP_MIX_LOP3_2R_IMAD_2R,     // 4 reads per pair (lop3 with an immediate)
    OP_MIX_SHA_LIKE,            // 3 SHF(1r) + 2 LOP3(3r) + 4 IMAD(2r imm-one) per 9: close to the tuned round mix
    OP_COUNT
};

const char* const OP_NAMES[OP_COUNT] = {
    "IADD3", "LOP3", "SHF.rot", "SHF.shr", "PRMT", "IMAD", "IMAD.add(c[1])", "IMAD.HI", "IMAD.WIDE",
    "mix SHF+IMAD 1:1", "mix LOP3+IMAD 1:1", "mix SHF,LOP3,IMAD 2:1", "mix SHF,SHF,LOP3,IMAD 3:1",
    "mix SHF+IMAD.HI 1:1", "mix SHF+LOP3 1:1",
    "LOP3 3 reg reads", "IMAD 3 reg reads", "IMAD 2 reg reads (imm)", "IADD3 3 reg reads",
    "mix LOP3(3r)+IMAD(3r)", "mix LOP3(3r)+IMAD(2r)", "mix SHF(1r)+IMAD(3r)", "mix SHF(1r)+IMAD(2r)",
    "mix LOP3(2r)+IMAD(2r)", "mix sha-like 3SHF+2LOP3+4IMAD(2r)"};

struct MbArgs {
    uint32_t m;      // opaque multiplier (a power of two at run time; unknown to ptxas)
    uint32_t one;    // opaque 1
    uint32_t b, c;   // opaque operands
    uint32_t iters;
    uint32_t magic;
    unsigned long long* cycles;   // [gridDim.x]
    unsigned long long* ns;       // [gridDim.x]
    uint32_t* sink;
};

#define A_IADD3(x) asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(b))  /* ptxas folds repeated adds: kept only as a curiosity */
#define A_LOP3(x) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(b), "r"(c))
#define A_ROT(x) asm volatile("shf.r.wrap.b32 %0, %0, %0, 7;" : "+r"(x))
#define A_SHR(x) asm volatile("shf.r.clamp.b32 %0, %0, %1, 3;" : "+r"(x) : "r"(b))
#define A_PRMT(x) asm volatile("prmt.b32 %0, %0, %1, 0x2103;" : "+r"(x) : "r"(b))
#define A_IMAD(x) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(m), "r"(b))
#define A_IMADADD(x) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x) : "r"(b), "r"(one))
#define A_IMADHI(x) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x) : "r"(m), "r"(b))
#define A_IMADWIDE(x, y)                                                                     \
    asm volatile("{.reg .u64 t; mad.wide.u32 t, %0, %2, %3; mov.b64 {%0, %1}, t;}"           \
                 : "+r"(x), "+r"(y) : "r"(m), "l"(wide_c))

// rotating-operand forms: X is updated from itself and two (one) other live accumulators
#define R_LOP3(x, y, z) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(y), "r"(z))
#define R_LOP3I(x, y) asm volatile("lop3.b32 %0, %0, %1, 0x5a5a5a5a, 0x96;" : "+r"(x) : "r"(y))
#define R_IMAD3(x, y, z) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(x) : "r"(y), "r"(z))
#define R_IMAD2(x, y) asm volatile("mad.lo.u32 %0, %1, 0x9e3779b1, %0;" : "+r"(x) : "r"(y))
#define R_IADD3(x, y, z) asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(x) : "r"(y), "r"(z))
#define R_ROT(x) asm volatile("shf.r.wrap.b32 %0, %0, %0, 7;" : "+r"(x))

// 8 chains x 4 = 32 ops per REP4 invocation
#define REP8(OPM) OPM(a0); OPM(a1); OPM(a2); OPM(a3); OPM(a4); OPM(a5); OPM(a6); OPM(a7)

template <int OP>
__global__ void __launch_bounds__(256) mb_kernel(const __grid_constant__ MbArgs A) {
    const uint32_t m = A.m, one = A.one, b = A.b, c = A.c;
    const unsigned long long wide_c = ((unsigned long long)A.c << 32) | A.b;
    uint32_t a0 = threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6,
             a7 = a0 + 7;
    (void)m; (void)one; (void)c; (void)wide_c;
    __syncthreads();
    unsigned long long t0 = clock64(), n0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(n0));
    for (uint32_t it = 0; it < A.iters; it++) {
#pragma unroll
        for (int r = 0; r < 16; r++) {  // 16 x 16 = 256 ops per trip
            if (OP == OP_IADD3) { REP8(A_IADD3); REP8(A_IADD3); }
            if (OP == OP_LOP3) { REP8(A_LOP3); REP8(A_LOP3); }
            if (OP == OP_SHF_ROT) { REP8(A_ROT); REP8(A_ROT); }
            if (OP == OP_SHR) { REP8(A_SHR); REP8(A_SHR); }
            if (OP == OP_PRMT) { REP8(A_PRMT); REP8(A_PRMT); }
            if (OP == OP_IMAD) { REP8(A_IMAD); REP8(A_IMAD); }
            if (OP == OP_IMAD_ADD) { REP8(A_IMADADD); REP8(A_IMADADD); }
            if (OP == OP_IMAD_HI) { REP8(A_IMADHI); REP8(A_IMADHI); }
            if (OP == OP_IMAD_WIDE) {
                A_IMADWIDE(a0, a1); A_IMADWIDE(a2, a3); A_IMADWIDE(a4, a5); A_IMADWIDE(a6, a7);
                A_IMADWIDE(a0, a1); A_IMADWIDE(a2, a3); A_IMADWIDE(a4, a5); A_IMADWIDE(a6, a7);
                A_IMADWIDE(a0, a1); A_IMADWIDE(a2, a3); A_IMADWIDE(a4, a5); A_IMADWIDE(a6, a7);
                A_IMADWIDE(a0, a1); A_IMADWIDE(a2, a3); A_IMADWIDE(a4, a5); A_IMADWIDE(a6, a7);
            }
            if (OP == OP_MIX_SHF_IMAD) {
                A_ROT(a0); A_IMADADD(a1); A_ROT(a2); A_IMADADD(a3); A_ROT(a4); A_IMADADD(a5); A_ROT(a6); A_IMADADD(a7);
                A_IMADADD(a0); A_ROT(a1); A_IMADADD(a2); A_ROT(a3); A_IMADADD(a4); A_ROT(a5); A_IMADADD(a6); A_ROT(a7);
            }
            if (OP == OP_MIX_LOP3_IMAD) {
                A_LOP3(a0); A_IMADADD(a1); A_LOP3(a2); A_IMADADD(a3); A_LOP3(a4); A_IMADADD(a5); A_LOP3(a6); A_IMADADD(a7);
                A_IMADADD(a0); A_LOP3(a1); A_IMADADD(a2); A_LOP3(a3); A_IMADADD(a4); A_LOP3(a5); A_IMADADD(a6); A_LOP3(a7);
            }
            if (OP == OP_MIX_2ALU_1IMAD) {  // 18 ops: 12 alu + 6 imad (counted exactly below)
                A_ROT(a0); A_LOP3(a1); A_IMADADD(a2); A_ROT(a3); A_LOP3(a4); A_IMADADD(a5);
                A_ROT(a6); A_LOP3(a7); A_IMADADD(a0); A_ROT(a1); A_LOP3(a2); A_IMADADD(a3);
                A_ROT(a4); A_LOP3(a5); A_IMADADD(a6); A_ROT(a7); A_LOP3(a0); A_IMADADD(a1);
            }
            if (OP == OP_MIX_3ALU_1IMAD) {  // 16 ops: 12 alu + 4 imad
                A_ROT(a0); A_ROT(a1); A_LOP3(a2); A_IMADADD(a3); A_ROT(a4); A_ROT(a5); A_LOP3(a6); A_IMADADD(a7);
                A_ROT(a2); A_ROT(a3); A_LOP3(a0); A_IMADADD(a1); A_ROT(a6); A_ROT(a7); A_LOP3(a4); A_IMADADD(a5);
            }
            if (OP == OP_MIX_SHF_IMADHI) {
                A_ROT(a0); A_IMADHI(a1); A_ROT(a2); A_IMADHI(a3); A_ROT(a4); A_IMADHI(a5); A_ROT(a6); A_IMADHI(a7);
                A_IMADHI(a0); A_ROT(a1); A_IMADHI(a2); A_ROT(a3); A_IMADHI(a4); A_ROT(a5); A_IMADHI(a6); A_ROT(a7);
            }
            if (OP == OP_LOP3_3R) {
                R_LOP3(a0, a1, a2); R_LOP3(a3, a4, a5); R_LOP3(a6, a7, a0); R_LOP3(a1, a2, a3);
                R_LOP3(a4, a5, a6); R_LOP3(a7, a0, a1); R_LOP3(a2, a3, a4); R_LOP3(a5, a6, a7);
                R_LOP3(a0, a3, a6); R_LOP3(a1, a4, a7); R_LOP3(a2, a5, a0); R_LOP3(a3, a6, a1);
                R_LOP3(a4, a7, a2); R_LOP3(a5, a0, a3); R_LOP3(a6, a1, a4); R_LOP3(a7, a2, a5);
            }
            if (OP == OP_IMAD_3R) {
                R_IMAD3(a0, a1, a2); R_IMAD3(a3, a4, a5); R_IMAD3(a6, a7, a0); R_IMAD3(a1, a2, a3);
                R_IMAD3(a4, a5, a6); R_IMAD3(a7, a0, a1); R_IMAD3(a2, a3, a4); R_IMAD3(a5, a6, a7);
                R_IMAD3(a0, a3, a6); R_IMAD3(a1, a4, a7); R_IMAD3(a2, a5, a0); R_IMAD3(a3, a6, a1);
                R_IMAD3(a4, a7, a2); R_IMAD3(a5, a0, a3); R_IMAD3(a6, a1, a4); R_IMAD3(a7, a2, a5);
            }
            if (OP == OP_IMAD_2R_IMM) {
                R_IMAD2(a0, a1); R_IMAD2(a2, a3); R_IMAD2(a4, a5); R_IMAD2(a6, a7);
                R_IMAD2(a1, a2); R_IMAD2(a3, a4); R_IMAD2(a5, a6); R_IMAD2(a7, a0);
                R_IMAD2(a0, a3); R_IMAD2(a2, a5); R_IMAD2(a4, a7); R_IMAD2(a6, a1);
                R_IMAD2(a1, a4); R_IMAD2(a3, a6); R_IMAD2(a5, a0); R_IMAD2(a7, a2);
            }
            if (OP == OP_IADD3_3R) {
                R_IADD3(a0, a1, a2); R_IADD3(a3, a4, a5); R_IADD3(a6, a7, a0); R_IADD3(a1, a2, a3);
                R_IADD3(a4, a5, a6); R_IADD3(a7, a0, a1); R_IADD3(a2, a3, a4); R_IADD3(a5, a6, a7);
                R_IADD3(a0, a3, a6); R_IADD3(a1, a4, a7); R_IADD3(a2, a5, a0); R_IADD3(a3, a6, a1);
                R_IADD3(a4, a7, a2); R_IADD3(a5, a0, a3); R_IADD3(a6, a1, a4); R_IADD3(a7, a2, a5);
            }
            if (OP == OP_MIX_LOP3_3R_IMAD_3R) {
                R_LOP3(a0, a1, a2); R_IMAD3(a3, a4, a5); R_LOP3(a6, a7, a0); R_IMAD3(a1, a2, a3);
                R_LOP3(a4, a5, a6); R_IMAD3(a7, a0, a1); R_LOP3(a2, a3, a4); R_IMAD3(a5, a6, a7);
                R_LOP3(a0, a3, a6); R_IMAD3(a1, a4, a7); R_LOP3(a2, a5, a0); R_IMAD3(a3, a6, a1);
                R_LOP3(a4, a7, a2); R_IMAD3(a5, a0, a3); R_LOP3(a6, a1, a4); R_IMAD3(a7, a2, a5);
            }
            if (OP == OP_MIX_LOP3_3R_IMAD_2R) {
                R_LOP3(a0, a1, a2); R_IMAD2(a3, a4); R_LOP3(a6, a7, a0); R_IMAD2(a1, a2);
                R_LOP3(a4, a5, a6); R_IMAD2(a7, a0); R_LOP3(a2, a3, a4); R_IMAD2(a5, a6);
                R_LOP3(a0, a3, a6); R_IMAD2(a1, a4); R_LOP3(a2, a5, a0); R_IMAD2(a3, a6);
                R_LOP3(a4, a7, a2); R_IMAD2(a5, a0); R_LOP3(a6, a1, a4); R_IMAD2(a7, a2);
            }
            if (OP == OP_MIX_SHF_1R_IMAD_3R) {
                R_ROT(a0); R_IMAD3(a3, a4, a5); R_ROT(a6); R_IMAD3(a1, a2, a3);
                R_ROT(a4); R_IMAD3(a7, a0, a1); R_ROT(a2); R_IMAD3(a5, a6, a7);
                R_ROT(a0); R_IMAD3(a1, a4, a7); R_ROT(a2); R_IMAD3(a3, a6, a1);
                R_ROT(a4); R_IMAD3(a5, a0, a3); R_ROT(a6); R_IMAD3(a7, a2, a5);
            }
            if (OP == OP_MIX_SHF_1R_IMAD_2R) {
                R_ROT(a0); R_IMAD2(a3, a4); R_ROT(a6); R_IMAD2(a1, a2);
                R_ROT(a4); R_IMAD2(a7, a0); R_ROT(a2); R_IMAD2(a5, a6);
                R_ROT(a0); R_IMAD2(a1, a4); R_ROT(a2); R_IMAD2(a3, a6);
                R_ROT(a4); R_IMAD2(a5, a0); R_ROT(a6); R_IMAD2(a7, a2);
            }
            if (OP == OP_MIX_LOP3_2R_IMAD_2R) {
                R_LOP3I(a0, a1); R_IMAD2(a3, a4); R_LOP3I(a6, a7); R_IMAD2(a1, a2);
                R_LOP3I(a4, a5); R_IMAD2(a7, a0); R_LOP3I(a2, a3); R_IMAD2(a5, a6);
                R_LOP3I(a0, a3); R_IMAD2(a1, a4); R_LOP3I(a2, a5); R_IMAD2(a3, a6);
                R_LOP3I(a4, a7); R_IMAD2(a5, a0); R_LOP3I(a6, a1); R_IMAD2(a7, a2);
            }
            if (OP == OP_MIX_SHA_LIKE) {  // 18 ops: 6 SHF + 4 LOP3(3r) + 8 IMAD(2r)
                R_ROT(a0); R_IMAD2(a1, a2); R_ROT(a3); R_IMAD2(a4, a5); R_ROT(a6); R_IMAD2(a7, a0);
                R_LOP3(a0, a3, a6); R_IMAD2(a2, a1); R_LOP3(a1, a4, a7); R_IMAD2(a5, a3);
                R_ROT(a2); R_IMAD2(a6, a4); R_ROT(a5); R_IMAD2(a0, a7); R_ROT(a7); R_IMAD2(a3, a1);
                R_LOP3(a2, a5, a7); R_LOP3(a4, a6, a0);
            }
            if (OP == OP_MIX_SHF_LOP3) {
                A_ROT(a0); A_LOP3(a1); A_ROT(a2); A_LOP3(a3); A_ROT(a4); A_LOP3(a5); A_ROT(a6); A_LOP3(a7);
                A_LOP3(a0); A_ROT(a1); A_LOP3(a2); A_ROT(a3); A_LOP3(a4); A_ROT(a5); A_LOP3(a6); A_ROT(a7);
            }
        }
    }
    unsigned long long t1 = clock64(), n1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(n1));
    uint32_t x = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
    if (x == A.magic) A.sink[0] = x;
    if (threadIdx.x == 0) {
        A.cycles[blockIdx.x] = t1 - t0;
        A.ns[blockIdx.x] = n1 - n0;
    }
}

int ops_per_trip(int op) {
    if (op == OP_MIX_2ALU_1IMAD) return 16 * 18;
    if (op == OP_MIX_3ALU_1IMAD) return 16 * 16;
    if (op == OP_MIX_SHA_LIKE) return 16 * 18;
    return 256;
}
int fma_ops_per_trip(int op) {
    switch (op) {
        case OP_IMAD: case OP_IMAD_ADD: case OP_IMAD_HI: case OP_IMAD_WIDE: return 256;
        case OP_MIX_SHF_IMAD: case OP_MIX_LOP3_IMAD: case OP_MIX_SHF_IMADHI: return 128;
        case OP_IMAD_3R: case OP_IMAD_2R_IMM: return 256;
        case OP_MIX_LOP3_3R_IMAD_3R: case OP_MIX_LOP3_3R_IMAD_2R: case OP_MIX_SHF_1R_IMAD_3R:
        case OP_MIX_SHF_1R_IMAD_2R: case OP_MIX_LOP3_2R_IMAD_2R: return 128;
        case OP_MIX_SHA_LIKE: return 16 * 8;
        case OP_MIX_2ALU_1IMAD: return 16 * 6;
        case OP_MIX_3ALU_1IMAD: return 16 * 4;
        default: return 0;
    }
}

template <int OP>
cudaError_t run_one(const MbArgs& a, int grid, cudaStream_t s) {
    mb_kernel<OP><<<grid, 256, 0, s>>>(a);
    return cudaGetLastError();
}

template <int OP>
struct Runner {
    static cudaError_t go(int op, const MbArgs& a, int grid, cudaStream_t s) {
        if (op == OP) return run_one<OP>(a, grid, s);
        return Runner<OP + 1>::go(op, a, grid, s);
    }
};
template <>
struct Runner<OP_COUNT> {
    static cudaError_t go(int, const MbArgs&, int, cudaStream_t) { return cudaErrorInvalidValue; }
};

}  // namespace

extern "C" int upow_microbench_count(void) { return OP_COUNT; }

extern "C" int upow_microbench(int device, int op, uint32_t iters, upow_microbench_result* out) {
    if (!out || op < 0 || op >= OP_COUNT) return UPOW_E_ARG;
    memset(out, 0, sizeof *out);
    strncpy(out->name, OP_NAMES[op], sizeof out->name - 1);
    if (cudaSetDevice(device) != cudaSuccess) return UPOW_E_CUDA;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return UPOW_E_CUDA;
    const int blocks_per_sm = prop.maxThreadsPerMultiProcessor / 256;
    const int grid = prop.multiProcessorCount * blocks_per_sm;
    if (iters == 0) iters = 2000;
    unsigned long long *d_cycles = 0, *d_ns = 0;
    uint32_t* d_sink = 0;
    int rc = UPOW_E_CUDA;
    cudaEvent_t e0 = 0, e1 = 0;
    unsigned long long* h = 0;
    do {
        if (cudaMalloc(&d_cycles, sizeof(unsigned long long) * grid) != cudaSuccess) break;
        if (cudaMalloc(&d_ns, sizeof(unsigned long long) * grid) != cudaSuccess) break;
        if (cudaMalloc(&d_sink, 64) != cudaSuccess) break;
        if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) break;
        MbArgs a;
        a.m = 1u << 13; a.one = 1; a.b = 0x9e3779b9u; a.c = 0x7f4a7c15u; a.iters = iters; a.magic = 0xdeadbeefu;
        a.cycles = d_cycles; a.ns = d_ns; a.sink = d_sink;
        // warm-up launch (clocks ramp, instruction cache), then the timed one
        MbArgs warm = a;
        warm.iters = iters / 4 + 1;
        if (Runner<0>::go(op, warm, grid, 0) != cudaSuccess) break;
        cudaEventRecord(e0, 0);
        if (Runner<0>::go(op, a, grid, 0) != cudaSuccess) break;
        cudaEventRecord(e1, 0);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        h = (unsigned long long*)malloc(sizeof(unsigned long long) * grid * 2);
        if (cudaMemcpy(h, d_cycles, sizeof(unsigned long long) * grid, cudaMemcpyDeviceToHost) != cudaSuccess) break;
        if (cudaMemcpy(h + grid, d_ns, sizeof(unsigned long long) * grid, cudaMemcpyDeviceToHost) != cudaSuccess) break;
        unsigned long long max_cyc = 0, max_ns = 0;
        double sum_cyc = 0;
        for (int i = 0; i < grid; i++) {
            if (h[i] > max_cyc) max_cyc = h[i];
            if (h[grid + i] > max_ns) max_ns = h[grid + i];
            sum_cyc += (double)h[i];
        }
        const double warps_per_sm = blocks_per_sm * 8.0;
        const double warp_insts_per_sm = warps_per_sm * (double)iters * ops_per_trip(op);
        out->sm_count = prop.multiProcessorCount;
        out->kernel_ms = ms;
        out->max_block_cycles = (double)max_cyc;
        out->mean_block_cycles = sum_cyc / grid;
        out->warp_inst_per_clk_per_sm = warp_insts_per_sm / (double)max_cyc;
        out->lanes_per_clk_per_sm = out->warp_inst_per_clk_per_sm * 32.0;
        out->fma_fraction = (double)fma_ops_per_trip(op) / ops_per_trip(op);
        out->sm_clock_mhz = max_ns ? (double)max_cyc / (double)max_ns * 1000.0 : 0.0;
        rc = UPOW_OK;
    } while (0);
    if (rc != UPOW_OK) fprintf(stderr, "upow_microbench: CUDA error: %s\n", cudaGetErrorString(cudaGetLastError()));
    free(h);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    cudaFree(d_cycles); cudaFree(d_ns); cudaFree(d_sink);
    return rc;
}

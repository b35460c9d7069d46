// Host side of libupow_b200.so: contexts, template slots, the tile scheduler that keeps the
// exact-minimum guarantee across GPUs, and the extern "C" entry points of include/upow_b200.h.
// No PyTorch, no NCCL: the nonce space shards into independent ranges; the only
// cross-GPU signal is a host-side "found in tile k" marker plus a per-device abort word.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/upow_b200.h"
#include "kernels.cuh"

using namespace upow;

namespace {

constexpr uint32_t ENUM_CAP = 1u << 22;          // device-side capacity for sweep results (16 MiB)
constexpr uint64_t NONCE_SPACE = 1ull << 32;
constexpr uint32_t L2_FLUSH_BYTES = 256u << 20;  // > 126 MB L2

std::string g_init_error;
std::mutex g_init_mu;

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Slot {
    bool used = false;
    uint8_t prefix[136];
    int prefix_len = 0;
    uint8_t prev_hash[32];
    int n_int = 0, charset_count = 16;
};

struct Device {
    int id = -1;
    int sm_count = 0;
    int grid_blocks = 0;   // persistent grid of the strided kernel
    int variant_grid[16] = {0};
    int batch_grid = 0;
    int strided_tuned_grid = 0;
    cudaDeviceProp prop;
    cudaStream_t stream = nullptr, ctrl = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    DevResult* d_res = nullptr;
    uint32_t* d_out = nullptr;
    uint32_t* d_digest = nullptr;
    // pinned host mirrors
    DevResult* h_res = nullptr;
    uint32_t* h_digest = nullptr;
    unsigned int* h_one = nullptr;
    void* d_flush = nullptr;
    // batch
    BatchJob* d_jobs = nullptr;
    uint32_t* d_active = nullptr;
    unsigned long long* d_best = nullptr;
    uint32_t* d_bdigest = nullptr;
    uint32_t batch_cap = 0;
    // block verifier
    VerifyRecord* d_vrecs = nullptr;
    uint8_t* d_vvalid = nullptr;
    uint32_t* d_vdig = nullptr;
    uint32_t verify_cap = 0;
};

}  // namespace

struct upow_ctx {
    std::vector<Device> devs;
    Slot slots[UPOW_MAX_SLOTS];
    std::atomic<int> cancel{0};
    std::mutex err_mu;
    std::string err;
    std::mutex api_mu;  // one search at a time per context
    int variant = -1;   // index into VARIANTS
};

namespace {

void set_err(upow_ctx* ctx, const std::string& s) {
    if (ctx) {
        std::lock_guard<std::mutex> g(ctx->err_mu);
        ctx->err = s;
    } else {
        std::lock_guard<std::mutex> g(g_init_mu);
        g_init_error = s;
    }
}

#define CU_TRY(ctx, expr)                                                                            \
    do {                                                                                             \
        cudaError_t e_ = (expr);                                                                     \
        if (e_ != cudaSuccess) {                                                                     \
            set_err(ctx, std::string(#expr) + ": " + cudaGetErrorString(e_));                        \
            return UPOW_E_CUDA;                                                                      \
        }                                                                                            \
    } while (0)

int init_device(upow_ctx* ctx, Device& d, int id) {
    d.id = id;
    CU_TRY(ctx, cudaSetDevice(id));
    CU_TRY(ctx, cudaGetDeviceProperties(&d.prop, id));
    d.sm_count = d.prop.multiProcessorCount;
    uint32_t k[64];
    for (int i = 0; i < 64; i++) k[i] = Kc(i);
    CU_TRY(ctx, cudaMemcpyToSymbol(K_DEV, k, sizeof k));
    CU_TRY(ctx, cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
    CU_TRY(ctx, cudaStreamCreateWithFlags(&d.ctrl, cudaStreamNonBlocking));
    CU_TRY(ctx, cudaEventCreate(&d.ev0));
    CU_TRY(ctx, cudaEventCreate(&d.ev1));
    CU_TRY(ctx, cudaMalloc(&d.d_res, sizeof(DevResult)));
    CU_TRY(ctx, cudaMalloc(&d.d_out, sizeof(uint32_t) * ENUM_CAP));
    CU_TRY(ctx, cudaMalloc(&d.d_digest, 32));
    CU_TRY(ctx, cudaMallocHost(&d.h_res, sizeof(DevResult)));
    CU_TRY(ctx, cudaMallocHost(&d.h_digest, 32));
    CU_TRY(ctx, cudaMallocHost(&d.h_one, sizeof(unsigned int)));
    *d.h_one = 1;
    int per_sm = 0;
    CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, search_strided_kernel<LAYOUT_V1, MODE_MIN>, 256, 0));
    if (per_sm < 1) per_sm = 1;
    d.grid_blocks = d.sm_count * per_sm;
    return UPOW_OK;
}

void destroy_device(Device& d) {
    if (d.id < 0) return;
    cudaSetDevice(d.id);
    if (d.stream) cudaStreamDestroy(d.stream);
    if (d.ctrl) cudaStreamDestroy(d.ctrl);
    if (d.ev0) cudaEventDestroy(d.ev0);
    if (d.ev1) cudaEventDestroy(d.ev1);
    cudaFree(d.d_res); cudaFree(d.d_out); cudaFree(d.d_digest); cudaFree(d.d_flush);
    cudaFree(d.d_jobs); cudaFree(d.d_active); cudaFree(d.d_best); cudaFree(d.d_bdigest);
    cudaFree(d.d_vrecs); cudaFree(d.d_vvalid); cudaFree(d.d_vdig);
    if (d.h_res) cudaFreeHost(d.h_res);
    if (d.h_digest) cudaFreeHost(d.h_digest);
    if (d.h_one) cudaFreeHost(d.h_one);
}

// Waits for ev on the device while watching the cancel flag / deadline; raises the device
// abort word (through the control stream) when either fires.  Returns 0, UPOW_CANCELLED or
// UPOW_EXHAUSTED (deadline), or UPOW_E_CUDA.
int wait_event(upow_ctx* ctx, Device& d, cudaEvent_t ev, double deadline) {
    int status = 0;
    bool aborted = false;
    for (;;) {
        cudaError_t q = cudaEventQuery(ev);
        if (q == cudaSuccess) return status;
        if (q != cudaErrorNotReady) {
            set_err(ctx, std::string("cudaEventQuery: ") + cudaGetErrorString(q));
            return UPOW_E_CUDA;
        }
        if (!aborted) {
            if (ctx->cancel.load(std::memory_order_relaxed)) status = UPOW_CANCELLED;
            else if (deadline > 0 && now_s() > deadline) status = UPOW_EXHAUSTED;
            if (status) {
                cudaMemcpyAsync(&d.d_res->abort_flag, d.h_one, sizeof(unsigned int), cudaMemcpyHostToDevice, d.ctrl);
                aborted = true;
            }
        }
        std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
}

struct TileOut {
    double sm_mhz = 0;
    bool found = false;
    uint32_t nonce = 0;
    uint64_t hashes = 0;
    double kernel_ms = 0;
    uint32_t count = 0;  // ENUM
    int status = 0;      // 0, UPOW_CANCELLED, UPOW_EXHAUSTED(deadline)
};

// ---- kernel variants -------------------------------------------------------------------
// Variant 0 is the plain-CUDA kernel (compiler-chosen pipes); the others are the
// hand-specialised loops of hot_loop.h with different shares of work on the fma pipe.
typedef void (*SearchFn)(const SearchArgs);
struct Variant {
    const char* name;
    SearchFn fn[2][2];  // [layout-1][mode]
    int block;          // threads per block (the kernels are warp-granular, any multiple of 32 works)
};
#define UPOW_PLAIN_VARIANT(NAME)                                                                     \
    { NAME, { { search_kernel<LAYOUT_V1, MODE_MIN, 1>, search_kernel<LAYOUT_V1, MODE_ENUM, 1> },      \
              { search_kernel<LAYOUT_V2, MODE_MIN, 1>, search_kernel<LAYOUT_V2, MODE_ENUM, 1> } }, 256 }
#define UPOW_TUNED_VARIANT(NAME, TN, MINB)                                                           \
    { NAME, { { search_kernel_tuned<LAYOUT_V1, MODE_MIN, TN, MINB>, search_kernel_tuned<LAYOUT_V1, MODE_ENUM, TN, MINB> }, \
              { search_kernel_tuned<LAYOUT_V2, MODE_MIN, TN, MINB>, search_kernel_tuned<LAYOUT_V2, MODE_ENUM, TN, MINB> } }, 256 }
#define UPOW_TUNED_VARIANT_B(NAME, TN, MINB, BLOCK)                                                  \
    { NAME, { { search_kernel_tuned<LAYOUT_V1, MODE_MIN, TN, MINB, BLOCK>, search_kernel_tuned<LAYOUT_V1, MODE_ENUM, TN, MINB, BLOCK> }, \
              { search_kernel_tuned<LAYOUT_V2, MODE_MIN, TN, MINB, BLOCK>, search_kernel_tuned<LAYOUT_V2, MODE_ENUM, TN, MINB, BLOCK> } }, BLOCK }
#define UPOW_TUNED_VARIANT_N(NAME, TN, MINB, BLOCK, NPT)                                             \
    { NAME, { { search_kernel_tuned<LAYOUT_V1, MODE_MIN, TN, MINB, BLOCK, NPT>, search_kernel_tuned<LAYOUT_V1, MODE_ENUM, TN, MINB, BLOCK, NPT> }, \
              { search_kernel_tuned<LAYOUT_V2, MODE_MIN, TN, MINB, BLOCK, NPT>, search_kernel_tuned<LAYOUT_V2, MODE_ENUM, TN, MINB, BLOCK, NPT> } }, BLOCK }
#define TN(...) Tune<__VA_ARGS__>
const Variant VARIANTS[] = {
    UPOW_PLAIN_VARIANT("plain: nvcc-chosen pipes"),
    UPOW_TUNED_VARIANT("add+shr on fma, 3 blocks/SM", TN(1, 1, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr on fma, 4 blocks/SM", TN(1, 1, 0, 0, 0, 0), 4),
    UPOW_TUNED_VARIANT("add+shr+rot25 on 1/4 rounds", TN(1, 1, 4, 1, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr+rot25 on 2/4 rounds", TN(1, 1, 4, 2, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr+rot25 on all rounds", TN(1, 1, 4, 4, 0, 0), 3),
    UPOW_TUNED_VARIANT("add only on fma", TN(1, 0, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr, 7-add rounds", TN(1, 1, 0, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add only, 7-add rounds", TN(1, 0, 0, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add, 7-add, shr via IMAD.HI", TN(1, 2, 0, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add, 7-add, shr via IMAD.HI on 2/4 words", TN(1, 2, 0, 0, 0, 0, 0, 2), 3),
    UPOW_TUNED_VARIANT("add, 7-add, shr via IMAD.WIDE on 2/4 words", TN(1, 1, 0, 0, 0, 0, 0, 2), 3),
    UPOW_TUNED_VARIANT("add only, 7-add rounds, 4 blocks/SM", TN(1, 0, 0, 0, 0, 0, 0), 4),
    UPOW_TUNED_VARIANT_B("add only, 7-add, 128-thread blocks x7 (28 warps/SM)", TN(1, 0, 0, 0, 0, 0, 0), 7, 128),
    UPOW_TUNED_VARIANT_N("add only, 7-add, 2 nonces per lane, 128-thread blocks x4", TN(1, 0, 0, 0, 0, 0, 0), 4, 128, 2),
};
constexpr int N_VARIANTS = sizeof(VARIANTS) / sizeof(VARIANTS[0]);
constexpr int DEFAULT_VARIANT = 8;

cudaError_t launch_search(const Device& d, const SearchArgs& a, int mode, int variant) {
    SearchFn fn = VARIANTS[variant].fn[a.tail.layout == LAYOUT_V1 ? 0 : 1][mode == MODE_MIN ? 0 : 1];
    int grid = d.variant_grid[variant];
    fn<<<grid, VARIANTS[variant].block, 0, d.stream>>>(a);
    return cudaGetLastError();
}
// variant 0 keeps the plain-CUDA strided kernel (independent implementation); every other variant
// uses the pipe-assigned build
template <int MODE>
cudaError_t launch_strided(const Device& d, const StridedArgs& a, bool tuned) {
    const int grid = tuned ? d.strided_tuned_grid : d.grid_blocks;
    if (a.tail.layout == LAYOUT_V1) {
        if (tuned) search_strided_kernel<LAYOUT_V1, MODE, true><<<grid, 256, 0, d.stream>>>(a);
        else search_strided_kernel<LAYOUT_V1, MODE, false><<<grid, 256, 0, d.stream>>>(a);
    } else {
        if (tuned) search_strided_kernel<LAYOUT_V2, MODE, true><<<grid, 256, 0, d.stream>>>(a);
        else search_strided_kernel<LAYOUT_V2, MODE, false><<<grid, 256, 0, d.stream>>>(a);
    }
    return cudaGetLastError();
}

// persistent grid of every variant: SM count x resident blocks per SM of its v1 ENUM kernel
int init_variant_grids(upow_ctx* ctx, Device& d) {
    static_assert(N_VARIANTS <= 16, "variant_grid too small");
    for (int v = 0; v < N_VARIANTS; v++) {
        int per_sm = 0;
        CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, VARIANTS[v].fn[0][1], VARIANTS[v].block, 0));
        if (per_sm < 1) per_sm = 1;
        d.variant_grid[v] = d.sm_count * per_sm;
    }
    int per_sm = 0;
    CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, batch_kernel<LAYOUT_V1>, 256, 0));
    d.batch_grid = d.sm_count * (per_sm < 1 ? 1 : per_sm);
    CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, search_strided_kernel<LAYOUT_V1, MODE_MIN, true>, 256, 0));
    d.strided_tuned_grid = d.sm_count * (per_sm < 1 ? 1 : per_sm);
    return UPOW_OK;
}

// One launch over the contiguous nonce range [lo, hi) of one (template, timestamp).
// MODE_ENUM copies up to out_cap nonces into out_host (unsorted).
int run_contiguous(upow_ctx* ctx, Device& d, const Tail& tail, const Rule& rule, uint64_t lo, uint64_t hi, int mode,
                   double deadline, TileOut* out, std::vector<uint32_t>* out_host) {
    CU_TRY(ctx, cudaSetDevice(d.id));
    SearchArgs a;
    memset(&a, 0, sizeof a);
    a.tail = tail;
    a.rule = rule;
    a.nonce_lo = lo;
    a.nonce_hi = hi;
    a.h_lo = (uint32_t)(lo >> 16);
    const uint32_t h_hi = (uint32_t)((hi + 0xffff) >> 16);
    a.h_count = h_hi - a.h_lo;
    // items of up to 64 H values; small ranges get smaller items so every warp has work
    uint32_t hc = 64;
    const uint32_t n_warps = d.variant_grid[ctx->variant] * (VARIANTS[ctx->variant].block / 32);
    while (hc > 1 && (uint64_t)((a.h_count + hc - 1) / hc) * 2048 < 4ull * n_warps) hc >>= 1;
    a.hc = hc;
    a.n_items = ((a.h_count + hc - 1) / hc) * 2048;
    a.res = d.d_res;
    a.out = d.d_out;
    a.cap = ENUM_CAP;
    fill_opaque(&a.q);
    if (tail.layout == LAYOUT_V2) prepare_v2_consts(tail, &a.v2);
    DevResult init;
    memset(&init, 0, sizeof init);
    init.best = ~0ull;
    *d.h_res = init;
    CU_TRY(ctx, cudaMemcpyAsync(d.d_res, d.h_res, sizeof(DevResult), cudaMemcpyHostToDevice, d.stream));
    CU_TRY(ctx, cudaEventRecord(d.ev0, d.stream));
    CU_TRY(ctx, launch_search(d, a, mode, ctx->variant));
    CU_TRY(ctx, cudaEventRecord(d.ev1, d.stream));
    int st = wait_event(ctx, d, d.ev1, deadline);
    if (st < 0) return st;
    CU_TRY(ctx, cudaMemcpyAsync(d.h_res, d.d_res, sizeof(DevResult), cudaMemcpyDeviceToHost, d.stream));
    CU_TRY(ctx, cudaStreamSynchronize(d.stream));
    float ms = 0;
    CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    out->status = st;
    out->kernel_ms = ms;
    out->hashes = d.h_res->items_done * 32ull * hc;
    out->found = d.h_res->best != ~0ull;
    out->nonce = (uint32_t)d.h_res->best;
    out->count = d.h_res->count;
    out->sm_mhz = d.h_res->clk_ns ? (double)d.h_res->clk_cycles / (double)d.h_res->clk_ns * 1000.0 : 0.0;
    if (mode == MODE_ENUM && out_host) {
        const uint64_t n = std::min<uint64_t>(out->count, ENUM_CAP);
        out_host->resize(n);
        if (n) CU_TRY(ctx, cudaMemcpy(out_host->data(), d.d_out, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    }
    return UPOW_OK;
}

// One launch over steps [s_lo, s_hi) of the progression start + i*stride.
int run_strided(upow_ctx* ctx, Device& d, const Tail& tail, const Rule& rule, uint64_t start, uint64_t stride,
                uint64_t s_lo, uint64_t s_hi, double deadline, TileOut* out) {
    CU_TRY(ctx, cudaSetDevice(d.id));
    StridedArgs a;
    memset(&a, 0, sizeof a);
    a.tail = tail;
    a.rule = rule;
    a.start = start + s_lo * stride;
    a.stride = stride;
    a.n_steps = s_hi - s_lo;
    uint32_t ic = 64;
    const bool tuned = ctx->variant != 0;
    const uint32_t n_warps = (tuned ? d.strided_tuned_grid : d.grid_blocks) * 8;
    while (ic > 1 && (a.n_steps + 32ull * ic - 1) / (32ull * ic) < 4ull * n_warps) ic >>= 1;
    a.ic = ic;
    a.n_items = (uint32_t)((a.n_steps + 32ull * ic - 1) / (32ull * ic));
    a.res = d.d_res;
    a.out = d.d_out;
    a.cap = ENUM_CAP;
    DevResult init;
    memset(&init, 0, sizeof init);
    init.best = ~0ull;
    *d.h_res = init;
    CU_TRY(ctx, cudaMemcpyAsync(d.d_res, d.h_res, sizeof(DevResult), cudaMemcpyHostToDevice, d.stream));
    CU_TRY(ctx, cudaEventRecord(d.ev0, d.stream));
    fill_opaque(&a.q);
    if (tail.layout == LAYOUT_V2) prepare_v2_consts(tail, &a.v2);
    CU_TRY(ctx, launch_strided<MODE_MIN>(d, a, tuned));
    CU_TRY(ctx, cudaEventRecord(d.ev1, d.stream));
    int st = wait_event(ctx, d, d.ev1, deadline);
    if (st < 0) return st;
    CU_TRY(ctx, cudaMemcpyAsync(d.h_res, d.d_res, sizeof(DevResult), cudaMemcpyDeviceToHost, d.stream));
    CU_TRY(ctx, cudaStreamSynchronize(d.stream));
    float ms = 0;
    CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    out->status = st;
    out->kernel_ms = ms;
    out->hashes = d.h_res->items_done * 32ull * ic;
    out->found = d.h_res->best != ~0ull;
    out->nonce = (uint32_t)d.h_res->best;
    return UPOW_OK;
}

int device_digest(upow_ctx* ctx, Device& d, const Tail& tail, uint32_t nonce, uint8_t out[32]) {
    CU_TRY(ctx, cudaSetDevice(d.id));
    digest_kernel<<<1, 32, 0, d.stream>>>(tail, nonce, d.d_digest);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaMemcpyAsync(d.h_digest, d.d_digest, 32, cudaMemcpyDeviceToHost, d.stream));
    CU_TRY(ctx, cudaStreamSynchronize(d.stream));
    for (int i = 0; i < 8; i++) store_be32(out + 4 * i, d.h_digest[i]);
    return UPOW_OK;
}

struct Prepared {
    Tail tail;
    Rule rule;
    uint8_t prefix[136];
    int prefix_len;
};

bool prepare_slot(const Slot& s, uint32_t ts_add, Prepared* p) {
    memcpy(p->prefix, s.prefix, sizeof p->prefix);
    p->prefix_len = s.prefix_len;
    if (ts_add) set_prefix_timestamp(p->prefix, s.prefix_len, get_prefix_timestamp(s.prefix, s.prefix_len) + ts_add);
    if (!prepare_tail(p->prefix, s.prefix_len, &p->tail)) return false;
    return prepare_rule(s.prev_hash, s.n_int, s.charset_count, p->tail.mid, &p->rule);
}

// Host-side cross-check of a device result: full SHA-256 of prefix || nonce and the rule.
bool host_verify(const Prepared& p, uint32_t nonce, const uint8_t dev_digest[32]) {
    uint8_t msg[140], dg[32];
    memcpy(msg, p.prefix, p.prefix_len);
    msg[p.prefix_len] = nonce; msg[p.prefix_len + 1] = nonce >> 8;
    msg[p.prefix_len + 2] = nonce >> 16; msg[p.prefix_len + 3] = nonce >> 24;
    host_sha256(msg, p.prefix_len + 4, dg);
    if (memcmp(dg, dev_digest, 32) != 0) return false;
    uint32_t w[8];
    for (int i = 0; i < 8; i++) w[i] = load_be32(dg + 4 * i);
    return rule_accepts(p.rule, w);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C-ABI
// ------------------------------------------------------------------------------------------
extern "C" const char* upow_version(void) { return "upow_b200 0.1 (sm_100a)"; }

extern "C" const char* upow_last_error(const upow_ctx* ctx) {
    // a per-thread copy: the string inside the context may be rewritten by another thread at any time
    thread_local std::string copy;
    if (ctx) {
        std::lock_guard<std::mutex> g(const_cast<upow_ctx*>(ctx)->err_mu);
        copy = ctx->err;
    } else {
        std::lock_guard<std::mutex> g(g_init_mu);
        copy = g_init_error;
    }
    return copy.c_str();
}

extern "C" int upow_init(int n_devices, const int* device_ids, upow_ctx** out_ctx) {
    if (!out_ctx || n_devices < 0 || n_devices > UPOW_MAX_DEVICES) return UPOW_E_ARG;
    *out_ctx = nullptr;
    int visible = 0;
    cudaError_t e = cudaGetDeviceCount(&visible);
    if (e != cudaSuccess || visible < 1) {
        set_err(nullptr, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                             " (upow_b200 has no CPU fallback)");
        return UPOW_E_CUDA;
    }
    if (n_devices == 0) n_devices = std::min(visible, UPOW_MAX_DEVICES);
    upow_ctx* ctx = new (std::nothrow) upow_ctx;
    if (!ctx) return UPOW_E_INTERNAL;
    ctx->devs.resize(n_devices);
    for (int i = 0; i < n_devices; i++) {
        int id = device_ids ? device_ids[i] : i;
        if (id < 0 || id >= visible) {
            set_err(nullptr, "device id out of range");
            upow_destroy(ctx);
            return UPOW_E_ARG;
        }
        int rc = init_device(ctx, ctx->devs[i], id);
        if (rc == UPOW_OK) rc = init_variant_grids(ctx, ctx->devs[i]);
        if (rc != UPOW_OK) {
            set_err(nullptr, ctx->err);
            upow_destroy(ctx);
            return rc;
        }
    }
    ctx->variant = DEFAULT_VARIANT;
    if (const char* v = getenv("UPOW_VARIANT")) {
        int k = atoi(v);
        if (k >= 0 && k < N_VARIANTS) ctx->variant = k;
    }
    *out_ctx = ctx;
    return UPOW_OK;
}

extern "C" int upow_destroy(upow_ctx* ctx) {
    if (!ctx) return UPOW_E_ARG;
    for (auto& d : ctx->devs) destroy_device(d);
    delete ctx;
    return UPOW_OK;
}

extern "C" int upow_device_count(const upow_ctx* ctx) { return ctx ? (int)ctx->devs.size() : UPOW_E_ARG; }

extern "C" int upow_get_device_info(const upow_ctx* ctx, int index, upow_device_info* out) {
    if (!ctx || !out || index < 0 || index >= (int)ctx->devs.size()) return UPOW_E_ARG;
    const Device& d = ctx->devs[index];
    memset(out, 0, sizeof *out);
    memcpy(out->name, d.prop.name, sizeof out->name - 1);
    out->device_id = d.id;
    out->sm_count = d.sm_count;
    out->cc_major = d.prop.major;
    out->cc_minor = d.prop.minor;
    out->max_sm_clock_khz = d.prop.clockRate;
    out->search_grid_blocks = d.variant_grid[ctx->variant];
    out->total_mem = d.prop.totalGlobalMem;
    return UPOW_OK;
}

extern "C" int upow_set_template(upow_ctx* ctx, int slot, const uint8_t* prefix, int prefix_len,
                                 const uint8_t* prev_hash32, int n_int, int charset_count) {
    if (!ctx) return UPOW_E_ARG;
    if (slot < 0 || slot >= UPOW_MAX_SLOTS || !prefix || !prev_hash32 || (prefix_len != 134 && prefix_len != 104)) {
        set_err(ctx, "upow_set_template: bad slot, NULL pointer or prefix_len not 104/134");
        return UPOW_E_ARG;
    }
    Slot s;
    s.used = true;
    memset(s.prefix, 0, sizeof s.prefix);
    memcpy(s.prefix, prefix, prefix_len);
    s.prefix_len = prefix_len;
    memcpy(s.prev_hash, prev_hash32, 32);
    s.n_int = n_int;
    s.charset_count = charset_count;
    Prepared p;
    if (!prepare_slot(s, 0, &p)) {
        set_err(ctx, "upow_set_template: bad rule (n_int < 0, charset_count outside 1..16, or n_int > 63 with a charset limit)");
        return UPOW_E_ARG;
    }
    std::lock_guard<std::mutex> g(ctx->api_mu);
    ctx->slots[slot] = s;
    return UPOW_OK;
}

extern "C" int upow_variant_count(void) { return N_VARIANTS; }
extern "C" const char* upow_variant_name(int v) { return v >= 0 && v < N_VARIANTS ? VARIANTS[v].name : ""; }
extern "C" int upow_get_variant(const upow_ctx* ctx) { return ctx ? ctx->variant : UPOW_E_ARG; }
extern "C" int upow_set_variant(upow_ctx* ctx, int v) {
    if (!ctx || v < 0 || v >= N_VARIANTS) return UPOW_E_ARG;
    std::lock_guard<std::mutex> g(ctx->api_mu);
    ctx->variant = v;
    return UPOW_OK;
}

extern "C" int upow_cancel(upow_ctx* ctx) {
    if (!ctx) return UPOW_E_ARG;
    ctx->cancel.store(1, std::memory_order_relaxed);
    return UPOW_OK;
}

extern "C" int upow_search_min(upow_ctx* ctx, int slot, uint64_t start_nonce, uint64_t stride,
                               const upow_limits* limits, upow_result* result) {
    if (!ctx || !result) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        memset(result, 0, sizeof *result);
        if (slot < 0 || slot >= UPOW_MAX_SLOTS || !ctx->slots[slot].used || stride < 1) {
            set_err(ctx, "upow_search_min: unknown slot or stride < 1");
            return UPOW_E_ARG;
        }
        const Slot s = ctx->slots[slot];
        upow_limits lim;
        memset(&lim, 0, sizeof lim);
        if (limits) lim = *limits;
        const uint64_t max_nonce = lim.max_nonce ? lim.max_nonce : NONCE_SPACE;
        if (max_nonce > NONCE_SPACE) {
            set_err(ctx, "upow_search_min: max_nonce > 2^32");
            return UPOW_E_ARG;
        }
        const double t_begin = now_s();
        const double deadline = lim.time_limit_s > 0 ? t_begin + lim.time_limit_s : 0;
        ctx->cancel.store(0, std::memory_order_relaxed);
        const int n_dev = (int)ctx->devs.size();

        // ---- tile space: (rollover index, tile index inside the timestamp), lexicographic ----
        const bool contiguous = stride == 1;
        uint64_t units_per_ts;  // nonces (contiguous) or steps (strided) per timestamp
        if (start_nonce >= max_nonce) units_per_ts = 0;
        else units_per_ts = contiguous ? max_nonce - start_nonce : (max_nonce - start_nonce + stride - 1) / stride;
        uint64_t tile = lim.tile_nonces;
        if (tile == 0) tile = n_dev > 1 ? (1ull << 28) : NONCE_SPACE;
        tile = std::max<uint64_t>((tile + 0xffff) & ~0xffffull, 1ull << 16);
        // contiguous tiles are aligned to multiples of `tile` so that H ranges never overlap
        const uint64_t first_tile = contiguous ? start_nonce / tile : 0;
        const uint64_t tiles_per_ts =
            units_per_ts == 0 ? 0 : (contiguous ? (max_nonce + tile - 1) / tile - first_tile : (units_per_ts + tile - 1) / tile);
        const uint64_t n_ts = (uint64_t)lim.max_rollovers + 1;
        const uint64_t n_tiles = tiles_per_ts * n_ts;

        std::mutex mu;
        uint64_t next_tile = 0;
        uint64_t found_tile = ~0ull;
        uint32_t found_nonce = 0;
        int final_status = 0;  // first non-zero status/error seen
        std::vector<uint64_t> dev_hashes(n_dev, 0);
        std::vector<double> dev_ms(n_dev, 0.0);
        std::vector<uint64_t> dev_tile(n_dev, ~0ull);   // tile each device is working on (under mu)
        std::atomic<uint32_t> launches{0};

        auto worker = [&](int di) {
            Device& d = ctx->devs[di];
            uint64_t cached_ts = ~0ull;
            Prepared p;
            for (;;) {
                uint64_t t;
                {
                    std::lock_guard<std::mutex> lk(mu);
                    if (final_status != 0 || next_tile >= n_tiles || next_tile > found_tile) return;
                    if (ctx->cancel.load(std::memory_order_relaxed)) { final_status = UPOW_CANCELLED; return; }
                    if (deadline > 0 && now_s() > deadline) { final_status = UPOW_EXHAUSTED; return; }
                    t = next_tile++;
                    dev_tile[di] = t;
                }
                const uint64_t ts_idx = t / tiles_per_ts, k = t % tiles_per_ts;
                if (ts_idx != cached_ts) {
                    if (!prepare_slot(s, (uint32_t)ts_idx, &p)) {
                        std::lock_guard<std::mutex> lk(mu);
                        final_status = UPOW_E_INTERNAL;
                        return;
                    }
                    cached_ts = ts_idx;
                }
                TileOut o;
                int rc;
                if (contiguous) {
                    const uint64_t lo = std::max<uint64_t>(start_nonce, (first_tile + k) * tile);
                    const uint64_t hi = std::min<uint64_t>(max_nonce, (first_tile + k + 1) * tile);
                    rc = run_contiguous(ctx, d, p.tail, p.rule, lo, hi, MODE_MIN, deadline, &o, nullptr);
                } else {
                    const uint64_t s_lo = k * tile, s_hi = std::min<uint64_t>(units_per_ts, (k + 1) * tile);
                    rc = run_strided(ctx, d, p.tail, p.rule, start_nonce, stride, s_lo, s_hi, deadline, &o);
                }
                launches.fetch_add(1);
                std::lock_guard<std::mutex> lk(mu);
                if (rc != UPOW_OK) { final_status = rc; return; }
                dev_hashes[di] += o.hashes;
                dev_ms[di] += o.kernel_ms;
                if (di == 0 && o.sm_mhz > 0) result->sm_clock_mhz = o.sm_mhz;
                dev_tile[di] = ~0ull;
                if (t > found_tile) continue;   // a tile behind an earlier find: stopped early on purpose, result unused
                if (o.found && t < found_tile) {
                    found_tile = t;
                    found_nonce = o.nonce;
                    // tiles after this one can no longer matter: stop the GPUs that are still working on them
                    for (int oj = 0; oj < n_dev; oj++) {
                        if (oj == di || dev_tile[oj] == ~0ull || dev_tile[oj] <= t) continue;
                        Device& od = ctx->devs[oj];
                        if (cudaSetDevice(od.id) == cudaSuccess)
                            cudaMemcpyAsync(&od.d_res->abort_flag, od.h_one, sizeof(unsigned int), cudaMemcpyHostToDevice, od.ctrl);
                    }
                    cudaSetDevice(d.id);
                }
                // an aborted tile (cancel / deadline) may have skipped nonces: its "not found" is not final
                if (o.status != 0 && final_status == 0) final_status = o.status;
            }
        };
        if (n_tiles > 0) {
            if (n_dev == 1) {
                worker(0);
            } else {
                std::vector<std::thread> th;
                for (int i = 0; i < n_dev; i++) th.emplace_back(worker, i);
                for (auto& t : th) t.join();
            }
        }
        // abort words travel on the control streams: make sure none is still in flight when the next call
        // resets the device-side result block
        for (auto& d : ctx->devs)
            if (cudaSetDevice(d.id) == cudaSuccess) cudaStreamSynchronize(d.ctrl);
        result->launches = launches.load();
        for (int i = 0; i < n_dev; i++) {
            result->per_device_hashes[i] = dev_hashes[i];
            result->per_device_kernel_ms[i] = dev_ms[i];
            result->hashes_done += dev_hashes[i];
            result->kernel_ms = std::max(result->kernel_ms, dev_ms[i]);
        }
        int rc = UPOW_EXHAUSTED;
        if (final_status < 0) {
            rc = final_status;
        } else if (final_status != 0) {
            rc = final_status;  // cancelled or timed out: any partial find is NOT reported as the minimum
        } else if (found_tile != ~0ull) {
            const uint64_t ts_idx = found_tile / tiles_per_ts;
            Prepared p;
            prepare_slot(s, (uint32_t)ts_idx, &p);
            int drc = device_digest(ctx, ctx->devs[0], p.tail, found_nonce, result->digest);
            if (drc != UPOW_OK) return drc;
            if (!host_verify(p, found_nonce, result->digest)) {
                set_err(ctx, "device result failed the host SHA-256 / rule cross-check");
                return UPOW_E_INTERNAL;
            }
            result->found = 1;
            result->rollovers = (uint32_t)ts_idx;
            result->timestamp = get_prefix_timestamp(p.prefix, p.prefix_len);
            result->nonce = found_nonce;
            rc = UPOW_OK;
        }
        result->elapsed_s = now_s() - t_begin;
        return rc;
    } catch (const std::exception& e) {
        set_err(ctx, std::string("exception: ") + e.what());
        return UPOW_E_INTERNAL;
    } catch (...) {
        set_err(ctx, "unknown exception");
        return UPOW_E_INTERNAL;
    }
}

extern "C" int upow_sweep(upow_ctx* ctx, int slot, uint64_t nonce_lo, uint64_t nonce_hi, uint32_t* out_nonces,
                          uint64_t cap, uint64_t* out_count, upow_result* stats) {
    if (!ctx || !out_count || (cap && !out_nonces)) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        *out_count = 0;
        if (stats) memset(stats, 0, sizeof *stats);
        if (slot < 0 || slot >= UPOW_MAX_SLOTS || !ctx->slots[slot].used || nonce_hi > NONCE_SPACE || nonce_lo > nonce_hi) {
            set_err(ctx, "upow_sweep: unknown slot or bad range");
            return UPOW_E_ARG;
        }
        const double t_begin = now_s();
        ctx->cancel.store(0, std::memory_order_relaxed);
        Prepared p;
        if (!prepare_slot(ctx->slots[slot], 0, &p)) return UPOW_E_INTERNAL;
        const int n_dev = (int)ctx->devs.size();
        // contiguous shares, boundaries aligned to 2^16 so no H value is evaluated twice
        std::vector<uint64_t> cut(n_dev + 1);
        for (int i = 0; i <= n_dev; i++) {
            uint64_t c = nonce_lo + (nonce_hi - nonce_lo) * i / n_dev;
            if (i != 0 && i != n_dev) c = (c + 0xffff) & ~0xffffull;
            cut[i] = std::min(std::max(c, nonce_lo), nonce_hi);
        }
        std::vector<std::vector<uint32_t>> parts(n_dev);
        std::vector<TileOut> outs(n_dev);
        std::vector<int> rcs(n_dev, UPOW_OK);
        auto worker = [&](int di) {
            if (cut[di] >= cut[di + 1]) return;
            rcs[di] = run_contiguous(ctx, ctx->devs[di], p.tail, p.rule, cut[di], cut[di + 1], MODE_ENUM, 0, &outs[di],
                                     &parts[di]);
        };
        if (n_dev == 1) worker(0);
        else {
            std::vector<std::thread> th;
            for (int i = 0; i < n_dev; i++) th.emplace_back(worker, i);
            for (auto& t : th) t.join();
        }
        int rc = UPOW_OK;
        uint64_t total = 0;
        bool dev_overflow = false;
        std::vector<uint32_t> all;
        for (int i = 0; i < n_dev; i++) {
            if (rcs[i] != UPOW_OK) return rcs[i];
            if (outs[i].status == UPOW_CANCELLED) rc = UPOW_CANCELLED;
            total += outs[i].count;
            if (outs[i].count > ENUM_CAP) dev_overflow = true;
            all.insert(all.end(), parts[i].begin(), parts[i].end());
            if (stats) {
                stats->per_device_hashes[i] = outs[i].hashes;
                stats->per_device_kernel_ms[i] = outs[i].kernel_ms;
                stats->hashes_done += outs[i].hashes;
                stats->kernel_ms = std::max(stats->kernel_ms, (double)outs[i].kernel_ms);
                stats->launches++;
                if (i == 0) stats->sm_clock_mhz = outs[i].sm_mhz;
            }
        }
        std::sort(all.begin(), all.end());
        const uint64_t n_copy = std::min<uint64_t>(all.size(), cap);
        if (n_copy) memcpy(out_nonces, all.data(), n_copy * sizeof(uint32_t));
        *out_count = total;
        if (stats) stats->elapsed_s = now_s() - t_begin;
        if (rc == UPOW_OK && (dev_overflow || total > cap)) {
            set_err(ctx, "upow_sweep: more valid nonces than the output buffer holds");
            rc = UPOW_E_OVERFLOW;
        }
        return rc;
    } catch (const std::exception& e) {
        set_err(ctx, std::string("exception: ") + e.what());
        return UPOW_E_INTERNAL;
    } catch (...) {
        set_err(ctx, "unknown exception");
        return UPOW_E_INTERNAL;
    }
}

// Shared front half of upow_search_batch / upow_search_any: validate and prepare every template.
namespace {
int prepare_templates(upow_ctx* ctx, uint32_t n_templates, const upow_template* templates, std::vector<Prepared>* prep,
                      const char* who) {
    prep->resize(n_templates);
    std::atomic<int> rc{UPOW_OK};
    std::atomic<uint32_t> bad{0};
    auto range = [&](uint32_t lo, uint32_t hi) {
        for (uint32_t i = lo; i < hi && rc.load() == UPOW_OK; i++) {
            Slot s;
            s.used = true;
            if (templates[i].prefix_len != 134 && templates[i].prefix_len != 104) { bad = i; rc = UPOW_E_ARG; return; }
            memcpy(s.prefix, templates[i].prefix, sizeof s.prefix);
            s.prefix_len = (int)templates[i].prefix_len;
            memcpy(s.prev_hash, templates[i].prev_hash, 32);
            s.n_int = templates[i].n_int;
            s.charset_count = templates[i].charset_count;
            if (!prepare_slot(s, 0, &(*prep)[i])) { bad = i; rc = UPOW_E_ARG; return; }
        }
    };
    // midstates of thousands of templates: spread over a few host threads
    const uint32_t n_thr = n_templates >= 1024 ? std::min<uint32_t>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;
    if (n_thr == 1) range(0, n_templates);
    else {
        std::vector<std::thread> th;
        for (uint32_t t = 0; t < n_thr; t++)
            th.emplace_back(range, (uint32_t)((uint64_t)n_templates * t / n_thr), (uint32_t)((uint64_t)n_templates * (t + 1) / n_thr));
        for (auto& t : th) t.join();
    }
    if (rc.load() != UPOW_OK)
        set_err(ctx, std::string(who) + ": template " + std::to_string(bad.load()) + ": prefix_len not 104/134 or bad rule");
    return rc.load();
}
}  // namespace

namespace {

int ensure_batch_capacity(upow_ctx* ctx, Device& d, uint32_t n) {
    if (n <= d.batch_cap) return UPOW_OK;
    CU_TRY(ctx, cudaSetDevice(d.id));
    cudaFree(d.d_jobs); cudaFree(d.d_active); cudaFree(d.d_best); cudaFree(d.d_bdigest);
    d.d_jobs = nullptr; d.d_active = nullptr; d.d_best = nullptr; d.d_bdigest = nullptr;
    d.batch_cap = 0;
    CU_TRY(ctx, cudaMalloc(&d.d_jobs, sizeof(BatchJob) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_active, sizeof(uint32_t) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_best, sizeof(unsigned long long) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_bdigest, 32 * n));
    d.batch_cap = n;
    return UPOW_OK;
}

// expected hashes until a template's first valid nonce: 16^n_target * 16 / charset
double expected_hashes(const Rule& r) {
    double e = 1;
    for (uint32_t q = 0; q < std::min<uint32_t>(r.n_target, 16); q++) e *= 16;
    return e * 16.0 / r.charset_count;
}

// The batch engine behind upow_search_batch and upow_search_any.
//
// Every GPU holds the whole job table.  The search advances in ROUNDS: in round r every template that
// has no valid nonce yet is searched over the next h_round H values (h_round * 2^16 nonces), and the
// still-active templates are dealt out evenly over the GPUs again at the start of every round -- a
// static template-to-GPU assignment leaves the unlucky GPU with several per cent more work (the search
// lengths are geometric).  All launches of a round are issued from this thread and then awaited
// together.  stop_at_first: return after the first round in which anything was found.
// best[i] = minimum valid nonce of template i or ~0ull; digests (8 words per template) from device 0.
int run_batch_rounds(upow_ctx* ctx, const std::vector<Prepared>& prep, uint64_t max_nonce, bool stop_at_first,
                     std::vector<unsigned long long>* best_out, std::vector<uint32_t>* digests_out,
                     upow_result* stats, uint32_t* first_round_out) {
    const uint32_t n = (uint32_t)prep.size();
    const int n_dev = (int)ctx->devs.size();
    best_out->assign(n, ~0ull);
    digests_out->assign((size_t)n * 8, 0);
    if (first_round_out) *first_round_out = 0xffffffffu;
    if (n == 0) return UPOW_OK;
    std::vector<BatchJob> jobs(n);
    bool has_v1 = false, has_v2 = false;
    for (uint32_t i = 0; i < n; i++) {
        jobs[i].tail = prep[i].tail;
        jobs[i].rule = prep[i].rule;
        (jobs[i].tail.layout == LAYOUT_V1 ? has_v1 : has_v2) = true;
    }
    for (auto& d : ctx->devs) {
        int rc = ensure_batch_capacity(ctx, d, n);
        if (rc != UPOW_OK) return rc;
        CU_TRY(ctx, cudaSetDevice(d.id));
        CU_TRY(ctx, cudaMemcpyAsync(d.d_jobs, jobs.data(), sizeof(BatchJob) * n, cudaMemcpyHostToDevice, d.stream));
        CU_TRY(ctx, cudaMemsetAsync(d.d_best, 0xff, sizeof(unsigned long long) * n, d.stream));
    }
    std::vector<unsigned long long>& best = *best_out;
    std::vector<uint32_t> active(n);
    for (uint32_t i = 0; i < n; i++) active[i] = i;
    std::vector<std::vector<unsigned long long>> dev_best(n_dev, std::vector<unsigned long long>(n));
    std::vector<uint64_t> dev_hashes(n_dev, 0);
    std::vector<double> dev_ms(n_dev, 0.0);
    uint32_t launches = 0;
    const uint32_t h_total = (uint32_t)((max_nonce + 0xffff) >> 16);
    uint32_t h_done = 0, round_no = 0;
    uint64_t fixed_round = 0;
    if (stop_at_first) {
        // first success over all templates ~ the easiest rule's expectation spread over n templates
        double e_min = expected_hashes(jobs[0].rule);
        for (auto& j : jobs) e_min = std::min(e_min, expected_hashes(j.rule));
        fixed_round = (uint64_t)(e_min / 4 / 65536.0 / n);
    }
    while (!active.empty() && h_done < h_total) {
        if (ctx->cancel.load(std::memory_order_relaxed)) return UPOW_CANCELLED;
        // round size: about a quarter of the expected remaining search length, but never so small that a
        // GPU gets fewer than 16 items per warp
        uint64_t h_round = fixed_round;
        if (!stop_at_first) {
            double expect = 0;
            for (uint32_t idx : active) expect = std::max(expect, expected_hashes(jobs[idx].rule));
            h_round = (uint64_t)(expect / 4 / 65536.0);
        }
        const uint64_t per_dev = (active.size() + n_dev - 1) / n_dev;
        const uint64_t h_min = ((uint64_t)16 * ctx->devs[0].batch_grid * 8 + 2047) / 2048 / per_dev + 1;
        if (!stop_at_first || round_no == 0) h_round = std::max<uint64_t>(std::max<uint64_t>(h_round, h_min), 1);
        if (stop_at_first) fixed_round = h_round;   // every round of an "any" search has the same size
        h_round = std::min<uint64_t>(h_round, h_total - h_done);
        uint32_t hc = 64;
        while (hc > 1 && hc > h_round) hc >>= 1;
        // deal the active templates out: device di takes active[cut[di] .. cut[di+1])
        std::vector<size_t> cut(n_dev + 1);
        for (int di = 0; di <= n_dev; di++) cut[di] = active.size() * di / n_dev;
        for (int di = 0; di < n_dev; di++) {
            Device& d = ctx->devs[di];
            const size_t cnt = cut[di + 1] - cut[di];
            if (cnt == 0) continue;
            CU_TRY(ctx, cudaSetDevice(d.id));
            BatchArgs a;
            memset(&a, 0, sizeof a);
            a.jobs = d.d_jobs; a.active = d.d_active; a.best = d.d_best; a.res = d.d_res;
            a.nonce_hi = max_nonce;
            a.n_active = (uint32_t)cnt;
            a.h_lo = h_done; a.h_count = (uint32_t)h_round; a.hc = hc;
            a.items_per_job = (uint32_t)((h_round + hc - 1) / hc) * 2048;
            fill_opaque(&a.q);
            DevResult init;
            memset(&init, 0, sizeof init);
            init.best = ~0ull;
            *d.h_res = init;
            CU_TRY(ctx, cudaMemcpyAsync(d.d_res, d.h_res, sizeof(DevResult), cudaMemcpyHostToDevice, d.stream));
            CU_TRY(ctx, cudaMemcpyAsync(d.d_active, active.data() + cut[di], sizeof(uint32_t) * cnt, cudaMemcpyHostToDevice, d.stream));
            CU_TRY(ctx, cudaEventRecord(d.ev0, d.stream));
            if (has_v1) { batch_kernel<LAYOUT_V1><<<d.batch_grid, 256, 0, d.stream>>>(a); launches++; }
            if (has_v1 && has_v2)  // both launches walk the same item list: rewind the queue in between
                CU_TRY(ctx, cudaMemsetAsync(&d.d_res->next_item, 0, sizeof(unsigned long long), d.stream));
            if (has_v2) { batch_kernel<LAYOUT_V2><<<d.batch_grid, 256, 0, d.stream>>>(a); launches++; }
            CU_TRY(ctx, cudaGetLastError());
            CU_TRY(ctx, cudaEventRecord(d.ev1, d.stream));
            // (no device-to-host copy here: into pageable memory it would block this thread until the kernel
            // has finished and serialise the GPUs; results are fetched in the wait loop below)
        }
        bool cancelled = false;
        for (int di = 0; di < n_dev; di++) {
            Device& d = ctx->devs[di];
            if (cut[di + 1] == cut[di]) continue;
            CU_TRY(ctx, cudaSetDevice(d.id));
            int st = wait_event(ctx, d, d.ev1, 0);
            if (st < 0) return st;
            if (st == UPOW_CANCELLED) cancelled = true;
            CU_TRY(ctx, cudaMemcpyAsync(dev_best[di].data(), d.d_best, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost, d.stream));
            CU_TRY(ctx, cudaMemcpyAsync(d.h_res, d.d_res, sizeof(DevResult), cudaMemcpyDeviceToHost, d.stream));
            CU_TRY(ctx, cudaStreamSynchronize(d.stream));
            float ms = 0;
            CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev0, d.ev1));
            dev_ms[di] += ms;
            dev_hashes[di] += d.h_res->items_done * 32ull * hc;
            for (size_t k = cut[di]; k < cut[di + 1]; k++) best[active[k]] = dev_best[di][active[k]];
        }
        if (cancelled) return UPOW_CANCELLED;
        std::vector<uint32_t> still;
        for (uint32_t idx : active)
            if (best[idx] == ~0ull) still.push_back(idx);
        const bool found_any = still.size() != active.size();
        active.swap(still);
        h_done += (uint32_t)h_round;
        if (found_any && first_round_out && *first_round_out == 0xffffffffu) *first_round_out = round_no;
        round_no++;
        if (stop_at_first && found_any) break;
    }
    // digests of all winners on device 0
    Device& d0 = ctx->devs[0];
    CU_TRY(ctx, cudaSetDevice(d0.id));
    CU_TRY(ctx, cudaMemcpyAsync(d0.d_best, best.data(), sizeof(unsigned long long) * n, cudaMemcpyHostToDevice, d0.stream));
    batch_digest_kernel<<<(n + 127) / 128, 128, 0, d0.stream>>>(d0.d_jobs, d0.d_best, n, d0.d_bdigest);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaMemcpyAsync(digests_out->data(), d0.d_bdigest, 32 * (size_t)n, cudaMemcpyDeviceToHost, d0.stream));
    CU_TRY(ctx, cudaStreamSynchronize(d0.stream));
    if (stats) {
        for (int di = 0; di < n_dev; di++) {
            stats->per_device_hashes[di] = dev_hashes[di];
            stats->per_device_kernel_ms[di] = dev_ms[di];
            stats->hashes_done += dev_hashes[di];
            stats->kernel_ms = std::max(stats->kernel_ms, dev_ms[di]);
        }
        stats->launches = launches;
    }
    return UPOW_OK;
}
}  // namespace

extern "C" int upow_search_batch(upow_ctx* ctx, uint32_t n_templates, const upow_template* templates,
                                 uint64_t max_nonce, upow_batch_item* results, upow_result* stats) {
    if (!ctx || (n_templates && (!templates || !results))) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        if (stats) memset(stats, 0, sizeof *stats);
        if (max_nonce == 0) max_nonce = NONCE_SPACE;
        if (max_nonce > NONCE_SPACE) return UPOW_E_ARG;
        const double t_begin = now_s();
        ctx->cancel.store(0, std::memory_order_relaxed);
        std::vector<Prepared> prep;
        int prc = prepare_templates(ctx, n_templates, templates, &prep, "upow_search_batch");
        if (prc != UPOW_OK) return prc;
        std::vector<unsigned long long> best;
        std::vector<uint32_t> dig;
        int rc = run_batch_rounds(ctx, prep, max_nonce, false, &best, &dig, stats, nullptr);
        if (rc != UPOW_OK) return rc;
        // every result is cross-checked with the host SHA-256 (spread over a few host threads)
        std::atomic<int> vrc{UPOW_OK};
        auto fill = [&](uint32_t lo, uint32_t hi) {
            for (uint32_t i = lo; i < hi; i++) {
                memset(&results[i], 0, sizeof results[i]);
                if (best[i] == ~0ull) continue;
                results[i].found = 1;
                results[i].nonce = (uint32_t)best[i];
                for (int w = 0; w < 8; w++) store_be32(results[i].digest + 4 * w, dig[(size_t)i * 8 + w]);
                if (!host_verify(prep[i], results[i].nonce, results[i].digest)) {
                    set_err(ctx, "batch result " + std::to_string(i) + " failed the host cross-check");
                    vrc = UPOW_E_INTERNAL;
                }
            }
        };
        const uint32_t n_thr = n_templates >= 1024 ? std::min<uint32_t>(8, std::max(1u, std::thread::hardware_concurrency())) : 1;
        if (n_thr == 1) fill(0, n_templates);
        else {
            std::vector<std::thread> th;
            for (uint32_t t = 0; t < n_thr; t++)
                th.emplace_back(fill, (uint32_t)((uint64_t)n_templates * t / n_thr), (uint32_t)((uint64_t)n_templates * (t + 1) / n_thr));
            for (auto& t : th) t.join();
        }
        if (stats) stats->elapsed_s = now_s() - t_begin;
        return vrc.load();
    } catch (const std::exception& e) {
        set_err(ctx, std::string("exception: ") + e.what());
        return UPOW_E_INTERNAL;
    } catch (...) {
        set_err(ctx, "unknown exception");
        return UPOW_E_INTERNAL;
    }
}

extern "C" int upow_search_any(upow_ctx* ctx, uint32_t n_templates, const upow_template* templates, uint64_t max_nonce,
                               uint32_t* out_index, upow_batch_item* out_item, upow_result* stats) {
    if (!ctx || !out_index || !out_item || (n_templates && !templates)) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        if (stats) memset(stats, 0, sizeof *stats);
        memset(out_item, 0, sizeof *out_item);
        *out_index = 0;
        if (max_nonce == 0) max_nonce = NONCE_SPACE;
        if (max_nonce > NONCE_SPACE) return UPOW_E_ARG;
        if (n_templates == 0) return UPOW_EXHAUSTED;
        const double t_begin = now_s();
        ctx->cancel.store(0, std::memory_order_relaxed);
        std::vector<Prepared> prep;
        int prc = prepare_templates(ctx, n_templates, templates, &prep, "upow_search_any");
        if (prc != UPOW_OK) return prc;
        std::vector<unsigned long long> best;
        std::vector<uint32_t> dig;
        uint32_t first_round = 0xffffffffu;
        int rc = run_batch_rounds(ctx, prep, max_nonce, true, &best, &dig, stats, &first_round);
        if (stats) stats->elapsed_s = now_s() - t_begin;
        if (rc != UPOW_OK) return rc;
        // the search stopped after the first round with a find: every recorded minimum belongs to that round,
        // and the round boundaries do not depend on the number of GPUs -- take the lowest template index
        uint32_t win = 0xffffffffu;
        for (uint32_t i = 0; i < n_templates; i++)
            if (best[i] != ~0ull) { win = i; break; }
        if (win == 0xffffffffu) return UPOW_EXHAUSTED;
        out_item->found = 1;
        out_item->nonce = (uint32_t)best[win];
        for (int w = 0; w < 8; w++) store_be32(out_item->digest + 4 * w, dig[(size_t)win * 8 + w]);
        if (!host_verify(prep[win], out_item->nonce, out_item->digest)) {
            set_err(ctx, "upow_search_any: result failed the host cross-check");
            return UPOW_E_INTERNAL;
        }
        *out_index = win;
        return UPOW_OK;
    } catch (const std::exception& e) {
        set_err(ctx, std::string("exception: ") + e.what());
        return UPOW_E_INTERNAL;
    } catch (...) {
        set_err(ctx, "unknown exception");
        return UPOW_E_INTERNAL;
    }
}

namespace {
static_assert(sizeof(upow_block) == sizeof(VerifyRecord), "upow_block and VerifyRecord must have one layout");

int verify_on_device(upow_ctx* ctx, Device& d, const upow_block* blocks, uint32_t n, uint8_t* out_valid,
                     uint8_t* out_digests, double* kernel_ms) {
    if (n == 0) return UPOW_OK;
    CU_TRY(ctx, cudaSetDevice(d.id));
    if (n > d.verify_cap) {
        cudaFree(d.d_vrecs); cudaFree(d.d_vvalid); cudaFree(d.d_vdig);
        d.d_vrecs = nullptr; d.d_vvalid = nullptr; d.d_vdig = nullptr; d.verify_cap = 0;
        CU_TRY(ctx, cudaMalloc(&d.d_vrecs, sizeof(VerifyRecord) * (size_t)n));
        CU_TRY(ctx, cudaMalloc(&d.d_vvalid, n));
        CU_TRY(ctx, cudaMalloc(&d.d_vdig, 32 * (size_t)n));
        d.verify_cap = n;
    }
    CU_TRY(ctx, cudaMemcpyAsync(d.d_vrecs, blocks, sizeof(VerifyRecord) * (size_t)n, cudaMemcpyHostToDevice, d.stream));
    CU_TRY(ctx, cudaEventRecord(d.ev0, d.stream));
    verify_kernel<<<(n + VERIFY_THREADS - 1) / VERIFY_THREADS, VERIFY_THREADS, 0, d.stream>>>(
        d.d_vrecs, n, d.d_vvalid, out_digests ? d.d_vdig : nullptr);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaEventRecord(d.ev1, d.stream));
    CU_TRY(ctx, cudaMemcpyAsync(out_valid, d.d_vvalid, n, cudaMemcpyDeviceToHost, d.stream));
    std::vector<uint32_t> words;
    if (out_digests) {
        words.resize((size_t)n * 8);
        CU_TRY(ctx, cudaMemcpyAsync(words.data(), d.d_vdig, 32 * (size_t)n, cudaMemcpyDeviceToHost, d.stream));
    }
    CU_TRY(ctx, cudaStreamSynchronize(d.stream));
    if (out_digests)
        for (size_t i = 0; i < words.size(); i++) store_be32(out_digests + 4 * i, words[i]);
    float ms = 0;
    CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev0, d.ev1));
    *kernel_ms = ms;
    return UPOW_OK;
}
}  // namespace

extern "C" int upow_verify_blocks(upow_ctx* ctx, uint32_t n_blocks, const upow_block* blocks, uint8_t* out_valid,
                                  uint8_t* out_digests, upow_result* stats) {
    if (!ctx || (n_blocks && (!blocks || !out_valid))) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        if (stats) memset(stats, 0, sizeof *stats);
        const double t_begin = now_s();
        const int n_dev = (int)ctx->devs.size();
        std::vector<int> rcs(n_dev, UPOW_OK);
        std::vector<double> ms(n_dev, 0.0);
        std::vector<uint32_t> cut(n_dev + 1);
        for (int i = 0; i <= n_dev; i++) cut[i] = (uint32_t)((uint64_t)n_blocks * i / n_dev);
        auto worker = [&](int di) {
            const uint32_t lo = cut[di], cnt = cut[di + 1] - cut[di];
            rcs[di] = verify_on_device(ctx, ctx->devs[di], blocks + lo, cnt, out_valid + lo,
                                       out_digests ? out_digests + 32 * (size_t)lo : nullptr, &ms[di]);
        };
        if (n_dev == 1) worker(0);
        else {
            std::vector<std::thread> th;
            for (int i = 0; i < n_dev; i++) th.emplace_back(worker, i);
            for (auto& t : th) t.join();
        }
        for (int di = 0; di < n_dev; di++) {
            if (rcs[di] != UPOW_OK) return rcs[di];
            if (stats) {
                stats->per_device_hashes[di] = cut[di + 1] - cut[di];
                stats->per_device_kernel_ms[di] = ms[di];
                stats->hashes_done += cut[di + 1] - cut[di];
                stats->kernel_ms = std::max(stats->kernel_ms, ms[di]);
                if (cut[di + 1] > cut[di]) stats->launches++;
            }
        }
        if (stats) stats->elapsed_s = now_s() - t_begin;
        return UPOW_OK;
    } catch (const std::exception& e) {
        set_err(ctx, std::string("exception: ") + e.what());
        return UPOW_E_INTERNAL;
    } catch (...) {
        set_err(ctx, "unknown exception");
        return UPOW_E_INTERNAL;
    }
}


namespace {
// messages [m_lo, m_hi) on one device; device buffers are per call (sizes vary widely)
int sha_batch_on_device(upow_ctx* ctx, Device& d, const uint8_t* data, const uint64_t* offsets, uint32_t m_lo,
                        uint32_t m_hi, uint8_t* out, double* kernel_ms) {
    const uint32_t n = m_hi - m_lo;
    if (n == 0) return UPOW_OK;
    CU_TRY(ctx, cudaSetDevice(d.id));
    const uint64_t b_lo = offsets[m_lo], b_hi = offsets[m_hi];
    const uint64_t bytes = b_hi - b_lo;
    uint8_t* d_data = nullptr;
    unsigned long long* d_off = nullptr;
    uint32_t* d_out = nullptr;
    int rc = UPOW_E_CUDA;
    std::vector<unsigned long long> rel(n + 1);
    for (uint32_t i = 0; i <= n; i++) rel[i] = offsets[m_lo + i] - b_lo;
    std::vector<uint32_t> words((size_t)n * 8);
    do {
        cudaError_t e;
        if ((e = cudaMalloc(&d_data, bytes + 16)) != cudaSuccess) { set_err(ctx, cudaGetErrorString(e)); break; }
        if ((e = cudaMalloc(&d_off, sizeof(unsigned long long) * (n + 1))) != cudaSuccess) { set_err(ctx, cudaGetErrorString(e)); break; }
        if ((e = cudaMalloc(&d_out, 32 * (size_t)n)) != cudaSuccess) { set_err(ctx, cudaGetErrorString(e)); break; }
        if (bytes) cudaMemcpyAsync(d_data, data + b_lo, bytes, cudaMemcpyHostToDevice, d.stream);
        cudaMemsetAsync(d_data + bytes, 0, 16, d.stream);
        cudaMemcpyAsync(d_off, rel.data(), sizeof(unsigned long long) * (n + 1), cudaMemcpyHostToDevice, d.stream);
        cudaEventRecord(d.ev0, d.stream);
        sha256_batch_kernel<<<(n + 127) / 128, 128, 0, d.stream>>>(d_data, d_off, n, d_out);
        if ((e = cudaGetLastError()) != cudaSuccess) { set_err(ctx, cudaGetErrorString(e)); break; }
        cudaEventRecord(d.ev1, d.stream);
        cudaMemcpyAsync(words.data(), d_out, 32 * (size_t)n, cudaMemcpyDeviceToHost, d.stream);
        if ((e = cudaStreamSynchronize(d.stream)) != cudaSuccess) { set_err(ctx, cudaGetErrorString(e)); break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, d.ev0, d.ev1);
        *kernel_ms = ms;
        for (size_t i = 0; i < words.size(); i++) store_be32(out + 32 * (size_t)m_lo + 4 * i, words[i]);
        rc = UPOW_OK;
    } while (0);
    cudaFree(d_data); cudaFree(d_off); cudaFree(d_out);
    return rc;
}
}  // namespace

extern "C" int upow_sha256_batch(upow_ctx* ctx, uint32_t n, const uint8_t* data, const uint64_t* offsets, uint8_t* out,
                                 upow_result* stats) {
    if (!ctx || (n && (!offsets || !out))) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        if (stats) memset(stats, 0, sizeof *stats);
        if (n == 0) return UPOW_OK;
        for (uint32_t i = 0; i < n; i++)
            if (offsets[i + 1] < offsets[i]) { set_err(ctx, "upow_sha256_batch: offsets must be non-decreasing"); return UPOW_E_ARG; }
        if (offsets[n] > offsets[0] && !data) return UPOW_E_ARG;
        const double t_begin = now_s();
        const int n_dev = (int)ctx->devs.size();
        std::vector<uint32_t> cut(n_dev + 1);
        for (int i = 0; i <= n_dev; i++) cut[i] = (uint32_t)((uint64_t)n * i / n_dev);
        std::vector<int> rcs(n_dev, UPOW_OK);
        std::vector<double> ms(n_dev, 0.0);
        auto worker = [&](int di) {
            rcs[di] = sha_batch_on_device(ctx, ctx->devs[di], data, offsets, cut[di], cut[di + 1], out, &ms[di]);
        };
        if (n_dev == 1) worker(0);
        else {
            std::vector<std::thread> th;
            for (int i = 0; i < n_dev; i++) th.emplace_back(worker, i);
            for (auto& t : th) t.join();
        }
        for (int di = 0; di < n_dev; di++) {
            if (rcs[di] != UPOW_OK) return rcs[di];
            if (stats) {
                stats->per_device_hashes[di] = cut[di + 1] - cut[di];
                stats->per_device_kernel_ms[di] = ms[di];
                stats->hashes_done += cut[di + 1] - cut[di];
                stats->kernel_ms = std::max(stats->kernel_ms, ms[di]);
                if (cut[di + 1] > cut[di]) stats->launches++;
            }
        }
        if (stats) stats->elapsed_s = now_s() - t_begin;
        return UPOW_OK;
    } catch (const std::exception& e) {
        set_err(ctx, std::string("exception: ") + e.what());
        return UPOW_E_INTERNAL;
    } catch (...) {
        set_err(ctx, "unknown exception");
        return UPOW_E_INTERNAL;
    }
}

extern "C" int upow_flush_l2(upow_ctx* ctx) {
    if (!ctx) return UPOW_E_ARG;
    for (auto& d : ctx->devs) {
        CU_TRY(ctx, cudaSetDevice(d.id));
        if (!d.d_flush) CU_TRY(ctx, cudaMalloc(&d.d_flush, L2_FLUSH_BYTES));
        CU_TRY(ctx, cudaMemsetAsync(d.d_flush, 0x5a, L2_FLUSH_BYTES, d.stream));
    }
    for (auto& d : ctx->devs) {
        CU_TRY(ctx, cudaSetDevice(d.id));
        CU_TRY(ctx, cudaStreamSynchronize(d.stream));
    }
    return UPOW_OK;
}

extern "C" int upow_selftest(int device) {
    if (cudaSetDevice(device) != cudaSuccess) return UPOW_E_CUDA;
    uint32_t k[64];
    for (int i = 0; i < 64; i++) k[i] = Kc(i);
    if (cudaMemcpyToSymbol(K_DEV, k, sizeof k) != cudaSuccess) return UPOW_E_CUDA;
    // FIPS 180-4 known answers: "", "abc", the 56-byte two-block message, and a 138-byte block-shaped message
    const char* msgs[4] = {"", "abc", "abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq", nullptr};
    uint8_t block_msg[138];
    for (int i = 0; i < 138; i++) block_msg[i] = (uint8_t)(i * 7 + 3);
    const uint8_t kat[3][32] = {
        {0xe3, 0xb0, 0xc4, 0x42, 0x98, 0xfc, 0x1c, 0x14, 0x9a, 0xfb, 0xf4, 0xc8, 0x99, 0x6f, 0xb9, 0x24,
         0x27, 0xae, 0x41, 0xe4, 0x64, 0x9b, 0x93, 0x4c, 0xa4, 0x95, 0x99, 0x1b, 0x78, 0x52, 0xb8, 0x55},
        {0xba, 0x78, 0x16, 0xbf, 0x8f, 0x01, 0xcf, 0xea, 0x41, 0x41, 0x40, 0xde, 0x5d, 0xae, 0x22, 0x23,
         0xb0, 0x03, 0x61, 0xa3, 0x96, 0x17, 0x7a, 0x9c, 0xb4, 0x10, 0xff, 0x61, 0xf2, 0x00, 0x15, 0xad},
        {0x24, 0x8d, 0x6a, 0x61, 0xd2, 0x06, 0x38, 0xb8, 0xe5, 0xc0, 0x26, 0x93, 0x0c, 0x3e, 0x60, 0x39,
         0xa3, 0x3c, 0xe4, 0x59, 0x64, 0xff, 0x21, 0x67, 0xf6, 0xec, 0xed, 0xd4, 0x19, 0xdb, 0x06, 0xc1}};
    std::vector<uint32_t> blocks, n_blocks, offsets;
    for (int m = 0; m < 4; m++) {
        const uint8_t* p = m < 3 ? (const uint8_t*)msgs[m] : block_msg;
        size_t len = m < 3 ? strlen(msgs[m]) : sizeof block_msg;
        std::vector<uint8_t> padded(p, p + len);
        padded.push_back(0x80);
        while (padded.size() % 64 != 56) padded.push_back(0);
        for (int i = 7; i >= 0; i--) padded.push_back((uint8_t)(((uint64_t)len * 8) >> (8 * i)));
        offsets.push_back((uint32_t)(blocks.size() / 16));
        n_blocks.push_back((uint32_t)(padded.size() / 64));
        for (size_t i = 0; i < padded.size(); i += 4) blocks.push_back(load_be32(&padded[i]));
    }
    uint32_t *d_blocks = 0, *d_nb = 0, *d_off = 0, *d_out = 0, *d_a = 0;
    int rc = UPOW_E_CUDA;
    do {
        if (cudaMalloc(&d_blocks, blocks.size() * 4) != cudaSuccess) break;
        if (cudaMalloc(&d_nb, 16) != cudaSuccess || cudaMalloc(&d_off, 16) != cudaSuccess) break;
        if (cudaMalloc(&d_out, 4 * 32) != cudaSuccess || cudaMalloc(&d_a, 4) != cudaSuccess) break;
        cudaMemcpy(d_blocks, blocks.data(), blocks.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d_nb, n_blocks.data(), 16, cudaMemcpyHostToDevice);
        cudaMemcpy(d_off, offsets.data(), 16, cudaMemcpyHostToDevice);
        selftest_kernel<<<1, 32>>>(d_blocks, d_nb, d_off, 4, d_out, d_a);
        if (cudaDeviceSynchronize() != cudaSuccess) break;
        uint32_t out[32], a64;
        if (cudaMemcpy(out, d_out, sizeof out, cudaMemcpyDeviceToHost) != cudaSuccess) break;
        if (cudaMemcpy(&a64, d_a, 4, cudaMemcpyDeviceToHost) != cudaSuccess) break;
        rc = UPOW_OK;
        for (int m = 0; m < 4; m++) {
            uint8_t dg[32], want[32];
            for (int i = 0; i < 8; i++) store_be32(dg + 4 * i, out[m * 8 + i]);
            if (m < 3) memcpy(want, kat[m], 32);
            else host_sha256(block_msg, sizeof block_msg, want);
            if (memcmp(dg, want, 32) != 0) rc = UPOW_E_INTERNAL;
        }
        if (a64 != load_be32(kat[0])) rc = UPOW_E_INTERNAL;  // unrolled hot-path rounds on the "" block
    } while (0);
    cudaFree(d_blocks); cudaFree(d_nb); cudaFree(d_off); cudaFree(d_out); cudaFree(d_a);
    return rc;
}

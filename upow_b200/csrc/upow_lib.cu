// Host side of libupow_b200.so: contexts, template slots, the launch pipeline that keeps the
// exact-minimum guarantee across GPUs, the batch engine, and the extern "C" entry points of
// include/upow_b200.h.
// No PyTorch, no NCCL: the nonce space shards into independent lgroup classes; the only
// cross-GPU signals are one 8-byte "limit" word per launch (written by the finder into the other GPUs'
// memory over NVLink peer access, or relayed by the host where peer access is missing) and a per-device
// kill word the host writes for cancel / deadline / "a find made the later launches pointless".
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <memory>
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
constexpr int DEPTH = 4;                         // result slots per GPU (a ring)
constexpr int INFLIGHT = 3;                      // launches queued per GPU: one running, one filling its tail, one behind
constexpr int KILL_STAGES = 16;
constexpr uint32_t LGROUPS = 2048;               // 65536 L values / 32 lanes

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

constexpr int NLANES = 4;
constexpr size_t LANE_BYTES = 8u << 20;
struct Lane {
    cudaStream_t st = nullptr;
    uint8_t* h[2] = {nullptr, nullptr};   // pinned
    uint8_t* d[2] = {nullptr, nullptr};   // device
    cudaEvent_t done[2] = {}, k0[2] = {}, k1[2] = {};
};

struct Device {
    int id = -1;
    int index = 0;         // position inside the context
    int sm_count = 0;
    int grid_blocks = 0;   // persistent grid of the plain strided kernel
    int variant_grid[16] = {0};
    int batch_grid = 0;
    int strided_tuned_grid = 0;
    cudaDeviceProp prop;
    cudaStream_t stream[2] = {nullptr, nullptr}, ctrl = nullptr;
    cudaEvent_t ev0[DEPTH] = {}, ev1[DEPTH] = {}, ev_done[DEPTH] = {}, ev_begin = nullptr, ev_aux[2] = {};
    DevResult* d_res = nullptr;          // [DEPTH]
    DevResult* h_res = nullptr;          // pinned mirror [DEPTH]
    DevResult* h_init = nullptr;         // pinned, constant: the state every launch starts from
    unsigned long long* d_kill = nullptr;
    unsigned long long* h_stage = nullptr;   // pinned [KILL_STAGES]
    std::mutex kill_mu;
    uint32_t stage_seq = 0;
    uint32_t serial = 0;                 // last launch serial handed out on this device
    uint32_t slot_seq = 0;
    bool peer_ok[UPOW_MAX_DEVICES] = {}; // may kernels on this device write into device j's memory?
    uint32_t* d_out = nullptr;
    void* d_flush = nullptr;
    // batch
    BatchJob* d_jobs = nullptr;
    uint32_t* d_active[2] = {nullptr, nullptr};
    unsigned long long* d_limit = nullptr;
    unsigned long long* d_limit_in = nullptr;
    uint32_t* d_bdigest = nullptr;
    BatchCtl* d_ctl = nullptr;
    BatchCtl* h_ctl = nullptr;           // pinned [2]
    uint32_t batch_cap = 0;
    // block verifier / batched hashing: staging lanes (allocated on first use, fixed size)
    Lane lanes[NLANES];
    bool lanes_ready = false;
    Device() = default;
    Device(const Device&) = delete;
    Device& operator=(const Device&) = delete;
};

struct Watchdog {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    bool quit = false, armed = false;
    double deadline = 0;
    std::function<void(int)> fire;   // called once, with UPOW_CANCELLED or UPOW_EXHAUSTED (deadline)
};

}  // namespace

struct upow_ctx {
    std::vector<std::unique_ptr<Device>> devs;
    Slot slots[UPOW_MAX_SLOTS];
    std::atomic<int> cancel{0};   // sticky: stays set until a call consumes it or upow_clear_cancel
    std::atomic<unsigned long long> progress{0};   // hashes of the running (or last) search, updated as launches retire
    std::mutex err_mu;
    std::string err;
    std::mutex api_mu;  // one search at a time per context
    int variant = -1;   // index into VARIANTS
    bool full_p2p = true;
    Watchdog wd;
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

// ---- watchdog: one thread per context; while a call is armed it looks at the cancel flag and the
// deadline every 200 us and fires the call's stop action once.  The calling thread itself sleeps in
// cudaEventSynchronize, so a finished launch is seen at once instead of after a poll interval. ----
void watchdog_main(upow_ctx* ctx) {
    Watchdog& w = ctx->wd;
    std::unique_lock<std::mutex> lk(w.mu);
    for (;;) {
        w.cv.wait(lk, [&] { return w.quit || w.armed; });
        if (w.quit) return;
        while (w.armed && !w.quit) {
            int status = 0;
            if (ctx->cancel.load(std::memory_order_relaxed)) status = UPOW_CANCELLED;
            else if (w.deadline > 0 && now_s() > w.deadline) status = UPOW_EXHAUSTED;
            if (status) {
                if (w.fire) w.fire(status);   // under w.mu: disarm() cannot return while this runs
                w.armed = false;
                break;
            }
            w.cv.wait_for(lk, std::chrono::microseconds(200));
        }
    }
}
void watchdog_arm(upow_ctx* ctx, double deadline, std::function<void(int)> fire) {
    Watchdog& w = ctx->wd;
    {
        std::lock_guard<std::mutex> g(w.mu);
        w.deadline = deadline;
        w.fire = std::move(fire);
        w.armed = true;
    }
    w.cv.notify_all();
}
void watchdog_disarm(upow_ctx* ctx) {
    Watchdog& w = ctx->wd;
    std::lock_guard<std::mutex> g(w.mu);
    w.armed = false;
    w.fire = nullptr;
}
struct WatchdogScope {
    upow_ctx* ctx;
    WatchdogScope(upow_ctx* c, double deadline, std::function<void(int)> fire) : ctx(c) {
        watchdog_arm(c, deadline, std::move(fire));
    }
    ~WatchdogScope() { watchdog_disarm(ctx); }
};

// Launches with serial in [lo, hi] on this device leave at their next item.
void write_kill(Device& d, uint32_t lo, uint32_t hi) {
    std::lock_guard<std::mutex> g(d.kill_mu);
    if (cudaSetDevice(d.id) != cudaSuccess) return;
    unsigned long long* stage = &d.h_stage[d.stage_seq++ % KILL_STAGES];
    *stage = kill_range(lo, hi);
    cudaMemcpyAsync(d.d_kill, stage, sizeof *stage, cudaMemcpyHostToDevice, d.ctrl);
}

int init_device(upow_ctx* ctx, Device& d, int id, int index) {
    d.id = id;
    d.index = index;
    CU_TRY(ctx, cudaSetDevice(id));
    CU_TRY(ctx, cudaGetDeviceProperties(&d.prop, id));
    d.sm_count = d.prop.multiProcessorCount;
    for (int k = 0; k < 2; k++) CU_TRY(ctx, cudaStreamCreateWithFlags(&d.stream[k], cudaStreamNonBlocking));
    CU_TRY(ctx, cudaStreamCreateWithFlags(&d.ctrl, cudaStreamNonBlocking));
    for (int k = 0; k < DEPTH; k++) {
        CU_TRY(ctx, cudaEventCreate(&d.ev0[k]));
        CU_TRY(ctx, cudaEventCreate(&d.ev1[k]));
        CU_TRY(ctx, cudaEventCreateWithFlags(&d.ev_done[k], cudaEventDisableTiming));
    }
    CU_TRY(ctx, cudaEventCreate(&d.ev_begin));
    for (int k = 0; k < 2; k++) CU_TRY(ctx, cudaEventCreate(&d.ev_aux[k]));
    CU_TRY(ctx, cudaMalloc(&d.d_res, sizeof(DevResult) * DEPTH));
    CU_TRY(ctx, cudaMalloc(&d.d_kill, sizeof(unsigned long long)));
    CU_TRY(ctx, cudaMalloc(&d.d_out, sizeof(uint32_t) * ENUM_CAP));
    CU_TRY(ctx, cudaMallocHost(&d.h_res, sizeof(DevResult) * DEPTH));
    CU_TRY(ctx, cudaMallocHost(&d.h_init, sizeof(DevResult)));
    CU_TRY(ctx, cudaMallocHost(&d.h_stage, sizeof(unsigned long long) * KILL_STAGES));
    CU_TRY(ctx, cudaMallocHost(&d.h_ctl, sizeof(BatchCtl) * 2));
    memset(d.h_init, 0, sizeof(DevResult));
    d.h_init->best = NO_NONCE;
    d.h_init->limit = NO_NONCE;
    const unsigned long long none = KILL_NONE;
    CU_TRY(ctx, cudaMemcpy(d.d_kill, &none, sizeof none, cudaMemcpyHostToDevice));
    int per_sm = 0;
    CU_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, search_strided_kernel<LAYOUT_V1, MODE_MIN>, 256, 0));
    if (per_sm < 1) per_sm = 1;
    d.grid_blocks = d.sm_count * per_sm;
    return UPOW_OK;
}

void destroy_device(Device& d) {
    if (d.id < 0) return;
    cudaSetDevice(d.id);
    cudaDeviceSynchronize();
    for (int k = 0; k < 2; k++) if (d.stream[k]) cudaStreamDestroy(d.stream[k]);
    if (d.ctrl) cudaStreamDestroy(d.ctrl);
    for (int k = 0; k < DEPTH; k++) {
        if (d.ev0[k]) cudaEventDestroy(d.ev0[k]);
        if (d.ev1[k]) cudaEventDestroy(d.ev1[k]);
        if (d.ev_done[k]) cudaEventDestroy(d.ev_done[k]);
    }
    if (d.ev_begin) cudaEventDestroy(d.ev_begin);
    for (int k = 0; k < 2; k++) if (d.ev_aux[k]) cudaEventDestroy(d.ev_aux[k]);
    cudaFree(d.d_res); cudaFree(d.d_kill); cudaFree(d.d_out); cudaFree(d.d_flush);
    cudaFree(d.d_jobs); cudaFree(d.d_active[0]); cudaFree(d.d_active[1]); cudaFree(d.d_limit); cudaFree(d.d_limit_in);
    cudaFree(d.d_bdigest); cudaFree(d.d_ctl);
    for (Lane& ln : d.lanes) {
        if (ln.st) cudaStreamDestroy(ln.st);
        for (int b = 0; b < 2; b++) {
            cudaFree(ln.d[b]);
            if (ln.h[b]) cudaFreeHost(ln.h[b]);
            if (ln.done[b]) cudaEventDestroy(ln.done[b]);
            if (ln.k0[b]) cudaEventDestroy(ln.k0[b]);
            if (ln.k1[b]) cudaEventDestroy(ln.k1[b]);
        }
    }
    if (d.h_res) cudaFreeHost(d.h_res);
    if (d.h_init) cudaFreeHost(d.h_init);
    if (d.h_stage) cudaFreeHost(d.h_stage);
    if (d.h_ctl) cudaFreeHost(d.h_ctl);
}

// ---- kernel variants -------------------------------------------------------------------
// Variant 0 is the plain-CUDA kernel (compiler-chosen pipes); the others are the hand-specialised loops of
// hot_loop.h with different shares of work on the fma pipe.  The shipped library carries the plain kernel
// and the fastest build; -DUPOW_ALL_VARIANTS adds the measured losers (profiles/r01_variants.json).
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
#ifndef UPOW_MINB
#define UPOW_MINB 3   // resident 256-thread blocks per SM the default build is compiled for
#endif
const Variant VARIANTS[] = {
    UPOW_PLAIN_VARIANT("plain: nvcc-chosen pipes"),
#ifdef UPOW_ALL_VARIANTS
    UPOW_TUNED_VARIANT("add+shr on fma, 3 blocks/SM", TN(1, 1, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr on fma, 4 blocks/SM", TN(1, 1, 0, 0, 0, 0), 4),
    UPOW_TUNED_VARIANT("add+shr+rot25 on 1/4 rounds", TN(1, 1, 4, 1, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr+rot25 on 2/4 rounds", TN(1, 1, 4, 2, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr+rot25 on all rounds", TN(1, 1, 4, 4, 0, 0), 3),
    UPOW_TUNED_VARIANT("add only on fma", TN(1, 0, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add+shr, 7-add rounds", TN(1, 1, 0, 0, 0, 0, 0), 3),
#endif
    UPOW_TUNED_VARIANT("add only, 7-add rounds", TN(1, 0, 0, 0, 0, 0, 0), UPOW_MINB),
#ifdef UPOW_ALL_VARIANTS
    UPOW_TUNED_VARIANT("add, 7-add, shr via IMAD.HI", TN(1, 2, 0, 0, 0, 0, 0), 3),
    UPOW_TUNED_VARIANT("add, 7-add, shr via IMAD.HI on 2/4 words", TN(1, 2, 0, 0, 0, 0, 0, 2), 3),
    UPOW_TUNED_VARIANT("add, 7-add, shr via IMAD.WIDE on 2/4 words", TN(1, 1, 0, 0, 0, 0, 0, 2), 3),
    UPOW_TUNED_VARIANT("add only, 7-add rounds, 4 blocks/SM", TN(1, 0, 0, 0, 0, 0, 0), 4),
    UPOW_TUNED_VARIANT_B("add only, 7-add, 128-thread blocks x7 (28 warps/SM)", TN(1, 0, 0, 0, 0, 0, 0), 7, 128),
    UPOW_TUNED_VARIANT_N("add only, 7-add, 2 nonces per lane, 128-thread blocks x4", TN(1, 0, 0, 0, 0, 0, 0), 4, 128, 2),
#endif
};
constexpr int N_VARIANTS = sizeof(VARIANTS) / sizeof(VARIANTS[0]);
#ifdef UPOW_ALL_VARIANTS
constexpr int DEFAULT_VARIANT = 8;
#else
constexpr int DEFAULT_VARIANT = 1;
#endif

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

// Can kernels on device a write into memory of device b?  Same physical device: always.  Otherwise NVLink /
// PCIe peer access, enabled here once per pair.
void init_peer_access(upow_ctx* ctx) {
    const int n = (int)ctx->devs.size();
    ctx->full_p2p = true;
    for (int a = 0; a < n; a++) {
        Device& da = *ctx->devs[a];
        for (int b = 0; b < n; b++) {
            if (a == b) continue;
            Device& db = *ctx->devs[b];
            // UPOW_NO_P2P=1: never enable peer access between different GPUs; UPOW_NO_P2P=all: also treat two
            // slots on the SAME GPU as unreachable (lets a one-GPU box test the host relay of finds)
            const char* no_p2p = getenv("UPOW_NO_P2P");
            bool ok = da.id == db.id && !(no_p2p && strcmp(no_p2p, "all") == 0);
            if (!ok && da.id != db.id && no_p2p == nullptr) {
                int can = 0;
                if (cudaDeviceCanAccessPeer(&can, da.id, db.id) == cudaSuccess && can && cudaSetDevice(da.id) == cudaSuccess) {
                    cudaError_t e = cudaDeviceEnablePeerAccess(db.id, 0);
                    ok = e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled;
                    cudaGetLastError();
                }
            }
            da.peer_ok[b] = ok;
            if (!ok) ctx->full_p2p = false;
        }
    }
}

// expected hashes until a template's first valid nonce: 16^n_target * 16 / charset
double expected_hashes(const Rule& r) {
    double e = 1;
    for (uint32_t q = 0; q < std::min<uint32_t>(r.n_target, 16); q++) e *= 16;
    return e * 16.0 / r.charset_count;
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

// ---- one launch ------------------------------------------------------------------------
struct Launch {
    int dev = -1;      // index into ctx->devs
    int slot = 0;
    uint32_t serial = 0;
};

// Reserves the next result slot and launch serial of a device (nothing is enqueued yet).
Launch reserve_launch(Device& d) {
    Launch l;
    l.dev = d.index;
    l.slot = (int)(d.slot_seq++ % DEPTH);
    l.serial = ++d.serial;
    return l;
}

void fill_sink(upow_ctx* ctx, Device& d, const Launch& l, const std::vector<Launch>* siblings, HitSink* sink, uint32_t* out) {
    sink->res = d.d_res + l.slot;
    sink->out = out;
    sink->cap = out ? ENUM_CAP : 0;
    sink->n_peers = 0;
    if (!siblings) return;
    for (const Launch& o : *siblings) {
        if (o.dev == l.dev || !d.peer_ok[o.dev]) continue;
        sink->peer_limit[sink->n_peers++] = &(ctx->devs[o.dev]->d_res + o.slot)->limit;
    }
}

// Enqueues: slot reset, kernel, result copy-back.  `fn_launch` issues the kernel on the given stream.
template <class F>
int enqueue_launch(upow_ctx* ctx, Device& d, const Launch& l, F&& fn_launch) {
    CU_TRY(ctx, cudaSetDevice(d.id));
    cudaStream_t st = d.stream[l.slot & 1];
    CU_TRY(ctx, cudaMemcpyAsync(d.d_res + l.slot, d.h_init, sizeof(DevResult), cudaMemcpyHostToDevice, st));
    CU_TRY(ctx, cudaEventRecord(d.ev0[l.slot], st));
    fn_launch(st);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaEventRecord(d.ev1[l.slot], st));
    CU_TRY(ctx, cudaMemcpyAsync(d.h_res + l.slot, d.d_res + l.slot, sizeof(DevResult), cudaMemcpyDeviceToHost, st));
    CU_TRY(ctx, cudaEventRecord(d.ev_done[l.slot], st));
    return UPOW_OK;
}

struct LaunchOut {
    DevResult res;
    float kernel_ms = 0;
    double sm_mhz = 0;
};
int await_launch(upow_ctx* ctx, Device& d, const Launch& l, LaunchOut* out) {
    CU_TRY(ctx, cudaSetDevice(d.id));
    CU_TRY(ctx, cudaEventSynchronize(d.ev_done[l.slot]));
    out->res = d.h_res[l.slot];
    CU_TRY(ctx, cudaEventElapsedTime(&out->kernel_ms, d.ev0[l.slot], d.ev1[l.slot]));
    out->sm_mhz = out->res.clk_ns ? (double)out->res.clk_cycles / (double)out->res.clk_ns * 1000.0 : 0.0;
    return UPOW_OK;
}

// hc: H values per item.  Large items amortise the per-item set-up; small ranges get smaller items so that
// every resident warp has work.
uint32_t pick_hc(uint32_t h_count, uint32_t lg_count, uint32_t n_warps) {
    uint32_t hc = 64;
    while (hc > 1 && (uint64_t)((h_count + hc - 1) / hc) * lg_count < 4ull * n_warps) hc >>= 1;
    return hc;
}

// Arguments of a contiguous launch over nonces [lo, hi) of one (template, timestamp), restricted to the lgroup
// class of device d (lg_first = its index, lg_step = number of devices sharing the range).
void contiguous_args(upow_ctx* ctx, const Device& d, const Prepared& p, uint64_t lo, uint64_t hi, uint32_t lg_first,
                     uint32_t lg_step, SearchArgs* a, uint32_t hc_cap = 64) {
    memset(a, 0, sizeof *a);
    a->tail = p.tail;
    a->rule = p.rule;
    a->nonce_lo = lo;
    a->nonce_hi = hi;
    a->h_lo = (uint32_t)(lo >> 16);
    a->h_count = (uint32_t)((hi + 0xffff) >> 16) - a->h_lo;
    a->lg_first = lg_first;
    a->lg_step = lg_step;
    a->lg_count = lg_first < LGROUPS ? (LGROUPS - lg_first + lg_step - 1) / lg_step : 0;
    const uint32_t n_warps = d.variant_grid[ctx->variant] * (VARIANTS[ctx->variant].block / 32);
    a->hc = std::min(pick_hc(a->h_count, a->lg_count, n_warps), std::max(hc_cap, 1u));
    a->n_items = ((a->h_count + a->hc - 1) / a->hc) * a->lg_count;
    a->kill.word = d.d_kill;
    fill_opaque(&a->q);
    if (p.tail.layout == LAYOUT_V2) prepare_v2_consts(p.tail, &a->v2);
}

void launch_search(const Device& d, const SearchArgs& a, int mode, int variant, cudaStream_t st) {
    SearchFn fn = VARIANTS[variant].fn[a.tail.layout == LAYOUT_V1 ? 0 : 1][mode == MODE_MIN ? 0 : 1];
    fn<<<d.variant_grid[variant], VARIANTS[variant].block, 0, st>>>(a);
}
// variant 0 keeps the plain-CUDA strided kernel (independent implementation); every other variant
// uses the pipe-assigned build
void launch_strided(const Device& d, const StridedArgs& a, bool tuned, cudaStream_t st) {
    const int grid = tuned ? d.strided_tuned_grid : d.grid_blocks;
    if (a.tail.layout == LAYOUT_V1) {
        if (tuned) search_strided_kernel<LAYOUT_V1, MODE_MIN, true><<<grid, 256, 0, st>>>(a);
        else search_strided_kernel<LAYOUT_V1, MODE_MIN, false><<<grid, 256, 0, st>>>(a);
    } else {
        if (tuned) search_strided_kernel<LAYOUT_V2, MODE_MIN, true><<<grid, 256, 0, st>>>(a);
        else search_strided_kernel<LAYOUT_V2, MODE_MIN, false><<<grid, 256, 0, st>>>(a);
    }
}

// A call was cancelled: the flag is consumed by the call that reports it.
int finish_status(upow_ctx* ctx, int rc) {
    if (rc == UPOW_CANCELLED) ctx->cancel.store(0, std::memory_order_relaxed);
    return rc;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C-ABI
// ------------------------------------------------------------------------------------------
#ifdef UPOW_CHECKED
extern "C" const char* upow_version(void) { return "upow_b200 0.2 (sm_100a, checked build: device-side index asserts)"; }
#else
extern "C" const char* upow_version(void) { return "upow_b200 0.2 (sm_100a)"; }
#endif

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
    try {
        ctx->variant = DEFAULT_VARIANT;
        for (int i = 0; i < n_devices; i++) {
            int id = device_ids ? device_ids[i] : i;
            if (id < 0 || id >= visible) {
                set_err(nullptr, "device id out of range");
                upow_destroy(ctx);
                return UPOW_E_ARG;
            }
            ctx->devs.emplace_back(new Device);
            int rc = init_device(ctx, *ctx->devs[i], id, i);
            if (rc == UPOW_OK) rc = init_variant_grids(ctx, *ctx->devs[i]);
            if (rc != UPOW_OK) {
                set_err(nullptr, ctx->err);
                upow_destroy(ctx);
                return rc;
            }
        }
        init_peer_access(ctx);
        if (const char* v = getenv("UPOW_VARIANT")) {
            int k = atoi(v);
            if (k >= 0 && k < N_VARIANTS) ctx->variant = k;
        }
        ctx->wd.th = std::thread(watchdog_main, ctx);
    } catch (const std::exception& ex) {
        set_err(nullptr, std::string("upow_init: ") + ex.what());
        upow_destroy(ctx);
        return UPOW_E_INTERNAL;
    }
    *out_ctx = ctx;
    return UPOW_OK;
}

extern "C" int upow_destroy(upow_ctx* ctx) {
    if (!ctx) return UPOW_E_ARG;
    if (ctx->wd.th.joinable()) {
        {
            std::lock_guard<std::mutex> g(ctx->wd.mu);
            ctx->wd.quit = true;
        }
        ctx->wd.cv.notify_all();
        ctx->wd.th.join();
    }
    for (auto& d : ctx->devs) destroy_device(*d);
    delete ctx;
    return UPOW_OK;
}

extern "C" int upow_device_count(const upow_ctx* ctx) { return ctx ? (int)ctx->devs.size() : UPOW_E_ARG; }

extern "C" int upow_get_device_info(const upow_ctx* ctx, int index, upow_device_info* out) {
    if (!ctx || !out || index < 0 || index >= (int)ctx->devs.size()) return UPOW_E_ARG;
    const Device& d = *ctx->devs[index];
    memset(out, 0, sizeof *out);
    memcpy(out->name, d.prop.name, sizeof out->name - 1);
    out->device_id = d.id;
    out->sm_count = d.sm_count;
    out->cc_major = d.prop.major;
    out->cc_minor = d.prop.minor;
    out->max_sm_clock_khz = d.prop.clockRate;
    out->search_grid_blocks = d.variant_grid[ctx->variant];
    out->total_mem = d.prop.totalGlobalMem;
    int peers = 0;
    for (int j = 0; j < (int)ctx->devs.size(); j++) peers += (j != index && d.peer_ok[j]);
    out->peer_devices = peers;
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
extern "C" int upow_default_variant(void) { return DEFAULT_VARIANT; }
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
extern "C" int upow_get_progress(const upow_ctx* ctx, uint64_t* hashes_done) {
    if (!ctx || !hashes_done) return UPOW_E_ARG;
    *hashes_done = ctx->progress.load(std::memory_order_relaxed);
    return UPOW_OK;
}
extern "C" int upow_clear_cancel(upow_ctx* ctx) {
    if (!ctx) return UPOW_E_ARG;
    ctx->cancel.store(0, std::memory_order_relaxed);
    return UPOW_OK;
}

namespace {

// One unit of a search: a piece of the (timestamp, nonce) order that is launched as a whole.
//   contiguous: H range of one timestamp, run by EVERY device (each its lgroup class);
//   strided:    step range of one timestamp, run by ONE device.
struct Unit {
    uint32_t ts_idx = 0;
    uint64_t lo = 0, hi = 0;   // contiguous: nonce bounds; strided: step bounds
    int only_dev = -1;
    std::vector<Launch> launches;
};

int search_min_impl(upow_ctx* ctx, const Slot& s, uint64_t start_nonce, uint64_t stride, const upow_limits& lim,
                    upow_result* result) {
    const uint64_t max_nonce = lim.max_nonce ? lim.max_nonce : NONCE_SPACE;
    const double t_begin = now_s();
    const double deadline = lim.time_limit_s > 0 ? t_begin + lim.time_limit_s : 0;
    const int n_dev = (int)ctx->devs.size();
    ctx->progress.store(0, std::memory_order_relaxed);
    const bool contiguous = stride == 1;
    const bool tuned_strided = ctx->variant != 0;
    // timestamp of rollover k: ts0 + k, or ts0 - k (mod 2^32 either way)
    auto ts_add = [&](uint32_t k) { return lim.rollover_backwards ? 0u - k : k; };

    Prepared p0;
    if (!prepare_slot(s, 0, &p0)) return UPOW_E_INTERNAL;
    const double e_h = expected_hashes(p0.rule) / 65536.0;   // expected search length in H values

    // ---- the unit sequence, lexicographic in (timestamp, nonce) ----
    uint64_t units_per_ts_total = 0;   // nonces (contiguous) or steps (strided) per timestamp
    if (start_nonce < max_nonce)
        units_per_ts_total = contiguous ? max_nonce - start_nonce : (max_nonce - start_nonce + stride - 1) / stride;
    const uint64_t n_ts = (uint64_t)lim.max_rollovers + 1;
    // contiguous: the first unit of a timestamp covers about 4x the expected search length, later ones grow
    // eightfold -- an easy difficulty is answered by a launch of microseconds, a hard one by one launch per
    // timestamp.  tile_nonces != 0 fixes the granule instead.  strided: fixed tiles of steps.
    uint64_t first_granule, growth;
    if (contiguous) {
        if (lim.tile_nonces) { first_granule = std::max<uint64_t>((lim.tile_nonces + 0xffff) >> 16, 1) << 16; growth = 1; }
        else {
            double g = std::max(4.0, 4.0 * e_h);
            uint64_t gh = 4;
            while ((double)gh < g && gh < 65536) gh <<= 1;
            first_granule = gh << 16; growth = 8;
        }
    } else {
        first_granule = lim.tile_nonces ? lim.tile_nonces : (n_dev > 1 ? (1ull << 26) : NONCE_SPACE);
        first_granule = std::max<uint64_t>(first_granule, 1ull << 10);
        growth = 1;
    }
    // The warp scheduler does not advance all warps evenly, so when a find is published some warps are still on
    // early items while others have run ahead by a few items: short items keep that spread (work done beyond
    // the answer) small where the whole search is short; long searches get the full 64-H items.
    uint32_t hc_cap = 64;
    while (hc_cap > 1 && (double)hc_cap * 32.0 > e_h) hc_cap >>= 1;
    uint32_t cur_ts = 0;
    uint64_t cur_pos = 0, cur_granule = first_granule;   // position inside the timestamp (nonces past start / steps)
    int rr_dev = 0;
    auto next_unit = [&](Unit* u) -> bool {
        if (units_per_ts_total == 0) return false;
        if (cur_pos >= units_per_ts_total) { cur_ts++; cur_pos = 0; cur_granule = first_granule; }
        if (cur_ts >= n_ts) return false;
        u->ts_idx = cur_ts;
        if (contiguous) {
            // granules end on multiples of 2^16 so that no H value is shared by two units
            uint64_t lo = start_nonce + cur_pos;
            uint64_t hi = std::min<uint64_t>(max_nonce, ((lo >> 16) << 16) + cur_granule);
            u->lo = lo; u->hi = hi; u->only_dev = -1;
            cur_pos = hi - start_nonce;
        } else {
            u->lo = cur_pos; u->hi = std::min(units_per_ts_total, cur_pos + cur_granule);
            u->only_dev = rr_dev; rr_dev = (rr_dev + 1) % n_dev;
            cur_pos = u->hi;
        }
        if (growth > 1 && cur_granule < NONCE_SPACE) cur_granule *= growth;
        return true;
    };

    // ---- pipeline state ----
    std::mutex launch_mu;                 // launches vs. the watchdog's kill writes
    std::atomic<int> stop_status{0};      // UPOW_CANCELLED / UPOW_EXHAUSTED(deadline) once the watchdog fired
    std::vector<uint32_t> first_serial(n_dev, 0);
    for (int i = 0; i < n_dev; i++) first_serial[i] = ctx->devs[i]->serial + 1;
    WatchdogScope wd(ctx, deadline, [&](int status) {
        std::lock_guard<std::mutex> lk(launch_mu);
        stop_status.store(status);
        for (int i = 0; i < n_dev; i++)
            if (ctx->devs[i]->serial >= first_serial[i]) write_kill(*ctx->devs[i], first_serial[i], ctx->devs[i]->serial);
    });

    std::vector<Unit> ring(contiguous ? INFLIGHT : INFLIGHT * n_dev);
    const size_t window = ring.size();
    uint64_t launched = 0, retired = 0;
    bool more = true, found = false;
    uint32_t found_ts = 0, found_nonce = 0;
    uint32_t found_digest[8] = {0};
    int rc_err = UPOW_OK;
    uint32_t cached_ts = ~0u;
    Prepared p;
    std::vector<bool> dev_used(n_dev, false);
    std::vector<int> last_slot(n_dev, -1);

    auto launch_unit = [&](Unit& u) -> int {
        std::lock_guard<std::mutex> lk(launch_mu);
        if (stop_status.load()) return 1;   // nothing is launched after the watchdog fired
        if (u.ts_idx != cached_ts) {
            if (!prepare_slot(s, ts_add(u.ts_idx), &p)) return UPOW_E_INTERNAL;
            cached_ts = u.ts_idx;
        }
        u.launches.clear();
        for (int di = 0; di < n_dev; di++)
            if (u.only_dev < 0 || u.only_dev == di) u.launches.push_back(reserve_launch(*ctx->devs[di]));
        for (const Launch& l : u.launches) {
            Device& d = *ctx->devs[l.dev];
            if (!dev_used[l.dev]) {
                CU_TRY(ctx, cudaSetDevice(d.id));
                CU_TRY(ctx, cudaEventRecord(d.ev_begin, d.stream[l.slot & 1]));
                dev_used[l.dev] = true;
            }
            int rc;
            if (contiguous) {
                SearchArgs a;
                contiguous_args(ctx, d, p, u.lo, u.hi, (uint32_t)l.dev, (uint32_t)n_dev, &a, hc_cap);
                fill_sink(ctx, d, l, &u.launches, &a.sink, nullptr);
                a.kill.serial = l.serial;
                rc = enqueue_launch(ctx, d, l, [&](cudaStream_t st) { launch_search(d, a, MODE_MIN, ctx->variant, st); });
            } else {
                StridedArgs a;
                memset(&a, 0, sizeof a);
                a.tail = p.tail;
                a.rule = p.rule;
                a.start = start_nonce + u.lo * stride;
                a.stride = stride;
                a.n_steps = u.hi - u.lo;
                uint32_t ic = 64;
                const uint32_t n_warps = (tuned_strided ? d.strided_tuned_grid : d.grid_blocks) * 8;
                while (ic > 1 && (a.n_steps + 32ull * ic - 1) / (32ull * ic) < 4ull * n_warps) ic >>= 1;
                a.ic = ic;
                a.n_items = (uint32_t)((a.n_steps + 32ull * ic - 1) / (32ull * ic));
                fill_sink(ctx, d, l, nullptr, &a.sink, nullptr);
                a.kill.word = d.d_kill;
                a.kill.serial = l.serial;
                fill_opaque(&a.q);
                if (p.tail.layout == LAYOUT_V2) prepare_v2_consts(p.tail, &a.v2);
                rc = enqueue_launch(ctx, d, l, [&](cudaStream_t st) { launch_strided(d, a, tuned_strided, st); });
            }
            if (rc != UPOW_OK) return rc;
            result->launches++;
        }
        return UPOW_OK;
    };

    for (;;) {
        // With growing units the first one covers four times the expected search length: it is launched alone
        // (anything queued behind it would hash for nothing until the host has seen the find); after it,
        // launches are pipelined.
        const size_t depth = (growth > 1 && retired == 0) ? 1 : window;
        while (more && !found && rc_err == UPOW_OK && !stop_status.load() && launched - retired < depth) {
            Unit& u = ring[launched % window];
            if (!next_unit(&u)) { more = false; break; }
            int rc = launch_unit(u);
            if (rc == 1) { u.launches.clear(); more = false; break; }
            if (rc != UPOW_OK) { rc_err = rc; more = false; }
            launched++;   // even a failed unit is retired below so that whatever part of it was enqueued is awaited
        }
        if (retired == launched) break;
        Unit& u = ring[retired % window];
        bool unit_found = false;
        uint32_t unit_nonce = 0;
        const uint32_t* unit_digest = nullptr;
        for (const Launch& l : u.launches) {
            Device& d = *ctx->devs[l.dev];
            LaunchOut o;
            int rc = await_launch(ctx, d, l, &o);
            if (rc != UPOW_OK) { if (rc_err == UPOW_OK) rc_err = rc; continue; }
            last_slot[l.dev] = l.slot;
            result->per_device_hashes[l.dev] += o.res.h_steps * 32ull;
            ctx->progress.fetch_add(o.res.h_steps * 32ull, std::memory_order_relaxed);
            if (l.dev == 0 && o.sm_mhz > 0) result->sm_clock_mhz = o.sm_mhz;
            if (!found && o.res.best != NO_NONCE && (!unit_found || (uint32_t)o.res.best < unit_nonce)) {
                unit_found = true;
                unit_nonce = (uint32_t)o.res.best;
                unit_digest = d.h_res[l.slot].digest;
            }
        }
        retired++;
        if (unit_found && !found && !stop_status.load()) {
            // a unit whose launches were cut short by cancel / deadline may have skipped nonces: only a unit
            // that ran undisturbed yields the minimum
            found = true;
            found_ts = u.ts_idx;
            found_nonce = unit_nonce;
            memcpy(found_digest, unit_digest, sizeof found_digest);
            // everything queued behind this unit is pointless now
            std::lock_guard<std::mutex> lk(launch_mu);
            for (int di = 0; di < n_dev; di++) {
                uint32_t lo = 0;
                for (uint64_t k = retired; k < launched; k++)
                    for (const Launch& l : ring[k % window].launches)
                        if (l.dev == di && (lo == 0 || l.serial < lo)) lo = l.serial;
                if (lo) write_kill(*ctx->devs[di], lo, ctx->devs[di]->serial);
            }
        }
    }

    for (int i = 0; i < n_dev; i++) {
        Device& d = *ctx->devs[i];
        if (dev_used[i] && last_slot[i] >= 0 && cudaSetDevice(d.id) == cudaSuccess) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, d.ev_begin, d.ev1[last_slot[i]]) == cudaSuccess) result->per_device_kernel_ms[i] = ms;
        }
        result->hashes_done += result->per_device_hashes[i];
        result->kernel_ms = std::max(result->kernel_ms, result->per_device_kernel_ms[i]);
    }
    int rc = UPOW_EXHAUSTED;
    const int st = stop_status.load();
    if (rc_err != UPOW_OK) rc = rc_err;
    else if (found) {
        Prepared pf;
        prepare_slot(s, ts_add(found_ts), &pf);
        for (int i = 0; i < 8; i++) store_be32(result->digest + 4 * i, found_digest[i]);
        if (!host_verify(pf, found_nonce, result->digest)) {
            set_err(ctx, "device result failed the host SHA-256 / rule cross-check");
            return UPOW_E_INTERNAL;
        }
        result->found = 1;
        result->rollovers = found_ts;
        result->timestamp = get_prefix_timestamp(pf.prefix, pf.prefix_len);
        result->nonce = found_nonce;
        rc = UPOW_OK;
    } else if (st != 0) rc = st;  // cancelled or timed out: a partial search proves nothing
    result->elapsed_s = now_s() - t_begin;
    return rc;
}

}  // namespace

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
        if (lim.max_nonce > NONCE_SPACE) {
            set_err(ctx, "upow_search_min: max_nonce > 2^32");
            return UPOW_E_ARG;
        }
        if (ctx->cancel.load(std::memory_order_relaxed)) return finish_status(ctx, UPOW_CANCELLED);  // cancelled before it began
        return finish_status(ctx, search_min_impl(ctx, s, start_nonce, stride, lim, result));
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
        if (ctx->cancel.load(std::memory_order_relaxed)) return finish_status(ctx, UPOW_CANCELLED);
        const double t_begin = now_s();
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
        std::mutex launch_mu;
        std::atomic<int> stop_status{0};
        std::vector<Launch> launches(n_dev);
        std::vector<bool> used(n_dev, false);
        WatchdogScope wd(ctx, 0, [&](int status) {
            std::lock_guard<std::mutex> lk(launch_mu);
            stop_status.store(status);
            for (int i = 0; i < n_dev; i++)
                if (used[i]) write_kill(*ctx->devs[i], launches[i].serial, launches[i].serial);
        });
        int rc = UPOW_OK;
        for (int di = 0; di < n_dev && rc == UPOW_OK; di++) {
            if (cut[di] >= cut[di + 1]) continue;
            std::lock_guard<std::mutex> lk(launch_mu);
            if (stop_status.load()) break;
            Device& d = *ctx->devs[di];
            launches[di] = reserve_launch(d);
            SearchArgs a;
            contiguous_args(ctx, d, p, cut[di], cut[di + 1], 0, 1, &a);
            fill_sink(ctx, d, launches[di], nullptr, &a.sink, d.d_out);
            a.kill.serial = launches[di].serial;
            rc = enqueue_launch(ctx, d, launches[di], [&](cudaStream_t st) { launch_search(d, a, MODE_ENUM, ctx->variant, st); });
            used[di] = true;
        }
        uint64_t total = 0;
        bool dev_overflow = false;
        std::vector<uint32_t> all;
        for (int di = 0; di < n_dev; di++) {
            if (!used[di]) continue;
            Device& d = *ctx->devs[di];
            LaunchOut o;
            int arc = await_launch(ctx, d, launches[di], &o);
            if (arc != UPOW_OK) { if (rc == UPOW_OK) rc = arc; continue; }
            total += o.res.count;
            if (o.res.count > ENUM_CAP) dev_overflow = true;
            const uint64_t n = std::min<uint64_t>(o.res.count, ENUM_CAP);
            if (n) {
                const size_t at = all.size();
                all.resize(at + n);
                cudaError_t e = cudaMemcpyAsync(all.data() + at, d.d_out, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, d.stream[launches[di].slot & 1]);
                if (e == cudaSuccess) e = cudaStreamSynchronize(d.stream[launches[di].slot & 1]);
                if (e != cudaSuccess) { set_err(ctx, cudaGetErrorString(e)); if (rc == UPOW_OK) rc = UPOW_E_CUDA; }
            }
            if (stats) {
                stats->per_device_hashes[di] = o.res.h_steps * 32ull;
                stats->per_device_kernel_ms[di] = o.kernel_ms;
                stats->hashes_done += o.res.h_steps * 32ull;
                stats->kernel_ms = std::max(stats->kernel_ms, (double)o.kernel_ms);
                stats->launches++;
                if (di == 0) stats->sm_clock_mhz = o.sm_mhz;
            }
        }
        if (rc != UPOW_OK) return rc;
        if (stop_status.load() == UPOW_CANCELLED) rc = UPOW_CANCELLED;
        std::sort(all.begin(), all.end());
        const uint64_t n_copy = std::min<uint64_t>(all.size(), cap);
        if (n_copy) memcpy(out_nonces, all.data(), n_copy * sizeof(uint32_t));
        *out_count = total;
        if (stats) stats->elapsed_s = now_s() - t_begin;
        if (rc == UPOW_OK && (dev_overflow || total > cap)) {
            set_err(ctx, "upow_sweep: more valid nonces than the output buffer holds");
            rc = UPOW_E_OVERFLOW;
        }
        return finish_status(ctx, rc);
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

int ensure_batch_capacity(upow_ctx* ctx, Device& d, uint32_t n) {
    if (n <= d.batch_cap && d.d_ctl) return UPOW_OK;
    CU_TRY(ctx, cudaSetDevice(d.id));
    cudaFree(d.d_jobs); cudaFree(d.d_active[0]); cudaFree(d.d_active[1]); cudaFree(d.d_limit); cudaFree(d.d_limit_in);
    cudaFree(d.d_bdigest); cudaFree(d.d_ctl);
    d.d_jobs = nullptr; d.d_active[0] = d.d_active[1] = nullptr; d.d_limit = d.d_limit_in = nullptr; d.d_bdigest = nullptr;
    d.d_ctl = nullptr;
    d.batch_cap = 0;
    CU_TRY(ctx, cudaMalloc(&d.d_jobs, sizeof(BatchJob) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_active[0], sizeof(uint32_t) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_active[1], sizeof(uint32_t) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_limit, sizeof(unsigned long long) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_limit_in, sizeof(unsigned long long) * n));
    CU_TRY(ctx, cudaMalloc(&d.d_bdigest, 32 * (size_t)n));
    CU_TRY(ctx, cudaMalloc(&d.d_ctl, sizeof(BatchCtl)));
    d.batch_cap = n;
    return UPOW_OK;
}

// limit[i] = min(limit[i], incoming[i]): how finds travel between GPUs that have no peer access (host relay)
__global__ void merge_limits_kernel(unsigned long long* limit, const unsigned long long* incoming, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && incoming[i] != NO_NONCE) atomicMin(&limit[i], incoming[i]);
}

struct Round {
    uint32_t h_lo, h_count, hc;
};

// The batch engine behind upow_search_batch and upow_search_any.
//
// Every GPU holds the whole job table and searches ITS lgroup class of every template, so all GPUs do the
// same work whatever the luck of the individual templates.  The search advances in ROUNDS (H ranges that are
// the same for every template and do not depend on the number of GPUs); one kernel launch per round, plus a
// small kernel that drops the solved templates from the device-resident work list.  Rounds are queued in
// chunks with no host round trip in between; the host only looks at a snapshot of the device state once per
// chunk, one chunk late, to see whether anything is left to do.  Finds travel between the GPUs as 8-byte
// atomicMin writes into each other's limit[] arrays.
// stop_at_first: rounds after the first one in which anything was found do nothing.
// best[i] = minimum valid nonce of template i or NO_NONCE; digests (8 words per template) from device 0.
int run_batch(upow_ctx* ctx, const std::vector<Prepared>& prep, uint64_t max_nonce, bool stop_at_first,
              std::vector<unsigned long long>* best_out, std::vector<uint32_t>* digests_out, upow_result* stats,
              uint32_t* first_round_out, std::vector<Round>* rounds_out) {
    const uint32_t n = (uint32_t)prep.size();
    const int n_dev = (int)ctx->devs.size();
    best_out->assign(n, NO_NONCE);
    digests_out->assign((size_t)n * 8, 0);
    if (first_round_out) *first_round_out = 0xffffffffu;
    if (n == 0) return UPOW_OK;
    std::vector<BatchJob> jobs(n);
    bool has_v1 = false, has_v2 = false;
    double e_min = 1e300;
    for (uint32_t i = 0; i < n; i++) {
        jobs[i].tail = prep[i].tail;
        jobs[i].rule = prep[i].rule;
        if (jobs[i].tail.layout == LAYOUT_V2) prepare_v2_consts(jobs[i].tail, &jobs[i].v2);
        else memset(&jobs[i].v2, 0, sizeof jobs[i].v2);
        (jobs[i].tail.layout == LAYOUT_V1 ? has_v1 : has_v2) = true;
        e_min = std::min(e_min, expected_hashes(jobs[i].rule));
    }
    std::vector<uint32_t> iota(n);
    for (uint32_t i = 0; i < n; i++) iota[i] = i;

    std::mutex launch_mu;
    std::atomic<int> stop_status{0};
    std::vector<uint32_t> first_serial(n_dev);
    for (int i = 0; i < n_dev; i++) first_serial[i] = ctx->devs[i]->serial + 1;
    WatchdogScope wd(ctx, 0, [&](int status) {
        std::lock_guard<std::mutex> lk(launch_mu);
        stop_status.store(status);
        for (int i = 0; i < n_dev; i++)
            if (ctx->devs[i]->serial >= first_serial[i]) write_kill(*ctx->devs[i], first_serial[i], ctx->devs[i]->serial);
    });

    BatchCtl ctl0;
    memset(&ctl0, 0, sizeof ctl0);
    ctl0.n_active[0] = n;
    ctl0.first_find_round = 0xffffffffu;
    for (auto& dp : ctx->devs) {
        Device& d = *dp;
        int rc = ensure_batch_capacity(ctx, d, n);
        if (rc != UPOW_OK) return rc;
        CU_TRY(ctx, cudaSetDevice(d.id));
        cudaStream_t st = d.stream[0];
        CU_TRY(ctx, cudaMemcpyAsync(d.d_jobs, jobs.data(), sizeof(BatchJob) * n, cudaMemcpyHostToDevice, st));
        CU_TRY(ctx, cudaMemcpyAsync(d.d_active[0], iota.data(), sizeof(uint32_t) * n, cudaMemcpyHostToDevice, st));
        CU_TRY(ctx, cudaMemsetAsync(d.d_limit, 0xff, sizeof(unsigned long long) * n, st));
        CU_TRY(ctx, cudaMemcpyAsync(d.d_ctl, &ctl0, sizeof ctl0, cudaMemcpyHostToDevice, st));
    }
    // every GPU's arrays must be in place before any kernel may write a find into them
    for (auto& dp : ctx->devs) {
        CU_TRY(ctx, cudaSetDevice(dp->id));
        CU_TRY(ctx, cudaStreamSynchronize(dp->stream[0]));
        CU_TRY(ctx, cudaEventRecord(dp->ev_begin, dp->stream[0]));
    }

    // round sizes: a quarter of the easiest template's expected search length, doubling every 8 rounds
    const uint32_t h_total = (uint32_t)((max_nonce + 0xffff) >> 16);
    uint32_t base = 1;
    while ((double)base * 4 * 65536.0 < e_min && base < 1024) base <<= 1;
    // item length inside a round: short items keep the warps of one template close together, so little is
    // hashed beyond a template's answer; long items amortise the per-item set-up (measured: profiles/r02_batch_items.json)
    uint32_t hc_max = 16;
    if (const char* e = getenv("UPOW_BATCH_HC")) hc_max = (uint32_t)std::max(1, atoi(e));
    if (const char* e = getenv("UPOW_BATCH_ROUND")) base = (uint32_t)std::max(1, atoi(e));
    // how often a warp re-reads its template's limit inside an item (H values): every hash evaluated after a
    // template's answer is wasted, and with thousands of templates those tails add up
    uint32_t poll = 2;   // measured (profiles/r02_batch_items.json): 16-H items polled every 2 H give the shortest cfg-5 run
    if (const char* e = getenv("UPOW_BATCH_POLL")) poll = (uint32_t)std::max(1, atoi(e));
    while (poll & (poll - 1)) poll &= poll - 1;   // power of two
    std::vector<Round>& rounds = *rounds_out;
    rounds.clear();
    constexpr int CHUNK = 8;
    uint32_t h_done = 0, parity = 0;
    int chunk_no = 0;
    uint32_t snap_parity[2] = {0, 0};   // which n_active entry is current in each snapshot
    bool finished = false;
    std::vector<unsigned long long> merged, tmp;
    while (!finished) {
        const uint32_t chunk_first_round = (uint32_t)rounds.size();
        {
            std::lock_guard<std::mutex> lk(launch_mu);
            if (stop_status.load()) break;
            for (int k = 0; k < CHUNK && h_done < h_total; k++) {
                Round r;
                const uint32_t ri = (uint32_t)rounds.size();
                uint64_t hr = (uint64_t)base << std::min<uint32_t>(ri / 8, 12);
                hr = std::min<uint64_t>(std::min<uint64_t>(hr, 4096), h_total - h_done);
                r.h_lo = h_done; r.h_count = (uint32_t)hr;
                r.hc = (uint32_t)std::min<uint64_t>(hc_max, hr);
                rounds.push_back(r);
                for (auto& dp : ctx->devs) {
                    Device& d = *dp;
                    CU_TRY(ctx, cudaSetDevice(d.id));
                    cudaStream_t st = d.stream[0];
                    BatchArgs a;
                    memset(&a, 0, sizeof a);
                    a.jobs = d.d_jobs; a.active = d.d_active[parity]; a.limit = d.d_limit; a.ctl = d.d_ctl;
                    for (auto& op : ctx->devs) {
                        if (op->index == d.index || !d.peer_ok[op->index]) continue;
                        a.peer_limit[a.n_peers] = op->d_limit;
                        a.peer_first_find[a.n_peers] = &op->d_ctl->first_find_round;
                        a.n_peers++;
                    }
                    a.nonce_hi = max_nonce;
                    a.n_jobs = n;
                    a.parity = parity; a.round = ri; a.stop_at_first = stop_at_first ? 1 : 0;
                    a.h_lo = r.h_lo; a.h_count = r.h_count; a.hc = r.hc;
                    a.lg_first = (uint32_t)d.index; a.lg_step = (uint32_t)n_dev;
                    a.lg_count = (LGROUPS - a.lg_first + a.lg_step - 1) / a.lg_step;
                    a.poll_mask = poll - 1;
                    a.kill.word = d.d_kill;
                    fill_opaque(&a.q);
                    if (has_v1) {
                        a.kill.serial = ++d.serial;
                        batch_kernel<LAYOUT_V1><<<d.batch_grid, 256, 0, st>>>(a);
                        if (stats) stats->launches++;
                    }
                    if (has_v1 && has_v2)  // both launches walk the same item list: rewind the queue in between
                        CU_TRY(ctx, cudaMemsetAsync(&d.d_ctl->next_item, 0, sizeof(unsigned long long), st));
                    if (has_v2) {
                        a.kill.serial = ++d.serial;
                        batch_kernel<LAYOUT_V2><<<d.batch_grid, 256, 0, st>>>(a);
                        if (stats) stats->launches++;
                    }
                    CU_TRY(ctx, cudaMemsetAsync(&d.d_ctl->n_active[parity ^ 1], 0, sizeof(unsigned int), st));
                    batch_compact_kernel<<<(n + 255) / 256, 256, 0, st>>>(d.d_active[parity], d.d_active[parity ^ 1], d.d_limit,
                                                                      d.d_ctl, parity, r.h_lo + r.h_count, n);
                    CU_TRY(ctx, cudaGetLastError());
                }
                parity ^= 1;
                h_done += (uint32_t)hr;
            }
            snap_parity[chunk_no & 1] = parity;
            for (auto& dp : ctx->devs) {
                Device& d = *dp;
                CU_TRY(ctx, cudaSetDevice(d.id));
                CU_TRY(ctx, cudaMemcpyAsync(&d.h_ctl[chunk_no & 1], d.d_ctl, sizeof(BatchCtl), cudaMemcpyDeviceToHost, d.stream[0]));
                CU_TRY(ctx, cudaEventRecord(d.ev_aux[chunk_no & 1], d.stream[0]));
            }
        }
        const bool last_chunk = h_done >= h_total;
        // look at the PREVIOUS chunk's snapshot (this one keeps the GPUs busy meanwhile); after the last chunk,
        // at this one's as well
        for (int look = chunk_no - 1; look <= chunk_no; look++) {
            if (look < 0 || (look == chunk_no && !last_chunk)) continue;
            uint32_t active_max = 0, ffr = 0xffffffffu;
            for (auto& dp : ctx->devs) {
                Device& d = *dp;
                CU_TRY(ctx, cudaSetDevice(d.id));
                CU_TRY(ctx, cudaEventSynchronize(d.ev_aux[look & 1]));
                const BatchCtl& c = d.h_ctl[look & 1];
                active_max = std::max(active_max, c.n_active[snap_parity[look & 1]]);
                ffr = std::min(ffr, c.first_find_round);
            }
            const uint32_t rounds_seen = look == chunk_no ? (uint32_t)rounds.size() : chunk_first_round;
            if (active_max == 0 || (stop_at_first && ffr < rounds_seen) || (look == chunk_no && last_chunk)) finished = true;
            if (!finished && !ctx->full_p2p) {
                // no peer access somewhere: the host carries the finds around, one chunk late
                merged.assign(n, NO_NONCE);
                tmp.resize(n);
                for (auto& dp : ctx->devs) {
                    CU_TRY(ctx, cudaSetDevice(dp->id));
                    CU_TRY(ctx, cudaMemcpyAsync(tmp.data(), dp->d_limit, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost, dp->ctrl));
                    CU_TRY(ctx, cudaStreamSynchronize(dp->ctrl));
                    for (uint32_t i = 0; i < n; i++) merged[i] = std::min(merged[i], tmp[i]);
                }
                for (auto& dp : ctx->devs) {
                    CU_TRY(ctx, cudaSetDevice(dp->id));
                    CU_TRY(ctx, cudaMemcpyAsync(dp->d_limit_in, merged.data(), sizeof(unsigned long long) * n, cudaMemcpyHostToDevice, dp->ctrl));
                    merge_limits_kernel<<<(n + 255) / 256, 256, 0, dp->ctrl>>>(dp->d_limit, dp->d_limit_in, n);
                    CU_TRY(ctx, cudaStreamSynchronize(dp->ctrl));
                }
            }
            if (finished) break;
        }
        chunk_no++;
    }
    // drain (queued rounds beyond the end find nothing to do and leave at once), collect
    std::vector<unsigned long long>& best = *best_out;
    uint32_t ffr = 0xffffffffu;
    tmp.resize(n);
    for (int di = 0; di < n_dev; di++) {
        Device& d = *ctx->devs[di];
        CU_TRY(ctx, cudaSetDevice(d.id));
        CU_TRY(ctx, cudaEventRecord(d.ev1[0], d.stream[0]));
        CU_TRY(ctx, cudaMemcpyAsync(&d.h_ctl[0], d.d_ctl, sizeof(BatchCtl), cudaMemcpyDeviceToHost, d.stream[0]));
        CU_TRY(ctx, cudaMemcpyAsync(tmp.data(), d.d_limit, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost, d.stream[0]));
        CU_TRY(ctx, cudaStreamSynchronize(d.stream[0]));
        for (uint32_t i = 0; i < n; i++) best[i] = std::min(best[i], tmp[i]);
        ffr = std::min(ffr, d.h_ctl[0].first_find_round);
        float ms = 0;
        CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev_begin, d.ev1[0]));
        if (stats) {
            stats->per_device_hashes[di] = d.h_ctl[0].h_steps * 32ull;
            stats->per_device_kernel_ms[di] = ms;
            stats->hashes_done += stats->per_device_hashes[di];
            stats->kernel_ms = std::max(stats->kernel_ms, (double)ms);
        }
    }
    if (stop_status.load() == UPOW_CANCELLED) return UPOW_CANCELLED;
    if (first_round_out) *first_round_out = ffr;
    // digests of all winners on device 0 (its limit[] is overwritten with the merged minima first)
    Device& d0 = *ctx->devs[0];
    CU_TRY(ctx, cudaSetDevice(d0.id));
    CU_TRY(ctx, cudaMemcpyAsync(d0.d_limit, best.data(), sizeof(unsigned long long) * n, cudaMemcpyHostToDevice, d0.stream[0]));
    batch_digest_kernel<<<(n + 127) / 128, 128, 0, d0.stream[0]>>>(d0.d_jobs, d0.d_limit, n, d0.d_bdigest);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaMemcpyAsync(digests_out->data(), d0.d_bdigest, 32 * (size_t)n, cudaMemcpyDeviceToHost, d0.stream[0]));
    CU_TRY(ctx, cudaStreamSynchronize(d0.stream[0]));
    for (uint32_t i = 0; i < n; i++)
        if (best[i] != NO_NONCE) best[i] -= 1;   // limit = nonce + 1
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
        if (ctx->cancel.load(std::memory_order_relaxed)) return finish_status(ctx, UPOW_CANCELLED);
        const double t_begin = now_s();
        std::vector<Prepared> prep;
        int prc = prepare_templates(ctx, n_templates, templates, &prep, "upow_search_batch");
        if (prc != UPOW_OK) return prc;
        std::vector<unsigned long long> best;
        std::vector<uint32_t> dig;
        std::vector<Round> rounds;
        int rc = run_batch(ctx, prep, max_nonce, false, &best, &dig, stats, nullptr, &rounds);
        if (rc != UPOW_OK) return finish_status(ctx, rc);
        // every result is cross-checked with the host SHA-256 (spread over a few host threads)
        std::atomic<int> vrc{UPOW_OK};
        auto fill = [&](uint32_t lo, uint32_t hi) {
            for (uint32_t i = lo; i < hi; i++) {
                memset(&results[i], 0, sizeof results[i]);
                if (best[i] == NO_NONCE) continue;
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
        if (ctx->cancel.load(std::memory_order_relaxed)) return finish_status(ctx, UPOW_CANCELLED);
        const double t_begin = now_s();
        std::vector<Prepared> prep;
        int prc = prepare_templates(ctx, n_templates, templates, &prep, "upow_search_any");
        if (prc != UPOW_OK) return prc;
        std::vector<unsigned long long> best;
        std::vector<uint32_t> dig;
        std::vector<Round> rounds;
        uint32_t first_round = 0xffffffffu;
        int rc = run_batch(ctx, prep, max_nonce, true, &best, &dig, stats, &first_round, &rounds);
        if (stats) stats->elapsed_s = now_s() - t_begin;
        if (rc != UPOW_OK) return finish_status(ctx, rc);
        if (first_round == 0xffffffffu || first_round >= rounds.size()) return UPOW_EXHAUSTED;
        // every GPU completed the first round with a find (later rounds may have run on a GPU that heard of
        // the find late: their results are ignored).  The round boundaries do not depend on the number of
        // GPUs, so neither does the answer: the lowest template index with a valid nonce inside that round.
        const unsigned long long r_lo = (unsigned long long)rounds[first_round].h_lo << 16;
        const unsigned long long r_hi = r_lo + ((unsigned long long)rounds[first_round].h_count << 16);
        uint32_t win = 0xffffffffu;
        for (uint32_t i = 0; i < n_templates; i++)
            if (best[i] != NO_NONCE && best[i] >= r_lo && best[i] < r_hi) { win = i; break; }
        if (win == 0xffffffffu) {
            set_err(ctx, "upow_search_any: no result inside the first round with a find");
            return UPOW_E_INTERNAL;
        }
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

// ---- staged transfers for the two bulk entry points (block verifier, batched hashing) ----------------
// Caller memory is ordinary pageable memory: a single cudaMemcpy of it runs at the speed of ONE host
// thread copying into the driver's staging buffer (~10 GB/s), far below the link.  Here every GPU has
// NLANES lanes, each a stream with two pinned + two device buffers; a lane is driven by its own host
// thread, which copies chunk k+1 into pinned memory while chunk k is on the wire or in the kernel.
// All buffers are allocated once per device (grow-never: fixed size) on the first bulk call.
int ensure_lanes(upow_ctx* ctx, Device& d) {
    if (d.lanes_ready) return UPOW_OK;
    CU_TRY(ctx, cudaSetDevice(d.id));
    for (int l = 0; l < NLANES; l++) {
        Lane& ln = d.lanes[l];
        CU_TRY(ctx, cudaStreamCreateWithFlags(&ln.st, cudaStreamNonBlocking));
        for (int b = 0; b < 2; b++) {
            CU_TRY(ctx, cudaMallocHost(&ln.h[b], LANE_BYTES));
            CU_TRY(ctx, cudaMalloc(&ln.d[b], LANE_BYTES));
            CU_TRY(ctx, cudaEventCreateWithFlags(&ln.done[b], cudaEventDisableTiming));
            CU_TRY(ctx, cudaEventCreate(&ln.k0[b]));
            CU_TRY(ctx, cudaEventCreate(&ln.k1[b]));
        }
    }
    d.lanes_ready = true;
    return UPOW_OK;
}

// Runs `n_chunks` chunks through the lanes of one device.  stage_in(c, pinned) fills the pinned buffer and
// returns the bytes to send; run(c, dev_buf, stream) launches the kernel; out_bytes(c, &offset) says which
// part of the device buffer comes back; stage_out(c, pinned + offset) hands it to the caller.
struct ChunkPlan {
    uint32_t n_chunks = 0;
    std::function<size_t(uint32_t, uint8_t*)> stage_in;
    std::function<void(uint32_t, uint8_t*, cudaStream_t)> run;
    std::function<size_t(uint32_t, size_t*)> out_bytes;
    std::function<void(uint32_t, const uint8_t*)> stage_out;
};

int run_lane(upow_ctx* ctx, Device& d, int lane, int n_lanes, const ChunkPlan& plan, double* kernel_ms) {
    CU_TRY(ctx, cudaSetDevice(d.id));
    Lane& ln = d.lanes[lane];
    int64_t pending[2] = {-1, -1};
    auto finish = [&](int b) -> int {
        if (pending[b] < 0) return UPOW_OK;
        CU_TRY(ctx, cudaEventSynchronize(ln.done[b]));
        size_t off = 0;
        plan.out_bytes((uint32_t)pending[b], &off);
        plan.stage_out((uint32_t)pending[b], ln.h[b] + off);
        float ms = 0;
        CU_TRY(ctx, cudaEventElapsedTime(&ms, ln.k0[b], ln.k1[b]));
        *kernel_ms += ms;
        pending[b] = -1;
        return UPOW_OK;
    };
    int k = 0;
    for (uint32_t c = (uint32_t)lane; c < plan.n_chunks; c += (uint32_t)n_lanes, k++) {
        const int b = k & 1;
        int rc = finish(b);
        if (rc != UPOW_OK) return rc;
        const size_t in_bytes = plan.stage_in(c, ln.h[b]);
        if (in_bytes) CU_TRY(ctx, cudaMemcpyAsync(ln.d[b], ln.h[b], in_bytes, cudaMemcpyHostToDevice, ln.st));
        CU_TRY(ctx, cudaEventRecord(ln.k0[b], ln.st));
        plan.run(c, ln.d[b], ln.st);
        CU_TRY(ctx, cudaGetLastError());
        CU_TRY(ctx, cudaEventRecord(ln.k1[b], ln.st));
        size_t off = 0;
        const size_t ob = plan.out_bytes(c, &off);
        if (ob) CU_TRY(ctx, cudaMemcpyAsync(ln.h[b] + off, ln.d[b] + off, ob, cudaMemcpyDeviceToHost, ln.st));
        CU_TRY(ctx, cudaEventRecord(ln.done[b], ln.st));
        pending[b] = c;
    }
    for (int b = 0; b < 2; b++) {
        int rc = finish((k + b) & 1);   // oldest first
        if (rc != UPOW_OK) return rc;
    }
    return UPOW_OK;
}

// All lanes of all devices; plans[di] describes device di's share.
int run_plans(upow_ctx* ctx, const std::vector<ChunkPlan>& plans, std::vector<double>* kernel_ms) {
    const int n_dev = (int)ctx->devs.size();
    kernel_ms->assign(n_dev, 0.0);
    for (int di = 0; di < n_dev; di++)
        if (plans[di].n_chunks) {
            int rc = ensure_lanes(ctx, *ctx->devs[di]);
            if (rc != UPOW_OK) return rc;
        }
    struct Job { int di, lane, n_lanes; double ms = 0; int rc = UPOW_OK; };
    std::vector<Job> jobs;
    for (int di = 0; di < n_dev; di++) {
        const int n_lanes = (int)std::min<uint32_t>(NLANES, plans[di].n_chunks);
        for (int l = 0; l < n_lanes; l++) jobs.push_back({di, l, n_lanes});
    }
    auto work = [&](Job& j) { j.rc = run_lane(ctx, *ctx->devs[j.di], j.lane, j.n_lanes, plans[j.di], &j.ms); };
    if (jobs.size() == 1) work(jobs[0]);
    else {
        std::vector<std::thread> th;
        for (auto& j : jobs) th.emplace_back(work, std::ref(j));
        for (auto& t : th) t.join();
    }
    for (auto& j : jobs) {
        if (j.rc != UPOW_OK) return j.rc;
        (*kernel_ms)[j.di] += j.ms;
    }
    return UPOW_OK;
}

// UPOW_BULK_PLAIN=1 selects the plain-CUDA compression in the two bulk kernels (for comparison runs)
bool bulk_tuned() {
    static const bool tuned = getenv("UPOW_BULK_PLAIN") == nullptr;
    return tuned;
}
void launch_verify(const VerifyRecord* recs, uint32_t n, uint8_t* valid, uint32_t* digests, cudaStream_t st) {
    Opaque q;
    fill_opaque(&q);
    const uint32_t grid = (n + VERIFY_THREADS - 1) / VERIFY_THREADS;
    if (bulk_tuned()) verify_kernel<true><<<grid, VERIFY_THREADS, 0, st>>>(recs, n, valid, digests, q);
    else verify_kernel<false><<<grid, VERIFY_THREADS, 0, st>>>(recs, n, valid, digests, q);
}
void launch_sha_batch(const uint8_t* data, const unsigned long long* offsets, uint32_t n, uint32_t* out,
                      unsigned long long readable, cudaStream_t st) {
    Opaque q;
    fill_opaque(&q);
    if (bulk_tuned()) sha256_batch_kernel<true><<<(n + 127) / 128, 128, 0, st>>>(data, offsets, n, out, readable, q);
    else sha256_batch_kernel<false><<<(n + 127) / 128, 128, 0, st>>>(data, offsets, n, out, readable, q);
}

constexpr uint32_t VERIFY_CHUNK = 32768;   // blocks per chunk: 6 MiB of records + 32 KiB verdicts + 1 MiB digests
static_assert((size_t)VERIFY_CHUNK * (sizeof(VerifyRecord) + 1 + 32) <= LANE_BYTES, "verify chunk does not fit a lane buffer");
}  // namespace

extern "C" int upow_verify_blocks(upow_ctx* ctx, uint32_t n_blocks, const upow_block* blocks, uint8_t* out_valid,
                                  uint8_t* out_digests, upow_result* stats) {
    if (!ctx || (n_blocks && (!blocks || !out_valid))) return UPOW_E_ARG;
    try {
        std::lock_guard<std::mutex> g(ctx->api_mu);
        if (stats) memset(stats, 0, sizeof *stats);
        const double t_begin = now_s();
        const int n_dev = (int)ctx->devs.size();
        std::vector<uint32_t> cut(n_dev + 1);
        for (int i = 0; i <= n_dev; i++) cut[i] = (uint32_t)((uint64_t)n_blocks * i / n_dev);
        std::vector<ChunkPlan> plans(n_dev);
        for (int di = 0; di < n_dev; di++) {
            const uint32_t lo = cut[di], cnt = cut[di + 1] - cut[di];
            ChunkPlan& pl = plans[di];
            pl.n_chunks = (cnt + VERIFY_CHUNK - 1) / VERIFY_CHUNK;
            auto span = [lo, cnt](uint32_t c, uint32_t* first) {
                *first = lo + c * VERIFY_CHUNK;
                return std::min<uint32_t>(VERIFY_CHUNK, lo + cnt - *first);
            };
            // device / pinned buffer: [records | verdicts | digests]
            const size_t off_valid = (size_t)VERIFY_CHUNK * sizeof(VerifyRecord), off_dig = off_valid + VERIFY_CHUNK;
            pl.stage_in = [=](uint32_t c, uint8_t* pinned) {
                uint32_t first;
                const uint32_t m = span(c, &first);
                memcpy(pinned, blocks + first, (size_t)m * sizeof(VerifyRecord));
                return (size_t)m * sizeof(VerifyRecord);
            };
            pl.run = [=](uint32_t c, uint8_t* dev, cudaStream_t st) {
                uint32_t first;
                const uint32_t m = span(c, &first);
                launch_verify(reinterpret_cast<const VerifyRecord*>(dev), m, dev + off_valid,
                              out_digests ? reinterpret_cast<uint32_t*>(dev + off_dig) : nullptr, st);
            };
            pl.out_bytes = [=](uint32_t c, size_t* off) {
                uint32_t first;
                const uint32_t m = span(c, &first);
                *off = off_valid;
                return out_digests ? (size_t)VERIFY_CHUNK + 32 * (size_t)m : (size_t)m;
            };
            pl.stage_out = [=](uint32_t c, const uint8_t* pinned) {
                uint32_t first;
                const uint32_t m = span(c, &first);
                memcpy(out_valid + first, pinned, m);
                if (out_digests) memcpy(out_digests + 32 * (size_t)first, pinned + VERIFY_CHUNK, 32 * (size_t)m);
            };
        }
        std::vector<double> ms;
        int rc = run_plans(ctx, plans, &ms);
        if (rc != UPOW_OK) return rc;
        for (int di = 0; di < n_dev; di++) {
            if (stats) {
                stats->per_device_hashes[di] = cut[di + 1] - cut[di];
                stats->per_device_kernel_ms[di] = ms[di];
                stats->hashes_done += cut[di + 1] - cut[di];
                stats->kernel_ms = std::max(stats->kernel_ms, ms[di]);
                stats->launches += plans[di].n_chunks;
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

extern "C" int upow_verify_blocks_dev(upow_ctx* ctx, int device_index, uint32_t n_blocks, const upow_block* d_blocks,
                                      uint8_t* d_valid, uint8_t* d_digests, upow_result* stats) {
    if (!ctx || device_index < 0 || device_index >= (int)ctx->devs.size() || (n_blocks && (!d_blocks || !d_valid))) return UPOW_E_ARG;
    std::lock_guard<std::mutex> g(ctx->api_mu);
    if (stats) memset(stats, 0, sizeof *stats);
    if (n_blocks == 0) return UPOW_OK;
    const double t_begin = now_s();
    Device& d = *ctx->devs[device_index];
    CU_TRY(ctx, cudaSetDevice(d.id));
    CU_TRY(ctx, cudaEventRecord(d.ev_aux[0], d.stream[0]));
    launch_verify(reinterpret_cast<const VerifyRecord*>(d_blocks), n_blocks, d_valid, reinterpret_cast<uint32_t*>(d_digests),
                  d.stream[0]);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaEventRecord(d.ev_aux[1], d.stream[0]));
    CU_TRY(ctx, cudaStreamSynchronize(d.stream[0]));
    float ms = 0;
    CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev_aux[0], d.ev_aux[1]));
    if (stats) {
        stats->per_device_hashes[device_index] = stats->hashes_done = n_blocks;
        stats->per_device_kernel_ms[device_index] = stats->kernel_ms = ms;
        stats->launches = 1;
        stats->elapsed_s = now_s() - t_begin;
    }
    return UPOW_OK;
}

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
        // chunks: runs of messages whose bytes + offsets + digests fit one lane buffer.
        // buffer layout: [digests 32 m | offsets 8 (m + 1) | data ... + 16 bytes of readable padding]
        struct Chunk { uint32_t m_lo, m_hi; };
        std::vector<Chunk> chunks;
        constexpr uint32_t MAX_MSGS = 1u << 16;
        for (uint32_t i = 0; i < n;) {
            uint32_t j = i;
            size_t need = 16 + 8;
            while (j < n && j - i < MAX_MSGS) {
                const size_t add = 32 + 8 + (size_t)(offsets[j + 1] - offsets[j]);
                if (need + add > LANE_BYTES) break;
                need += add;
                j++;
            }
            if (j == i) {
                set_err(ctx, "upow_sha256_batch: message " + std::to_string(i) + " is larger than the " +
                                 std::to_string(LANE_BYTES >> 20) + " MiB staging buffer");
                return UPOW_E_ARG;
            }
            chunks.push_back({i, j});
            i = j;
        }
        // chunks are dealt to the devices in contiguous runs
        std::vector<uint32_t> cut(n_dev + 1);
        for (int i = 0; i <= n_dev; i++) cut[i] = (uint32_t)((uint64_t)chunks.size() * i / n_dev);
        std::vector<ChunkPlan> plans(n_dev);
        for (int di = 0; di < n_dev; di++) {
            ChunkPlan& pl = plans[di];
            const uint32_t c0 = cut[di];
            pl.n_chunks = cut[di + 1] - cut[di];
            const Chunk* ch = chunks.data();
            pl.stage_in = [=](uint32_t c, uint8_t* pinned) {
                const Chunk k = ch[c0 + c];
                const uint32_t m = k.m_hi - k.m_lo;
                unsigned long long* rel = reinterpret_cast<unsigned long long*>(pinned + 32 * (size_t)m);
                const uint64_t b0 = offsets[k.m_lo];
                for (uint32_t i = 0; i <= m; i++) rel[i] = offsets[k.m_lo + i] - b0;
                uint8_t* dst = pinned + 32 * (size_t)m + 8 * (size_t)(m + 1);
                const size_t bytes = (size_t)(offsets[k.m_hi] - b0);
                if (bytes) memcpy(dst, data + b0, bytes);
                memset(dst + bytes, 0, 16);
                return 32 * (size_t)m + 8 * (size_t)(m + 1) + bytes + 16;
            };
            pl.run = [=](uint32_t c, uint8_t* dev, cudaStream_t st) {
                const Chunk k = ch[c0 + c];
                const uint32_t m = k.m_hi - k.m_lo;
                launch_sha_batch(dev + 32 * (size_t)m + 8 * (size_t)(m + 1), reinterpret_cast<const unsigned long long*>(dev + 32 * (size_t)m),
                                 m, reinterpret_cast<uint32_t*>(dev), (unsigned long long)(offsets[k.m_hi] - offsets[k.m_lo]) + 16, st);
            };
            pl.out_bytes = [=](uint32_t c, size_t* off) {
                *off = 0;
                return 32 * (size_t)(ch[c0 + c].m_hi - ch[c0 + c].m_lo);
            };
            pl.stage_out = [=](uint32_t c, const uint8_t* pinned) {
                const Chunk k = ch[c0 + c];
                memcpy(out + 32 * (size_t)k.m_lo, pinned, 32 * (size_t)(k.m_hi - k.m_lo));
            };
        }
        std::vector<double> ms;
        int rc = run_plans(ctx, plans, &ms);
        if (rc != UPOW_OK) return rc;
        for (int di = 0; di < n_dev; di++) {
            if (!stats || cut[di] == cut[di + 1]) continue;
            const uint64_t msgs = chunks[cut[di + 1] - 1].m_hi - chunks[cut[di]].m_lo;
            stats->per_device_hashes[di] = msgs;
            stats->per_device_kernel_ms[di] = ms[di];
            stats->hashes_done += msgs;
            stats->kernel_ms = std::max(stats->kernel_ms, ms[di]);
            stats->launches += plans[di].n_chunks;
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

extern "C" int upow_sha256_batch_dev(upow_ctx* ctx, int device_index, uint32_t n, const uint8_t* d_data,
                                     const uint64_t* d_offsets, uint8_t* d_out, upow_result* stats) {
    if (!ctx || device_index < 0 || device_index >= (int)ctx->devs.size() || (n && (!d_data || !d_offsets || !d_out))) return UPOW_E_ARG;
    std::lock_guard<std::mutex> g(ctx->api_mu);
    if (stats) memset(stats, 0, sizeof *stats);
    if (n == 0) return UPOW_OK;
    const double t_begin = now_s();
    Device& d = *ctx->devs[device_index];
    CU_TRY(ctx, cudaSetDevice(d.id));
    CU_TRY(ctx, cudaEventRecord(d.ev_aux[0], d.stream[0]));
    launch_sha_batch(d_data, reinterpret_cast<const unsigned long long*>(d_offsets), n, reinterpret_cast<uint32_t*>(d_out), ~0ull,
                     d.stream[0]);
    CU_TRY(ctx, cudaGetLastError());
    CU_TRY(ctx, cudaEventRecord(d.ev_aux[1], d.stream[0]));
    CU_TRY(ctx, cudaStreamSynchronize(d.stream[0]));
    float ms = 0;
    CU_TRY(ctx, cudaEventElapsedTime(&ms, d.ev_aux[0], d.ev_aux[1]));
    if (stats) {
        stats->per_device_hashes[device_index] = stats->hashes_done = n;
        stats->per_device_kernel_ms[device_index] = stats->kernel_ms = ms;
        stats->launches = 1;
        stats->elapsed_s = now_s() - t_begin;
    }
    return UPOW_OK;
}

extern "C" int upow_flush_l2(upow_ctx* ctx) {
    if (!ctx) return UPOW_E_ARG;
    for (auto& dp : ctx->devs) {
        Device& d = *dp;
        CU_TRY(ctx, cudaSetDevice(d.id));
        if (!d.d_flush) CU_TRY(ctx, cudaMalloc(&d.d_flush, L2_FLUSH_BYTES));
        CU_TRY(ctx, cudaMemsetAsync(d.d_flush, 0x5a, L2_FLUSH_BYTES, d.stream[0]));
    }
    for (auto& dp : ctx->devs) {
        Device& d = *dp;
        CU_TRY(ctx, cudaSetDevice(d.id));
        CU_TRY(ctx, cudaStreamSynchronize(d.stream[0]));
    }
    return UPOW_OK;
}

extern "C" int upow_selftest(int device) {
    if (cudaSetDevice(device) != cudaSuccess) return UPOW_E_CUDA;
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

// upow_sweep_noalloc: the sweep of upow_sweep for callers that must not allocate.
//
// Who that is: a host program that has already handed all of the GPU's memory to its own arena
// (an inference server, a database, another miner), so that a cudaMalloc from a plug-in either
// fails or steals from that arena; a latency-bound caller that cannot accept the device-wide
// synchronisation cudaMalloc/cudaFree imply; code running between driver calls where creating
// streams, events or buffers is forbidden.  upow_init allocates result slots, a 16 MiB hit
// buffer, pinned mirrors, streams and events per GPU; this entry point needs none of them.
//
// Its device-side scratch -- the work-queue / hit-count block and the buffer the kernel appends
// valid nonces to -- lives in ONE file-scope static __device__ array of fixed size
// (UPOW_NOALLOC_SCRATCH_BYTES, about 256 KiB).  Every GPU that loads the library carries that array
// until the process exits, whether or not this entry point is ever called.
//
// Rules that follow from the design:
//   * capacity is fixed: at most UPOW_NOALLOC_MAX_HITS nonces per call; a larger `cap` is refused;
//   * the array is shared by all calls on a device and still holds the previous call's data, so
//     each call zeroes the control block before launching (the hit buffer needs no clearing: only
//     entries below the final count are read);
//   * two calls whose GPU work overlaps would trample each other's scratch.  Calls made through
//     this entry point are serialised per device inside the process and run on the legacy
//     default stream, synchronously; anything else that touches the array concurrently races;
//   * the FIRST call on a device looks up the array's device address (cudaGetSymbolAddress) and
//     uploads the SHA-256 round constants; make that first call before any stream capture begins.
//
// In host code the array's name is only a handle, not a device pointer: the address handed to
// the kernel is the one cudaGetSymbolAddress returns for the device that is current.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <string>

#include "../../include/upow_b200.h"
#include "hot_loop.h"
#include "sha256_core.h"
#include "template_prep.h"

// kernels.cuh also defines non-template kernels and a __constant__ table, and upow_lib.cu already
// owns one copy of those symbols.  This file takes its own, separately named copy so that the
// existing translation unit stays exactly as it is.
namespace noalloc_tu {
using namespace ::upow;
#include "kernels.cuh"
}  // namespace noalloc_tu

using namespace upow;
using noalloc_tu::upow::DevResult;
using noalloc_tu::upow::SearchArgs;
using noalloc_tu::upow::search_kernel_tuned;
using noalloc_tu::upow::MODE_ENUM;

namespace {

constexpr size_t CONTROL_BYTES = 256;  // DevResult lives here (64 B used), keeps the hit buffer 256-aligned
static_assert(sizeof(DevResult) <= CONTROL_BYTES, "control block too small");
static_assert(UPOW_NOALLOC_SCRATCH_BYTES == CONTROL_BYTES + 4ull * UPOW_NOALLOC_MAX_HITS, "header and scratch disagree");

static __device__ __align__(256) unsigned char g_noalloc_scratch[UPOW_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES_SEEN = 64;
struct PerDevice {
    std::mutex mu;
    bool ready = false;
    unsigned char* scratch = nullptr;  // device address of g_noalloc_scratch on this device
    int grid = 0;
};
PerDevice g_dev[MAX_DEVICES_SEEN];
std::mutex g_err_mu;
std::string g_err;

using NoAllocTune = Tune<1, 0, 0, 0, 0, 0, 0>;  // same build as the library's default variant

void set_error(const std::string& s) {
    std::lock_guard<std::mutex> g(g_err_mu);
    g_err = s;
}

#define NA_TRY(expr)                                                                      \
    do {                                                                                  \
        cudaError_t e_ = (expr);                                                          \
        if (e_ != cudaSuccess) {                                                          \
            set_error(std::string("upow_sweep_noalloc: " #expr ": ") + cudaGetErrorString(e_)); \
            return UPOW_E_CUDA;                                                           \
        }                                                                                 \
    } while (0)

// One-time set-up for the device that is current: device address of the scratch array, round
// constants of this translation unit, persistent grid size.  Allocates nothing.
int setup_current_device(PerDevice& pd) {
    void* p = nullptr;
    NA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    uint32_t k[64];
    for (int i = 0; i < 64; i++) k[i] = Kc(i);
    NA_TRY(cudaMemcpyToSymbol(noalloc_tu::upow::K_DEV, k, sizeof k));
    int dev = 0, sms = 0, per_sm = 0;
    NA_TRY(cudaGetDevice(&dev));
    NA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    NA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
        &per_sm, search_kernel_tuned<LAYOUT_V1, MODE_ENUM, NoAllocTune, 3>, 256, 0));
    pd.scratch = static_cast<unsigned char*>(p);
    pd.grid = sms * (per_sm < 1 ? 1 : per_sm);
    pd.ready = true;
    return UPOW_OK;
}

}  // namespace

extern "C" const char* upow_noalloc_last_error(void) {
    thread_local std::string copy;   // per-thread copy: another thread may rewrite g_err at any time
    std::lock_guard<std::mutex> g(g_err_mu);
    copy = g_err;
    return copy.c_str();
}

extern "C" int upow_sweep_noalloc(int device, const uint8_t* prefix, int prefix_len, const uint8_t* prev_hash32,
                                  int n_int, int charset_count, uint64_t nonce_lo, uint64_t nonce_hi,
                                  uint32_t* out_nonces, uint64_t cap, uint64_t* out_count) {
    if (!prefix || !prev_hash32 || !out_count || (cap && !out_nonces) || device < 0 || device >= MAX_DEVICES_SEEN) {
        set_error("upow_sweep_noalloc: NULL pointer or device index out of range");
        return UPOW_E_ARG;
    }
    *out_count = 0;
    if (cap > UPOW_NOALLOC_MAX_HITS) {
        set_error("upow_sweep_noalloc: cap exceeds UPOW_NOALLOC_MAX_HITS (" + std::to_string(UPOW_NOALLOC_MAX_HITS) +
                  "), the fixed size of the static device scratch; split the range or use upow_sweep");
        return UPOW_E_ARG;
    }
    if (nonce_hi > (1ull << 32) || nonce_lo > nonce_hi) {
        set_error("upow_sweep_noalloc: bad nonce range");
        return UPOW_E_ARG;
    }
    Tail tail;
    Rule rule;
    if (!prepare_tail(prefix, prefix_len, &tail) || !prepare_rule(prev_hash32, n_int, charset_count, tail.mid, &rule)) {
        set_error("upow_sweep_noalloc: prefix_len must be 104 or 134 and the rule must be valid");
        return UPOW_E_ARG;
    }
    int prev_device = 0;
    NA_TRY(cudaGetDevice(&prev_device));
    NA_TRY(cudaSetDevice(device));
    PerDevice& pd = g_dev[device];
    std::lock_guard<std::mutex> lock(pd.mu);  // calls on one device never overlap inside this process
    int rc = UPOW_OK;
    do {
        if (!pd.ready && (rc = setup_current_device(pd)) != UPOW_OK) break;
        DevResult* res = reinterpret_cast<DevResult*>(pd.scratch);
        uint32_t* hits = reinterpret_cast<uint32_t*>(pd.scratch + CONTROL_BYTES);

        SearchArgs a;
        memset(&a, 0, sizeof a);
        a.tail = tail;
        a.rule = rule;
        a.nonce_lo = nonce_lo;
        a.nonce_hi = nonce_hi;
        a.h_lo = (uint32_t)(nonce_lo >> 16);
        a.h_count = (uint32_t)((nonce_hi + 0xffff) >> 16) - a.h_lo;
        uint32_t hc = 64;
        while (hc > 1 && (uint64_t)((a.h_count + hc - 1) / hc) * 2048 < 4ull * pd.grid * 8) hc >>= 1;
        a.hc = hc;
        a.n_items = ((a.h_count + hc - 1) / hc) * 2048;
        a.res = res;
        a.out = hits;
        a.cap = UPOW_NOALLOC_MAX_HITS;
        fill_opaque(&a.q);
        if (tail.layout == LAYOUT_V2) prepare_v2_consts(tail, &a.v2);

        cudaError_t e;
        // the control block still holds the previous call's counters: start from zero
        if ((e = cudaMemsetAsync(res, 0, CONTROL_BYTES, 0)) != cudaSuccess) { rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break; }
        if (a.n_items) {
            if (tail.layout == LAYOUT_V1)
                search_kernel_tuned<LAYOUT_V1, MODE_ENUM, NoAllocTune, 3><<<pd.grid, 256, 0, 0>>>(a);
            else
                search_kernel_tuned<LAYOUT_V2, MODE_ENUM, NoAllocTune, 3><<<pd.grid, 256, 0, 0>>>(a);
            if ((e = cudaGetLastError()) != cudaSuccess) { rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break; }
        }
        DevResult host_res;
        if ((e = cudaMemcpy(&host_res, res, sizeof host_res, cudaMemcpyDeviceToHost)) != cudaSuccess) {
            rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break;
        }
        const uint64_t found = host_res.count;
        const uint64_t n_copy = std::min<uint64_t>(std::min<uint64_t>(found, UPOW_NOALLOC_MAX_HITS), cap);
        if (n_copy && (e = cudaMemcpy(out_nonces, hits, n_copy * sizeof(uint32_t), cudaMemcpyDeviceToHost)) != cudaSuccess) {
            rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break;
        }
        *out_count = found;
        if (found > cap) {
            // which nonces made it into the first `cap` slots depends on arrival order: report, do not pretend
            set_error("upow_sweep_noalloc: more valid nonces than cap");
            rc = UPOW_E_OVERFLOW;
            break;
        }
        std::sort(out_nonces, out_nonces + n_copy);  // in place, no heap
    } while (0);
    cudaSetDevice(prev_device);
    return rc;
}

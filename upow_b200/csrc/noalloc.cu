// upow_sweep_noalloc: the sweep of upow_sweep for callers that must not allocate.
//
// Who that is: a host program that has already handed all of the GPU's memory to its own arena
// (an inference server, a database, another miner), so that a cudaMalloc from a plug-in either
// fails or steals from that arena; a latency-bound caller that cannot accept the device-wide
// synchronisation cudaMalloc/cudaFree imply; code running between driver calls where creating
// streams, events or buffers is forbidden.  upow_init allocates result slots, a 16 MiB hit
// buffer, pinned mirrors, streams and events per GPU; this entry point needs none of them.
//
// Its device-side scratch -- the work-queue / hit-count block, the stop word and the buffer the
// kernel appends valid nonces to -- lives in ONE file-scope static __device__ array of fixed size
// (UPOW_NOALLOC_SCRATCH_BYTES, about 256 KiB).  It belongs to the library's device code image: a GPU
// holds it from the moment the CUDA runtime loads that image on it (with the default lazy loading,
// at this entry point's first call on the device) until the process exits.
//
// Rules that follow from the design:
//   * capacity is fixed: at most UPOW_NOALLOC_MAX_HITS nonces per call; a larger `cap` is refused;
//   * the array is shared by all calls on a device and still holds the previous call's data, so
//     each call zeroes the control block before launching (the hit buffer needs no clearing: only
//     entries below the final count are read);
//   * two calls whose GPU work overlaps would trample each other's scratch.  Calls made through
//     this entry point are serialised per device inside the process and each one waits for its own
//     GPU work (on the calling thread's per-thread default stream, which needs no creation and does
//     not synchronise with the host program's other streams); anything else that touches the array
//     concurrently races;
//   * every call looks up the array's device address (cudaGetSymbolAddress: a table lookup once the
//     code image is loaded, and never stale after a device reset); the FIRST call on a device also
//     sizes the grid and makes the runtime load the kernel and materialise the array, which takes
//     device memory from the driver once: make that first call before the host program claims all
//     free memory, and before any stream capture begins.
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
#include "search_kernels.cuh"

using namespace upow;

namespace {

// layout of the scratch: [0,128) DevResult, [128,136) kill word (never matches: nothing cancels this call),
// [256, ...) hit buffer
constexpr size_t CONTROL_BYTES = 256;
constexpr size_t KILL_OFFSET = 128;
static_assert(sizeof(DevResult) <= KILL_OFFSET, "control block too small");
static_assert(UPOW_NOALLOC_SCRATCH_BYTES == CONTROL_BYTES + 4ull * UPOW_NOALLOC_MAX_HITS, "header and scratch disagree");

static __device__ __align__(256) unsigned char g_noalloc_scratch[UPOW_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES_SEEN = 64;
struct PerDevice {
    std::mutex mu;
    int grid = 0;                      // persistent grid (SMs x resident blocks), sized at the first call
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

// For the device that is current: device address of the scratch array (looked up on every call) and, once,
// the persistent grid size.  Allocates nothing itself (the runtime loads the code image here if it has not yet).
int resolve_current_device(PerDevice& pd, unsigned char** scratch) {
    void* p = nullptr;
    NA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    *scratch = static_cast<unsigned char*>(p);
    if (pd.grid == 0) {
        int dev = 0, sms = 0, per_sm = 0;
        NA_TRY(cudaGetDevice(&dev));
        NA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        NA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(
            &per_sm, search_kernel_tuned<LAYOUT_V1, MODE_ENUM, NoAllocTune, 3>, 256, 0));
        pd.grid = sms * (per_sm < 1 ? 1 : per_sm);
    }
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
        unsigned char* scratch = nullptr;
        if ((rc = resolve_current_device(pd, &scratch)) != UPOW_OK) break;
        DevResult* res = reinterpret_cast<DevResult*>(scratch);
        unsigned long long* kill = reinterpret_cast<unsigned long long*>(scratch + KILL_OFFSET);
        uint32_t* hits = reinterpret_cast<uint32_t*>(scratch + CONTROL_BYTES);
        const cudaStream_t st = cudaStreamPerThread;

        SearchArgs a;
        memset(&a, 0, sizeof a);
        a.tail = tail;
        a.rule = rule;
        a.nonce_lo = nonce_lo;
        a.nonce_hi = nonce_hi;
        a.h_lo = (uint32_t)(nonce_lo >> 16);
        a.h_count = (uint32_t)((nonce_hi + 0xffff) >> 16) - a.h_lo;
        a.lg_first = 0;
        a.lg_step = 1;
        a.lg_count = 2048;
        uint32_t hc = 64;
        while (hc > 1 && (uint64_t)((a.h_count + hc - 1) / hc) * 2048 < 4ull * pd.grid * 8) hc >>= 1;
        a.hc = hc;
        a.n_items = ((a.h_count + hc - 1) / hc) * 2048;
        a.sink.res = res;
        a.sink.out = hits;
        a.sink.cap = UPOW_NOALLOC_MAX_HITS;
        a.sink.n_peers = 0;
        a.kill.word = kill;
        a.kill.serial = 1;
        fill_opaque(&a.q);
        if (tail.layout == LAYOUT_V2) prepare_v2_consts(tail, &a.v2);

        cudaError_t e;
        // the control block still holds the previous call's counters: start from zero (a zero kill word means
        // "serials 0..0", which the serial 1 of this launch never matches)
        if ((e = cudaMemsetAsync(scratch, 0, CONTROL_BYTES, st)) != cudaSuccess) { rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break; }
        if (a.n_items) {
            if (tail.layout == LAYOUT_V1)
                search_kernel_tuned<LAYOUT_V1, MODE_ENUM, NoAllocTune, 3><<<pd.grid, 256, 0, st>>>(a);
            else
                search_kernel_tuned<LAYOUT_V2, MODE_ENUM, NoAllocTune, 3><<<pd.grid, 256, 0, st>>>(a);
            if ((e = cudaGetLastError()) != cudaSuccess) { rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break; }
        }
        DevResult host_res;
        if ((e = cudaMemcpyAsync(&host_res, res, sizeof host_res, cudaMemcpyDeviceToHost, st)) != cudaSuccess ||
            (e = cudaStreamSynchronize(st)) != cudaSuccess) {
            rc = UPOW_E_CUDA; set_error(cudaGetErrorString(e)); break;
        }
        const uint64_t found = host_res.count;
        const uint64_t n_copy = std::min<uint64_t>(std::min<uint64_t>(found, UPOW_NOALLOC_MAX_HITS), cap);
        if (n_copy && ((e = cudaMemcpyAsync(out_nonces, hits, n_copy * sizeof(uint32_t), cudaMemcpyDeviceToHost, st)) != cudaSuccess ||
                       (e = cudaStreamSynchronize(st)) != cudaSuccess)) {
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

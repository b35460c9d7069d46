/*
 * upow_b200 -- C-ABI of the B200 nonce-search library (libupow_b200.so).
 *
 * The reference (upowai/upow) has no FFI or plugin interface: its hot path is the Python
 * loop in miner.py:83-98.  This header is the boundary a maintainer binds with ctypes
 * (see INTEGRATION.md); every entry point names the reference code it replaces.
 *
 * Conventions
 *   - every function returns int: UPOW_OK (0) = done / found, UPOW_EXHAUSTED (1) = range
 *     exhausted or time limit reached (the reference's `search_complete = False`,
 *     miner.py:96-98, and its OverflowError at nonce >= 2^32, miner.py:87),
 *     UPOW_CANCELLED (2), negative = error (text via upow_last_error).
 *   - the caller owns every buffer; the library returns no pointers except the opaque ctx.
 *   - nothing here calls exit()/abort() or throws across the boundary.
 *   - there is NO CPU fallback: without a CUDA device upow_init fails with UPOW_E_CUDA.
 */
#ifndef UPOW_B200_H
#define UPOW_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UPOW_MAX_DEVICES 8
#define UPOW_MAX_SLOTS 64

#define UPOW_OK 0
#define UPOW_EXHAUSTED 1
#define UPOW_CANCELLED 2
#define UPOW_E_ARG (-1)      /* bad argument, e.g. prefix_len not 104/134 */
#define UPOW_E_CUDA (-2)     /* CUDA error; text in upow_last_error */
#define UPOW_E_OVERFLOW (-3) /* sweep found more nonces than the caller's buffer holds */
#define UPOW_E_INTERNAL (-4) /* device result failed the host-side cross-check */

typedef struct upow_ctx upow_ctx;

/* Search limits.  Zero-initialise for the reference's behaviour (one pass over [start, 2^32)). */
typedef struct upow_limits {
    uint64_t max_nonce;      /* exclusive end of the nonce space per timestamp; 0 means 2^32 */
    uint32_t max_rollovers;  /* how many times "next timestamp, nonce = start" may be applied when the
                                nonce space is exhausted (the reference instead dies with OverflowError
                                and is restarted with a fresh timestamp, miner.py:87,149-156) */
    uint32_t rollover_backwards; /* 0: next timestamp = timestamp + 1 -- the order of the synthetic benchmark
                                (SURVEY 8(d) cfg 4); blocks found that way carry a timestamp AHEAD of the wall
                                clock and a node rejects them as "timestamp in the future" until the clock has
                                caught up (upow/manager.py:456-460).  1: next timestamp = timestamp - 1, what a
                                live miner needs: the node accepts last_block.timestamp < ts <= now
                                (upow/manager.py:445-460), so start at now and allow at most
                                now - last_block.timestamp - 1 rollovers */
    double time_limit_s;     /* <= 0: none.  The reference gives up after 90 s (miner.py:96-98) */
    uint64_t tile_nonces;    /* work granule handed to one GPU; 0 = library default */
} upow_limits;

typedef struct upow_result {
    int32_t found;           /* 1 when nonce/timestamp/digest are valid */
    uint32_t rollovers;      /* timestamp increments applied to reach the result */
    uint32_t timestamp;      /* timestamp of the winning candidate */
    uint32_t nonce;          /* lowest valid nonce in search order */
    uint8_t digest[32];      /* sha256(prefix(timestamp) || nonce_le32), computed on the device */
    uint64_t hashes_done;    /* hashes actually evaluated, all devices */
    double elapsed_s;        /* host wall-clock of the call */
    double kernel_ms;        /* device time of the search kernels (CUDA events), max over devices */
    uint64_t per_device_hashes[UPOW_MAX_DEVICES];
    double per_device_kernel_ms[UPOW_MAX_DEVICES];
    double sm_clock_mhz;     /* SM clock observed by the kernels (clock64 / globaltimer), device 0 */
    uint32_t launches;       /* search-kernel launches issued */
    uint32_t reserved0;
} upow_result;

typedef struct upow_device_info {
    char name[64];
    int32_t device_id;
    int32_t sm_count;
    int32_t cc_major, cc_minor;
    int32_t max_sm_clock_khz;
    int32_t search_grid_blocks;   /* persistent grid used by the search kernels */
    uint64_t total_mem;
    int32_t peer_devices;         /* how many of the context's other GPUs this one reaches by peer access (finds
                                     travel GPU to GPU; without it the host relays them) */
    int32_t reserved0;
} upow_device_info;

typedef struct upow_microbench_result {
    char name[48];
    int32_t sm_count;
    int32_t reserved0;
    double kernel_ms;
    double max_block_cycles, mean_block_cycles;
    double warp_inst_per_clk_per_sm;  /* issued warp instructions per SM clock */
    double lanes_per_clk_per_sm;      /* = 32 x the above */
    double fma_fraction;              /* share of the ops that run on the fma pipe */
    double sm_clock_mhz;              /* clock64 / globaltimer during the run */
} upow_microbench_result;

/* One template of a batch (upow_search_batch). */
typedef struct upow_template {
    uint8_t prefix[136];     /* block_prefix bytes (miner.py:74-82), 134 or 104 used */
    uint32_t prefix_len;
    uint8_t prev_hash[32];   /* bytes.fromhex(last_block["hash"]) */
    int32_t n_int;           /* int(difficulty) */
    int32_t charset_count;   /* ceil(16*(1-difficulty%1)), 16 when the fraction is 0 */
} upow_template;

typedef struct upow_batch_item {
    int32_t found;
    uint32_t nonce;
    uint8_t digest[32];
} upow_batch_item;

/* One finished block for upow_verify_blocks (192 bytes). */
typedef struct upow_block {
    uint8_t content[140];    /* bytes.fromhex(block_content): 138 or 108 used (upow/manager.py:401-419) */
    uint32_t content_len;
    uint8_t prev_hash[32];   /* bytes.fromhex(last_block["hash"]) the block must chain to */
    int32_t n_int;           /* floor(difficulty) */
    int32_t charset_count;   /* ceil(16*(1-difficulty%1)), 16 when the fraction is 0 */
    uint32_t reserved[2];
} upow_block;

/* Creates a context driving n_devices GPUs (device_ids == NULL: devices 0..n-1; n_devices == 0:
 * every visible device, at most UPOW_MAX_DEVICES).  Replaces the worker-pool set-up of
 * miner.py:139-148 (one OS process per worker): one calling thread queues launches on every GPU, the GPUs
 * share each nonce space (GPU g of N takes the nonces whose bits 5..15 are congruent to g mod N) and tell
 * each other about finds through peer memory where peer access exists. */
int upow_init(int n_devices, const int* device_ids, upow_ctx** out_ctx);
int upow_destroy(upow_ctx* ctx);
int upow_device_count(const upow_ctx* ctx);
int upow_get_device_info(const upow_ctx* ctx, int index, upow_device_info* out);

/* Text of the last error on this context (or of the last failed upow_init when ctx == NULL). */
const char* upow_last_error(const upow_ctx* ctx);

/* Registers a block template in `slot`.
 *   prefix / prefix_len : block_prefix exactly as miner.py:74-82 builds it (134 or 104 bytes);
 *                         the library copies it and derives midstate + tail words (the part of
 *                         hashlib.sha256 at miner.py:52,61 that does not depend on the nonce).
 *   prev_hash32, n_int, charset_count : the rule of miner.py:43-49 -- the digest must start with
 *                         the last n_int hex chars of prev_hash (n_int == 0: all 64, Python's
 *                         s[-0:]), and hex char number n_int must be < charset_count. */
int upow_set_template(upow_ctx* ctx, int slot, const uint8_t* prefix, int prefix_len, const uint8_t* prev_hash32,
                      int n_int, int charset_count);

/* The hot loop of miner.py:83-98: first valid nonce of start_nonce, start_nonce+stride, ...
 * below limits->max_nonce, i.e. the MINIMUM valid nonce of that progression, exact on any
 * number of GPUs.  With rollover enabled the search order is lexicographic (timestamp, nonce).
 * limits may be NULL. */
int upow_search_min(upow_ctx* ctx, int slot, uint64_t start_nonce, uint64_t stride, const upow_limits* limits,
                    upow_result* result);

/* Every valid nonce in [nonce_lo, nonce_hi) (what the miner.py:86-89 predicate accepts), sorted
 * ascending into out_nonces[0..min(count,cap)); *out_count is the true count.  stats may be NULL. */
int upow_sweep(upow_ctx* ctx, int slot, uint64_t nonce_lo, uint64_t nonce_hi, uint32_t* out_nonces, uint64_t cap,
               uint64_t* out_count, upow_result* stats);

/* n_templates independent searches (minimum valid nonce in [0, max_nonce) of each template); every GPU
 * holds all per-template midstates and searches its share of every template's nonces.  Corresponds to
 * running miner.py:83-98 once per template.  stats may be NULL. */
int upow_search_batch(upow_ctx* ctx, uint32_t n_templates, const upow_template* templates, uint64_t max_nonce,
                      upow_batch_item* results, upow_result* stats);

/* The first success over many templates: every template advances through the nonce space in rounds
 * (nonce ranges that are the same for all templates and do not depend on the number of GPUs); the call returns after the earliest round in which any template has a valid nonce, with
 * the lowest template index among that round's finds (so the answer does not depend on the number
 * of GPUs) and that template's minimum nonce.  This is what a miner wants when one template's 2^32
 * nonces per second are not enough work for a B200: vary timestamp / transaction subset / address
 * (the freedom the node leaves, upow/manager.py:365-378,445-460, upow/node/main.py:580-600) and
 * race the templates.  UPOW_EXHAUSTED when no template has a valid nonce below max_nonce. */
int upow_search_any(upow_ctx* ctx, uint32_t n_templates, const upow_template* templates, uint64_t max_nonce,
                    uint32_t* out_index, upow_batch_item* out_item, upow_result* stats);

/* Batch form of the node's check_block_is_valid (upow/manager.py:130-151): sha256(content) and the
 * difficulty rule against prev_hash, for n blocks at once (chain sync verifies them one by one,
 * upow/node/main.py:97-150).  out_valid[i] = 1 valid, 0 not valid, 0xff content_len is neither 138
 * nor 108; out_digests (may be NULL) receives 32 bytes per block.  Blocks are split over the GPUs. */
int upow_verify_blocks(upow_ctx* ctx, uint32_t n_blocks, const upow_block* blocks, uint8_t* out_valid,
                       uint8_t* out_digests, upow_result* stats);
/* The same for records that already live in the memory of the context's GPU number device_index: all three
 * pointers are DEVICE pointers there (d_digests may be NULL), nothing is copied. */
int upow_verify_blocks_dev(upow_ctx* ctx, int device_index, uint32_t n_blocks, const upow_block* d_blocks,
                           uint8_t* d_valid, uint8_t* d_digests, upow_result* stats);

/* SHA-256 of n variable-length messages: message i is data[offsets[i] .. offsets[i+1]) (offsets has
 * n+1 entries, non-decreasing); out receives 32 bytes per message.  This is the per-transaction hashing
 * of upow/node/main.py:692 and upow/manager.py:365-378 (merkle tree) and Transaction.hash
 * (upow/upow_transactions/transaction.py:85-88) in one call.  Messages are split over the GPUs and staged
 * through pinned buffers in chunks; a single message must fit one 8 MiB staging buffer (UPOW_E_ARG otherwise;
 * the _dev variant has no such limit). */
int upow_sha256_batch(upow_ctx* ctx, uint32_t n, const uint8_t* data, const uint64_t* offsets, uint8_t* out,
                      upow_result* stats);
/* The same for data that already lives on the context's GPU number device_index: d_data, d_offsets and
 * d_out are DEVICE pointers there; d_data must be 4-byte aligned and readable 8 bytes past offsets[n]. */
int upow_sha256_batch_dev(upow_ctx* ctx, int device_index, uint32_t n, const uint8_t* d_data, const uint64_t* d_offsets,
                          uint8_t* d_out, upow_result* stats);

/* The library carries several builds of the search kernel that differ only in how the integer
 * work is spread over the SM's alu and fma issue pipes (DESIGN.md "pipe balance"); results are
 * identical.  Variant 0 is the plain-CUDA kernel.  The default is the fastest one measured on
 * B200; env UPOW_VARIANT overrides it at upow_init. */
int upow_variant_count(void);
int upow_default_variant(void);
const char* upow_variant_name(int variant);
int upow_get_variant(const upow_ctx* ctx);
int upow_set_variant(upow_ctx* ctx, int variant);

/* No counterpart in the reference.  The sweep of upow_sweep for callers that must not allocate:
 * a host whose own arena already owns all GPU memory, a caller that cannot take the device-wide
 * stall of cudaMalloc/cudaFree, code where creating streams, events or buffers is off limits.
 * Needs no upow_ctx.  The device scratch (work queue, hit counter, stop word, hit buffer) is a fixed
 * static array of UPOW_NOALLOC_SCRATCH_BYTES inside the library's device code image; a GPU holds it
 * from the moment the CUDA runtime loads that image there (with lazy loading: this entry point's first
 * call on the device) until the process exits, so:
 *   - cap may not exceed UPOW_NOALLOC_MAX_HITS (UPOW_E_ARG otherwise);
 *   - all calls on a device share the array: the library serialises its own calls per device, runs
 *     each on the calling thread's per-thread default stream (no stream is created, the host's other
 *     streams are not synchronised) and waits for it, and clears the control block at the start of each
 *     call; GPU work from elsewhere that overlaps a call on the same array races;
 *   - every call looks the array's device address up again (never stale after a device reset); the
 *     FIRST call on a device also makes the runtime load the kernel and materialise the array, which
 *     takes device memory once: issue it before the host claims all free memory and before any stream
 *     capture starts.
 * Calls no cudaMalloc/cudaFree, creates no stream or event.  Output: valid nonces of
 * [nonce_lo, nonce_hi), sorted ascending; *out_count is the true count; UPOW_E_OVERFLOW if it exceeds
 * cap.  Error text: upow_noalloc_last_error(). */
#define UPOW_NOALLOC_MAX_HITS 65536u
#define UPOW_NOALLOC_SCRATCH_BYTES (256u + 4u * UPOW_NOALLOC_MAX_HITS)
int upow_sweep_noalloc(int device, const uint8_t* prefix, int prefix_len, const uint8_t* prev_hash32, int n_int,
                       int charset_count, uint64_t nonce_lo, uint64_t nonce_hi, uint32_t* out_nonces, uint64_t cap,
                       uint64_t* out_count);
const char* upow_noalloc_last_error(void);

/* Stops a running search from another thread or a signal handler (only sets a flag; the
 * reference's equivalent is the parent terminating its workers, miner.py:155-156).  The flag is STICKY: a
 * cancel that arrives while no search is running is not lost -- the next search / sweep / batch call on the
 * context returns UPOW_CANCELLED at once.  The call that reports UPOW_CANCELLED consumes the flag;
 * upow_clear_cancel drops a pending one (call it when a fresh template has been fetched). */
int upow_cancel(upow_ctx* ctx);
int upow_clear_cancel(upow_ctx* ctx);

/* Hashes evaluated so far by the running (or the last) upow_search_min call, from any thread; it advances
 * whenever a launch retires.  This is what the reference's progress line prints every 5 000 000 nonces
 * (miner.py:89-95). */
int upow_get_progress(const upow_ctx* ctx, uint64_t* hashes_done);

/* SHA-256 known-answer tests through the device code paths on `device`. 0 = all passed. */
int upow_selftest(int device);

/* Issue-rate microbenchmarks used for the roofline denominators (bench.py). */
int upow_microbench_count(void);
int upow_microbench(int device, int op, uint32_t iters, upow_microbench_result* out);

/* Writes `bytes` of device memory (> L2) so the next timed step starts with a cold L2. */
int upow_flush_l2(upow_ctx* ctx);

const char* upow_version(void);

#ifdef __cplusplus
}
#endif
#endif /* UPOW_B200_H */

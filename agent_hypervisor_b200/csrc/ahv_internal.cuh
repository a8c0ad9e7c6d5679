// ahv_internal.cuh - context, error plumbing and launch helpers shared by the .cu files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/ahv.h"

#define AHV_MAX_TIMERS 64
#define AHV_PIPE_EVENTS 72       // 64 for the chunks of ahv_audit_sessions + 8 for ahv_upload_fence
#define AHV_STATUS_PACK_OVERFLOW 2u // ahv_pack_deltas_dev: the pre-images need more bytes than the arena the caller gave
#define AHV_STATUS_MAX_LEAVES 1u   // a _dev tree entry was given a max_session_leaves smaller than a session really has

struct ahv_ctx {
    int device;
    cudaStream_t stream;
    cudaStream_t copy_stream;     // H2D of the pipelined host entry point
    cudaStream_t d2h_stream;      // D2H of the pipelined host entry point (PCIe is full duplex)
    cudaEvent_t ev_pipe[AHV_PIPE_EVENTS];   // chunk hand-off events of the pipelined host entry (copy-in done / hashed)
    int sm_count;
    int clock_khz;
    uint64_t total_mem;
    // device scratch: one growable slab, bump-allocated inside each _dev entry point
    uint8_t *scratch;
    uint64_t scratch_cap;
    uint64_t scratch_used;
    // device staging for the host-pointer entry points (inputs/outputs of the _dev call)
    uint8_t *io;
    uint64_t io_cap;
    uint64_t io_used;
    uint64_t launches;
    uint32_t upload_fences;    // ahv_upload_fence calls so far (picks one of eight events)
    cudaEvent_t t_start[AHV_MAX_TIMERS], t_stop[AHV_MAX_TIMERS];
    // tuning options (ahv_set_option)
    int64_t opt_sort;          // 1: length-sort variable-length batches (default 1)
    int64_t opt_block;         // threads per block for SHA kernels (default 128)
    int64_t opt_fixed_variant; // 0: direct LDG.128, 1: smem-staged (cp.async)
    int64_t opt_blocking_sync; // 1: ahv_audit_sessions sleeps on an event instead of spinning while the pipeline drains (default 0)
    int64_t opt_sigma_rows;    // sigma_eff, tiled kernel: rows per tile (power of two, 1..128); 0 = from the average row length
    uint64_t sigma_edges_hint; // set by the host entry around its call of the _dev entry: number of vouch records (0 = unknown)
    int64_t opt_chain_prefix;  // chained capture: hash the blocks in front of the parent field in a parallel pre-pass: 0 never, 1 up to 2^16 sessions (default), 2 always
    int64_t opt_slash_slice;   // k_slash_lanes: bytes of shared memory per warp; 0 = chosen per batch (default)
    int64_t opt_slash_variant; // 0: sessions that fit are staged in shared memory (default), 1: always global memory
    int64_t opt_sigma_variant; // 1: warp-tiled sigma_eff kernel (default), 0: one row per thread straight from global memory
    int64_t opt_sigma_exact;   // 1: every row summed in the reference's order (bit-identical); 0 (default): rows longer than opt_sigma_long_cut take the fixed-shape parallel sum
    int64_t opt_sigma_long_cut;// row length (edges) above which the parallel sum is used (default 512)
    int64_t opt_sigma_heavy;   // tiles with more edges than this go through the shared queue (default 3072)
    int64_t opt_sigma_share;   // 1 (default): heavy tiles of the tiled sigma_eff kernel go through a queue shared by the CTAs
    int64_t opt_sigma_chunk;   // edges per shared-memory chunk of the tiled sigma_eff kernel: 2048 (default) or 1024
    int64_t opt_chain_lanes;   // sessions per warp in the chained-capture kernel (0 = 32)
    int64_t opt_compress;      // 1: scaled-sum compression (adds on the FMA pipe, default), 0: plain additions
    int64_t opt_top_max;       // level width at which the in-CTA tree kernel takes over (0 = auto from the session count, else 2..256; <= 64 uses the warp-per-session finish)
    int64_t opt_peak_ctas;     // CTAs of 256 threads per SM in the ahv_int_peak kernels (default 4)
    int64_t opt_leaf_timer;    // >0: bracket every leaf-hash launch with timer slots base..base+15 (round robin)
    uint64_t leaf_launches;    // leaf-hash kernel launches so far (indexes the round robin)
    uint8_t *sigma_work;       // queue + claim flags of the tiled sigma_eff kernel (ahv_liability.cu)
    uint64_t sigma_work_tiles;
    uint32_t sigma_tick_epoch; // number of the last streaming sigma_eff launch (which of its two tile counters is live)
    uint32_t sigma_epoch;      // number of the last tiled sigma_eff launch (work-queue generation)
    void *comm;                // ncclComm_t of this rank (ahv_comm_init), NULL in a world of one
    int comm_world, comm_rank;
    uint32_t *status_h, *status_d; // sticky device-side error flags (mapped pinned host word and its device alias); ahv_sync reports them
    uint8_t *one_out;          // 64 B of mapped pinned host memory: where ahv_sha256_one's kernel writes its digest
};

namespace ahv {

void set_error(const char *fmt, ...);

#define AHV_CUDA(call)                                                                          \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess) {                                                                \
            ahv::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return (e_ == cudaErrorMemoryAllocation) ? AHV_E_NOMEM : AHV_E_CUDA;                \
        }                                                                                       \
    } while (0)

#define AHV_CHECK_LAUNCH(ctx)                                                                   \
    do {                                                                                        \
        (ctx)->launches++;                                                                      \
        cudaError_t e_ = cudaGetLastError();                                                    \
        if (e_ != cudaSuccess) {                                                                \
            ahv::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(e_), __FILE__, __LINE__); \
            return AHV_E_CUDA;                                                                  \
        }                                                                                       \
    } while (0)

#define AHV_REQUIRE(cond, msg)                                                                  \
    do {                                                                                        \
        if (!(cond)) {                                                                          \
            ahv::set_error("invalid argument: %s", msg);                                        \
            return AHV_E_INVALID;                                                               \
        }                                                                                       \
    } while (0)

#define AHV_TRY(expr)                                                                           \
    do {                                                                                        \
        int rc_ = (expr);                                                                       \
        if (rc_ != AHV_OK) return rc_;                                                          \
    } while (0)

// scratch slab --------------------------------------------------------------
int scratch_reserve(ahv_ctx *ctx, uint64_t bytes);   // grow (may sync); resets the bump pointer
void *scratch_take(ahv_ctx *ctx, uint64_t bytes);    // 256-byte aligned; NULL if reserve was too small
int io_reserve(ahv_ctx *ctx, uint64_t bytes);        // same, for the host entry points' staging slab
void *io_take(ahv_ctx *ctx, uint64_t bytes);
static inline uint64_t pad256(uint64_t b) { return (b + 255u) & ~(uint64_t)255u; }

int use_device(ahv_ctx *ctx);
int check_status(ahv_ctx *ctx);                       // sticky device-side error flags -> error code (clears them)
void comm_release(ahv_ctx *ctx);                     // ahv_comm.cu: destroys the communicator, if any
// ahv_slashing.cu: the slashing pass on device arrays; n_records = rowptr[n_agents] when the caller knows it, else UINT64_MAX
int slash_cascade_launch(ahv_ctx *ctx, const struct ahv_slash_batch *b, uint64_t n_records);

static inline unsigned grid_for(uint64_t threads, unsigned block)
{
    uint64_t g = (threads + block - 1) / block;
    return (unsigned)(g ? g : 1);
}

}  // namespace ahv

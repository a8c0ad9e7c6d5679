// ahv_ctx.cu - context lifetime, memory helpers, timers, error text.
#include <stdarg.h>

#include "ahv_internal.cuh"

namespace ahv {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
}

int use_device(ahv_ctx *ctx)
{
    if (!ctx) {
        set_error("null context");
        return AHV_E_INVALID;
    }
    AHV_CUDA(cudaSetDevice(ctx->device));
    return AHV_OK;
}

int scratch_reserve(ahv_ctx *ctx, uint64_t bytes)
{
    bytes = pad256(bytes) + 256;
    if (bytes > ctx->scratch_cap) {
        AHV_CUDA(cudaStreamSynchronize(ctx->stream));
        AHV_CUDA(cudaStreamSynchronize(ctx->copy_stream));
        if (ctx->scratch) AHV_CUDA(cudaFree(ctx->scratch));
        ctx->scratch = nullptr;
        ctx->scratch_cap = 0;
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("device scratch allocation of %llu bytes failed: %s", (unsigned long long)bytes,
                      cudaGetErrorString(e));
            return AHV_E_NOMEM;
        }
        ctx->scratch = (uint8_t *)p;
        ctx->scratch_cap = bytes;
    }
    ctx->scratch_used = 0;
    return AHV_OK;
}

void *scratch_take(ahv_ctx *ctx, uint64_t bytes)
{
    uint64_t b = pad256(bytes ? bytes : 1);
    if (ctx->scratch_used + b > ctx->scratch_cap) return nullptr;
    void *p = ctx->scratch + ctx->scratch_used;
    ctx->scratch_used += b;
    return p;
}

int io_reserve(ahv_ctx *ctx, uint64_t bytes)
{
    bytes = pad256(bytes) + 256;
    if (bytes > ctx->io_cap) {
        AHV_CUDA(cudaStreamSynchronize(ctx->stream));
        AHV_CUDA(cudaStreamSynchronize(ctx->copy_stream));
        if (ctx->io) AHV_CUDA(cudaFree(ctx->io));
        ctx->io = nullptr;
        ctx->io_cap = 0;
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("device staging allocation of %llu bytes failed: %s", (unsigned long long)bytes,
                      cudaGetErrorString(e));
            return AHV_E_NOMEM;
        }
        ctx->io = (uint8_t *)p;
        ctx->io_cap = bytes;
    }
    ctx->io_used = 0;
    return AHV_OK;
}

// Device-side sanity flags raised by kernels since the last check (sticky until reported once).
int check_status(ahv_ctx *ctx)
{
    const uint32_t st = __atomic_exchange_n(ctx->status_h, 0u, __ATOMIC_ACQ_REL);
    if (st & AHV_STATUS_MAX_LEAVES) {
        set_error("max_session_leaves was smaller than a session's real leaf count: the affected sessions' roots were "
                  "written as zeros; pass the true maximum (or 0 to let the library find it)");
        return AHV_E_INVALID;
    }
    if (st & AHV_STATUS_PACK_OVERFLOW) {
        set_error("ahv_pack_deltas_dev: the pre-images need more bytes than arena_capacity; nothing was written");
        return AHV_E_TOOLARGE;
    }
    return AHV_OK;
}

void *io_take(ahv_ctx *ctx, uint64_t bytes)
{
    uint64_t b = pad256(bytes ? bytes : 1);
    if (ctx->io_used + b > ctx->io_cap) return nullptr;
    void *p = ctx->io + ctx->io_used;
    ctx->io_used += b;
    return p;
}

}  // namespace ahv

using namespace ahv;

extern "C" {

int ahv_abi_version(void) { return AHV_ABI_VERSION; }

const char *ahv_last_error(void) { return g_err; }

int ahv_create(int device, ahv_ctx **out_ctx)
{
    if (!out_ctx) {
        set_error("out_ctx is null");
        return AHV_E_INVALID;
    }
    *out_ctx = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device available (%s); this library has no CPU path",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return AHV_E_NODEVICE;
    }
    if (device < 0 || device >= ndev) {
        set_error("device ordinal %d out of range (0..%d)", device, ndev - 1);
        return AHV_E_INVALID;
    }
    AHV_CUDA(cudaSetDevice(device));
    ahv_ctx *c = new ahv_ctx();
    memset(c, 0, sizeof *c);
    c->device = device;
    cudaDeviceProp prop;
    AHV_CUDA(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    c->clock_khz = prop.clockRate;
    c->total_mem = prop.totalGlobalMem;
    AHV_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    AHV_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    AHV_CUDA(cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
    for (int i = 0; i < AHV_PIPE_EVENTS; ++i) AHV_CUDA(cudaEventCreateWithFlags(&c->ev_pipe[i], cudaEventDisableTiming));
    for (int i = 0; i < AHV_MAX_TIMERS; ++i) {
        AHV_CUDA(cudaEventCreate(&c->t_start[i]));
        AHV_CUDA(cudaEventCreate(&c->t_stop[i]));
    }
    {
        void *h = nullptr, *dv = nullptr;
        AHV_CUDA(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
        memset(h, 0, 64);
        AHV_CUDA(cudaHostGetDevicePointer(&dv, h, 0));
        c->status_h = (uint32_t *)h;
        c->status_d = (uint32_t *)dv;
    }
    c->comm_world = 1;
    c->opt_sort = 1;
    c->opt_block = 128;
    c->opt_fixed_variant = 0;
    c->opt_top_max = 0;
    c->opt_compress = 1;
    c->opt_sigma_variant = 1;
    c->opt_sigma_exact = 0;
    c->opt_sigma_long_cut = 64;
    c->opt_sigma_chunk = 1440;
    c->opt_sigma_share = 1;
    c->opt_sigma_heavy = 3072;
    c->opt_slash_variant = 0;
    c->opt_slash_slice = 0;
    c->opt_chain_prefix = 1;
    c->opt_sigma_rows = 0;
    c->sigma_edges_hint = 0;
    c->opt_blocking_sync = 0;
    *out_ctx = c;
    return AHV_OK;
}

void ahv_destroy(ahv_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->copy_stream);
    cudaStreamSynchronize(ctx->d2h_stream);
    comm_release(ctx);
    if (ctx->scratch) cudaFree(ctx->scratch);
    if (ctx->io) cudaFree(ctx->io);
    if (ctx->sigma_work) cudaFree(ctx->sigma_work);
    if (ctx->one_out) cudaFreeHost(ctx->one_out);
    if (ctx->status_h) cudaFreeHost(ctx->status_h);
    for (int i = 0; i < AHV_PIPE_EVENTS; ++i) cudaEventDestroy(ctx->ev_pipe[i]);
    for (int i = 0; i < AHV_MAX_TIMERS; ++i) {
        cudaEventDestroy(ctx->t_start[i]);
        cudaEventDestroy(ctx->t_stop[i]);
    }
    cudaStreamDestroy(ctx->stream);
    cudaStreamDestroy(ctx->copy_stream);
    cudaStreamDestroy(ctx->d2h_stream);
    delete ctx;
}

int ahv_sync(ahv_ctx *ctx)
{
    AHV_TRY(use_device(ctx));
    AHV_CUDA(cudaStreamSynchronize(ctx->copy_stream));
    AHV_CUDA(cudaStreamSynchronize(ctx->stream));
    AHV_CUDA(cudaStreamSynchronize(ctx->d2h_stream));
    return check_status(ctx);
}

int ahv_device_info(ahv_ctx *ctx, int *sm_count, int *sm_clock_khz, uint64_t *total_mem)
{
    AHV_TRY(use_device(ctx));
    if (sm_count) *sm_count = ctx->sm_count;
    if (sm_clock_khz) *sm_clock_khz = ctx->clock_khz;
    if (total_mem) *total_mem = ctx->total_mem;
    return AHV_OK;
}

void *ahv_stream(ahv_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }

uint64_t ahv_launch_count(ahv_ctx *ctx) { return ctx ? ctx->launches : 0; }

int ahv_dev_alloc(ahv_ctx *ctx, uint64_t bytes, void **out_dptr)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(out_dptr, "out_dptr is null");
    *out_dptr = nullptr;
    cudaError_t e = cudaMalloc(out_dptr, bytes ? bytes : 1);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("cudaMalloc(%llu) failed: %s", (unsigned long long)bytes, cudaGetErrorString(e));
        return AHV_E_NOMEM;
    }
    return AHV_OK;
}

int ahv_dev_free(ahv_ctx *ctx, void *dptr)
{
    AHV_TRY(use_device(ctx));
    if (dptr) AHV_CUDA(cudaFree(dptr));
    return AHV_OK;
}

int ahv_host_alloc(ahv_ctx *ctx, uint64_t bytes, void **out_hptr)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(out_hptr, "out_hptr is null");
    *out_hptr = nullptr;
    cudaError_t e = cudaHostAlloc(out_hptr, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("cudaHostAlloc(%llu) failed: %s", (unsigned long long)bytes, cudaGetErrorString(e));
        return AHV_E_NOMEM;
    }
    return AHV_OK;
}

int ahv_host_alloc_flags(ahv_ctx *ctx, uint64_t bytes, uint32_t flags, void **out_hptr)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(out_hptr, "out_hptr is null");
    AHV_REQUIRE((flags & ~(uint32_t)AHV_HOST_WRITE_COMBINED) == 0, "unknown host allocation flag");
    *out_hptr = nullptr;
    const unsigned f = (flags & AHV_HOST_WRITE_COMBINED) ? cudaHostAllocWriteCombined : cudaHostAllocDefault;
    cudaError_t e = cudaHostAlloc(out_hptr, bytes ? bytes : 1, f);
    if (e != cudaSuccess) {
        cudaGetLastError();
        set_error("cudaHostAlloc(%llu, flags %u) failed: %s", (unsigned long long)bytes, flags, cudaGetErrorString(e));
        return AHV_E_NOMEM;
    }
    return AHV_OK;
}

int ahv_host_free(ahv_ctx *ctx, void *hptr)
{
    AHV_TRY(use_device(ctx));
    if (hptr) AHV_CUDA(cudaFreeHost(hptr));
    return AHV_OK;
}

int ahv_memcpy_h2d(ahv_ctx *ctx, void *dptr, const void *hptr, uint64_t bytes)
{
    AHV_TRY(use_device(ctx));
    if (bytes) AHV_CUDA(cudaMemcpyAsync(dptr, hptr, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return AHV_OK;
}

int ahv_upload_async(ahv_ctx *ctx, void *dptr, const void *hptr, uint64_t bytes)
{
    AHV_TRY(use_device(ctx));
    if (bytes) AHV_CUDA(cudaMemcpyAsync(dptr, hptr, bytes, cudaMemcpyHostToDevice, ctx->copy_stream));
    return AHV_OK;
}

int ahv_upload_fence(ahv_ctx *ctx)
{
    AHV_TRY(use_device(ctx));
    cudaEvent_t ev = ctx->ev_pipe[AHV_PIPE_EVENTS - 1 - (ctx->upload_fences++ & 7u)];     // (ahv_audit_sessions uses the low slots)
    AHV_CUDA(cudaEventRecord(ev, ctx->copy_stream));
    AHV_CUDA(cudaStreamWaitEvent(ctx->stream, ev, 0));
    return AHV_OK;
}

int ahv_memcpy_d2h(ahv_ctx *ctx, void *hptr, const void *dptr, uint64_t bytes)
{
    AHV_TRY(use_device(ctx));
    if (bytes) AHV_CUDA(cudaMemcpyAsync(hptr, dptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    return AHV_OK;
}

int ahv_memset_dev(ahv_ctx *ctx, void *dptr, int value, uint64_t bytes)
{
    AHV_TRY(use_device(ctx));
    if (bytes) AHV_CUDA(cudaMemsetAsync(dptr, value, bytes, ctx->stream));
    return AHV_OK;
}

int ahv_timer_start(ahv_ctx *ctx, int slot)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(slot >= 0 && slot < AHV_MAX_TIMERS, "timer slot out of range");
    AHV_CUDA(cudaEventRecord(ctx->t_start[slot], ctx->stream));
    return AHV_OK;
}

int ahv_timer_stop(ahv_ctx *ctx, int slot)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(slot >= 0 && slot < AHV_MAX_TIMERS, "timer slot out of range");
    AHV_CUDA(cudaEventRecord(ctx->t_stop[slot], ctx->stream));
    return AHV_OK;
}

int ahv_timer_elapsed_ms(ahv_ctx *ctx, int slot, float *out_ms)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(slot >= 0 && slot < AHV_MAX_TIMERS && out_ms, "bad timer arguments");
    AHV_CUDA(cudaEventSynchronize(ctx->t_stop[slot]));
    AHV_CUDA(cudaEventElapsedTime(out_ms, ctx->t_start[slot], ctx->t_stop[slot]));
    return AHV_OK;
}

int ahv_timer_wait(ahv_ctx *ctx, int slot)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(slot >= 0 && slot < AHV_MAX_TIMERS, "timer slot out of range");
    AHV_CUDA(cudaEventSynchronize(ctx->t_stop[slot]));
    return check_status(ctx);
}

int ahv_set_option(ahv_ctx *ctx, const char *key, int64_t value)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(key, "key is null");
    if (!strcmp(key, "sort")) ctx->opt_sort = value;
    else if (!strcmp(key, "block")) {
        AHV_REQUIRE(value == 64 || value == 128 || value == 256, "block must be 64/128/256");
        ctx->opt_block = value;
    } else if (!strcmp(key, "fixed_variant")) ctx->opt_fixed_variant = value;
    else if (!strcmp(key, "blocking_sync")) ctx->opt_blocking_sync = value != 0;
    else if (!strcmp(key, "slash_variant")) {
        AHV_REQUIRE(value >= 0 && value <= 2, "slash_variant must be 0, 1 or 2");
        ctx->opt_slash_variant = value;
    } else if (!strcmp(key, "sigma_rows")) {
        AHV_REQUIRE(value == 0 || (value >= 1 && value <= 128 && (value & (value - 1)) == 0), "sigma_rows must be 0 or a power of two up to 128");
        ctx->opt_sigma_rows = value;
    } else if (!strcmp(key, "chain_prefix")) {
        AHV_REQUIRE(value >= 0 && value <= 2, "chain_prefix must be 0, 1 or 2");
        ctx->opt_chain_prefix = value;
    } else if (!strcmp(key, "slash_slice")) {
        AHV_REQUIRE(value == 0 || (value >= 1024 && value <= 10752 && value % 16 == 0),
                    "slash_slice must be 0 (chosen per batch) or a multiple of 16 in 1024..10752");
        ctx->opt_slash_slice = value;
    }
    else if (!strcmp(key, "sigma_variant")) {
        AHV_REQUIRE(value >= 0 && value <= 2, "sigma_variant must be 0, 1 or 2");
        ctx->opt_sigma_variant = value;
    } else if (!strcmp(key, "sigma_exact")) {
        AHV_REQUIRE(value == 0 || value == 1, "sigma_exact must be 0 or 1");
        ctx->opt_sigma_exact = value;
    } else if (!strcmp(key, "sigma_long_cut")) {
        AHV_REQUIRE(value >= 4 && value <= 0x7fffffff, "sigma_long_cut must be 4..2^31-1");
        ctx->opt_sigma_long_cut = value;
    } else if (!strcmp(key, "sigma_share")) {
        AHV_REQUIRE(value == 0 || value == 1, "sigma_share must be 0 or 1");
        ctx->opt_sigma_share = value;
    } else if (!strcmp(key, "sigma_heavy")) {
        AHV_REQUIRE(value >= 256 && value <= 0x7fffffff, "sigma_heavy must be 256..2^31-1");
        ctx->opt_sigma_heavy = value;
    } else if (!strcmp(key, "sigma_chunk")) {
        AHV_REQUIRE(value == 1024 || value == 1296 || value == 1440 || value == 1536 || value == 2048, "sigma_chunk must be 1024, 1296, 1440, 1536 or 2048");
        ctx->opt_sigma_chunk = value;
    } else if (!strcmp(key, "chain_lanes")) {
        AHV_REQUIRE(value >= 0 && value <= 32, "chain_lanes must be 0 (auto) or 1..32");
        ctx->opt_chain_lanes = value;
    } else if (!strcmp(key, "compress")) {
        AHV_REQUIRE(value == 0 || value == 1, "compress must be 0 or 1");
        ctx->opt_compress = value;
    } else if (!strcmp(key, "top_max")) {
        AHV_REQUIRE(value == 0 || (value >= 2 && value <= 256), "top_max must be 0 (auto) or 2..256");
        ctx->opt_top_max = value;
    } else if (!strcmp(key, "peak_ctas")) {
        AHV_REQUIRE(value >= 1 && value <= 8, "peak_ctas must be 1..8");
        ctx->opt_peak_ctas = value;
    } else if (!strcmp(key, "leaf_timer")) {
        AHV_REQUIRE(value == 0 || (value > 0 && value + 16 <= AHV_MAX_TIMERS), "leaf_timer base must leave 16 slots");
        ctx->opt_leaf_timer = value;
        ctx->leaf_launches = 0;
    } else {
        set_error("unknown option '%s'", key);
        return AHV_E_INVALID;
    }
    return AHV_OK;
}

}  // extern "C"

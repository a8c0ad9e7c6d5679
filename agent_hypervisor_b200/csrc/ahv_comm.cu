// ahv_comm.cu - the one exchange step of the multi-GPU audit path: every rank's Merkle
// roots (32 B per session) and verdicts (1 B per session) are collected with ONE ncclAllGather issued on
// the context's stream, so a step needs no host synchronisation.  One process per GPU; the communicator
// belongs to the ahv_ctx.  The reference has no communication layer (it is a single-process Python
// library), so there is no reference line to cite here.
//
// NCCL is opened at run time (dlopen of libnccl.so.2, or the path in AHV_NCCL_LIB) the first time a
// communicator is asked for: single-GPU users need no NCCL on their machine, and a process that has
// already loaded an NCCL (e.g. through another framework) shares that copy instead of getting a second one.
#include <dlfcn.h>
#include <nccl.h>       // types and prototypes only; no link-time dependency
#include <stdlib.h>

#include <mutex>

#include "ahv_internal.cuh"

using namespace ahv;

namespace {

struct NcclApi {
    void *handle = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    int version = 0;
};

NcclApi g_nccl;
std::mutex g_nccl_lock;

int nccl_load()
{
    std::lock_guard<std::mutex> guard(g_nccl_lock);
    if (g_nccl.handle) return AHV_OK;
    const char *names[3] = {getenv("AHV_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    char tried[400] = "";
    for (const char *nm : names) {
        if (!nm || !*nm) continue;
        h = dlopen(nm, RTLD_NOW | RTLD_LOCAL);
        if (h) break;
        const char *why = dlerror();
        size_t used = strlen(tried);
        snprintf(tried + used, sizeof tried - used, "%s%s: %s", used ? "; " : "", nm, why ? why : "?");
    }
    if (!h) {
        set_error("NCCL is not available (%s); set AHV_NCCL_LIB to the path of libnccl.so.2", tried);
        return AHV_E_NODEVICE;
    }
    NcclApi api;
    api.handle = h;
#define AHV_SYM(field, name)                                             \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(h, name));  \
    if (!api.field) {                                                    \
        set_error("NCCL library lacks %s", name);                        \
        dlclose(h);                                                      \
        return AHV_E_NODEVICE;                                           \
    }
    AHV_SYM(GetVersion, "ncclGetVersion")
    AHV_SYM(GetUniqueId, "ncclGetUniqueId")
    AHV_SYM(CommInitRank, "ncclCommInitRank")
    AHV_SYM(CommDestroy, "ncclCommDestroy")
    AHV_SYM(AllGather, "ncclAllGather")
    AHV_SYM(GetErrorString, "ncclGetErrorString")
#undef AHV_SYM
    api.GetVersion(&api.version);
    g_nccl = api;
    return AHV_OK;
}

#define AHV_NCCL(call)                                                                                     \
    do {                                                                                                   \
        ncclResult_t r_ = (call);                                                                          \
        if (r_ != ncclSuccess) {                                                                           \
            set_error("%s failed: %s (%s:%d)", #call, g_nccl.GetErrorString(r_), __FILE__, __LINE__);      \
            return AHV_E_CUDA;                                                                             \
        }                                                                                                  \
    } while (0)

// One rank's block of the gathered result: 32*S bytes of roots, then S verdict bytes, padded to 16 bytes.
__host__ __device__ inline uint64_t gather_stride(uint64_t S) { return (33ull * S + 15ull) & ~15ull; }

// roots move as 16-byte words, verdicts as bytes; one launch
__global__ void __launch_bounds__(256)
k_pack_results(const uint8_t *__restrict__ roots, const uint8_t *__restrict__ verdict, uint64_t S, uint8_t *__restrict__ block)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 2 * S) reinterpret_cast<uint4 *>(block)[i] = __ldg(reinterpret_cast<const uint4 *>(roots) + i);
    if (i < S) block[32 * S + i] = verdict ? verdict[i] : (uint8_t)0;
    const uint64_t stride = gather_stride(S);
    if (i < stride - 33 * S) block[33 * S + i] = 0;          // padding bytes: defined content on every rank
}

}  // namespace

namespace ahv {

void comm_release(ahv_ctx *ctx)
{
    if (ctx && ctx->comm && g_nccl.handle) g_nccl.CommDestroy((ncclComm_t)ctx->comm);
    if (ctx) {
        ctx->comm = nullptr;
        ctx->comm_world = 1;
        ctx->comm_rank = 0;
    }
}

}  // namespace ahv

extern "C" {

int ahv_comm_unique_id(uint8_t *out_id)
{
    if (!out_id) {
        set_error("invalid argument: out_id is null");
        return AHV_E_INVALID;
    }
    AHV_TRY(nccl_load());
    static_assert(sizeof(ncclUniqueId) == AHV_COMM_ID_BYTES, "NCCL unique id size");
    ncclUniqueId id;
    AHV_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(out_id, &id, sizeof id);
    return AHV_OK;
}

int ahv_comm_init(ahv_ctx *ctx, int world, int rank, const uint8_t *id)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(world >= 1 && rank >= 0 && rank < world, "need 0 <= rank < world");
    AHV_REQUIRE(!ctx->comm, "this context already has a communicator");
    if (world == 1) {
        ctx->comm_world = 1;
        ctx->comm_rank = 0;
        return AHV_OK;
    }
    AHV_REQUIRE(id, "unique id is null");
    AHV_TRY(nccl_load());
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof uid);
    ncclComm_t comm = nullptr;
    AHV_NCCL(g_nccl.CommInitRank(&comm, world, uid, rank));       // collective: returns once every rank has joined
    ctx->comm = comm;
    ctx->comm_world = world;
    ctx->comm_rank = rank;
    return AHV_OK;
}

int ahv_comm_destroy(ahv_ctx *ctx)
{
    AHV_TRY(use_device(ctx));
    AHV_CUDA(cudaStreamSynchronize(ctx->stream));
    comm_release(ctx);
    return AHV_OK;
}

int ahv_comm_info(ahv_ctx *ctx, int *world, int *rank, int *nccl_version)
{
    AHV_TRY(use_device(ctx));
    if (world) *world = ctx->comm_world > 0 ? ctx->comm_world : 1;
    if (rank) *rank = ctx->comm_rank;
    if (nccl_version) *nccl_version = g_nccl.version;
    return AHV_OK;
}

int ahv_comm_all_gather_dev(ahv_ctx *ctx, const void *d_send, void *d_recv, uint64_t bytes_per_rank)
{
    AHV_TRY(use_device(ctx));
    if (bytes_per_rank == 0) return AHV_OK;
    AHV_REQUIRE(d_send && d_recv, "send / recv buffer is null");
    if (!ctx->comm) {                                             // a world of one
        if (d_send != d_recv) AHV_CUDA(cudaMemcpyAsync(d_recv, d_send, bytes_per_rank, cudaMemcpyDeviceToDevice, ctx->stream));
        return AHV_OK;
    }
    AHV_NCCL(g_nccl.AllGather(d_send, d_recv, bytes_per_rank, ncclUint8, (ncclComm_t)ctx->comm, ctx->stream));
    return AHV_OK;
}

int ahv_comm_all_gather(ahv_ctx *ctx, const void *send, void *recv, uint64_t bytes_per_rank)
{
    AHV_TRY(use_device(ctx));
    if (bytes_per_rank == 0) return AHV_OK;
    AHV_REQUIRE(send && recv, "send / recv buffer is null");
    const uint64_t world = ctx->comm ? (uint64_t)ctx->comm_world : 1;
    AHV_TRY(io_reserve(ctx, pad256(bytes_per_rank) + pad256(world * bytes_per_rank)));
    uint8_t *d_send = (uint8_t *)io_take(ctx, bytes_per_rank), *d_recv = (uint8_t *)io_take(ctx, world * bytes_per_rank);
    if (!d_send || !d_recv) {
        set_error("internal: staging slab too small");
        return AHV_E_NOMEM;
    }
    AHV_CUDA(cudaMemcpyAsync(d_send, send, bytes_per_rank, cudaMemcpyHostToDevice, ctx->stream));
    AHV_TRY(ahv_comm_all_gather_dev(ctx, d_send, d_recv, bytes_per_rank));
    AHV_CUDA(cudaMemcpyAsync(recv, d_recv, world * bytes_per_rank, cudaMemcpyDeviceToHost, ctx->stream));
    AHV_CUDA(cudaStreamSynchronize(ctx->stream));
    return AHV_OK;
}

uint64_t ahv_gather_stride(uint64_t n_sessions) { return gather_stride(n_sessions); }

int ahv_gather_results_dev(ahv_ctx *ctx, const uint8_t *d_roots, const uint8_t *d_verdict, uint64_t S, uint8_t *d_all)
{
    AHV_TRY(use_device(ctx));
    if (S == 0) return AHV_OK;
    AHV_REQUIRE(d_roots && d_all, "roots / out buffer is null");
    AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_roots) & 15u) == 0 && (reinterpret_cast<uintptr_t>(d_all) & 15u) == 0,
                "roots and the gathered buffer must be 16-byte aligned");
    AHV_REQUIRE(S < 0x7fffffffull, "too many sessions per rank");
    const uint64_t stride = gather_stride(S);
    uint8_t *mine = d_all + (uint64_t)ctx->comm_rank * stride;
    k_pack_results<<<grid_for(2 * S, 256), 256, 0, ctx->stream>>>(d_roots, d_verdict, S, mine);
    AHV_CHECK_LAUNCH(ctx);
    if (ctx->comm)                                                // in place: this rank's block already sits in d_all
        AHV_NCCL(g_nccl.AllGather(mine, d_all, stride, ncclUint8, (ncclComm_t)ctx->comm, ctx->stream));
    return AHV_OK;
}

}  // extern "C"

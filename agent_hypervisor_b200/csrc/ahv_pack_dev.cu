// ahv_pack_dev.cu - the canonical packer on the device (widening of SURVEY.md 8f row 1): the same emitter as
// the host packer (pack_core.cuh: byte-identical to json.dumps(..., sort_keys=True).encode() of
// audit/delta.py:43-62), one thread per delta, so that a replay uploads COLUMNS (the fields, < 200 bytes per
// delta) instead of pre-images (~330 bytes) and no host core serialises anything.
//   pass 1  k_pack_sizes : every delta's length and the position of its parent_hash value
//   scan    k_scan_*     : lengths -> offsets (tile totals, their prefix sums, tiles with their carry: three launches;
//                          one CTA alone for short arrays)
//   pass 2  k_pack_write : the bytes, into the arena the hashing kernels read (staged per warp in shared memory,
//                          written out with 16-byte stores)
// HBM-bound in principle (columns in, pre-images out); in practice bound by the serial, divergent emission of
// each lane's delta (a few instructions per output byte).
#include "ahv_internal.cuh"
#include "pack_core.cuh"
#include "scan.cuh"

using namespace ahv::scan;

using namespace ahv;
using namespace ahv_pack;

namespace {

__global__ void __launch_bounds__(256)
k_pack_sizes(const Cols c, uint64_t *__restrict__ offsets, uint32_t *__restrict__ ppos)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= c.n) return;
    SinkT<false> s(nullptr);
    const uint64_t pp = emit_delta(s, c, i);
    offsets[i + 1] = s.n;
    if (i == 0) offsets[0] = 0;
    // a position past 4 GiB cannot be stored: UINT32_MAX - 1 is never a valid one, so it reads as "other"
    if (ppos) ppos[i] = pp == 0xffffffffull ? 0xffffffffu : (uint32_t)(pp < 0xfffffffeull ? pp : 0xfffffffeull);
}

// Pass 2.  The 32 deltas of a warp are neighbours in the arena, so the warp's output is ONE contiguous byte range:
// every lane serialises its delta into the warp's shared-memory slot (byte stores are cheap there), then the
// warp copies the range out with 16-byte stores (the slot is offset by the range's misalignment, so 16-byte
// pieces of shared memory line up with 16-byte pieces of the arena).  A warp whose 32 pre-images exceed the slot
// (very long strings) writes straight to global memory instead.
constexpr uint32_t PACK_WARP_SLOT = 12288;      // bytes of shared memory per warp (+16 for the alignment shift)
constexpr uint32_t PACK_THREADS = 128;

__global__ void __launch_bounds__(PACK_THREADS)
k_pack_write(const Cols c, const uint64_t *__restrict__ offsets, uint8_t *__restrict__ arena, uint64_t capacity,
             uint32_t *__restrict__ status)
{
    extern __shared__ __align__(16) uint8_t pack_smem[];
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint64_t i0 = i - lane;
    if (i0 >= c.n) return;                                   // whole warp out of range
    if (offsets[c.n] > capacity) {                           // nothing is written past the caller's arena
        if (i == 0) atomicOr(status, AHV_STATUS_PACK_OVERFLOW);
        return;
    }
    const uint64_t i1 = (i0 + 32 < c.n) ? i0 + 32 : c.n;
    const uint64_t base = offsets[i0], end = offsets[i1];
    const uint32_t shift = (uint32_t)((reinterpret_cast<uintptr_t>(arena) + base) & 15u);
    uint8_t *slot = pack_smem + (size_t)wid * (PACK_WARP_SLOT + 16);
    if (end - base > PACK_WARP_SLOT) {                       // too long for the slot: direct
        if (i < c.n) {
            SinkT<true> s(arena + offsets[i]);
            emit_delta(s, c, i);
        }
        return;
    }
    if (i < c.n) {
        SinkT<true> s(slot + shift + (offsets[i] - base));
        emit_delta(s, c, i);
    }
    __syncwarp();
    const uint32_t total = (uint32_t)(end - base);
    uint8_t *g = arena + base - shift;                       // 16-byte aligned; slot[k] <-> g[k]
    const uint32_t span = shift + total;                     // bytes [shift, span) of the slot are data
    for (uint32_t k = lane * 16u; k < span; k += 32u * 16u) {
        if (k >= shift && k + 16u <= span) {
            *reinterpret_cast<uint4 *>(g + k) = *reinterpret_cast<const uint4 *>(slot + k);
        } else {                                             // first / last piece: only the bytes that are ours
            const uint32_t lo = k < shift ? shift : k, hi = (k + 16u < span) ? k + 16u : span;
            for (uint32_t q = lo; q < hi; ++q) g[q] = slot[q];
        }
    }
}

}  // namespace

extern "C" int ahv_pack_deltas_dev(ahv_ctx *ctx, const ahv_delta_columns *cols, uint64_t *d_out_offsets,
                                   uint8_t *d_out_arena, uint64_t arena_capacity, uint64_t *out_bytes,
                                   uint32_t *d_out_parent_pos, uint32_t flags)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(cols && d_out_offsets, "cols / out_offsets is null");
    Cols c;
    c.n = cols->n;
    c.delta_id = {cols->delta_id_blob, cols->delta_id_off, nullptr};
    c.session_id = {cols->session_id_blob, cols->session_id_off, nullptr};
    c.agent_did = {cols->agent_did_blob, cols->agent_did_off, nullptr};
    c.turn_id = cols->turn_id;
    c.ts_us = cols->timestamp_us;
    c.ts_off = cols->timestamp_utcoffset_s;
    c.parent_kind = cols->parent_kind;
    c.parent_digest = cols->parent_digest;
    c.parent_other = {cols->parent_other_blob, cols->parent_other_off, nullptr};
    c.change_rowptr = cols->change_rowptr;
    c.ch_path = {cols->change_path_blob, cols->change_path_off, nullptr};
    c.ch_op = {cols->change_op_blob, cols->change_op_off, nullptr};
    c.ch_content = {cols->change_content_blob, cols->change_content_off, cols->change_content_null};
    c.ch_prev = {cols->change_prev_blob, cols->change_prev_off, cols->change_prev_null};
    const uint64_t n = c.n;
    AHV_REQUIRE(n < 0xffffffffull, "too many deltas (limit 2^32-2 per call)");
    if (n == 0) {
        AHV_CUDA(cudaMemsetAsync(d_out_offsets, 0, 8, ctx->stream));
        if (out_bytes) *out_bytes = 0;
        return AHV_OK;
    }
    AHV_REQUIRE(c.turn_id && c.ts_us && c.ts_off && c.parent_kind && c.change_rowptr && c.delta_id.off && c.session_id.off &&
                c.agent_did.off, "a required column is null");
    if (!((flags & AHV_PACK_OFFSETS_VALID) && d_out_arena)) {
        k_pack_sizes<<<grid_for(n, 256), 256, 0, ctx->stream>>>(c, d_out_offsets, d_out_parent_pos);
        AHV_CHECK_LAUNCH(ctx);
        AHV_TRY(scan_u64_inplace(ctx, d_out_offsets, n, 1, 0));
    }
    if (d_out_arena) {
        const size_t smem = (size_t)(PACK_THREADS / 32) * (PACK_WARP_SLOT + 16);
        static bool attr_set[64] = {};                        // per device: function attributes belong to a device's context
        const int dev = ctx->device & 63;
        if (!attr_set[dev]) {
            AHV_CUDA(cudaFuncSetAttribute(k_pack_write, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set[dev] = true;
        }
        k_pack_write<<<grid_for(n, PACK_THREADS), PACK_THREADS, smem, ctx->stream>>>(c, d_out_offsets, d_out_arena, arena_capacity,
                                                                                      ctx->status_d);
        AHV_CHECK_LAUNCH(ctx);
    }
    if (out_bytes) {                                         // the caller wants the size on the host: one synchronisation
        AHV_CUDA(cudaMemcpyAsync(out_bytes, d_out_offsets + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
        AHV_CUDA(cudaStreamSynchronize(ctx->stream));
        if (d_out_arena && *out_bytes > arena_capacity) {
            __atomic_and_fetch(ctx->status_h, ~AHV_STATUS_PACK_OVERFLOW, __ATOMIC_ACQ_REL);
            set_error("arena capacity %llu < %llu bytes needed", (unsigned long long)arena_capacity, (unsigned long long)*out_bytes);
            return AHV_E_TOOLARGE;
        }
    }
    return AHV_OK;
}

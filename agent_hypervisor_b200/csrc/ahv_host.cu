// ahv_host.cu - host-pointer entry points: stage caller buffers on the device, run the
// device-pointer entry point, copy results back, synchronise.  These are what the Python
// shim (agent_hypervisor_b200/_native.py) binds; the reference-side call each replaces is
// cited in include/ahv.h.
#include <vector>

#include "ahv_internal.cuh"

using namespace ahv;

namespace {

struct Stager {
    ahv_ctx *ctx;
    int rc = AHV_OK;
    explicit Stager(ahv_ctx *c) : ctx(c) {}

    template <class T>
    T *in(const T *h, uint64_t count)
    {
        if (rc != AHV_OK || !h || count == 0) return nullptr;
        T *d = (T *)io_take(ctx, count * sizeof(T));
        if (!d) {
            set_error("internal: staging slab too small");
            rc = AHV_E_NOMEM;
            return nullptr;
        }
        cudaError_t e = cudaMemcpyAsync(d, h, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream);
        if (e != cudaSuccess) {
            set_error("H2D copy failed: %s", cudaGetErrorString(e));
            rc = AHV_E_CUDA;
        }
        return d;
    }
    template <class T>
    T *out(T *h, uint64_t count)
    {
        if (rc != AHV_OK || !h || count == 0) return nullptr;
        T *d = (T *)io_take(ctx, count * sizeof(T));
        if (!d) {
            set_error("internal: staging slab too small");
            rc = AHV_E_NOMEM;
        }
        return d;
    }
    template <class T>
    void back(T *h, const T *d, uint64_t count)
    {
        if (rc != AHV_OK || !h || !d || count == 0) return;
        cudaError_t e = cudaMemcpyAsync(h, d, count * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream);
        if (e != cudaSuccess) {
            set_error("D2H copy failed: %s", cudaGetErrorString(e));
            rc = AHV_E_CUDA;
        }
    }
    int finish()
    {
        if (rc != AHV_OK) return rc;
        AHV_CUDA(cudaStreamSynchronize(ctx->stream));
        return check_status(ctx);
    }
};

uint64_t p256(uint64_t b) { return pad256(b ? b : 1); }

// message count and arena size from (offsets | fixed_len)
// offsets[lo .. hi] non-decreasing and inside an arena of `bytes` bytes
int check_offsets(const uint64_t *offsets, uint64_t lo, uint64_t hi, uint64_t bytes)
{
    // every message length is offsets[i+1] - offsets[i]: one decreasing pair would wrap to ~2^64 bytes
    uint64_t bad = 0;
    for (uint64_t i = lo; i < hi; ++i) bad |= (uint64_t)(offsets[i + 1] < offsets[i]);
    AHV_REQUIRE(!bad && offsets[hi] <= bytes, "offsets must be non-decreasing");
    return AHV_OK;
}

int arena_extent(const uint64_t *offsets, uint64_t n, uint64_t fixed_len, uint64_t *bytes, bool check = true)
{
    if (offsets) {
        *bytes = offsets[n];
        if (check) AHV_TRY(check_offsets(offsets, 0, n, *bytes));
    } else {
        *bytes = n * fixed_len;
    }
    return AHV_OK;
}

}  // namespace

namespace {

// row pointers of a chunk of sessions, made relative to the chunk's first leaf
__global__ void __launch_bounds__(256)
k_rebase_u64(uint64_t *__restrict__ a, uint64_t n, uint64_t base)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] -= base;
}

}  // namespace

extern "C" {

int ahv_sha256_batch(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t n, uint64_t fixed_len,
                     uint8_t *out_digests)
{
    AHV_TRY(use_device(ctx));
    if (n == 0) return AHV_OK;
    AHV_REQUIRE(out_digests, "out_digests is null");
    AHV_REQUIRE(offsets || fixed_len > 0, "offsets is NULL and fixed_len is 0");
    uint64_t ab = 0;
    AHV_TRY(arena_extent(offsets, n, fixed_len, &ab));
    AHV_REQUIRE(arena || ab == 0, "arena is null");
    AHV_TRY(io_reserve(ctx, p256(ab + 16) + p256(8 * (n + 1)) + p256(32 * n)));
    Stager st(ctx);
    uint8_t *d_arena = (uint8_t *)io_take(ctx, ab + 16);       // slack: aligned reads may touch the last word
    if (ab) AHV_CUDA(cudaMemcpyAsync(d_arena, arena, ab, cudaMemcpyHostToDevice, ctx->stream));
    uint64_t *d_off = st.in(offsets, offsets ? n + 1 : 0);
    uint8_t *d_out = st.out(out_digests, 32 * n);
    if (st.rc != AHV_OK) return st.rc;
    AHV_TRY(ahv_sha256_batch_dev(ctx, d_arena, ab, d_off, n, fixed_len, d_out));
    st.back(out_digests, d_out, 32 * n);
    return st.finish();
}

int ahv_merkle_roots(ahv_ctx *ctx, const uint8_t *digests, const uint64_t *rowptr, uint64_t S, uint8_t *out_roots,
                     uint8_t *out_empty)
{
    AHV_TRY(use_device(ctx));
    if (S == 0) return AHV_OK;
    AHV_REQUIRE(rowptr && out_roots, "session_rowptr / out_roots is null");
    const uint64_t n = rowptr[S] - rowptr[0];
    AHV_REQUIRE(rowptr[0] == 0, "session_rowptr[0] must be 0");
    AHV_REQUIRE(digests || n == 0, "digests is null");
    uint64_t maxl = 0;
    for (uint64_t s = 0; s < S; ++s) {
        AHV_REQUIRE(rowptr[s + 1] >= rowptr[s], "session_rowptr must be non-decreasing");
        uint64_t c = rowptr[s + 1] - rowptr[s];
        if (c > maxl) maxl = c;
    }
    AHV_TRY(io_reserve(ctx, p256(32 * n) + p256(8 * (S + 1)) + p256(32 * S) + p256(S)));
    Stager st(ctx);
    uint8_t *d_dig = st.in(digests, 32 * n);
    uint64_t *d_rp = st.in(rowptr, S + 1);
    uint8_t *d_roots = st.out(out_roots, 32 * S);
    uint8_t *d_empty = (uint8_t *)io_take(ctx, S);
    if (st.rc != AHV_OK) return st.rc;
    if (!d_dig) d_dig = d_roots;   // n == 0: never dereferenced, keeps alignment checks happy
    AHV_TRY(ahv_merkle_roots_dev(ctx, d_dig, d_rp, S, n, maxl ? maxl : 1, d_roots, d_empty));
    st.back(out_roots, d_roots, 32 * S);
    st.back(out_empty, d_empty, S);
    return st.finish();
}

static int session_shape(const uint64_t *rowptr, uint64_t S, uint64_t uniform, uint64_t *n, uint64_t *maxl)
{
    if (uniform) {
        *n = S * uniform;
        *maxl = uniform;
        return AHV_OK;
    }
    AHV_REQUIRE(rowptr || S == 0, "session_rowptr is NULL and uniform_leaves is 0");
    *n = S ? rowptr[S] : 0;
    *maxl = 0;
    if (S) AHV_REQUIRE(rowptr[0] == 0, "session_rowptr[0] must be 0");
    for (uint64_t s = 0; s < S; ++s) {
        AHV_REQUIRE(rowptr[s + 1] >= rowptr[s], "session_rowptr must be non-decreasing");
        uint64_t c = rowptr[s + 1] - rowptr[s];
        if (c > *maxl) *maxl = c;
    }
    if (*maxl == 0) *maxl = 1;
    return AHV_OK;
}

int ahv_commit(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t fixed_len,
               const uint64_t *rowptr, uint64_t S, uint64_t uniform, uint8_t *out_digests, uint8_t *out_roots,
               uint8_t *out_empty)
{
    AHV_TRY(use_device(ctx));
    if (S == 0) return AHV_OK;
    AHV_REQUIRE(out_roots, "out_roots is null");
    uint64_t n = 0, maxl = 0, ab = 0;
    AHV_TRY(session_shape(rowptr, S, uniform, &n, &maxl));
    AHV_REQUIRE(offsets || fixed_len > 0 || n == 0, "offsets is NULL and fixed_len is 0");
    AHV_TRY(arena_extent(offsets, n, fixed_len, &ab));
    AHV_REQUIRE(arena || ab == 0, "arena is null");
    AHV_TRY(io_reserve(ctx, p256(ab + 16) + p256(8 * (n + 1)) + p256(8 * (S + 1)) + p256(32 * n) + p256(32 * S) + p256(S)));
    Stager st(ctx);
    uint8_t *d_arena = (uint8_t *)io_take(ctx, ab + 16);
    if (ab) AHV_CUDA(cudaMemcpyAsync(d_arena, arena, ab, cudaMemcpyHostToDevice, ctx->stream));
    uint64_t *d_off = st.in(offsets, offsets ? n + 1 : 0);
    uint64_t *d_rp = uniform ? nullptr : st.in(rowptr, S + 1);
    uint8_t *d_dig = st.out(out_digests, 32 * n);
    uint8_t *d_roots = st.out(out_roots, 32 * S);
    uint8_t *d_empty = (uint8_t *)io_take(ctx, S);
    if (st.rc != AHV_OK) return st.rc;
    AHV_TRY(ahv_commit_dev(ctx, d_arena, ab, d_off, fixed_len, d_rp, S, n, uniform, maxl, d_dig, d_roots, d_empty));
    st.back(out_digests, d_dig, 32 * n);
    st.back(out_roots, d_roots, 32 * S);
    st.back(out_empty, d_empty, S);
    return st.finish();
}

int ahv_verify_chains(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t fixed_len,
                      const uint8_t *parents, const uint8_t *kind, const uint64_t *rowptr, uint64_t S,
                      uint64_t uniform, const uint8_t *expected_roots, uint8_t *out_verdict, int64_t *out_first_bad,
                      uint8_t *out_digests, uint8_t *out_roots, uint8_t *out_match)
{
    AHV_TRY(use_device(ctx));
    if (S == 0) return AHV_OK;
    uint64_t n = 0, maxl = 0, ab = 0;
    AHV_TRY(session_shape(rowptr, S, uniform, &n, &maxl));
    AHV_REQUIRE(offsets || fixed_len > 0 || n == 0, "offsets is NULL and fixed_len is 0");
    AHV_REQUIRE((parents && kind) || n == 0, "parent_digests / parent_kind is null");
    AHV_REQUIRE(!out_match || expected_roots, "out_root_match needs expected_roots");
    AHV_TRY(arena_extent(offsets, n, fixed_len, &ab));
    AHV_REQUIRE(arena || ab == 0, "arena is null");
    AHV_TRY(io_reserve(ctx, p256(ab + 16) + p256(8 * (n + 1)) + p256(8 * (S + 1)) + 2 * p256(32 * n) + p256(n) +
                                3 * p256(32 * S) + 2 * p256(S) + p256(8 * S)));
    Stager st(ctx);
    uint8_t *d_arena = (uint8_t *)io_take(ctx, ab + 16);
    if (ab) AHV_CUDA(cudaMemcpyAsync(d_arena, arena, ab, cudaMemcpyHostToDevice, ctx->stream));
    uint64_t *d_off = st.in(offsets, offsets ? n + 1 : 0);
    uint64_t *d_rp = uniform ? nullptr : st.in(rowptr, S + 1);
    uint8_t *d_par = st.in(parents, 32 * n);
    uint8_t *d_kind = st.in(kind, n);
    uint8_t *d_exp = st.in(expected_roots, 32 * S);
    uint8_t *d_verdict = st.out(out_verdict, S);
    int64_t *d_fb = st.out(out_first_bad, S);
    uint8_t *d_dig = st.out(out_digests, 32 * n);
    uint8_t *d_roots = st.out(out_roots, 32 * S);
    uint8_t *d_match = st.out(out_match, S);
    if (st.rc != AHV_OK) return st.rc;
    if (n == 0) {   // only empty sessions: verify_chain's loop body never runs -> True
        if (out_verdict) memset(out_verdict, 1, S);
        if (out_first_bad) for (uint64_t s = 0; s < S; ++s) out_first_bad[s] = -1;
        if (out_roots) memset(out_roots, 0, 32 * S);
        if (out_match) memset(out_match, 0, S);
        return AHV_OK;
    }
    AHV_TRY(ahv_verify_chains_dev(ctx, d_arena, ab, d_off, fixed_len, d_par, d_kind, d_rp, S, n, uniform, maxl, d_exp,
                                  d_verdict, d_fb, d_dig, d_roots, d_match));
    st.back(out_verdict, d_verdict, S);
    st.back(out_first_bad, d_fb, S);
    st.back(out_digests, d_dig, 32 * n);
    st.back(out_roots, d_roots, 32 * S);
    st.back(out_match, d_match, S);
    return st.finish();
}

// Commit + verify of one batch on ONE upload: leaf digests and Merkle roots (audit/delta.py:63,
// 117-134), then a full verify pass (re-hash, parent links, root rebuild and compare against the
// roots just committed; audit/delta.py:136-152, audit/commitment.py:59-64).  A batch of at least 64
// sessions and 8 MB is cut into session chunks and pipelined: chunk k+1 is copied in while chunk k is
// hashed and chunk k-1's results are copied out (three streams, PCIe full duplex); ragged sessions get
// per-chunk row pointers relative to the chunk's first leaf.
int ahv_audit_sessions(ahv_ctx *ctx, const uint8_t *arena, const uint64_t *offsets, uint64_t fixed_len,
                       const uint8_t *parents, const uint8_t *kind, const uint32_t *parent_pos,
                       const uint64_t *rowptr, uint64_t S, uint64_t uniform, const uint8_t *expected_roots,
                       uint8_t *out_digests, uint8_t *out_roots, uint8_t *out_empty, uint8_t *out_verdict,
                       int64_t *out_first_bad, uint8_t *out_match)
{
    AHV_TRY(use_device(ctx));
    if (S == 0) return AHV_OK;
    AHV_REQUIRE(out_roots && out_verdict, "out_roots / out_verdict is null");
    uint64_t n = 0, maxl = 0, ab = 0;
    AHV_TRY(session_shape(rowptr, S, uniform, &n, &maxl));
    AHV_REQUIRE(offsets || fixed_len > 0 || n == 0, "offsets is NULL and fixed_len is 0");
    AHV_REQUIRE((parents && kind) || parent_pos || n == 0,
                "give parent_digests + parent_kind, or parent_pos (positions of the parent values in the pre-images)");
    const bool from_pos = !(parents && kind);
    // (the offsets of a pipelined call are checked chunk by chunk, while the chunk before is on its way up:
    // 10^7 offsets take a core 6 ms, an eighth of the upload they would otherwise wait in front of)
    AHV_TRY(arena_extent(offsets, n, fixed_len, &ab, /*check=*/false));
    AHV_REQUIRE(arena || ab == 0, "arena is null");
    if (n == 0) {   // only empty sessions
        memset(out_roots, 0, 32 * S);
        if (out_empty) memset(out_empty, 1, S);
        memset(out_verdict, 1, S);
        if (out_first_bad) for (uint64_t s = 0; s < S; ++s) out_first_bad[s] = -1;
        if (out_match) memset(out_match, 0, S);
        return AHV_OK;
    }
    AHV_TRY(io_reserve(ctx, p256(ab + 16) + p256(8 * (n + 1)) + p256(8 * (S + 1)) + 2 * p256(32 * n) + p256(n) + p256(4 * n) +
                                3 * p256(32 * S) + 3 * p256(S) + p256(8 * S) + p256(8 * (S + 64))));
    uint8_t *d_arena = (uint8_t *)io_take(ctx, ab + 16);
    uint32_t *d_pos = from_pos ? (uint32_t *)io_take(ctx, 4 * n) : nullptr;
    uint64_t *d_off = offsets ? (uint64_t *)io_take(ctx, 8 * (n + 1)) : nullptr;
    uint64_t *d_rp = uniform ? nullptr : (uint64_t *)io_take(ctx, 8 * (S + 1));
    uint8_t *d_par = (uint8_t *)io_take(ctx, 32 * n);
    uint8_t *d_kind = (uint8_t *)io_take(ctx, n);
    uint8_t *d_dig = (uint8_t *)io_take(ctx, 32 * n);
    uint8_t *d_roots = (uint8_t *)io_take(ctx, 32 * S);
    uint8_t *d_roots_v = (uint8_t *)io_take(ctx, 32 * S);
    uint8_t *d_exp = expected_roots ? (uint8_t *)io_take(ctx, 32 * S) : nullptr;   // stored commitments to verify against
    uint8_t *d_empty = (uint8_t *)io_take(ctx, S);
    uint8_t *d_verdict = (uint8_t *)io_take(ctx, S);
    uint8_t *d_match = (uint8_t *)io_take(ctx, S);
    int64_t *d_fb = (int64_t *)io_take(ctx, 8 * S);
    if (!d_arena || (from_pos && !d_pos) || (offsets && !d_off) || (!uniform && !d_rp) || !d_par || !d_kind || !d_dig || !d_roots ||
        !d_roots_v || (expected_roots && !d_exp) || !d_empty || !d_verdict || !d_match || !d_fb) {
        set_error("internal: staging slab too small");
        return AHV_E_NOMEM;
    }

    const bool pipelined = S >= 64 && ab >= (8ull << 20);      // worth overlapping: many sessions, megabytes to move
    if (!pipelined) {
        if (offsets) AHV_TRY(check_offsets(offsets, 0, n, ab));
        cudaStream_t st = ctx->stream;
        AHV_CUDA(cudaMemcpyAsync(d_arena, arena, ab, cudaMemcpyHostToDevice, st));
        if (offsets) AHV_CUDA(cudaMemcpyAsync(d_off, offsets, 8 * (n + 1), cudaMemcpyHostToDevice, st));
        if (!uniform) AHV_CUDA(cudaMemcpyAsync(d_rp, rowptr, 8 * (S + 1), cudaMemcpyHostToDevice, st));
        if (from_pos) {
            AHV_CUDA(cudaMemcpyAsync(d_pos, parent_pos, 4 * n, cudaMemcpyHostToDevice, st));
            AHV_TRY(ahv_extract_parents_dev(ctx, d_arena, ab, d_off, fixed_len, d_pos, n, d_par, d_kind));
        } else {
            AHV_CUDA(cudaMemcpyAsync(d_par, parents, 32 * n, cudaMemcpyHostToDevice, st));
            AHV_CUDA(cudaMemcpyAsync(d_kind, kind, n, cudaMemcpyHostToDevice, st));
        }
        if (d_exp) AHV_CUDA(cudaMemcpyAsync(d_exp, expected_roots, 32 * S, cudaMemcpyHostToDevice, st));
        AHV_TRY(ahv_commit_dev(ctx, d_arena, ab, d_off, fixed_len, d_rp, S, n, uniform, maxl, d_dig, d_roots, d_empty));
        AHV_TRY(ahv_verify_chains_dev(ctx, d_arena, ab, d_off, fixed_len, d_par, d_kind, d_rp, S, n, uniform, maxl,
                                      d_exp ? d_exp : d_roots, d_verdict, d_fb, nullptr, d_roots_v, d_match));
        if (out_digests) AHV_CUDA(cudaMemcpyAsync(out_digests, d_dig, 32 * n, cudaMemcpyDeviceToHost, st));
        AHV_CUDA(cudaMemcpyAsync(out_roots, d_roots, 32 * S, cudaMemcpyDeviceToHost, st));
        if (out_empty) AHV_CUDA(cudaMemcpyAsync(out_empty, d_empty, S, cudaMemcpyDeviceToHost, st));
        AHV_CUDA(cudaMemcpyAsync(out_verdict, d_verdict, S, cudaMemcpyDeviceToHost, st));
        if (out_first_bad) AHV_CUDA(cudaMemcpyAsync(out_first_bad, d_fb, 8 * S, cudaMemcpyDeviceToHost, st));
        if (out_match) AHV_CUDA(cudaMemcpyAsync(out_match, d_match, S, cudaMemcpyDeviceToHost, st));
        AHV_CUDA(cudaStreamSynchronize(st));
        return AHV_OK;
    }

    // ---- pipelined: sessions are independent, so each chunk is a complete sub-batch
    constexpr int MAXC = 32;
    int nchunk = (int)(S / 512);
    if (nchunk < 2) nchunk = 2;
    if (nchunk > MAXC) nchunk = MAXC;
    const uint64_t per = (S + nchunk - 1) / nchunk;
    static_assert(2 * MAXC + 8 <= AHV_PIPE_EVENTS, "the context's event pool must cover every chunk");
    struct Drain {                        // on every exit path the three streams have drained before the buffers are reused
        ahv_ctx *ctx;
        ~Drain()
        {
            cudaStreamSynchronize(ctx->copy_stream);
            cudaStreamSynchronize(ctx->stream);
            cudaStreamSynchronize(ctx->d2h_stream);
        }
    } drain{ctx};
    cudaEvent_t *ev_in = ctx->ev_pipe, *ev_done = ctx->ev_pipe + MAXC;     // made once in ahv_create
    // Ragged sessions: chunk k's tree works on ITS digests, so it needs row pointers relative to its first
    // leaf; they are laid out chunk after chunk (each with its own end entry) and uploaded with the chunk.
    // (rebased on the device, k_rebase_u64: the caller's row pointers go up as they are - no O(sessions) loop and
    // no pageable staging buffer on the host, which cost up to half the call at millions of tiny sessions)
    uint64_t *d_rel = nullptr;
    if (!uniform) {
        d_rel = (uint64_t *)io_take(ctx, 8 * (S + (size_t)nchunk));
        if (!d_rel) {
            set_error("internal: staging slab too small");
            return AHV_E_NOMEM;
        }
    }
    // (the first chunk may grow the scratch slab, which synchronises once; later chunks reuse it)
    int rc = AHV_OK;
    for (int k = 0; k < nchunk && rc == AHV_OK; ++k) {
        const uint64_t s0 = (uint64_t)k * per, s1 = (s0 + per < S) ? s0 + per : S;
        if (s0 >= s1) { nchunk = k; break; }
        const uint64_t Sc = s1 - s0;
        const uint64_t l0 = uniform ? s0 * uniform : rowptr[s0], l1 = uniform ? s1 * uniform : rowptr[s1], nc = l1 - l0;
        if (offsets && (rc = check_offsets(offsets, l0, l1, ab)) != AHV_OK) break;   // nothing of this chunk has been touched yet
        const uint64_t b0 = offsets ? offsets[l0] : l0 * fixed_len, b1 = offsets ? offsets[l1] : l1 * fixed_len, bc = b1 - b0;
        cudaStream_t hs = ctx->copy_stream;
        if (bc) AHV_CUDA(cudaMemcpyAsync(d_arena + b0, arena + b0, bc, cudaMemcpyHostToDevice, hs));
        if (offsets) {
            // entry l0 is already there (it is the previous chunk's last entry), except for the first chunk
            const uint64_t f = k == 0 ? l0 : l0 + 1;
            if (l1 + 1 > f) AHV_CUDA(cudaMemcpyAsync(d_off + f, offsets + f, 8 * (l1 + 1 - f), cudaMemcpyHostToDevice, hs));
        }
        uint64_t *d_rp_k = nullptr;
        if (!uniform) {
            d_rp_k = d_rel + s0 + k;
            AHV_CUDA(cudaMemcpyAsync(d_rp_k, rowptr + s0, 8 * (Sc + 1), cudaMemcpyHostToDevice, hs));
            if (l0) {
                k_rebase_u64<<<grid_for(Sc + 1, 256), 256, 0, hs>>>(d_rp_k, Sc + 1, l0);
                AHV_CUDA(cudaGetLastError());
            }
        }
        if (nc) {
            if (from_pos) {
                AHV_CUDA(cudaMemcpyAsync(d_pos + l0, parent_pos + l0, 4 * nc, cudaMemcpyHostToDevice, hs));
            } else {
                AHV_CUDA(cudaMemcpyAsync(d_par + 32 * l0, parents + 32 * l0, 32 * nc, cudaMemcpyHostToDevice, hs));
                AHV_CUDA(cudaMemcpyAsync(d_kind + l0, kind + l0, nc, cudaMemcpyHostToDevice, hs));
            }
        }
        if (d_exp) AHV_CUDA(cudaMemcpyAsync(d_exp + 32 * s0, expected_roots + 32 * s0, 32 * Sc, cudaMemcpyHostToDevice, hs));
        AHV_CUDA(cudaEventRecord(ev_in[k], hs));
        AHV_CUDA(cudaStreamWaitEvent(ctx->stream, ev_in[k], 0));
        // byte offsets stay absolute (base d_arena); with fixed-size deltas the chunk's own base is passed
        const uint8_t *a_k = offsets ? d_arena : d_arena + b0;
        const uint64_t *off_k = offsets ? d_off + l0 : nullptr;
        if (from_pos && nc)
            rc = ahv_extract_parents_dev(ctx, a_k, ab, off_k, fixed_len, d_pos + l0, nc, d_par + 32 * l0, d_kind + l0);
        if (rc != AHV_OK) break;
        rc = ahv_commit_dev(ctx, a_k, ab, off_k, fixed_len, d_rp_k, Sc, nc, uniform, maxl, d_dig + 32 * l0,
                            d_roots + 32 * s0, d_empty + s0);
        if (rc == AHV_OK)
            rc = ahv_verify_chains_dev(ctx, a_k, ab, off_k, fixed_len, d_par + 32 * l0, d_kind + l0, d_rp_k, Sc, nc, uniform,
                                       maxl, (d_exp ? d_exp : d_roots) + 32 * s0, d_verdict + s0, d_fb + s0, nullptr,
                                       d_roots_v + 32 * s0, d_match + s0);
        if (rc != AHV_OK) break;
        AHV_CUDA(cudaEventRecord(ev_done[k], ctx->stream));
        cudaStream_t ds = ctx->d2h_stream;
        AHV_CUDA(cudaStreamWaitEvent(ds, ev_done[k], 0));
        if (out_digests && nc) AHV_CUDA(cudaMemcpyAsync(out_digests + 32 * l0, d_dig + 32 * l0, 32 * nc, cudaMemcpyDeviceToHost, ds));
        AHV_CUDA(cudaMemcpyAsync(out_roots + 32 * s0, d_roots + 32 * s0, 32 * Sc, cudaMemcpyDeviceToHost, ds));
        if (out_empty) AHV_CUDA(cudaMemcpyAsync(out_empty + s0, d_empty + s0, Sc, cudaMemcpyDeviceToHost, ds));
        AHV_CUDA(cudaMemcpyAsync(out_verdict + s0, d_verdict + s0, Sc, cudaMemcpyDeviceToHost, ds));
        if (out_first_bad) AHV_CUDA(cudaMemcpyAsync(out_first_bad + s0, d_fb + s0, 8 * Sc, cudaMemcpyDeviceToHost, ds));
        if (out_match) AHV_CUDA(cudaMemcpyAsync(out_match + s0, d_match + s0, Sc, cudaMemcpyDeviceToHost, ds));
    }
    // Option blocking_sync: wait without spinning when the host cores are busy packing the next batch (the copy-out
    // stream is the last link of every chunk's chain, so one blocking event on it covers all three streams).
    if (rc == AHV_OK && ctx->opt_blocking_sync) {
        cudaEvent_t ev_end;
        if (cudaEventCreateWithFlags(&ev_end, cudaEventBlockingSync | cudaEventDisableTiming) == cudaSuccess) {
            if (cudaEventRecord(ev_end, ctx->d2h_stream) == cudaSuccess) cudaEventSynchronize(ev_end);
            cudaEventDestroy(ev_end);
        }
    }
    cudaStreamSynchronize(ctx->copy_stream);
    cudaStreamSynchronize(ctx->stream);
    cudaError_t e = cudaStreamSynchronize(ctx->d2h_stream);
    if (rc != AHV_OK) return rc;
    if (e != cudaSuccess) {
        set_error("pipelined audit failed: %s", cudaGetErrorString(e));
        return AHV_E_CUDA;
    }
    return AHV_OK;
}

int ahv_chain_capture(ahv_ctx *ctx, uint8_t *arena, const uint64_t *offsets, const uint32_t *hexpos,
                      const uint64_t *rowptr, uint64_t S, uint8_t *out_digests, uint8_t *out_parents,
                      uint8_t *out_kind)
{
    AHV_TRY(use_device(ctx));
    if (S == 0) return AHV_OK;
    AHV_REQUIRE(arena && offsets && hexpos && rowptr, "arena / offsets / hexpos / session_rowptr is null");
    uint64_t n = 0, maxl = 0;
    AHV_TRY(session_shape(rowptr, S, 0, &n, &maxl));
    if (n == 0) return AHV_OK;
    const uint64_t ab = offsets[n];
    for (uint64_t i = 0; i < n; ++i)
        if (hexpos[i] != 0xffffffffu)
            AHV_REQUIRE((uint64_t)hexpos[i] + 64 <= offsets[i + 1] - offsets[i], "hexpos outside its message");
    AHV_TRY(io_reserve(ctx, p256(ab + 16) + p256(8 * (n + 1)) + p256(4 * n) + p256(8 * (S + 1)) + 2 * p256(32 * n) + p256(n)));
    Stager st(ctx);
    uint8_t *d_arena = (uint8_t *)io_take(ctx, ab + 16);
    AHV_CUDA(cudaMemcpyAsync(d_arena, arena, ab, cudaMemcpyHostToDevice, ctx->stream));
    uint64_t *d_off = st.in(offsets, n + 1);
    uint32_t *d_hp = st.in(hexpos, n);
    uint64_t *d_rp = st.in(rowptr, S + 1);
    uint8_t *d_dig = st.out(out_digests, 32 * n);
    uint8_t *d_par = st.out(out_parents, 32 * n);
    uint8_t *d_kind = st.out(out_kind, n);
    if (st.rc != AHV_OK) return st.rc;
    AHV_TRY(ahv_chain_capture_dev(ctx, d_arena, ab, d_off, d_hp, d_rp, S, n, d_dig, d_par, d_kind));
    st.back(arena, d_arena, ab);
    st.back(out_digests, d_dig, 32 * n);
    st.back(out_parents, d_par, 32 * n);
    st.back(out_kind, d_kind, n);
    return st.finish();
}

int ahv_sigma_eff_rings(ahv_ctx *ctx, const uint32_t *rowptr, const double *bonded, const uint8_t *active,
                        const int64_t *expiry, int64_t now, const double *sigma, const double *omega,
                        const uint8_t *consensus, const uint8_t *current_ring, uint64_t A, double *out_sigma,
                        uint8_t *out_ring, uint8_t *out_demote, uint8_t *out_near)
{
    AHV_TRY(use_device(ctx));
    if (A == 0) return AHV_OK;
    AHV_REQUIRE(rowptr && sigma && omega && out_sigma, "rowptr / sigma / omega / out_sigma_eff is null");
    const uint64_t E = rowptr[A];
    AHV_REQUIRE(E == 0 || (bonded && active && expiry), "edge arrays are null");
    AHV_TRY(io_reserve(ctx, p256(4 * (A + 1)) + p256(8 * E) + p256(E) + p256(8 * E) + 3 * p256(8 * A) + 5 * p256(A)));
    Stager st(ctx);
    uint32_t *d_rp = st.in(rowptr, A + 1);
    double *d_b = st.in(bonded, E);
    uint8_t *d_a = st.in(active, E);
    int64_t *d_e = st.in(expiry, E);
    double *d_s = st.in(sigma, A);
    double *d_o = st.in(omega, A);
    uint8_t *d_c = st.in(consensus, A);
    uint8_t *d_r = st.in(current_ring, A);
    double *d_os = st.out(out_sigma, A);
    uint8_t *d_or = st.out(out_ring, A);
    uint8_t *d_od = st.out(out_demote, A);
    uint8_t *d_on = st.out(out_near, A);
    if (st.rc != AHV_OK) return st.rc;
    ctx->sigma_edges_hint = E;                    // known here: the tiled kernel sizes its tiles for the average row
    const int rc_dev = ahv_sigma_eff_rings_dev(ctx, d_rp, d_b, d_a, d_e, now, d_s, d_o, d_c, d_r, A, d_os, d_or, d_od, d_on);
    ctx->sigma_edges_hint = 0;
    AHV_TRY(rc_dev);
    st.back(out_sigma, d_os, A);
    st.back(out_ring, d_or, A);
    st.back(out_demote, d_od, A);
    st.back(out_near, d_on, A);
    return st.finish();
}

int ahv_exposure_sums(ahv_ctx *ctx, const uint32_t *rowptr, const double *bonded, const uint8_t *active,
                      const int64_t *expiry, int64_t now, uint64_t n_rows, int compensated, double *out_sum)
{
    AHV_TRY(use_device(ctx));
    if (n_rows == 0) return AHV_OK;
    AHV_REQUIRE(rowptr && out_sum, "rowptr / out_sum is null");
    const uint64_t E = rowptr[n_rows];
    AHV_REQUIRE(E == 0 || (bonded && active && expiry), "edge arrays are null");
    AHV_TRY(io_reserve(ctx, p256(4 * (n_rows + 1)) + 2 * p256(8 * E) + p256(E) + p256(8 * n_rows)));
    Stager st(ctx);
    uint32_t *d_rp = st.in(rowptr, n_rows + 1);
    double *d_b = st.in(bonded, E);
    uint8_t *d_a = st.in(active, E);
    int64_t *d_e = st.in(expiry, E);
    double *d_o = st.out(out_sum, n_rows);
    if (st.rc != AHV_OK) return st.rc;
    AHV_TRY(ahv_exposure_sums_dev(ctx, d_rp, d_b, d_a, d_e, now, n_rows, compensated, d_o));
    st.back(out_sum, d_o, n_rows);
    return st.finish();
}

int ahv_slash_cascade(ahv_ctx *ctx, const ahv_slash_batch *b)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(b, "batch is null");
    const uint64_t S = b->n_sessions, A = b->n_agents;
    if (S == 0) return AHV_OK;
    AHV_REQUIRE(b->rowptr && b->session_agent_rowptr && b->event_rowptr, "rowptr / session_agent_rowptr / event_rowptr is null");
    AHV_REQUIRE(b->session_agent_rowptr[S] == A, "session_agent_rowptr[n_sessions] != n_agents");
    AHV_REQUIRE(A < 0xffffffffull, "more than 2^32 - 2 agent rows");
    const uint64_t E = b->rowptr[A], NE = b->event_rowptr[S];
    AHV_REQUIRE(E == 0 || (b->voucher && b->active && b->expiry && b->clip_log), "voucher / active / expiry / clip_log is null");
    AHV_REQUIRE(NE == 0 || (b->event_agent && b->event_sigma && b->event_omega), "event arrays are null");
    AHV_REQUIRE(A == 0 || (b->scores && b->touched), "scores / touched is null");
    AHV_REQUIRE((NE + E == 0 || b->slash_log) && b->slash_count && b->clip_count && b->status,
                "slash_log / slash_count / clip_count / status is null");
    for (uint64_t s = 0; s < S; ++s)
        AHV_REQUIRE(b->session_agent_rowptr[s] <= b->session_agent_rowptr[s + 1] && b->event_rowptr[s] <= b->event_rowptr[s + 1],
                    "session_agent_rowptr / event_rowptr must not decrease");
    AHV_TRY(io_reserve(ctx, p256(4 * (A + 1)) + p256(4 * E) + p256(E) + p256(8 * E) + 2 * p256(8 * (S + 1)) + p256(4 * NE) +
                                2 * p256(8 * NE) + p256(NE) + p256(8 * A) + p256(A) + p256(sizeof(ahv_slash_record) * (NE + E)) +
                                p256(sizeof(ahv_clip_record) * E) + 2 * p256(4 * S) + p256(S) + 4096));
    Stager st(ctx);
    ahv_slash_batch d = *b;
    d.rowptr = st.in(b->rowptr, A + 1);
    d.voucher = st.in(b->voucher, E);
    d.active = st.in(b->active, E);
    d.expiry = st.in(b->expiry, E);
    d.session_agent_rowptr = st.in(b->session_agent_rowptr, S + 1);
    d.event_rowptr = st.in(b->event_rowptr, S + 1);
    d.event_agent = st.in(b->event_agent, NE);
    d.event_sigma = st.in(b->event_sigma, NE);
    d.event_omega = st.in(b->event_omega, NE);
    d.event_depth = st.in(b->event_depth, NE);
    d.scores = st.in(b->scores, A);
    d.touched = st.out(b->touched, A);
    d.slash_log = st.out(b->slash_log, NE + E);
    d.clip_log = st.out(b->clip_log, E);
    d.slash_count = st.out(b->slash_count, S);
    d.clip_count = st.out(b->clip_count, S);
    d.status = st.out(b->status, S);
    if (st.rc != AHV_OK) return st.rc;
    if (A) AHV_CUDA(cudaMemsetAsync(d.touched, 0, A, ctx->stream));
    AHV_TRY(slash_cascade_launch(ctx, &d, E));      // E known here: the lane kernel's slice is sized for the average session
    st.back(b->active, d.active, E);
    st.back(b->scores, d.scores, A);
    st.back(b->touched, d.touched, A);
    st.back(b->slash_log, d.slash_log, NE + E);
    st.back(b->clip_log, d.clip_log, E);
    st.back(b->slash_count, d.slash_count, S);
    st.back(b->clip_count, d.clip_count, S);
    st.back(b->status, d.status, S);
    return st.finish();
}

int ahv_ring_check(ahv_ctx *ctx, const uint8_t *agent_ring, const uint8_t *flags, const double *sigma,
                   const uint8_t *consensus, const uint8_t *witness, uint64_t n, uint8_t *out_allowed,
                   uint8_t *out_reason, uint8_t *out_required)
{
    AHV_TRY(use_device(ctx));
    if (n == 0) return AHV_OK;
    AHV_REQUIRE(agent_ring && flags && sigma && consensus && witness, "an input array is null");
    AHV_TRY(io_reserve(ctx, 7 * p256(n) + p256(8 * n)));
    Stager st(ctx);
    uint8_t *d_ar = st.in(agent_ring, n);
    uint8_t *d_fl = st.in(flags, n);
    double *d_s = st.in(sigma, n);
    uint8_t *d_c = st.in(consensus, n);
    uint8_t *d_w = st.in(witness, n);
    uint8_t *d_al = st.out(out_allowed, n);
    uint8_t *d_re = st.out(out_reason, n);
    uint8_t *d_rq = st.out(out_required, n);
    if (st.rc != AHV_OK) return st.rc;
    AHV_TRY(ahv_ring_check_dev(ctx, d_ar, d_fl, d_s, d_c, d_w, n, d_al, d_re, d_rq));
    st.back(out_allowed, d_al, n);
    st.back(out_reason, d_re, n);
    st.back(out_required, d_rq, n);
    return st.finish();
}

}  // extern "C"

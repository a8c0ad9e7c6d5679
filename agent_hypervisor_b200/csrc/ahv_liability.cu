// ahv_liability.cu - batched sigma_eff over a CSR vouch graph + ring classification,
// and the batched RingEnforcer gate.  HBM-bound passes (fp64 compare/add only).
//
// Reference behaviour (paths under /root/reference/src/hypervisor):
//   sigma_eff        liability/vouching.py:128-151 (+ :34-38 expiry, :186-197 active filter)
//   ring             models.py:34-42 ; should_demote rings/enforcer.py:134-137
//   required_ring    models.py:122-132 ; gate rings/enforcer.py:44-128
#include "ahv_internal.cuh"
#include "liability_core.cuh"

using namespace ahv;

namespace {

// One thread per agent row; the row is summed sequentially in insertion order so the
// result is bit-identical to the reference's Python loop (vouching.py:146-148).
__global__ void __launch_bounds__(256)
k_sigma_eff_rings(const uint32_t *__restrict__ rowptr, const double *__restrict__ bonded,
                  const uint8_t *__restrict__ active, const int64_t *__restrict__ expiry, int64_t now,
                  const double *__restrict__ sigma, const double *__restrict__ omega,
                  const uint8_t *__restrict__ consensus, const uint8_t *__restrict__ current_ring, uint64_t A,
                  double *__restrict__ out_sigma, uint8_t *__restrict__ out_ring, uint8_t *__restrict__ out_demote,
                  uint8_t *__restrict__ out_near)
{
    uint64_t a = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= A) return;
    const uint32_t e0 = rowptr[a], e1 = rowptr[a + 1];
    double contrib = 0.0;
    for (uint32_t e = e0; e < e1; ++e)
        contrib = vouch_accumulate(contrib, bonded[e], active[e], expiry[e], now);
    const double se = sigma_eff_of(sigma[a], omega[a], contrib);
    out_sigma[a] = se;
    if (out_ring) out_ring[a] = (uint8_t)ring_from_sigma(se, consensus[a] != 0);
    if (out_demote) out_demote[a] = (uint8_t)(ring_from_sigma(se, false) > (int)current_ring[a]);
    if (out_near) out_near[a] = (uint8_t)near_threshold(se);
}

// Tiled form (default): a CTA owns 256 consecutive agent rows.  Their edges are one contiguous
// CSR range, streamed through shared memory in chunks with fully coalesced loads (phase 1: each
// edge becomes its contribution, bonded or +0.0 when inactive/expired); then every thread adds
// up the slice of its own row inside the chunk, in CSR order (phase 2).  Order of additions per
// row is exactly the reference's, so results stay bit-identical; x + (+0.0) == x for every x the
// running sum can hold (it starts at +0.0 and can never become -0.0).
//
// A row of d edges is d dependent fp64 additions (8.1 cycles each), so one very long row holds its
// whole tile back: 14 017 edges = 58 us at best, longer than the rest of a 10^7-edge graph takes.
// With PAR, rows longer than `long_cut` edges are therefore summed by the whole CTA in a FIXED shape
// that depends on the row alone (not on tiling, chunking or scheduling, so it is the same run to run):
//     partial[p] = live contributions at row positions p, p+256, p+512, ... added in that order   (p = 0..255)
//     warp w     = xor-butterfly over partial[32w .. 32w+31] with strides 16, 8, 4, 2, 1
//     row sum    = ((((((w0 + w1) + w2) + w3) + w4) + w5) + w6) + w7
// which differs from the reference's left-to-right sum by a few ulp (north_star allows 1e-6
// relative); agents whose sigma_eff lies within 1e-6 of a ring threshold are flagged in out_near
// either way.  Rows up to `long_cut` keep the reference's order and stay bit-identical; option
// sigma_exact=1 keeps it for every row.
constexpr int SIG_ROWS = 256;
constexpr int SIG_CHUNK = 2048;
constexpr int SIG_WARPS = SIG_ROWS / 32;

// Running sum of one row.  Plain: left-to-right fp64 additions (the explicit += loop of
// vouching.py:146-148, and Python's sum() up to CPython 3.11).  Compensated: CPython >= 3.12's
// sum() of floats (Neumaier's variant of Kahan-Babuska, Python/bltinmodule.c), which is what
// get_total_exposure / total_exposure evaluate to under the interpreter the reference runs on.
template <bool COMPENSATED> struct RowSum;
template <> struct RowSum<false> {
    double f = 0.0;
    __device__ __forceinline__ void add(double x) { f = __dadd_rn(f, x); }
    __device__ __forceinline__ double result() const { return f; }
};
template <> struct RowSum<true> {
    double f = 0.0, c = 0.0;
    __device__ __forceinline__ void add(double x)
    {
        const double t = __dadd_rn(f, x);
        if (fabs(f) >= fabs(x)) c = __dadd_rn(c, __dadd_rn(__dsub_rn(f, t), x));
        else c = __dadd_rn(c, __dadd_rn(__dsub_rn(x, t), f));
        f = t;
    }
    __device__ __forceinline__ double result() const
    {
        return (c != 0.0 && isfinite(c)) ? __dadd_rn(f, c) : f;
    }
};

// MODE 0: sigma_eff + rings.  MODE 1: plain live sum of each row.  MODE 2: compensated live sum.
// PAR (MODE 0 only): rows of more than long_cut edges take the fixed-shape parallel sum.
template <int MODE, bool PAR>
__global__ void __launch_bounds__(SIG_ROWS)
k_sigma_eff_tiled(const uint32_t *__restrict__ rowptr, const double *__restrict__ bonded,
                  const uint8_t *__restrict__ active, const int64_t *__restrict__ expiry, int64_t now,
                  const double *__restrict__ sigma, const double *__restrict__ omega,
                  const uint8_t *__restrict__ consensus, const uint8_t *__restrict__ current_ring, uint64_t A,
                  double *__restrict__ out_sigma, uint8_t *__restrict__ out_ring, uint8_t *__restrict__ out_demote,
                  uint8_t *__restrict__ out_near, uint32_t long_cut)
{
    __shared__ double v[SIG_CHUNK];
    __shared__ double wsum[SIG_WARPS];
    __shared__ uint32_t wcount[SIG_WARPS];
    __shared__ uint32_t long_r0[SIG_ROWS], long_r1[SIG_ROWS];
    __shared__ uint16_t long_owner[SIG_ROWS];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const uint64_t a0 = (uint64_t)blockIdx.x * SIG_ROWS;
    const uint64_t a = a0 + tid;
    const bool valid = a < A;
    const uint64_t a_last = (a0 + SIG_ROWS < A) ? a0 + SIG_ROWS : A;
    const uint32_t e_begin = rowptr[a0], e_end = rowptr[a_last];
    const uint32_t my0 = valid ? rowptr[a] : e_end, my1 = valid ? rowptr[a + 1] : e_end;
    RowSum<MODE == 2> acc;

    // list of this tile's long rows, in row order (usually empty)
    const bool is_long = PAR && (my1 - my0 > long_cut);
    uint32_t n_long = 0;
    if (PAR) {
        const uint32_t b = __ballot_sync(0xffffffffu, is_long);
        if (lane == 0) wcount[wid] = __popc(b);
        __syncthreads();
        uint32_t before = 0;
#pragma unroll
        for (int w = 0; w < SIG_WARPS; ++w) {
            const uint32_t c = wcount[w];
            if (w < (int)wid) before += c;
            n_long += c;
        }
        if (is_long) {
            const uint32_t k = before + __popc(b & ((1u << lane) - 1u));
            long_r0[k] = my0;
            long_r1[k] = my1;
            long_owner[k] = (uint16_t)tid;
        }
        // (made visible by the barrier after phase 1 of the first chunk: a long row has edges, so there is one)
    }
    uint32_t li = 0;          // next long row (uniform over the CTA)
    double part = 0.0;        // this thread's strided partial of long row li

    for (uint32_t c0 = e_begin; c0 < e_end; c0 += SIG_CHUNK) {
        const uint32_t c1 = (e_end - c0 > (uint32_t)SIG_CHUNK) ? c0 + SIG_CHUNK : e_end;
        // phase 1: issue every load of this thread's 8 edges before using any (memory-level parallelism)
        double b_[SIG_CHUNK / SIG_ROWS];
        int64_t x_[SIG_CHUNK / SIG_ROWS];
        uint8_t a_[SIG_CHUNK / SIG_ROWS];
#pragma unroll
        for (int k = 0; k < SIG_CHUNK / SIG_ROWS; ++k) {
            const uint32_t e = c0 + tid + k * SIG_ROWS;
            const bool in = e < c1;
            b_[k] = in ? __ldg(bonded + e) : 0.0;
            x_[k] = in ? __ldg(expiry + e) : AHV_EXPIRY_NONE;
            a_[k] = in ? __ldg(active + e) : (uint8_t)0;
        }
#pragma unroll
        for (int k = 0; k < SIG_CHUNK / SIG_ROWS; ++k)
            v[tid + k * SIG_ROWS] = vouch_accumulate(0.0, b_[k], a_[k], x_[k], now);
        __syncthreads();
        // phase 2: own row, reference order; values fetched four at a time ahead of the dependent additions
        if (!is_long) {
            const uint32_t lo = my0 > c0 ? my0 : c0, hi = my1 < c1 ? my1 : c1;
            uint32_t e = lo;
            for (; e + 4 <= hi; e += 4) {
                const double x0 = v[e - c0], x1 = v[e - c0 + 1], x2 = v[e - c0 + 2], x3 = v[e - c0 + 3];
                acc.add(x0); acc.add(x1); acc.add(x2); acc.add(x3);
            }
            for (; e < hi; ++e) acc.add(v[e - c0]);
        }
        if (PAR) {
            while (li < n_long) {
                const uint32_t r0 = long_r0[li], r1 = long_r1[li];
                if (r0 >= c1) break;                                   // starts in a later chunk
                const uint32_t lo = r0 > c0 ? r0 : c0, hi = r1 < c1 ? r1 : c1;
                // row positions congruent to tid (mod 256) inside [lo, hi)
                for (uint32_t e = lo + ((tid - (lo - r0)) & (SIG_ROWS - 1)); e < hi; e += SIG_ROWS)
                    part = __dadd_rn(part, v[e - c0]);
                if (r1 > c1) break;                                    // continues in the next chunk
                double x = part;
#pragma unroll
                for (int o = 16; o; o >>= 1) x = __dadd_rn(x, __shfl_xor_sync(0xffffffffu, x, o));
                if (lane == 0) wsum[wid] = x;
                __syncthreads();
                if (tid == long_owner[li]) {
                    double t = wsum[0];
#pragma unroll
                    for (int w = 1; w < SIG_WARPS; ++w) t = __dadd_rn(t, wsum[w]);
                    acc.f = t;
                }
                __syncthreads();
                part = 0.0;
                ++li;
            }
        }
        __syncthreads();
    }
    if (!valid) return;
    const double contrib = acc.result();
    if (MODE != 0) {
        out_sigma[a] = contrib;
        return;
    }
    const double se = sigma_eff_of(sigma[a], omega[a], contrib);
    out_sigma[a] = se;
    if (out_ring) out_ring[a] = (uint8_t)ring_from_sigma(se, consensus[a] != 0);
    if (out_demote) out_demote[a] = (uint8_t)(ring_from_sigma(se, false) > (int)current_ring[a]);
    if (out_near) out_near[a] = (uint8_t)near_threshold(se);
}

__global__ void __launch_bounds__(256)
k_ring_check(const uint8_t *__restrict__ agent_ring, const uint8_t *__restrict__ flags,
             const double *__restrict__ sigma, const uint8_t *__restrict__ consensus,
             const uint8_t *__restrict__ witness, uint64_t n, uint8_t *__restrict__ allowed,
             uint8_t *__restrict__ reason, uint8_t *__restrict__ required)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int req = required_ring(flags[i]);
    int code = gate_reason(agent_ring[i], req, sigma[i], consensus[i] != 0, witness[i] != 0);
    if (allowed) allowed[i] = (uint8_t)(code == AHV_REASON_GRANTED);
    if (reason) reason[i] = (uint8_t)code;
    if (required) required[i] = (uint8_t)req;
}

}  // namespace

extern "C" {

int ahv_sigma_eff_rings_dev(ahv_ctx *ctx, const uint32_t *d_rowptr, const double *d_bonded, const uint8_t *d_active,
                            const int64_t *d_expiry, int64_t now, const double *d_sigma, const double *d_omega,
                            const uint8_t *d_consensus, const uint8_t *d_current_ring, uint64_t A,
                            double *d_out_sigma, uint8_t *d_out_ring, uint8_t *d_out_demote, uint8_t *d_out_near)
{
    AHV_TRY(use_device(ctx));
    if (A == 0) return AHV_OK;
    AHV_REQUIRE(d_rowptr && d_sigma && d_omega && d_out_sigma, "rowptr / sigma / omega / out_sigma_eff is null");
    AHV_REQUIRE(!d_out_ring || d_consensus, "out_ring needs consensus");
    AHV_REQUIRE(!d_out_demote || d_current_ring, "out_demote needs current_ring");
    AHV_REQUIRE(A < 0xffffffffull, "too many agents (limit 2^32-2)");
    if (ctx->opt_sigma_variant == 0)
        k_sigma_eff_rings<<<grid_for(A, 256), 256, 0, ctx->stream>>>(d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma,
                                                                     d_omega, d_consensus, d_current_ring, A, d_out_sigma,
                                                                     d_out_ring, d_out_demote, d_out_near);
    else if (ctx->opt_sigma_exact)
        k_sigma_eff_tiled<0, false><<<grid_for(A, SIG_ROWS), SIG_ROWS, 0, ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, 0u);
    else
        k_sigma_eff_tiled<0, true><<<grid_for(A, SIG_ROWS), SIG_ROWS, 0, ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

int ahv_exposure_sums_dev(ahv_ctx *ctx, const uint32_t *d_rowptr, const double *d_bonded, const uint8_t *d_active,
                          const int64_t *d_expiry, int64_t now, uint64_t n_rows, int compensated, double *d_out_sum)
{
    AHV_TRY(use_device(ctx));
    if (n_rows == 0) return AHV_OK;
    AHV_REQUIRE(d_rowptr && d_out_sum, "rowptr / out_sum is null");
    AHV_REQUIRE(n_rows < 0xffffffffull, "too many rows (limit 2^32-2)");
    const unsigned g = grid_for(n_rows, SIG_ROWS);
    if (compensated)
        k_sigma_eff_tiled<2, false><<<g, SIG_ROWS, 0, ctx->stream>>>(d_rowptr, d_bonded, d_active, d_expiry, now, nullptr,
                                                                     nullptr, nullptr, nullptr, n_rows, d_out_sum, nullptr,
                                                                     nullptr, nullptr, 0u);
    else
        k_sigma_eff_tiled<1, false><<<g, SIG_ROWS, 0, ctx->stream>>>(d_rowptr, d_bonded, d_active, d_expiry, now, nullptr,
                                                                     nullptr, nullptr, nullptr, n_rows, d_out_sum, nullptr,
                                                                     nullptr, nullptr, 0u);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

int ahv_ring_check_dev(ahv_ctx *ctx, const uint8_t *d_agent_ring, const uint8_t *d_flags, const double *d_sigma,
                       const uint8_t *d_consensus, const uint8_t *d_witness, uint64_t n, uint8_t *d_allowed,
                       uint8_t *d_reason, uint8_t *d_required)
{
    AHV_TRY(use_device(ctx));
    if (n == 0) return AHV_OK;
    AHV_REQUIRE(d_agent_ring && d_flags && d_sigma && d_consensus && d_witness, "an input array is null");
    k_ring_check<<<grid_for(n, 256), 256, 0, ctx->stream>>>(d_agent_ring, d_flags, d_sigma, d_consensus, d_witness, n,
                                                            d_allowed, d_reason, d_required);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

}  // extern "C"

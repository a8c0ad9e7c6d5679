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

// Tiled form (default): a CTA of 128 threads owns 128 consecutive agent rows.  Their edges are one
// contiguous CSR range; it arrives in shared memory in chunks of up to 1440 edges by THREE bulk
// asynchronous copies per chunk (cp.async.bulk, the 1-D TMA path: one thread issues them, the bytes land
// without passing through registers and complete on an mbarrier), so the copy costs next to no issue
// slots and eight CTAs fit an SM, their copy and sum phases overlapping.  Most tiles fit one chunk: one
// wait, then every thread adds up its own row in CSR order.  The order of additions per row is exactly
// the reference's (vouching.py:146-148), so results are bit-identical.  A record that is not live is
// skipped, which equals adding +0.0: the running sum starts at +0.0 and can never become -0.0.
// The bulk copies move whole 16-edge groups (16-byte alignment of all three arrays): a chunk grid starts
// at the 16-edge boundary at or below the tile's first edge (those extra edges lie inside the arrays),
// and the last, partial group of the tile is fetched with ordinary loads so nothing past the arrays'
// ends is ever read.
//
// A row of d edges is d dependent fp64 additions, so one long row holds its whole tile back (a row of
// 14 017 edges alone costs more than the rest of a 10^7-edge graph).  With PAR, rows longer than
// `long_cut` edges are therefore summed by the WARP that owns them (all 32 lanes on one row at a time, found
// with a ballot, no shared list and no CTA barrier), in a FIXED shape that depends on the row alone
// (not on tiling, chunking or scheduling, so it is the same run to run):
//     partial[p] = live amounts at row positions p, p+32, p+64, ... added in that order   (p = 0..31)
//     row sum    = xor-butterfly over the 32 partials with strides 16, 8, 4, 2, 1
//                  (every lane ends with the same value; lane 0's is used)
// which differs from the reference's left-to-right sum by a few ulp (north_star allows 1e-6 relative);
// agents whose sigma_eff lies within 1e-6 of a ring threshold are flagged in out_near either way.  Rows
// up to `long_cut` keep the reference's order and stay bit-identical; option sigma_exact=1 keeps it for
// every row.
constexpr int SIG_ROWS = 128;
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

template <int CHUNK>
struct alignas(128) SigStage {         // one chunk of edges as it lies in global memory
    double b[CHUNK];                   // after the conversion pass: the amount each record contributes
    int64_t x[CHUNK];
    uint8_t a[CHUNK];
};

__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint32_t mbar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(bytes), "r"(mbar) : "memory");
}

// amount a record contributes: bonded when live (vouching.py:34-38,186-197), else +0.0.
// "expiry is not None and now > expiry" with None encoded as INT64_MAX is just "now > expiry" (nothing is greater
// than INT64_MAX), so live = active and expiry >= now: one 64-bit compare per record.
static_assert(AHV_EXPIRY_NONE == INT64_MAX, "the live test relies on the encoding of 'no expiry'");
__device__ __forceinline__ double live_amount(double b, uint32_t active, int64_t x, int64_t now)
{
    return (active != 0u && x >= now) ? b : 0.0;
}

// Work sharing between the CTAs of one launch.  A tile whose rows hold far more edges than a chunk (one row of
// 14 017 records = 7 chunks, one after the other) takes several times longer than an ordinary tile; if its
// CTA is scheduled late, the whole launch waits for it.  So the CTAs of the first wave look through ALL
// tiles' edge counts while their own first chunk is in flight (a handful of loads each) and put the heavy
// ones in a queue; a CTA that has finished its own tile takes heavy tiles from the queue, so they start
// within the first microseconds.  Every heavy tile is claimed (atomic exchange with this launch's number) by
// whoever gets to it first - its own CTA or one that took it from the queue - so it is processed exactly
// once whatever the timing; should a CTA ever read a queue slot before its writer has filled it, it finds an
// older launch's tile number there, and processing an ordinary tile twice only writes the same values twice.
// Launches alternate between two queues, each launch zeroing the other one's counters, so nothing needs
// clearing between launches and a graph without heavy tiles pays one extra load per CTA.  None of this
// changes any result.
struct SigWork {
    uint32_t *counters;        // [2][2]: per queue, tail and head
    uint32_t *queue;           // [2][cap]; NULL = no work sharing
    uint32_t *claimed;         // per tile: number of the launch that claimed it last
    uint32_t cap;
    uint32_t epoch;            // this launch's number (never 0); its queue is epoch & 1
    uint32_t heavy;            // a tile with more edges than this goes through the queue
};
constexpr uint32_t SIG_NONE = 0xffffffffu;
constexpr uint32_t SIG_SCAN_CTAS = 512;         // CTAs that share the look-through (all resident in the first wave)
constexpr uint32_t SIG_QUEUE_CAP = 4096;
constexpr uint32_t SIG_AHEAD = 1024;            // look-ahead distance of the L2 prefetches, in tiles (about one wave of CTAs)

// MODE 0: sigma_eff + rings.  MODE 1: plain live sum of each row.  MODE 2: compensated live sum.
// PAR (MODE 0 only): rows of more than long_cut edges take the fixed-shape warp sum.
template <int MODE, bool PAR, int SIG_CHUNK>
__global__ void __launch_bounds__(SIG_ROWS, SIG_CHUNK == 2048 ? 6 : (SIG_CHUNK >= 1440 ? 8 : (SIG_CHUNK == 1296 ? 9 : 10)))
k_sigma_eff_tiled(const uint32_t *__restrict__ rowptr, const double *__restrict__ bonded,
                  const uint8_t *__restrict__ active, const int64_t *__restrict__ expiry, int64_t now,
                  const double *__restrict__ sigma, const double *__restrict__ omega,
                  const uint8_t *__restrict__ consensus, const uint8_t *__restrict__ current_ring, uint64_t A,
                  double *__restrict__ out_sigma, uint8_t *__restrict__ out_ring, uint8_t *__restrict__ out_demote,
                  uint8_t *__restrict__ out_near, uint32_t long_cut, SigWork wk, uint32_t rows_log2)
{
    extern __shared__ __align__(128) uint8_t sig_smem[];
    SigStage<SIG_CHUNK> &st = *reinterpret_cast<SigStage<SIG_CHUNK> *>(sig_smem);
    // Rows per tile: 128 (one per thread) unless the graph's rows are long on average - then a tile of 128 rows would
    // be dozens of chunks walked by one CTA, and a small graph would not even fill the SMs - in which case a tile
    // has R = 2^rows_log2 rows, owned by every (128 / R)-th thread so that they stay spread over the four warps.
    const uint32_t R = 1u << rows_log2, own_shift = 7u - rows_log2;
    __shared__ __align__(8) uint64_t mbar_storage;
    __shared__ uint32_t s_word;
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&mbar_storage);
    const uint32_t ntiles = gridDim.x;
    if (tid == 0) {
        asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(mbar), "r"(1u) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    uint32_t *const q_cnt = wk.queue ? wk.counters + 2 * (wk.epoch & 1u) : nullptr;
    uint32_t *const q_ent = wk.queue ? wk.queue + (size_t)wk.cap * (wk.epoch & 1u) : nullptr;
    if (wk.queue && blockIdx.x == 0 && tid < 2) wk.counters[2 * ((wk.epoch & 1u) ^ 1u) + tid] = 0u;   // the next launch's queue
    bool first = true;
    __syncthreads();                                      // mbarrier initialised
    uint32_t phase = 0;                                   // chunks this CTA has waited for so far (mbarrier parity)
    uint32_t tile = blockIdx.x;
    while (tile != SIG_NONE) {
        const uint64_t a0 = (uint64_t)tile * R;
        const uint64_t a = a0 + (tid >> own_shift);
        const bool valid = (tid & ((1u << own_shift) - 1u)) == 0u && a < A;
        const uint64_t a_last = (a0 + R < A) ? a0 + R : A;
        const uint32_t e_begin = __ldg(rowptr + a0), e_end = __ldg(rowptr + a_last);
        const uint32_t my0 = valid ? __ldg(rowptr + a) : 0u, my1 = valid ? __ldg(rowptr + a + 1) : 0u;
        // Look-ahead through L2 (first tile of each CTA; CTAs start roughly in tile order): the row pointers of
        // the tile SIG_AHEAD tiles further on were put into L2 by an earlier CTA, so they arrive quickly; when this
        // tile is done, that tile's edges and per-agent inputs are sent for (bulk L2 prefetch), and the row pointers
        // of the tile 2 * SIG_AHEAD further on.  A CTA then finds its inputs in L2 instead of waiting for DRAM
        // three times in a row.
        uint32_t f0 = 0, f1 = 0;
        const uint64_t tile_f = (uint64_t)tile + SIG_AHEAD;
        if (first && tid == 0 && tile_f < ntiles) {
            const uint64_t r0 = tile_f * R, r1 = (r0 + R < A) ? r0 + R : A;
            f0 = __ldg(rowptr + r0);
            f1 = __ldg(rowptr + r1);
        }
        // how many heavy tiles are queued: asked for now, looked at when this tile is done
        uint32_t q_tail = 0;
        if (wk.queue && tid == 0) q_tail = *reinterpret_cast<volatile uint32_t *>(&q_cnt[0]);
        bool take = true;
        if (wk.queue && e_end - e_begin > wk.heavy) {          // heavy: first come, first served
            if (tid == 0) s_word = atomicExch(&wk.claimed[tile], wk.epoch);
            __syncthreads();
            take = (s_word != wk.epoch);
            __syncthreads();                                          // s_word is reused below
        }
        if (take) {
            // chunk grid from the 16-edge boundary at or below the first edge; copies issued by thread 0
            const uint32_t base = e_begin & ~15u;
            const uint32_t nchunk = (e_end > base) ? (e_end - base + SIG_CHUNK - 1) / SIG_CHUNK : 0;
            auto issue = [&](uint32_t k) {                            // thread 0 only
                const uint32_t c0 = base + k * SIG_CHUNK;
                const uint32_t cnt = (e_end - c0 > (uint32_t)SIG_CHUNK) ? (uint32_t)SIG_CHUNK : e_end - c0;
                const uint32_t cnt16 = cnt & ~15u;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(cnt16 * 17u) : "memory");
                if (cnt16) {
                    bulk_g2s(st.b, bonded + c0, cnt16 * 8u, mbar);
                    bulk_g2s(st.x, expiry + c0, cnt16 * 8u, mbar);
                    bulk_g2s(st.a, active + c0, cnt16, mbar);
                }
            };
            if (tid == 0 && nchunk) issue(0);
            // look-through (first tile of the first SIG_SCAN_CTAS CTAs, while the chunk is in flight):
            // CTA b checks tiles b + i * SIG_SCAN_CTAS, one per thread
            if (wk.queue && first && blockIdx.x < SIG_SCAN_CTAS) {
                for (uint32_t t = blockIdx.x + tid * SIG_SCAN_CTAS; t < ntiles; t += SIG_ROWS * SIG_SCAN_CTAS) {
                    const uint64_t r0 = (uint64_t)t * R, r1 = (r0 + R < A) ? r0 + R : A;
                    if (__ldg(rowptr + r1) - __ldg(rowptr + r0) > wk.heavy) {
                        const uint32_t slot = atomicAdd(&q_cnt[0], 1u);
                        if (slot < wk.cap) q_ent[slot] = t;
                    }
                }
            }

            // per-agent inputs are requested now, long before they are needed
            double sg = 0.0, om = 0.0;
            uint8_t cons = 0, ring_in = 0;
            if (MODE == 0 && valid) {
                sg = __ldg(sigma + a);
                om = __ldg(omega + a);
                if (out_ring) cons = __ldg(consensus + a);
                if (out_demote) ring_in = __ldg(current_ring + a);
            }

            RowSum<MODE == 2> acc;
            // long rows are taken by the warp that owns them, one after the other, all 32 lanes on each: found with
            // a ballot, the row bounds passed round with shuffles - no shared list, no CTA barrier for them
            const bool is_long = PAR && (my1 - my0 > long_cut);
            uint32_t long_mask = PAR ? __ballot_sync(0xffffffffu, is_long) : 0u;     // the warp's long rows still to finish
            double part = 0.0;        // this lane's strided partial of the long row in progress

            for (uint32_t k = 0; k < nchunk; ++k, ++phase) {
                const uint32_t c0 = base + k * SIG_CHUNK;
                const uint32_t c1 = (e_end - c0 > (uint32_t)SIG_CHUNK) ? c0 + SIG_CHUNK : e_end;
                // the last, partial 16-edge group of the tile: ordinary loads (a bulk copy would read past e_end)
                const uint32_t cnt16 = (c1 - c0) & ~15u;
                if (tid < (c1 - c0) - cnt16) {
                    const uint32_t i = cnt16 + tid;
                    st.b[i] = __ldg(bonded + c0 + i);
                    st.x[i] = __ldg(expiry + c0 + i);
                    st.a[i] = __ldg(active + c0 + i);
                }
                {   // wait for this chunk's bytes
                    uint32_t done = 0;
                    while (!done)
                        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                                     "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(mbar), "r"(phase & 1u) : "memory");
                }
                if (cnt16 != c1 - c0) __syncthreads();            // the tail elements were written by other threads
                // conversion pass, all lanes busy: b[i] <- the amount record i contributes (two records per step: 16-byte
                // shared-memory accesses at a 16-byte stride; four per step was measured slower - bank conflicts)
                for (uint32_t i = 2 * tid; i < c1 - c0; i += 2 * SIG_ROWS) {
                    double2 b2 = *reinterpret_cast<const double2 *>(&st.b[i]);
                    const longlong2 x2 = *reinterpret_cast<const longlong2 *>(&st.x[i]);
                    const uint32_t a2 = *reinterpret_cast<const uint16_t *>(&st.a[i]);
                    b2.x = live_amount(b2.x, a2 & 0xffu, x2.x, now);
                    b2.y = live_amount(b2.y, a2 >> 8, x2.y, now);         // (one slot past an odd count: never read)
                    *reinterpret_cast<double2 *>(&st.b[i]) = b2;
                }
                __syncthreads();
                if (!is_long) {                                   // own row, reference order
                    const uint32_t lo = my0 > c0 ? my0 : c0, hi = my1 < c1 ? my1 : c1;
                    uint32_t e = lo;
                    for (; e + 4 <= hi; e += 4) {                 // operands fetched ahead of the dependent additions
                        const double x0 = st.b[e - c0], x1 = st.b[e - c0 + 1], x2 = st.b[e - c0 + 2], x3 = st.b[e - c0 + 3];
                        acc.add(x0); acc.add(x1); acc.add(x2); acc.add(x3);
                    }
                    for (; e < hi; ++e) acc.add(st.b[e - c0]);
                }
                if (PAR) {
                    while (long_mask) {                           // in row order
                        const int owner = __ffs(long_mask) - 1;
                        const uint32_t r0 = __shfl_sync(0xffffffffu, my0, owner), r1 = __shfl_sync(0xffffffffu, my1, owner);
                        if (r0 >= c1) break;                                   // starts in a later chunk
                        const uint32_t lo = r0 > c0 ? r0 : c0, hi = r1 < c1 ? r1 : c1;
                        // row positions congruent to lane (mod 32) inside [lo, hi)
                        for (uint32_t e = lo + ((lane - (lo - r0)) & 31u); e < hi; e += 32)
                            part = __dadd_rn(part, st.b[e - c0]);
                        if (r1 > c1) break;                                    // continues in the next chunk
                        double x = part;
#pragma unroll
                        for (int o = 16; o; o >>= 1) x = __dadd_rn(x, __shfl_xor_sync(0xffffffffu, x, o));
                        if ((int)lane == owner) acc.f = x;
                        part = 0.0;
                        long_mask &= long_mask - 1u;
                    }
                }
                if (k + 1 < nchunk) {
                    __syncthreads();                              // every warp is done with the stage
                    if (tid == 0) issue(k + 1);
                }
            }
            if (valid) {
                const double contrib = acc.result();
                if (MODE != 0) {
                    out_sigma[a] = contrib;
                } else {
                    const double se = sigma_eff_of(sg, om, contrib);
                    out_sigma[a] = se;
                    if (out_ring) out_ring[a] = (uint8_t)ring_from_sigma(se, cons != 0);
                    if (out_demote) out_demote[a] = (uint8_t)(ring_from_sigma(se, false) > (int)ring_in);
                    if (out_near) out_near[a] = (uint8_t)near_threshold(se);
                }
            }
        }
        if (first && tid < 8) {
            const uint64_t tile_ff = (uint64_t)tile + 2 * SIG_AHEAD;
            if (tid < 4 && tile_ff < ntiles) {                        // 128 row pointers = four 128-byte lines
                const uint64_t r = tile_ff * R + 32u * tid;
                if (r <= A) asm volatile("prefetch.global.L2 [%0];" ::"l"(rowptr + r));
            }
            if (tile_f < ntiles) {
                const uint64_t r0 = tile_f * R;
                if (tid == 0 && f1 > f0) {
                    const uint32_t p0 = f0 & ~15u;
                    uint32_t cnt = (f1 - p0) & ~15u;
                    if (cnt > 4096u) cnt = 4096u;                     // a heavy tile's later chunks are left to themselves
                    if (cnt) {
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(bonded + p0), "r"(cnt * 8u) : "memory");
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(expiry + p0), "r"(cnt * 8u) : "memory");
                        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(active + p0), "r"(cnt) : "memory");
                    }
                }
                if (MODE == 0 && tid >= 4) {                          // sigma and omega of 128 agents: 8 lines each
                    const uint64_t r = r0 + 32u * (tid - 4u);         // two lines per thread and array
                    if (r < A) {
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(sigma + r));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(omega + r));
                    }
                    if (r + 16 < A) {
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(sigma + r + 16));
                        asm volatile("prefetch.global.L2 [%0];" ::"l"(omega + r + 16));
                    }
                }
            }
        }
        first = false;
        if (!wk.queue) break;
        // next: a heavy tile from the queue, if any is waiting (thread 0 knows; nothing queued = one barrier and out)
        if (!__syncthreads_or(tid == 0 && q_tail != 0u)) break;      // (also: everyone is done with the stage and the lists)
        if (tid == 0) {
            uint32_t next = SIG_NONE;
            const uint32_t tail = min(q_tail, wk.cap);
            if (tail && *reinterpret_cast<volatile uint32_t *>(&q_cnt[1]) < tail) {
                const uint32_t slot = atomicAdd(&q_cnt[1], 1u);
                if (slot < tail) next = *reinterpret_cast<volatile uint32_t *>(&q_ent[slot]);
                if (next >= ntiles) next = SIG_NONE;      // a slot nobody has ever written
            }
            s_word = next;
        }
        __syncthreads();
        tile = s_word;
        __syncthreads();
    }
}

// Streaming form (option sigma_variant = 2; an experiment kept for comparison, measured SLOWER than the tiled
// form: 53 vs 39 us on the uniform config-4 graph, 76 vs 57 us on the Zipf one - with two stages per CTA only
// four CTAs = 16 warps fit an SM, and the conversion / summing phases, which are chains of dependent
// shared-memory loads and fp64 additions, then have too few warps to overlap with): the same tiles, sums and
// results as k_sigma_eff_tiled, but the
// CTAs are persistent (four per SM) and each runs a two-stage pipeline: while the 128 threads convert and sum
// the chunk in one shared-memory stage, the bulk copies of the NEXT chunk - the same tile's, or the first one
// of the CTA's next tile - are already filling the other stage, and the row pointers of the tiles after that
// are on their way into registers.  Tiles are handed out by an atomic counter (a CTA that meets a heavy tile
// simply comes back for fewer), fetched three tiles ahead so that no one waits for the counter.  A CTA pays
// the DRAM latency of row pointers -> edges -> per-agent inputs once, at its start, instead of once per tile.
constexpr int SIG_SCH = 1536;                    // edges per stage
constexpr uint32_t SIG_STREAM_CTAS_PER_SM = 4;

template <bool PAR>
__global__ void __launch_bounds__(SIG_ROWS, SIG_STREAM_CTAS_PER_SM)
k_sigma_eff_stream(const uint32_t *__restrict__ rowptr, const double *__restrict__ bonded,
                   const uint8_t *__restrict__ active, const int64_t *__restrict__ expiry, int64_t now,
                   const double *__restrict__ sigma, const double *__restrict__ omega,
                   const uint8_t *__restrict__ consensus, const uint8_t *__restrict__ current_ring, uint64_t A,
                   double *__restrict__ out_sigma, uint8_t *__restrict__ out_ring, uint8_t *__restrict__ out_demote,
                   uint8_t *__restrict__ out_near, uint32_t long_cut, uint32_t ntiles, uint32_t *tick, uint32_t epoch)
{
    extern __shared__ __align__(128) uint8_t sig_smem[];
    SigStage<SIG_SCH> *stages = reinterpret_cast<SigStage<SIG_SCH> *>(sig_smem);
    __shared__ __align__(8) uint64_t mbar_storage[2];
    __shared__ double long_sum[SIG_ROWS];
    __shared__ uint32_t long_r0[SIG_ROWS], long_r1[SIG_ROWS];
    __shared__ uint16_t long_owner[SIG_ROWS];
    __shared__ uint32_t wcount[SIG_WARPS];
    __shared__ uint32_t s_ids[4];                         // tile numbers of this CTA's tiles j, j+1, j+2, j+3 (ring)
    const uint32_t tid = threadIdx.x, lane = tid & 31u, wid = tid >> 5;
    const uint32_t mbar0 = (uint32_t)__cvta_generic_to_shared(&mbar_storage[0]);
    uint32_t *const ctr = tick + (epoch & 1u);            // tiles handed out beyond the first gridDim.x
    if (tid == 0) {
        asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(mbar0), "r"(1u) : "memory");
        asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(mbar0 + 8u), "r"(1u) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (blockIdx.x == 0) tick[(epoch & 1u) ^ 1u] = 0u;            // the next launch's counter
        s_ids[0] = blockIdx.x;
        s_ids[1] = gridDim.x + atomicAdd(ctr, 1u);
        s_ids[2] = gridDim.x + atomicAdd(ctr, 1u);
    }
    __syncthreads();

    auto tile_end_row = [&](uint32_t t) { const uint64_t r = (uint64_t)t * SIG_ROWS + SIG_ROWS; return r < A ? r : A; };
    // thread 0: the bulk copies of chunk k of a tile whose edges are [e0, e1), into stage s
    auto issue = [&](uint32_t e0, uint32_t e1, uint32_t k, uint32_t s) {
        const uint32_t mb = mbar0 + 8u * s;
        const uint32_t base = e0 & ~15u;
        uint32_t cnt16 = 0, c0 = base + k * SIG_SCH;
        if (e1 > c0) {
            const uint32_t cnt = (e1 - c0 > (uint32_t)SIG_SCH) ? (uint32_t)SIG_SCH : e1 - c0;
            cnt16 = cnt & ~15u;
        }
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(cnt16 * 17u) : "memory");
        if (cnt16) {
            bulk_g2s(stages[s].b, bonded + c0, cnt16 * 8u, mb);
            bulk_g2s(stages[s].x, expiry + c0, cnt16 * 8u, mb);
            bulk_g2s(stages[s].a, active + c0, cnt16, mb);
        }
    };
    auto nchunks = [&](uint32_t e0, uint32_t e1) {        // at least one (an empty tile is one empty chunk)
        const uint32_t base = e0 & ~15u;
        const uint32_t c = (e1 > base) ? (e1 - base + SIG_SCH - 1) / SIG_SCH : 0u;
        return c ? c : 1u;
    };

    // prologue: tile 0 of this CTA (its bounds are waited for here, once), bounds of tile 1 requested
    uint32_t id = s_ids[0], id1 = s_ids[1], id2 = s_ids[2];
    uint32_t cb0 = __ldg(rowptr + (uint64_t)id * SIG_ROWS), cb1 = __ldg(rowptr + tile_end_row(id));
    uint32_t nb0 = 0, nb1 = 0;
    if (id1 < ntiles) {
        nb0 = __ldg(rowptr + (uint64_t)id1 * SIG_ROWS);
        nb1 = __ldg(rowptr + tile_end_row(id1));
    }
    uint32_t my0 = 0, my1 = 0;
    {
        const uint64_t a = (uint64_t)id * SIG_ROWS + tid;
        if (a < A) {
            my0 = __ldg(rowptr + a);
            my1 = __ldg(rowptr + a + 1);
        }
    }
    if (tid == 0) issue(cb0, cb1, 0, 0);
    uint32_t n = 0;                                       // chunks consumed so far: stage n & 1, parity (n >> 1) & 1

    for (uint32_t j = 0;; ++j) {
        // ---- requests for later tiles (consumed one tile from now)
        uint32_t n2b0 = 0, n2b1 = 0, nmy0 = 0, nmy1 = 0;
        if (id2 < ntiles) {
            n2b0 = __ldg(rowptr + (uint64_t)id2 * SIG_ROWS);
            n2b1 = __ldg(rowptr + tile_end_row(id2));
        }
        if (id1 < ntiles) {
            const uint64_t a = (uint64_t)id1 * SIG_ROWS + tid;
            if (a < A) {
                nmy0 = __ldg(rowptr + a);
                nmy1 = __ldg(rowptr + a + 1);
            }
        }
        if (tid == 0) s_ids[(j + 3u) & 3u] = (id2 < ntiles) ? gridDim.x + atomicAdd(ctr, 1u) : SIG_NONE;
        // ---- this tile
        const uint64_t a = (uint64_t)id * SIG_ROWS + tid;
        const bool valid = a < A;
        double sg = 0.0, om = 0.0;
        uint8_t cons = 0, ring_in = 0;
        if (valid) {
            sg = __ldg(sigma + a);
            om = __ldg(omega + a);
            if (out_ring) cons = __ldg(consensus + a);
            if (out_demote) ring_in = __ldg(current_ring + a);
        }
        const uint32_t e_end = cb1, base = cb0 & ~15u, nchunk = nchunks(cb0, cb1);
        RowSum<false> acc;
        const bool is_long = PAR && (my1 - my0 > long_cut);
        uint32_t n_long = 0;
        if (PAR) {
            const uint32_t bal = __ballot_sync(0xffffffffu, is_long);
            if (lane == 0) wcount[wid] = __popc(bal);
            __syncthreads();
            uint32_t before = 0;
#pragma unroll
            for (int w = 0; w < SIG_WARPS; ++w) {
                const uint32_t c = wcount[w];
                if (w < (int)wid) before += c;
                n_long += c;
            }
            if (is_long) {
                const uint32_t k = before + __popc(bal & ((1u << lane) - 1u));
                long_r0[k] = my0;
                long_r1[k] = my1;
                long_owner[k] = (uint16_t)tid;
            }
            // (read after the first chunk's conversion barrier)
        }
        uint32_t lk = wid;
        double part = 0.0;
        for (uint32_t k = 0; k < nchunk; ++k, ++n) {
            SigStage<SIG_SCH> &st = stages[n & 1u];
            // the next chunk's copies go out now, into the stage the previous chunk has left
            if (tid == 0) {
                if (k + 1 < nchunk) issue(cb0, cb1, k + 1, (n + 1u) & 1u);
                else if (id1 < ntiles) issue(nb0, nb1, 0, (n + 1u) & 1u);
            }
            const uint32_t c0 = base + k * SIG_SCH;
            const uint32_t c1 = (e_end <= c0) ? c0 : ((e_end - c0 > (uint32_t)SIG_SCH) ? c0 + SIG_SCH : e_end);
            const uint32_t cnt16 = (c1 - c0) & ~15u;
            if (tid < (c1 - c0) - cnt16) {                            // last, partial 16-edge group: ordinary loads
                const uint32_t i = cnt16 + tid;
                st.b[i] = __ldg(bonded + c0 + i);
                st.x[i] = __ldg(expiry + c0 + i);
                st.a[i] = __ldg(active + c0 + i);
            }
            {
                uint32_t done = 0;
                const uint32_t mb = mbar0 + 8u * (n & 1u);
                while (!done)
                    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                                 "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(mb), "r"((n >> 1) & 1u) : "memory");
            }
            if (cnt16 != c1 - c0) __syncthreads();
            for (uint32_t i = 2 * tid; i < c1 - c0; i += 2 * SIG_ROWS) {
                double2 b2 = *reinterpret_cast<const double2 *>(&st.b[i]);
                const longlong2 x2 = *reinterpret_cast<const longlong2 *>(&st.x[i]);
                const uint32_t a2 = *reinterpret_cast<const uint16_t *>(&st.a[i]);
                b2.x = live_amount(b2.x, a2 & 0xffu, x2.x, now);
                b2.y = live_amount(b2.y, a2 >> 8, x2.y, now);
                *reinterpret_cast<double2 *>(&st.b[i]) = b2;
            }
            __syncthreads();
            if (!is_long) {
                const uint32_t lo = my0 > c0 ? my0 : c0, hi = my1 < c1 ? my1 : c1;
                uint32_t e = lo;
                for (; e + 4 <= hi; e += 4) {
                    const double x0 = st.b[e - c0], x1 = st.b[e - c0 + 1], x2 = st.b[e - c0 + 2], x3 = st.b[e - c0 + 3];
                    acc.add(x0); acc.add(x1); acc.add(x2); acc.add(x3);
                }
                for (; e < hi; ++e) acc.add(st.b[e - c0]);
            }
            if (PAR) {
                while (lk < n_long) {
                    const uint32_t r0 = long_r0[lk], r1 = long_r1[lk];
                    if (r0 >= c1) break;
                    const uint32_t lo = r0 > c0 ? r0 : c0, hi = r1 < c1 ? r1 : c1;
                    for (uint32_t e = lo + ((lane - (lo - r0)) & 31u); e < hi; e += 32)
                        part = __dadd_rn(part, st.b[e - c0]);
                    if (r1 > c1) break;
                    double x = part;
#pragma unroll
                    for (int o = 16; o; o >>= 1) x = __dadd_rn(x, __shfl_xor_sync(0xffffffffu, x, o));
                    if (lane == 0) long_sum[long_owner[lk]] = x;
                    part = 0.0;
                    lk += SIG_WARPS;
                }
            }
            __syncthreads();                              // the stage is free for the chunk after next; long_sum is complete
        }
        if (valid) {
            if (is_long) acc.f = long_sum[tid];
            const double se = sigma_eff_of(sg, om, acc.result());
            out_sigma[a] = se;
            if (out_ring) out_ring[a] = (uint8_t)ring_from_sigma(se, cons != 0);
            if (out_demote) out_demote[a] = (uint8_t)(ring_from_sigma(se, false) > (int)ring_in);
            if (out_near) out_near[a] = (uint8_t)near_threshold(se);
        }
        if (id1 >= ntiles) break;
        // ---- on to the CTA's next tile
        id = id1; id1 = id2; id2 = s_ids[(j + 3u) & 3u];                 // written before this tile's barriers
        cb0 = nb0; cb1 = nb1; nb0 = n2b0; nb1 = n2b1;
        my0 = nmy0; my1 = nmy1;
    }
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

namespace {

// the queue / claim flags of the tiled sigma_eff kernel: allocated once per context (regrown when a larger
// graph arrives), left in its empty state by every launch
int sigma_work(ahv_ctx *ctx, uint64_t ntiles, SigWork *out)
{
    const uint64_t off_queue = 256, off_claimed = off_queue + pad256(2 * 4ull * SIG_QUEUE_CAP);
    if (ntiles > ctx->sigma_work_tiles) {
        AHV_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->sigma_work) AHV_CUDA(cudaFree(ctx->sigma_work));
        ctx->sigma_work = nullptr;
        ctx->sigma_work_tiles = 0;
        const uint64_t tiles = ntiles + ntiles / 4 + 1024;
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, off_claimed + 4ull * tiles);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("device allocation for the sigma_eff work queue failed: %s", cudaGetErrorString(e));
            return AHV_E_NOMEM;
        }
        ctx->sigma_work = (uint8_t *)p;
        ctx->sigma_work_tiles = tiles;
        AHV_CUDA(cudaMemsetAsync(ctx->sigma_work, 0, off_claimed + 4ull * tiles, ctx->stream));
        AHV_CUDA(cudaMemsetAsync(ctx->sigma_work + off_queue, 0xff, 2 * 4ull * SIG_QUEUE_CAP, ctx->stream));
        ctx->sigma_epoch = 0;
        ctx->sigma_tick_epoch = 0;
    }
    if (++ctx->sigma_epoch == 0) ctx->sigma_epoch = 1;    // 0 is what a never-claimed tile holds
    out->counters = reinterpret_cast<uint32_t *>(ctx->sigma_work);
    out->queue = reinterpret_cast<uint32_t *>(ctx->sigma_work + off_queue);
    out->claimed = reinterpret_cast<uint32_t *>(ctx->sigma_work + off_claimed);
    out->cap = SIG_QUEUE_CAP;
    out->epoch = ctx->sigma_epoch;
    out->heavy = (uint32_t)ctx->opt_sigma_heavy;
    return AHV_OK;
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
    // the tiled kernel moves 16-edge groups with bulk copies: all three edge arrays must be 16-byte aligned
    // (anything from cudaMalloc / ahv_dev_alloc is); otherwise every row is summed by its own thread, in the
    // reference's order
    const bool aligned = ((reinterpret_cast<uintptr_t>(d_bonded) | reinterpret_cast<uintptr_t>(d_expiry) |
                           reinterpret_cast<uintptr_t>(d_active)) & 15u) == 0;
    SigWork wk = {nullptr, nullptr, nullptr, 0, 0, 0};
    // rows per tile of the tiled kernel: option sigma_rows, else from the average row length when the caller knows the
    // record count (the host entry does), else 128
    uint32_t rl = 7;
    if (ctx->opt_sigma_rows) {
        rl = 0;
        while ((2u << rl) <= (uint32_t)ctx->opt_sigma_rows) ++rl;
    } else if (ctx->sigma_edges_hint) {
        const uint64_t avg = (ctx->sigma_edges_hint + A - 1) / A;
        while (rl > 0 && ((uint64_t)avg << rl) > 2048) --rl;
    }
    const uint64_t R = 1ull << rl;
    if (ctx->opt_sigma_variant == 1 && aligned && ctx->opt_sigma_share) AHV_TRY(sigma_work(ctx, grid_for(A, (unsigned)R), &wk));
    if (ctx->opt_sigma_variant == 2 && aligned) {
        AHV_TRY(sigma_work(ctx, grid_for(A, SIG_ROWS), &wk));
        const uint32_t ntiles = grid_for(A, SIG_ROWS);
        const uint32_t resident = (uint32_t)ctx->sm_count * SIG_STREAM_CTAS_PER_SM;
        const uint32_t g = ntiles < resident ? ntiles : resident;
        const size_t smem = 2 * sizeof(SigStage<SIG_SCH>);
        const uint32_t tick_epoch = ++ctx->sigma_tick_epoch;       // streaming launches alternate between two tile counters
        static bool attr_seen[64][2] = {};                    // per device: function attributes belong to a device's context
        const int which = ctx->opt_sigma_exact ? 0 : 1;
        bool *attr_set = attr_seen[ctx->device & 63];
        if (!attr_set[which]) {
            if (which) AHV_CUDA(cudaFuncSetAttribute(k_sigma_eff_stream<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            else AHV_CUDA(cudaFuncSetAttribute(k_sigma_eff_stream<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set[which] = true;
        }
        if (which)
            k_sigma_eff_stream<true><<<g, SIG_ROWS, smem, ctx->stream>>>(
                d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
                d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut, ntiles, wk.counters + 4, tick_epoch);
        else
            k_sigma_eff_stream<false><<<g, SIG_ROWS, smem, ctx->stream>>>(
                d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
                d_out_ring, d_out_demote, d_out_near, 0u, ntiles, wk.counters + 4, tick_epoch);
    } else if (ctx->opt_sigma_variant == 0 || !aligned)
        k_sigma_eff_rings<<<grid_for(A, 256), 256, 0, ctx->stream>>>(d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma,
                                                                     d_omega, d_consensus, d_current_ring, A, d_out_sigma,
                                                                     d_out_ring, d_out_demote, d_out_near);
    else if (ctx->opt_sigma_exact)
        k_sigma_eff_tiled<0, false, 2048><<<grid_for(A, (unsigned)R), SIG_ROWS, sizeof(SigStage<2048>), ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, 0u, wk, rl);
    else if (ctx->opt_sigma_chunk == 1024)
        k_sigma_eff_tiled<0, true, 1024><<<grid_for(A, (unsigned)R), SIG_ROWS, sizeof(SigStage<1024>), ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut, wk, rl);
    else if (ctx->opt_sigma_chunk == 1296)
        k_sigma_eff_tiled<0, true, 1296><<<grid_for(A, (unsigned)R), SIG_ROWS, sizeof(SigStage<1296>), ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut, wk, rl);
    else if (ctx->opt_sigma_chunk == 1536)
        k_sigma_eff_tiled<0, true, 1536><<<grid_for(A, (unsigned)R), SIG_ROWS, sizeof(SigStage<1536>), ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut, wk, rl);
    else if (ctx->opt_sigma_chunk != 2048)
        k_sigma_eff_tiled<0, true, 1440><<<grid_for(A, (unsigned)R), SIG_ROWS, sizeof(SigStage<1440>), ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut, wk, rl);
    else
        k_sigma_eff_tiled<0, true, 2048><<<grid_for(A, (unsigned)R), SIG_ROWS, sizeof(SigStage<2048>), ctx->stream>>>(
            d_rowptr, d_bonded, d_active, d_expiry, now, d_sigma, d_omega, d_consensus, d_current_ring, A, d_out_sigma,
            d_out_ring, d_out_demote, d_out_near, (uint32_t)ctx->opt_sigma_long_cut, wk, rl);
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
    AHV_REQUIRE(((reinterpret_cast<uintptr_t>(d_bonded) | reinterpret_cast<uintptr_t>(d_expiry) |
                  reinterpret_cast<uintptr_t>(d_active)) & 15u) == 0, "bonded / active / expiry must be 16-byte aligned");
    const unsigned g = grid_for(n_rows, SIG_ROWS);
    if (compensated)
        k_sigma_eff_tiled<2, false, 2048><<<g, SIG_ROWS, sizeof(SigStage<2048>), ctx->stream>>>(d_rowptr, d_bonded, d_active, d_expiry, now, nullptr,
                                                                     nullptr, nullptr, nullptr, n_rows, d_out_sum, nullptr,
                                                                     nullptr, nullptr, 0u, SigWork{nullptr, nullptr, nullptr, 0, 0, 0}, 7u);
    else
        k_sigma_eff_tiled<1, false, 2048><<<g, SIG_ROWS, sizeof(SigStage<2048>), ctx->stream>>>(d_rowptr, d_bonded, d_active, d_expiry, now, nullptr,
                                                                     nullptr, nullptr, nullptr, n_rows, d_out_sum, nullptr,
                                                                     nullptr, nullptr, 0u, SigWork{nullptr, nullptr, nullptr, 0, 0, 0}, 7u);
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

// Collateral slashing with bounded cascade (reference: liability/slashing.py:53-141) for many
// sessions in one launch.
//
// One warp per session.  A slash of vouchee v (slashing.py:62-141):
//   scores[v] = 0.0                                                         (:88)
//   for every live vouch record e of row v, in insertion order              (:90-92, vouching.py:153-155)
//       before = scores[voucher[e]]; after = max(before * (1.0 - omega), 0.05)     (:94-96)
//       scores[voucher[e]] = after; log the clip; release the bond (active[e] = 0)  (:97-108)
//   then, while depth < 2, for each of THOSE clips in order with after < 0.05 + 0.01 whose voucher
//   still has a live vouch record: slash that voucher one level deeper      (:122-138)
// The 32 lanes take 32 records of the row at a time.  Records of one chunk that name the same voucher
// are the only ones that depend on each other; a lane of rank r inside such a group applies the clip
// r times to get its `before` (the function is deterministic, so this equals the sequential result)
// and the last lane of the group stores the score.  Chunks run in order, so later chunks read what
// earlier ones wrote.  The cascade is the reference's depth-first order with an explicit two-entry
// stack over the clip log, which doubles as the output.
//
// The cascade is a chain of dependent look-ups (row bounds -> records -> scores -> next row ...), a DRAM
// round trip each when done in global memory.  The warp kernel has two views of a session: staged in its
// warp's slice of shared memory when it fits (13 B per agent + 17 B per record; slash_variant 2 only, see
// below) and on global memory (GlobalView), where rows are long enough to keep the lanes busy and the
// number of resident warps hides the latency.  Both views produce the same records.
//
// Small and mid-size sessions (a handful of records per row) would leave most of a warp's lanes idle, so
// k_slash_lanes runs first: a warp takes 32 consecutive sessions, cuts them into runs that fit its slice of
// shared memory (11 B per agent + 2 B per record), stages a run's contiguous rows, records and scores with
// coalesced reads, then every lane walks ITS session serially out of shared memory (LaneRun in
// slashing_core.cuh) and the warp writes back what changed.  Sessions so large that fewer than two fit a
// slice are listed for k_slash_cascade, which then only visits that list, on global memory.
#include "ahv_internal.cuh"
#include "slashing_core.cuh"

namespace ahv {
namespace {

using namespace ahv::slashing;

// Session state in global memory.  Agents and records are addressed by their index inside the session.
struct GlobalView {
    const uint32_t *rowptr;       // + a_lo
    const uint32_t *voucher;      // + e_base
    uint8_t *active;              // + e_base
    const int64_t *expiry;        // + e_base
    double *scores;               // + a_lo
    uint8_t *touched;             // + a_lo
    const ahv_clip_record *cl;    // the session's clip records
    int64_t now;
    uint32_t a_lo, n_agents, e_base;

    __device__ __forceinline__ uint32_t row_begin(uint32_t v) const { return rowptr[v] - e_base; }
    __device__ __forceinline__ uint32_t row_end(uint32_t v) const { return rowptr[v + 1] - e_base; }
    // F_LIVE / F_BAD_VOUCHER of record e and its voucher (vouching.py:34-38,191-196)
    __device__ __forceinline__ uint32_t flags(uint32_t e, uint32_t &u) const
    {
        const int64_t x = expiry[e];
        if (active[e] == 0 || (x != AHV_EXPIRY_NONE && now > x)) return 0;
        u = voucher[e] - a_lo;                    // wraps far out of range for a voucher below a_lo
        return u < n_agents ? F_LIVE : F_BAD_VOUCHER;
    }
    __device__ __forceinline__ void release(uint32_t e) { active[e] = 0; }
    __device__ __forceinline__ double score(uint32_t u) const { return scores[u]; }
    __device__ __forceinline__ void set_score(uint32_t u, double x)
    {
        scores[u] = x;
        touched[u] = 1;
    }
    __device__ __forceinline__ void note_clip(uint32_t, uint32_t, double) {}
    __device__ __forceinline__ double clip_after(uint32_t c) const { return cl[c].sigma_after; }
    __device__ __forceinline__ uint32_t clip_voucher(uint32_t c) const { return voucher[cl[c].edge - e_base] - a_lo; }
};

// Session state staged in the warp's slice of shared memory.
struct SharedView {
    double *s_score;              // [A]
    double *s_after;              // [E]  sigma_after of clip c
    uint32_t *s_rowptr;           // [A+1] record indices inside the session
    uint32_t *s_voucher;          // [E]  agent index inside the session
    uint32_t *s_clip_voucher;     // [E]  voucher of clip c
    uint8_t *s_flags;             // [E]
    uint8_t *s_touched;           // [A]

    static __device__ __forceinline__ uint64_t bytes(uint32_t A, uint32_t E) { return 13ull * A + 17ull * E + 4u; }
    __device__ __forceinline__ void carve(uint8_t *base, uint32_t A, uint32_t E)
    {
        s_score = reinterpret_cast<double *>(base);
        s_after = s_score + A;
        s_rowptr = reinterpret_cast<uint32_t *>(s_after + E);
        s_voucher = s_rowptr + (A + 1);
        s_clip_voucher = s_voucher + E;
        s_flags = reinterpret_cast<uint8_t *>(s_clip_voucher + E);
        s_touched = s_flags + E;
    }
    __device__ __forceinline__ uint32_t row_begin(uint32_t v) const { return s_rowptr[v]; }
    __device__ __forceinline__ uint32_t row_end(uint32_t v) const { return s_rowptr[v + 1]; }
    __device__ __forceinline__ uint32_t flags(uint32_t e, uint32_t &u) const
    {
        u = s_voucher[e];
        return s_flags[e] & (F_LIVE | F_BAD_VOUCHER);
    }
    __device__ __forceinline__ void release(uint32_t e) { s_flags[e] = F_WAS_LIVE; }
    __device__ __forceinline__ double score(uint32_t u) const { return s_score[u]; }
    __device__ __forceinline__ void set_score(uint32_t u, double x)
    {
        s_score[u] = x;
        s_touched[u] = 1;
    }
    __device__ __forceinline__ void note_clip(uint32_t c, uint32_t u, double after)
    {
        s_clip_voucher[c] = u;
        s_after[c] = after;
    }
    __device__ __forceinline__ double clip_after(uint32_t c) const { return s_after[c]; }
    __device__ __forceinline__ uint32_t clip_voucher(uint32_t c) const { return s_clip_voucher[c]; }
};

struct SessionLog {
    ahv_slash_record *sl;
    ahv_clip_record *cl;
    uint32_t n_slash, n_clip;
    uint32_t a_lo, e_base;
    uint32_t status;
};

// get_vouchers_for(u) non-empty?  (slashing.py:126-129)
template <class View>
__device__ bool row_any_live(const View &m, uint32_t u, int lane)
{
    const uint32_t end = m.row_end(u);
    for (uint32_t base = m.row_begin(u); base < end; base += 32) {
        const uint32_t e = base + lane;
        uint32_t w;
        const bool l = e < end && (m.flags(e, w) & F_LIVE);
        if (__any_sync(FULL, l)) return true;
    }
    return false;
}

// One SlashingEngine.slash call without its cascade; returns the index of its record.
template <class View>
__device__ uint32_t slash_once(View &m, SessionLog &g, int lane, uint32_t v, double sigma_before, double keep, uint32_t depth,
                               int32_t parent, uint32_t ev)
{
    if (lane == 0) m.set_score(v, 0.0);           // :88
    __syncwarp();
    const uint32_t me = g.n_slash++;
    const uint32_t c0 = g.n_clip;
    const uint32_t end = m.row_end(v);
    for (uint32_t base = m.row_begin(v); base < end; base += 32) {
        const uint32_t e = base + lane;
        uint32_t u = 0;
        const uint32_t fl = e < end ? m.flags(e, u) : 0u;
        if (fl & F_BAD_VOUCHER) g.status |= 2;    // a live record whose voucher is outside the session: left alone
        const bool live = (fl & F_LIVE) != 0;
        const unsigned mask = __ballot_sync(FULL, live);
        // lanes naming the same voucher form a group; idle lanes get a key no live lane can have
        const unsigned peers = __match_any_sync(FULL, live ? (unsigned long long)u : ((1ull << 32) | (unsigned)lane));
        double before = 0.0, after = 0.0;
        if (live) {
            const int rank = __popc(peers & ((1u << lane) - 1u));
            before = m.score(u);                  // :94 (a missing key is a dense 0.0)
            for (int r = 0; r < rank; ++r) before = clip_once(before, keep);
            after = clip_once(before, keep);
        }
        __syncwarp();                             // every lane has read its score before any lane writes
        if (live) {
            if ((peers >> lane) == 1u) m.set_score(u, after);       // last lane of its group, :97
            m.release(e);                         // :108 release_bond
            const uint32_t c = g.n_clip + __popc(mask & ((1u << lane) - 1u));
            m.note_clip(c, u, after);
            ahv_clip_record *q = g.cl + c;        // :99-105
            q->edge = g.e_base + e;
            q->slash = me;
            q->sigma_before = before;
            q->sigma_after = after;
        }
        g.n_clip += __popc(mask);
        __syncwarp();
    }
    if (lane == 0) {
        ahv_slash_record *r = g.sl + me;          // :110-119
        r->agent = g.a_lo + v;
        r->parent = parent;
        r->event = ev;
        r->depth = depth;
        r->clip_begin = c0;
        r->clip_count = g.n_clip - c0;
        r->sigma_before = sigma_before;
    }
    return me;
}

// Every event of one session, in order, with the reference's depth-first cascade.
template <class View>
__device__ void run_events(const SlashArgs &k, View &m, SessionLog &g, int lane, uint64_t s, uint32_t n_agents)
{
    const double thresh = add_rn(SIGMA_FLOOR, 0.01);                      // :124
    for (uint64_t ev = k.ev_rp[s]; ev < k.ev_rp[s + 1]; ++ev) {
        const uint32_t v0 = k.ev_agent[ev] - g.a_lo;
        if (v0 >= n_agents) {                     // an event naming an agent of another session
            g.status |= 1;
            continue;
        }
        const double keep = add_rn(1.0, -k.ev_omega[ev]);                // (1.0 - risk_weight), :95
        const uint32_t depth0 = k.ev_depth ? k.ev_depth[ev] : 0u;         // the caller's cascade_depth argument (:70)
        uint32_t f_cur[MAX_CASCADE_DEPTH], f_end[MAX_CASCADE_DEPTH], f_id[MAX_CASCADE_DEPTH];
        f_cur[0] = g.n_clip;
        f_id[0] = slash_once(m, g, lane, v0, k.ev_sigma[ev], keep, depth0, -1, (uint32_t)ev);
        f_end[0] = g.n_clip;
        int d = depth0 < (uint32_t)MAX_CASCADE_DEPTH ? 0 : -1;
        while (d >= 0) {                          // :122-138, depth first, clips in order
            if (f_cur[d] == f_end[d]) {
                --d;
                continue;
            }
            const uint32_t c = f_cur[d]++;
            const double after = m.clip_after(c);
            if (!(after < thresh)) continue;
            const uint32_t u = m.clip_voucher(c);
            if (!row_any_live(m, u, lane)) continue;
            const uint32_t c0 = g.n_clip;
            const uint32_t id = slash_once(m, g, lane, u, after, keep, depth0 + d + 1, (int32_t)f_id[d], (uint32_t)ev);
            if (depth0 + d + 1 < (uint32_t)MAX_CASCADE_DEPTH) {
                ++d;
                f_cur[d] = c0;
                f_end[d] = g.n_clip;
                f_id[d] = id;
            }
        }
    }
}

__global__ void __launch_bounds__(32 * LANE_WARPS)
k_slash_lanes(const SlashArgs k)
{
    extern __shared__ __align__(16) uint8_t slash_smem[];
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const uint64_t n_groups = (k.n_sessions + 31) / 32;
    const uint64_t n_warps = (uint64_t)gridDim.x * LANE_WARPS;
    for (uint64_t grp = (uint64_t)blockIdx.x * LANE_WARPS + wib; grp < n_groups; grp += n_warps) {
        const uint64_t s0 = grp * 32;
        const uint32_t n = (uint32_t)(k.n_sessions - s0 < 32 ? k.n_sessions - s0 : 32);
        // lane i looks after session s0 + i: its first agent row / record / event, and the ones after it
        const uint64_t s = s0 + ((uint32_t)lane < n ? lane : n - 1);
        const uint32_t my_a0 = (uint32_t)k.sess_rp[s], my_a1 = (uint32_t)k.sess_rp[s + 1];
        const uint64_t ev_lo = k.ev_rp[s], ev_hi = k.ev_rp[s + 1];
        const uint32_t my_e0 = k.rowptr[my_a0], my_e1 = k.rowptr[my_a1];
        uint32_t p = 0;
        while (p < n) {
            const uint32_t a_lo = __shfl_sync(FULL, my_a0, p), e_lo = __shfl_sync(FULL, my_e0, p);
            // a session of which fewer than LANE_MIN_SESSIONS would fit a slice is k_slash_cascade's (a whole
            // warp for it)
            if (lane_group_bytes(__shfl_sync(FULL, my_a1, p) - a_lo, __shfl_sync(FULL, my_e1, p) - e_lo) >
                k.lane_slice / LANE_MIN_SESSIONS) {
                if (lane == 0) k.big_sessions[1 + atomicAdd(k.big_sessions, 1u)] = (uint32_t)(s0 + p);
                ++p;
                continue;
            }
            // the longest run of sessions p .. q-1 whose state fits the slice (sizes grow with q: a prefix)
            const bool fit = (uint32_t)lane >= p && (uint32_t)lane < n &&
                             lane_group_bytes(my_a1 - a_lo, my_e1 - e_lo) <= k.lane_slice;
            const uint32_t cnt = __popc(__ballot_sync(FULL, fit));
            const uint32_t q = p + cnt;
            const uint32_t A = __shfl_sync(FULL, my_a1, q - 1) - a_lo, E = __shfl_sync(FULL, my_e1, q - 1) - e_lo;
            LaneRun r;
            stage_run(r, slash_smem + (size_t)wib * k.lane_slice, k, a_lo, A, e_lo, E, (uint32_t)lane, 32u);
            __syncwarp();
            if ((uint32_t)lane >= p && (uint32_t)lane < q) run_session(r, k, s, a_lo, e_lo, my_a0, my_a1, my_e0, ev_lo, ev_hi);
            __syncwarp();
            write_back_run(r, k, a_lo, A, e_lo, E, (uint32_t)lane, 32u);
            __syncwarp();                         // the slice is reused by the next run
            p = q;
        }
    }
}

template <bool STAGE>
__global__ void __launch_bounds__(32 * SLASH_WARPS)
k_slash_cascade(const SlashArgs k, const int from_list)
{
    extern __shared__ __align__(16) uint8_t slash_smem[];
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const uint64_t warp = (uint64_t)blockIdx.x * SLASH_WARPS + wib;
    const uint64_t n_warps = (uint64_t)gridDim.x * SLASH_WARPS;
    // with the lane kernel in front, only the sessions of the groups it listed are left
    const uint64_t n_items = from_list ? (uint64_t)k.big_sessions[0] : k.n_sessions;
    for (uint64_t w = warp; w < n_items; w += n_warps) {
        const uint64_t s = from_list ? (uint64_t)k.big_sessions[1 + w] : w;
        const uint32_t a_lo = (uint32_t)k.sess_rp[s];
        const uint32_t A = (uint32_t)k.sess_rp[s + 1] - a_lo;
        const uint32_t e_base = k.rowptr[a_lo];
        const uint32_t E = k.rowptr[a_lo + A] - e_base;
        SessionLog g;
        g.a_lo = a_lo;
        g.e_base = e_base;
        g.sl = k.slash_log + (k.ev_rp[s] + e_base);
        g.cl = k.clip_log + e_base;
        g.n_slash = g.n_clip = 0;
        g.status = 0;
        if (k.ev_rp[s] == k.ev_rp[s + 1]) {       // nothing to do for this session
        } else if (STAGE && SharedView::bytes(A, E) <= SLASH_SMEM_PER_WARP) {
            SharedView m;
            m.carve(slash_smem + (size_t)wib * SLASH_SMEM_PER_WARP, A, E);
#pragma unroll 4
            for (uint32_t a = lane; a <= A; a += 32) m.s_rowptr[a] = k.rowptr[a_lo + a] - e_base;
#pragma unroll 4
            for (uint32_t a = lane; a < A; a += 32) {
                m.s_score[a] = k.scores[a_lo + a];
                m.s_touched[a] = 0;
            }
#pragma unroll 8
            for (uint32_t e = lane; e < E; e += 32) {
                const uint32_t u = k.voucher[e_base + e] - a_lo;
                const uint32_t live = live_flag(k.active[e_base + e], k.expiry[e_base + e], k.now);
                m.s_voucher[e] = u;
                m.s_flags[e] = !live ? 0 : (u < A ? (F_LIVE | F_WAS_LIVE) : F_BAD_VOUCHER);
            }
            __syncwarp();
            run_events(k, m, g, lane, s, A);
            __syncwarp();
            for (uint32_t a = lane; a < A; a += 32)
                if (m.s_touched[a]) {
                    k.scores[a_lo + a] = m.s_score[a];
                    k.touched[a_lo + a] = 1;
                }
            for (uint32_t e = lane; e < E; e += 32)
                if (m.s_flags[e] == F_WAS_LIVE) k.active[e_base + e] = 0;
            __syncwarp();                         // the slice is reused by this warp's next session
        } else {
            GlobalView m;
            m.rowptr = k.rowptr + a_lo;
            m.voucher = k.voucher + e_base;
            m.active = k.active + e_base;
            m.expiry = k.expiry + e_base;
            m.scores = k.scores + a_lo;
            m.touched = k.touched + a_lo;
            m.cl = g.cl;
            m.now = k.now;
            m.a_lo = a_lo;
            m.n_agents = A;
            m.e_base = e_base;
            run_events(k, m, g, lane, s, A);
        }
        g.status |= __reduce_or_sync(FULL, g.status);
        if (lane == 0) {
            k.slash_count[s] = g.n_slash;
            k.clip_count[s] = g.n_clip;
            k.status[s] = (uint8_t)g.status;
        }
    }
}

}  // namespace
}  // namespace ahv

using namespace ahv;

namespace {

// Bytes of shared memory per warp for k_slash_lanes.  A warp fills its slice with as many consecutive sessions
// as fit (32 at most) and gives each a lane; the slices of the resident warps share the SM's 227 KB.  Measured
// (tools/slash_lab.py --slice): 7040 B lets 8 CTAs live on an SM (the register limit) and is the best
// size whenever it still gives every lane a session, and again for sessions so large that a warp holds only a
// few whatever the slice; in between the slice grows just enough for 32 average sessions (6 or 5 CTAs per SM).
uint32_t auto_lane_slice(uint64_t n_sessions, uint64_t n_agents, uint64_t n_records)
{
    constexpr uint64_t SMALL = 7040, MID = 9216;
    if (n_records == UINT64_MAX || n_sessions == 0) return LANE_SMEM_PER_WARP;   // records not known to the host
    const uint64_t avg = lane_group_bytes((n_agents + n_sessions - 1) / n_sessions, (n_records + n_sessions - 1) / n_sessions);
    if (32 * avg <= SMALL) return (uint32_t)SMALL;
    if (32 * avg <= MID) return (uint32_t)MID;
    if (32 * avg <= LANE_SMEM_PER_WARP) return LANE_SMEM_PER_WARP;
    if (4 * avg <= SMALL) return (uint32_t)SMALL;
    return LANE_SMEM_PER_WARP;
}

}  // namespace

namespace ahv {

int slash_cascade_launch(ahv_ctx *ctx, const ahv_slash_batch *b, uint64_t n_records)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(b, "batch is null");
    if (b->n_sessions == 0) return AHV_OK;
    AHV_REQUIRE(b->rowptr && b->session_agent_rowptr && b->event_rowptr, "rowptr / session_agent_rowptr / event_rowptr is null");
    AHV_REQUIRE((b->n_agents == 0 || (b->scores && b->touched)) && b->slash_count && b->clip_count && b->status,
                "scores / touched / slash_count / clip_count / status is null");
    AHV_REQUIRE(b->n_agents < 0xffffffffull, "more than 2^32 - 2 agent rows");
    SlashArgs k;
    k.rowptr = b->rowptr; k.voucher = b->voucher; k.active = b->active; k.expiry = b->expiry; k.now = b->now;
    k.sess_rp = b->session_agent_rowptr; k.ev_rp = b->event_rowptr; k.ev_agent = b->event_agent;
    k.ev_sigma = b->event_sigma; k.ev_omega = b->event_omega; k.ev_depth = b->event_depth; k.n_sessions = b->n_sessions;
    k.scores = b->scores; k.touched = b->touched; k.slash_log = b->slash_log; k.clip_log = b->clip_log;
    k.slash_count = b->slash_count; k.clip_count = b->clip_count; k.status = b->status;
    k.big_sessions = nullptr;
    const uint64_t want = (b->n_sessions + SLASH_WARPS - 1) / SLASH_WARPS;
    const uint64_t cap = (uint64_t)ctx->sm_count * 16;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    if (ctx->opt_slash_variant == 1) {            // everything on global memory, warp per session (kept for A/B runs)
        k_slash_cascade<false><<<grid, 32 * SLASH_WARPS, 0, ctx->stream>>>(k, 0);
        AHV_CHECK_LAUNCH(ctx);
        return AHV_OK;
    }
    const int lanes = ctx->opt_slash_variant == 0;                       // 2: warp per session only, staged where it fits
    if (lanes) {
        AHV_REQUIRE(b->n_sessions < 0xffffffffull, "more than 2^32 - 2 sessions");
        AHV_TRY(scratch_reserve(ctx, 4 * (b->n_sessions + 1) + 256));
        k.big_sessions = (uint32_t *)scratch_take(ctx, 4 * (b->n_sessions + 1));
        AHV_REQUIRE(k.big_sessions, "internal: scratch too small for the session list");
        AHV_CUDA(cudaMemsetAsync(k.big_sessions, 0, 4, ctx->stream));
        const uint64_t groups = (b->n_sessions + 31) / 32;
        const uint64_t gwant = (groups + LANE_WARPS - 1) / LANE_WARPS;
        k.lane_slice = ctx->opt_slash_slice ? (uint32_t)ctx->opt_slash_slice : auto_lane_slice(b->n_sessions, b->n_agents, n_records);
        const uint64_t lcap = (uint64_t)ctx->sm_count * 32;
        k_slash_lanes<<<(unsigned)(gwant < lcap ? gwant : lcap), 32 * LANE_WARPS, LANE_WARPS * k.lane_slice, ctx->stream>>>(k);
        AHV_CHECK_LAUNCH(ctx);
    }
    // What the lane kernel leaves are sessions of several KB, nearly always too large for this kernel's 6 KB slices:
    // they walk global memory, where the number of resident warps hides the look-up latency, so no shared memory is
    // reserved (256 agents x 2048 records: 2.80 -> 2.16 ms).  The staged form remains for slash_variant 2.
    if (lanes) k_slash_cascade<false><<<grid, 32 * SLASH_WARPS, 0, ctx->stream>>>(k, 1);
    else k_slash_cascade<true><<<grid, 32 * SLASH_WARPS, SLASH_WARPS * SLASH_SMEM_PER_WARP, ctx->stream>>>(k, 0);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

}  // namespace ahv

extern "C" int ahv_slash_cascade_dev(ahv_ctx *ctx, const ahv_slash_batch *b) { return slash_cascade_launch(ctx, b, UINT64_MAX); }

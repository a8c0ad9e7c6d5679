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
#include "ahv_internal.cuh"

namespace ahv {
namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr double SIGMA_FLOOR = 0.05;              // slashing.py:54
constexpr int MAX_CASCADE_DEPTH = 2;              // slashing.py:53
constexpr int SLASH_WARPS = 4;

struct SlashArgs {
    const uint32_t *rowptr;
    const uint32_t *voucher;
    uint8_t *active;
    const int64_t *expiry;
    int64_t now;
    const uint64_t *sess_rp;
    const uint64_t *ev_rp;
    const uint32_t *ev_agent;
    const double *ev_sigma;
    const double *ev_omega;
    const uint8_t *ev_depth;
    uint64_t n_sessions;
    double *scores;
    uint8_t *touched;
    ahv_slash_record *slash_log;
    ahv_clip_record *clip_log;
    uint32_t *slash_count;
    uint32_t *clip_count;
    uint8_t *status;
};

__device__ __forceinline__ double clip_once(double x, double keep)
{
    const double n = __dmul_rn(x, keep);          // :95
    return (SIGMA_FLOOR > n) ? SIGMA_FLOOR : n;   // :96  Python max(n, FLOOR): a NaN n stays
}

__device__ __forceinline__ bool edge_live(const SlashArgs &k, uint64_t e)
{
    const int64_t x = k.expiry[e];                // vouching.py:34-38
    return k.active[e] != 0 && !(x != AHV_EXPIRY_NONE && k.now > x);
}

// get_vouchers_for(u) non-empty?  (slashing.py:126-129)
__device__ bool row_any_live(const SlashArgs &k, uint32_t u, int lane)
{
    const uint64_t end = k.rowptr[u + 1];
    for (uint64_t base = k.rowptr[u]; base < end; base += 32) {
        const uint64_t e = base + lane;
        const bool l = e < end && edge_live(k, e);
        if (__any_sync(FULL, l)) return true;
    }
    return false;
}

struct SessionLog {
    ahv_slash_record *sl;
    ahv_clip_record *cl;
    uint64_t e_base;          // first vouch record of the session (clip records hold global record indices)
    uint32_t n_slash, n_clip;
    uint32_t a_lo, a_hi;
    uint32_t status;
};

// One SlashingEngine.slash call without its cascade; returns the index of its record.
__device__ uint32_t slash_once(const SlashArgs &k, SessionLog &g, int lane, uint32_t v, double sigma_before, double keep,
                               uint32_t depth, int32_t parent, uint32_t ev)
{
    if (lane == 0) {
        k.scores[v] = 0.0;                        // :88
        k.touched[v] = 1;
    }
    __syncwarp();
    const uint32_t me = g.n_slash++;
    const uint32_t c0 = g.n_clip;
    const uint64_t end = k.rowptr[v + 1];
    for (uint64_t base = k.rowptr[v]; base < end; base += 32) {
        const uint64_t e = base + lane;
        bool live = e < end && edge_live(k, e);
        uint32_t u = 0;
        if (live) {
            u = k.voucher[e];
            if (u < g.a_lo || u >= g.a_hi) {      // a voucher outside the session: refuse the record
                live = false;
                g.status |= 2;
            }
        }
        const unsigned mask = __ballot_sync(FULL, live);
        // lanes naming the same voucher form a group; idle lanes get a key no live lane can have
        const unsigned peers = __match_any_sync(FULL, live ? (unsigned long long)u : ((1ull << 32) | (unsigned)lane));
        double before = 0.0, after = 0.0;
        if (live) {
            const int rank = __popc(peers & ((1u << lane) - 1u));
            before = k.scores[u];                 // :94 (a missing key is a dense 0.0)
            for (int r = 0; r < rank; ++r) before = clip_once(before, keep);
            after = clip_once(before, keep);
        }
        __syncwarp();                             // every lane has read its score before any lane writes
        if (live) {
            if ((peers >> lane) == 1u) {          // last lane of its group
                k.scores[u] = after;              // :97
                k.touched[u] = 1;
            }
            k.active[e] = 0;                      // :108 release_bond
            ahv_clip_record *q = g.cl + (g.n_clip + __popc(mask & ((1u << lane) - 1u)));
            q->edge = (uint32_t)e;
            q->slash = me;
            q->sigma_before = before;
            q->sigma_after = after;
        }
        g.n_clip += __popc(mask);
        __syncwarp();
    }
    g.status |= __reduce_or_sync(FULL, g.status);
    if (lane == 0) {
        ahv_slash_record *r = g.sl + me;          // :110-119
        r->agent = v;
        r->parent = parent;
        r->event = ev;
        r->depth = depth;
        r->clip_begin = c0;
        r->clip_count = g.n_clip - c0;
        r->sigma_before = sigma_before;
    }
    return me;
}

__global__ void __launch_bounds__(32 * SLASH_WARPS)
k_slash_cascade(const SlashArgs k)
{
    const int lane = threadIdx.x & 31;
    const uint64_t warp = (uint64_t)blockIdx.x * SLASH_WARPS + (threadIdx.x >> 5);
    const uint64_t n_warps = (uint64_t)gridDim.x * SLASH_WARPS;
    const double thresh = __dadd_rn(SIGMA_FLOOR, 0.01);                  // :124
    for (uint64_t s = warp; s < k.n_sessions; s += n_warps) {
        SessionLog g;
        g.a_lo = (uint32_t)k.sess_rp[s];
        g.a_hi = (uint32_t)k.sess_rp[s + 1];
        g.e_base = k.rowptr[g.a_lo];
        g.sl = k.slash_log + (k.ev_rp[s] + g.e_base);
        g.cl = k.clip_log + g.e_base;
        g.n_slash = g.n_clip = 0;
        g.status = 0;
        for (uint64_t ev = k.ev_rp[s]; ev < k.ev_rp[s + 1]; ++ev) {
            const uint32_t v0 = k.ev_agent[ev];
            if (v0 < g.a_lo || v0 >= g.a_hi) {    // an event naming an agent of another session
                g.status |= 1;
                continue;
            }
            const double keep = __dsub_rn(1.0, k.ev_omega[ev]);          // (1.0 - risk_weight), :95
            uint32_t f_cur[MAX_CASCADE_DEPTH], f_end[MAX_CASCADE_DEPTH], f_id[MAX_CASCADE_DEPTH];
            const uint32_t depth0 = k.ev_depth ? k.ev_depth[ev] : 0u;     // the caller's cascade_depth argument (:70)
            f_cur[0] = g.n_clip;
            f_id[0] = slash_once(k, g, lane, v0, k.ev_sigma[ev], keep, depth0, -1, (uint32_t)ev);
            f_end[0] = g.n_clip;
            int d = depth0 < (uint32_t)MAX_CASCADE_DEPTH ? 0 : -1;
            while (d >= 0) {                      // :122-138, depth first, clips in order
                if (f_cur[d] == f_end[d]) {
                    --d;
                    continue;
                }
                const ahv_clip_record *q = g.cl + f_cur[d]++;
                const double after = q->sigma_after;
                if (!(after < thresh)) continue;
                const uint32_t u = k.voucher[q->edge];
                if (!row_any_live(k, u, lane)) continue;
                const uint32_t c0 = g.n_clip;
                const uint32_t id = slash_once(k, g, lane, u, after, keep, depth0 + d + 1, (int32_t)f_id[d], (uint32_t)ev);
                if (depth0 + d + 1 < (uint32_t)MAX_CASCADE_DEPTH) {
                    ++d;
                    f_cur[d] = c0;
                    f_end[d] = g.n_clip;
                    f_id[d] = id;
                }
            }
        }
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

extern "C" int ahv_slash_cascade_dev(ahv_ctx *ctx, const ahv_slash_batch *b)
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
    const uint64_t want = (b->n_sessions + SLASH_WARPS - 1) / SLASH_WARPS;
    const uint64_t cap = (uint64_t)ctx->sm_count * 16;
    k_slash_cascade<<<(unsigned)(want < cap ? want : cap), 32 * SLASH_WARPS, 0, ctx->stream>>>(k);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

// slashing_core.cuh - the per-session arithmetic and the serial (lane-per-session) walker of the
// slashing pass, host+device so the CPU-only test harness can run the same code (tests/hostemu); the
// product has no CPU path.  Reference: liability/slashing.py:53-141.
#pragma once
#include <stdint.h>

#include "../../include/ahv.h"

#if defined(__CUDACC__)
#define AHV_SL_HD __host__ __device__
#define AHV_SL_HDI __host__ __device__ __forceinline__
#else
#define AHV_SL_HD
#define AHV_SL_HDI inline
#endif

namespace ahv {
namespace slashing {

constexpr unsigned FULL = 0xffffffffu;
constexpr double SIGMA_FLOOR = 0.05;              // slashing.py:54
constexpr int MAX_CASCADE_DEPTH = 2;              // slashing.py:53
constexpr int SLASH_WARPS = 4;
constexpr uint32_t SLASH_SMEM_PER_WARP = 6144;    // bytes of staged session state per warp
constexpr int LANE_WARPS = 4;                     // k_slash_lanes: warps per CTA
constexpr uint32_t LANE_SMEM_PER_WARP = 10752;    // bytes of staged state for a run of sessions (5 CTAs per SM)
static_assert(LANE_WARPS * LANE_SMEM_PER_WARP <= 48 * 1024, "k_slash_lanes must fit the default dynamic shared-memory limit");
constexpr int LANE_CLIP_CACHE = 12;
constexpr int LANE_MIN_SESSIONS = 2;              // fewer sessions than this in a run: a warp per session instead

AHV_SL_HDI uint64_t lane_group_bytes(uint64_t agents, uint64_t records)
{
    return 11 * agents + 2 * records + 16;      // < 8191 agents and < 65536 records: 16-bit words inside a group
}

constexpr uint8_t F_LIVE = 1, F_WAS_LIVE = 2, F_BAD_VOUCHER = 4;
// The lane walker keeps ONE 16-bit word per staged record: the voucher's run-local agent index in the low 10
// bits (a run holds fewer than 1000 agents; R_OUTSIDE = not an agent of the run) and three flags above it.
constexpr uint16_t R_VOUCHER = 0x03ff, R_OUTSIDE = 0x03ff;
constexpr uint16_t R_LIVE = 0x0400, R_WAS_LIVE = 0x0800;
constexpr uint16_t R_CASCADE = 0x1000;    // the clip of this record left its voucher under the cascade threshold
constexpr uint16_t R_ROW_DEAD = 0x2000;   // on the FIRST record of a row: no record of the row can be live any more
static_assert(LANE_SMEM_PER_WARP / 11 < R_OUTSIDE, "run-local agent indices must fit 10 bits");

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
    uint32_t *big_sessions;       // scratch: [0] = count, [1..] = sessions k_slash_lanes left to k_slash_cascade
    uint32_t lane_slice = LANE_SMEM_PER_WARP;   // k_slash_lanes: bytes of staged state per warp (<= LANE_SMEM_PER_WARP)
};

AHV_SL_HDI double clip_once(double x, double keep)
{
#if defined(__CUDA_ARCH__)
    const double n = __dmul_rn(x, keep);          // :95
#else
    const volatile double n0 = x * keep;
    const double n = n0;
#endif
    return (SIGMA_FLOOR > n) ? SIGMA_FLOOR : n;   // :96  Python max(n, FLOOR): a NaN n stays
}

AHV_SL_HDI double add_rn(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    const volatile double s = a + b;
    return s;
#endif
}

// vouching.py:34-38,191-196 on values already loaded (bitwise on purpose: both loads are always issued,
// neither waits for the other)
AHV_SL_HDI uint32_t live_flag(uint8_t active, int64_t expiry, int64_t now)
{
    const uint32_t expired = (uint32_t)(expiry != AHV_EXPIRY_NONE) & (uint32_t)(now > expiry);
    return (uint32_t)(active != 0) & (expired ^ 1u);
}

// One lane = one session, serial, on the group's staged state.  Indices: agents and records are
// group-local in shared memory; `la` is the first agent of this lane's session, `ln` their number.
//
// The cascade (slashing.py:122-138) walks the clips of a slash in order and slashes the vouchers whose
// recorded sigma_after fell under the threshold.  Here a slash leaves that decision on the record it clipped
// (R_CASCADE), so the walk re-reads the ROW it has just released instead of the clip log: the released
// records of the row, in order, ARE the slash's clips (slash(v) releases every live record of row v, so
// only the first slash of v that finds live records produces clips, and nothing else touches the row).
// MAX_CASCADE_DEPTH is 2, so the reference's explicit stack becomes two nested loops with scalar cursors.
struct LaneRun {
    double *s_score;
    uint16_t *s_rowptr;           // group-local record index
    uint16_t *s_rec;              // voucher (group-local agent index) | R_* flags
    uint8_t *s_touched;
    ahv_slash_record *sl;
    ahv_clip_record *cl;
    ahv_clip_record *cl_next;     // cl + n_clip
    uint32_t la, ln;              // this session's agents inside the group
    uint32_t ga_lo, ge_lo;        // the group's first agent row / first record (global indices)
    uint32_t n_slash, n_clip, status;
    // sigma_after of the first LANE_CLIP_CACHE clips of the slash at cascade level 0 / 1 (the sigma_before of
    // a cascaded slash), kept in (L1-resident) local memory instead of reading the clip records back
    double c_after[MAX_CASCADE_DEPTH][LANE_CLIP_CACHE];

    // slashing.py:126-129.  A row found (or left) without a live record stays that way: remembered on the
    // row's first record.  At a high risk weight most cascade candidates are agents slashed a moment ago;
    // without the mark each of them costs a scan of the whole row.
    AHV_SL_HDI bool any_live(uint32_t v)
    {
        const uint32_t e0 = s_rowptr[la + v], e1 = s_rowptr[la + v + 1];
        if (e0 == e1 || (s_rec[e0] & R_ROW_DEAD)) return false;
        for (uint32_t e = e0; e < e1; ++e) {
            const uint32_t x = s_rec[e];
            if ((x & R_LIVE) && (x & R_VOUCHER) - la < ln) return true;
        }
        s_rec[e0] |= R_ROW_DEAD;
        return false;
    }
    // frame: the cascade level (0 / 1) whose clips this slash produces, or -1 when it will not cascade
    AHV_SL_HDI uint32_t slash(uint32_t v, double sigma_before, double keep, double thresh, uint32_t depth, int32_t parent,
                              uint32_t ev, int frame)
    {
        s_score[la + v] = 0.0;                    // :88
        s_touched[la + v] = 1;
        const uint32_t me = n_slash++, c0 = n_clip;
        const uint32_t e0 = s_rowptr[la + v], e1 = s_rowptr[la + v + 1];
        for (uint32_t e = e0; e < e1; ++e) {                                     // :92
            const uint32_t x = s_rec[e];
            const uint32_t u = (x & R_VOUCHER) - la;
            if (!(x & R_LIVE)) continue;
            if (u >= ln) {
                status |= 2;                                                    // voucher outside the session
                continue;
            }
            const double before = s_score[la + u];                              // :94
            const double after = clip_once(before, keep);                       // :95-96
            s_score[la + u] = after;                                            // :97 (write_back_run marks u touched)
            // :108 release_bond, and :124 decided now: will this clip ask for a cascade?
            s_rec[e] = (uint16_t)((x & ~(uint32_t)R_LIVE) | (after < thresh ? (uint32_t)R_CASCADE : 0u));
            if (frame >= 0 && n_clip - c0 < (uint32_t)LANE_CLIP_CACHE) c_after[frame][n_clip - c0] = after;
            ahv_clip_record *q = cl_next++;                                     // :99-105
            ++n_clip;
            q->edge = ge_lo + e;
            q->slash = me;
            q->sigma_before = before;
            q->sigma_after = after;
        }
        if (e0 < e1) s_rec[e0] |= R_ROW_DEAD;
        ahv_slash_record *r = sl + me;            // :110-119
        r->agent = ga_lo + la + v;
        r->parent = parent;
        r->event = ev;
        r->depth = depth;
        r->clip_begin = c0;
        r->clip_count = n_clip - c0;
        r->sigma_before = sigma_before;
        return me;
    }
    // the o-th clip's sigma_after of the slash whose clips start at c0 (cascade level `frame`)
    AHV_SL_HDI double clip_after(int frame, uint32_t c0, uint32_t o) const
    {
        return o < (uint32_t)LANE_CLIP_CACHE ? c_after[frame][o] : cl[c0 + o].sigma_after;   // a long row: own earlier stores
    }
    static AHV_SL_HDI bool released(uint32_t x) { return (x & (uint32_t)(R_LIVE | R_WAS_LIVE)) == (uint32_t)R_WAS_LIVE; }

    AHV_SL_HDI void events(const SlashArgs &k, uint64_t ev_lo, uint64_t ev_hi)
    {
        const double thresh = add_rn(SIGMA_FLOOR, 0.01);                        // :124
        static_assert(MAX_CASCADE_DEPTH == 2, "the cascade below is written out for two levels");
        for (uint64_t ev = ev_lo; ev < ev_hi; ++ev) {
            const uint32_t v0 = k.ev_agent[ev] - (ga_lo + la);
            if (v0 >= ln) {
                status |= 1;
                continue;
            }
            const double keep = add_rn(1.0, -k.ev_omega[ev]);                   // (1.0 - risk_weight), :95
            const uint32_t depth0 = k.ev_depth ? k.ev_depth[ev] : 0u;
            const bool level0 = depth0 < (uint32_t)MAX_CASCADE_DEPTH;
            const uint32_t c0_0 = n_clip;
            const uint32_t id0 = slash(v0, k.ev_sigma[ev], keep, thresh, depth0, -1, (uint32_t)ev, level0 ? 0 : -1);
            if (!level0 || n_clip == c0_0) continue;
            const bool level1 = depth0 + 1 < (uint32_t)MAX_CASCADE_DEPTH;
            uint32_t o0 = 0;
            for (uint32_t e = s_rowptr[la + v0], e_end = s_rowptr[la + v0 + 1]; e < e_end; ++e) {   // :122-138, level 0
                const uint32_t x = s_rec[e];
                if (!released(x)) continue;
                const uint32_t o = o0++;
                if (!(x & R_CASCADE)) continue;
                const uint32_t u = (x & R_VOUCHER) - la;
                if (!any_live(u)) continue;
                const uint32_t c0_1 = n_clip;
                const uint32_t id1 = slash(u, clip_after(0, c0_0, o), keep, thresh, depth0 + 1, (int32_t)id0, (uint32_t)ev, level1 ? 1 : -1);
                if (!level1 || n_clip == c0_1) continue;
                uint32_t o1 = 0;
                for (uint32_t f = s_rowptr[la + u], f_end = s_rowptr[la + u + 1]; f < f_end; ++f) {   // level 1
                    const uint32_t y = s_rec[f];
                    if (!released(y)) continue;
                    const uint32_t p = o1++;
                    if (!(y & R_CASCADE)) continue;
                    const uint32_t w = (y & R_VOUCHER) - la;
                    if (!any_live(w)) continue;
                    slash(w, clip_after(1, c0_1, p), keep, thresh, depth0 + 2, (int32_t)id1, (uint32_t)ev, -1);
                }
            }
        }
    }
};

// Carve a slice for a run of A agents / E records (all sessions p .. q-1 of a group) and fill it from the
// batch: row pointers and voucher indices become 16-bit run-local ones, live = active and not expired.
// `lane` / `stride`: 0..31 / 32 for a warp, 0 / 1 for one host thread.
AHV_SL_HDI void stage_run(LaneRun &r, uint8_t *base, const SlashArgs &k, uint32_t a_lo, uint32_t A, uint32_t e_lo, uint32_t E,
                          uint32_t lane, uint32_t stride)
{
    r.s_score = reinterpret_cast<double *>(base);
    r.s_rowptr = reinterpret_cast<uint16_t *>(r.s_score + A);
    r.s_rec = r.s_rowptr + (A + 1);
    r.s_touched = reinterpret_cast<uint8_t *>(r.s_rec + E);
#pragma unroll 8
    for (uint32_t a = lane; a <= A; a += stride) r.s_rowptr[a] = (uint16_t)(k.rowptr[a_lo + a] - e_lo);
#pragma unroll 8
    for (uint32_t a = lane; a < A; a += stride) {
        r.s_score[a] = k.scores[a_lo + a];
        r.s_touched[a] = 0;
    }
#pragma unroll 8
    for (uint32_t e = lane; e < E; e += stride) {
        // a voucher outside the run becomes R_OUTSIDE: outside every session of the run as well
        const uint32_t u = k.voucher[e_lo + e] - a_lo;
        r.s_rec[e] = (uint16_t)((u < A ? u : (uint32_t)R_OUTSIDE) |
                                (live_flag(k.active[e_lo + e], k.expiry[e_lo + e], k.now) ? (R_LIVE | R_WAS_LIVE) : 0));
    }
}

// What changed goes back: released bonds, touched scores.  The walker marks only the agents it slashed; the
// vouchers it clipped are the vouchers of the released records, marked here (by all lanes, in parallel).
// The caller synchronises the lanes before this (the walkers are done) - and once inside, between the two loops.
AHV_SL_HDI void write_back_run(const LaneRun &r, const SlashArgs &k, uint32_t a_lo, uint32_t A, uint32_t e_lo, uint32_t E,
                               uint32_t lane, uint32_t stride)
{
    for (uint32_t e = lane; e < E; e += stride) {
        const uint32_t x = r.s_rec[e];
        if ((x & (R_LIVE | R_WAS_LIVE)) == R_WAS_LIVE) {
            k.active[e_lo + e] = 0;
            r.s_touched[x & R_VOUCHER] = 1;
        }
    }
#if defined(__CUDA_ARCH__)
    __syncwarp();
#endif
    for (uint32_t a = lane; a < A; a += stride)
        if (r.s_touched[a]) {
            k.scores[a_lo + a] = r.s_score[a];
            k.touched[a_lo + a] = 1;
        }
}

// One session of a staged run (agents my_a0 .. my_a1, first record my_e0, events ev_lo .. ev_hi).
AHV_SL_HDI void run_session(LaneRun &r, const SlashArgs &k, uint64_t s, uint32_t a_lo, uint32_t e_lo, uint32_t my_a0,
                            uint32_t my_a1, uint32_t my_e0, uint64_t ev_lo, uint64_t ev_hi)
{
    r.la = my_a0 - a_lo;
    r.ln = my_a1 - my_a0;
    r.ga_lo = a_lo;
    r.ge_lo = e_lo;
    r.sl = k.slash_log + (ev_lo + my_e0);
    r.cl = r.cl_next = k.clip_log + my_e0;
    r.n_slash = r.n_clip = r.status = 0;
    r.events(k, ev_lo, ev_hi);
    k.slash_count[s] = r.n_slash;
    k.clip_count[s] = r.n_clip;
    k.status[s] = (uint8_t)r.status;
}

}  // namespace slashing
}  // namespace ahv

// Host emulation of the per-thread device functions (TEST INFRASTRUCTURE ONLY).
// Compiles the same sha256_core.cuh the kernels use, as plain C++, so logic
// errors are caught in the CPU-only container before a GPU run.  Never loaded by
// the product package.
#include <cstring>
#include "../../agent_hypervisor_b200/csrc/sha256_lab.cuh"

extern "C" {
void emu_sha256(const uint8_t *p, uint64_t len, uint8_t *out32)
{
    uint32_t st[8];
    ahv::sha256_message(p, len, st);
    alignas(16) uint8_t tmp[32];
    ahv::store_digest(tmp, st);
    memcpy(out32, tmp, 32);
}
void emu_node(const uint8_t *l32, const uint8_t *r32, uint8_t *out32)
{
    alignas(16) uint8_t a[32], b[32], o[32];
    memcpy(a, l32, 32);
    memcpy(b, r32, 32);
    uint32_t l[8], r[8], d[8];
    ahv::load_digest(a, l);
    ahv::load_digest(b, r);
    ahv::merkle_node(l, r, d);
    ahv::store_digest(o, d);
    memcpy(out32, o, 32);
}
int emu_digest_equal(const uint8_t *mem32, const uint8_t *other32)
{
    alignas(16) uint8_t a[32], b[32];
    memcpy(a, mem32, 32);
    memcpy(b, other32, 32);
    uint32_t st[8];
    ahv::load_digest(b, st);
    return ahv::digest_equal_mem(a, st) ? 1 : 0;
}
}

#include "../../agent_hypervisor_b200/csrc/liability_core.cuh"
extern "C" {
double emu_sigma_row(const double *bonded, const uint8_t *active, const int64_t *expiry, int n, int64_t now,
                     double sigma, double omega)
{
    double c = 0.0;
    for (int e = 0; e < n; ++e) c = ahv::vouch_accumulate(c, bonded[e], active[e], expiry[e], now);
    return ahv::sigma_eff_of(sigma, omega, c);
}
int emu_ring(double se, int consensus) { return ahv::ring_from_sigma(se, consensus != 0); }
int emu_gate(int agent_ring, int flags, double s, int consensus, int witness, int *req_out)
{
    int req = ahv::required_ring((uint8_t)flags);
    *req_out = req;
    return ahv::gate_reason(agent_ring, req, s, consensus != 0, witness != 0);
}
int emu_near(double se) { return ahv::near_threshold(se); }
}

extern "C" void emu_compress_variants(const uint32_t *block16, uint32_t *out40)
{
    const ahv::ShaMults mu = ahv::make_sha_mults();
    for (int v = 0; v < 5; ++v) {
        uint32_t st[8], w[16];
        ahv::sha256_init(st);
        for (int i = 0; i < 16; ++i) w[i] = block16[i];
        if (v == 0) ahv::sha256_compress(st, w);
        else if (v == 1) ahv::sha256_compress_imad(st, w, mu);
        else if (v == 2) ahv::sha256_compress_sel<0x1FFFu>(st, w, mu);
        else if (v == 3) ahv::sha256_compress_scaled(st, w, mu.c);
        else { const ahv::KW64 kct = ahv::make_kc_table(); ahv::sha256_compress_scaled_t<5>(st, w, mu.c, kct); }
        for (int i = 0; i < 8; ++i) out40[8 * v + i] = st[i];
    }
}

extern "C" {
void emu_sha256_scaled(const uint8_t *p, uint64_t len, uint8_t *out32)
{
    ahv::CompressScaled cp;
    cp.rc = ahv::SHA_SC;
    uint32_t st[8];
    ahv::sha256_message(p, len, st, cp);
    alignas(16) uint8_t tmp[32];
    ahv::store_digest(tmp, st);
    memcpy(out32, tmp, 32);
}
void emu_node_scaled(const uint8_t *l32, const uint8_t *r32, uint8_t *out32)
{
    ahv::CompressScaled cp;
    cp.rc = ahv::SHA_SC;
    alignas(16) uint8_t a[32], b[32], o[32];
    memcpy(a, l32, 32);
    memcpy(b, r32, 32);
    uint32_t l[8], r[8], d[8];
    ahv::load_digest(a, l);
    ahv::load_digest(b, r);
    ahv::merkle_node(l, r, d, cp);
    ahv::store_digest(o, d);
    memcpy(out32, o, 32);
}
// fixed 256-byte message through the pad-table path (what k_sha256_fixed does)
void emu_sha256_fixed256(const uint8_t *p, int scaled, uint8_t *out32)
{
    uint32_t st[8];
    ahv::sha256_init(st);
    ahv::CompressScaled cs;
    cs.rc = ahv::SHA_SC;
    ahv::CompressNative cn;
    for (int b = 0; b < 4; ++b) {
        uint32_t w[16];
        for (int j = 0; j < 16; ++j) {
            const uint8_t *q = p + 64 * b + 4 * j;
            w[j] = ((uint32_t)q[0] << 24) | ((uint32_t)q[1] << 16) | ((uint32_t)q[2] << 8) | q[3];
        }
        if (scaled) cs.block(st, w); else cn.block(st, w);
    }
    if (scaled) {
        const ahv::KW64 kw = ahv::make_kw_pad_scaled(2048);
        cs.pad(st, kw);
    } else {
        const ahv::KW64 kw = ahv::make_kw_pad(2048);
        cn.pad(st, kw);
    }
    alignas(16) uint8_t tmp[32];
    ahv::store_digest(tmp, st);
    memcpy(out32, tmp, 32);
}
}

extern "C" void emu_compress_search(const uint32_t *block16, uint32_t *out16)
{
    for (int v = 0; v < 2; ++v) {
        uint32_t st[8], w[16];
        ahv::sha256_init(st);
        for (int i = 0; i < 16; ++i) w[i] = block16[i];
        if (v == 0) ahv::sha256_compress(st, w);
        else ahv::sha256_compress_scaled_v<3 + 16 + 32>(st, w, ahv::SHA_SC);
        for (int i = 0; i < 8; ++i) out16[8 * v + i] = st[i];
    }
}

// ---- slashing: the lane-per-session form (k_slash_lanes) run by one host thread -------------------
// Same staging (16-bit run-local indices), same LaneRun walker, same run-cutting policy as the kernel;
// sessions the kernel would hand to k_slash_cascade are reported in out_big (count returned).
#include <vector>
#include "../../agent_hypervisor_b200/csrc/slashing_core.cuh"
extern "C" uint64_t emu_slash_lanes(const ahv_slash_batch *b, uint32_t *out_big)
{
    using namespace ahv::slashing;
    SlashArgs k;
    k.rowptr = b->rowptr; k.voucher = b->voucher; k.active = b->active; k.expiry = b->expiry; k.now = b->now;
    k.sess_rp = b->session_agent_rowptr; k.ev_rp = b->event_rowptr; k.ev_agent = b->event_agent;
    k.ev_sigma = b->event_sigma; k.ev_omega = b->event_omega; k.ev_depth = b->event_depth; k.n_sessions = b->n_sessions;
    k.scores = b->scores; k.touched = b->touched; k.slash_log = b->slash_log; k.clip_log = b->clip_log;
    k.slash_count = b->slash_count; k.clip_count = b->clip_count; k.status = b->status; k.big_sessions = nullptr;
    std::vector<double> slice(LANE_SMEM_PER_WARP / 8 + 2);
    uint64_t n_big = 0;
    for (uint64_t s0 = 0; s0 < k.n_sessions; s0 += 32) {
        const uint32_t n = (uint32_t)(k.n_sessions - s0 < 32 ? k.n_sessions - s0 : 32);
        uint32_t p = 0;
        while (p < n) {
            const uint32_t a_lo = (uint32_t)k.sess_rp[s0 + p], e_lo = k.rowptr[a_lo];
            const uint32_t a_p1 = (uint32_t)k.sess_rp[s0 + p + 1];
            if (lane_group_bytes(a_p1 - a_lo, k.rowptr[a_p1] - e_lo) > LANE_SMEM_PER_WARP / LANE_MIN_SESSIONS) {
                out_big[n_big++] = (uint32_t)(s0 + p);
                ++p;
                continue;
            }
            uint32_t cnt = 0;
            while (p + cnt < n) {
                const uint32_t a1 = (uint32_t)k.sess_rp[s0 + p + cnt + 1];
                if (lane_group_bytes(a1 - a_lo, k.rowptr[a1] - e_lo) > LANE_SMEM_PER_WARP) break;
                ++cnt;
            }
            const uint32_t q = p + cnt;
            const uint32_t a_hi = (uint32_t)k.sess_rp[s0 + q];
            const uint32_t A = a_hi - a_lo, E = k.rowptr[a_hi] - e_lo;
            LaneRun r;
            stage_run(r, reinterpret_cast<uint8_t *>(slice.data()), k, a_lo, A, e_lo, E, 0u, 1u);
            for (uint32_t i = p; i < q; ++i) {
                const uint64_t s = s0 + i;
                const uint32_t my_a0 = (uint32_t)k.sess_rp[s], my_a1 = (uint32_t)k.sess_rp[s + 1];
                run_session(r, k, s, a_lo, e_lo, my_a0, my_a1, k.rowptr[my_a0], k.ev_rp[s], k.ev_rp[s + 1]);
            }
            write_back_run(r, k, a_lo, A, e_lo, E, 0u, 1u);
            p = q;
        }
    }
    return n_big;
}

// The rolled compression bodies the tree kernels use (sha256_compress_scaled_rolled for the two hex blocks,
// sha256_compress_kw_scaled_rolled for the padding block): one Merkle node from two raw digests.
extern "C" void emu_node_rolled(const uint8_t *l32, const uint8_t *r32, uint8_t *out32, int variant)
{
    alignas(16) uint8_t a[32], b[32], o[32];
    memcpy(a, l32, 32);
    memcpy(b, r32, 32);
    uint32_t l[8], r[8], st[8], w[16];
    ahv::load_digest(a, l);
    ahv::load_digest(b, r);
    ahv::sha256_init(st);
    for (int half = 0; half < 2; ++half) {
        for (int i = 0; i < 8; ++i) ahv::hex_expand(half ? r[i] : l[i], w[2 * i], w[2 * i + 1]);
        if (variant == 16) ahv::sha256_compress_scaled_rolled<16>(st, w, ahv::SHA_SC);
        else if (variant == 32) ahv::sha256_compress_scaled_rolled<32>(st, w, ahv::SHA_SC);
        else ahv::sha256_compress_scaled_rolled<48>(st, w, ahv::SHA_SC);
    }
    static constexpr ahv::KW64 kwc = ahv::make_kw_pad_scaled(1024);
    ahv::sha256_compress_kw_scaled_rolled(st, kwc.v, ahv::SHA_SC);
    ahv::store_digest(o, st);
    memcpy(out32, o, 32);
}

/*
 * Plain-C CPU restatement of the audit / liability hot path.
 *
 * TEST INFRASTRUCTURE ONLY: linked by nothing under agent_hypervisor_b200/.
 * Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.
 *
 * Reference lines followed (paths under /root/reference/src/hypervisor):
 *   orc_sha256*          audit/delta.py:63   hashlib.sha256(payload).hexdigest()
 *                        (third party: CPython hashlib / OpenSSL; algorithm =
 *                        FIPS 180-4 SHA-256, restated here from the standard)
 *   orc_merkle_roots     audit/delta.py:117-134
 *   orc_verify_chains    audit/delta.py:136-152 (effective rule, see below)
 *   orc_sigma_eff_rings  liability/vouching.py:34-38,128-151,186-197,
 *                        models.py:34-42, rings/enforcer.py:134-137
 *   orc_ring_check       models.py:122-132, rings/enforcer.py:44-128
 *   orc_slash_cascade    liability/slashing.py:53-141 (+ vouching.py:153-155,168-174)
 *
 * Parity is pinned: tests/test_oracle_golden.py checks every function against
 * vectors generated from the imported reference (tests/golden/make_golden.py)
 * and against NIST known answers.
 *
 * Build: make -C oracle   (gcc -O2 -fopenmp -shared)
 */
#include <stdint.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ SHA-256 */
static const uint32_t K256[64] = {
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u,
    0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
    0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
    0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,
    0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};

static inline uint32_t ror32(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

static void compress(uint32_t st[8], const uint8_t blk[64])
{
    uint32_t w[64];
    for (int t = 0; t < 16; ++t)
        w[t] = ((uint32_t)blk[4 * t] << 24) | ((uint32_t)blk[4 * t + 1] << 16) |
               ((uint32_t)blk[4 * t + 2] << 8) | (uint32_t)blk[4 * t + 3];
    for (int t = 16; t < 64; ++t) {
        uint32_t s0 = ror32(w[t - 15], 7) ^ ror32(w[t - 15], 18) ^ (w[t - 15] >> 3);
        uint32_t s1 = ror32(w[t - 2], 17) ^ ror32(w[t - 2], 19) ^ (w[t - 2] >> 10);
        w[t] = w[t - 16] + s0 + w[t - 7] + s1;
    }
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    for (int t = 0; t < 64; ++t) {
        uint32_t S1 = ror32(e, 6) ^ ror32(e, 11) ^ ror32(e, 25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = h + S1 + ch + K256[t] + w[t];
        uint32_t S0 = ror32(a, 2) ^ ror32(a, 13) ^ ror32(a, 22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        h = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

void orc_sha256(const uint8_t *msg, uint64_t len, uint8_t out[32])
{
    uint32_t st[8] = {0x6a09e667u, 0xbb67ae85u, 0x3c6ef372u, 0xa54ff53au,
                      0x510e527fu, 0x9b05688cu, 0x1f83d9abu, 0x5be0cd19u};
    uint64_t full = len / 64;
    for (uint64_t b = 0; b < full; ++b) compress(st, msg + 64 * b);
    uint8_t tail[128];
    uint64_t rem = len - 64 * full;
    memset(tail, 0, sizeof tail);
    if (rem) memcpy(tail, msg + 64 * full, rem);
    tail[rem] = 0x80;
    int nb = (rem + 9 > 64) ? 2 : 1;
    uint64_t bits = len * 8;
    for (int i = 0; i < 8; ++i) tail[64 * nb - 1 - i] = (uint8_t)(bits >> (8 * i));
    for (int b = 0; b < nb; ++b) compress(st, tail + 64 * b);
    for (int i = 0; i < 8; ++i) {
        out[4 * i] = (uint8_t)(st[i] >> 24);
        out[4 * i + 1] = (uint8_t)(st[i] >> 16);
        out[4 * i + 2] = (uint8_t)(st[i] >> 8);
        out[4 * i + 3] = (uint8_t)st[i];
    }
}

/* delta.py:63 over a packed arena */
void orc_sha256_batch(const uint8_t *arena, const uint64_t *offsets, uint64_t n, uint8_t *out)
{
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t i = 0; i < (int64_t)n; ++i)
        orc_sha256(arena + offsets[i], offsets[i + 1] - offsets[i], out + 32 * i);
}

/* number of SHA-256 compressions for a len-byte message */
uint64_t orc_count_blocks(const uint64_t *offsets, uint64_t n)
{
    uint64_t c = 0;
    for (uint64_t i = 0; i < n; ++i) c += (offsets[i + 1] - offsets[i] + 9 + 63) / 64;
    return c;
}

/* ------------------------------------------------------------------ Merkle */
static void hex_of(const uint8_t d[32], uint8_t out[64])
{
    static const char H[] = "0123456789abcdef";
    for (int i = 0; i < 32; ++i) {
        out[2 * i] = (uint8_t)H[d[i] >> 4];
        out[2 * i + 1] = (uint8_t)H[d[i] & 15];
    }
}

/* delta.py:125-134 for one session; lvl is caller scratch of n*32 bytes */
static void merkle_one(const uint8_t *leaves, uint64_t n, uint8_t *lvl, uint8_t root[32])
{
    memcpy(lvl, leaves, 32 * n);
    while (n > 1) {                                         /* :125 */
        uint64_t m = 0;
        for (uint64_t i = 0; i < n; i += 2) {               /* :127 */
            const uint8_t *l = lvl + 32 * i;
            const uint8_t *r = (i + 1 < n) ? lvl + 32 * (i + 1) : l;   /* :129 */
            uint8_t msg[128];
            hex_of(l, msg);
            hex_of(r, msg + 64);
            uint8_t d[32];
            orc_sha256(msg, 128, d);                        /* :130 */
            memcpy(lvl + 32 * m, d, 32);
            ++m;
        }
        n = m;
    }
    memcpy(root, lvl, 32);
}

/* out_empty[s]=1 when the session has no deltas (reference returns None, :119-120) */
int orc_merkle_roots(const uint8_t *digests, const uint64_t *rowptr, uint64_t n_sessions,
                     uint8_t *out_roots, uint8_t *out_empty)
{
    int err = 0;
#pragma omp parallel
    {
        uint8_t *buf = NULL;
        uint64_t cap = 0;
#pragma omp for schedule(dynamic, 16)
        for (int64_t s = 0; s < (int64_t)n_sessions; ++s) {
            uint64_t n = rowptr[s + 1] - rowptr[s];
            if (n == 0) {
                out_empty[s] = 1;
                memset(out_roots + 32 * s, 0, 32);
                continue;
            }
            out_empty[s] = 0;
            if (n > cap) {
                buf = (uint8_t *)realloc(buf, 32 * n);
                cap = n;
                if (!buf) { err = 1; cap = 0; continue; }
            }
            merkle_one(digests + 32 * rowptr[s], n, buf, out_roots + 32 * s);
        }
        free(buf);
    }
    return err;
}

/* ------------------------------------------------------------------ verify
 * delta.py:136-152.  compute_hash() at :140 overwrites delta.delta_hash, so the
 * comparison at :141 never fails; the effective verdict is
 *   kind[first] == NONE  and  for i>first: kind[i]==HEX and parent[i]==sha256(preimage[i-1]).
 * parent_kind: 0 None, 1 64-char lowercase hex (raw bytes in parents), 2 other string.
 * out_digests (optional) receives every recomputed leaf digest.
 */
void orc_verify_chains(const uint8_t *arena, const uint64_t *offsets, const uint8_t *parents,
                       const uint8_t *parent_kind, const uint64_t *rowptr, uint64_t n_sessions,
                       uint8_t *out_verdict, int64_t *out_first_bad, uint8_t *out_digests)
{
#pragma omp parallel for schedule(dynamic, 16)
    for (int64_t s = 0; s < (int64_t)n_sessions; ++s) {
        uint8_t prev[32], cur[32];
        out_verdict[s] = 1;
        out_first_bad[s] = -1;
        for (uint64_t i = rowptr[s]; i < rowptr[s + 1]; ++i) {
            orc_sha256(arena + offsets[i], offsets[i + 1] - offsets[i], cur);
            if (out_digests) memcpy(out_digests + 32 * i, cur, 32);
            int ok;
            if (i == rowptr[s]) ok = parent_kind[i] == 0;
            else ok = parent_kind[i] == 1 && memcmp(parents + 32 * i, prev, 32) == 0;
            if (!ok && out_verdict[s]) {
                out_verdict[s] = 0;
                out_first_bad[s] = (int64_t)(i - rowptr[s]);
                if (!out_digests) break;   /* reference returns here; keep hashing only to fill digests */
            }
            memcpy(prev, cur, 32);
        }
    }
}

/* ------------------------------------------------------------------ sigma_eff + rings */
static inline int ring_of(double s, int consensus)
{   /* models.py:34-42 */
    if (s > 0.95 && consensus) return 1;
    else if (s > 0.60) return 2;
    return 3;
}

void orc_sigma_eff_rings(const uint32_t *rowptr, const double *bonded, const uint8_t *active,
                         const int64_t *expiry, int64_t now, const double *sigma, const double *omega,
                         const uint8_t *consensus, const uint8_t *current_ring, uint64_t n_agents,
                         double *out_sigma, uint8_t *out_ring, uint8_t *out_demote, uint8_t *out_near)
{
#pragma omp parallel for schedule(static)
    for (int64_t a = 0; a < (int64_t)n_agents; ++a) {
        volatile double contrib = 0.0;                       /* vouching.py:146 */
        for (uint32_t e = rowptr[a]; e < rowptr[a + 1]; ++e) {
            int expired = (expiry[e] != INT64_MAX) && (now > expiry[e]);   /* :34-38 */
            if (active[e] && !expired) contrib = contrib + bonded[e];      /* :147-148 */
        }
        volatile double prod = omega[a] * contrib;            /* :150, two roundings (no FMA) */
        double se = sigma[a] + prod;
        if (1.0 < se) se = 1.0;                               /* :151 Python min(): NaN stays */
        out_sigma[a] = se;
        out_ring[a] = (uint8_t)ring_of(se, consensus[a] != 0);
        out_demote[a] = (uint8_t)(ring_of(se, 0) > (int)current_ring[a]);  /* enforcer.py:134-137 */
        out_near[a] = (uint8_t)((fabs(se - 0.60) <= 1e-6 * 0.60) || (fabs(se - 0.95) <= 1e-6 * 0.95));
    }
}

/* Python sum() over the live bonded amounts of each row (vouching.py:157-166, liability/__init__.py:64-66).
 * compensated != 0: CPython >= 3.12 (Neumaier, Python/bltinmodule.c); else plain left-to-right. */
void orc_exposure_sums(const uint32_t *rowptr, const double *bonded, const uint8_t *active, const int64_t *expiry,
                       int64_t now, uint64_t n_rows, int compensated, double *out_sum)
{
    for (uint64_t r = 0; r < n_rows; ++r) {
        volatile double f = 0.0, c = 0.0;
        for (uint32_t e = rowptr[r]; e < rowptr[r + 1]; ++e) {
            int expired = (expiry[e] != INT64_MAX) && (now > expiry[e]);
            if (!active[e] || expired) continue;
            double x = bonded[e];
            if (compensated) {
                volatile double t = f + x;
                if (fabs(f) >= fabs(x)) { volatile double u = f - t; u = u + x; c = c + u; }
                else { volatile double u = x - t; u = u + f; c = c + u; }
                f = t;
            } else {
                f = f + x;
            }
        }
        if (compensated && c != 0.0 && isfinite(c)) f = f + c;
        out_sum[r] = f;
    }
}

/* ---- SlashingEngine.slash (liability/slashing.py:62-141) over the CSR of vouch records by vouchee ----
 * Log layouts are the ones include/ahv.h declares (ahv_slash_record 32 B, ahv_clip_record 24 B). */
typedef struct { uint32_t agent; int32_t parent; uint32_t event; uint32_t depth; uint32_t clip_begin; uint32_t clip_count;
                 double sigma_before; } orc_slash_rec;
typedef struct { uint32_t edge; uint32_t slash; double before; double after; } orc_clip_rec;

typedef struct {
    const uint32_t *rowptr, *voucher; uint8_t *active; const int64_t *expiry; int64_t now;
    double *scores; uint8_t *touched; orc_slash_rec *sl; orc_clip_rec *cl; uint32_t n_slash, n_clip;
} slash_state;

static int slash_live(const slash_state *st, uint32_t e)       /* vouching.py:34-38,191-196 */
{
    return st->active[e] && !(st->expiry[e] != INT64_MAX && st->now > st->expiry[e]);
}

static void slash_one(slash_state *st, uint32_t v, double sigma_before, double omega, uint32_t depth, int32_t parent,
                      uint32_t ev)
{
    st->scores[v] = 0.0;                                        /* :88 */
    st->touched[v] = 1;
    const uint32_t me = st->n_slash++, c0 = st->n_clip;
    for (uint32_t e = st->rowptr[v]; e < st->rowptr[v + 1]; ++e) {   /* :92 */
        if (!slash_live(st, e)) continue;
        const uint32_t u = st->voucher[e];
        const double before = st->scores[u];                    /* :94 */
        volatile double keep = 1.0 - omega;
        volatile double nw = before * keep;                     /* :95 */
        double after = (0.05 > nw) ? 0.05 : nw;                 /* :96 Python max(new, FLOOR): NaN stays */
        st->scores[u] = after;                                  /* :97 */
        st->touched[u] = 1;
        orc_clip_rec *q = &st->cl[st->n_clip++];
        q->edge = e; q->slash = me; q->before = before; q->after = after;
        st->active[e] = 0;                                      /* :108 release_bond */
    }
    const uint32_t c1 = st->n_clip;
    orc_slash_rec *r = &st->sl[me];
    r->agent = v; r->parent = parent; r->event = ev; r->depth = depth; r->clip_begin = c0; r->clip_count = c1 - c0;
    r->sigma_before = sigma_before;
    if (depth < 2) {                                            /* :122 MAX_CASCADE_DEPTH */
        volatile double th = 0.05 + 0.01;                       /* :124 */
        for (uint32_t c = c0; c < c1; ++c) {
            const double after = st->cl[c].after;
            if (!(after < th)) continue;
            const uint32_t u = st->voucher[st->cl[c].edge];
            int any = 0;
            for (uint32_t e = st->rowptr[u]; e < st->rowptr[u + 1] && !any; ++e) any = slash_live(st, e);
            if (any) slash_one(st, u, after, omega, depth + 1, (int32_t)me, ev);   /* :126-138 */
        }
    }
}

void orc_slash_cascade(const uint32_t *rowptr, const uint32_t *voucher, uint8_t *active, const int64_t *expiry, int64_t now,
                       const uint64_t *session_agent_rowptr, const uint64_t *event_rowptr, const uint32_t *event_agent,
                       const double *event_sigma, const double *event_omega, const uint8_t *event_depth, uint64_t n_sessions,
                       double *scores,
                       uint8_t *touched, void *slash_log, void *clip_log, uint32_t *out_n_slash, uint32_t *out_n_clip)
{
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t s = 0; s < (int64_t)n_sessions; ++s) {
        const uint64_t a_lo = session_agent_rowptr[s];
        const uint64_t e_base = rowptr[a_lo];
        slash_state st = {rowptr, voucher, active, expiry, now, scores, touched,
                          (orc_slash_rec *)slash_log + event_rowptr[s] + e_base, (orc_clip_rec *)clip_log + e_base, 0, 0};
        for (uint64_t ev = event_rowptr[s]; ev < event_rowptr[s + 1]; ++ev)
            slash_one(&st, event_agent[ev], event_sigma[ev], event_omega[ev], event_depth ? event_depth[ev] : 0, -1, (uint32_t)ev);
        out_n_slash[s] = st.n_slash;
        out_n_clip[s] = st.n_clip;
    }
}

/* action_flags: bit0 is_admin, bit1 is_read_only, bit2 reversibility==NONE */
void orc_ring_check(const uint8_t *agent_ring, const uint8_t *action_flags, const double *sigma,
                    const uint8_t *consensus, const uint8_t *witness, uint64_t n,
                    uint8_t *out_allowed, uint8_t *out_reason, uint8_t *out_required)
{
    for (uint64_t i = 0; i < n; ++i) {
        int f = action_flags[i], req;
        if (f & 1) req = 0;                                   /* models.py:125-126 */
        else if ((f & 4) && !(f & 2)) req = 1;                /* :127-128 */
        else if (f & 2) req = 3;                              /* :129-130 */
        else req = 2;
        double s = sigma[i];
        int reason;
        if (req == 0 && !witness[i]) reason = 1;              /* enforcer.py:61-70 */
        else if (req == 1 && s < 0.95) reason = 2;            /* :74-84 */
        else if (req == 1 && !consensus[i]) reason = 3;       /* :85-93 */
        else if (req == 2 && s < 0.60) reason = 4;            /* :96-107 */
        else if ((int)agent_ring[i] > req) reason = 5;        /* :110-120 */
        else reason = 0;
        out_reason[i] = (uint8_t)reason;
        out_allowed[i] = (uint8_t)(reason == 0);
        out_required[i] = (uint8_t)req;
    }
}

int orc_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void orc_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

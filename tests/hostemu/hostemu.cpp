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

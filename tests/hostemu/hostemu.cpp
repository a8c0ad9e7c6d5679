// Host emulation of the per-thread device functions (TEST INFRASTRUCTURE ONLY).
// Compiles the same sha256_core.cuh the kernels use, as plain C++, so logic
// errors are caught in the CPU-only container before a GPU run.  Never loaded by
// the product package.
#include <cstring>
#include "../../agent_hypervisor_b200/csrc/sha256_core.cuh"

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

// liability_core.cuh - per-element sigma_eff / ring / gate arithmetic (host+device so the
// CPU-only test harness can run the same code; the product has no CPU path).
#pragma once
#include <stdint.h>

#include "../../include/ahv.h"

#if defined(__CUDACC__)
#define AHV_HD __host__ __device__ __forceinline__
#else
#define AHV_HD inline
#endif

namespace ahv {

// vouching.py:34-38,186-197 : record counts iff active and not expired (now > expiry)
AHV_HD double vouch_accumulate(double contrib, double bonded, uint8_t active, int64_t expiry, int64_t now)
{
    const bool expired = (expiry != AHV_EXPIRY_NONE) && (now > expiry);
    if (active && !expired) {
#if defined(__CUDA_ARCH__)
        contrib = __dadd_rn(contrib, bonded);      // vouching.py:147-148, one rounding per add
#else
        contrib = contrib + bonded;
#endif
    }
    return contrib;
}

// vouching.py:150-151 : sigma + (omega * sum) with two roundings (never fused), then
// Python min(x, 1.0): returns 1.0 only when 1.0 < x, so NaN stays NaN.
AHV_HD double sigma_eff_of(double sigma, double omega, double contrib)
{
#if defined(__CUDA_ARCH__)
    double se = __dadd_rn(sigma, __dmul_rn(omega, contrib));
#else
    volatile double prod = omega * contrib;
    double se = sigma + prod;
#endif
    return (1.0 < se) ? 1.0 : se;
}

// models.py:34-42
AHV_HD int ring_from_sigma(double se, bool consensus)
{
    if (se > 0.95 && consensus) return 1;
    else if (se > 0.60) return 2;
    return 3;
}

AHV_HD int near_threshold(double se)
{
    double d2 = se - 0.60, d1 = se - 0.95;
    if (d2 < 0) d2 = -d2;
    if (d1 < 0) d1 = -d1;
    return (d2 <= 1e-6 * 0.60) || (d1 <= 1e-6 * 0.95);
}

// models.py:122-132
AHV_HD int required_ring(uint8_t flags)
{
    if (flags & AHV_ACT_ADMIN) return 0;
    else if ((flags & AHV_ACT_REV_NONE) && !(flags & AHV_ACT_READ_ONLY)) return 1;
    else if (flags & AHV_ACT_READ_ONLY) return 3;
    return 2;
}

// rings/enforcer.py:58-128, tests in the reference's order
AHV_HD int gate_reason(int agent_ring, int req, double s, bool consensus, bool witness)
{
    if (req == 0 && !witness) return AHV_REASON_NEEDS_WITNESS;
    if (req == 1 && s < 0.95) return AHV_REASON_RING1_SIGMA;
    if (req == 1 && !consensus) return AHV_REASON_NEEDS_CONSENSUS;
    if (req == 2 && s < 0.60) return AHV_REASON_RING2_SIGMA;
    if (agent_ring > req) return AHV_REASON_RING_INSUFFICIENT;
    return AHV_REASON_GRANTED;
}

}  // namespace ahv

// sha256_lab.cuh - experimental SHA-256 compression formulations used only by the integer-pipe lab
// (csrc/ahv_peak.cu, tools/lab*.py) and the host emulation tests.  Nothing here is on the product path;
// the formulation that won is sha256_compress_scaled in sha256_core.cuh.  See profiles/r01/int_pipe_lab.md.
#pragma once
#include "sha256_core.cuh"

namespace ahv {

// ---------------------------------------------------------------- FMA-pipe variants
// The ALU pipe (SHF/LOP3/IADD3/PRMT, 64 lanes/clk/SM) is the binding resource of SHA-256 on
// sm_100; IMAD issues on the FMA pipe and co-issues with it.  These variants move work there.
// Multipliers arrive at run time (kernel parameter) so ptxas cannot turn them back into
// shifts/adds.
struct ShaMults {
    uint32_t one;              // 1          : a*1 + b      = add
    uint32_t m25, m14, m29;    // sigma0     : x*2^25 -> (x<<25 | x>>7), x*2^14 -> rotr18, hi(x*2^29) = x>>3
    uint32_t m15, m13, m22;    // sigma1     : rotr17, rotr19, x>>10
    uint32_t c;                // SHA_SC (scale constant of sha256_compress_scaled) as a run-time value
};

inline ShaMults make_sha_mults()
{
    ShaMults m = {1u, 1u << 25, 1u << 14, 1u << 29, 1u << 15, 1u << 13, 1u << 22, 0x9E3779B1u};
    return m;
}

AHV_HD uint32_t madd(uint32_t a, uint32_t b, uint32_t one)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
#else
    return a * one + b;
#endif
}

// xor of the two halves of x * m (m = 2^(32-n)) is rotr(x, n)
AHV_HD void mulwide(uint32_t x, uint32_t m, uint32_t &lo, uint32_t &hi)
{
#if defined(__CUDA_ARCH__)
    unsigned long long p;
    asm("mul.wide.u32 %0, %1, %2;" : "=l"(p) : "r"(x), "r"(m));
    lo = (uint32_t)p;
    hi = (uint32_t)(p >> 32);
#else
    uint64_t p = (uint64_t)x * m;
    lo = (uint32_t)p;
    hi = (uint32_t)(p >> 32);
#endif
}

AHV_HD uint32_t mulhi(uint32_t x, uint32_t m)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(d) : "r"(x), "r"(m));
    return d;
#else
    return (uint32_t)(((uint64_t)x * m) >> 32);
#endif
}

// x >> N issued as IMAD.HI with the immediate 2^(32-N): same value, FMA pipe instead of ALU pipe
// (ptxas keeps the multiply: it only strength-reduces the low half of a product).
template <int N>
AHV_HD uint32_t shr_fma(uint32_t x)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(d) : "r"(x), "n"(1u << (32 - N)));
    return d;
#else
    return x >> N;
#endif
}
AHV_HD uint32_t sig0_f(uint32_t x) { return rotr32(x, 7) ^ rotr32(x, 18) ^ shr_fma<3>(x); }
AHV_HD uint32_t sig1_f(uint32_t x) { return rotr32(x, 17) ^ rotr32(x, 19) ^ shr_fma<10>(x); }

AHV_HD uint32_t sig0_wide(uint32_t x, const ShaMults &mu)
{
    uint32_t l1, h1, l2, h2;
    mulwide(x, mu.m25, l1, h1);
    mulwide(x, mu.m14, l2, h2);
    return (l1 ^ h1 ^ l2) ^ (h2 ^ mulhi(x, mu.m29));
}

AHV_HD uint32_t sig1_wide(uint32_t x, const ShaMults &mu)
{
    uint32_t l1, h1, l2, h2;
    mulwide(x, mu.m15, l1, h1);
    mulwide(x, mu.m13, l2, h2);
    return (l1 ^ h1 ^ l2) ^ (h2 ^ mulhi(x, mu.m22));
}

// a + b as IMAD.IADD with an immediate multiplier (two register reads, FMA pipe)
AHV_HD uint32_t madd1(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mad.lo.u32 %0, %1, 1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
#else
    return a + b;
#endif
}

// Add-site selector: bit set -> that addition is issued as IMAD.IADD (FMA pipe), clear -> plain
// C addition (the compiler fuses neighbours into IADD3 on the ALU pipe).
//   round:    bit0 w+K   bit1 h+(w+K)   bit2 +Ch   bit3 +Sig1   bit4 d+=t1   bit5 Sig0+Maj   bit6 t1+t2
//   schedule: bit7 w+sig1   bit8 +w[i-7]   bit9 +sig0      bit10: sigmas through IMAD.WIDE/HI
//   bit11: state feed-forward adds      bit12: schedule's plain right shifts as IMAD.HI
template <uint32_t MASK, int BIT>
AHV_HD uint32_t add_site(uint32_t a, uint32_t b)
{
    return (MASK >> BIT) & 1u ? madd1(a, b) : a + b;
}

#define AHV_ROUND_SEL(a, b, c, d, e, f, g, h, k, w)                                        \
    {                                                                                      \
        uint32_t t1_ = add_site<MASK, 1>(h, add_site<MASK, 0>(w, k));                      \
        t1_ = add_site<MASK, 2>(t1_, Ch(e, f, g));                                         \
        t1_ = add_site<MASK, 3>(t1_, Sig1(e));                                             \
        (d) = add_site<MASK, 4>(d, t1_);                                                   \
        (h) = add_site<MASK, 6>(t1_, add_site<MASK, 5>(Sig0(a), Maj(a, b, c)));            \
    }

template <uint32_t MASK>
AHV_HD void sha256_compress_sel(uint32_t st[8], uint32_t w[16], const ShaMults &mu)
{
    const uint32_t K[64] = {AHV_K256_LIST};
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        if (t >= 16) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j;
                const bool wide = (MASK >> 10) & 1u;
                const bool shr_hi = (MASK >> 12) & 1u;     // plain right shifts as IMAD.HI (FMA pipe)
                const uint32_t x2 = w[(i - 2) & 15], x15 = w[(i - 15) & 15];
                const uint32_t s1v = wide ? sig1_wide(x2, mu)
                                          : (shr_hi ? (rotr32(x2, 17) ^ rotr32(x2, 19) ^ mulhi(x2, mu.m22)) : sig1(x2));
                const uint32_t s0v = wide ? sig0_wide(x15, mu)
                                          : (shr_hi ? (rotr32(x15, 7) ^ rotr32(x15, 18) ^ mulhi(x15, mu.m29)) : sig0(x15));
                w[i & 15] = add_site<MASK, 9>(add_site<MASK, 8>(add_site<MASK, 7>(w[i & 15], s1v), w[(i - 7) & 15]), s0v);
            }
        }
        AHV_ROUND_SEL(a, b, c, d, e, f, g, h, K[t + 0], w[(t + 0) & 15])
        AHV_ROUND_SEL(h, a, b, c, d, e, f, g, K[t + 1], w[(t + 1) & 15])
        AHV_ROUND_SEL(g, h, a, b, c, d, e, f, K[t + 2], w[(t + 2) & 15])
        AHV_ROUND_SEL(f, g, h, a, b, c, d, e, K[t + 3], w[(t + 3) & 15])
        AHV_ROUND_SEL(e, f, g, h, a, b, c, d, K[t + 4], w[(t + 4) & 15])
        AHV_ROUND_SEL(d, e, f, g, h, a, b, c, K[t + 5], w[(t + 5) & 15])
        AHV_ROUND_SEL(c, d, e, f, g, h, a, b, K[t + 6], w[(t + 6) & 15])
        AHV_ROUND_SEL(b, c, d, e, f, g, h, a, K[t + 7], w[(t + 7) & 15])
    }
    st[0] = add_site<MASK, 11>(st[0], a); st[1] = add_site<MASK, 11>(st[1], b);
    st[2] = add_site<MASK, 11>(st[2], c); st[3] = add_site<MASK, 11>(st[3], d);
    st[4] = add_site<MASK, 11>(st[4], e); st[5] = add_site<MASK, 11>(st[5], f);
    st[6] = add_site<MASK, 11>(st[6], g); st[7] = add_site<MASK, 11>(st[7], h);
}

#define AHV_ROUND_SC(a, b, c, d, e, f, g, h, k, w)                                 \
    {                                                                              \
        uint32_t s_;                                                               \
        if (VAR & 32) {                                                            \
            s_ = imad_rk((h) + (w) + (k), rc_, 0u);      /* one IADD3 (reg+reg+imm), one IMAD */ \
        } else if (VAR & 4) {                                                      \
            s_ = imad_c(w, imad_rk(h, rc_, kct_.v[ti_]));                          \
        } else if (VAR & 8) {                                                      \
            s_ = imad_c((w) + (k), imad_c(h, rc_));                                \
        } else if (VAR & 2) {                                                      \
            s_ = imad_c((w) + (k), imad_rk(h, rc_, 0u));                           \
        } else {                                                                   \
            s_ = imad_rk(h, rc_, (k) * SHA_SC);                                    \
            s_ = imad_c(w, s_);                                                    \
        }                                                                          \
        s_ = imad_c(Ch(e, f, g), s_);                                              \
        s_ = imad_c(Sig1(e), s_);                                                  \
        (d) = imad_ci(s_, d);                                                      \
        (h) = imad_ci(imad_c(Maj(a, b, c), s_), Sig0(a));                          \
    }

// rc must hold SHA_SC at run time (kernel parameter), so that the K*C immediates can ride along.
template <int VAR, class KCT>
AHV_HD void sha256_compress_scaled_t(uint32_t st[8], uint32_t w[16], uint32_t rc_, const KCT &kct_)
{
    const uint32_t K[64] = {AHV_K256_LIST};
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        if (t >= 16) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j;
                if (VAR & 1) {
                    // w[i] = w[i-16] + (w[i-7] + sig1 + sig0): inner sum scaled, unscale adds w[i-16]
                    uint32_t z = (VAR & 8) ? imad_c(w[(i - 7) & 15], rc_)       // run-time ZERO addend
                                           : imad_rk(w[(i - 7) & 15], rc_, 0u); // run-time C: ptxas must not fold C*CI
                    z = imad_c((VAR & 16) ? sig1_f(w[(i - 2) & 15]) : sig1(w[(i - 2) & 15]), z);
                    z = imad_c((VAR & 16) ? sig0_f(w[(i - 15) & 15]) : sig0(w[(i - 15) & 15]), z);
                    w[i & 15] = imad_ci(z, w[i & 15]);
                } else {
                    w[i & 15] += sig1(w[(i - 2) & 15]) + w[(i - 7) & 15] + sig0(w[(i - 15) & 15]);
                }
            }
        }
        { const int ti_ = t + 0; AHV_ROUND_SC(a, b, c, d, e, f, g, h, K[t + 0], w[(t + 0) & 15]) }
        { const int ti_ = t + 1; AHV_ROUND_SC(h, a, b, c, d, e, f, g, K[t + 1], w[(t + 1) & 15]) }
        { const int ti_ = t + 2; AHV_ROUND_SC(g, h, a, b, c, d, e, f, K[t + 2], w[(t + 2) & 15]) }
        { const int ti_ = t + 3; AHV_ROUND_SC(f, g, h, a, b, c, d, e, K[t + 3], w[(t + 3) & 15]) }
        { const int ti_ = t + 4; AHV_ROUND_SC(e, f, g, h, a, b, c, d, K[t + 4], w[(t + 4) & 15]) }
        { const int ti_ = t + 5; AHV_ROUND_SC(d, e, f, g, h, a, b, c, K[t + 5], w[(t + 5) & 15]) }
        { const int ti_ = t + 6; AHV_ROUND_SC(c, d, e, f, g, h, a, b, K[t + 6], w[(t + 6) & 15]) }
        { const int ti_ = t + 7; AHV_ROUND_SC(b, c, d, e, f, g, h, a, K[t + 7], w[(t + 7) & 15]) }
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

struct NoKCT {
    uint32_t v[1];
};

template <int VAR>
AHV_HD void sha256_compress_scaled_v(uint32_t st[8], uint32_t w[16], uint32_t rc_)
{
    const NoKCT none = {{0u}};
    sha256_compress_scaled_t<VAR>(st, w, rc_, none);
}


// K[t]*C table for the VAR&4 form (K folded into the chain's first IMAD as a constant-bank addend)
constexpr KW64 make_kc_table()
{
    KW64 r = {};
    for (int t = 0; t < 64; ++t) r.v[t] = K256_TABLE[t] * SHA_SC;
    return r;
}

// Search space of equivalent scaled-sum formulations (tools/lab4.py): ptxas' scheduling makes
// algebraically identical orderings differ by several percent, so the order is chosen by measurement.
//   A: which term starts the t1 chain with the run-time factor   0 h, 1 w(+K), 2 Ch, 3 Sig1
//   B: where K is added                                           0 on w, 1 on h
//   C: a' = (t1 + Maj)*.. + Sig0  (0)  or  (t1 + Sig0)*.. + Maj  (1)
//   D: schedule chain starts with   0 w[i-7], 1 sig1, 2 sig0
template <int A, int B, int C, int D>
AHV_HD void sha256_compress_search(uint32_t st[8], uint32_t w[16], uint32_t rc_)
{
    const uint32_t K[64] = {AHV_K256_LIST};
    uint32_t v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = st[i];
#pragma unroll
    for (int t = 0; t < 64; ++t) {
        if (t >= 16) {
            const uint32_t t7 = w[(t - 7) & 15], s1v = sig1(w[(t - 2) & 15]), s0v = sig0(w[(t - 15) & 15]);
            uint32_t z;
            if (D == 0) z = imad_c(s0v, imad_c(s1v, imad_rk(t7, rc_, 0u)));
            else if (D == 1) z = imad_c(s0v, imad_c(t7, imad_rk(s1v, rc_, 0u)));
            else z = imad_c(s1v, imad_c(t7, imad_rk(s0v, rc_, 0u)));
            w[t & 15] = imad_ci(z, w[t & 15]);
        }
        // rotating roles: a=v[(0-t)&7] ... h=v[(7-t)&7]
        uint32_t &a = v[(0 - t) & 7], &b = v[(1 - t) & 7], &c = v[(2 - t) & 7], &d = v[(3 - t) & 7];
        uint32_t &e = v[(4 - t) & 7], &f = v[(5 - t) & 7], &g = v[(6 - t) & 7], &h = v[(7 - t) & 7];
        const uint32_t hh = (B == 1) ? h + K[t] : h;
        const uint32_t ww = (B == 0) ? w[t & 15] + K[t] : w[t & 15];
        const uint32_t chv = Ch(e, f, g), s1e = Sig1(e);
        uint32_t s;
        if (A == 0) s = imad_c(s1e, imad_c(chv, imad_c(ww, imad_rk(hh, rc_, 0u))));
        else if (A == 1) s = imad_c(s1e, imad_c(chv, imad_c(hh, imad_rk(ww, rc_, 0u))));
        else if (A == 2) s = imad_c(s1e, imad_c(ww, imad_c(hh, imad_rk(chv, rc_, 0u))));
        else s = imad_c(chv, imad_c(ww, imad_c(hh, imad_rk(s1e, rc_, 0u))));
        d = imad_ci(s, d);
        if (C == 0) h = imad_ci(imad_c(Maj(a, b, c), s), Sig0(a));
        else h = imad_ci(imad_c(Sig0(a), s), Maj(a, b, c));
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) st[i] += v[i];
}

// 3-register IMAD form (multiplier in a register): measured slower (dispatch stalls), kept for the lab
#define AHV_ROUND_IMAD(a, b, c, d, e, f, g, h, k, w)                       \
    {                                                                      \
        uint32_t t1_ = madd(h, madd(w, k, one_), one_);                    \
        t1_ = madd(t1_, Ch(e, f, g), one_);                                \
        t1_ = madd(t1_, Sig1(e), one_);                                    \
        (d) = madd(d, t1_, one_);                                          \
        (h) = madd(t1_, madd(Sig0(a), Maj(a, b, c), one_), one_);          \
    }

AHV_HD void sha256_compress_imad(uint32_t st[8], uint32_t w[16], const ShaMults &mu)
{
    const uint32_t K[64] = {AHV_K256_LIST};
    const uint32_t one_ = mu.one;
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        if (t >= 16) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j;
                w[i & 15] = madd(madd(madd(w[i & 15], sig1(w[(i - 2) & 15]), one_), w[(i - 7) & 15], one_),
                                 sig0(w[(i - 15) & 15]), one_);
            }
        }
        AHV_ROUND_IMAD(a, b, c, d, e, f, g, h, K[t + 0], w[(t + 0) & 15])
        AHV_ROUND_IMAD(h, a, b, c, d, e, f, g, K[t + 1], w[(t + 1) & 15])
        AHV_ROUND_IMAD(g, h, a, b, c, d, e, f, K[t + 2], w[(t + 2) & 15])
        AHV_ROUND_IMAD(f, g, h, a, b, c, d, e, K[t + 3], w[(t + 3) & 15])
        AHV_ROUND_IMAD(e, f, g, h, a, b, c, d, K[t + 4], w[(t + 4) & 15])
        AHV_ROUND_IMAD(d, e, f, g, h, a, b, c, K[t + 5], w[(t + 5) & 15])
        AHV_ROUND_IMAD(c, d, e, f, g, h, a, b, K[t + 6], w[(t + 6) & 15])
        AHV_ROUND_IMAD(b, c, d, e, f, g, h, a, K[t + 7], w[(t + 7) & 15])
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// (experiment, not used by the kernels: see CompressScaled::node_block) The same compression with its last
// rounds rolled: rounds 0 .. STRAIGHT-1 are straight-line code with
// immediate constants, the rest runs 16 rounds per iteration (after 16 rounds both the register roles and
// the 16-word schedule window are back in place) with K read from a table by index.  Smaller code for the
// tree kernels, whose node would otherwise not fit the 32 KB instruction cache.  Same arithmetic.
template <int STRAIGHT>
AHV_HD void sha256_compress_scaled_rolled(uint32_t st[8], uint32_t w[16], uint32_t rc_)
{
    static_assert(STRAIGHT % 16 == 0 && STRAIGHT >= 16 && STRAIGHT <= 64, "whole 16-round groups");
    static constexpr uint32_t K[64] = {AHV_K256_LIST};
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#define AHV_SCHED8(base)                                                           \
    _Pragma("unroll") for (int j_ = 0; j_ < 8; ++j_) {                             \
        const int i_ = (base) + j_;                                                \
        uint32_t z_ = imad_rk(w[(i_ - 7) & 15], rc_, 0u);                          \
        z_ = imad_c(sig1(w[(i_ - 2) & 15]), z_);                                   \
        z_ = imad_c(sig0(w[(i_ - 15) & 15]), z_);                                  \
        w[i_ & 15] = imad_ci(z_, w[i_ & 15]);                                      \
    }
#define AHV_ROUNDS8(kp, wb)                                                        \
    AHV_ROUND_SCP(a, b, c, d, e, f, g, h, (kp)[0], w[((wb) + 0) & 15])             \
    AHV_ROUND_SCP(h, a, b, c, d, e, f, g, (kp)[1], w[((wb) + 1) & 15])             \
    AHV_ROUND_SCP(g, h, a, b, c, d, e, f, (kp)[2], w[((wb) + 2) & 15])             \
    AHV_ROUND_SCP(f, g, h, a, b, c, d, e, (kp)[3], w[((wb) + 3) & 15])             \
    AHV_ROUND_SCP(e, f, g, h, a, b, c, d, (kp)[4], w[((wb) + 4) & 15])             \
    AHV_ROUND_SCP(d, e, f, g, h, a, b, c, (kp)[5], w[((wb) + 5) & 15])             \
    AHV_ROUND_SCP(c, d, e, f, g, h, a, b, (kp)[6], w[((wb) + 6) & 15])             \
    AHV_ROUND_SCP(b, c, d, e, f, g, h, a, (kp)[7], w[((wb) + 7) & 15])
#pragma unroll
    for (int t = 0; t < STRAIGHT; t += 8) {
        if (t >= 16) { AHV_SCHED8(t) }
        AHV_ROUNDS8(K + t, t)
    }
#pragma unroll 1
    for (int t = STRAIGHT; t < 64; t += 16) {
        const uint32_t *k = K + t;
        AHV_SCHED8(0)
        AHV_ROUNDS8(k, 0)
        AHV_SCHED8(8)
        AHV_ROUNDS8(k + 8, 8)
    }
#undef AHV_SCHED8
#undef AHV_ROUNDS8
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}


}  // namespace ahv

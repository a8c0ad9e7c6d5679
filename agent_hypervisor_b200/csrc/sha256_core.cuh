// sha256_core.cuh - per-thread SHA-256 building blocks shared by every audit kernel.
//
// All functions are __host__ __device__ so that tests/hostemu can run the very same
// per-thread logic on the CPU (test infrastructure only; the product has no CPU path).
//
// Reference behaviour implemented here (paths under /root/reference/src/hypervisor):
//   sha256 of a byte string        audit/delta.py:63  (hashlib.sha256(payload).hexdigest())
//   node = sha256(hex(l)+hex(r))   audit/delta.py:130
//
// Layout notes: a digest "in registers" is the 8 SHA-256 state words (big-endian
// interpretation of the 32 digest bytes).  In memory a digest is the 32 raw bytes.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define AHV_HD __host__ __device__ __forceinline__
#else
#define AHV_HD inline
#endif

namespace ahv {

// ---------------------------------------------------------------- primitives
AHV_HD uint32_t rotr32(uint32_t x, int n)
{
#if defined(__CUDA_ARCH__)
    return __funnelshift_r(x, x, n);   // SHF.R.W
#else
    return (x >> n) | (x << (32 - n));
#endif
}

// byte permute: result byte k = byte (sel>>4k & 7) of the 8-byte pool {hi:lo}={b,a}
AHV_HD uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
#if defined(__CUDA_ARCH__)
    return __byte_perm(a, b, sel);
#else
    uint64_t pool = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int k = 0; k < 4; ++k) {
        uint32_t s = (sel >> (4 * k)) & 7u;
        r |= (uint32_t)((pool >> (8 * s)) & 0xffu) << (8 * k);
    }
    return r;
#endif
}

AHV_HD uint32_t bswap32(uint32_t x) { return prmt(x, 0u, 0x0123u); }

AHV_HD uint32_t Sig0(uint32_t x) { return rotr32(x, 2) ^ rotr32(x, 13) ^ rotr32(x, 22); }
AHV_HD uint32_t Sig1(uint32_t x) { return rotr32(x, 6) ^ rotr32(x, 11) ^ rotr32(x, 25); }
AHV_HD uint32_t sig0(uint32_t x) { return rotr32(x, 7) ^ rotr32(x, 18) ^ (x >> 3); }
AHV_HD uint32_t sig1(uint32_t x) { return rotr32(x, 17) ^ rotr32(x, 19) ^ (x >> 10); }
AHV_HD uint32_t Ch(uint32_t e, uint32_t f, uint32_t g) { return g ^ (e & (f ^ g)); }          // one LOP3
AHV_HD uint32_t Maj(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (c & (a | b)); }   // one LOP3

struct KW64 {
    uint32_t v[64];
};

constexpr uint32_t K256_TABLE[64] = {
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u,
    0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u,
    0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau,
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u,
    0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u,
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u,
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u,
    0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u};

constexpr uint32_t c_rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }

// K[t] + W[t] for a block that holds only padding: 0x80 then zeros then the bit length.
// (block 3 of a 128-byte node message: bitlen = 1024; block 5 of a 256-byte delta: 2048)
constexpr KW64 make_kw_pad(uint64_t bitlen)
{
    uint32_t w[64] = {};
    w[0] = 0x80000000u;
    w[14] = (uint32_t)(bitlen >> 32);
    w[15] = (uint32_t)bitlen;
    for (int t = 16; t < 64; ++t) {
        uint32_t s0 = c_rotr(w[t - 15], 7) ^ c_rotr(w[t - 15], 18) ^ (w[t - 15] >> 3);
        uint32_t s1 = c_rotr(w[t - 2], 17) ^ c_rotr(w[t - 2], 19) ^ (w[t - 2] >> 10);
        w[t] = w[t - 16] + s0 + w[t - 7] + s1;
    }
    KW64 r = {};
    for (int t = 0; t < 64; ++t) r.v[t] = w[t] + K256_TABLE[t];
    return r;
}

AHV_HD void sha256_init(uint32_t st[8])
{
    st[0] = 0x6a09e667u; st[1] = 0xbb67ae85u; st[2] = 0x3c6ef372u; st[3] = 0xa54ff53au;
    st[4] = 0x510e527fu; st[5] = 0x9b05688cu; st[6] = 0x1f83d9abu; st[7] = 0x5be0cd19u;
}

#define AHV_K256_LIST                                                                                       \
    0x428a2f98u, 0x71374491u, 0xb5c0fbcfu, 0xe9b5dba5u, 0x3956c25bu, 0x59f111f1u, 0x923f82a4u, 0xab1c5ed5u, \
    0xd807aa98u, 0x12835b01u, 0x243185beu, 0x550c7dc3u, 0x72be5d74u, 0x80deb1feu, 0x9bdc06a7u, 0xc19bf174u, \
    0xe49b69c1u, 0xefbe4786u, 0x0fc19dc6u, 0x240ca1ccu, 0x2de92c6fu, 0x4a7484aau, 0x5cb0a9dcu, 0x76f988dau, \
    0x983e5152u, 0xa831c66du, 0xb00327c8u, 0xbf597fc7u, 0xc6e00bf3u, 0xd5a79147u, 0x06ca6351u, 0x14292967u, \
    0x27b70a85u, 0x2e1b2138u, 0x4d2c6dfcu, 0x53380d13u, 0x650a7354u, 0x766a0abbu, 0x81c2c92eu, 0x92722c85u, \
    0xa2bfe8a1u, 0xa81a664bu, 0xc24b8b70u, 0xc76c51a3u, 0xd192e819u, 0xd6990624u, 0xf40e3585u, 0x106aa070u, \
    0x19a4c116u, 0x1e376c08u, 0x2748774cu, 0x34b0bcb5u, 0x391c0cb3u, 0x4ed8aa4au, 0x5b9cca4fu, 0x682e6ff3u, \
    0x748f82eeu, 0x78a5636fu, 0x84c87814u, 0x8cc70208u, 0x90befffau, 0xa4506cebu, 0xbef9a3f7u, 0xc67178f2u

// One SHA-256 round on rotating names (no register moves after unrolling).
#define AHV_ROUND(a, b, c, d, e, f, g, h, kw)                              \
    {                                                                      \
        uint32_t t1_ = (h) + Sig1(e) + Ch(e, f, g) + (kw);                 \
        (d) += t1_;                                                        \
        (h) = t1_ + Sig0(a) + Maj(a, b, c);                                \
    }

// Compress one 64-byte block given as 16 big-endian words.  w[] is consumed
// (rolling 16-word schedule kept in registers).
AHV_HD void sha256_compress(uint32_t st[8], uint32_t w[16])
{
    const uint32_t K[64] = {AHV_K256_LIST};
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        if (t >= 16) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j;
                w[i & 15] += sig1(w[(i - 2) & 15]) + w[(i - 7) & 15] + sig0(w[(i - 15) & 15]);
            }
        }
        AHV_ROUND(a, b, c, d, e, f, g, h, K[t + 0] + w[(t + 0) & 15])
        AHV_ROUND(h, a, b, c, d, e, f, g, K[t + 1] + w[(t + 1) & 15])
        AHV_ROUND(g, h, a, b, c, d, e, f, K[t + 2] + w[(t + 2) & 15])
        AHV_ROUND(f, g, h, a, b, c, d, e, K[t + 3] + w[(t + 3) & 15])
        AHV_ROUND(e, f, g, h, a, b, c, d, K[t + 4] + w[(t + 4) & 15])
        AHV_ROUND(d, e, f, g, h, a, b, c, K[t + 5] + w[(t + 5) & 15])
        AHV_ROUND(c, d, e, f, g, h, a, b, K[t + 6] + w[(t + 6) & 15])
        AHV_ROUND(b, c, d, e, f, g, h, a, K[t + 7] + w[(t + 7) & 15])
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// Compress a block whose K+W values are known constants (padding-only block).
// kw may live in registers, constant memory or kernel parameters.
template <class KWT>
AHV_HD void sha256_compress_kw(uint32_t st[8], const KWT &kw)
{
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        AHV_ROUND(a, b, c, d, e, f, g, h, kw.v[t + 0])
        AHV_ROUND(h, a, b, c, d, e, f, g, kw.v[t + 1])
        AHV_ROUND(g, h, a, b, c, d, e, f, kw.v[t + 2])
        AHV_ROUND(f, g, h, a, b, c, d, e, kw.v[t + 3])
        AHV_ROUND(e, f, g, h, a, b, c, d, kw.v[t + 4])
        AHV_ROUND(d, e, f, g, h, a, b, c, kw.v[t + 5])
        AHV_ROUND(c, d, e, f, g, h, a, b, kw.v[t + 6])
        AHV_ROUND(b, c, d, e, f, g, h, a, kw.v[t + 7])
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// ---------------------------------------------------------------- scaled-sum compression
// ptxas turns "a*1+b" back into an add and then places it itself (mostly IADD3 on the ALU
// pipe, which SHA-256 already saturates).  To move the additions to the FMA pipe for real,
// sums are accumulated in a scaled domain: with C odd and C*CI = 1 (mod 2^32),
//     (x1 + x2 + ... + xn) * C   is a chain of   acc = x_i * C + acc      (IMAD, immediate C)
// and the final  acc * CI + y  both unscales and adds one more term.  Every step is an IMAD
// with an immediate multiplier (two register reads), which co-issues with the ALU pipe at
// ~0.96 IPC (measured), and none of them can be rewritten as an add.
constexpr uint32_t SHA_SC = 0x9E3779B1u;
constexpr uint32_t inv32(uint32_t c)
{
    uint32_t x = c;                         // Newton: x <- x * (2 - c*x), doubles correct bits
    for (int i = 0; i < 5; ++i) x *= 2u - c * x;
    return x;
}
constexpr uint32_t SHA_SCI = inv32(SHA_SC);
static_assert(SHA_SC * SHA_SCI == 1u, "scale constant must be invertible");

AHV_HD uint32_t imad_c(uint32_t a, uint32_t b)        // a*C + b
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mad.lo.u32 %0, %1, 0x9E3779B1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
#else
    return a * SHA_SC + b;
#endif
}

AHV_HD uint32_t imad_ci(uint32_t a, uint32_t b)       // a*CI + b
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "n"(SHA_SCI), "r"(b));
    return d;
#else
    return a * SHA_SCI + b;
#endif
}

AHV_HD uint32_t imad_rk(uint32_t a, uint32_t rc, uint32_t kc)   // a*rc + kc, rc = C at run time, kc immediate
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(rc), "r"(kc));
    return d;
#else
    return a * rc + kc;
#endif
}

AHV_HD uint32_t mul_c(uint32_t a)                     // a*C
{
#if defined(__CUDA_ARCH__)
    uint32_t d;
    asm("mul.lo.u32 %0, %1, 0x9E3779B1;" : "=r"(d) : "r"(a));
    return d;
#else
    return a * SHA_SC;
#endif
}

// t1*C = (h + K + w + Ch + Sig1)*C ; e' = d + t1 ; a' = t1 + Maj + Sig0
// One round in the scaled domain.  t1*C = (h + Ch + Sig1 + (w+K))*C is a chain of IMADs whose first
// step multiplies by the run-time copy of C; e' = d + t1 and a' = t1 + Maj + Sig0 unscale with CI.
// (w + K stays one ALU add: folding it into an IMAD immediate was measured slower.)
#define AHV_ROUND_SCP(a, b, c, d, e, f, g, h, k, w)                                \
    {                                                                              \
        uint32_t s_ = imad_c((w) + (k), imad_rk(h, rc_, 0u));                      \
        s_ = imad_c(Ch(e, f, g), s_);                                              \
        s_ = imad_c(Sig1(e), s_);                                                  \
        (d) = imad_ci(s_, d);                                                      \
        (h) = imad_ci(imad_c(Maj(a, b, c), s_), Sig0(a));                          \
    }

// Production scaled-sum compression (the formulation chosen by tools/lab3.py / lab4.py).
// rc_ must hold SHA_SC at run time (kernel parameter).
AHV_HD void sha256_compress_scaled(uint32_t st[8], uint32_t w[16], uint32_t rc_)
{
    const uint32_t K[64] = {AHV_K256_LIST};
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        if (t >= 16) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = t + j;
                // w[i] = w[i-16] + (w[i-7] + sig1 + sig0): inner sum scaled, the unscale adds w[i-16]
                uint32_t z = imad_rk(w[(i - 7) & 15], rc_, 0u);
                z = imad_c(sig1(w[(i - 2) & 15]), z);
                z = imad_c(sig0(w[(i - 15) & 15]), z);
                w[i & 15] = imad_ci(z, w[i & 15]);
            }
        }
        AHV_ROUND_SCP(a, b, c, d, e, f, g, h, K[t + 0], w[(t + 0) & 15])
        AHV_ROUND_SCP(h, a, b, c, d, e, f, g, K[t + 1], w[(t + 1) & 15])
        AHV_ROUND_SCP(g, h, a, b, c, d, e, f, K[t + 2], w[(t + 2) & 15])
        AHV_ROUND_SCP(f, g, h, a, b, c, d, e, K[t + 3], w[(t + 3) & 15])
        AHV_ROUND_SCP(e, f, g, h, a, b, c, d, K[t + 4], w[(t + 4) & 15])
        AHV_ROUND_SCP(d, e, f, g, h, a, b, c, K[t + 5], w[(t + 5) & 15])
        AHV_ROUND_SCP(c, d, e, f, g, h, a, b, K[t + 6], w[(t + 6) & 15])
        AHV_ROUND_SCP(b, c, d, e, f, g, h, a, K[t + 7], w[(t + 7) & 15])
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// (K+W)*C table for a padding-only block, for sha256_compress_kw_scaled
constexpr KW64 make_kw_pad_scaled(uint64_t bitlen)
{
    KW64 r = make_kw_pad(bitlen);
    for (int t = 0; t < 64; ++t) r.v[t] *= SHA_SC;
    return r;
}

#define AHV_ROUND_SCKW(a, b, c, d, e, f, g, h, kwc)                                \
    {                                                                              \
        uint32_t s_ = imad_rk(h, rc_, 0u) + (kwc);                                 \
        s_ = imad_c(Ch(e, f, g), s_);                                              \
        s_ = imad_c(Sig1(e), s_);                                                  \
        (d) = imad_ci(s_, d);                                                      \
        (h) = imad_ci(imad_c(Maj(a, b, c), s_), Sig0(a));                          \
    }

// Padding-only block in the scaled-sum formulation; kwc.v[t] = (K[t] + W[t]) * C.
template <class KWT>
AHV_HD void sha256_compress_kw_scaled(uint32_t st[8], const KWT &kwc, uint32_t rc_)
{
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll
    for (int t = 0; t < 64; t += 8) {
        AHV_ROUND_SCKW(a, b, c, d, e, f, g, h, kwc.v[t + 0])
        AHV_ROUND_SCKW(h, a, b, c, d, e, f, g, kwc.v[t + 1])
        AHV_ROUND_SCKW(g, h, a, b, c, d, e, f, kwc.v[t + 2])
        AHV_ROUND_SCKW(f, g, h, a, b, c, d, e, kwc.v[t + 3])
        AHV_ROUND_SCKW(e, f, g, h, a, b, c, d, kwc.v[t + 4])
        AHV_ROUND_SCKW(d, e, f, g, h, a, b, c, kwc.v[t + 5])
        AHV_ROUND_SCKW(c, d, e, f, g, h, a, b, kwc.v[t + 6])
        AHV_ROUND_SCKW(b, c, d, e, f, g, h, a, kwc.v[t + 7])
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// The same block with its 64 rounds rolled into 8 iterations of 8 (the register roles are back in place
// after 8 rounds), the constants read from a table by index: ~140 instructions of code instead of ~1000.
// For kernels whose straight-line code would otherwise spill out of the 32 KB instruction cache.
AHV_HD void sha256_compress_kw_scaled_rolled(uint32_t st[8], const uint32_t *kwc, uint32_t rc_)
{
    uint32_t a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
#pragma unroll 1
    for (int t = 0; t < 64; t += 8) {
        AHV_ROUND_SCKW(a, b, c, d, e, f, g, h, kwc[t + 0])
        AHV_ROUND_SCKW(h, a, b, c, d, e, f, g, kwc[t + 1])
        AHV_ROUND_SCKW(g, h, a, b, c, d, e, f, kwc[t + 2])
        AHV_ROUND_SCKW(f, g, h, a, b, c, d, e, kwc[t + 3])
        AHV_ROUND_SCKW(e, f, g, h, a, b, c, d, kwc[t + 4])
        AHV_ROUND_SCKW(d, e, f, g, h, a, b, c, kwc[t + 5])
        AHV_ROUND_SCKW(c, d, e, f, g, h, a, b, kwc[t + 6])
        AHV_ROUND_SCKW(b, c, d, e, f, g, h, a, kwc[t + 7])
    }
    st[0] += a; st[1] += b; st[2] += c; st[3] += d; st[4] += e; st[5] += f; st[6] += g; st[7] += h;
}

// Compression policies: which formulation a kernel uses.  Native = plain additions (ptxas keeps
// most of them on the ALU pipe); Scaled = additions as IMAD chains on the FMA pipe (rc = SHA_SC
// passed at run time).  Both give identical results (tests/test_hostemu.py, GPU parity tests).
struct CompressNative {
    AHV_HD void block(uint32_t st[8], uint32_t w[16]) const { sha256_compress(st, w); }
    AHV_HD void node_block(uint32_t st[8], uint32_t w[16]) const { sha256_compress(st, w); }
    template <class KWT>
    AHV_HD void pad(uint32_t st[8], const KWT &kw) const { sha256_compress_kw(st, kw); }
    AHV_HD void node_pad(uint32_t st[8]) const
    {
        constexpr KW64 kw = make_kw_pad(1024);
        sha256_compress_kw(st, kw);
    }
    static constexpr bool scaled = false;
};

struct CompressScaled {
    uint32_t rc;
    AHV_HD void block(uint32_t st[8], uint32_t w[16]) const { sha256_compress_scaled(st, w, rc); }
    // a hex block of a Merkle node.  Rolling THIS body as well (sha256_lab.cuh: sha256_compress_scaled_rolled)
    // brings a node to 27 KB / 19 KB of code, inside the instruction cache, but measured slower: 9.34 / 9.36
    // ms per bench step against 9.27 - the rolled rounds lose their immediate constants.
    AHV_HD void node_block(uint32_t st[8], uint32_t w[16]) const { sha256_compress_scaled(st, w, rc); }
    // (the leaf kernels keep this block straight-line: 4 of their 5 compressions run the same 24 KB body, so
    // instruction fetch is not their problem - rolling it measured 5 % slower, 2.96 vs 2.81 ms per launch)
    template <class KWT>
    AHV_HD void pad(uint32_t st[8], const KWT &kwc) const { sha256_compress_kw_scaled(st, kwc, rc); }
    AHV_HD void node_pad(uint32_t st[8]) const
    {
        // rolled: with the straight-line version a node is 48 KB of code, 16 KB past the instruction cache
        // (ncu: no_instruction 2.1 stalls per issue, ALU pipe 89 %); 34 KB with this one, 4 % faster
        static constexpr KW64 kwc = make_kw_pad_scaled(1024);
        sha256_compress_kw_scaled_rolled(st, kwc.v, rc);
    }
    static constexpr bool scaled = true;
};

// ---------------------------------------------------------------- hex expansion
// 4 nibble-valued bytes -> 4 lowercase hex ASCII bytes.
AHV_HD uint32_t nib2ascii(uint32_t x)
{
    uint32_t m = ((x + 0x06060606u) >> 4) & 0x01010101u;   // 1 where nibble >= 10
    return x + 0x30303030u + m * 0x27u;                    // '0'+n  or  'a'+n-10
}

// one state word (8 nibbles) -> two big-endian message words of ASCII hex
AHV_HD void hex_expand(uint32_t v, uint32_t &w0, uint32_t &w1)
{
    uint32_t hi = nib2ascii((v >> 4) & 0x0f0f0f0fu);   // chars of nibbles 7,5,3,1
    uint32_t lo = nib2ascii(v & 0x0f0f0f0fu);          // chars of nibbles 6,4,2,0
    w0 = prmt(hi, lo, 0x3726u);                        // n7 n6 n5 n4
    w1 = prmt(hi, lo, 0x1504u);                        // n3 n2 n1 n0
}

// audit/delta.py:130 : sha256((left_hex + right_hex).encode()) on state-word digests.
template <class P>
AHV_HD void merkle_node(const uint32_t l[8], const uint32_t r[8], uint32_t out[8], const P &cp)
{
    uint32_t st[8], w[16];
    sha256_init(st);
    // one rolled loop over the two hex blocks: a single copy of the compression code keeps the
    // kernel inside the instruction cache (an unrolled node is ~85 KB of straight-line code and
    // stalls on instruction fetch: "no_instruction" 6.8 per issue in ncu, profiles/r01)
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
#pragma unroll
        for (int i = 0; i < 8; ++i) hex_expand(half ? r[i] : l[i], w[2 * i], w[2 * i + 1]);
        cp.node_block(st, w);
    }
    cp.node_pad(st);
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = st[i];
}

AHV_HD void merkle_node(const uint32_t l[8], const uint32_t r[8], uint32_t out[8])
{
    merkle_node(l, r, out, CompressNative());
}

// ---------------------------------------------------------------- digest <-> memory
struct alignas(16) U4 {
    uint32_t x, y, z, w;
};

// state words -> the 32 raw digest bytes at a 16-byte aligned address
AHV_HD void store_digest(uint8_t *dst, const uint32_t st[8])
{
    U4 *p = reinterpret_cast<U4 *>(dst);
    U4 a = {bswap32(st[0]), bswap32(st[1]), bswap32(st[2]), bswap32(st[3])};
    U4 b = {bswap32(st[4]), bswap32(st[5]), bswap32(st[6]), bswap32(st[7])};
    p[0] = a;
    p[1] = b;
}

AHV_HD void load_digest(const uint8_t *src, uint32_t st[8])
{
    const U4 *p = reinterpret_cast<const U4 *>(src);
    U4 a = p[0], b = p[1];
    st[0] = bswap32(a.x); st[1] = bswap32(a.y); st[2] = bswap32(a.z); st[3] = bswap32(a.w);
    st[4] = bswap32(b.x); st[5] = bswap32(b.y); st[6] = bswap32(b.z); st[7] = bswap32(b.w);
}

AHV_HD bool digest_equal_mem(const uint8_t *src, const uint32_t st[8])
{
    uint32_t p[8];
    load_digest(src, p);
    uint32_t diff = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) diff |= p[i] ^ st[i];
    return diff == 0;
}

// Same node hash with the children still in memory (raw 32-byte digests): the right child is
// fetched only when its block is due, which keeps 8 registers free in the level kernel.
template <class P>
AHV_HD void merkle_node_mem(const uint8_t *lp, const uint8_t *rp, uint32_t out[8], const P &cp)
{
    uint32_t st[8], w[16];
    sha256_init(st);
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
        uint32_t x[8];
        load_digest(half ? rp : lp, x);
#pragma unroll
        for (int i = 0; i < 8; ++i) hex_expand(x[i], w[2 * i], w[2 * i + 1]);
        cp.node_block(st, w);
    }
    cp.node_pad(st);
#pragma unroll
    for (int i = 0; i < 8; ++i) out[i] = st[i];
}

// ---------------------------------------------------------------- whole messages
// Word loader over an arbitrarily aligned byte string.  Reads only 4-byte aligned
// words that contain at least one byte of [p, p+len); never touches memory past the
// aligned word holding the last message byte.
struct MsgReader {
    const uint32_t *q;   // aligned base
    uint32_t a;          // p & 3
    uint32_t sel;        // PRMT selector doing alignment + big-endian swap in one op
    AHV_HD void init(const uint8_t *p)
    {
        a = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 3u);
        q = reinterpret_cast<const uint32_t *>(p - a);
        sel = 0x0123u + 0x1111u * a;
    }
};

// 17 raw aligned words covering one 64-byte block at any byte alignment: four 16-byte loads when the
// block is 16-byte aligned, else 4-byte loads (the 17th only when the block really spills into it).
AHV_HD void load_block_raw(const uint32_t *qb, bool aligned16, uint32_t a, uint32_t x[17])
{
    if (aligned16) {
        const U4 *q4 = reinterpret_cast<const U4 *>(qb);
        const U4 v0 = q4[0], v1 = q4[1], v2 = q4[2], v3 = q4[3];
        x[0] = v0.x; x[1] = v0.y; x[2] = v0.z; x[3] = v0.w;
        x[4] = v1.x; x[5] = v1.y; x[6] = v1.z; x[7] = v1.w;
        x[8] = v2.x; x[9] = v2.y; x[10] = v2.z; x[11] = v2.w;
        x[12] = v3.x; x[13] = v3.y; x[14] = v3.z; x[15] = v3.w;
        x[16] = 0u;
    } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = qb[j];
        x[16] = a ? qb[16] : 0u;
    }
}

// SHA-256 of len bytes at p (any alignment).  One compress copy: full blocks take the fast word
// path (next block's raw words are fetched before the current block is compressed), the final
// one or two blocks build padding per word.
//
// sha256_message_from continues from the state after the first b0 (full) blocks - the chained capture
// hashes the blocks in front of the parent field ahead of time (sha256_prefix), in parallel over all deltas.
template <class P>
AHV_HD void sha256_message_from(const uint8_t *p, uint64_t len, uint32_t st[8], uint64_t b0, const P &cp)
{
    MsgReader rd;
    rd.init(p);
    const uint64_t nfull = len >> 6;
    const uint32_t rem = (uint32_t)(len & 63u);
    const uint64_t nblk = nfull + ((rem + 9u > 64u) ? 2u : 1u);
    const bool aligned16 = ((reinterpret_cast<uintptr_t>(p) & 15u) == 0);
    uint32_t x[17];
    if (b0 < nfull) load_block_raw(rd.q + 16 * b0, aligned16, rd.a, x);
    for (uint64_t b = b0; b < nblk; ++b) {
        uint32_t w[16];
        const uint32_t *qb = rd.q + 16 * b;
        if (b < nfull) {
#pragma unroll
            for (int j = 0; j < 16; ++j) w[j] = prmt(x[j], x[j + 1], rd.sel);
            if (b + 1 < nfull) load_block_raw(qb + 16, aligned16, rd.a, x);
        } else {
            // tail: byte position of word j relative to the start of the tail data
            const uint32_t tb = (uint32_t)(b - nfull) * 64u;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const uint32_t pos = tb + 4u * j;
                uint32_t v = 0;
                if (pos < rem) {
                    uint32_t x0 = qb[j];
                    uint32_t x1 = (rd.a && pos + 4u - rd.a < rem) ? qb[j + 1] : 0u;
                    v = prmt(x0, x1, rd.sel);
                    const uint32_t nvalid = rem - pos;             // >= 1
                    if (nvalid < 4u) {
                        v &= 0xffffffffu << (32u - 8u * nvalid);
                        v |= 0x80000000u >> (8u * nvalid);
                    }
                } else if (pos == rem) {
                    v = 0x80000000u;
                }
                w[j] = v;
            }
            if (b == nblk - 1) {
                const uint64_t bits = len << 3;
                w[14] = (uint32_t)(bits >> 32);
                w[15] = (uint32_t)bits;
            }
        }
        cp.block(st, w);
    }
}

template <class P>
AHV_HD void sha256_message(const uint8_t *p, uint64_t len, uint32_t st[8], const P &cp)
{
    sha256_init(st);
    sha256_message_from(p, len, st, 0, cp);
}

// State after the first nb full blocks of the message at p (nb * 64 <= its length).
template <class P>
AHV_HD void sha256_prefix(const uint8_t *p, uint64_t nb, uint32_t st[8], const P &cp)
{
    MsgReader rd;
    rd.init(p);
    sha256_init(st);
    const bool aligned16 = ((reinterpret_cast<uintptr_t>(p) & 15u) == 0);
    uint32_t x[17];
    if (nb) load_block_raw(rd.q, aligned16, rd.a, x);
    for (uint64_t b = 0; b < nb; ++b) {
        uint32_t w[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = prmt(x[j], x[j + 1], rd.sel);
        if (b + 1 < nb) load_block_raw(rd.q + 16 * (b + 1), aligned16, rd.a, x);
        cp.block(st, w);
    }
}

AHV_HD void sha256_message(const uint8_t *p, uint64_t len, uint32_t st[8])
{
    sha256_message(p, len, st, CompressNative());
}

}  // namespace ahv

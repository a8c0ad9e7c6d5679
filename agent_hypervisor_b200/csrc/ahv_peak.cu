// ahv_peak.cu - integer-pipe ceiling microbenchmarks (INT_PEAK must be measured).
// Register-only kernels, no memory traffic; every SM gets 4 x 256 threads.
//   mode 0: SHF + LOP3 only (instructions that can only issue on the ALU pipe)
//   mode 1: the SHA-256 round function on constant K+W, two per-thread states
//           (6 SHF : 4 LOP3 : 4 add per round; the compiler places the adds)
//   mode 2: full compressions (message schedule included) on register-resident blocks
//   mode 3: IMAD only (multiplier from a kernel argument)            - FMA-pipe integer rate
//   mode 4: SHF + LOP3 + 2 IMAD per slot                             - do the two pipes co-issue?
//   mode 5: IMAD.WIDE only        mode 6: IMAD.HI only
//   mode 7: SHF + LOP3 + IMAD.WIDE per slot
//   mode 8: full compressions with every add issued as IMAD (sha256_compress_imad)
//   mode 9: mode 8 + message-schedule sigmas through IMAD.WIDE (sha256_compress_wide)
#include "ahv_internal.cuh"
#include "sha256_lab.cuh"

using namespace ahv;

namespace {

constexpr int PEAK_UNROLL = 8;

__global__ void __launch_bounds__(256) k_peak_alu(int iters, uint32_t *out)
{
    uint32_t a[PEAK_UNROLL], b = threadIdx.x * 2654435761u + 1u, c = blockIdx.x * 40503u + 7u;
#pragma unroll
    for (int k = 0; k < PEAK_UNROLL; ++k) a[k] = b + k * 0x9e3779b9u;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < PEAK_UNROLL; ++k) {
            uint32_t t;
            asm("shf.r.wrap.b32 %0, %1, %2, 7;" : "=r"(t) : "r"(a[k]), "r"(a[(k + 1) % PEAK_UNROLL]));
            asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[k]) : "r"(t), "r"(b), "r"(c));
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < PEAK_UNROLL; ++k) r ^= a[k];
    if (r == 0x12345678u) out[0] = r;   // practically never; keeps the chain live
}

// MODE: 3 imad, 4 alu+2imad, 5 wide, 6 hi, 7 alu+wide
template <int MODE>
__global__ void __launch_bounds__(256) k_peak_mix(int iters, uint32_t m, uint32_t *out)
{
    uint32_t a[PEAK_UNROLL], f[PEAK_UNROLL], g[PEAK_UNROLL];
    uint32_t b = threadIdx.x * 2654435761u + 1u, c = blockIdx.x * 40503u + 7u;
#pragma unroll
    for (int k = 0; k < PEAK_UNROLL; ++k) {
        a[k] = b + k * 0x9e3779b9u;
        f[k] = c + k * 0x85ebca6bu + threadIdx.x;
        g[k] = b ^ (k * 0xc2b2ae35u);
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < PEAK_UNROLL; ++k) {
            if (MODE == 4 || MODE == 7) {
                uint32_t t;
                asm("shf.r.wrap.b32 %0, %1, %2, 7;" : "=r"(t) : "r"(a[k]), "r"(a[(k + 1) % PEAK_UNROLL]));
                asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[k]) : "r"(t), "r"(b), "r"(c));
            }
            if (MODE == 3 || MODE == 4) {
                asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(f[k]) : "r"(f[k]), "r"(m), "r"(g[k]));
                asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(g[k]) : "r"(g[k]), "r"(m), "r"(f[(k + 1) % PEAK_UNROLL]));
            }
            if (MODE == 5 || MODE == 7) {
                uint64_t p;
                asm("mul.wide.u32 %0, %1, %2;" : "=l"(p) : "r"(f[k]), "r"(m));
                f[k] = (uint32_t)p;
                g[k] = (uint32_t)(p >> 32) + g[k];   // consume both halves (one extra add, counted below)
            }
            if (MODE == 10) {
                uint32_t t;
                asm("shf.r.wrap.b32 %0, %1, %2, 7;" : "=r"(t) : "r"(a[k]), "r"(a[(k + 1) % PEAK_UNROLL]));
                asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[k]) : "r"(t), "r"(b), "r"(c));
                asm("mad.lo.u32 %0, %1, 0x9E3779B1, %2;" : "=r"(f[k]) : "r"(f[k]), "r"(g[k]));
                asm("mad.lo.u32 %0, %1, 0x85EBCA6B, %2;" : "=r"(g[k]) : "r"(g[k]), "r"(f[(k + 1) % PEAK_UNROLL]));
            }
            if (MODE == 11) {   // 3 ALU : 1 IMAD(imm)
                uint32_t t;
                asm("shf.r.wrap.b32 %0, %1, %2, 7;" : "=r"(t) : "r"(a[k]), "r"(a[(k + 1) % PEAK_UNROLL]));
                asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[k]) : "r"(t), "r"(b), "r"(c));
                asm("shf.r.wrap.b32 %0, %1, %2, 11;" : "=r"(g[k]) : "r"(g[k]), "r"(a[k]));
                asm("mad.lo.u32 %0, %1, 0x9E3779B1, %2;" : "=r"(f[k]) : "r"(f[k]), "r"(g[k]));
            }
            if (MODE == 6) {
                asm("mul.hi.u32 %0, %1, %2;" : "=r"(f[k]) : "r"(f[k] | 0x80000000u), "r"(m));
            }
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < PEAK_UNROLL; ++k) r ^= a[k] ^ f[k] ^ g[k];
    if (r == 0x12345678u) out[0] = r;
}

__global__ void __launch_bounds__(256) k_peak_rounds(int iters, uint32_t *out)
{
    uint32_t s0[8], s1[8];
    sha256_init(s0);
    sha256_init(s1);
    s0[0] ^= threadIdx.x;
    s1[0] ^= threadIdx.x * 0x9e3779b9u + 0x55u;    // per-thread: keeps it off the uniform datapath
    constexpr KW64 kw = make_kw_pad(2048);
    for (int it = 0; it < iters; ++it) {
        sha256_compress_kw(s0, kw);
        sha256_compress_kw(s1, kw);
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) r ^= s0[k] ^ s1[k];
    if (r == 0x12345678u) out[0] = r;
}

// compress with a compile-time add-site mask (see sha256_compress_sel); NST independent states
template <uint32_t MASK, int NST>
__global__ void __launch_bounds__(256) k_peak_sel(int iters, const __grid_constant__ ShaMults mu, uint32_t *out)
{
    uint32_t st[NST][8];
#pragma unroll
    for (int q = 0; q < NST; ++q) {
        sha256_init(st[q]);
        st[q][0] ^= threadIdx.x * (q + 1);
        st[q][1] ^= blockIdx.x + q;
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int q = 0; q < NST; ++q) {
            uint32_t w[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) w[k] = st[q][k & 7] ^ (0x01010101u * k);
            sha256_compress_sel<MASK>(st[q], w, mu);
        }
    }
    uint32_t r = 0;
#pragma unroll
    for (int q = 0; q < NST; ++q)
#pragma unroll
        for (int k = 0; k < 8; ++k) r ^= st[q][k];
    if (r == 0x12345678u) out[0] = r;
}

// scaled compress with K*C taken from a kernel-parameter table (constant-bank operand)
template <int VAR>
__global__ void __launch_bounds__(256) k_peak_kct(int iters, const __grid_constant__ ShaMults mu,
                                                  const __grid_constant__ KW64 kct, uint32_t *out)
{
    uint32_t st[8];
    sha256_init(st);
    st[0] ^= threadIdx.x;
    st[1] ^= blockIdx.x;
    for (int it = 0; it < iters; ++it) {
        uint32_t w[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) w[k] = st[k & 7] ^ (0x01010101u * k);
        sha256_compress_scaled_t<VAR>(st, w, mu.c, kct);
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) r ^= st[k];
    if (r == 0x12345678u) out[0] = r;
}

// dependent fp64 additions: latency of the serial chain a sequential row sum is made of
__global__ void __launch_bounds__(256) k_peak_dadd_chain(int iters, double x, double *out)
{
    double acc = (double)threadIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) acc = __dadd_rn(acc, x);
    }
    if (acc == -1.0) out[0] = acc;
}

template <int A, int B, int C, int D>
__global__ void __launch_bounds__(256) k_peak_search(int iters, uint32_t rc, uint32_t *out)
{
    uint32_t st[8];
    sha256_init(st);
    st[0] ^= threadIdx.x;
    st[1] ^= blockIdx.x;
    for (int it = 0; it < iters; ++it) {
        uint32_t w[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) w[k] = st[k & 7] ^ (0x01010101u * k);
        sha256_compress_search<A, B, C, D>(st, w, rc);
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) r ^= st[k];
    if (r == 0x12345678u) out[0] = r;
}

template <int VARIANT>
__global__ void __launch_bounds__(256) k_peak_compress(int iters, const __grid_constant__ ShaMults mu, uint32_t *out)
{
    uint32_t st[8];
    sha256_init(st);
    st[0] ^= threadIdx.x;
    st[1] ^= blockIdx.x;
    for (int it = 0; it < iters; ++it) {
        uint32_t w[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) w[k] = st[k & 7] ^ (0x01010101u * k);
        if (VARIANT == 0) sha256_compress(st, w);
        else if (VARIANT == 1) sha256_compress_imad(st, w, mu);
        else if (VARIANT == 2) sha256_compress_sel<0xFFEu>(st, w, mu);
        else if (VARIANT == 3) sha256_compress_scaled_v<1>(st, w, mu.c);
        else if (VARIANT == 4) sha256_compress_scaled_v<0>(st, w, mu.c);
        else if (VARIANT == 5) sha256_compress_scaled_v<2>(st, w, mu.c);
        else if (VARIANT == 6) sha256_compress_scaled_v<3>(st, w, mu.c);
        else if (VARIANT == 7) sha256_compress_scaled_v<8 + 1>(st, w, mu.one - 1u);   // run-time zero addend
        else if (VARIANT == 8) sha256_compress_scaled_v<8>(st, w, mu.one - 1u);
        else if (VARIANT == 9) sha256_compress_scaled_v<3 + 16>(st, w, mu.c);    // + plain shifts as IMAD.HI
        else if (VARIANT == 10) sha256_compress_scaled_v<3 + 32>(st, w, mu.c);   // h+K+w as one IADD3
        else sha256_compress_scaled_v<3 + 16 + 32>(st, w, mu.c);                 // both
    }
    uint32_t r = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) r ^= st[k];
    if (r == 0x12345678u) out[0] = r;
}

}  // namespace

extern "C" int ahv_int_peak(ahv_ctx *ctx, int mode, int iters, double *out_instr_per_s, float *out_ms)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(((mode >= 0 && mode <= 23) || (mode >= 100 && mode < 100 + 18) || (mode >= 200 && mode < 200 + 18) || (mode >= 300 && mode < 348)) &&
                    iters > 0 && out_instr_per_s, "bad ahv_int_peak arguments");
    AHV_TRY(scratch_reserve(ctx, 256));
    uint32_t *d = (uint32_t *)scratch_take(ctx, 8);
    const unsigned grid = (unsigned)ctx->sm_count * (unsigned)(ctx->opt_peak_ctas > 0 ? ctx->opt_peak_ctas : 4), block = 256;
    const ShaMults mu = make_sha_mults();
    cudaEvent_t e0, e1;
    AHV_CUDA(cudaEventCreate(&e0));
    AHV_CUDA(cudaEventCreate(&e1));
    double per_thread = 0;
    for (int rep = 0; rep < 2; ++rep) {      // first repetition warms up
        AHV_CUDA(cudaEventRecord(e0, ctx->stream));
        const double slots = (double)iters * PEAK_UNROLL;
        switch (mode) {
        case 0: k_peak_alu<<<grid, block, 0, ctx->stream>>>(iters, d); per_thread = slots * 2.0; break;
        case 1: k_peak_rounds<<<grid, block, 0, ctx->stream>>>(iters, d); per_thread = (double)iters * 2.0 * (64.0 * 14.0 + 8.0); break;
        case 2: k_peak_compress<0><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 3: k_peak_mix<3><<<grid, block, 0, ctx->stream>>>(iters, 1u, d); per_thread = slots * 2.0; break;
        case 4: k_peak_mix<4><<<grid, block, 0, ctx->stream>>>(iters, 1u, d); per_thread = slots * 4.0; break;
        case 5: k_peak_mix<5><<<grid, block, 0, ctx->stream>>>(iters, 3u, d); per_thread = slots * 2.0; break;
        case 6: k_peak_mix<6><<<grid, block, 0, ctx->stream>>>(iters, 0xfffffff1u, d); per_thread = slots * 2.0; break;
        case 7: k_peak_mix<7><<<grid, block, 0, ctx->stream>>>(iters, 3u, d); per_thread = slots * 4.0; break;
        case 8: k_peak_compress<1><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 10: k_peak_mix<10><<<grid, block, 0, ctx->stream>>>(iters, 1u, d); per_thread = slots * 4.0; break;
        case 11: k_peak_mix<11><<<grid, block, 0, ctx->stream>>>(iters, 1u, d); per_thread = slots * 4.0; break;
        case 12: k_peak_compress<3><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 13: k_peak_compress<4><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 14: k_peak_compress<5><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 15: k_peak_compress<6><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 16: { const KW64 kct = make_kc_table(); k_peak_kct<4><<<grid, block, 0, ctx->stream>>>(iters, mu, kct, d); per_thread = (double)iters * 1400.0; break; }
        case 17: { const KW64 kct = make_kc_table(); k_peak_kct<5><<<grid, block, 0, ctx->stream>>>(iters, mu, kct, d); per_thread = (double)iters * 1400.0; break; }
        case 18: k_peak_compress<7><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 19: k_peak_compress<8><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 20:   // one warp per SM: nothing hides the latency; "instr/s" = additions per second per thread x threads
            k_peak_dadd_chain<<<(unsigned)ctx->sm_count, 32, 0, ctx->stream>>>(iters, 1e-9, (double *)d);
            per_thread = (double)iters * 16.0 * ((double)ctx->sm_count * 32.0) / ((double)grid * block);
            break;
        case 300: k_peak_search<0, 0, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 301: k_peak_search<0, 0, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 302: k_peak_search<0, 0, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 303: k_peak_search<0, 0, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 304: k_peak_search<0, 0, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 305: k_peak_search<0, 0, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 306: k_peak_search<0, 1, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 307: k_peak_search<0, 1, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 308: k_peak_search<0, 1, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 309: k_peak_search<0, 1, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 310: k_peak_search<0, 1, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 311: k_peak_search<0, 1, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 312: k_peak_search<1, 0, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 313: k_peak_search<1, 0, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 314: k_peak_search<1, 0, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 315: k_peak_search<1, 0, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 316: k_peak_search<1, 0, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 317: k_peak_search<1, 0, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 318: k_peak_search<1, 1, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 319: k_peak_search<1, 1, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 320: k_peak_search<1, 1, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 321: k_peak_search<1, 1, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 322: k_peak_search<1, 1, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 323: k_peak_search<1, 1, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 324: k_peak_search<2, 0, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 325: k_peak_search<2, 0, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 326: k_peak_search<2, 0, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 327: k_peak_search<2, 0, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 328: k_peak_search<2, 0, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 329: k_peak_search<2, 0, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 330: k_peak_search<2, 1, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 331: k_peak_search<2, 1, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 332: k_peak_search<2, 1, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 333: k_peak_search<2, 1, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 334: k_peak_search<2, 1, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 335: k_peak_search<2, 1, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 336: k_peak_search<3, 0, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 337: k_peak_search<3, 0, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 338: k_peak_search<3, 0, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 339: k_peak_search<3, 0, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 340: k_peak_search<3, 0, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 341: k_peak_search<3, 0, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 342: k_peak_search<3, 1, 0, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 343: k_peak_search<3, 1, 0, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 344: k_peak_search<3, 1, 0, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 345: k_peak_search<3, 1, 1, 0><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 346: k_peak_search<3, 1, 1, 1><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 347: k_peak_search<3, 1, 1, 2><<<grid, block, 0, ctx->stream>>>(iters, mu.c, d); per_thread = (double)iters * 1400.0; break;
        case 21: k_peak_compress<9><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 22: k_peak_compress<10><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 23: k_peak_compress<11><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
        case 9: k_peak_compress<2><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break;
#define SEL_CASE(idx, mask)                                                                                              \
        case 100 + idx: k_peak_sel<mask, 1><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 1400.0; break; \
        case 200 + idx: k_peak_sel<mask, 2><<<grid, block, 0, ctx->stream>>>(iters, mu, d); per_thread = (double)iters * 2800.0; break;
        SEL_CASE(0, 0x000u)    // all native
        SEL_CASE(1, 0x010u)    // d += t1
        SEL_CASE(2, 0x070u)    // d += t1, Sig0+Maj, t1+t2
        SEL_CASE(3, 0x07Eu)    // every round add except w+K
        SEL_CASE(4, 0x380u)    // schedule adds only
        SEL_CASE(5, 0x3F0u)    // schedule + {d, Sig0+Maj, t1+t2}
        SEL_CASE(6, 0xBFEu)    // everything except w+K
        SEL_CASE(7, 0xBFFu)    // everything
        SEL_CASE(8, 0x390u)    // schedule + d
        SEL_CASE(9, 0x00Eu)    // t1 chain only
        SEL_CASE(10, 0xFFEu)   // everything except w+K, sigmas via WIDE/HI
        SEL_CASE(11, 0x400u)   // only sigmas via WIDE/HI
        SEL_CASE(12, 0x060u)   // Sig0+Maj, t1+t2
        SEL_CASE(13, 0x3E0u)   // schedule + {Sig0+Maj, t1+t2}
        SEL_CASE(14, 0x200u)   // one schedule add
        SEL_CASE(15, 0x180u)   // two schedule adds
        SEL_CASE(16, 0x1000u)  // schedule right shifts as IMAD.HI
        SEL_CASE(17, 0x1380u)  // + schedule adds
        default: break;
        }
        AHV_CHECK_LAUNCH(ctx);
        AHV_CUDA(cudaEventRecord(e1, ctx->stream));
        AHV_CUDA(cudaEventSynchronize(e1));
    }
    float ms = 0;
    AHV_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    *out_instr_per_s = per_thread * (double)grid * block / (ms * 1e-3);
    if (out_ms) *out_ms = ms;
    return AHV_OK;
}

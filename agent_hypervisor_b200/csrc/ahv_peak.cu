// ahv_peak.cu - integer-pipe ceiling microbenchmark (INT_PEAK must be measured).
// Register-only kernels, no memory traffic; every SM gets 4 x 256 threads.
//   mode 0: SHF + LOP3 only (instructions that can only issue on the ALU pipe)
//   mode 1: the SHA-256 round function on constant K+W (6 SHF : 4 LOP3 : 4 add per round);
//           the compiler is free to place the adds on the ALU (IADD3) or FMA (IMAD) pipe
//   mode 2: full compressions (message schedule included) on register-resident blocks
#include "ahv_internal.cuh"
#include "sha256_core.cuh"

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

__global__ void __launch_bounds__(256) k_peak_rounds(int iters, uint32_t *out)
{
    uint32_t s0[8], s1[8];
    sha256_init(s0);
    sha256_init(s1);
    s0[0] ^= threadIdx.x;
    s1[0] ^= blockIdx.x + 0x55u;
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

__global__ void __launch_bounds__(256) k_peak_compress(int iters, uint32_t *out)
{
    uint32_t st[8];
    sha256_init(st);
    st[0] ^= threadIdx.x;
    st[1] ^= blockIdx.x;
    for (int it = 0; it < iters; ++it) {
        uint32_t w[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) w[k] = st[k & 7] ^ (0x01010101u * k);
        sha256_compress(st, w);
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
    AHV_REQUIRE(mode >= 0 && mode <= 2 && iters > 0 && out_instr_per_s, "bad ahv_int_peak arguments");
    AHV_TRY(scratch_reserve(ctx, 256));
    uint32_t *d = (uint32_t *)scratch_take(ctx, 4);
    const unsigned grid = (unsigned)ctx->sm_count * 4, block = 256;
    cudaEvent_t e0, e1;
    AHV_CUDA(cudaEventCreate(&e0));
    AHV_CUDA(cudaEventCreate(&e1));
    double per_thread = 0;
    for (int rep = 0; rep < 2; ++rep) {      // first repetition warms up
        AHV_CUDA(cudaEventRecord(e0, ctx->stream));
        if (mode == 0) {
            k_peak_alu<<<grid, block, 0, ctx->stream>>>(iters, d);
            per_thread = (double)iters * PEAK_UNROLL * 2.0;
        } else if (mode == 1) {
            k_peak_rounds<<<grid, block, 0, ctx->stream>>>(iters, d);
            per_thread = (double)iters * 2.0 * (64.0 * 14.0 + 8.0);
        } else {
            k_peak_compress<<<grid, block, 0, ctx->stream>>>(iters, d);
            per_thread = (double)iters * 1400.0;    // algorithmic count of SURVEY 8d
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

// ahv_audit.cu - SHA-256 leaf hashing, Merkle level reduction, chain verification and
// session-parallel chained capture for sm_100a, plus their C-ABI entry points.
//
// Reference behaviour (paths under /root/reference/src/hypervisor):
//   leaf hash        audit/delta.py:63
//   Merkle root      audit/delta.py:117-134
//   verify_chain     audit/delta.py:136-152
//   capture chaining audit/delta.py:101-114
//
// All of this is integer-pipe work (SHF/LOP3/IADD3): no tensor cores, HBM traffic
// is a few percent of peak.  One message (or one tree node) per thread, schedule in
// registers, 16-byte vector loads; see DESIGN.md for the per-kernel rooflines.
#include "ahv_internal.cuh"
#include "scan.cuh"

using namespace ahv::scan;
#include "sha256_core.cuh"

using namespace ahv;

namespace {

constexpr int NBINS = 1024;          // length-sort bins (block counts, clamped)
constexpr unsigned long long FB_NONE = ~0ull;

// ------------------------------------------------------------------ session map
struct SessMap {
    const uint64_t *rowptr;   // S+1 entries, or nullptr when uniform
    uint64_t S;
    uint32_t uniform;         // leaves per session when uniform, else 0
};

__device__ __forceinline__ uint64_t sess_of(const SessMap &sm, uint64_t i, uint64_t *start)
{
    if (sm.uniform) {
        uint32_t s = (uint32_t)i / sm.uniform;
        *start = (uint64_t)s * sm.uniform;
        return s;
    }
    // last s with rowptr[s] <= i  (empty sessions share a start; pick the non-empty one)
    uint64_t lo = 0, hi = sm.S;      // invariant: rowptr[lo] <= i < rowptr[hi]
    while (hi - lo > 1) {
        uint64_t mid = (lo + hi) >> 1;
        if (sm.rowptr[mid] <= i) lo = mid; else hi = mid;
    }
    *start = sm.rowptr[lo];
    return lo;
}

struct VerifyArgs {
    const uint8_t *parents;         // n*32
    const uint8_t *kind;            // n
    const uint8_t *first;           // n flags (ragged) or nullptr (uniform)
    unsigned long long *first_bad;  // S, initialised to FB_NONE
    SessMap sm;
};

__device__ __forceinline__ void report_bad(const VerifyArgs &va, uint64_t j)
{
    uint64_t start;
    uint64_t s = sess_of(va.sm, j, &start);
    atomicMin(&va.first_bad[s], (unsigned long long)(j - start));
}

// audit/delta.py:145-150 for the pair (delta i, delta i+1), given delta i's fresh digest
__device__ __forceinline__ void verify_epilogue(const VerifyArgs &va, uint64_t i, uint64_t n,
                                                const uint32_t st[8])
{
    const bool first_i = va.sm.uniform ? ((uint32_t)i % va.sm.uniform == 0) : (va.first[i] != 0);
    if (first_i && va.kind[i] != AHV_PARENT_NONE) report_bad(va, i);     // :145-147
    const uint64_t j = i + 1;
    if (j < n) {
        const bool first_j = va.sm.uniform ? ((uint32_t)j % va.sm.uniform == 0) : (va.first[j] != 0);
        if (!first_j) {
            bool ok = (va.kind[j] == AHV_PARENT_HEX) && digest_equal_mem(va.parents + 32 * j, st);  // :149
            if (!ok) report_bad(va, j);
        }
    }
}

// compression policy selected at compile time per kernel instance (see sha256_core.cuh)
template <bool SC> struct Pol;
template <> struct Pol<false> {
    typedef CompressNative type;
    static __device__ __forceinline__ CompressNative make(uint32_t) { return CompressNative(); }
};
template <> struct Pol<true> {
    typedef CompressScaled type;
    static __device__ __forceinline__ CompressScaled make(uint32_t rc)
    {
        CompressScaled p;
        p.rc = rc;
        return p;
    }
};

// ------------------------------------------------------------------ leaf kernels
// General path: any length, any alignment, optional length-sorted permutation.
template <bool VERIFY, bool SC>
__global__ void __launch_bounds__(256)
k_sha256_var(const uint8_t *__restrict__ arena, const uint64_t *__restrict__ offsets, uint64_t fixed_len,
             uint64_t n, const uint32_t *__restrict__ perm, uint8_t *__restrict__ out, uint32_t rc, VerifyArgs va)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint64_t i = perm ? perm[t] : t;
    uint64_t off, len;
    if (offsets) {
        off = offsets[i];
        len = offsets[i + 1] - off;
    } else {
        off = i * fixed_len;
        len = fixed_len;
    }
    uint32_t st[8];
    sha256_message(arena + off, len, st, cp);
    if (out) store_digest(out + 32 * i, st);
    if (VERIFY) verify_epilogue(va, i, n, st);
}

__device__ __forceinline__ U4 ldg_u4(const U4 *p)
{
    uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
    U4 r = {v.x, v.y, v.z, v.w};
    return r;
}

__device__ __forceinline__ void words_from_u4(uint32_t w[16], const U4 &v0, const U4 &v1, const U4 &v2, const U4 &v3)
{
    w[0] = bswap32(v0.x); w[1] = bswap32(v0.y); w[2] = bswap32(v0.z); w[3] = bswap32(v0.w);
    w[4] = bswap32(v1.x); w[5] = bswap32(v1.y); w[6] = bswap32(v1.z); w[7] = bswap32(v1.w);
    w[8] = bswap32(v2.x); w[9] = bswap32(v2.y); w[10] = bswap32(v2.z); w[11] = bswap32(v2.w);
    w[12] = bswap32(v3.x); w[13] = bswap32(v3.y); w[14] = bswap32(v3.z); w[15] = bswap32(v3.w);
}

// Fixed-size path: every message is nb*64 bytes at stride nb*64, arena 16-byte aligned.
// The padding block is the same for every message, so its K+W values arrive as a
// kernel parameter (constant-bank operands, no schedule work).  Next block's 64 bytes
// are prefetched into registers while the current block is compressed.
template <bool VERIFY, bool SC>
__global__ void __launch_bounds__(256)
k_sha256_fixed(const uint8_t *__restrict__ arena, uint32_t nb, uint64_t n, uint8_t *__restrict__ out,
               const __grid_constant__ KW64 kwpad, uint32_t rc, VerifyArgs va)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const U4 *p = reinterpret_cast<const U4 *>(arena + i * (uint64_t)nb * 64u);
    uint32_t st[8];
    sha256_init(st);
    U4 v0 = ldg_u4(p), v1 = ldg_u4(p + 1), v2 = ldg_u4(p + 2), v3 = ldg_u4(p + 3);
#pragma unroll 1
    for (uint32_t b = 0; b < nb; ++b) {
        uint32_t w[16];
        words_from_u4(w, v0, v1, v2, v3);
        if (b + 1 < nb) {
            p += 4;
            v0 = ldg_u4(p); v1 = ldg_u4(p + 1); v2 = ldg_u4(p + 2); v3 = ldg_u4(p + 3);
        }
        cp.block(st, w);
    }
    cp.pad(st, kwpad);
    if (out) store_digest(out + 32 * i, st);
    if (VERIFY) verify_epilogue(va, i, n, st);
}

// Fixed-size path, shared-memory staged: each warp copies its 32 messages' next 64-byte
// blocks with fully coalesced 16-byte cp.async (512 contiguous bytes per instruction
// when messages are 64 bytes; whole 128-byte lines otherwise), double buffered, then
// every lane reads its own block from a padded (conflict-free) slot.
template <bool VERIFY, bool SC>
__global__ void __launch_bounds__(256)
k_sha256_fixed_smem(const uint8_t *__restrict__ arena, uint32_t nb, uint64_t n, uint8_t *__restrict__ out,
                    const __grid_constant__ KW64 kwpad, uint32_t rc, VerifyArgs va)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    // per warp: 2 stages x 32 lanes x 80 bytes (64 data + 16 pad -> LDS.128 conflict-free)
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint8_t *wbase = smem + warp * (2u * 32u * 80u);
    const uint64_t warp_first = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane;
    if (warp_first >= n) return;
    const uint64_t i = warp_first + lane;
    const uint32_t nmsg = (uint32_t)min((uint64_t)32, n - warp_first);
    const uint64_t msg_bytes = (uint64_t)nb * 64u;
    const uint8_t *gwarp = arena + warp_first * msg_bytes;

    // stage s <- block b of all nmsg messages: 4 chunks of 16 B per message.
    // chunk id c in [0, nmsg*4): message m = c>>2, part k = c&3 ; lanes take c = lane + 32*r
    auto issue = [&](uint32_t b, uint32_t s) {
#pragma unroll
        for (uint32_t r = 0; r < 4; ++r) {
            uint32_t c = lane + 32u * r;
            uint32_t m = c >> 2, k = c & 3u;
            if (m < nmsg) {
                const uint8_t *src = gwarp + (uint64_t)m * msg_bytes + (uint64_t)b * 64u + k * 16u;
                uint32_t dst = (uint32_t)__cvta_generic_to_shared(wbase + s * (32u * 80u) + m * 80u + k * 16u);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src));
            }
        }
        asm volatile("cp.async.commit_group;\n" ::);
    };

    uint32_t st[8];
    sha256_init(st);
    issue(0, 0);
#pragma unroll 1
    for (uint32_t b = 0; b < nb; ++b) {
        if (b + 1 < nb) {
            issue(b + 1, (b + 1) & 1u);
            asm volatile("cp.async.wait_group 1;\n" ::);
        } else {
            asm volatile("cp.async.wait_group 0;\n" ::);
        }
        __syncwarp();
        const U4 *sp = reinterpret_cast<const U4 *>(wbase + (b & 1u) * (32u * 80u) + lane * 80u);
        U4 v0 = sp[0], v1 = sp[1], v2 = sp[2], v3 = sp[3];
        __syncwarp();     // all lanes have read stage b&1 before it is refilled two iterations later
        uint32_t w[16];
        words_from_u4(w, v0, v1, v2, v3);
        cp.block(st, w);
    }
    if (i < n) {
        cp.pad(st, kwpad);
        if (out) store_digest(out + 32 * i, st);
        if (VERIFY) verify_epilogue(va, i, n, st);
    }
}

// Fixed-size path, TMA-staged (option fixed_variant=2): every lane issues ONE bulk asynchronous copy
// (cp.async.bulk, the 1-D TMA path: UBLKCP in SASS) of its whole message into a padded shared-memory
// slot; the warp's 32 copies complete on one mbarrier (complete_tx byte counting), then each lane
// reads its message with conflict-free LDS.128 (slot stride = message + 16 bytes).  No load
// instructions in the hash loop at all.  Measured equal to the direct-load kernel within noise:
// the kernel is bound by the ALU pipe, not by how the bytes arrive.
template <bool VERIFY, bool SC>
__global__ void __launch_bounds__(128)
k_sha256_fixed_tma(const uint8_t *__restrict__ arena, uint32_t nb, uint64_t n, uint8_t *__restrict__ out,
                   const __grid_constant__ KW64 kwpad, uint32_t rc, VerifyArgs va)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    extern __shared__ __align__(128) uint8_t smem[];
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t msg_bytes = nb * 64u, slot = msg_bytes + 16u;
    const uint64_t warp_first = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane;
    if (warp_first >= n) return;
    const uint64_t i = warp_first + lane;
    const uint32_t nmsg = (uint32_t)min((uint64_t)32, n - warp_first);
    uint8_t *wbase = smem + 128 + (size_t)warp * 32u * slot;
    const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(smem + warp * 16u);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared.b64 [%0], %1;" ::"r"(mbar), "r"(1u) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (lane == 0)
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(nmsg * msg_bytes) : "memory");
    __syncwarp();
    if (lane < nmsg) {
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(wbase + lane * slot);
        const uint8_t *src = arena + i * (uint64_t)msg_bytes;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(msg_bytes), "r"(mbar) : "memory");
    }
    {   // all lanes wait for phase 0 of the warp's barrier
        uint32_t done = 0;
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                         "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(mbar), "r"(0u) : "memory");
        }
    }
    if (i >= n) return;
    uint32_t st[8];
    sha256_init(st);
    const U4 *sp = reinterpret_cast<const U4 *>(wbase + lane * slot);
#pragma unroll 1
    for (uint32_t b = 0; b < nb; ++b) {
        uint32_t w[16];
        words_from_u4(w, sp[4 * b], sp[4 * b + 1], sp[4 * b + 2], sp[4 * b + 3]);
        cp.block(st, w);
    }
    cp.pad(st, kwpad);
    if (out) store_digest(out + 32 * i, st);
    if (VERIFY) verify_epilogue(va, i, n, st);
}

// ------------------------------------------------------------------ length sort
__device__ __forceinline__ uint32_t bin_of(uint64_t len)
{
    uint64_t nb = (len + 9u + 63u) >> 6;
    return (uint32_t)(nb < (uint64_t)NBINS ? nb : (uint64_t)(NBINS - 1));
}

__global__ void k_len_hist(const uint64_t *__restrict__ offsets, uint64_t n, uint32_t *__restrict__ hist)
{
    __shared__ uint32_t sh[NBINS];
    for (int k = threadIdx.x; k < NBINS; k += blockDim.x) sh[k] = 0;
    __syncthreads();
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x)
        atomicAdd(&sh[bin_of(offsets[i + 1] - offsets[i])], 1u);
    __syncthreads();
    for (int k = threadIdx.x; k < NBINS; k += blockDim.x)
        if (sh[k]) atomicAdd(&hist[k], sh[k]);
}

// cursor[b] = number of messages in bins > b  (longest first).  One CTA, one thread per bin: thread t takes bin
// NBINS-1-t, so the exclusive prefix sum over t is the count in the longer bins.  (One thread walking the 1024
// bins took 25 us - a dependent global load per bin - which showed in every chunk of a replay.)
__global__ void __launch_bounds__(NBINS)
k_bin_scan(const uint32_t *__restrict__ hist, uint32_t *__restrict__ cursor)
{
    static_assert(NBINS == 1024, "one warp of warp totals");
    __shared__ uint32_t warp_sum[32];
    const uint32_t t = threadIdx.x, lane = t & 31u, wid = t >> 5;
    const uint32_t b = (uint32_t)NBINS - 1u - t;
    const uint32_t v = hist[b];
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= (uint32_t)o) x += y;
    }
    if (lane == 31) warp_sum[wid] = x;
    __syncthreads();
    if (wid == 0) {
        uint32_t ws = warp_sum[lane];
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, ws, o);
            if (lane >= (uint32_t)o) ws += y;
        }
        warp_sum[lane] = ws;
    }
    __syncthreads();
    cursor[b] = (wid ? warp_sum[wid - 1] : 0u) + x - v;
}

__global__ void k_len_scatter(const uint64_t *__restrict__ offsets, uint64_t n, uint32_t *__restrict__ cursor,
                              uint32_t *__restrict__ perm)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = i < n;
    uint32_t bin = valid ? bin_of(offsets[i + 1] - offsets[i]) : 0xffffffffu;
    // warp-aggregated atomics: one atomicAdd per distinct bin per warp
    uint32_t peers = __match_any_sync(0xffffffffu, bin);
    uint32_t leader = __ffs(peers) - 1;
    uint32_t lane = threadIdx.x & 31u;
    uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    uint32_t base = 0;
    if (valid && lane == leader) base = atomicAdd(&cursor[bin], (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (valid) perm[base + rank] = (uint32_t)i;
}

// ------------------------------------------------------------------ session helpers
__global__ void k_mark_first(const uint64_t *__restrict__ rowptr, uint64_t S, uint8_t *__restrict__ first)
{
    uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < S && rowptr[s + 1] > rowptr[s]) first[rowptr[s]] = 1;
}

__global__ void k_fill_u64(unsigned long long *p, uint64_t n, unsigned long long v)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

__global__ void k_max_session(const uint64_t *__restrict__ rowptr, uint64_t S, unsigned long long *out)
{
    uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long v = (s < S) ? (rowptr[s + 1] - rowptr[s]) : 0;
    for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0 && v) atomicMax(out, v);
}

// Many sessions: the per-level node counts are written by one thread per (session, level) and turned into row
// pointers by the tiled scan (scan.cuh); k_level_rowptrs below - one CTA per level - is for short session lists
// (10^6 sessions: 1.04 ms there).
__global__ void __launch_bounds__(256)
k_level_counts(const uint64_t *__restrict__ rowptr, uint64_t S, uint64_t *__restrict__ lvl_rp /* L*(S+1) */)
{
    const uint32_t l = blockIdx.y + 1;
    uint64_t *rp = lvl_rp + (uint64_t)blockIdx.y * (S + 1);
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const uint64_t nn = rowptr[s + 1] - rowptr[s];
    rp[s + 1] = (nn + ((1ull << l) - 1ull)) >> l;
    if (s == 0) rp[0] = 0;
}

// per-level CSR: lvl_rp[l-1][s] = sum_{s'<s} ceil(n_s' / 2^l) for l = 1..L ; one block per level
__global__ void __launch_bounds__(1024)
k_level_rowptrs(const uint64_t *__restrict__ rowptr, uint64_t S, uint64_t *__restrict__ lvl_rp /* L*(S+1) */)
{
    const uint32_t l = blockIdx.x + 1;
    uint64_t *rp = lvl_rp + (uint64_t)blockIdx.x * (S + 1);
    __shared__ uint64_t warp_sum[32];
    __shared__ uint64_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    for (uint64_t base = 0; base < S; base += blockDim.x) {
        uint64_t s = base + threadIdx.x;
        uint64_t c = 0;
        if (s < S) {
            uint64_t nn = rowptr[s + 1] - rowptr[s];
            c = (nn + ((1ull << l) - 1ull)) >> l;
        }
        uint64_t x = c;                                  // inclusive warp scan
        for (int o = 1; o < 32; o <<= 1) {
            uint64_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= (uint32_t)o) x += y;
        }
        if (lane == 31) warp_sum[wid] = x;
        __syncthreads();
        if (wid == 0) {
            uint64_t ws = (lane < (blockDim.x >> 5)) ? warp_sum[lane] : 0;
            for (int o = 1; o < 32; o <<= 1) {
                uint64_t y = __shfl_up_sync(0xffffffffu, ws, o);
                if (lane >= (uint32_t)o) ws += y;
            }
            warp_sum[lane] = ws;                         // inclusive over warps
        }
        __syncthreads();
        uint64_t before = carry + (wid ? warp_sum[wid - 1] : 0) + (x - c);
        if (s < S) rp[s] = before;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = before + c;
        __syncthreads();
    }
    if (threadIdx.x == 0) rp[S] = carry;
}

// ------------------------------------------------------------------ Merkle level step
// One thread per output node (audit/delta.py:127-131).  Sessions that are already down
// to one node carry it forward unchanged (the reference loop stops at len == 1, :125).
template <bool UNIFORM, bool SC>
__global__ void __launch_bounds__(128, 8)
k_merkle_level(const uint8_t *__restrict__ in, uint8_t *__restrict__ out, uint64_t S,
               uint32_t c_in_u, uint32_t c_out_u,                       // uniform counts per session
               const uint64_t *__restrict__ rp_in, const uint64_t *__restrict__ rp_out, uint32_t rc)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t in0, k, c_in;
    if (UNIFORM) {
        if (j >= S * c_out_u) return;
        uint32_t s = (uint32_t)j / c_out_u;
        k = (uint32_t)j - s * c_out_u;
        in0 = (uint64_t)s * c_in_u;
        c_in = c_in_u;
    } else {
        if (j >= rp_out[S]) return;
        uint64_t lo = 0, hi = S;                         // rp_out[lo] <= j < rp_out[hi]
        while (hi - lo > 1) {
            uint64_t mid = (lo + hi) >> 1;
            if (rp_out[mid] <= j) lo = mid; else hi = mid;
        }
        k = j - rp_out[lo];
        in0 = rp_in[lo];
        c_in = rp_in[lo + 1] - in0;
    }
    const uint8_t *lp = in + 32 * (in0 + 2 * k);
    if (c_in == 1) {
        const uint4 *src = reinterpret_cast<const uint4 *>(lp);
        uint4 *dst = reinterpret_cast<uint4 *>(out + 32 * j);
        dst[0] = src[0];
        dst[1] = src[1];
        return;
    }
    const uint8_t *rp = (2 * k + 1 < c_in) ? lp + 32 : lp;    // odd level: duplicate the last node (:129)
    uint32_t d[8];
    merkle_node_mem(lp, rp, d, cp);
    store_digest(out + 32 * j, d);
}

// Finish a session's tree inside one CTA once a level has <= TOP_MAX nodes: the nodes are
// pair-hashed level by level in shared memory.  Cuts the launch count for deep trees and
// is the whole tree for small sessions.
constexpr uint32_t TOP_MAX = 256;
template <bool UNIFORM, bool SC>
__global__ void __launch_bounds__(128)
k_merkle_top(const uint8_t *__restrict__ in, uint64_t S, uint32_t c_in_u, const uint64_t *__restrict__ rp_in,
             uint8_t *__restrict__ roots, uint8_t *__restrict__ empty, uint32_t rc, uint32_t *__restrict__ status)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    __shared__ uint32_t buf[2][TOP_MAX][8 + 1];   // +1 word pad against bank conflicts
    const uint64_t s = blockIdx.x;
    if (s >= S) return;
    uint64_t in0;
    uint32_t c;
    if (UNIFORM) {
        in0 = s * c_in_u;
        c = c_in_u;
    } else {
        in0 = rp_in[s];
        c = (uint32_t)(rp_in[s + 1] - in0);
    }
    if (c == 0 || c > TOP_MAX) {
        // c > TOP_MAX: the caller of a _dev entry understated max_session_leaves, so fewer level steps ran than this
        // session needs; it gets a zero root and the context's sticky error flag is raised (reported by ahv_sync)
        if (threadIdx.x < 32) roots[32 * s + threadIdx.x] = 0;
        if (threadIdx.x == 0 && empty) empty[s] = (c == 0) ? 1 : 0;
        if (threadIdx.x == 0 && c) atomicOr(status, AHV_STATUS_MAX_LEAVES);
        return;
    }
    if (threadIdx.x == 0 && empty) empty[s] = 0;
    if (c == 1) {
        if (threadIdx.x < 8)
            reinterpret_cast<uint32_t *>(roots + 32 * s)[threadIdx.x] =
                reinterpret_cast<const uint32_t *>(in + 32 * in0)[threadIdx.x];
        return;
    }
    // stage the level in shared memory, then pair-hash level by level (one copy of the node code)
    uint32_t cur = 0;
    for (uint32_t k = threadIdx.x; k < c; k += blockDim.x) {
        uint32_t d[8];
        load_digest(in + 32 * (in0 + k), d);
#pragma unroll
        for (int q = 0; q < 8; ++q) buf[cur][k][q] = d[q];
    }
    __syncthreads();
    while (c > 1) {
        const uint32_t c_out = (c + 1) >> 1;
        for (uint32_t k = threadIdx.x; k < c_out; k += blockDim.x) {
            uint32_t l[8], r[8], d[8];
            const uint32_t kr = (2 * k + 1 < c) ? 2 * k + 1 : 2 * k;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                l[q] = buf[cur][2 * k][q];
                r[q] = buf[cur][kr][q];
            }
            merkle_node(l, r, d, cp);
#pragma unroll
            for (int q = 0; q < 8; ++q) buf[cur ^ 1][k][q] = d[q];
        }
        cur ^= 1;
        c = c_out;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        uint32_t d[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) d[q] = buf[cur][0][q];
        store_digest(roots + 32 * s, d);
    }
}

// Same finish, one THREAD per session, for levels of <= 2 nodes: what is left when there are so many sessions
// that the global level steps run to the end (tree_top = 2).  A warp per session would hash one node with one
// lane (10^6 sessions: 5.5 ms; 0.2 ms here).
template <bool UNIFORM, bool SC>
__global__ void __launch_bounds__(128)
k_merkle_top_lane(const uint8_t *__restrict__ in, uint64_t S, uint32_t c_in_u, const uint64_t *__restrict__ rp_in,
                  uint8_t *__restrict__ roots, uint8_t *__restrict__ empty, uint32_t rc, uint32_t *__restrict__ status)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    uint64_t in0;
    uint32_t c;
    if (UNIFORM) {
        in0 = s * c_in_u;
        c = c_in_u;
    } else {
        in0 = rp_in[s];
        c = (uint32_t)(rp_in[s + 1] - in0);
    }
    uint32_t d[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (c == 0 || c > 2u) {                               // c > 2: see k_merkle_top
        store_digest(roots + 32 * s, d);
        if (empty) empty[s] = (c == 0) ? 1 : 0;
        if (c) atomicOr(status, AHV_STATUS_MAX_LEAVES);
        return;
    }
    if (empty) empty[s] = 0;
    uint32_t l[8];
    load_digest(in + 32 * in0, l);
    if (c == 1) {
        store_digest(roots + 32 * s, l);
        return;
    }
    uint32_t r[8];
    load_digest(in + 32 * (in0 + 1), r);
    merkle_node(l, r, d, cp);
    store_digest(roots + 32 * s, d);
}

// Same finish, one WARP per session (4 sessions per CTA) for levels of <= 64 nodes: no idle
// warps, no block-wide barriers; the level is pair-hashed in the warp's shared-memory slot.
template <bool UNIFORM, bool SC>
__global__ void __launch_bounds__(128)
k_merkle_top_warp(const uint8_t *__restrict__ in, uint64_t S, uint32_t c_in_u, const uint64_t *__restrict__ rp_in,
                  uint8_t *__restrict__ roots, uint8_t *__restrict__ empty, uint32_t rc, uint32_t *__restrict__ status)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    __shared__ uint32_t buf[4][2][64][8 + 1];
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint64_t s = (uint64_t)blockIdx.x * 4 + wid;
    if (s >= S) return;
    uint64_t in0;
    uint32_t c;
    if (UNIFORM) {
        in0 = s * c_in_u;
        c = c_in_u;
    } else {
        in0 = rp_in[s];
        c = (uint32_t)(rp_in[s + 1] - in0);
    }
    if (c == 0 || c > 64u) {                              // c > 64: see k_merkle_top
        roots[32 * s + lane] = 0;
        if (lane == 0 && empty) empty[s] = (c == 0) ? 1 : 0;
        if (lane == 0 && c) atomicOr(status, AHV_STATUS_MAX_LEAVES);
        return;
    }
    if (lane == 0 && empty) empty[s] = 0;
    if (c == 1) {
        if (lane < 8)
            reinterpret_cast<uint32_t *>(roots + 32 * s)[lane] = reinterpret_cast<const uint32_t *>(in + 32 * in0)[lane];
        return;
    }
    uint32_t cur = 0;
    for (uint32_t k = lane; k < c; k += 32) {            // c <= 64
        uint32_t d[8];
        load_digest(in + 32 * (in0 + k), d);
#pragma unroll
        for (int q = 0; q < 8; ++q) buf[wid][cur][k][q] = d[q];
    }
    __syncwarp();
    while (c > 1) {
        const uint32_t c_out = (c + 1) >> 1;             // <= 32: one node per lane
        if (lane < c_out) {
            uint32_t l[8], r[8], d[8];
            const uint32_t kr = (2 * lane + 1 < c) ? 2 * lane + 1 : 2 * lane;
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                l[q] = buf[wid][cur][2 * lane][q];
                r[q] = buf[wid][cur][kr][q];
            }
            merkle_node(l, r, d, cp);
#pragma unroll
            for (int q = 0; q < 8; ++q) buf[wid][cur ^ 1][lane][q] = d[q];
        }
        cur ^= 1;
        c = c_out;
        __syncwarp();
    }
    if (lane == 0) {
        uint32_t d[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) d[q] = buf[wid][cur][0][q];
        store_digest(roots + 32 * s, d);
    }
}

// ------------------------------------------------------------------ verify finalisation
__global__ void k_verify_finalize(const unsigned long long *__restrict__ fb, uint64_t S,
                                  uint8_t *__restrict__ verdict, int64_t *__restrict__ first_bad)
{
    uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    unsigned long long v = fb[s];
    if (verdict) verdict[s] = (v == FB_NONE) ? 1 : 0;
    if (first_bad) first_bad[s] = (v == FB_NONE) ? -1 : (int64_t)v;
}

// audit/commitment.py:59-64 on raw digests; an empty session has no root -> no match
__global__ void k_root_compare(const uint8_t *__restrict__ roots, const uint8_t *__restrict__ expected,
                               const uint8_t *__restrict__ empty, uint64_t S, uint8_t *__restrict__ match)
{
    uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const uint4 *a = reinterpret_cast<const uint4 *>(roots + 32 * s);
    const uint4 *b = reinterpret_cast<const uint4 *>(expected + 32 * s);
    uint4 a0 = a[0], a1 = a[1], b0 = b[0], b1 = b[1];
    uint32_t diff = (a0.x ^ b0.x) | (a0.y ^ b0.y) | (a0.z ^ b0.z) | (a0.w ^ b0.w) |
                    (a1.x ^ b1.x) | (a1.y ^ b1.y) | (a1.z ^ b1.z) | (a1.w ^ b1.w);
    match[s] = (diff == 0 && !(empty && empty[s])) ? 1 : 0;
}

// ------------------------------------------------------------------ parents from the pre-images
// The claimed parent of a delta is already inside its canonical pre-image ("parent_hash": "<64 hex>"
// or null, audit/delta.py:61).  Given the byte position of that value, this fills the parent side
// arrays on the device, so a host batch needs 4 bytes per delta instead of 33.
__global__ void __launch_bounds__(256)
k_extract_parents(const uint8_t *__restrict__ arena, const uint64_t *__restrict__ offsets, uint64_t fixed_len,
                  const uint32_t *__restrict__ pos, uint64_t n, uint8_t *__restrict__ parents,
                  uint8_t *__restrict__ kind)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t off, len;
    if (offsets) {
        off = offsets[i];
        len = offsets[i + 1] - off;
    } else {
        off = i * fixed_len;
        len = fixed_len;
    }
    const uint32_t p = pos[i];
    uint32_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint8_t k = AHV_PARENT_NONE;
    if (p != 0xffffffffu) {
        k = AHV_PARENT_OTHER;
        // the value must be exactly "<64 lowercase hex>": quotes on both sides, 64 valid characters
        if (p >= 1 && (uint64_t)p + 65 <= len) {
            const uint8_t *b = arena + off + p;
            bool ok = (b[-1] == '"') && (b[64] == '"');
#pragma unroll 4
            for (int q = 0; q < 32 && ok; ++q) {
                const uint32_t hi = b[2 * q], lo = b[2 * q + 1];
                const uint32_t vh = (hi - '0' < 10u) ? hi - '0' : ((hi - 'a' < 6u) ? hi - 'a' + 10 : 0xffu);
                const uint32_t vl = (lo - '0' < 10u) ? lo - '0' : ((lo - 'a' < 6u) ? lo - 'a' + 10 : 0xffu);
                ok = (vh | vl) < 16u;
                w[q >> 2] |= ((vh << 4) | (vl & 15u)) << (8 * (q & 3));     // memory byte order
            }
            if (ok) k = AHV_PARENT_HEX;
        }
        if (k != AHV_PARENT_HEX) {
#pragma unroll
            for (int q = 0; q < 8; ++q) w[q] = 0;
        }
    }
    uint4 *dst = reinterpret_cast<uint4 *>(parents + 32 * i);
    dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
    dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
    kind[i] = k;
}

// ------------------------------------------------------------------ one message, lowest latency
// A single capture (audit/delta.py:85-115 -> :63) is one short message: two copies and a kernel through the
// batch entry point cost ~30 us of which the hashing is 2.  Here the message travels INSIDE the launch (a
// __grid_constant__ kernel parameter, up to AHV_ONE_MAX bytes) and the digest is written straight into
// mapped pinned host memory, so the whole call is one launch and one stream synchronisation.
struct OneMsg {
    uint32_t len;
    uint32_t pad[3];
    uint8_t data[AHV_ONE_MAX + 16];               // slack: aligned word reads may touch the word after the last byte
};

template <bool SC>
__global__ void __launch_bounds__(32)
k_sha256_one(const __grid_constant__ OneMsg m, uint8_t *out_mapped, uint32_t rc)
{
    if (threadIdx.x != 0) return;
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    uint32_t st[8];
    sha256_message(m.data, m.len, st, cp);
    store_digest(out_mapped, st);
    __threadfence_system();
}

// ------------------------------------------------------------------ chained capture
// One thread per session walks its chain (audit/delta.py:101-114): hash delta i, write the
// 64 hex characters into delta i+1's pre-image, continue.  Plain (coherent) loads only:
// the thread re-reads bytes it has just written.
// Blocks of a delta that lie wholly in front of its parent field do not depend on the chain.
__device__ __forceinline__ uint64_t chain_prefix_blocks(uint32_t hp, uint64_t len)
{
    if (hp == 0xffffffffu) return 0;
    const uint64_t a = hp >> 6, b = len >> 6;
    return a < b ? a : b;
}

// Pre-pass of the chained capture, one thread per delta: the SHA-256 state after those blocks, left in the
// delta's slot of out_digests (raw state words; k_chain_capture reads it before it stores the digest there).
template <bool SC>
__global__ void __launch_bounds__(128)
k_chain_prefix(const uint8_t *__restrict__ arena, const uint64_t *__restrict__ offsets, const uint32_t *__restrict__ hexpos,
               uint64_t n, uint8_t *__restrict__ out_digests, uint32_t rc)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t off = offsets[i];
    const uint64_t nb = chain_prefix_blocks(hexpos[i], offsets[i + 1] - off);
    if (nb == 0) return;
    uint32_t st[8];
    sha256_prefix(arena + off, nb, st, cp);
    U4 *o = reinterpret_cast<U4 *>(out_digests + 32 * i);
    o[0] = U4{st[0], st[1], st[2], st[3]};
    o[1] = U4{st[4], st[5], st[6], st[7]};
}

template <bool SC, bool PREFIX>
__global__ void __launch_bounds__(128)
k_chain_capture(uint8_t *arena, const uint64_t *__restrict__ offsets, const uint32_t *__restrict__ hexpos,
                const uint64_t *__restrict__ rowptr, uint64_t S, uint8_t *out_digests,
                uint8_t *__restrict__ out_parents, uint8_t *__restrict__ out_kind, uint32_t rc, uint32_t lanes_per_warp)
{
    const typename Pol<SC>::type cp = Pol<SC>::make(rc);
    // lanes_per_warp < 32 spreads the sessions over more warps (only the first lanes of each carry one): an
    // experiment, see ahv_chain_capture_dev
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = (uint32_t)(t & 31u);
    if (lane >= lanes_per_warp) return;
    const uint64_t s = (t >> 5) * lanes_per_warp + lane;
    if (s >= S) return;
    uint32_t prev[8];
    const uint64_t i_begin = rowptr[s], i_end = rowptr[s + 1];
    if (i_begin >= i_end) return;
    // the next delta's extent and patch position are fetched while the current one is hashed: they do not depend
    // on the chain, only the bytes do
    uint64_t off = offsets[i_begin], off_next = offsets[i_begin + 1];
    uint32_t hp = hexpos[i_begin];
    U4 mid0 = {0, 0, 0, 0}, mid1 = {0, 0, 0, 0};       // PREFIX: the state k_chain_prefix left for delta i
    if (PREFIX && chain_prefix_blocks(hp, off_next - off)) {
        const U4 *m = reinterpret_cast<const U4 *>(out_digests + 32 * i_begin);
        mid0 = m[0];
        mid1 = m[1];
    }
    for (uint64_t i = i_begin; i < i_end; ++i) {
        const uint64_t len = off_next - off;
        uint64_t off_after = 0;
        uint32_t hp_next = 0xffffffffu;
        U4 nmid0 = {0, 0, 0, 0}, nmid1 = {0, 0, 0, 0};
        if (i + 1 < i_end) {
            off_after = __ldg(offsets + i + 2);
            hp_next = __ldg(hexpos + i + 1);
            if (PREFIX && chain_prefix_blocks(hp_next, off_after - off_next)) {   // chain-independent as well: fetched a delta ahead
                const U4 *m = reinterpret_cast<const U4 *>(out_digests + 32 * (i + 1));
                nmid0 = m[0];
                nmid1 = m[1];
            }
        }
        if (i > i_begin && hp != 0xffffffffu) {
            uint8_t *dst = arena + off + hp;
            if ((reinterpret_cast<uintptr_t>(dst) & 3u) == 0) {            // 16 word stores instead of 64 byte stores
                uint32_t *d4 = reinterpret_cast<uint32_t *>(dst);
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    uint32_t w0, w1;
                    hex_expand(prev[q], w0, w1);
                    d4[2 * q] = bswap32(w0);                               // first character at the lowest address
                    d4[2 * q + 1] = bswap32(w1);
                }
            } else {
                // the field starts r = 1..3 bytes into a word (the usual case: its position moves with the digits of
                // delta_id and turn_id): 15 whole words, each made of two neighbouring hex words, and the 4 bytes at
                // the two ragged ends - not 64 byte stores, which kept the load/store unit busy for 3.5 us per delta
                const uint32_t r = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 3u);
                uint32_t *d4 = reinterpret_cast<uint32_t *>(dst - r);
                uint32_t m[16];                                            // the field as little-endian words
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    uint32_t w0, w1;
                    hex_expand(prev[q], w0, w1);
                    m[2 * q] = bswap32(w0);
                    m[2 * q + 1] = bswap32(w1);
                }
#pragma unroll
                for (int k = 1; k < 16; ++k) d4[k] = __funnelshift_l(m[k - 1], m[k], 8u * r);
#pragma unroll
                for (uint32_t t = 0; t < 3; ++t) {
                    if (t < 4u - r) dst[t] = (uint8_t)(m[0] >> (8u * t));
                    if (t < r) dst[64u - r + t] = (uint8_t)(m[15] >> (8u * (4u - r + t)));
                }
            }
            if (out_parents) store_digest(out_parents + 32 * i, prev);
            if (out_kind) out_kind[i] = AHV_PARENT_HEX;
        } else {
            if (out_parents) {
                uint32_t z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                store_digest(out_parents + 32 * i, z);
            }
            if (out_kind) out_kind[i] = AHV_PARENT_NONE;
        }
        const uint64_t nb = PREFIX ? chain_prefix_blocks(hp, len) : 0;
        if (nb) {
            prev[0] = mid0.x; prev[1] = mid0.y; prev[2] = mid0.z; prev[3] = mid0.w;
            prev[4] = mid1.x; prev[5] = mid1.y; prev[6] = mid1.z; prev[7] = mid1.w;
            sha256_message_from(arena + off, len, prev, nb, cp);
        } else {
            sha256_message(arena + off, len, prev, cp);
        }
        if (out_digests) store_digest(out_digests + 32 * i, prev);
        off = off_next;
        off_next = off_after;
        hp = hp_next;
        mid0 = nmid0;
        mid1 = nmid1;
    }
}

// ------------------------------------------------------------------ host-side drivers
int check_sizes(uint64_t n, uint64_t n_sessions)
{
    if (n >= 0xffffffffull || n_sessions >= 0xffffffffull) {
        set_error("batch too large: n=%llu sessions=%llu (limit 2^32-2 per call)", (unsigned long long)n,
                  (unsigned long long)n_sessions);
        return AHV_E_TOOLARGE;
    }
    return AHV_OK;
}

// scratch bytes the leaf stage needs
uint64_t leaf_scratch_bytes(uint64_t n, bool sorted)
{
    return sorted ? pad256(4ull * n) + 2 * pad256(4ull * NBINS) : 0;
}

// Enqueue leaf hashing (optionally with the verify epilogue).  Scratch must be reserved.
int run_leaves(ahv_ctx *ctx, const uint8_t *d_arena, const uint64_t *d_offsets, uint64_t fixed_len, uint64_t n,
               uint8_t *d_out, const VerifyArgs *va)
{
    if (n == 0) return AHV_OK;
    const unsigned block = (unsigned)ctx->opt_block;
    VerifyArgs dummy = {};
    const VerifyArgs &v = va ? *va : dummy;
    const bool fixed_fast = (!d_offsets && fixed_len > 0 && (fixed_len & 63u) == 0 &&
                             (reinterpret_cast<uintptr_t>(d_arena) & 15u) == 0 && fixed_len / 64 < 0xffffffffull);
    const int tslot = ctx->opt_leaf_timer ? (int)(ctx->opt_leaf_timer + (ctx->leaf_launches++ & 15u)) : -1;
    const bool sc = ctx->opt_compress != 0;
    const uint32_t rc = SHA_SC;
    if (fixed_fast) {
        const KW64 kw = sc ? make_kw_pad_scaled(fixed_len * 8) : make_kw_pad(fixed_len * 8);
        const uint32_t nb = (uint32_t)(fixed_len / 64);
        if (tslot >= 0) AHV_CUDA(cudaEventRecord(ctx->t_start[tslot], ctx->stream));
        if (ctx->opt_fixed_variant == 1) {
            size_t smem = (size_t)(block / 32) * 2 * 32 * 80;
#define AHV_LAUNCH_LEAF(KERNEL, SMEM, ...)                                                                  \
    do {                                                                                                   \
        if (va && sc) KERNEL<true, true><<<grid_for(n, block), block, SMEM, ctx->stream>>>(__VA_ARGS__);    \
        else if (va) KERNEL<true, false><<<grid_for(n, block), block, SMEM, ctx->stream>>>(__VA_ARGS__);    \
        else if (sc) KERNEL<false, true><<<grid_for(n, block), block, SMEM, ctx->stream>>>(__VA_ARGS__);    \
        else KERNEL<false, false><<<grid_for(n, block), block, SMEM, ctx->stream>>>(__VA_ARGS__);           \
    } while (0)
            AHV_LAUNCH_LEAF(k_sha256_fixed_smem, smem, d_arena, nb, n, d_out, kw, rc, v);
        } else if (ctx->opt_fixed_variant == 2 && fixed_len <= 320) {
            const unsigned block = 128;                       // shadows: this kernel is built for 4 warps
            size_t smem = 128 + (size_t)4 * 32 * (fixed_len + 16);
            AHV_LAUNCH_LEAF(k_sha256_fixed_tma, smem, d_arena, nb, n, d_out, kw, rc, v);
        } else {
            AHV_LAUNCH_LEAF(k_sha256_fixed, 0, d_arena, nb, n, d_out, kw, rc, v);
        }
        AHV_CHECK_LAUNCH(ctx);
        if (tslot >= 0) AHV_CUDA(cudaEventRecord(ctx->t_stop[tslot], ctx->stream));
        return AHV_OK;
    }
    const uint32_t *perm = nullptr;
    if (d_offsets && ctx->opt_sort && n > 1024) {
        uint32_t *p = (uint32_t *)scratch_take(ctx, 4ull * n);
        uint32_t *hist = (uint32_t *)scratch_take(ctx, 4ull * NBINS);
        uint32_t *cursor = (uint32_t *)scratch_take(ctx, 4ull * NBINS);
        if (!p || !hist || !cursor) {
            set_error("internal: scratch too small for length sort");
            return AHV_E_NOMEM;
        }
        AHV_CUDA(cudaMemsetAsync(hist, 0, 4ull * NBINS, ctx->stream));
        uint64_t gmax = (uint64_t)ctx->sm_count * 8, gfull = grid_for(n, 256);
        unsigned g = (unsigned)(gfull < gmax ? gfull : gmax);
        k_len_hist<<<g, 256, 0, ctx->stream>>>(d_offsets, n, hist);
        AHV_CHECK_LAUNCH(ctx);
        k_bin_scan<<<1, NBINS, 0, ctx->stream>>>(hist, cursor);
        AHV_CHECK_LAUNCH(ctx);
        k_len_scatter<<<grid_for(n, 256), 256, 0, ctx->stream>>>(d_offsets, n, cursor, p);
        AHV_CHECK_LAUNCH(ctx);
        perm = p;
    }
    if (tslot >= 0) AHV_CUDA(cudaEventRecord(ctx->t_start[tslot], ctx->stream));
    AHV_LAUNCH_LEAF(k_sha256_var, 0, d_arena, d_offsets, fixed_len, n, perm, d_out, rc, v);
    AHV_CHECK_LAUNCH(ctx);
    if (tslot >= 0) AHV_CUDA(cudaEventRecord(ctx->t_stop[tslot], ctx->stream));
    return AHV_OK;
}

// levels reduced by global level steps before the in-CTA top kernel takes over
// Level width at which a session's tree is finished inside one warp / CTA.  Dense global level
// steps run at full ALU rate as long as a level still has tens of thousands of nodes over all
// sessions, so with many sessions the hand-over comes late (2 nodes), with few it comes early (64).
uint64_t tree_top(const ahv_ctx *ctx, uint64_t S)
{
    if (ctx->opt_top_max >= 2 && ctx->opt_top_max <= (int64_t)TOP_MAX) return (uint64_t)ctx->opt_top_max;
    uint64_t t = 16384 / (S ? S : 1);
    return t < 2 ? 2 : (t > 64 ? 64 : t);
}

uint32_t global_levels(const ahv_ctx *ctx, uint64_t S, uint64_t max_leaves)
{
    const uint64_t top = tree_top(ctx, S);
    uint32_t l = 0;
    while (((max_leaves + ((1ull << l) - 1)) >> l) > top) ++l;
    return l;
}

uint64_t tree_scratch_bytes(const ahv_ctx *ctx, uint64_t n, uint64_t S, uint64_t max_leaves, bool uniform)
{
    const uint32_t G = global_levels(ctx, S, max_leaves);
    if (G == 0) return 0;
    uint64_t b = 2 * pad256(32ull * (n / 2 + S + 1));
    if (!uniform) b += pad256(8ull * G * (S + 1)) + scan_scratch_bytes(S, G) + 256;
    return b;
}

// Enqueue the tree: digests (n leaves in S sessions) -> roots, empty flags.  Scratch reserved.
int run_tree(ahv_ctx *ctx, const uint8_t *d_digests, const uint64_t *d_rowptr, uint64_t S, uint64_t n,
             uint64_t uniform, uint64_t max_leaves, uint8_t *d_roots, uint8_t *d_empty)
{
    if (S == 0) return AHV_OK;
    const uint32_t G = global_levels(ctx, S, max_leaves);
    const uint8_t *cur = d_digests;
    const uint64_t *rp_cur = d_rowptr;
    uint32_t c_cur = (uint32_t)uniform;
    const bool sc = ctx->opt_compress != 0;
    const uint32_t rc = SHA_SC;
    if (G > 0) {
        uint8_t *bufA = (uint8_t *)scratch_take(ctx, 32ull * (n / 2 + S + 1));
        uint8_t *bufB = (uint8_t *)scratch_take(ctx, 32ull * (n / 2 + S + 1));
        uint64_t *lvl_rp = nullptr;
        if (!uniform) lvl_rp = (uint64_t *)scratch_take(ctx, 8ull * G * (S + 1));
        if (!bufA || !bufB || (!uniform && !lvl_rp)) {
            set_error("internal: scratch too small for tree levels");
            return AHV_E_NOMEM;
        }
        if (!uniform && S > 8ull * SCAN_TILE) {
            k_level_counts<<<dim3(grid_for(S, 256), G), 256, 0, ctx->stream>>>(d_rowptr, S, lvl_rp);
            AHV_CHECK_LAUNCH(ctx);
            AHV_TRY(scan_u64_inplace(ctx, lvl_rp, S, G, S + 1, /*reserve=*/false));
        } else if (!uniform) {
            k_level_rowptrs<<<G, 1024, 0, ctx->stream>>>(d_rowptr, S, lvl_rp);
            AHV_CHECK_LAUNCH(ctx);
        }
        for (uint32_t l = 1; l <= G; ++l) {
            uint8_t *dst = (l & 1) ? bufA : bufB;
            if (uniform) {
                uint32_t c_out = (c_cur + 1) >> 1;
                uint64_t total = S * (uint64_t)c_out;
                if (sc) k_merkle_level<true, true><<<grid_for(total, 128), 128, 0, ctx->stream>>>(cur, dst, S, c_cur, c_out, nullptr, nullptr, rc);
                else k_merkle_level<true, false><<<grid_for(total, 128), 128, 0, ctx->stream>>>(cur, dst, S, c_cur, c_out, nullptr, nullptr, rc);
                c_cur = c_out;
            } else {
                const uint64_t *rp_out = lvl_rp + (uint64_t)(l - 1) * (S + 1);
                uint64_t bound = ((n + ((1ull << l) - 1)) >> l) + S;   // >= sum ceil(n_s/2^l)
                if (sc) k_merkle_level<false, true><<<grid_for(bound, 128), 128, 0, ctx->stream>>>(cur, dst, S, 0, 0, rp_cur, rp_out, rc);
                else k_merkle_level<false, false><<<grid_for(bound, 128), 128, 0, ctx->stream>>>(cur, dst, S, 0, 0, rp_cur, rp_out, rc);
                rp_cur = rp_out;
            }
            AHV_CHECK_LAUNCH(ctx);
            cur = dst;
        }
    }
    const uint64_t top = tree_top(ctx, S);
    if (top > 64 && S > 0x7fffffffull) {                  // one CTA per session: gridDim.x is limited to 2^31 - 1
        set_error("more than 2^31 - 1 sessions with an in-CTA tree finish (top_max > 64); use top_max <= 64 or split the batch");
        return AHV_E_TOOLARGE;
    }
    if (top <= 2) {
        const unsigned g = grid_for(S, 128);
        if (uniform && sc) k_merkle_top_lane<true, true><<<g, 128, 0, ctx->stream>>>(cur, S, c_cur, nullptr, d_roots, d_empty, rc, ctx->status_d);
        else if (uniform) k_merkle_top_lane<true, false><<<g, 128, 0, ctx->stream>>>(cur, S, c_cur, nullptr, d_roots, d_empty, rc, ctx->status_d);
        else if (sc) k_merkle_top_lane<false, true><<<g, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, rc, ctx->status_d);
        else k_merkle_top_lane<false, false><<<g, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, rc, ctx->status_d);
    } else if (top <= 64) {
        const unsigned g4 = (unsigned)((S + 3) / 4);
        if (uniform && sc) k_merkle_top_warp<true, true><<<g4, 128, 0, ctx->stream>>>(cur, S, c_cur, nullptr, d_roots, d_empty, rc, ctx->status_d);
        else if (uniform) k_merkle_top_warp<true, false><<<g4, 128, 0, ctx->stream>>>(cur, S, c_cur, nullptr, d_roots, d_empty, rc, ctx->status_d);
        else if (sc) k_merkle_top_warp<false, true><<<g4, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, rc, ctx->status_d);
        else k_merkle_top_warp<false, false><<<g4, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, rc, ctx->status_d);
    } else
    if (uniform && sc) k_merkle_top<true, true><<<(unsigned)S, 128, 0, ctx->stream>>>(cur, S, c_cur, nullptr, d_roots, d_empty, rc, ctx->status_d);
    else if (uniform) k_merkle_top<true, false><<<(unsigned)S, 128, 0, ctx->stream>>>(cur, S, c_cur, nullptr, d_roots, d_empty, rc, ctx->status_d);
    else if (sc) k_merkle_top<false, true><<<(unsigned)S, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, rc, ctx->status_d);
    else k_merkle_top<false, false><<<(unsigned)S, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, rc, ctx->status_d);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

// max session size from a device rowptr (synchronises)
int device_max_session(ahv_ctx *ctx, const uint64_t *d_rowptr, uint64_t S, uint64_t *out)
{
    unsigned long long *d = (unsigned long long *)scratch_take(ctx, 8);
    if (!d) {
        set_error("internal: scratch too small");
        return AHV_E_NOMEM;
    }
    AHV_CUDA(cudaMemsetAsync(d, 0, 8, ctx->stream));
    k_max_session<<<grid_for(S, 256), 256, 0, ctx->stream>>>(d_rowptr, S, d);
    AHV_CHECK_LAUNCH(ctx);
    unsigned long long h = 0;
    AHV_CUDA(cudaMemcpyAsync(&h, d, 8, cudaMemcpyDeviceToHost, ctx->stream));
    AHV_CUDA(cudaStreamSynchronize(ctx->stream));
    *out = h;
    return AHV_OK;
}

struct Shape {
    uint64_t n, S, uniform, max_leaves;
};

// validate the (offsets | fixed_len) x (rowptr | uniform) combination
int check_shape(const void *offsets, uint64_t fixed_len, const void *rowptr, uint64_t S, uint64_t n,
                uint64_t uniform)
{
    AHV_REQUIRE(offsets || fixed_len > 0 || n == 0, "offsets is NULL and fixed_len is 0");
    AHV_REQUIRE(rowptr || uniform > 0 || S == 0, "session_rowptr is NULL and uniform_leaves is 0");
    if (uniform) AHV_REQUIRE(S * uniform == n, "n_sessions * uniform_leaves != n_leaves");
    return check_sizes(n, S);
}

}  // namespace

// =================================================================== C ABI (device pointers)
extern "C" {

int ahv_sha256_batch_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                         uint64_t n, uint64_t fixed_len, uint8_t *d_out)
{
    AHV_TRY(use_device(ctx));
    (void)arena_bytes;
    AHV_REQUIRE(d_out || n == 0, "out_digests is null");
    AHV_REQUIRE(d_offsets || fixed_len > 0 || n == 0, "offsets is NULL and fixed_len is 0");
    AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_out) & 15u) == 0, "out_digests must be 16-byte aligned");
    AHV_TRY(check_sizes(n, 0));
    const bool sorted = d_offsets && ctx->opt_sort && n > 1024;
    AHV_TRY(scratch_reserve(ctx, leaf_scratch_bytes(n, sorted)));
    return run_leaves(ctx, d_arena, d_offsets, fixed_len, n, d_out, nullptr);
}

int ahv_merkle_roots_dev(ahv_ctx *ctx, const uint8_t *d_digests, const uint64_t *d_rowptr, uint64_t S,
                         uint64_t n, uint64_t max_leaves, uint8_t *d_roots, uint8_t *d_empty)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(d_roots || S == 0, "out_roots is null");
    AHV_REQUIRE(d_rowptr || S == 0, "session_rowptr is null");
    AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_digests) & 15u) == 0 && (reinterpret_cast<uintptr_t>(d_roots) & 15u) == 0,
                "digest buffers must be 16-byte aligned");
    AHV_TRY(check_sizes(n, S));
    if (S == 0) return AHV_OK;
    if (max_leaves == 0) {
        AHV_TRY(scratch_reserve(ctx, 256));
        AHV_TRY(device_max_session(ctx, d_rowptr, S, &max_leaves));
    }
    AHV_TRY(scratch_reserve(ctx, tree_scratch_bytes(ctx, n, S, max_leaves, false)));
    return run_tree(ctx, d_digests, d_rowptr, S, n, 0, max_leaves, d_roots, d_empty);
}

int ahv_commit_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                   uint64_t fixed_len, const uint64_t *d_rowptr, uint64_t S, uint64_t n, uint64_t uniform,
                   uint64_t max_leaves, uint8_t *d_digests, uint8_t *d_roots, uint8_t *d_empty)
{
    AHV_TRY(use_device(ctx));
    (void)arena_bytes;
    AHV_TRY(check_shape(d_offsets, fixed_len, d_rowptr, S, n, uniform));
    AHV_REQUIRE(d_roots || S == 0, "out_roots is null");
    if (uniform) max_leaves = uniform;
    if (max_leaves == 0 && S) {
        AHV_TRY(scratch_reserve(ctx, 256));
        AHV_TRY(device_max_session(ctx, d_rowptr, S, &max_leaves));
    }
    const bool sorted = d_offsets && ctx->opt_sort && n > 1024;
    uint64_t need = leaf_scratch_bytes(n, sorted) + tree_scratch_bytes(ctx, n, S, max_leaves, uniform != 0) +
                    (d_digests ? 0 : pad256(32ull * n));
    AHV_TRY(scratch_reserve(ctx, need));
    uint8_t *dig = d_digests;
    if (!dig && n) {
        dig = (uint8_t *)scratch_take(ctx, 32ull * n);
        if (!dig) {
            set_error("internal: scratch too small for digests");
            return AHV_E_NOMEM;
        }
    }
    AHV_TRY(run_leaves(ctx, d_arena, d_offsets, fixed_len, n, dig, nullptr));
    return run_tree(ctx, dig, d_rowptr, S, n, uniform, max_leaves, d_roots, d_empty);
}

int ahv_verify_chains_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                          uint64_t fixed_len, const uint8_t *d_parents, const uint8_t *d_kind,
                          const uint64_t *d_rowptr, uint64_t S, uint64_t n, uint64_t uniform, uint64_t max_leaves,
                          const uint8_t *d_expected, uint8_t *d_verdict, int64_t *d_first_bad, uint8_t *d_digests,
                          uint8_t *d_roots, uint8_t *d_match)
{
    AHV_TRY(use_device(ctx));
    (void)arena_bytes;
    AHV_TRY(check_shape(d_offsets, fixed_len, d_rowptr, S, n, uniform));
    AHV_REQUIRE((d_parents && d_kind) || n == 0, "parent_digests / parent_kind is null");
    AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_parents) & 15u) == 0, "parent_digests must be 16-byte aligned");
    AHV_REQUIRE(!d_match || d_expected, "out_root_match needs expected_roots");
    if (S == 0) return AHV_OK;
    const bool want_tree = d_expected || d_roots;
    if (uniform) max_leaves = uniform;
    if (want_tree && max_leaves == 0) {
        AHV_TRY(scratch_reserve(ctx, 256));
        AHV_TRY(device_max_session(ctx, d_rowptr, S, &max_leaves));
    }
    const bool sorted = d_offsets && ctx->opt_sort && n > 1024;
    const bool need_dig = want_tree && !d_digests;
    uint64_t need = leaf_scratch_bytes(n, sorted) + pad256(8ull * S) + (uniform ? 0 : pad256(n)) +
                    (want_tree ? tree_scratch_bytes(ctx, n, S, max_leaves, uniform != 0) + pad256(32ull * S) + pad256(S) : 0) +
                    (need_dig ? pad256(32ull * n) : 0);
    AHV_TRY(scratch_reserve(ctx, need));
    unsigned long long *fb = (unsigned long long *)scratch_take(ctx, 8ull * S);
    uint8_t *first = nullptr;
    if (!uniform && n) first = (uint8_t *)scratch_take(ctx, n);
    uint8_t *dig = d_digests;
    if (need_dig && n) dig = (uint8_t *)scratch_take(ctx, 32ull * n);
    if (!fb || (!uniform && n && !first) || (need_dig && n && !dig)) {
        set_error("internal: scratch too small for verify");
        return AHV_E_NOMEM;
    }
    k_fill_u64<<<grid_for(S, 256), 256, 0, ctx->stream>>>(fb, S, FB_NONE);
    AHV_CHECK_LAUNCH(ctx);
    if (first) {
        AHV_CUDA(cudaMemsetAsync(first, 0, n, ctx->stream));
        k_mark_first<<<grid_for(S, 256), 256, 0, ctx->stream>>>(d_rowptr, S, first);
        AHV_CHECK_LAUNCH(ctx);
    }
    VerifyArgs va;
    va.parents = d_parents;
    va.kind = d_kind;
    va.first = first;
    va.first_bad = fb;
    va.sm.rowptr = d_rowptr;
    va.sm.S = S;
    va.sm.uniform = (uint32_t)uniform;
    AHV_TRY(run_leaves(ctx, d_arena, d_offsets, fixed_len, n, dig, &va));
    k_verify_finalize<<<grid_for(S, 256), 256, 0, ctx->stream>>>(fb, S, d_verdict, d_first_bad);
    AHV_CHECK_LAUNCH(ctx);
    if (want_tree) {
        uint8_t *roots = d_roots ? d_roots : (uint8_t *)scratch_take(ctx, 32ull * S);
        uint8_t *empty = (uint8_t *)scratch_take(ctx, S);
        if (!roots || !empty) {
            set_error("internal: scratch too small for verify roots");
            return AHV_E_NOMEM;
        }
        AHV_TRY(run_tree(ctx, dig, d_rowptr, S, n, uniform, max_leaves, roots, empty));
        if (d_match) {
            k_root_compare<<<grid_for(S, 256), 256, 0, ctx->stream>>>(roots, d_expected, empty, S, d_match);
            AHV_CHECK_LAUNCH(ctx);
        }
    }
    return AHV_OK;
}

int ahv_sha256_one(ahv_ctx *ctx, const uint8_t *msg, uint64_t len, uint8_t *out_digest)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(out_digest && (msg || len == 0), "msg / out_digest is null");
    if (len > AHV_ONE_MAX) {                      // too long to ride in the launch: the batch path
        const uint64_t offs[2] = {0, len};
        return ahv_sha256_batch(ctx, msg, offs, 1, 0, out_digest);
    }
    if (!ctx->one_out) {
        void *h = nullptr;
        AHV_CUDA(cudaHostAlloc(&h, 64, cudaHostAllocMapped));
        ctx->one_out = (uint8_t *)h;
    }
    void *d_out = nullptr;
    AHV_CUDA(cudaHostGetDevicePointer(&d_out, ctx->one_out, 0));
    OneMsg m;
    m.len = (uint32_t)len;
    if (len) memcpy(m.data, msg, len);
    // plain additions here: one lane's dependency chain is what counts, not the pipe balance of a full chip
    // (1.7 us per block against 2.2 us with the scaled-sum form)
    k_sha256_one<false><<<1, 32, 0, ctx->stream>>>(m, (uint8_t *)d_out, SHA_SC);
    AHV_CHECK_LAUNCH(ctx);
    AHV_CUDA(cudaStreamSynchronize(ctx->stream));
    memcpy(out_digest, ctx->one_out, 32);
    return AHV_OK;
}

int ahv_extract_parents_dev(ahv_ctx *ctx, const uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                            uint64_t fixed_len, const uint32_t *d_parent_pos, uint64_t n, uint8_t *d_out_parents,
                            uint8_t *d_out_kind)
{
    AHV_TRY(use_device(ctx));
    (void)arena_bytes;
    if (n == 0) return AHV_OK;
    AHV_REQUIRE(d_parent_pos && d_out_parents && d_out_kind, "parent_pos / out_parent_digests / out_parent_kind is null");
    AHV_REQUIRE(d_offsets || fixed_len > 0, "offsets is NULL and fixed_len is 0");
    AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_out_parents) & 15u) == 0, "out_parent_digests must be 16-byte aligned");
    AHV_TRY(check_sizes(n, 0));
    k_extract_parents<<<grid_for(n, 256), 256, 0, ctx->stream>>>(d_arena, d_offsets, fixed_len, d_parent_pos, n, d_out_parents,
                                                                 d_out_kind);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

int ahv_chain_capture_dev(ahv_ctx *ctx, uint8_t *d_arena, uint64_t arena_bytes, const uint64_t *d_offsets,
                          const uint32_t *d_hexpos, const uint64_t *d_rowptr, uint64_t S, uint64_t n,
                          uint8_t *d_digests, uint8_t *d_parents, uint8_t *d_kind)
{
    AHV_TRY(use_device(ctx));
    (void)arena_bytes;
    AHV_REQUIRE((d_offsets && d_hexpos && d_rowptr) || n == 0, "offsets / hexpos / session_rowptr is null");
    AHV_TRY(check_sizes(n, S));
    if (S == 0 || n == 0) return AHV_OK;
    // one lane per session: with fewer lanes than the chip can keep busy a lane's dependency chain is what
    // counts, and the plain-addition form has the shorter one.
    // Sessions per warp: as few as keep the number of warps within the number of warp schedulers (4 per SM).  A
    // warp instruction occupies its scheduler's integer pipe for two cycles whatever the number of active lanes, so
    // thinner warps do not run their chains faster by themselves - two warps on one scheduler take turns, which
    // doubles the time (10^4 sessions x 1000: 9.3 ms with 17 lanes = 589 warps, 14.1 with 16 = 625 warps on 592
    // schedulers) - but every load and store of a thinner warp touches fewer lines (10^3 sessions: 8.7 ms with 32
    // lanes, 7.7 with 2).  Option chain_lanes fixes the number.
    const uint64_t schedulers = 4ull * (uint64_t)ctx->sm_count;
    uint64_t lpw = ctx->opt_chain_lanes ? (uint64_t)ctx->opt_chain_lanes : (S + schedulers - 1) / schedulers;
    lpw = lpw < 1 ? 1 : (lpw > 32 ? 32 : lpw);
    const uint64_t warps = (S + lpw - 1) / lpw;
    const unsigned g = grid_for(warps * 32, 64);
    // The blocks in front of a delta's parent field do not depend on the chain: hashed first, all deltas in
    // parallel, their state parked in the delta's digest slot (so only when digests are asked for).  For the
    // 256-byte deltas of config 2 that is one compression of five, for benchmark-shaped deltas (the field
    // sits behind the change list) three of six.
    // Only while the pass is bound by the length of a chain: with more than 2^16 sessions every scheduler has several
    // warps, the compressions themselves bound it, and the pre-pass is extra traffic (10^5 x 100: 4.3 ms without, 4.9 with).
    const bool prefix = d_digests && (ctx->opt_chain_prefix == 2 || (ctx->opt_chain_prefix == 1 && S <= 65536));
    if (prefix) {
        AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_digests) & 15u) == 0, "out_digests must be 16-byte aligned");
        if (ctx->opt_compress)
            k_chain_prefix<true><<<grid_for(n, 128), 128, 0, ctx->stream>>>(d_arena, d_offsets, d_hexpos, n, d_digests, SHA_SC);
        else
            k_chain_prefix<false><<<grid_for(n, 128), 128, 0, ctx->stream>>>(d_arena, d_offsets, d_hexpos, n, d_digests, SHA_SC);
        AHV_CHECK_LAUNCH(ctx);
    }
#define AHV_CHAIN_LAUNCH(SC, PF)                                                                                          \
    k_chain_capture<SC, PF><<<g, 64, 0, ctx->stream>>>(d_arena, d_offsets, d_hexpos, d_rowptr, S, d_digests, d_parents, d_kind, \
                                                       SHA_SC, (uint32_t)lpw)
    if (ctx->opt_compress && S > 65536) {
        if (prefix) AHV_CHAIN_LAUNCH(true, true);
        else AHV_CHAIN_LAUNCH(true, false);
    } else {
        if (prefix) AHV_CHAIN_LAUNCH(false, true);
        else AHV_CHAIN_LAUNCH(false, false);
    }
#undef AHV_CHAIN_LAUNCH
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

}  // extern "C"

// =================================================================== allocation-free Merkle roots
//
// ahv_merkle_roots_noalloc_dev: same result as ahv_merkle_roots_dev, for callers that must not
// allocate device memory while they run.  Two kinds of caller are in that position:
//   * a service that shares the GPU with another framework whose allocator already owns the rest
//     of HBM (a fixed budget carved up at start-up): a cudaMalloc issued later, when a bigger
//     batch happens to arrive, can fail or fragment the other tenant's pool;
//   * a latency-sensitive loop: cudaMalloc / cudaFree synchronise the whole device, so the
//     growable slab behind ahv_merkle_roots_dev can stall every other stream the first time a
//     larger batch shows up.
// Here the level buffers and per-level row pointers live in ONE fixed array inside the library
// image (AHV_NOALLOC_SCRATCH_BYTES = 8 MiB).  Every GPU that loads the library holds that array
// from load until the process exits, whether or not this entry point is ever called.
//
// Rules for users:
//   * Batches whose scratch need exceeds the array are refused with AHV_E_TOOLARGE before anything
//     is launched.  The need is 2 * 32 * (n_leaves/2 + n_sessions + 1) bytes for the two level
//     buffers plus 8 * (n_sessions + 1) bytes per level reduced outside the in-CTA finish, each
//     part rounded up to 256 bytes; roughly 32 bytes per leaf + 72 bytes per session, i.e. about
//     250 000 leaves in a handful of sessions.  max_session_leaves must be given (> 0): finding it
//     would need a device-to-host read back.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress,
//     with that device current) and remembers it; make that first call before any stream capture
//     begins on that device.  No call, first or later, uses cudaMalloc, cudaFree or any other
//     allocating routine.
//   * All calls on a device share the one array, and it still holds whatever the previous call left
//     there.  Nothing in it has to start at zero: every byte a call reads it has written earlier in
//     the same call (each level is produced in full before the next level reads it), so no clearing
//     is needed between calls.  For the same reason two calls whose kernels overlap on the GPU
//     (different streams, contexts or host threads) race on the array; run them one after another.
namespace {

constexpr uint64_t AHV_NOALLOC_SCRATCH_BYTES = 8ull << 20;
constexpr int AHV_NOALLOC_MAX_DEVICES = 64;

// In host code this name is only a handle for the runtime, never an address to hand to a kernel.
static __device__ __align__(256) uint8_t g_noalloc_scratch[AHV_NOALLOC_SCRATCH_BYTES];

// device address of g_noalloc_scratch on each device, looked up once per device
uint8_t *g_noalloc_addr[AHV_NOALLOC_MAX_DEVICES] = {};

int noalloc_base(ahv_ctx *ctx, uint8_t **out)
{
    if (ctx->device < 0 || ctx->device >= AHV_NOALLOC_MAX_DEVICES) {
        set_error("device ordinal %d not supported by the allocation-free entry point (limit %d)", ctx->device,
                  AHV_NOALLOC_MAX_DEVICES - 1);
        return AHV_E_INVALID;
    }
    uint8_t *p = __atomic_load_n(&g_noalloc_addr[ctx->device], __ATOMIC_ACQUIRE);
    if (!p) {
        // use_device() has already made ctx->device current on this thread
        void *sym = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);
        if (e != cudaSuccess || !sym) {
            cudaGetLastError();
            set_error("cudaGetSymbolAddress(static scratch) failed on device %d: %s", ctx->device,
                      cudaGetErrorString(e));
            return AHV_E_CUDA;
        }
        p = (uint8_t *)sym;
        __atomic_store_n(&g_noalloc_addr[ctx->device], p, __ATOMIC_RELEASE);   // same value from any thread
    }
    *out = p;
    return AHV_OK;
}

}  // namespace

extern "C" int ahv_merkle_roots_noalloc_dev(ahv_ctx *ctx, const uint8_t *d_digests, const uint64_t *d_rowptr,
                                            uint64_t S, uint64_t n, uint64_t max_leaves, uint8_t *d_roots,
                                            uint8_t *d_empty)
{
    AHV_TRY(use_device(ctx));
    AHV_REQUIRE(d_roots || S == 0, "out_roots is null");
    AHV_REQUIRE(d_rowptr || S == 0, "session_rowptr is null");
    AHV_REQUIRE((reinterpret_cast<uintptr_t>(d_digests) & 15u) == 0 && (reinterpret_cast<uintptr_t>(d_roots) & 15u) == 0,
                "digest buffers must be 16-byte aligned");
    AHV_REQUIRE(max_leaves > 0 || S == 0, "max_session_leaves must be given (> 0) for the allocation-free entry point");
    AHV_TRY(check_sizes(n, S));
    if (S == 0) return AHV_OK;

    const uint32_t G = global_levels(ctx, S, max_leaves);
    const uint64_t lvl_bytes = pad256(32ull * (n / 2 + S + 1));
    const uint64_t rp_bytes = G ? pad256(8ull * G * (S + 1)) : 0;
    const uint64_t need = G ? 2 * lvl_bytes + rp_bytes : 0;
    if (need > AHV_NOALLOC_SCRATCH_BYTES) {
        set_error("batch needs %llu bytes of scratch but the allocation-free entry point has a fixed %llu-byte array "
                  "(n_leaves=%llu, n_sessions=%llu); use ahv_merkle_roots_dev or split the batch",
                  (unsigned long long)need, (unsigned long long)AHV_NOALLOC_SCRATCH_BYTES, (unsigned long long)n,
                  (unsigned long long)S);
        return AHV_E_TOOLARGE;
    }
    if (S > 0x7fffffffull) {
        set_error("more than 2^31 - 1 sessions in one allocation-free call");
        return AHV_E_TOOLARGE;
    }
    uint8_t *base = nullptr;
    AHV_TRY(noalloc_base(ctx, &base));

    const uint8_t *cur = d_digests;
    const uint64_t *rp_cur = d_rowptr;
    if (G > 0) {
        uint8_t *bufA = base;
        uint8_t *bufB = base + lvl_bytes;
        uint64_t *lvl_rp = reinterpret_cast<uint64_t *>(base + 2 * lvl_bytes);
        k_level_rowptrs<<<G, 1024, 0, ctx->stream>>>(d_rowptr, S, lvl_rp);
        AHV_CHECK_LAUNCH(ctx);
        for (uint32_t l = 1; l <= G; ++l) {
            uint8_t *dst = (l & 1) ? bufA : bufB;
            const uint64_t *rp_out = lvl_rp + (uint64_t)(l - 1) * (S + 1);
            const uint64_t bound = ((n + ((1ull << l) - 1)) >> l) + S;
            k_merkle_level<false, true><<<grid_for(bound, 128), 128, 0, ctx->stream>>>(cur, dst, S, 0, 0, rp_cur, rp_out, SHA_SC);
            AHV_CHECK_LAUNCH(ctx);
            rp_cur = rp_out;
            cur = dst;
        }
    }
    k_merkle_top<false, true><<<(unsigned)S, 128, 0, ctx->stream>>>(cur, S, 0, rp_cur, d_roots, d_empty, SHA_SC, ctx->status_d);
    AHV_CHECK_LAUNCH(ctx);
    return AHV_OK;
}

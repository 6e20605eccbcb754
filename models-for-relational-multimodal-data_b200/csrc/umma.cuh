// tcgen05 (5th-gen tensor core) building blocks for the per-edge kernels, sm_100a inline PTX.
//
// Operand tiles live in shared memory in the no-swizzle "interleaved" core-matrix format
//     byte(r, c) = (c / 4) * (ROWS * 16) + r * 16 + (c % 4) * 4        (fp32 / tf32 elements)
// i.e. 16-byte column chunks form planes of ROWS x 16 B.  One physical tile is at the same
// time a K-major operand [rows x cols] (LBO = plane, SBO = 128 B) and an MN-major operand
// [cols x rows] (LBO = 128 B, SBO = plane), which is what lets the data GEMM (edges x H) and
// the weight-gradient GEMM (H x edges) share it.
//
// fp32 parity (1e-5) on tensor cores: 3xTF32.  x = hi + lo with hi = x rounded to tf32
// (low 13 mantissa bits zero) and lo = x - hi exactly; A*B ~= Alo*Bhi + Ahi*Blo + Ahi*Bhi
// accumulated in fp32 in tensor memory (small terms first).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200mp {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- tf32 split
__device__ __forceinline__ float tf32_hi(float x) {
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void split4(float4 v, float4& hi, float4& lo) {
    hi = make_float4(tf32_hi(v.x), tf32_hi(v.y), tf32_hi(v.z), tf32_hi(v.w));
    lo = make_float4(v.x - hi.x, v.y - hi.y, v.z - hi.z, v.w - hi.w);
}

// ---- descriptors
// shared-memory matrix descriptor, no swizzle, sm_100 version bit set
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout = 0) {
    uint64_t d = (uint64_t)layout << 61;
    d |= (uint64_t)((saddr >> 4) & 0x3FFFu);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// instruction descriptor: kind::tf32, fp32 accumulate
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ---- tensor memory
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {   // one full warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {     // the allocating warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// make generic-proxy shared-memory writes visible to the async proxy (the MMA's operand reads)
// one 128-byte line towards L2 (no register, no wait): rows a persistent CTA will gather a tile from now are first touches in
// DRAM (~1 us) unless announced
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t ok = 0;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
// arrives on the mbarrier once every previously issued MMA of this thread has completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- MMA issue (one thread).  D[tmem] (+)= A * B
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n}"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// ---- tensor memory <-> registers: 16 consecutive columns of this thread's lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])),
                   "r"(__float_as_uint(v[3])), "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])),
                   "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])), "r"(__float_as_uint(v[8])),
                   "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
                   "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])),
                   "r"(__float_as_uint(v[15])) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- tile geometry (fp32 elements, H columns)
template <int ROWS, int H>
struct TileFmt {
    static constexpr int PLANE = ROWS * 16;                  // bytes per 16-byte column chunk plane
    static constexpr int BYTES = PLANE * (H / 4);
    static constexpr int FLOATS = BYTES / 4;
    __device__ static __forceinline__ int off(int r, int chunk) { return chunk * (ROWS * 4) + r * 4; }   // float offset of (r, 4*chunk)
};

// Issue the tf32 MMAs of  D[ROWS x N] (+)= A[ROWS x K] * B^T  where A is a K-major tile view
// (rows x K) and B a tile [N x K]: K-major if !B_MN (B = W[n][k], out = A W^T) or, for B_MN,
// the MN-major view of a tile stored [k][n] (out = A W).  `first` = 0 overwrites D.
template <int ROWS, int NB, int K, bool B_MN>
__device__ __forceinline__ void issue_data_gemm(uint32_t d_tmem, uint32_t a_saddr, uint32_t b_saddr, uint32_t accumulate) {
    constexpr uint32_t idesc = idesc_tf32(ROWS, NB, 0, B_MN ? 1 : 0);
    constexpr uint32_t A_PLANE = ROWS * 16;
    // B as [N x K] K-major tile: planes over K chunks of N rows;  as MN-major view of a [K x N] tile:
    // planes over N chunks of K rows
    constexpr uint32_t B_PLANE = (B_MN ? K : NB) * 16;
#pragma unroll
    for (int ks = 0; ks < K / 8; ++ks) {
        const uint64_t ad = smem_desc(a_saddr + ks * 2 * A_PLANE, A_PLANE, 128);
        const uint64_t bd = B_MN ? smem_desc(b_saddr + ks * 128, 128, B_PLANE)
                                 : smem_desc(b_saddr + ks * 2 * B_PLANE, B_PLANE, 128);
        mma_ss(d_tmem, ad, bd, idesc, (ks > 0) ? 1u : accumulate);
    }
}

// Weight-gradient GEMM  D[MA x NB] (+)= G^T X  with G a tile [ROWS x MA] and X a tile [ROWS x NB]
// (both read through their MN-major views; the reduction runs over the ROWS edges of the tile).
template <int ROWS, int MA, int NB>
__device__ __forceinline__ void issue_wgrad_gemm(uint32_t d_tmem, uint32_t g_saddr, uint32_t x_saddr, uint32_t accumulate) {
    constexpr uint32_t idesc = idesc_tf32(MA, NB, 1, 1);
    constexpr uint32_t PLANE = ROWS * 16;
#pragma unroll
    for (int ks = 0; ks < ROWS / 8; ++ks) {
        const uint64_t ad = smem_desc(g_saddr + ks * 128, 128, PLANE);
        const uint64_t bd = smem_desc(x_saddr + ks * 128, 128, PLANE);
        mma_ss(d_tmem, ad, bd, idesc, (ks > 0) ? 1u : accumulate);
    }
}

// ---- activation tiles: "row32" format (layout type 1 = 128-byte swizzle with 32-byte atomicity)
// A [ROWS x H] fp32 tile is H/32 panels of [ROWS][32] floats, 128-byte rows, the 32-byte unit
// index XOR-ed with (row % 4):
//     byte(r, c) = (c / 32) * ROWS * 128 + r * 128 + ((((c % 32) / 8) ^ (r % 4)) * 32) + (c % 8) * 4
// Measured on B200 (scripts/umma_probe.py): kind::tf32 accepts MN-major operands ONLY with this
// layout type (types 0/2/4/6 read zeros), and the same bytes are also a valid K-major operand
// under layout type 1 (4-row groups at SBO = 512).  So one tile serves the data GEMM
// (edges x H, K-major) and the weight-gradient GEMM (H x edges, MN-major).  Panels must be
// 512-byte aligned (the XOR uses absolute address bits [7,9)).
template <int ROWS, int H>
struct RowFmt {
    static constexpr int PANEL = ROWS * 128;                 // bytes per 32-column panel
    static constexpr int BYTES = PANEL * (H / 32);
    static constexpr int FLOATS = BYTES / 4;
    // float offset of the 16-byte chunk j (columns 4j..4j+3) of row r
    __device__ static __forceinline__ int off(int r, int j) {
        return (j >> 3) * (PANEL / 4) + r * 32 + ((((j & 7) >> 1) ^ (r & 3)) << 3) + ((j & 1) << 2);
    }
};

// D[ROWS x NB] (+)= A * W^T : A = row32 tile [ROWS x K] (K-major view), W = interleaved K-major
// tile [NB x K] (layout type 0).
template <int ROWS, int NB, int K>
__device__ __forceinline__ void issue_data_gemm_row32(uint32_t d_tmem, uint32_t a_saddr, uint32_t w_saddr, uint32_t accumulate) {
    constexpr uint32_t idesc = idesc_tf32(ROWS, NB, 0, 0);
    constexpr uint32_t W_PLANE = NB * 16;
#pragma unroll
    for (int ks = 0; ks < K / 8; ++ks) {
        const uint64_t ad = smem_desc(a_saddr + (ks >> 2) * (ROWS * 128) + (ks & 3) * 32, ROWS * 128, 512, 1);
        const uint64_t bd = smem_desc(w_saddr + ks * 2 * W_PLANE, W_PLANE, 128, 0);
        mma_ss(d_tmem, ad, bd, idesc, (ks > 0) ? 1u : accumulate);
    }
}

// D[MA x NB] (+)= G^T X : G = row32 tile [ROWS x MA], X = row32 tile [ROWS x NB], both through
// their MN-major views; the reduction runs over the ROWS edges.
template <int ROWS, int MA, int NB>
__device__ __forceinline__ void issue_wgrad_gemm_row32(uint32_t d_tmem, uint32_t g_saddr, uint32_t x_saddr, uint32_t accumulate) {
    constexpr uint32_t idesc = idesc_tf32(MA, NB, 1, 1);
#pragma unroll
    for (int ks = 0; ks < ROWS / 8; ++ks) {
        const uint64_t ad = smem_desc(g_saddr + ks * 1024, ROWS * 128, 512, 1);
        const uint64_t bd = smem_desc(x_saddr + ks * 1024, ROWS * 128, 512, 1);
        mma_ss(d_tmem, ad, bd, idesc, (ks > 0) ? 1u : accumulate);
    }
}

}  // namespace umma
}  // namespace b200mp

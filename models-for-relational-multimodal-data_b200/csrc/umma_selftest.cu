// b200mp_umma_selftest: one-CTA checks of the tcgen05 building blocks in umma.cuh (descriptor
// encodings, K-major / MN-major views of the interleaved tile, M=64 accumulator layout, A from
// tensor memory, 3xTF32 accuracy).  Test infrastructure for tests/test_gpu_umma.py.
//   mode 0: Y[128x64] = X W^T   (A K-major, B K-major)            3xTF32
//   mode 1: Y[128x64] = X W     (A K-major, B MN-major view)      3xTF32
//   mode 2: Y[64x64]  = X^T G   (A, B MN-major views, M = 64)     3xTF32   (W = G [128x64])
//   mode 3: mode 0 with the A_lo operand read from tensor memory  3xTF32
//   mode 4: mode 0 single pass (hi*hi only) -- shows the tf32 error the split removes
#include "common.cuh"
#include "umma.cuh"

namespace b200mp {
using namespace umma;

// coalesced load of a [ROWS x 64] fp32 tile into hi / lo interleaved tiles (128 threads)
template <int ROWS>
__device__ __forceinline__ void selftest_load(float* hi, float* lo, const float* __restrict__ src, int tid) {
    const int w = tid >> 5, lane = tid & 31, r8 = lane & 7, c = 4 * w + (lane >> 3);
#pragma unroll 4
    for (int i = 0; i < ROWS / 8; ++i) {
        const int r = i * 8 + r8;
        float4 v = ldg4(src + (int64_t)r * 64 + c * 4), h, l;
        split4(v, h, l);
        st4(hi + TileFmt<ROWS, 64>::off(r, c), h);
        st4(lo + TileFmt<ROWS, 64>::off(r, c), l);
    }
}

// the same into row32 tiles (8 lanes per 128-byte half row)
template <int ROWS>
__device__ __forceinline__ void selftest_load_row32(float* hi, float* lo, const float* __restrict__ src, int tid) {
    const int j = tid & 15, r0 = tid >> 4;
#pragma unroll 4
    for (int i = 0; i < ROWS / 8; ++i) {
        const int r = i * 8 + r0;
        float4 v = ldg4(src + (int64_t)r * 64 + j * 4), h, l;
        split4(v, h, l);
        st4(hi + RowFmt<ROWS, 64>::off(r, j), h);
        st4(lo + RowFmt<ROWS, 64>::off(r, j), l);
    }
}

// probe variants of the MN-major descriptors (selftest only): explicit LBO / SBO
template <int ROWS, int NB, int K>
__device__ __forceinline__ void probe_data_gemm_bmn(uint32_t d, uint32_t a, uint32_t b, uint32_t acc, uint32_t kstep,
                                                    uint32_t lbo, uint32_t sbo) {
    constexpr uint32_t idesc = idesc_tf32(ROWS, NB, 0, 1);
    for (int ks = 0; ks < K / 8; ++ks)
        mma_ss(d, smem_desc(a + ks * 2 * ROWS * 16, ROWS * 16, 128), smem_desc(b + ks * kstep, lbo, sbo), idesc, ks > 0 ? 1u : acc);
}
template <int ROWS, int MA, int NB>
__device__ __forceinline__ void probe_wgrad(uint32_t d, uint32_t g, uint32_t x, uint32_t acc, uint32_t kstep, uint32_t lbo, uint32_t sbo) {
    constexpr uint32_t idesc = idesc_tf32(MA, NB, 1, 1);
    for (int ks = 0; ks < ROWS / 8; ++ks)
        mma_ss(d, smem_desc(g + ks * kstep, lbo, sbo), smem_desc(x + ks * kstep, lbo, sbo), idesc, ks > 0 ? 1u : acc);
}

__global__ void __launch_bounds__(128) umma_selftest_kernel(int mode_in, const float* __restrict__ X, const float* __restrict__ W,
                                                             float* __restrict__ Y) {
    const int mode = mode_in & 0xff, variant = mode_in >> 8;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    using TA = TileFmt<128, 64>;
    using TW = TileFmt<64, 64>;
    float* a_hi = smem;
    float* a_lo = a_hi + TA::FLOATS;
    float* b_hi = a_lo + TA::FLOATS;          // mode 2: a second 128-row tile (G)
    float* b_lo = b_hi + TA::FLOATS;
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;

    if (warp == 0) tmem_alloc(&tmem_base, 128);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (mode >= 5) selftest_load_row32<128>(a_hi, a_lo, X, tid);
    else selftest_load<128>(a_hi, a_lo, X, tid);
    if (mode == 6) selftest_load_row32<128>(b_hi, b_lo, W, tid);
    else if (mode == 2) selftest_load<128>(b_hi, b_lo, W, tid);
    else selftest_load<64>(b_hi, b_lo, W, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = tmem_base;
    const uint32_t lane_base = tm + ((uint32_t)(warp * 32) << 16);

    if (mode == 3) {          // stage A_lo in tensor memory columns [64, 128): lane = row, column = k
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float v[16];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4 t = lds4(a_lo + TA::off(tid, q * 4 + j));
                v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
            }
            tmem_st16(lane_base + 64 + q * 16, v);
        }
        tmem_st_wait();
        fence_before();
        __syncthreads();
        fence_after();
    }

    if (variant & 2) {        // prefill the accumulator so that "MMA wrote zeros" and "MMA did nothing" differ
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = 7.0f;
#pragma unroll
        for (int q = 0; q < 4; ++q) tmem_st16(lane_base + q * 16, v);
        tmem_st_wait();
        fence_before();
        __syncthreads();
        fence_after();
    }
    if (tid == 0 && (variant & 1)) {
        const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
        // variant bits [2,4): 0 = (lbo 128, sbo plane), 1 = swapped, 2 = lbo plane sbo 128 with k step = plane*2
        const int v = (variant >> 2) & 3;
        const uint32_t plane = (mode == 1 ? 64 : 128) * 16;
        const uint32_t lbo = v == 0 ? 128 : plane, sbo = v == 0 ? plane : 128, kstep = v == 2 ? 2 * plane : 128;
        if (mode == 1) {
            probe_data_gemm_bmn<128, 64, 64>(tm, al, bh, 0, kstep, lbo, sbo);
            probe_data_gemm_bmn<128, 64, 64>(tm, ah, bl, 1, kstep, lbo, sbo);
            probe_data_gemm_bmn<128, 64, 64>(tm, ah, bh, 1, kstep, lbo, sbo);
        } else {
            probe_wgrad<128, 64, 64>(tm, al, bh, 0, kstep, lbo, sbo);
            probe_wgrad<128, 64, 64>(tm, ah, bl, 1, kstep, lbo, sbo);
            probe_wgrad<128, 64, 64>(tm, ah, bh, 1, kstep, lbo, sbo);
        }
        mma_commit(&bar);
    } else if (tid == 0) {
        const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
        if (mode == 0) {
            issue_data_gemm<128, 64, 64, false>(tm, al, bh, 0);
            issue_data_gemm<128, 64, 64, false>(tm, ah, bl, 1);
            issue_data_gemm<128, 64, 64, false>(tm, ah, bh, 1);
        } else if (mode == 1) {
            issue_data_gemm<128, 64, 64, true>(tm, al, bh, 0);
            issue_data_gemm<128, 64, 64, true>(tm, ah, bl, 1);
            issue_data_gemm<128, 64, 64, true>(tm, ah, bh, 1);
        } else if (mode == 2) {
            issue_wgrad_gemm<128, 64, 64>(tm, al, bh, 0);
            issue_wgrad_gemm<128, 64, 64>(tm, ah, bl, 1);
            issue_wgrad_gemm<128, 64, 64>(tm, ah, bh, 1);
        } else if (mode == 3) {
            constexpr uint32_t idesc = idesc_tf32(128, 64, 0, 0);
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
                mma_ts(tm, tm + 64 + ks * 8, smem_desc(bh + ks * 2 * TW::PLANE, TW::PLANE, 128), idesc, ks > 0);
            issue_data_gemm<128, 64, 64, false>(tm, ah, bl, 1);
            issue_data_gemm<128, 64, 64, false>(tm, ah, bh, 1);
        } else if (mode == 5) {
            issue_data_gemm_row32<128, 64, 64>(tm, al, bh, 0);
            issue_data_gemm_row32<128, 64, 64>(tm, ah, bl, 1);
            issue_data_gemm_row32<128, 64, 64>(tm, ah, bh, 1);
        } else if (mode == 6) {
            issue_wgrad_gemm_row32<128, 64, 64>(tm, al, bh, 0);
            issue_wgrad_gemm_row32<128, 64, 64>(tm, ah, bl, 1);
            issue_wgrad_gemm_row32<128, 64, 64>(tm, ah, bh, 1);
        } else {
            issue_data_gemm<128, 64, 64, false>(tm, ah, bh, 0);
        }
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after();

    if (mode == 2 || mode == 6) {          // M = 64 accumulator: row m sits in lane (m % 16) + 32 * (m / 16)
        const int lane = tid & 31;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float v[16];
            tmem_ld16(lane_base + q * 16, v);
            if (lane < 16) {
                float* o = Y + (int64_t)(warp * 16 + lane) * 64 + q * 16;
#pragma unroll
                for (int j = 0; j < 16; ++j) o[j] = v[j];
            }
        }
    } else {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float v[16];
            tmem_ld16(lane_base + q * 16, v);
            float* o = Y + (int64_t)tid * 64 + q * 16;
#pragma unroll
            for (int j = 0; j < 16; ++j) o[j] = v[j];
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 128);
}

// Raw probe: shared-memory images and descriptor fields come from the host, one tf32 pass.
__global__ void __launch_bounds__(128) umma_probe_kernel(const float4* __restrict__ a_img, int a_vec, const float4* __restrict__ b_img,
                                                         int b_vec, uint64_t adesc, uint64_t bdesc, int a_kstep, int b_kstep,
                                                         int ksteps, uint32_t idesc, float* __restrict__ D) {
    extern __shared__ __align__(128) float smem_raw[];
    float4* sa = reinterpret_cast<float4*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float4* sb = sa + 2048;                      // 32 KB each
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base;
    const int tid = threadIdx.x, warp = tid >> 5;
    if (warp == 0) tmem_alloc(&tmem_base, 64);
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    for (int i = tid; i < 12800; i += 128) sa[i] = make_float4(0.f, 0.f, 0.f, 0.f);      // 200 KB: wild strides stay in bounds
    __syncthreads();
    for (int i = tid; i < a_vec; i += 128) sa[i] = a_img[i];
    for (int i = tid; i < b_vec; i += 128) sb[i] = b_img[i];
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = tmem_base;
    const uint32_t lane_base = tm + ((uint32_t)(warp * 32) << 16);
    {
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = -7.0f;
#pragma unroll
        for (int q = 0; q < 4; ++q) tmem_st16(lane_base + q * 16, v);
        tmem_st_wait();
        fence_before();
        __syncthreads();
        fence_after();
    }
    if (tid == 0) {
        const uint32_t a0 = smem_u32(sa), b0 = smem_u32(sb);
        for (int ks = 0; ks < ksteps; ++ks)
            mma_ss(tm, adesc | (uint64_t)(((a0 + ks * a_kstep) >> 4) & 0x3FFFu), bdesc | (uint64_t)(((b0 + ks * b_kstep) >> 4) & 0x3FFFu),
                   idesc, ks > 0);
        mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    fence_after();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float v[16];
        tmem_ld16(lane_base + q * 16, v);
#pragma unroll
        for (int j = 0; j < 16; ++j) D[(int64_t)tid * 64 + q * 16 + j] = v[j];
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 64);
}

}  // namespace b200mp

extern "C" int b200mp_umma_probe(const float* a_img, int a_bytes, const float* b_img, int b_bytes, uint64_t adesc, uint64_t bdesc,
                                 int a_kstep, int b_kstep, int ksteps, uint32_t idesc, float* D, void* stream) {
    using namespace b200mp;
    if (!a_img || !b_img || !D || a_bytes > 32768 || b_bytes > 32768 || (a_bytes & 15) || (b_bytes & 15)) return B200MP_EINVAL;
    const size_t smem = 12800 * 16 + 1024;
    B200MP_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>((const float4*)a_img, a_bytes / 16, (const float4*)b_img, b_bytes / 16,
                                                             adesc, bdesc, a_kstep, b_kstep, ksteps, idesc, D);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_umma_selftest(int mode, const float* X, const float* W, float* Y, void* stream) {
    using namespace b200mp;
    if (mode < 0 || (mode & 0xff) > 6 || !X || !W || !Y) return B200MP_EINVAL;
    const size_t smem = 4 * umma::TileFmt<128, 64>::BYTES + 1024;
    B200MP_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    umma_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(mode, X, W, Y);
    B200MP_LAUNCH_CHECK();
    return 0;
}

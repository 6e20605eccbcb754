// Shared pieces of the tensor-core (tcgen05, 3xTF32) per-edge kernels for H = 64.
//
// One CTA = 256 threads = one tile of TE = 128 consecutive edges (segment order) at a time:
//   - the threads load [TE x 64] fp32 rows coalesced (16 lanes per 256-byte row), split them into
//     tf32 hi / lo and store them as row32 tiles (umma.cuh) in shared memory;
//   - one thread issues the MMAs (accumulators in tensor memory), tcgen05.commit -> mbarrier;
//   - in every epilogue thread (t, ch) = (tid % 128, tid / 128) owns columns 32 ch .. 32 ch + 31 of
//     edge row t (tensor-memory lane t; warps w and w + 4 share a lane quarter).
// Weights sit in shared memory for the CTA's lifetime as interleaved K-major [N x K] tiles.
#pragma once
#include "common.cuh"
#include "umma.cuh"

namespace b200mp {
namespace ue {
using namespace umma;

constexpr int TE = 128;          // edges per tile = MMA M
constexpr int NTH = 256;         // threads per CTA
constexpr int HH = 64;           // hidden width these kernels are written for
using RT = RowFmt<TE, HH>;       // activation tile, 32 KB
using WT = TileFmt<HH, HH>;      // weight tile, 16 KB

// 3xTF32 data GEMM  D[TE x 64] (+)= X W^T   (X = row32 hi/lo tiles, W = interleaved hi/lo tiles)
__device__ __forceinline__ void gemm3_data(uint32_t d, const float* x_hi, const float* x_lo, const float* w_hi, const float* w_lo,
                                           uint32_t accumulate) {
    issue_data_gemm_row32<TE, HH, HH>(d, smem_u32(x_lo), smem_u32(w_hi), accumulate);
    issue_data_gemm_row32<TE, HH, HH>(d, smem_u32(x_hi), smem_u32(w_lo), 1);
    issue_data_gemm_row32<TE, HH, HH>(d, smem_u32(x_hi), smem_u32(w_hi), 1);
}
// 3xTF32 weight-gradient GEMM  D[64 x 64] (+)= G^T X  over the TE edges of the tile
__device__ __forceinline__ void gemm3_wgrad(uint32_t d, const float* g_hi, const float* g_lo, const float* x_hi, const float* x_lo,
                                            uint32_t accumulate) {
    issue_wgrad_gemm_row32<TE, HH, HH>(d, smem_u32(g_lo), smem_u32(x_hi), accumulate);
    issue_wgrad_gemm_row32<TE, HH, HH>(d, smem_u32(g_hi), smem_u32(x_lo), 1);
    issue_wgrad_gemm_row32<TE, HH, HH>(d, smem_u32(g_hi), smem_u32(x_hi), 1);
}

// Weight [64 x 64] row-major in global -> interleaved K-major B tile (hi / lo).
// TRANSPOSE = false: B[n][k] = W[n][k]   (out = X W^T);   true: B[n][k] = W[k][n]   (out = X W)
template <bool TRANSPOSE>
__device__ __forceinline__ void load_weight_tiles(float* w_hi, float* w_lo, const float* __restrict__ W, int tid) {
    for (int i = tid; i < HH * HH; i += NTH) {
        const int a = i / HH, b = i % HH;            // W[a][b]
        const int n = TRANSPOSE ? b : a, k = TRANSPOSE ? a : b;
        const float v = __ldg(W + i), h = tf32_hi(v);
        const int o = WT::off(n, k >> 2) + (k & 3);
        w_hi[o] = h;
        w_lo[o] = v - h;
    }
}

// Coalesced tile rows -> registers: 256 threads, thread handles the 16-byte chunk j = tid % 16 of
// rows r0 + 16 i (r0 = tid / 16, i < 8).  Loads are branch-free (row index clamped) so that all
// eight stay in flight; rows >= nrows are zeroed when the registers are stored to shared memory,
// not here (a select right after the load would wait for it and defeat the prefetch).
struct RowRegs { float4 v[8]; };
__device__ __forceinline__ void load_rows(RowRegs& R, const float* __restrict__ src, const int32_t* __restrict__ rowmap,
                                          int64_t row0, int nrows, int tid) {
    const int j = tid & 15, r0 = tid >> 4;
    int64_t rows[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int rc = min(i * 16 + r0, nrows - 1);
        rows[i] = rowmap ? (int64_t)__ldg(rowmap + row0 + rc) : row0 + rc;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) R.v[i] = ldg4(src + rows[i] * HH + j * 4);
}
// R = T1[i1[row]] + T2[i2[row]] (either table nullable), same thread mapping as load_rows.  The
// row indices come from SHARED memory (staged one tile ahead): with the index loads in global
// memory ptxas serialises index -> address -> row per row and every row pays a DRAM latency.
__device__ __forceinline__ void gather_rows(RowRegs& R, const float* __restrict__ T1, const int* i1,
                                            const float* __restrict__ T2, const int* i2, int tid) {
    const int j = tid & 15, r0 = tid >> 4;
    float4 t[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        R.v[i] = T1 ? ldg4(T1 + (int64_t)i1[i * 16 + r0] * HH + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        t[i] = T2 ? ldg4(T2 + (int64_t)i2[i * 16 + r0] * HH + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) R.v[i] = R.v[i] + t[i];
}
__device__ __forceinline__ void store_rows_split(const RowRegs& R, float* hi, float* lo, int nrows, int tid) {
    const int j = tid & 15, r0 = tid >> 4;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = i * 16 + r0;
        float4 h, l;
        split4(r < nrows ? R.v[i] : make_float4(0.f, 0.f, 0.f, 0.f), h, l);
        const int o = RT::off(r, j);
        st4(hi + o, h);
        st4(lo + o, l);
    }
}
__device__ __forceinline__ void store_rows_raw(const RowRegs& R, float* dst, int nrows, int tid) {
    const int j = tid & 15, r0 = tid >> 4;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = i * 16 + r0;
        st4(dst + RT::off(r, j), r < nrows ? R.v[i] : make_float4(0.f, 0.f, 0.f, 0.f));
    }
}

// Column sums over the 32 rows of a warp: v[i] = this row's value in column i (i < 32).  Returns
// in lane L the sum of column L over the warp's rows (31 shuffles: each step halves the columns
// a lane is responsible for).
__device__ __forceinline__ float warp_column_sums(float (&v)[32], int lane) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        const bool up = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const float keep = up ? v[i + s] : v[i];
            const float send = up ? v[i] : v[i + s];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    return v[0];
}

// The M = 64 accumulator (weight gradient [64 x 64]) -> this CTA's partial in global memory.
// Row m sits in tensor-memory lane (m % 16) + 32 (m / 16): warp w, lanes 0..15 own rows 16w..16w+15.
__device__ __forceinline__ void store_wgrad_partial(uint32_t tmem_cols, float* __restrict__ part, float scale, bool valid, int tid) {
    const int warp = tid >> 5, lane = tid & 31;
    if (warp >= 4) return;                           // warp-uniform: rows 16w..16w+15 belong to warps 0..3
    const uint32_t base = tmem_cols + ((uint32_t)(warp * 32) << 16);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        float v[16];
        if (valid) tmem_ld16(base + q * 16, v);      // uniform branch (whole CTA)
        if (lane < 16) {
            float* o = part + (warp * 16 + lane) * HH + q * 16;
#pragma unroll
            for (int c = 0; c < 16; c += 4)
                st4(o + c, valid ? make_float4(scale * v[c], scale * v[c + 1], scale * v[c + 2], scale * v[c + 3])
                                 : make_float4(0.f, 0.f, 0.f, 0.f));
        }
    }
}

struct TileSync {
    uint64_t bar[2];
    uint32_t tmem_base;
};

}  // namespace ue
}  // namespace b200mp

// Shared pieces of the tensor-core (tcgen05, 3xTF32) per-edge kernels for H = 64.
//
// One CTA = NTH threads (256 or 512) = one tile of TE = 128 consecutive edges (segment order) at a
// time:
//   - the threads load [TE x 64] fp32 rows coalesced (16 lanes per 256-byte row), split them into
//     tf32 hi / lo and store them as row32 tiles (umma.cuh) in shared memory;
//   - one thread issues the MMAs (accumulators in tensor memory), tcgen05.commit -> mbarrier;
//   - in every epilogue thread (t, ch) = (tid % 128, tid / 128) owns the CW = 64 / (NTH / 128)
//     columns CW ch .. CW ch + CW - 1 of edge row t (tensor-memory lane t; warps w, w + 4, ...
//     share a lane quarter).
// The kernels whose tiles fill shared memory (one CTA per SM) run 512 threads: their SIMT phases
// are latency bound (shared-memory round trips, tensor-memory loads, shuffles) and 16 warps hide
// twice as much of it as 8.
// Weights sit in shared memory for the CTA's lifetime as interleaved K-major [N x K] tiles.
#pragma once
#include <cuda_bf16.h>
#include "common.cuh"
#include "umma.cuh"

namespace b200mp {
namespace ue {
using namespace umma;

// ---- edge-tensor storage type: float, or __nv_bfloat16 (bf16 STORAGE of every [E,H] tensor, fp32 arithmetic inside the kernels:
// BASELINE.json config 3).  Four consecutive elements <-> float4.
using bf16 = __nv_bfloat16;
__device__ __forceinline__ float4 ldg_e4(const float* p) { return ldg4(p); }
__device__ __forceinline__ float4 ldg_e4(const bf16* p) {
    const uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
    return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                       __uint_as_float(r.y & 0xffff0000u));
}
__device__ __forceinline__ float4 ld_e4(const float* p) { return *reinterpret_cast<const float4*>(p); }      // plain (read-modify-write paths)
__device__ __forceinline__ float4 ld_e4(const bf16* p) {
    const uint2 r = *reinterpret_cast<const uint2*>(p);
    return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                       __uint_as_float(r.y & 0xffff0000u));
}
__device__ __forceinline__ void st_e4(float* p, float4 v) { st4(p, v); }
__device__ __forceinline__ void st_e4(bf16* p, float4 v) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<const uint32_t*>(&a);
    r.y = *reinterpret_cast<const uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = r;
}

constexpr int TE = 128;          // edges per tile = MMA M
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
template <bool TRANSPOSE, int NTH>
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

// Coalesced tile rows -> registers: thread (g, j) = (tid / 16, tid % 16) handles the 16-byte chunk j
// of the NP = 16 TE / NTH ADJACENT rows g NP + i (i < NP); the 16 lanes of a group read one whole
// 256-byte row per access whatever the row is, so global loads, stores and REDs issued in this
// mapping are coalesced (4 lines per warp request; the row-per-thread mapping of the tensor-memory
// epilogues would touch 32 lines per request and saturates the L1 pipe).  Adjacent rows of one
// thread mostly share their segment, which lets per-row node data be reused.  Loads are branch-free
// (row index clamped) so that all of them stay in flight; rows >= nrows are zeroed when the
// registers are stored to shared memory, not here (a select right after the load would wait for
// it and defeat the prefetch).
template <int NTH>
struct RowRegs {
    static constexpr int NP = TE * 16 / NTH;
    float4 v[NP];
};
template <int NTH, typename T>
__device__ __forceinline__ void load_rows(RowRegs<NTH>& R, const T* __restrict__ src, const int32_t* __restrict__ rowmap,
                                          int64_t row0, int nrows, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int j = tid & 15, r0 = tid >> 4;
    int64_t rows[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const int rc = min(r0 * NP + i, nrows - 1);
        rows[i] = rowmap ? (int64_t)__ldg(rowmap + row0 + rc) : row0 + rc;
    }
#pragma unroll
    for (int i = 0; i < NP; ++i) R.v[i] = ldg_e4(src + rows[i] * HH + j * 4);
}
// R = T1[i1[row]] + T2[i2[row]] (either table nullable), same thread mapping as load_rows.  The
// row indices come from SHARED memory (staged one tile ahead): with the index loads in global
// memory ptxas serialises index -> address -> row per row and every row pays a DRAM latency.
template <int NTH>
__device__ __forceinline__ void gather_rows(RowRegs<NTH>& R, const float* __restrict__ T1, const int* i1,
                                            const float* __restrict__ T2, const int* i2, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int j = tid & 15, r0 = tid >> 4;
    float4 t[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        R.v[i] = T1 ? ldg4(T1 + (int64_t)i1[r0 * NP + i] * HH + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        t[i] = T2 ? ldg4(T2 + (int64_t)i2[r0 * NP + i] * HH + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < NP; ++i) R.v[i] = R.v[i] + t[i];
}
// The same gather with the two tables kept apart (no add yet): lets the caller issue further loads
// before the first use of these registers forces a wait on the gathered rows.
template <int NTH>
__device__ __forceinline__ void gather_rows_issue(RowRegs<NTH>& R1, RowRegs<NTH>& R2, const float* __restrict__ T1, const int* i1,
                                                  const float* __restrict__ T2, const int* i2, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int j = tid & 15, r0 = tid >> 4;
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        R1.v[i] = T1 ? ldg4(T1 + (int64_t)i1[r0 * NP + i] * HH + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        R2.v[i] = T2 ? ldg4(T2 + (int64_t)i2[r0 * NP + i] * HH + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
// Row ids of a tile through an optional map, and rows from precomputed ids: with the ids fetched one
// tile earlier the row loads do not wait for an index load (one DRAM latency instead of two).
template <int NTH>
struct RowIds { int v[RowRegs<NTH>::NP]; };
template <int NTH>
__device__ __forceinline__ void load_row_ids(RowIds<NTH>& I, const int32_t* __restrict__ rowmap, int64_t row0, int nrows, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int r0 = tid >> 4;
#pragma unroll
    for (int i = 0; i < NP; ++i) I.v[i] = __ldg(rowmap + row0 + min(r0 * NP + i, nrows - 1));
}
template <int NTH, typename T>
__device__ __forceinline__ void load_rows_by_id(RowRegs<NTH>& R, const T* __restrict__ src, const RowIds<NTH>& I, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int j = tid & 15;
#pragma unroll
    for (int i = 0; i < NP; ++i) R.v[i] = ldg_e4(src + (int64_t)I.v[i] * HH + j * 4);
}
template <int NTH>
__device__ __forceinline__ void store_rows_split(const RowRegs<NTH>& R, float* hi, float* lo, int nrows, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int j = tid & 15, r0 = tid >> 4;
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const int r = r0 * NP + i;
        float4 h, l;
        split4(r < nrows ? R.v[i] : make_float4(0.f, 0.f, 0.f, 0.f), h, l);
        const int o = RT::off(r, j);
        st4(hi + o, h);
        st4(lo + o, l);
    }
}
template <int NTH>
__device__ __forceinline__ void store_rows_raw(const RowRegs<NTH>& R, float* dst, int nrows, int tid) {
    constexpr int NP = RowRegs<NTH>::NP;
    const int j = tid & 15, r0 = tid >> 4;
#pragma unroll
    for (int i = 0; i < NP; ++i) {
        const int r = r0 * NP + i;
        st4(dst + RT::off(r, j), r < nrows ? R.v[i] : make_float4(0.f, 0.f, 0.f, 0.f));
    }
}

// Staging tile for accumulator rows on their way from tensor memory (row-per-thread) to the
// coalesced row mapping above: [TE][64] fp32, 256-byte rows, 16-byte chunk index XOR-ed with
// (row % 8) -- conflict free for both sides.  Not an MMA operand.
__device__ __forceinline__ int stage_off(int r, int c) { return r * HH + (((c ^ r) & 7) << 2) + ((c & 8) << 2); }
// thread (t, ch): its CW = 64 / (NTH / TE) accumulator columns of row t -> staging tile
template <int NTH>
__device__ __forceinline__ void stage_from_tmem(uint32_t taddr, float* S, int t, int ch) {
    constexpr int CW = HH / (NTH / TE);
#pragma unroll
    for (int q = 0; q < CW / 16; ++q) {
        float v[16];
        tmem_ld16(taddr + q * 16, v);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj)
            st4(S + stage_off(t, ch * (CW / 4) + q * 4 + jj), f4(v[4 * jj], v[4 * jj + 1], v[4 * jj + 2], v[4 * jj + 3]));
    }
}

// Column sums over the 32 rows of a warp: v[i] = this row's value in column i (i < W, W = 16 or
// 32).  Returns in lane L the sum of column L % W over the warp's rows (W - 1 shuffles: each step
// halves the columns a lane is responsible for; W = 16 adds one shuffle across the half warps).
template <int W>
__device__ __forceinline__ float warp_column_sums(float (&v)[W], int lane) {
#pragma unroll
    for (int s = W / 2; s >= 1; s >>= 1) {
        const bool up = (lane & s) != 0;
#pragma unroll
        for (int i = 0; i < s; ++i) {
            const float keep = up ? v[i + s] : v[i];
            const float send = up ? v[i] : v[i + s];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, s);
        }
    }
    if (W == 16) v[0] += __shfl_xor_sync(0xffffffffu, v[0], 16);
    return v[0];
}
// Cross-warp finish of per-warp column sums: bsum[warp][32] holds warp_column_sums results of
// every warp (column group ch = warp / 4, CW columns per group); returns column c (c < 64).
template <int NTH>
__device__ __forceinline__ float column_total(const float* bsum, int c) {
    constexpr int CW = HH / (NTH / TE);
    const int w0 = (c / CW) * 4, l = c % CW;
    return bsum[w0 * 32 + l] + bsum[(w0 + 1) * 32 + l] + bsum[(w0 + 2) * 32 + l] + bsum[(w0 + 3) * 32 + l];
}

// The M = 64 accumulator (weight gradient [64 x 64]) -> this CTA's partial in global memory.
// Row m sits in tensor-memory lane (m % 16) + 32 (m / 16): warp w, lanes 0..15 own rows 16w..16w+15.
// accumulate: add to what an earlier drain left in `part`.  The weight-gradient accumulators are drained every WGRAD_FLUSH_EDGES
// edges: the error of a sum kept in ONE tensor-memory accumulator grows linearly with the number of MMAs added into it (the
// tensor core's fp32 accumulation does not round to nearest).  Measured on B200 against fp64, edge update's gW2 = g^T h over 1 M
// edges, relative L2: 2.3e-5 with 3 400 edges per accumulator chain, 1.2e-5 at 2 048, 6.4e-6 at 1 024, 3.1e-6 at 512, 1.5e-6 at
// 256 (kernel time +0 / +1 / +3 / +6 / +14 %).  1 024 keeps the gradient inside the 1e-5 tolerance for +3 %.
constexpr int WGRAD_FLUSH_EDGES = 1024;
__device__ __forceinline__ void store_wgrad_partial(uint32_t tmem_cols, float* part, float scale, bool valid, int tid,
                                                    bool accumulate = false) {
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
            for (int c = 0; c < 16; c += 4) {
                float4 w = valid ? make_float4(scale * v[c], scale * v[c + 1], scale * v[c + 2], scale * v[c + 3])
                                 : make_float4(0.f, 0.f, 0.f, 0.f);
                // later drains add with a vector RED: fire and forget (a load-add-store chain costs an L2 round trip per step while
                // the other warps wait at the next barrier: measured +9..13 % kernel time at a 1024-edge interval); every address
                // has ONE writer thread and its REDs arrive in program order, so the sum is still run-to-run identical
                if (accumulate) red_add4(o + c, w); else st4(o + c, w);
            }
        }
    }
}

// ---- two independent 64-edge pipelines per CTA ---------------------------------------------------
// ncu on the one-tile-per-CTA version: 38 % of the warp samples sat in the two mbarrier spins (the
// tensor pipe's 3xTF32 time, fully exposed) and shared memory is too full for a second 128-edge tile.
// So the CTA's 512 threads form TWO groups of 256, each running the whole tile pipeline on its own
// 64-edge tile (M = 64 MMAs) with its own buffers, mbarriers, named barrier and tensor-memory columns,
// sharing only the weight tiles: while one group waits for its MMAs the other runs its SIMT phases.
namespace dual {
constexpr int RG = 64;                    // edges per group tile
constexpr int NTG = 256;                  // threads per group
constexpr int NPG = RG * 16 / NTG;        // 4 adjacent rows per thread in the coalesced mapping
using RG_T = RowFmt<RG, HH>;              // 16 KB operand tile

struct Rows { float4 v[NPG]; };
__device__ __forceinline__ void group_sync(int grp) { asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "r"(NTG) : "memory"); }
template <typename T>
__device__ __forceinline__ void load_rows(Rows& R, const T* __restrict__ src, const int32_t* __restrict__ rowmap,
                                          int64_t row0, int nrows, int lt) {
    const int j = lt & 15, g = lt >> 4;
    int64_t rows[NPG];
#pragma unroll
    for (int i = 0; i < NPG; ++i) {
        const int rc = min(g * NPG + i, nrows - 1);
        rows[i] = rowmap ? (int64_t)__ldg(rowmap + row0 + rc) : row0 + rc;
    }
#pragma unroll
    for (int i = 0; i < NPG; ++i) R.v[i] = ldg_e4(src + rows[i] * HH + j * 4);
}
__device__ __forceinline__ void store_rows_split(const Rows& R, float* hi, float* lo, int nrows, int lt) {
    const int j = lt & 15, g = lt >> 4;
#pragma unroll
    for (int i = 0; i < NPG; ++i) {
        const int r = g * NPG + i;
        float4 h, l;
        split4(r < nrows ? R.v[i] : make_float4(0.f, 0.f, 0.f, 0.f), h, l);
        const int o = RG_T::off(r, j);
        st4(hi + o, h);
        st4(lo + o, l);
    }
}
// M = 64 accumulator [64 x 64] -> staging tile: row m sits in tensor-memory lane (m % 16) + 32 (m / 16),
// so warp (quarter q = warp % 4, column half = group-local warp / 4) reads rows 16 q .. 16 q + 15 in its
// lanes 0..15 (lanes 16..31 carry nothing).
__device__ __forceinline__ void stage_from_tmem64(uint32_t tcols, float* S, int lwarp, int lane) {
    const int q = lwarp & 3, half = lwarp >> 2;
    const uint32_t taddr = tcols + ((uint32_t)(q * 32) << 16) + half * 32;
    const int r = q * 16 + (lane & 15);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        float v[16];
        tmem_ld16(taddr + k * 16, v);
        if (lane < 16) {
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
                st4(S + stage_off(r, half * 8 + k * 4 + jj), f4(v[4 * jj], v[4 * jj + 1], v[4 * jj + 2], v[4 * jj + 3]));
        }
    }
}
__device__ __forceinline__ void gemm3_data64(uint32_t d, const float* x_hi, const float* x_lo, const float* w_hi, const float* w_lo,
                                             uint32_t accumulate) {
    issue_data_gemm_row32<RG, HH, HH>(d, smem_u32(x_lo), smem_u32(w_hi), accumulate);
    issue_data_gemm_row32<RG, HH, HH>(d, smem_u32(x_hi), smem_u32(w_lo), 1);
    issue_data_gemm_row32<RG, HH, HH>(d, smem_u32(x_hi), smem_u32(w_hi), 1);
}
__device__ __forceinline__ void gemm3_wgrad64(uint32_t d, const float* g_hi, const float* g_lo, const float* x_hi, const float* x_lo,
                                              uint32_t accumulate) {
    issue_wgrad_gemm_row32<RG, HH, HH>(d, smem_u32(g_lo), smem_u32(x_hi), accumulate);
    issue_wgrad_gemm_row32<RG, HH, HH>(d, smem_u32(g_hi), smem_u32(x_lo), 1);
    issue_wgrad_gemm_row32<RG, HH, HH>(d, smem_u32(g_hi), smem_u32(x_hi), 1);
}
// per-thread column partials -> this group's column sums (red: [NTG / 16][64] floats of the group's smem)
__device__ __forceinline__ void group_column_sums(float* red, float4 acc, float* __restrict__ out, float scale, int lt, int grp) {
    group_sync(grp);
    st4(red + (lt >> 4) * HH + (lt & 15) * 4, acc);
    group_sync(grp);
    if (lt < HH) {
        float sum = 0.f;
#pragma unroll 8
        for (int g = 0; g < NTG / 16; ++g) sum += red[g * HH + lt];
        out[lt] = scale * sum;
    }
}
}  // namespace dual

struct TileSync {
    uint64_t bar[2];
    uint32_t tmem_base;
};

}  // namespace ue
}  // namespace b200mp

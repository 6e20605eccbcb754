// Node-level linears on the tensor cores (tcgen05, 3xTF32), fp32 in and out, ONE pass over HBM:
//
//   b200mp_node_linear      Y[N, NO]  = X[N, K] W^T (+ b)        PNAConv post_nn over [N, 4H] aggregates
//                                                                (pna.py:71-79 -> PyG PNAConv.update) and
//                                                                its data gradient (W passed transposed)
//   b200mp_node_linear_wgrad  G[K, NO] = X[N, K]^T Y[N, NO]      its weight gradient (reduction over N)
//
// The library path these replace is split(x) + three TF32 GEMMs, each re-reading an [N, K]
// operand and read-modify-writing the output: ~6K + 5NO floats of HBM traffic per node row against
// K + NO here.
//
// Data kernel: one CTA = 256 threads = one tile of 128 node rows, two CTAs per SM (their phases
// overlap each other).  K is walked in chunks of 32 columns: the X chunk [128 x 32] is loaded
// coalesced (8 lanes per 128-byte row piece), split into tf32 hi / lo and stored as one row32 panel
// each (umma.cuh); the matching weight chunk [NO x 32] hi / lo arrives as ONE bulk copy
// (cp.async.bulk -> mbarrier) of a pre-formatted image (interleaved K-major tile, built once per
// call on the host side from the tiny weight matrix).  Accumulator [128 x NO] in tensor memory.
#include "umma_edge.cuh"

namespace b200mp {
using namespace ue;

namespace ng {
constexpr int NT2 = 256;                  // threads per CTA
constexpr int KCH = 32;                   // K columns per chunk = one row32 panel
using XP = RowFmt<TE, KCH>;               // 16 KB panel

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct NodeLinArgs {
    int64_t N; int K; int ldx;
    const float* X; const float* Wimg; const float* bias; float* Y;
    const float* X1; const float* X2;         // segmented input: columns 64 s .. 64 s + 63 of X come from X<s>, each [N, 64]
    int ysplit;                               // output as NO / 64 separate [N, 64] blocks (Y + g N 64) instead of [N, NO]
    int ldy;                                  // row stride of Y in floats (NO unless the call fills a column block of a wider output)
    float* colsum_part;                       // CS variant: [gridDim.x][K] column sums of X over the rows this CTA handled
};
// Row epilogue of b200mp_row_linear (the per-edge linears at H = 128 run on this kernel): for tile row k
//   z = X[xrow[k]] W^T + bias + [pre-addend rows];  relu: z = max(z, 0);  [mask rows]: z = z * [mask > 0]
//   Y[orow[k]] = beta z + [post-addend rows * coef]
// The addend / mask rows come through two generic slots (pointer, optional row map, role), fetched into registers BEFORE the
// accumulator is staged, so that their latency overlaps the tensor-memory reads instead of following them.  A post-addend may
// alias Y when its row map equals orow (each element is read and written by one thread).
struct EpiSlot {
    const float* p; const int32_t* idx;       // rows of NO floats; idx null = tile row k
    int role;                                 // 0 unused | 1 added before the activation | 2 mask | 3 added after, times coef
    float coef;
};
struct RowEpi {
    const int32_t* xrow;
    EpiSlot s0, s1;
    const int32_t* orow;
    float beta;
    int relu;
};
struct NoEpi {};
}  // namespace ng
using namespace ng;

// Wimg: [K / 32][2 (hi, lo)][8 (16-byte k chunk)][NO][4] floats
//
// Pipeline (one CTA per SM, everything double buffered): chunk c's weights are bulk-copied while chunk
// c - 1's MMAs run, and chunk c's X panel is split and stored while they run too -- the serial version
// (two CTAs per SM, single buffers) spent ~3.4 us per chunk for 0.6 us of tensor work, mostly waiting for
// the 48-64 KB weight copy from L2.
// CS: the column sums of X fall out of the pass as a by-product (per-CTA partials, fixed summation order): with X = the gradients
// [g_0 | g_1 | ..] of a group of linears, they ARE the bias gradients, which otherwise cost one more pass over every g_i.
template <int NO, bool EPI = false, bool CS = false>
__global__ void __launch_bounds__(NT2, 1) node_linear_umma_kernel(NodeLinArgs a, typename std::conditional<EPI, RowEpi, NoEpi>::type ep) {
    constexpr int CS_K = 192;                 // widest segmented input (three [N,64] blocks)
    __shared__ float cs_s[CS ? NT2 / 32 : 1][CS ? CS_K : 1];      // one row per warp: no atomics, fixed order
    constexpr int W_PART = NO * KCH;                       // floats per hi (or lo) weight chunk
    constexpr int W_BUF = 2 * W_PART;                      // hi + lo
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;
    float* Xb = smem;                         // [2][hi | lo] X panels, 64 KB; the first 32 KB double as the epilogue's staging tile
    float* Wb = Xb + 4 * XP::FLOATS;          // [2][hi | lo] weight chunks
    __shared__ uint64_t bar_w[2], bar_m[2];
    __shared__ uint32_t tmem_base;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int t = tid & (TE - 1), ch = tid >> 7;           // tensor-memory side: row t, 32 columns
    const int j8 = tid & 7, g8 = tid >> 3;                 // chunk loads: 16-byte piece j8 of rows 4 g8 .. 4 g8 + 3
    const int j = tid & 15, g = tid >> 4;                  // output rows: 16-byte chunk j of rows 8 g .. 8 g + 7
    constexpr uint32_t NCOLS = NO <= 64 ? 64 : NO <= 128 ? 128 : 256;
    if (warp == 0) tmem_alloc(&tmem_base, NCOLS);
    if (tid == 0) {
        mbar_init(&bar_w[0], 1); mbar_init(&bar_w[1], 1); mbar_init(&bar_m[0], 1); mbar_init(&bar_m[1], 1);
        mbar_fence_init();
    }
    if constexpr (CS) {
        for (int i = tid; i < (NT2 / 32) * CS_K; i += NT2) (&cs_s[0][0])[i] = 0.f;
    }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = tmem_base;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const int nchunks = a.K / KCH;
    const int64_t ntiles = (a.N + TE - 1) / TE;
    const int64_t my_tiles = (int64_t)blockIdx.x < ntiles ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total = my_tiles * nchunks;              // chunks this CTA walks, numbered gc = 0 .. total - 1

    // X chunks are prefetched PF chunks ahead in registers (PF = 3 measured no faster than 1: the loop is
    // bound by the issue -> commit -> wait chain of one chunk's MMAs, not by the loads)
    constexpr int PF = 1;
    float4 xr[PF][4];
    int64_t pf_tile = blockIdx.x;             // prefetch cursor: next chunk to load
    int pf_c = 0;
    auto load_next = [&](float4 (&dst)[4]) {
        if (pf_tile >= ntiles) return;
        const int64_t r0 = pf_tile * TE;
        const int nr = (int)min64(TE, a.N - r0);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int rc = min(g8 * 4 + i, nr - 1);
            int64_t xr_ = r0 + rc;
            if constexpr (EPI) { if (ep.xrow) xr_ = __ldg(ep.xrow + xr_); }
            const float* xb = a.X1 ? (pf_c < 2 ? a.X : pf_c < 4 ? a.X1 : a.X2) + xr_ * 64 + (pf_c & 1) * KCH
                                   : a.X + xr_ * a.ldx + pf_c * KCH;
            dst[i] = ldg4(xb + j8 * 4);
        }
        if (++pf_c == nchunks) { pf_c = 0; pf_tile += gridDim.x; }
    };
    auto issue_w = [&](int64_t gc_) {         // one thread: weights of global chunk gc_ -> buffer gc_ & 1
        const int b = (int)(gc_ & 1);
        mbar_expect_tx(&bar_w[b], W_BUF * 4);
        bulk_g2s(Wb + b * W_BUF, a.Wimg + (size_t)(gc_ % nchunks) * W_BUF, W_BUF * 4, &bar_w[b]);
    };
    if (total > 0) {
#pragma unroll
        for (int k = 0; k < PF; ++k) load_next(xr[k]);
        if (tid == 0) issue_w(0);
    }

    int64_t gc = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t row0 = tile * TE;
        const int nrows = (int)min64(TE, a.N - row0);
        int e_r0[8], e_r1[8], e_or[8];            // EPI: row ids of this thread's 8 output rows (clamped), fetched ahead of the K loop
        if constexpr (EPI) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int64_t k = row0 + min(g * 8 + i, nrows - 1);
                e_r0[i] = (ep.s0.role && ep.s0.idx) ? __ldg(ep.s0.idx + k) : (int)k;
                e_r1[i] = (ep.s1.role && ep.s1.idx) ? __ldg(ep.s1.idx + k) : (int)k;
                e_or[i] = ep.orow ? __ldg(ep.orow + k) : (int)k;
            }
        }
        for (int c = 0; c < nchunks; ++c, ++gc) {
            const int b = (int)(gc & 1);
            const uint32_t ph = (uint32_t)((gc >> 1) & 1);            // phase of buffer b's barriers at their use number gc / 2
            float* X_hi = Xb + b * 2 * XP::FLOATS;
            float* X_lo = X_hi + XP::FLOATS;
            if (gc >= 2) {                     // the MMAs of chunk gc - 2 read X buffer b
                mbar_wait(&bar_m[b], ph ^ 1);
                fence_after();
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = g8 * 4 + i;
                float4 h, l;
                split4(r < nrows ? xr[0][i] : f4(0.f, 0.f, 0.f, 0.f), h, l);
                const int o = XP::off(r, j8);
                st4(X_hi + o, h);
                st4(X_lo + o, l);
            }
            if constexpr (CS) {              // lanes (lane & 7) = piece j8, (lane >> 3) = the warp's four row groups
                float4 cs = f4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (g8 * 4 + i < nrows) cs = cs + xr[0][i];
#pragma unroll
                for (int off = 8; off <= 16; off <<= 1) {
                    cs.x += __shfl_xor_sync(0xffffffffu, cs.x, off);
                    cs.y += __shfl_xor_sync(0xffffffffu, cs.y, off);
                    cs.z += __shfl_xor_sync(0xffffffffu, cs.z, off);
                    cs.w += __shfl_xor_sync(0xffffffffu, cs.w, off);
                }
                if ((tid & 31) < 8) {
                    float* d = &cs_s[warp][c * KCH + j8 * 4];
                    d[0] += cs.x; d[1] += cs.y; d[2] += cs.z; d[3] += cs.w;
                }
            }
#pragma unroll
            for (int k = 0; k + 1 < PF; ++k) {
#pragma unroll
                for (int i = 0; i < 4; ++i) xr[k][i] = xr[k + 1][i];
            }
            load_next(xr[PF - 1]);
            if (tid == 0 && gc + 1 < total) {  // next chunk's weights -> the other buffer, free once chunk gc - 1's MMAs are done
                if (gc >= 1) mbar_wait(&bar_m[b ^ 1], (uint32_t)(((gc - 1) >> 1) & 1));
                issue_w(gc + 1);
            }
            fence_async_smem();
            fence_before();
            __syncthreads();
            fence_after();
            if (tid == 0) {
                mbar_wait(&bar_w[b], ph);
                const float* W_hi = Wb + b * W_BUF;
                const float* W_lo = W_hi + W_PART;
                issue_data_gemm_row32<TE, NO, KCH>(tm, smem_u32(X_lo), smem_u32(W_hi), c > 0);
                issue_data_gemm_row32<TE, NO, KCH>(tm, smem_u32(X_hi), smem_u32(W_lo), 1);
                issue_data_gemm_row32<TE, NO, KCH>(tm, smem_u32(X_hi), smem_u32(W_hi), 1);
                mma_commit(&bar_m[b]);
            }
        }
        // ---- epilogue: all MMAs of the tile done (the last two commits cover both X buffers)
        {
            const int64_t last = gc - 1;
            mbar_wait(&bar_m[last & 1], (uint32_t)((last >> 1) & 1));
            if (last >= 1) mbar_wait(&bar_m[(last - 1) & 1], (uint32_t)(((last - 1) >> 1) & 1));
            fence_after();
        }
        float* stage = Xb;
#pragma unroll 1
        for (int grp = 0; grp < NO / 64; ++grp) {
            float4 e_a0[8], e_a1[8];             // EPI: the slots' rows of this column group, in flight while the accumulator is staged
            if constexpr (EPI) {
                const int co = grp * 64 + j * 4;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (ep.s0.role) e_a0[i] = ld_e4(ep.s0.p + (int64_t)e_r0[i] * NO + co);
                    if (ep.s1.role) e_a1[i] = ld_e4(ep.s1.p + (int64_t)e_r1[i] * NO + co);
                }
            }
            stage_from_tmem<NT2>(tm + lane_off + grp * 64 + ch * 32, stage, t, ch);
            __syncthreads();
            const float4 b4 = a.bias ? ldg4(a.bias + grp * 64 + j * 4) : f4(0.f, 0.f, 0.f, 0.f);
            if constexpr (EPI) {
                const int co = grp * 64 + j * 4;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = g * 8 + i;
                    float4 v = lds4(stage + stage_off(r, j)) + b4;
                    if (ep.s0.role == 1) v = v + e_a0[i];
                    if (ep.s1.role == 1) v = v + e_a1[i];
                    if (ep.relu) v = relu4(v);
                    if (ep.s0.role == 2) v = f4(e_a0[i].x > 0.f ? v.x : 0.f, e_a0[i].y > 0.f ? v.y : 0.f, e_a0[i].z > 0.f ? v.z : 0.f, e_a0[i].w > 0.f ? v.w : 0.f);
                    if (ep.s1.role == 2) v = f4(e_a1[i].x > 0.f ? v.x : 0.f, e_a1[i].y > 0.f ? v.y : 0.f, e_a1[i].z > 0.f ? v.z : 0.f, e_a1[i].w > 0.f ? v.w : 0.f);
                    v = ep.beta * v;
                    if (ep.s0.role == 3) v = v + ep.s0.coef * e_a0[i];
                    if (ep.s1.role == 3) v = v + ep.s1.coef * e_a1[i];
                    if (r < nrows) st4(a.Y + (int64_t)e_or[i] * NO + co, v);
                }
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int r = g * 8 + i;
                    float* yo = a.ysplit ? a.Y + ((int64_t)grp * a.N + row0 + r) * 64 : a.Y + (row0 + r) * a.ldy + grp * 64;
                    if (r < nrows) st4(yo + j * 4, lds4(stage + stage_off(r, j)) + b4);
                }
            }
            __syncthreads();
        }
        fence_before();
        __syncthreads();
        fence_after();
    }
    fence_before();
    __syncthreads();
    if constexpr (CS) {
        if (tid < a.K) {
            float sum = 0.f;
#pragma unroll
            for (int w = 0; w < NT2 / 32; ++w) sum += cs_s[w][tid];
            a.colsum_part[(int64_t)blockIdx.x * a.K + tid] = sum;
        }
    }
    if (warp == 0) tmem_dealloc(tm, NCOLS);
}

// the CS variant: NO = 64 (the data gradient of a group of [*, 64] -> [*, 64] linears), at most n_part CTAs
static int launch_node_linear_colsum(const NodeLinArgs& a, int n_part, cudaStream_t st) {
    constexpr int NO = 64;
    const size_t smem = 4 * XP::BYTES + 2 * (2 * NO * KCH * 4) + 1024;
    auto kern = node_linear_umma_kernel<NO, false, true>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.N + TE - 1) / TE;
    int grid = persistent_grid((const void*)kern, smem, ntiles, NT2);
    if (grid > n_part) grid = n_part;
    kern<<<grid, NT2, smem, st>>>(a, NoEpi{});
    B200MP_LAUNCH_CHECK();
    return 0;
}

template <int NO>
static int launch_node_linear(const NodeLinArgs& a, cudaStream_t st) {
    const size_t smem = 4 * XP::BYTES + 2 * (2 * NO * KCH * 4) + 1024;    // X panels and weight chunks, double buffered
    auto kern = node_linear_umma_kernel<NO, false>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.N + TE - 1) / TE;
    const int grid = persistent_grid((const void*)kern, smem, ntiles, NT2);
    kern<<<grid, NT2, smem, st>>>(a, NoEpi{});
    B200MP_LAUNCH_CHECK();
    return 0;
}
template <int NO>
static int launch_row_linear(const NodeLinArgs& a, const RowEpi& ep, cudaStream_t st) {
    const size_t smem = 4 * XP::BYTES + 2 * (2 * NO * KCH * 4) + 1024;
    auto kern = node_linear_umma_kernel<NO, true>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.N + TE - 1) / TE;
    const int grid = persistent_grid((const void*)kern, smem, ntiles, NT2);
    kern<<<grid, NT2, smem, st>>>(a, ep);
    B200MP_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ weight gradient
// G[K, NO] = X[N, K]^T Y[N, NO]: CTA (bx, kb) accumulates rows 128 kb .. 128 kb + 127 of G over the
// 32-row tiles bx, bx + gridDim.x, ... in tensor memory (M = 128, N = NO) and writes one partial
// [128 x NO] at the end (summed by the caller: deterministic).  Both operands are read through the
// MN-major views of row32 panels (umma.cuh), the reduction runs over the node rows of the tile.
namespace ng {
constexpr int RW = 32;                    // node rows per tile
constexpr uint32_t WG_DRAIN_TILES = 16;   // 512 rows between drains of the weight-gradient accumulator (1024: 1.06e-5 of max|gW| over 300 k PNA gradient rows)
using WP = RowFmt<RW, 32>;                // 4 KB panel [32 rows x 32 columns]
struct NodeWgArgs {
    int64_t N; int ldx; int ldy;
    const float* X; const float* Y; float* part;   // part: [gridDim.x][K][NO]
    int K;
    const float* X1;                               // pair mode (K = 128): columns 64..127 of X come from X1, both at ldx
    const int32_t* xrow; const int32_t* yrow;      // storage rows of X / Y for reduction row r (null = r)
    int single;                                    // pair mode with ONE gradient: columns 64..127 of X are zeros (not read)
};
}  // namespace ng

template <int NO>
__global__ void __launch_bounds__(NT2, 2) node_wgrad_umma_kernel(NodeWgArgs a) {
    constexpr int XPN = 4, YPN = NO / 32;                  // panels per tile: X part (128 columns), Y
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;
    float* X_hi = smem;
    float* X_lo = X_hi + XPN * WP::FLOATS;
    float* Y_hi = X_lo + XPN * WP::FLOATS;
    float* Y_lo = Y_hi + YPN * WP::FLOATS;
    __shared__ uint64_t bar_m;
    __shared__ uint32_t tmem_base;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int piece = tid & 7, row = tid >> 3;             // 16-byte piece of one 128-byte panel row
    const int kb = blockIdx.y;
    if (warp == 0) tmem_alloc(&tmem_base, 256);
    if (tid == 0) { mbar_init(&bar_m, 1); mbar_fence_init(); }
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = tmem_base;
    const int64_t ntiles = (a.N + RW - 1) / RW;
    uint32_t it = 0;

    float4 xr[XPN], yr[YPN];
    bool ok_pf = false;
    auto prefetch = [&](int64_t tile_) {
        const int64_t r = tile_ * RW + row;
        ok_pf = r < a.N;
        const int64_t rc = ok_pf ? r : a.N - 1;
        const int64_t rx = a.xrow ? (int64_t)__ldg(a.xrow + rc) : rc, ry = a.yrow ? (int64_t)__ldg(a.yrow + rc) : rc;
#pragma unroll
        for (int p = 0; p < XPN; ++p)
            xr[p] = (a.single && p >= 2) ? f4(0.f, 0.f, 0.f, 0.f)
                  : ldg4((a.X1 && p >= 2) ? a.X1 + rx * a.ldx + (p - 2) * 32 + piece * 4
                                          : a.X + rx * a.ldx + kb * 128 + p * 32 + piece * 4);
#pragma unroll
        for (int p = 0; p < YPN; ++p) yr[p] = ldg4(a.Y + ry * a.ldy + p * 32 + piece * 4);
    };
    if ((int64_t)blockIdx.x < ntiles) prefetch(blockIdx.x);
    // accumulator [128 x NO] -> this CTA's partial (thread (t, ch) = row 128 kb + t, columns 32 ch + 64 q ..): at the end and every
    // WG_DRAIN_TILES tiles in between -- a sum kept in one tensor-memory accumulator loses accuracy linearly in the number of MMAs
    // added into it (umma_edge.cuh: WGRAD_FLUSH_EDGES has the measurement); `valid` false writes zeros, `accumulate` adds to the
    // partial an earlier drain left
    auto drain = [&](bool valid, bool accumulate) {
        const int t = tid & (TE - 1), ch = tid >> 7;
        float* out = a.part + ((int64_t)blockIdx.x * a.K + kb * 128 + t) * NO;
        const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
#pragma unroll 1
        for (int c0 = ch * 16; c0 < NO; c0 += 32) {
            float v[16];
            if (valid) tmem_ld16(tm + lane_off + c0, v);
#pragma unroll
            for (int c = 0; c < 16; c += 4) {
                float4 w = valid ? f4(v[c], v[c + 1], v[c + 2], v[c + 3]) : f4(0.f, 0.f, 0.f, 0.f);
                if (accumulate) red_add4(out + c0 + c, w); else st4(out + c0 + c, w);     // single writer per address: deterministic
            }
        }
    };

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int o = WP::off(row, piece);
        const bool ok = ok_pf;
#pragma unroll
        for (int p = 0; p < XPN; ++p) {
            float4 h, l;
            split4(ok ? xr[p] : f4(0.f, 0.f, 0.f, 0.f), h, l);
            st4(X_hi + p * WP::FLOATS + o, h);
            st4(X_lo + p * WP::FLOATS + o, l);
        }
#pragma unroll
        for (int p = 0; p < YPN; ++p) {
            float4 h, l;
            split4(ok ? yr[p] : f4(0.f, 0.f, 0.f, 0.f), h, l);
            st4(Y_hi + p * WP::FLOATS + o, h);
            st4(Y_lo + p * WP::FLOATS + o, l);
        }
        if (tile + gridDim.x < ntiles) prefetch(tile + gridDim.x);
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            issue_wgrad_gemm_row32<RW, 128, NO>(tm, smem_u32(X_lo), smem_u32(Y_hi), (it % WG_DRAIN_TILES) != 0);
            issue_wgrad_gemm_row32<RW, 128, NO>(tm, smem_u32(X_hi), smem_u32(Y_lo), 1);
            issue_wgrad_gemm_row32<RW, 128, NO>(tm, smem_u32(X_hi), smem_u32(Y_hi), 1);
            mma_commit(&bar_m);
        }
        mbar_wait(&bar_m, it & 1);
        fence_after();
        if ((it % WG_DRAIN_TILES) == WG_DRAIN_TILES - 1) drain(true, it >= WG_DRAIN_TILES);
    }
    if ((it % WG_DRAIN_TILES) != 0 || it < WG_DRAIN_TILES) drain((it % WG_DRAIN_TILES) != 0, it >= WG_DRAIN_TILES);
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 256);
}

template <int NO>
static int launch_node_wgrad(const NodeWgArgs& a, int n_part, cudaStream_t st) {
    const size_t smem = 2 * (4 + NO / 32) * WP::BYTES + 1024;
    auto kern = node_wgrad_umma_kernel<NO>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<dim3(n_part, a.K / 128), NT2, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200mp

extern "C" int b200mp_node_linear_ld(const float* X, int64_t N, int K, int ldx, const float* Wimg, int NO, const float* bias,
                                     float* Y, int ldy, void* stream) {
    using namespace b200mp;
    if (N < 0 || K <= 0 || K % 32 || ldx < K || ldx % 4 || ldy < NO || ldy % 4 || !X || !Wimg || !Y) return B200MP_EINVAL;
    if (N == 0) return 0;
    NodeLinArgs a{N, K, ldx, X, Wimg, bias, Y, nullptr, nullptr, 0, ldy, nullptr};
    cudaStream_t st = (cudaStream_t)stream;
    switch (NO) {
        case 64: return launch_node_linear<64>(a, st);
        case 128: return launch_node_linear<128>(a, st);
        case 192: return launch_node_linear<192>(a, st);
        case 256: return launch_node_linear<256>(a, st);
    }
    return B200MP_EINVAL;
}

extern "C" int b200mp_node_linear(const float* X, int64_t N, int K, int ldx, const float* Wimg, int NO, const float* bias,
                                  float* Y, void* stream) {
    return b200mp_node_linear_ld(X, N, K, ldx, Wimg, NO, bias, Y, NO, stream);
}

extern "C" int b200mp_node_linear_wgrad_rows(const float* X, const int32_t* xrow, int ldx, int K, const float* Y, const int32_t* yrow,
                                             int ldy, int NO, int64_t N, float* part, int n_part, void* stream) {
    using namespace b200mp;
    if (N <= 0 || K <= 0 || K % 128 || ldx < K || ldx % 4 || ldy < NO || ldy % 4 || n_part <= 0 || !X || !Y || !part)
        return B200MP_EINVAL;
    NodeWgArgs a{N, ldx, ldy, X, Y, part, K, nullptr, xrow, yrow, 0};
    cudaStream_t st = (cudaStream_t)stream;
    switch (NO) {
        case 64: return launch_node_wgrad<64>(a, n_part, st);
        case 128: return launch_node_wgrad<128>(a, n_part, st);
        case 192: return launch_node_wgrad<192>(a, n_part, st);
        case 256: return launch_node_wgrad<256>(a, n_part, st);
    }
    return B200MP_EINVAL;
}

extern "C" int b200mp_node_linear_wgrad(const float* X, int ldx, int K, const float* Y, int ldy, int NO, int64_t N,
                                        float* part, int n_part, void* stream) {
    return b200mp_node_linear_wgrad_rows(X, nullptr, ldx, K, Y, nullptr, ldy, NO, N, part, n_part, stream);
}

/* Weight gradients of two H = 64 node linears that share their input: part[p][0:64] = G0^T Y,
 * part[p][64:128] = G1^T Y over the rows handled by CTA p (G0, G1, Y: [N, 64] contiguous). */
extern "C" int b200mp_node_pair_wgrad(const float* G0, const float* G1, const float* Y, int64_t N, float* part, int n_part,
                                      void* stream) {
    using namespace b200mp;
    if (N <= 0 || n_part <= 0 || !G0 || !Y || !part) return B200MP_EINVAL;
    NodeWgArgs a{N, 64, 64, G0, Y, part, 128, G1 ? G1 : G0, nullptr, nullptr, G1 ? 0 : 1};   // G1 == NULL: part[p][64:128] = 0
    return launch_node_wgrad<64>(a, n_part, (cudaStream_t)stream);
}

/* H = 64 node linears sharing their rows: X = [X0 | X1 | X2] (nx blocks of [N,64]), Y = X W^T + bias with
 * NO in {64,128,192,256}; ysplit: Y as NO/64 separate [N,64] blocks. */
extern "C" int b200mp_node_linear64(const float* X0, const float* X1, const float* X2, int nx, int64_t N, const float* Wimg,
                                    int NO, const float* bias, float* Y, int ysplit, void* stream) {
    using namespace b200mp;
    if (N < 0 || nx < 1 || nx > 3 || !X0 || (nx > 1 && !X1) || (nx > 2 && !X2) || !Wimg || !Y) return B200MP_EINVAL;
    if (N == 0) return 0;
    NodeLinArgs a{N, 64 * nx, 64, X0, Wimg, bias, Y, nx > 1 ? X1 : X0, nx > 2 ? X2 : X0, ysplit, NO, nullptr};
    cudaStream_t st = (cudaStream_t)stream;
    switch (NO) {
        case 64: return launch_node_linear<64>(a, st);
        case 128: return launch_node_linear<128>(a, st);
        case 192: return launch_node_linear<192>(a, st);
        case 256: return launch_node_linear<256>(a, st);
    }
    return B200MP_EINVAL;
}

/* Y[orow[k]] = alpha R[rrow[k]] + beta act(X[xrow[k]] W^T + bias + T1[idx1[k]] + T2[idx2[k]]) * [mask[k] > 0] + R2[k] over N rows: the
 * per-edge linears of the message / edge-update forward and backward at widths the fused H = 64 kernels do not cover (H = 128:
 * pna.py:50 default).  Same kernel, tiles and operand image as b200mp_node_linear; K % 32 == 0, NO in {64,128,192,256}; T1 / T2 / mask
 * / R / R2 / Y rows are NO floats wide. */
extern "C" int b200mp_row_linear(const float* X, const int32_t* xrow, int64_t N, int K, int ldx, const float* Wimg, int NO,
                                 const float* bias, const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                                 int relu, const float* mask, const float* R, const int32_t* rrow, float alpha, float beta,
                                 const float* R2, float* Y, const int32_t* orow, void* stream) {
    using namespace b200mp;
    if (N < 0 || K <= 0 || K % 32 || ldx < K || ldx % 4 || !X || !Wimg || !Y || (T1 && !idx1) || (T2 && !idx2)) return B200MP_EINVAL;
    if (N == 0) return 0;
    NodeLinArgs a{N, K, ldx, X, Wimg, bias, Y, nullptr, nullptr, 0, NO, nullptr};
    EpiSlot slots[2] = {{nullptr, nullptr, 0, 0.f}, {nullptr, nullptr, 0, 0.f}};
    int ns = 0;
    auto add = [&](const float* p, const int32_t* idx, int role, float coef) {
        if (!p) return true;
        if (ns == 2) return false;
        slots[ns++] = EpiSlot{p, idx, role, coef};
        return true;
    };
    if (!add(T1, idx1, 1, 1.f) || !add(T2, idx2, 1, 1.f) || !add(mask, nullptr, 2, 1.f) || !add(R, rrow, 3, alpha) || !add(R2, nullptr, 3, 1.f))
        return B200MP_EINVAL;                 // at most two of T1 / T2 / mask / R / R2 per call
    RowEpi ep{xrow, slots[0], slots[1], orow, beta, relu};
    cudaStream_t st = (cudaStream_t)stream;
    switch (NO) {
        case 64: return launch_row_linear<64>(a, ep, st);
        case 128: return launch_row_linear<128>(a, ep, st);
        case 192: return launch_row_linear<192>(a, ep, st);
        case 256: return launch_row_linear<256>(a, ep, st);
    }
    return B200MP_EINVAL;
}

/* b200mp_node_linear64 with NO = 64 that also returns the column sums of X = [X0 | X1 | X2]: colsum_part [n_part][64 nx], row p = the
 * sums over the rows CTA p handled (rows p >= the launched grid are left untouched: zero them first); the caller adds the rows up.
 * With X the gradients of a group of linears this pass yields their data gradient AND their bias gradients. */
extern "C" int b200mp_node_linear64_colsum(const float* X0, const float* X1, const float* X2, int nx, int64_t N, const float* Wimg,
                                           const float* bias, float* Y, float* colsum_part, int n_part, void* stream) {
    using namespace b200mp;
    if (N <= 0 || nx < 1 || nx > 3 || !X0 || (nx > 1 && !X1) || (nx > 2 && !X2) || !Wimg || !Y || !colsum_part || n_part < 1)
        return B200MP_EINVAL;
    NodeLinArgs a{N, 64 * nx, 64, X0, Wimg, bias, Y, nx > 1 ? X1 : X0, nx > 2 ? X2 : X0, 0, 64, colsum_part};
    return launch_node_linear_colsum(a, n_part, (cudaStream_t)stream);
}

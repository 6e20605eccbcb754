// Tensor-core (tcgen05, 3xTF32) per-edge linear forward for H = 64 (b200mp_edge_linear_fwd):
// PNA / GINE message (modes 0 / 1) and the edge update (mode 2).  Same contract as the SIMT
// kernels in edge_fwd.cu, which remain the path for H = 16 / 32 / 128.
//
//   z = e W1^T + b1 + T1[idx1] + T2[idx2]     Z (data GEMM, tmem cols 0..63); the gathered node rows
//                                             are fetched coalesced while the MMAs run
//   mode 0: out1 = z     mode 1: out1 = relu(z)
//   mode 2: out1 = h = relu(z);  out2[row] = alpha e + beta (h W2^T + b2)     U (tmem cols 64..127)
#include "umma_edge.cuh"

namespace b200mp {
using namespace ue;

struct EdgeFwdUArgs {
    int64_t E;
    const float* e; const int32_t* eid;
    const float* W1; const float* b1;
    const float* T1; const int32_t* idx1;
    const float* T2; const int32_t* idx2;
    const float* W2; const float* b2;
    float alpha, beta;
    float* out1; float* out2;
};

template <int MODE>
__global__ void __launch_bounds__(NTH, MODE == 2 ? 1 : 2) edge_linear_fwd_umma_kernel(EdgeFwdUArgs a) {
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;   // keeps the shared address space
    float* E_hi = smem;                       // e tile; mode 2: later h
    float* E_lo = E_hi + RT::FLOATS;          // modes 0/1: later the gathered node rows
    float* ST = E_lo + RT::FLOATS;            // mode 2 only: gathered node rows, later alpha e
    float* W1_hi = MODE == 2 ? ST + RT::FLOATS : ST;
    float* W1_lo = W1_hi + WT::FLOATS;
    float* W2_hi = W1_lo + WT::FLOATS;        // mode 2 only
    float* W2_lo = W2_hi + WT::FLOATS;
    __shared__ TileSync sync;
    __shared__ int sidx[2][TE];               // gather indices of the tile (clamped rows)

    const int tid = threadIdx.x, warp = tid >> 5;
    const int t = tid & (TE - 1), ch = tid >> 7;           // epilogue: row t, columns 32 ch ..
    constexpr uint32_t NCOLS = MODE == 2 ? 128 : 64;
    if (warp == 0) tmem_alloc(&sync.tmem_base, NCOLS);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_init(&sync.bar[1], 1); mbar_fence_init(); }
    load_weight_tiles<false>(W1_hi, W1_lo, a.W1, tid);
    if (MODE == 2) load_weight_tiles<false>(W2_hi, W2_lo, a.W2, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t T_Z = tm + lane_off + ch * 32, T_U = tm + 64 + lane_off + ch * 32;
    float* gath = MODE == 2 ? ST : E_lo;

    const int64_t ntiles = (a.E + TE - 1) / TE;
    uint32_t it = 0;
    RowRegs R;
    int nrows_pf = 0, idx_pf = 0;             // thread tid: index (tid / 128) of row tid % 128, one tile ahead
    const int32_t* idx_src = ch == 0 ? a.idx1 : a.idx2;
    if ((int64_t)blockIdx.x < ntiles) {
        nrows_pf = (int)min64(TE, a.E - (int64_t)blockIdx.x * TE);
        load_rows(R, a.e, a.eid, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        if (idx_src) idx_pf = __ldg(idx_src + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
    }

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE;
        const int nrows = nrows_pf;
        const uint32_t par = it & 1;

        store_rows_split(R, E_hi, E_lo, nrows, tid);
        sidx[ch][t] = idx_pf;
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, E_hi, E_lo, W1_hi, W1_lo, 0);
            mma_commit(&sync.bar[0]);
        }
        gather_rows(R, a.T1, sidx[0], a.T2, sidx[1], tid);   // coalesced, in flight while the MMAs run
        if (MODE != 2) {                     // E_lo is still an MMA operand: wait before reusing it
            mbar_wait(&sync.bar[0], par);
            fence_after();
        }
        store_rows_raw(R, gath, nrows, tid);
        if (tile + gridDim.x < ntiles) {     // next tile's e rows in flight
            nrows_pf = (int)min64(TE, a.E - (tile + gridDim.x) * TE);
            load_rows(R, a.e, a.eid, (tile + gridDim.x) * TE, nrows_pf, tid);
            if (idx_src) idx_pf = __ldg(idx_src + (tile + gridDim.x) * TE + min(t, nrows_pf - 1));
        }
        if (MODE == 2) {
            mbar_wait(&sync.bar[0], par);
            fence_after();
        }
        __syncthreads();                     // gathered rows visible to their row threads

        // ---- epilogue 1: thread (t, ch) = columns 32 ch .. 32 ch + 31 of edge row t
        {
            const bool valid = t < nrows;
            float* o1 = a.out1 + (row0 + t) * HH + ch * 32;
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                float z[16];
                tmem_ld16(T_Z + q * 16, z);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int o = RT::off(t, ch * 8 + q * 4 + j);
                    const int c = q * 16 + j * 4;
                    float4 v = f4(z[4 * j], z[4 * j + 1], z[4 * j + 2], z[4 * j + 3]) + ldg4(a.b1 + ch * 32 + c) + lds4(gath + o);
                    if (MODE >= 1) v = relu4(v);
                    if (!valid) v = f4(0.f, 0.f, 0.f, 0.f);
                    if (valid) st4(o1 + c, v);
                    if (MODE == 2) {
                        const float4 ev = lds4(E_hi + o) + lds4(E_lo + o);
                        st4(ST + o, a.alpha * ev);
                        float4 hh, hl;
                        split4(v, hh, hl);
                        st4(E_hi + o, hh);
                        st4(E_lo + o, hl);
                    }
                }
            }
        }
        if (MODE == 2) {
            fence_async_smem();
            fence_before();
            __syncthreads();
            fence_after();
            if (tid == 0) {
                gemm3_data(tm + 64, E_hi, E_lo, W2_hi, W2_lo, 0);
                mma_commit(&sync.bar[1]);
            }
            int64_t row = row0 + t;
            if (a.eid && t < nrows) row = __ldg(a.eid + row0 + t);
            float* o2 = a.out2 + row * HH + ch * 32;
            mbar_wait(&sync.bar[1], par);
            fence_after();
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                float u[16];
                tmem_ld16(T_U + q * 16, u);
                if (t < nrows) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int c = q * 16 + j * 4;
                        const float4 r4 = lds4(ST + RT::off(t, ch * 8 + q * 4 + j)) +
                                          a.beta * (f4(u[4 * j], u[4 * j + 1], u[4 * j + 2], u[4 * j + 3]) + ldg4(a.b2 + ch * 32 + c));
                        st4(o2 + c, r4);
                    }
                }
            }
        }
        fence_before();
        __syncthreads();
        fence_after();
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, NCOLS);
}

template <int MODE>
static int launch_fwd_umma_mode(const EdgeFwdUArgs& a, cudaStream_t st) {
    const size_t smem = (MODE == 2 ? 3 * RT::BYTES + 4 * WT::BYTES : 2 * RT::BYTES + 2 * WT::BYTES) + 1024;
    auto kern = edge_linear_fwd_umma_kernel<MODE>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.E + TE - 1) / TE;
    int grid = persistent_grid((const void*)kern, smem, ntiles, NTH);
    kern<<<grid, NTH, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

int launch_edge_fwd_umma(int mode, int64_t E, const float* e, const int32_t* eid, const float* W1, const float* b1,
                         const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                         const float* W2, const float* b2, float alpha, float beta, float* out1, float* out2, cudaStream_t st) {
    EdgeFwdUArgs a{E, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out1, out2};
    switch (mode) {
        case 0: return launch_fwd_umma_mode<0>(a, st);
        case 1: return launch_fwd_umma_mode<1>(a, st);
        case 2: return launch_fwd_umma_mode<2>(a, st);
    }
    return B200MP_EINVAL;
}

}  // namespace b200mp

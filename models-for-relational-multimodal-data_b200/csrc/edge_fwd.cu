// Per-edge linear forward (b200mp_edge_linear_fwd): PNA / GINE message and the edge update.
// Persistent CTAs stream tiles of TE consecutive edges (segment order): the e-tile is loaded
// coalesced into shared memory, multiplied with the H x H weight held in shared memory for the
// CTA's lifetime, and the epilogue adds the gathered node rows, applies the activation and
// (edge update) runs the second linear + residual before anything is written back.
//
// Follows: PyG PNAConv.message / GINEConv.message as constructed at
//   /root/reference/src/nn/gnn/pna.py:71-74, gine.py:66-70  and the edge-update MLP at
//   gine.py:71-75,88-89, pna.py:80-84,97-99, src/nn/models/fused.py:215-219,254.
#include "common.cuh"

namespace b200mp {

struct EdgeFwdArgs {
    int64_t E;
    const float* e; const int32_t* eid;
    const float* W1; const float* b1;
    const float* T1; const int32_t* idx1;
    const float* T2; const int32_t* idx2;
    const float* W2; const float* b2;
    float alpha, beta;
    float* out1; float* out2; const int32_t* out_eid;
};

template <int H, int MODE>
__global__ void __launch_bounds__(NT) edge_linear_fwd_kernel(EdgeFwdArgs a) {
    using T = Tile<H>;
    extern __shared__ __align__(16) float smem[];
    float* Ws1 = smem;
    float* Xs = Ws1 + T::WS;
    float* Ws2 = Xs + T::XS;           // MODE 2 only
    float* Hs = Ws2 + T::WS;           // MODE 2 only

    const int tx = threadIdx.x % T::TX, ty = threadIdx.x / T::TX;
    const int c = tx * 4;
    load_weight<H, true>(Ws1, a.W1);
    if (MODE == 2) load_weight<H, true>(Ws2, a.W2);
    const float4 bias1 = ldg4(a.b1 + c);
    float4 bias2 = f4(0.f, 0.f, 0.f, 0.f);
    if (MODE == 2) bias2 = ldg4(a.b2 + c);

    const int64_t ntiles = (a.E + T::TE - 1) / T::TE;
    for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int64_t row0 = t * T::TE;
        const int nrows = (int)min64(T::TE, a.E - row0);
        __syncthreads();                       // previous tile fully consumed (also covers the weight load)
        load_tile<H>(Xs, a.e, a.eid, row0, nrows);
        __syncthreads();

        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        tile_gemm<H>(Xs, Ws1, acc, tx, ty);

#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty * 4 + i;
            float4 v = f4(0.f, 0.f, 0.f, 0.f);
            if (r < nrows) {
                const int64_t k = row0 + r;
                v = f4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]) + bias1;
                if (a.T1) v = v + ldg4(a.T1 + (int64_t)__ldg(a.idx1 + k) * H + c);
                if (a.T2) v = v + ldg4(a.T2 + (int64_t)__ldg(a.idx2 + k) * H + c);
                if (MODE >= 1) v = relu4(v);
                st4(a.out1 + k * H + c, v);
            }
            if (MODE == 2) st4(Hs + r * T::LDX + c, v);
        }

        if (MODE == 2) {
            __syncthreads();
            float acc2[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc2[i][j] = 0.f;
            tile_gemm<H>(Hs, Ws2, acc2, tx, ty);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = ty * 4 + i;
                if (r < nrows) {
                    const int64_t k = row0 + r;
                    const int64_t row = a.out_eid ? (int64_t)__ldg(a.out_eid + k) : k;
                    float4 u = f4(acc2[i][0], acc2[i][1], acc2[i][2], acc2[i][3]) + bias2;
                    float4 x = lds4(Xs + r * T::LDX + c);
                    st4(a.out2 + row * H + c, a.alpha * x + a.beta * u);
                }
            }
        }
    }
}

template <int H, int MODE>
static int launch_fwd(const EdgeFwdArgs& a, cudaStream_t st) {
    using T = Tile<H>;
    size_t smem = sizeof(float) * (size_t)(MODE == 2 ? 2 * (T::WS + T::XS) : (T::WS + T::XS));
    auto kern = edge_linear_fwd_kernel<H, MODE>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t ntiles = (a.E + T::TE - 1) / T::TE;
    int grid = persistent_grid((const void*)kern, smem, ntiles);
    kern<<<grid, NT, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

template <int H>
static int dispatch_mode(int mode, const EdgeFwdArgs& a, cudaStream_t st) {
    switch (mode) {
        case 0: return launch_fwd<H, 0>(a, st);
        case 1: return launch_fwd<H, 1>(a, st);
        case 2: return launch_fwd<H, 2>(a, st);
    }
    return B200MP_EINVAL;
}

int launch_edge_fwd_umma(int mode, int64_t E, const float* e, const int32_t* eid, const float* W1, const float* b1,
                         const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                         const float* W2, const float* b2, float alpha, float beta, float* out1, float* out2,
                         const int32_t* out_eid, cudaStream_t st);

}  // namespace b200mp

extern "C" int b200mp_edge_linear_fwd(int mode, int H, int64_t E,
                                      const float* e, const int32_t* eid,
                                      const float* W1, const float* b1,
                                      const float* T1, const int32_t* idx1,
                                      const float* T2, const int32_t* idx2,
                                      const float* W2, const float* b2, float alpha, float beta,
                                      float* out1, float* out2, const int32_t* out_eid, void* stream) {
    using namespace b200mp;
    if (E < 0 || mode < 0 || mode > 2) return B200MP_EINVAL;
    if (E == 0) return 0;
    if (!W1 || !b1 || !out1 || !e) return B200MP_EINVAL;
    if ((T1 && !idx1) || (T2 && !idx2)) return B200MP_EINVAL;
    if (mode == 2 && (!W2 || !b2 || !out2)) return B200MP_EINVAL;
    EdgeFwdArgs a{E, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out1, out2, out_eid};
    cudaStream_t st = (cudaStream_t)stream;
    if (H == 64 && use_tensor_cores())
        return launch_edge_fwd_umma(mode, E, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out1, out2, out_eid, st);
    switch (H) {
        case 16: return dispatch_mode<16>(mode, a, st);
        case 32: return dispatch_mode<32>(mode, a, st);
        case 64: return dispatch_mode<64>(mode, a, st);
        case 128: return dispatch_mode<128>(mode, a, st);
    }
    return B200MP_EINVAL;
}

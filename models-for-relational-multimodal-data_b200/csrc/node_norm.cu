// BatchNorm + ReLU + residual average over node rows, the tail of every backbone layer of the reference:
//     x' = (x + relu(BatchNorm(c))) / 2          (/root/reference/src/nn/gnn/gine.py:87, pna.py:96, training mode)
// with c = the conv output [N, H], batch statistics over the N rows (biased variance, torch semantics).  torch runs
// this as statistics + transform + relu + add + div (8 passes over [N, H]) and five more kernels backward; here:
//   forward   pass 1  column sums of c and c^2 (per-thread fp32 partials over <= a few hundred rows, fp64 across CTAs)
//             pass 2  read c, x -> write x'
//   backward  pass 1  column sums of gr and gr * chat,  gr = g / 2 where BatchNorm(c) > 0   (= the beta / gamma gradients)
//             pass 2  read g, c -> write g_c = gamma rstd (gr - mean(gr) - chat mean(gr chat)) and g_x = g / 2
// H / 4 lanes per row, 16-byte accesses, grids sized to the SM count.
#include "common.cuh"

namespace b200mp {

struct NodeNormArgs {
    int64_t N; int H;
    const float* c; const float* x;
    const float* gamma; const float* beta;
    const float* mean_rstd;            // [2, H]
};

__device__ __forceinline__ float4 mul4n(float4 a, float4 b) { return f4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }

// CTA-wide column reduction of two per-thread float4 partials into fp64 global sums [2, H]
template <int LPR>
__device__ __forceinline__ void cta_column_reduce(float4 a0, float4 a1, double* __restrict__ sums, int H) {
    __shared__ float4 red[2][256];
    red[0][threadIdx.x] = a0;
    red[1][threadIdx.x] = a1;
    __syncthreads();
    if (threadIdx.x < LPR) {
        double s0[4] = {0, 0, 0, 0}, s1[4] = {0, 0, 0, 0};
        for (int k = threadIdx.x; k < 256; k += LPR) {
            const float4 u = red[0][k], v = red[1][k];
            s0[0] += u.x; s0[1] += u.y; s0[2] += u.z; s0[3] += u.w;
            s1[0] += v.x; s1[1] += v.y; s1[2] += v.z; s1[3] += v.w;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            atomicAdd(sums + threadIdx.x * 4 + i, s0[i]);
            atomicAdd(sums + H + threadIdx.x * 4 + i, s1[i]);
        }
    }
}

template <int LPR>
__global__ void __launch_bounds__(256) bn_stats_kernel(NodeNormArgs a, double* __restrict__ sums) {
    constexpr int RPC = 256 / LPR;
    const int c4 = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    float4 s = f4(0.f, 0.f, 0.f, 0.f), q = s;
    for (int64_t r = (int64_t)blockIdx.x * RPC + rl; r < a.N; r += (int64_t)gridDim.x * RPC) {
        const float4 v = ldg4(a.c + r * a.H + c4 * 4);
        s = s + v;
        q = q + mul4n(v, v);
    }
    cta_column_reduce<LPR>(s, q, sums, a.H);
}

template <int LPR>
__global__ void __launch_bounds__(256) bn_relu_res_fwd_kernel(NodeNormArgs a, float* __restrict__ out) {
    constexpr int RPC = 256 / LPR;
    const int c4 = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    const float4 mean = ldg4(a.mean_rstd + c4 * 4), rstd = ldg4(a.mean_rstd + a.H + c4 * 4);
    const float4 sc = mul4n(rstd, ldg4(a.gamma + c4 * 4)), bt = ldg4(a.beta + c4 * 4);
    for (int64_t r = (int64_t)blockIdx.x * RPC + rl; r < a.N; r += (int64_t)gridDim.x * RPC) {
        const float4 v = ldg4(a.c + r * a.H + c4 * 4), xv = ldg4(a.x + r * a.H + c4 * 4);
        const float4 y = relu4(f4((v.x - mean.x) * sc.x + bt.x, (v.y - mean.y) * sc.y + bt.y, (v.z - mean.z) * sc.z + bt.z,
                                  (v.w - mean.w) * sc.w + bt.w));
        st4(out + r * a.H + c4 * 4, f4((xv.x + y.x) * 0.5f, (xv.y + y.y) * 0.5f, (xv.z + y.z) * 0.5f, (xv.w + y.w) * 0.5f));
    }
}

// gr = 0.5 g [BatchNorm(c) > 0], chat = (c - mean) rstd
__device__ __forceinline__ void bn_bwd_terms(float4 g, float4 v, float4 mean, float4 rstd, float4 gm, float4 bt, float4& gr, float4& ch) {
    ch = f4((v.x - mean.x) * rstd.x, (v.y - mean.y) * rstd.y, (v.z - mean.z) * rstd.z, (v.w - mean.w) * rstd.w);
    const float4 sc = mul4n(rstd, gm);                       // the forward's expression, so the mask is the forward's
    gr = f4((v.x - mean.x) * sc.x + bt.x > 0.f ? 0.5f * g.x : 0.f, (v.y - mean.y) * sc.y + bt.y > 0.f ? 0.5f * g.y : 0.f,
            (v.z - mean.z) * sc.z + bt.z > 0.f ? 0.5f * g.z : 0.f, (v.w - mean.w) * sc.w + bt.w > 0.f ? 0.5f * g.w : 0.f);
}

template <int LPR>
__global__ void __launch_bounds__(256) bn_relu_res_bwd_reduce_kernel(NodeNormArgs a, const float* __restrict__ g,
                                                                     double* __restrict__ sums) {
    constexpr int RPC = 256 / LPR;
    const int c4 = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    const float4 mean = ldg4(a.mean_rstd + c4 * 4), rstd = ldg4(a.mean_rstd + a.H + c4 * 4);
    const float4 gm = ldg4(a.gamma + c4 * 4), bt = ldg4(a.beta + c4 * 4);
    float4 s = f4(0.f, 0.f, 0.f, 0.f), q = s;
    for (int64_t r = (int64_t)blockIdx.x * RPC + rl; r < a.N; r += (int64_t)gridDim.x * RPC) {
        float4 gr, ch;
        bn_bwd_terms(ldg4(g + r * a.H + c4 * 4), ldg4(a.c + r * a.H + c4 * 4), mean, rstd, gm, bt, gr, ch);
        s = s + gr;
        q = q + mul4n(gr, ch);
    }
    cta_column_reduce<LPR>(s, q, sums, a.H);
}

template <int LPR>
__global__ void __launch_bounds__(256) bn_relu_res_bwd_apply_kernel(NodeNormArgs a, const float* __restrict__ g,
                                                                    const float* __restrict__ gsum,   // [2, H]: mean(gr), mean(gr chat)
                                                                    float* __restrict__ gc, float* __restrict__ gx) {
    constexpr int RPC = 256 / LPR;
    const int c4 = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    const float4 mean = ldg4(a.mean_rstd + c4 * 4), rstd = ldg4(a.mean_rstd + a.H + c4 * 4);
    const float4 gm = ldg4(a.gamma + c4 * 4), bt = ldg4(a.beta + c4 * 4);
    const float4 m1 = ldg4(gsum + c4 * 4), m2 = ldg4(gsum + a.H + c4 * 4);
    const float4 k = mul4n(gm, rstd);
    for (int64_t r = (int64_t)blockIdx.x * RPC + rl; r < a.N; r += (int64_t)gridDim.x * RPC) {
        const float4 gv = ldg4(g + r * a.H + c4 * 4);
        float4 gr, ch;
        bn_bwd_terms(gv, ldg4(a.c + r * a.H + c4 * 4), mean, rstd, gm, bt, gr, ch);
        st4(gc + r * a.H + c4 * 4, f4(k.x * (gr.x - m1.x - ch.x * m2.x), k.y * (gr.y - m1.y - ch.y * m2.y),
                                      k.z * (gr.z - m1.z - ch.z * m2.z), k.w * (gr.w - m1.w - ch.w * m2.w)));
        st4(gx + r * a.H + c4 * 4, f4(0.5f * gv.x, 0.5f * gv.y, 0.5f * gv.z, 0.5f * gv.w));
    }
}

static int node_norm_grid(int64_t N, int lpr) {
    const int64_t iters = (N + 256 / lpr - 1) / (256 / lpr);
    const int64_t cap = (int64_t)device_info().sms * 8;
    return (int)(iters < 1 ? 1 : iters < cap ? iters : cap);
}
static bool node_norm_ok(int64_t N, int H) { return N >= 0 && (H == 16 || H == 32 || H == 64 || H == 128); }

#define B200MP_NN_DISPATCH(KERNEL, ...)                                                   \
    switch (H / 4) {                                                                      \
        case 4: KERNEL<4><<<grid, 256, 0, st>>>(__VA_ARGS__); break;                      \
        case 8: KERNEL<8><<<grid, 256, 0, st>>>(__VA_ARGS__); break;                      \
        case 16: KERNEL<16><<<grid, 256, 0, st>>>(__VA_ARGS__); break;                    \
        default: KERNEL<32><<<grid, 256, 0, st>>>(__VA_ARGS__); break;                    \
    }

}  // namespace b200mp

extern "C" int b200mp_bn_stats(const float* c, int64_t N, int H, double* sums, void* stream) {
    using namespace b200mp;
    if (!node_norm_ok(N, H)) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!c || !sums) return B200MP_EINVAL;
    NodeNormArgs a{N, H, c, nullptr, nullptr, nullptr, nullptr};
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = node_norm_grid(N, H / 4);
    B200MP_NN_DISPATCH(bn_stats_kernel, a, sums)
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_bn_relu_res_fwd(const float* c, const float* x, int64_t N, int H, const float* gamma, const float* beta,
                                      const float* mean_rstd, float* out, void* stream) {
    using namespace b200mp;
    if (!node_norm_ok(N, H)) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!c || !x || !gamma || !beta || !mean_rstd || !out) return B200MP_EINVAL;
    NodeNormArgs a{N, H, c, x, gamma, beta, mean_rstd};
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = node_norm_grid(N, H / 4);
    B200MP_NN_DISPATCH(bn_relu_res_fwd_kernel, a, out)
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_bn_relu_res_bwd_reduce(const float* g, const float* c, int64_t N, int H, const float* gamma,
                                             const float* beta, const float* mean_rstd, double* sums, void* stream) {
    using namespace b200mp;
    if (!node_norm_ok(N, H)) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!g || !c || !gamma || !beta || !mean_rstd || !sums) return B200MP_EINVAL;
    NodeNormArgs a{N, H, c, nullptr, gamma, beta, mean_rstd};
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = node_norm_grid(N, H / 4);
    B200MP_NN_DISPATCH(bn_relu_res_bwd_reduce_kernel, a, g, sums)
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_bn_relu_res_bwd_apply(const float* g, const float* c, int64_t N, int H, const float* gamma,
                                            const float* beta, const float* mean_rstd, const float* gmeans, float* gc, float* gx,
                                            void* stream) {
    using namespace b200mp;
    if (!node_norm_ok(N, H)) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!g || !c || !gamma || !beta || !mean_rstd || !gmeans || !gc || !gx) return B200MP_EINVAL;
    NodeNormArgs a{N, H, c, nullptr, gamma, beta, mean_rstd};
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = node_norm_grid(N, H / 4);
    B200MP_NN_DISPATCH(bn_relu_res_bwd_apply_kernel, a, g, gmeans, gc, gx)
    B200MP_LAUNCH_CHECK();
    return 0;
}

// Per-edge backward kernels: b200mp_message_bwd (PNA / GINE message + aggregation backward)
// and b200mp_edge_update_bwd.  Same persistent tiling as the forward.  Per tile a CTA
//   - forms the per-edge gradient G in registers (from the saved message and node-level
//     coefficient rows, which are consecutive in segment order and therefore cache-resident),
//   - reduces it into the gathered node table with 16-byte vector REDs,
//   - multiplies it with the weight held in shared memory (data gradient) and
//   - accumulates G^T e into a register-resident weight gradient that is written once per CTA
//     as a partial (summed by the caller: deterministic, no atomics on parameters).
//
// Gradient rules restated from torch autograd over the PyG operators:
// min/max gradients are split evenly between ties; std has zero gradient where clamped/masked.
#include "common.cuh"

namespace b200mp {

// ------------------------------------------------------------------ message backward
struct MsgBwdArgs {
    int64_t E;
    const float* m; const float* e; const int32_t* eid;
    const float* node_in; const int32_t* seg; const int32_t* nbr;
    const float* W; float* g_e; int accumulate; float* gT1;
    float* gW_part; float* gb_part;
};

template <int H, int KIND>
__global__ void __launch_bounds__(NT) message_bwd_kernel(MsgBwdArgs a) {
    using T = Tile<H>;
    constexpr int MT = T::MT;
    extern __shared__ __align__(16) float smem[];
    float* Ws = smem;              // Ws[c][j] = W[c][j]  (G W)
    float* Xs = Ws + T::WS;        // e tile
    float* Gs = Xs + T::XS;        // G tile

    const int tx = threadIdx.x % T::TX, ty = threadIdx.x / T::TX;
    const int c = tx * 4;
    load_weight<H, false>(Ws, a.W);

    float accW[MT][MT];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < MT; ++j) accW[i][j] = 0.f;
    float accb = 0.f;

    const int64_t ntiles = (a.E + T::TE - 1) / T::TE;
    for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int64_t row0 = t * T::TE;
        const int nrows = (int)min64(T::TE, a.E - row0);
        __syncthreads();
        load_tile<H>(Xs, a.e, a.eid, row0, nrows);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty * 4 + i;
            float4 G = f4(0.f, 0.f, 0.f, 0.f);
            if (r < nrows) {
                const int64_t k = row0 + r;
                const float4 mv = ldg4(a.m + k * H + c);
                const int64_t d = __ldg(a.seg + k);
                if (KIND == 0) {
                    const float* q = a.node_in + d * 7 * H + c;
                    const float4 c0 = ldg4(q), c1 = ldg4(q + H), gmn = ldg4(q + 2 * H), mn = ldg4(q + 3 * H),
                                 gmx = ldg4(q + 4 * H), mx = ldg4(q + 5 * H), mu = ldg4(q + 6 * H);
                    G.x = fmaf(c1.x, mv.x - mu.x, c0.x) + (mv.x == mn.x ? gmn.x : 0.f) + (mv.x == mx.x ? gmx.x : 0.f);
                    G.y = fmaf(c1.y, mv.y - mu.y, c0.y) + (mv.y == mn.y ? gmn.y : 0.f) + (mv.y == mx.y ? gmx.y : 0.f);
                    G.z = fmaf(c1.z, mv.z - mu.z, c0.z) + (mv.z == mn.z ? gmn.z : 0.f) + (mv.z == mx.z ? gmx.z : 0.f);
                    G.w = fmaf(c1.w, mv.w - mu.w, c0.w) + (mv.w == mn.w ? gmn.w : 0.f) + (mv.w == mx.w ? gmx.w : 0.f);
                } else {
                    const float4 ga = ldg4(a.node_in + d * H + c);
                    G = f4(mv.x > 0.f ? ga.x : 0.f, mv.y > 0.f ? ga.y : 0.f, mv.z > 0.f ? ga.z : 0.f, mv.w > 0.f ? ga.w : 0.f);
                }
                if (a.gT1) red_add4(a.gT1 + (int64_t)__ldg(a.nbr + k) * H + c, G);
            }
            st4(Gs + r * T::LDX + c, G);
        }
        __syncthreads();

        tile_wgrad<H>(Gs, Xs, accW);
        if (threadIdx.x < H) {
            float s = 0.f;
            for (int r = 0; r < nrows; ++r) s += Gs[r * T::LDX + threadIdx.x];
            accb += s;
        }

        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        tile_gemm<H>(Gs, Ws, acc, tx, ty);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty * 4 + i;
            if (r < nrows) {
                const int64_t k = row0 + r;
                const int64_t row = a.eid ? (int64_t)__ldg(a.eid + k) : k;
                float* p = a.g_e + row * H + c;
                float4 o = f4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
                if (a.accumulate) o = o + *reinterpret_cast<const float4*>(p);
                st4(p, o);
            }
        }
    }
    store_wgrad<H>(a.gW_part + (int64_t)blockIdx.x * H * H, accW, 1.f);
    if (threadIdx.x < H) a.gb_part[(int64_t)blockIdx.x * H + threadIdx.x] = accb;
}

template <int H, int KIND>
static int launch_msg_bwd(const MsgBwdArgs& a, int n_part, cudaStream_t st) {
    using T = Tile<H>;
    size_t smem = sizeof(float) * (size_t)(T::WS + 2 * T::XS);
    auto kern = message_bwd_kernel<H, KIND>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<n_part, NT, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ edge-update backward
struct UpdBwdArgs {
    int64_t E;
    const float* g_out; const float* h; const float* e;
    const int32_t* src; const int32_t* dst;
    const float* W1; const float* W2; float alpha, beta;
    float* g_e; float* gP; float* gQ;
    float* gW1_part; float* gb1_part; float* gW2_part; float* gb2_part;
};

template <int H>
__global__ void __launch_bounds__(NT) edge_update_bwd_kernel(UpdBwdArgs a) {
    using T = Tile<H>;
    constexpr int MT = T::MT;
    extern __shared__ __align__(16) float smem[];
    float* Ws1 = smem;               // W1e direct
    float* Ws2 = Ws1 + T::WS;        // W2 direct
    float* Gs = Ws2 + T::WS;         // g_out tile (unscaled)
    float* Hs = Gs + T::XS;          // h tile
    float* Es = Hs + T::XS;          // e tile
    float* Ghs = Es + T::XS;         // Gh tile
    int* ds = reinterpret_cast<int*>(Ghs + T::XS);   // dst ids of the tile [TE]

    const int tx = threadIdx.x % T::TX, ty = threadIdx.x / T::TX;
    const int c = tx * 4;
    load_weight<H, false>(Ws1, a.W1);
    load_weight<H, false>(Ws2, a.W2);

    float accW1[MT][MT], accW2[MT][MT];
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < MT; ++j) { accW1[i][j] = 0.f; accW2[i][j] = 0.f; }
    float accb1 = 0.f, accb2 = 0.f;

    const int64_t ntiles = (a.E + T::TE - 1) / T::TE;
    for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int64_t row0 = t * T::TE;
        const int nrows = (int)min64(T::TE, a.E - row0);
        __syncthreads();
        load_tile<H>(Gs, a.g_out, nullptr, row0, nrows);
        load_tile<H>(Hs, a.h, nullptr, row0, nrows);
        load_tile<H>(Es, a.e, nullptr, row0, nrows);
        for (int r = threadIdx.x; r < T::TE; r += NT) ds[r] = r < nrows ? __ldg(a.dst + row0 + r) : -1;
        __syncthreads();

        // grad W2 / b2 (scaled by beta when stored)
        tile_wgrad<H>(Gs, Hs, accW2);
        if (threadIdx.x < H) {
            float s = 0.f;
            for (int r = 0; r < nrows; ++r) s += Gs[r * T::LDX + threadIdx.x];
            accb2 += s;
        }

        // Gh = (beta g_out W2) * [h > 0]
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        tile_gemm<H>(Gs, Ws2, acc, tx, ty);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty * 4 + i;
            float4 G = f4(0.f, 0.f, 0.f, 0.f);
            if (r < nrows) {
                const float4 hv = lds4(Hs + r * T::LDX + c);
                G = f4(hv.x > 0.f ? a.beta * acc[i][0] : 0.f, hv.y > 0.f ? a.beta * acc[i][1] : 0.f,
                       hv.z > 0.f ? a.beta * acc[i][2] : 0.f, hv.w > 0.f ? a.beta * acc[i][3] : 0.f);
                if (a.gP) red_add4(a.gP + (int64_t)__ldg(a.src + row0 + r) * H + c, G);
            }
            st4(Ghs + r * T::LDX + c, G);
        }
        __syncthreads();

        // grad Q[dst] (run-length reduced over equal consecutive dst) and grad b1
        if (threadIdx.x < H) {
            float run = 0.f, tot = 0.f;
            for (int r = 0; r < nrows; ++r) {
                const float v = Ghs[r * T::LDX + threadIdx.x];
                run += v;
                tot += v;
                if (r == nrows - 1 || ds[r + 1] != ds[r]) {
                    if (a.gQ) atomicAdd(a.gQ + (int64_t)ds[r] * H + threadIdx.x, run);
                    run = 0.f;
                }
            }
            accb1 += tot;
        }

        tile_wgrad<H>(Ghs, Es, accW1);

#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        tile_gemm<H>(Ghs, Ws1, acc, tx, ty);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = ty * 4 + i;
            if (r < nrows) {
                const float4 g = lds4(Gs + r * T::LDX + c);
                st4(a.g_e + (row0 + r) * H + c, a.alpha * g + f4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
            }
        }
    }
    store_wgrad<H>(a.gW1_part + (int64_t)blockIdx.x * H * H, accW1, 1.f);
    store_wgrad<H>(a.gW2_part + (int64_t)blockIdx.x * H * H, accW2, a.beta);
    if (threadIdx.x < H) {
        a.gb1_part[(int64_t)blockIdx.x * H + threadIdx.x] = accb1;
        a.gb2_part[(int64_t)blockIdx.x * H + threadIdx.x] = a.beta * accb2;
    }
}

template <int H>
static int launch_upd_bwd(const UpdBwdArgs& a, int n_part, cudaStream_t st) {
    using T = Tile<H>;
    size_t smem = sizeof(float) * (size_t)(2 * T::WS + 4 * T::XS) + sizeof(int) * T::TE;
    auto kern = edge_update_bwd_kernel<H>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<n_part, NT, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

int launch_msg_bwd_umma(int kind, int64_t E, const float* m, const float* e, const int32_t* eid, const float* node_in,
                        const int32_t* seg, const int32_t* nbr, const float* W, float* g_e, int accumulate, float* gT1,
                        float* gW_part, float* gb_part, int n_part, const float* g_acc, cudaStream_t st, float* G_out = nullptr);
int launch_upd_bwd_umma(int64_t E, const float* g_out, const float* h, const float* e, const int32_t* src, const int32_t* dst,
                        const float* W1, const float* W2, float alpha, float beta, float* g_e, float* gP, float* gQ,
                        float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part, int n_part,
                        const int32_t* g_eid, const int32_t* e_eid, cudaStream_t st, float* gh_out = nullptr);

}  // namespace b200mp

extern "C" int b200mp_message_bwd(int kind, int H, int64_t E,
                                  const float* m, const float* e, const int32_t* eid,
                                  const float* node_in, const int32_t* seg, const int32_t* nbr,
                                  const float* W, float* g_e, int accumulate, float* gT1,
                                  float* gW_part, float* gb_part, int n_part, const float* g_acc, void* stream) {
    using namespace b200mp;
    if (E < 0 || n_part <= 0 || !W || !gW_part || !gb_part) return B200MP_EINVAL;
    if (E > 0 && (!g_e || !m || !e || !node_in || !seg || (gT1 && !nbr))) return B200MP_EINVAL;
    MsgBwdArgs a{E, m, e, eid, node_in, seg, nbr, W, g_e, accumulate, gT1, gW_part, gb_part};
    cudaStream_t st = (cudaStream_t)stream;
#define B200MP_MB(HH) return kind == 0 ? launch_msg_bwd<HH, 0>(a, n_part, st) : launch_msg_bwd<HH, 1>(a, n_part, st)
    if (kind != 0 && kind != 1) return B200MP_EINVAL;
    if (H == 64 && use_tensor_cores())
        return launch_msg_bwd_umma(kind, E, m, e, eid, node_in, seg, nbr, W, g_e, accumulate, gT1, gW_part, gb_part, n_part, g_acc, st);
    if (g_acc) return B200MP_EINVAL;                             // tensor-core path only
    switch (H) { case 16: B200MP_MB(16); case 32: B200MP_MB(32); case 64: B200MP_MB(64); case 128: B200MP_MB(128); }
#undef B200MP_MB
    return B200MP_EINVAL;
}

extern "C" int b200mp_edge_update_bwd(int H, int64_t E,
                                      const float* g_out, const float* h, const float* e,
                                      const int32_t* src, const int32_t* dst,
                                      const float* W1, const float* W2, float alpha, float beta,
                                      float* g_e, float* gP, float* gQ,
                                      float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part,
                                      int n_part, const int32_t* g_eid, const int32_t* e_eid, void* stream) {
    using namespace b200mp;
    if (E < 0 || n_part <= 0 || !W1 || !W2 || !gW1_part || !gb1_part || !gW2_part || !gb2_part) return B200MP_EINVAL;
    if (E > 0 && (!g_e || !g_out || !h || !e || !dst || (gP && !src))) return B200MP_EINVAL;
    UpdBwdArgs a{E, g_out, h, e, src, dst, W1, W2, alpha, beta, g_e, gP, gQ, gW1_part, gb1_part, gW2_part, gb2_part};
    cudaStream_t st = (cudaStream_t)stream;
    if (H == 64 && use_tensor_cores())
        return launch_upd_bwd_umma(E, g_out, h, e, src, dst, W1, W2, alpha, beta, g_e, gP, gQ, gW1_part, gb1_part, gW2_part,
                                   gb2_part, n_part, g_eid, e_eid, st);
    if (g_eid || e_eid) return B200MP_EINVAL;                    // tensor-core path only
    switch (H) {
        case 16: return launch_upd_bwd<16>(a, n_part, st);
        case 32: return launch_upd_bwd<32>(a, n_part, st);
        case 64: return launch_upd_bwd<64>(a, n_part, st);
        case 128: return launch_upd_bwd<128>(a, n_part, st);
    }
    return B200MP_EINVAL;
}

// Atomic-free variants for H = 64 (tensor-core path): instead of scattering the per-edge gradient rows into the node tables with
// vector REDs, the kernels write them once as [E,H] (segment order) and the caller sums them per node over the CSC / CSR orders
// with the segmented-sum kernels -- bit-identical from run to run (SURVEY.md section 7 "atomic-free determinism both directions").
extern "C" int b200mp_message_bwd_rows(int kind, int64_t E, const float* m, const float* e, const int32_t* eid,
                                       const float* node_in, const int32_t* seg, const float* W, float* g_e, int accumulate,
                                       float* G_rows, float* gW_part, float* gb_part, int n_part, const float* g_acc, void* stream) {
    using namespace b200mp;
    if (E < 0 || n_part <= 0 || !W || !gW_part || !gb_part || (kind != 0 && kind != 1)) return B200MP_EINVAL;
    if (E > 0 && (!g_e || !m || !e || !node_in || !seg || !G_rows)) return B200MP_EINVAL;
    if (!use_tensor_cores()) return B200MP_EINVAL;
    return launch_msg_bwd_umma(kind, E, m, e, eid, node_in, seg, nullptr, W, g_e, accumulate, nullptr, gW_part, gb_part, n_part, g_acc,
                               (cudaStream_t)stream, G_rows);
}

extern "C" int b200mp_edge_update_bwd_rows(int64_t E, const float* g_out, const float* h, const float* e, const float* W1,
                                           const float* W2, float alpha, float beta, float* g_e, float* Gh_rows, float* gW1_part,
                                           float* gb1_part, float* gW2_part, float* gb2_part, int n_part, const int32_t* g_eid,
                                           const int32_t* e_eid, const int32_t* src, const int32_t* dst, void* stream) {
    using namespace b200mp;
    if (E < 0 || n_part <= 0 || !W1 || !W2 || !gW1_part || !gb1_part || !gW2_part || !gb2_part) return B200MP_EINVAL;
    if (E > 0 && (!g_e || !g_out || !h || !e || !Gh_rows || !src || !dst)) return B200MP_EINVAL;
    if (!use_tensor_cores()) return B200MP_EINVAL;
    return launch_upd_bwd_umma(E, g_out, h, e, src, dst, W1, W2, alpha, beta, g_e, nullptr, nullptr, gW1_part, gb1_part, gW2_part, gb2_part,
                               n_part, g_eid, e_eid, (cudaStream_t)stream, Gh_rows);
}

// ------------------------------------------------------------------ row passes of the H = 128 tensor-core composition
// At H = 128 the per-edge products run on the tcgen05 row-linear kernel (node_gemm_umma.cu: b200mp_row_linear) and the weight
// gradients on b200mp_node_linear_wgrad; what is left of the message backward is the per-edge gradient itself, one pass:
//   G[k] = rule(m[k], node_in[seg[k]])   (same rules as message_bwd_kernel above),   gT1[nbr[k]] += G[k]  (vector REDs, optional)
namespace b200mp {
template <int KIND>
__global__ void __launch_bounds__(256) message_grad_rows_kernel(const float* __restrict__ m, const float* __restrict__ node_in,
                                                                const int32_t* __restrict__ seg, const int32_t* __restrict__ nbr,
                                                                int64_t E, int H, float* __restrict__ G, float* __restrict__ gT1) {
    const int W4 = H / 4;
    const int64_t total = E * W4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / W4;
        const int c = (int)(i - k * W4) * 4;
        const float4 mv = ldg4(m + i * 4);
        const int64_t d = __ldg(seg + k);
        float4 g;
        if (KIND == 0) {
            const float* q = node_in + d * 7 * H + c;
            const float4 c0 = ldg4(q), c1 = ldg4(q + H), gmn = ldg4(q + 2 * H), mn = ldg4(q + 3 * H),
                         gmx = ldg4(q + 4 * H), mx = ldg4(q + 5 * H), mu = ldg4(q + 6 * H);
            g.x = fmaf(c1.x, mv.x - mu.x, c0.x) + (mv.x == mn.x ? gmn.x : 0.f) + (mv.x == mx.x ? gmx.x : 0.f);
            g.y = fmaf(c1.y, mv.y - mu.y, c0.y) + (mv.y == mn.y ? gmn.y : 0.f) + (mv.y == mx.y ? gmx.y : 0.f);
            g.z = fmaf(c1.z, mv.z - mu.z, c0.z) + (mv.z == mn.z ? gmn.z : 0.f) + (mv.z == mx.z ? gmx.z : 0.f);
            g.w = fmaf(c1.w, mv.w - mu.w, c0.w) + (mv.w == mn.w ? gmn.w : 0.f) + (mv.w == mx.w ? gmx.w : 0.f);
        } else {
            const float4 ga = ldg4(node_in + d * H + c);
            g = f4(mv.x > 0.f ? ga.x : 0.f, mv.y > 0.f ? ga.y : 0.f, mv.z > 0.f ? ga.z : 0.f, mv.w > 0.f ? ga.w : 0.f);
        }
        st4(G + i * 4, g);
        if (gT1) red_add4(gT1 + (int64_t)__ldg(nbr + k) * H + c, g);
    }
}

// out[idx[k]] += rows[k]   (vector REDs; the scatter half of the edge-update backward: gP[src] += Gh)
__global__ void __launch_bounds__(256) scatter_add_rows_kernel(const float* __restrict__ rows, const int32_t* __restrict__ idx,
                                                               int64_t n, int W4, float* __restrict__ out) {
    const int64_t total = n * W4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t k = i / W4;
        const int c = (int)(i - k * W4);
        red_add4(out + ((int64_t)__ldg(idx + k) * W4 + c) * 4, ldg4(rows + i * 4));
    }
}

static inline int rows_grid(int64_t items) {
    const int64_t want = (items + 255) / 256, cap = (int64_t)device_info().sms * 16;
    return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}
}  // namespace b200mp

extern "C" int b200mp_message_grad_rows(int kind, int H, int64_t E, const float* m, const float* node_in, const int32_t* seg,
                                        const int32_t* nbr, float* G, float* gT1, void* stream) {
    using namespace b200mp;
    if (E < 0 || H < 4 || (H & 3) || (kind != 0 && kind != 1)) return B200MP_EINVAL;
    if (E == 0) return 0;
    if (!m || !node_in || !seg || !G || (gT1 && !nbr)) return B200MP_EINVAL;
    const int grid = rows_grid(E * (H / 4));
    if (kind == 0) message_grad_rows_kernel<0><<<grid, 256, 0, (cudaStream_t)stream>>>(m, node_in, seg, nbr, E, H, G, gT1);
    else message_grad_rows_kernel<1><<<grid, 256, 0, (cudaStream_t)stream>>>(m, node_in, seg, nbr, E, H, G, gT1);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_scatter_add_rows(const float* rows, const int32_t* idx, int64_t n, int H, float* out, void* stream) {
    using namespace b200mp;
    if (n < 0 || H < 4 || (H & 3)) return B200MP_EINVAL;
    if (n == 0) return 0;
    if (!rows || !idx || !out) return B200MP_EINVAL;
    scatter_add_rows_kernel<<<rows_grid(n * (H / 4)), 256, 0, (cudaStream_t)stream>>>(rows, idx, n, H / 4, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

// Read-out side of the path: the first layer of the reference's edge read-out heads with
// its two node gathers, the ranking metric of link prediction, and the pooled write-back of the fused layer.
//
// Replaces (reference lines):
//   src/nn/gnn/decoder.py:17-21 (ClassifierHead.forward), :62-71 (LinkPredHead.forward): x[edge_index.T] -> relu -> cat
//     with the edge features -> Linear.  relu acts per element, so Linear([relu(x_s) | relu(x_d) | e]) =
//     Rs[src] + Rd[dst] + T with Rs = relu(x) Ws^T, Rd = relu(x) Wd^T (N rows) and T = e We^T + b (E rows): the
//     [E, 3H] concatenation is never built, the gathers move rows of the OUTPUT width;
//   src/utils/metric.py:11-52 (SSMetric.mrr): rank of every positive among its K negatives (per-sample numpy argsort
//     loop on the host in the reference);
//   src/nn/models/fused.py:256-268: unique / index_add_ / bincount mean of the target-edge node embeddings, averaged
//     into x_gnn in place.
// All of it is HBM-bound row traffic; no atomics anywhere (results are run-to-run identical).
#include "common.cuh"

namespace b200mp {

// out[e] = relu(Rs[src[e]] + Rd[dst[e]] + T[e])   (Rs / Rd / T nullable, at least one given); W floats per row, W % 4 == 0.
// One thread per 16-byte chunk: the W / 4 lanes of a row are adjacent, so row reads and writes are whole lines.
__global__ void __launch_bounds__(256) pair_gather_relu_kernel(const float* __restrict__ Rs, const float* __restrict__ Rd,
                                                               const float* __restrict__ T, const int64_t* __restrict__ src,
                                                               const int64_t* __restrict__ dst, int64_t E, int W4, int relu,
                                                               float* __restrict__ out) {
    const int64_t total = E * W4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = i / W4;
        const int j = (int)(i - e * W4);
        float4 v = T ? ldg4(T + i * 4) : f4(0.f, 0.f, 0.f, 0.f);
        if (Rs) v = v + ldg4(Rs + (__ldg(src + e) * W4 + j) * 4);
        if (Rd) v = v + ldg4(Rd + (__ldg(dst + e) * W4 + j) * 4);
        st4(out + i * 4, relu ? relu4(v) : v);
    }
}

// rank[b] = 1 + #{k : neg[b, k] > pos[b]} + #{k : neg[b, k] == pos[b]}: the position of the positive in the descending
// order of [neg_b, pos_b] with the positive LAST among equal scores -- what `(-combined).argsort().argsort()[-1] + 1`
// yields with a stable sort (metric.py:34-37; numpy's default sort leaves ties unspecified).  One warp per positive.
__global__ void __launch_bounds__(256) rank_positive_kernel(const float* __restrict__ pos, const float* __restrict__ neg, int64_t B,
                                                            int K, int32_t* __restrict__ rank) {
    const int lane = threadIdx.x & 31;
    for (int64_t b = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; b < B; b += ((int64_t)gridDim.x * blockDim.x) >> 5) {
        const float p = __ldg(pos + b);
        int c = 0;
        for (int k = lane; k < K; k += 32) c += __ldg(neg + b * K + k) >= p ? 1 : 0;
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) c += __shfl_xor_sync(0xffffffffu, c, s);
        if (lane == 0) rank[b] = c + 1;
    }
}

// Pooled write-back (fused.py:256-268).  The 2B (node, row-of-emb) pairs are sorted by node (ptr / perm from
// b200mp_csc_build over index = target_edge_index.flatten()):
//   out[n] = cnt_n > 0 ? (x[n] + mean_{k in seg n} emb[perm[k]]) / 2 : x[n]
// one thread per (node, 16-byte chunk), rows of a node summed in sorted order.
__global__ void __launch_bounds__(256) segment_mean_blend_fwd_kernel(const float* __restrict__ x, const float* __restrict__ emb,
                                                                     const int64_t* __restrict__ ptr, const int64_t* __restrict__ perm,
                                                                     int64_t N, int W4, float* __restrict__ out) {
    const int64_t total = N * W4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t n = i / W4;
        const int j = (int)(i - n * W4);
        const int64_t k0 = __ldg(ptr + n), k1 = __ldg(ptr + n + 1);
        float4 v = ldg4(x + i * 4);
        if (k1 > k0) {
            float4 s = f4(0.f, 0.f, 0.f, 0.f);
            for (int64_t k = k0; k < k1; ++k) s = s + ldg4(emb + (__ldg(perm + k) * W4 + j) * 4);
            const float inv = 1.f / (float)(k1 - k0);
            v = f4(0.5f * (v.x + s.x * inv), 0.5f * (v.y + s.y * inv), 0.5f * (v.z + s.z * inv), 0.5f * (v.w + s.w * inv));
        }
        st4(out + i * 4, v);
    }
}
// backward: g_x[n] = g[n] * (cnt_n > 0 ? 1/2 : 1);  g_emb[r] = g[index[r]] / (2 cnt_{index[r]})   (r < R rows of emb)
__global__ void __launch_bounds__(256) segment_mean_blend_bwd_kernel(const float* __restrict__ g, const int64_t* __restrict__ ptr,
                                                                     const int64_t* __restrict__ index, int64_t N, int64_t R, int W4,
                                                                     float* __restrict__ gx, float* __restrict__ gemb) {
    const int64_t tn = N * W4, total = tn + R * W4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        if (i < tn) {
            if (!gx) continue;
            const int64_t n = i / W4;
            const float sc = __ldg(ptr + n + 1) > __ldg(ptr + n) ? 0.5f : 1.f;
            st4(gx + i * 4, sc * ldg4(g + i * 4));
        } else {
            if (!gemb) continue;
            const int64_t q = i - tn, r = q / W4;
            const int j = (int)(q - r * W4);
            const int64_t n = __ldg(index + r);
            const float sc = 0.5f / (float)(__ldg(ptr + n + 1) - __ldg(ptr + n));
            st4(gemb + q * 4, sc * ldg4(g + (n * W4 + j) * 4));
        }
    }
}

static inline int flat_grid(int64_t items) {
    const int64_t want = (items + 255) / 256, cap = (int64_t)device_info().sms * 16;
    return (int)(want < 1 ? 1 : (want < cap ? want : cap));
}

}  // namespace b200mp

extern "C" int b200mp_pair_gather_act(const float* Rs, const float* Rd, const float* T, const int64_t* src, const int64_t* dst,
                                      int64_t E, int W, int relu, float* out, void* stream) {
    using namespace b200mp;
    if (E < 0 || W < 4 || (W & 3) || (!Rs && !Rd && !T)) return B200MP_EINVAL;
    if (E == 0) return 0;
    if (!out || (Rs && !src) || (Rd && !dst)) return B200MP_EINVAL;
    pair_gather_relu_kernel<<<flat_grid(E * (W / 4)), 256, 0, (cudaStream_t)stream>>>(Rs, Rd, T, src, dst, E, W / 4, relu, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_rank_positive(const float* pos, const float* neg, int64_t B, int K, int32_t* rank, void* stream) {
    using namespace b200mp;
    if (B < 0 || K < 0) return B200MP_EINVAL;
    if (B == 0) return 0;
    if (!pos || !rank || (K > 0 && !neg)) return B200MP_EINVAL;
    rank_positive_kernel<<<flat_grid(B * 32), 256, 0, (cudaStream_t)stream>>>(pos, neg, B, K, rank);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_segment_mean_blend_fwd(const float* x, const float* emb, const int64_t* ptr, const int64_t* perm, int64_t N,
                                             int W, float* out, void* stream) {
    using namespace b200mp;
    if (N < 0 || W < 4 || (W & 3)) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!x || !ptr || !out) return B200MP_EINVAL;
    segment_mean_blend_fwd_kernel<<<flat_grid(N * (W / 4)), 256, 0, (cudaStream_t)stream>>>(x, emb, ptr, perm, N, W / 4, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_segment_mean_blend_bwd(const float* g, const int64_t* ptr, const int64_t* index, int64_t N, int64_t R, int W,
                                             float* gx, float* gemb, void* stream) {
    using namespace b200mp;
    if (N < 0 || R < 0 || W < 4 || (W & 3)) return B200MP_EINVAL;
    if (N + R == 0) return 0;
    if (!g || !ptr || (R > 0 && gemb && !index)) return B200MP_EINVAL;
    segment_mean_blend_bwd_kernel<<<flat_grid((N + R) * (W / 4)), 256, 0, (cudaStream_t)stream>>>(g, ptr, index, N, R, W / 4, gx, gemb);
    B200MP_LAUNCH_CHECK();
    return 0;
}

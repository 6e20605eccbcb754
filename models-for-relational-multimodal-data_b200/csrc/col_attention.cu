// Column attention core of the FT-Transformer layers: the
// softmax(Q K^T / sqrt(d)) V of torch.nn.MultiheadAttention inside nn.TransformerEncoderLayer as the
// reference builds it (src/nn/models/tabgnn.py:199-209, fused.py:83-93,187-197, inteleaved.py:155:
// d_model = C, nhead = 8, batch_first) over [rows, S, C] column tokens, S = K + 1 <= 8, head_dim = C / 8 (1, 2, 4, 8 or 16).
// The projections around it (in_proj, out_proj, feed-forward) are [rows * S, 64] x 64 linears and run on
// the tcgen05 node kernels (functional.linear_group); this file is the part between them.
//
// The sequences are tiny (S <= 8, d <= 16), so nothing is tiled: eight lanes per (row, head), lane = token; the lanes
// stage their tokens' slices in shared memory (each operand leaves global memory once), the rest is registers.
// Attention-probability dropout (the layer's dropout applies to them in training) uses a counter hash
// of (seed, row, head, i, j), recomputed in backward -- no mask tensor.
#include <cstdlib>
#include "common.cuh"

namespace b200mp {

__device__ __forceinline__ uint32_t hash32(uint64_t z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return (uint32_t)((z ^ (z >> 31)) >> 32);
}
// keep-mask scale of probability (i, j): 0 or 1 / (1 - p)
__device__ __forceinline__ float drop_scale(uint64_t seed, int64_t row, int head, int i, int j, float p, float inv_keep) {
    if (p <= 0.f) return 1.f;
    const uint32_t h = hash32(seed ^ ((uint64_t)row * 0x100000001b3ull + (uint64_t)((head * 8 + i) * 8 + j)));
    return ((float)h * (1.0f / 4294967296.0f)) < p ? 0.f : inv_keep;
}

// HD contiguous floats to / from registers: 16-byte accesses when HD is a multiple of 4, else 8- or 4-byte
template <int HD>
__device__ __forceinline__ void ld_vec(float (&d)[HD], const float* __restrict__ p) {
    if constexpr (HD % 4 == 0) {
#pragma unroll
        for (int c = 0; c < HD; c += 4) { const float4 x = ldg4(p + c); d[c] = x.x; d[c + 1] = x.y; d[c + 2] = x.z; d[c + 3] = x.w; }
    } else if constexpr (HD % 2 == 0) {
#pragma unroll
        for (int c = 0; c < HD; c += 2) { const float2 x = __ldg(reinterpret_cast<const float2*>(p + c)); d[c] = x.x; d[c + 1] = x.y; }
    } else {
#pragma unroll
        for (int c = 0; c < HD; ++c) d[c] = __ldg(p + c);
    }
}
template <int HD>
__device__ __forceinline__ void st_vec(float* __restrict__ p, const float (&d)[HD]) {
    if constexpr (HD % 4 == 0) {
#pragma unroll
        for (int c = 0; c < HD; c += 4) st4(p + c, f4(d[c], d[c + 1], d[c + 2], d[c + 3]));
    } else if constexpr (HD % 2 == 0) {
#pragma unroll
        for (int c = 0; c < HD; c += 2) *reinterpret_cast<float2*>(p + c) = make_float2(d[c], d[c + 1]);
    } else {
#pragma unroll
        for (int c = 0; c < HD; ++c) p[c] = d[c];
    }
}
template <int HD>
__device__ __forceinline__ float dot(const float* __restrict__ a, const float* __restrict__ b) {
    float x[HD], y[HD], s = 0.f;
    ld_vec<HD>(x, a); ld_vec<HD>(y, b);
#pragma unroll
    for (int c = 0; c < HD; ++c) s = fmaf(x[c], y[c], s);
    return s;
}
// d += w * p[0..HD)
template <int HD>
__device__ __forceinline__ void axpy(float (&d)[HD], float w, const float* __restrict__ p) {
    float x[HD];
    ld_vec<HD>(x, p);
#pragma unroll
    for (int c = 0; c < HD; ++c) d[c] = fmaf(w, x[c], d[c]);
}

struct AttnArgs {
    int64_t R; int S; int C; int nhead;
    const float* q; const float* k; const float* v;       // [R, S, C] each
    float p; uint64_t seed;
};

// out[r, i, h*HD : (h+1)*HD] = sum_j drop(softmax_j(q_i . k_j / sqrt(HD))) v_j
// Eight lanes per (row, head), lane i = token i (lanes >= S idle): every lane stages its token's k and v slices in shared
// memory once, so a (row, head) reads each operand from global memory exactly once.
constexpr int ATT_BT = 128;               // threads per CTA = 16 (row, head) groups
template <int HD>
__global__ void __launch_bounds__(ATT_BT) col_attention_fwd_kernel(AttnArgs a, float* __restrict__ out) {
    constexpr int NG = ATT_BT / 8;
    constexpr int LD = HD % 4 == 0 ? HD : 4;
    constexpr int GS = 8 * LD + 8;            // group stride: +8 floats, so that the four groups of a warp read four different bank octets
    __shared__ __align__(16) float sk[NG][GS], sv[NG][GS];
    const int grp = threadIdx.x >> 3, i = threadIdx.x & 7;
    const int64_t gid = (int64_t)blockIdx.x * NG + grp;
    const bool live = gid < a.R * a.nhead && i < a.S;
    const int h = (int)(gid % a.nhead);
    const int64_t r = gid / a.nhead;
    const float scale = rsqrtf((float)HD), inv_keep = a.p > 0.f ? 1.f / (1.f - a.p) : 1.f;
    const int64_t o = live ? (r * a.S + i) * a.C + h * HD : 0;
    float qi[HD], tmp[HD];
    if (live) {
        ld_vec<HD>(qi, a.q + o);
        ld_vec<HD>(tmp, a.k + o);
#pragma unroll
        for (int c = 0; c < HD; ++c) sk[grp][i * LD + c] = tmp[c];
        ld_vec<HD>(tmp, a.v + o);
#pragma unroll
        for (int c = 0; c < HD; ++c) sv[grp][i * LD + c] = tmp[c];
    }
    __syncwarp();
    if (!live) return;
    float s[8], m = -INFINITY;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float d = 0.f;
        if (j < a.S) {
#pragma unroll
            for (int c = 0; c < HD; ++c) d = fmaf(qi[c], sk[grp][j * LD + c], d);
        }
        s[j] = j < a.S ? d * scale : -INFINITY;
        m = fmaxf(m, s[j]);
    }
    float l = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j] = j < a.S ? __expf(s[j] - m) : 0.f; l += s[j]; }
    const float inv_l = 1.f / l;
    float acc[HD];
#pragma unroll
    for (int c = 0; c < HD; ++c) acc[c] = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        if (j < a.S) {
            const float pj = s[j] * inv_l * drop_scale(a.seed, r, h, i, j, a.p, inv_keep);
#pragma unroll
            for (int c = 0; c < HD; ++c) acc[c] = fmaf(pj, sv[grp][j * LD + c], acc[c]);
        }
    }
    st_vec<HD>(out + o, acc);
}

// The same forward as a persistent loop: a CTA walks (row, head) groups gid, gid + stride, ... and the NEXT group's q / k / v slices are
// in flight (registers) while the current one is computed -- the one-group-per-CTA form above alternates a load phase with a
// compute phase and pays a CTA launch per 16 groups (524 k CTAs per 1 M rows).
template <int HD>
__global__ void __launch_bounds__(ATT_BT) col_attention_fwd_loop_kernel(AttnArgs a, float* __restrict__ out) {
    constexpr int NG = ATT_BT / 8;
    constexpr int LD = HD % 4 == 0 ? HD : 4;
    constexpr int GS = 8 * LD + 8;
    __shared__ __align__(16) float sk[NG][GS], sv[NG][GS];
    const int grp = threadIdx.x >> 3, i = threadIdx.x & 7;
    const int64_t groups = a.R * a.nhead, stride = (int64_t)gridDim.x * NG;
    const float scale = rsqrtf((float)HD), inv_keep = a.p > 0.f ? 1.f / (1.f - a.p) : 1.f;
    const bool tok = i < a.S;
    float qi[HD], ki[HD], vi[HD];
    int64_t gid = (int64_t)blockIdx.x * NG + grp;
    auto offset = [&](int64_t g_) { return ((g_ / a.nhead) * a.S + i) * a.C + (g_ % a.nhead) * HD; };
    if (tok && gid < groups) {
        const int64_t o = offset(gid);
        ld_vec<HD>(qi, a.q + o); ld_vec<HD>(ki, a.k + o); ld_vec<HD>(vi, a.v + o);
    }
    for (; gid - (grp & 3) < groups; gid += stride) {   // warp-uniform trip count: the four groups of a warp leave the loop together
        const int h = (int)(gid % a.nhead);
        const int64_t r = gid / a.nhead;
        const bool on = tok && gid < groups;
        if (on) {
#pragma unroll
            for (int c = 0; c < HD; ++c) { sk[grp][i * LD + c] = ki[c]; sv[grp][i * LD + c] = vi[c]; }
        }
        __syncwarp();
        float q0[HD];
#pragma unroll
        for (int c = 0; c < HD; ++c) q0[c] = qi[c];
        const int64_t nxt = gid + stride;
        if (tok && nxt < groups) {                  // next group's slices in flight during this group's arithmetic
            const int64_t o = offset(nxt);
            ld_vec<HD>(qi, a.q + o); ld_vec<HD>(ki, a.k + o); ld_vec<HD>(vi, a.v + o);
        }
        if (on) {
            float s[8], m = -INFINITY;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float d = 0.f;
                if (j < a.S) {
#pragma unroll
                    for (int c = 0; c < HD; ++c) d = fmaf(q0[c], sk[grp][j * LD + c], d);
                }
                s[j] = j < a.S ? d * scale : -INFINITY;
                m = fmaxf(m, s[j]);
            }
            float l = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) { s[j] = j < a.S ? __expf(s[j] - m) : 0.f; l += s[j]; }
            const float inv_l = 1.f / l;
            float acc[HD];
#pragma unroll
            for (int c = 0; c < HD; ++c) acc[c] = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < a.S) {
                    const float pj = s[j] * inv_l * drop_scale(a.seed, r, h, i, j, a.p, inv_keep);
#pragma unroll
                    for (int c = 0; c < HD; ++c) acc[c] = fmaf(pj, sv[grp][j * LD + c], acc[c]);
                }
            }
            st_vec<HD>(out + (r * a.S + i) * a.C + h * HD, acc);
        }
        __syncwarp();                               // the group's shared slices are rewritten by the next iteration
    }
}

// Backward: eight lanes per (row, head), lane i = token i (lanes >= S idle).  Each lane stages its token's q, k, v, dO
// slices in shared memory once, computes ROW i of the probabilities (P = softmax rows, Pd = drop(P),
// dPd_ij = dO_i . v_j, dP_ij = dPd_ij * dropscale_ij, D_i = sum_j P_ij dP_ij, dS_ij = P_ij (dP_ij - D_i)) and
// dQ_i = scale sum_j dS_ij k_j, leaves dS_i. and Pd_i. in shared memory, and after a warp barrier reads COLUMN i:
//   dK_i = scale sum_t dS_ti q_t ,  dV_i = sum_t Pd_ti dO_t
// (the first version recomputed every row in every lane: S times the loads and dot products; 88 -> see DESIGN.md)
template <int HD>
__global__ void __launch_bounds__(ATT_BT) col_attention_bwd_kernel(AttnArgs a, const float* __restrict__ g_out,
                                                                   float* __restrict__ gq, float* __restrict__ gk,
                                                                   float* __restrict__ gv) {
    constexpr int NG = ATT_BT / 8;
    constexpr int LD = HD % 4 == 0 ? HD : 4;              // 16-byte rows at least
    constexpr int GS = 8 * LD + 8;            // (as in the forward kernel: no bank conflicts between the groups of a warp)
    __shared__ __align__(16) float sq[NG][GS], sk[NG][GS], sv[NG][GS], sg[NG][GS];
    __shared__ float sds[NG][8][9], spd[NG][8][9];
    const int grp = threadIdx.x >> 3, i = threadIdx.x & 7;
    const int64_t gid = (int64_t)blockIdx.x * NG + grp;
    const bool live = gid < a.R * a.nhead && i < a.S;
    const int h = (int)(gid % a.nhead);
    const int64_t r = gid / a.nhead;
    const float scale = rsqrtf((float)HD), inv_keep = a.p > 0.f ? 1.f / (1.f - a.p) : 1.f;
    const int64_t o = live ? (r * a.S + i) * a.C + h * HD : 0;
    float qi[HD], gi[HD], tmp[HD];
    if (live) {
        ld_vec<HD>(qi, a.q + o);
        ld_vec<HD>(gi, g_out + o);
#pragma unroll
        for (int c = 0; c < HD; ++c) { sq[grp][i * LD + c] = qi[c]; sg[grp][i * LD + c] = gi[c]; }
        ld_vec<HD>(tmp, a.k + o);
#pragma unroll
        for (int c = 0; c < HD; ++c) sk[grp][i * LD + c] = tmp[c];
        ld_vec<HD>(tmp, a.v + o);
#pragma unroll
        for (int c = 0; c < HD; ++c) sv[grp][i * LD + c] = tmp[c];
    }
    __syncwarp();
    float dq[HD], dk[HD], dv[HD];
#pragma unroll
    for (int c = 0; c < HD; ++c) dq[c] = dk[c] = dv[c] = 0.f;
    if (live) {
        float s[8], dp[8], m = -INFINITY;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float d = 0.f, e = 0.f;
            if (j < a.S) {
#pragma unroll
                for (int c = 0; c < HD; ++c) { d = fmaf(qi[c], sk[grp][j * LD + c], d); e = fmaf(gi[c], sv[grp][j * LD + c], e); }
            }
            s[j] = j < a.S ? d * scale : -INFINITY;
            dp[j] = j < a.S ? e * drop_scale(a.seed, r, h, i, j, a.p, inv_keep) : 0.f;
            m = fmaxf(m, s[j]);
        }
        float l = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] = j < a.S ? __expf(s[j] - m) : 0.f; l += s[j]; }
        const float inv_l = 1.f / l;
        float D = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { s[j] *= inv_l; D = fmaf(s[j], dp[j], D); }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            if (j < a.S) {
                const float ds = s[j] * (dp[j] - D) * scale;
                sds[grp][i][j] = ds;
                spd[grp][i][j] = s[j] * drop_scale(a.seed, r, h, i, j, a.p, inv_keep);
#pragma unroll
                for (int c = 0; c < HD; ++c) dq[c] = fmaf(ds, sk[grp][j * LD + c], dq[c]);
            }
        }
    }
    __syncwarp();
    if (live) {
        for (int t = 0; t < a.S; ++t) {
            const float ds = sds[grp][t][i], pd = spd[grp][t][i];
#pragma unroll
            for (int c = 0; c < HD; ++c) { dk[c] = fmaf(ds, sq[grp][t * LD + c], dk[c]); dv[c] = fmaf(pd, sg[grp][t * LD + c], dv[c]); }
        }
        st_vec<HD>(gq + o, dq);
        st_vec<HD>(gk + o, dk);
        st_vec<HD>(gv + o, dv);
    }
}

}  // namespace b200mp

extern "C" int b200mp_col_attention_fwd(const float* q, const float* k, const float* v, int64_t R, int S, int C, int nhead,
                                        float p_drop, uint64_t seed, float* out, void* stream) {
    using namespace b200mp;
    if (R < 0 || S < 1 || S > 8 || nhead < 1 || nhead > 16 || C % nhead || p_drop < 0.f || p_drop >= 1.f) return B200MP_EINVAL;
    if (R == 0) return 0;
    if (!q || !k || !v || !out) return B200MP_EINVAL;
    AttnArgs a{R, S, C, nhead, q, k, v, p_drop, seed};
    const int64_t groups = R * nhead;
    const unsigned grid = (unsigned)((groups + ATT_BT / 8 - 1) / (ATT_BT / 8));
    cudaStream_t st = (cudaStream_t)stream;
    const char* lv = getenv("B200MP_ATT_LOOP");             // A/B switch: persistent forward with the next group's loads in flight
    if (lv && lv[0] == '1') {
        const int64_t cap = (int64_t)device_info().sms * 8;
        const unsigned lgrid = (unsigned)((int64_t)grid < cap ? (int64_t)grid : cap);
        switch (C / nhead) {
            case 1: col_attention_fwd_loop_kernel<1><<<lgrid, ATT_BT, 0, st>>>(a, out); break;
            case 2: col_attention_fwd_loop_kernel<2><<<lgrid, ATT_BT, 0, st>>>(a, out); break;
            case 4: col_attention_fwd_loop_kernel<4><<<lgrid, ATT_BT, 0, st>>>(a, out); break;
            case 8: col_attention_fwd_loop_kernel<8><<<lgrid, ATT_BT, 0, st>>>(a, out); break;
            case 16: col_attention_fwd_loop_kernel<16><<<lgrid, ATT_BT, 0, st>>>(a, out); break;
            default: return B200MP_EINVAL;
        }
        B200MP_LAUNCH_CHECK();
        return 0;
    }
    switch (C / nhead) {
        case 1: col_attention_fwd_kernel<1><<<grid, ATT_BT, 0, st>>>(a, out); break;
        case 2: col_attention_fwd_kernel<2><<<grid, ATT_BT, 0, st>>>(a, out); break;
        case 4: col_attention_fwd_kernel<4><<<grid, ATT_BT, 0, st>>>(a, out); break;
        case 8: col_attention_fwd_kernel<8><<<grid, ATT_BT, 0, st>>>(a, out); break;
        case 16: col_attention_fwd_kernel<16><<<grid, ATT_BT, 0, st>>>(a, out); break;
        default: return B200MP_EINVAL;
    }
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_col_attention_bwd(const float* q, const float* k, const float* v, const float* g_out, int64_t R, int S,
                                        int C, int nhead, float p_drop, uint64_t seed, float* gq, float* gk, float* gv,
                                        void* stream) {
    using namespace b200mp;
    if (R < 0 || S < 1 || S > 8 || nhead < 1 || nhead > 16 || C % nhead || p_drop < 0.f || p_drop >= 1.f) return B200MP_EINVAL;
    if (R == 0) return 0;
    if (!q || !k || !v || !g_out || !gq || !gk || !gv) return B200MP_EINVAL;
    AttnArgs a{R, S, C, nhead, q, k, v, p_drop, seed};
    const int64_t groups = R * nhead;
    const unsigned grid = (unsigned)((groups + ATT_BT / 8 - 1) / (ATT_BT / 8));
    cudaStream_t st = (cudaStream_t)stream;
    switch (C / nhead) {
        case 1: col_attention_bwd_kernel<1><<<grid, ATT_BT, 0, st>>>(a, g_out, gq, gk, gv); break;
        case 2: col_attention_bwd_kernel<2><<<grid, ATT_BT, 0, st>>>(a, g_out, gq, gk, gv); break;
        case 4: col_attention_bwd_kernel<4><<<grid, ATT_BT, 0, st>>>(a, g_out, gq, gk, gv); break;
        case 8: col_attention_bwd_kernel<8><<<grid, ATT_BT, 0, st>>>(a, g_out, gq, gk, gv); break;
        case 16: col_attention_bwd_kernel<16><<<grid, ATT_BT, 0, st>>>(a, g_out, gq, gk, gv); break;
        default: return B200MP_EINVAL;
    }
    B200MP_LAUNCH_CHECK();
    return 0;
}

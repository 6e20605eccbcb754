// Residual + dropout + LayerNorm over column tokens, the element-wise glue of the reference's post-norm
// nn.TransformerEncoderLayer and the LayerNorm after it (/root/reference/src/nn/models/tabgnn.py:199-209,219,
// fused.py:83-93,249, inteleaved.py:155,219):
//     z = x + drop(sub)        (sub == NULL: z = x, a plain LayerNorm)
//     y = (z - mean(z)) * rstd(z) * gamma + beta ,   rstd = 1 / sqrt(var + eps), biased variance (torch semantics)
// Rows are C = 16..128 floats: C / 4 lanes own one row (16 bytes each), the two row statistics are xor-shuffle
// reductions inside that lane group, so a warp handles 32 / (C / 4) rows per load instruction and every access is a
// full 16-byte vector.  torch's own kernels spend one CTA per row here and two more kernels on the gamma / beta
// gradients; this is one pass forward (read x, sub; write y, 8 bytes of statistics per row) and one pass backward
// (read g, x, sub; write g_x, g_sub), with the gamma / beta gradients as per-thread column partials reduced once per
// CTA.  The dropout mask is a counter hash of (seed, row, column quad), regenerated in backward.
#include "common.cuh"

namespace b200mp {

__device__ __forceinline__ uint32_t rn_hash(uint64_t z) {
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return (uint32_t)((z ^ (z >> 31)) >> 32);
}
// keep-scales of the four elements of column quad c4 of a row: 16 bits of two hashes each (p quantised to 2^-16)
__device__ __forceinline__ float4 drop4(uint64_t seed, int64_t row, int c4, float p, float inv_keep) {
    const uint64_t key = seed ^ ((uint64_t)row * 0x100000001b3ull + (uint64_t)c4);
    const uint32_t h1 = rn_hash(key), h2 = rn_hash(key ^ 0x5851f42d4c957f2dull);
    const uint32_t thr = (uint32_t)(p * 65536.f);
    return f4((h1 & 65535u) < thr ? 0.f : inv_keep, (h1 >> 16) < thr ? 0.f : inv_keep,
              (h2 & 65535u) < thr ? 0.f : inv_keep, (h2 >> 16) < thr ? 0.f : inv_keep);
}
__device__ __forceinline__ float4 mul4(float4 a, float4 b) { return f4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }

template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int off = LPR / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

struct RowNormArgs {
    int64_t rows; int C;
    const float* x; const float* sub; float p; uint64_t seed;
    const float* gamma; const float* beta; float eps;
    // extended forms (b200mp_add_dropout_layernorm_fwd_ex / _bwd_ex); the plain entry points leave them at {1, null, ...}:
    //   forward   y = yscale * LayerNorm(z) + bs0 * b0 + bs1 * b1      (b0 / b1 nullable: residual blends such as (t + layer(t)) / 2)
    //   backward  g = gs0 * g0 + gs1 * g1 + gs2 * g2                   (g1 / g2 nullable: a gradient that is still a sum of terms)
    float yscale; const float* b0; float bs0; const float* b1; float bs1;
    const float* g1; const float* g2; float gs0, gs1, gs2;
};

template <int LPR>
__global__ void __launch_bounds__(256) row_norm_fwd_kernel(RowNormArgs a, float* __restrict__ y, float* __restrict__ stats) {
    constexpr int RPC = 256 / LPR;                              // rows per CTA iteration
    const int c4 = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    const float4 gm = ldg4(a.gamma + c4 * 4), bt = ldg4(a.beta + c4 * 4);
    const float inv_keep = a.p > 0.f ? 1.f / (1.f - a.p) : 1.f, inv_c = 1.f / (float)a.C;
    const int64_t iters = (a.rows + RPC - 1) / RPC;
    for (int64_t it = blockIdx.x; it < iters; it += gridDim.x) {
        const int64_t r = it * RPC + rl;
        const bool live = r < a.rows;                            // dead lanes still take part in the shuffles
        float4 z = f4(0.f, 0.f, 0.f, 0.f);
        if (live) {
            z = ldg4(a.x + r * a.C + c4 * 4);
            if (a.sub) {
                float4 s = ldg4(a.sub + r * a.C + c4 * 4);
                if (a.p > 0.f) s = mul4(s, drop4(a.seed, r, c4, a.p, inv_keep));
                z = z + s;
            }
        }
        const float mean = group_sum<LPR>(z.x + z.y + z.z + z.w) * inv_c;
        const float4 d = f4(z.x - mean, z.y - mean, z.z - mean, z.w - mean);
        const float var = group_sum<LPR>(d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w) * inv_c;
        const float rstd = rsqrtf(var + a.eps);
        if (live) {
            float4 o = f4(d.x * rstd * gm.x + bt.x, d.y * rstd * gm.y + bt.y, d.z * rstd * gm.z + bt.z, d.w * rstd * gm.w + bt.w);
            if (a.b0 || a.yscale != 1.f) {
                o = a.yscale * o;
                if (a.b0) o = o + a.bs0 * ldg4(a.b0 + r * a.C + c4 * 4);
                if (a.b1) o = o + a.bs1 * ldg4(a.b1 + r * a.C + c4 * 4);
            }
            st4(y + r * a.C + c4 * 4, o);
            if (c4 == 0) *reinterpret_cast<float2*>(stats + r * 2) = make_float2(mean, rstd);
        }
    }
}

// g_z = rstd (gh - mean(gh) - zh mean(gh zh)),  gh = g gamma, zh = (z - mean) rstd;  g_x = g_z, g_sub = g_z * dropscale
template <int LPR>
__global__ void __launch_bounds__(256) row_norm_bwd_kernel(RowNormArgs a, const float* __restrict__ g,
                                                           const float* __restrict__ stats, float* __restrict__ gx,
                                                           float* __restrict__ gsub, float* __restrict__ ggamma,
                                                           float* __restrict__ gbeta) {
    constexpr int RPC = 256 / LPR;
    __shared__ float4 red[2][256];
    const int c4 = threadIdx.x % LPR, rl = threadIdx.x / LPR;
    const float4 gm = ldg4(a.gamma + c4 * 4);
    const float inv_keep = a.p > 0.f ? 1.f / (1.f - a.p) : 1.f, inv_c = 1.f / (float)a.C;
    float4 agm = f4(0.f, 0.f, 0.f, 0.f), abt = agm;
    const int64_t iters = (a.rows + RPC - 1) / RPC;
    for (int64_t it = blockIdx.x; it < iters; it += gridDim.x) {
        const int64_t r = it * RPC + rl;
        const bool live = r < a.rows;
        float4 z = f4(0.f, 0.f, 0.f, 0.f), gv = z, ds = f4(1.f, 1.f, 1.f, 1.f);
        float mean = 0.f, rstd = 0.f;
        if (live) {
            gv = ldg4(g + r * a.C + c4 * 4);
            if (a.g1 || a.gs0 != 1.f) {
                gv = a.gs0 * gv;
                if (a.g1) gv = gv + a.gs1 * ldg4(a.g1 + r * a.C + c4 * 4);
                if (a.g2) gv = gv + a.gs2 * ldg4(a.g2 + r * a.C + c4 * 4);
            }
            z = ldg4(a.x + r * a.C + c4 * 4);
            if (a.sub) {
                float4 s = ldg4(a.sub + r * a.C + c4 * 4);
                if (a.p > 0.f) { ds = drop4(a.seed, r, c4, a.p, inv_keep); s = mul4(s, ds); }
                z = z + s;
            }
            const float2 st = __ldg(reinterpret_cast<const float2*>(stats + r * 2));
            mean = st.x; rstd = st.y;
        }
        const float4 zh = f4((z.x - mean) * rstd, (z.y - mean) * rstd, (z.z - mean) * rstd, (z.w - mean) * rstd);
        const float4 gh = mul4(gv, gm);
        abt = abt + gv;
        agm = agm + mul4(gv, zh);
        const float m1 = group_sum<LPR>(gh.x + gh.y + gh.z + gh.w) * inv_c;
        const float m2 = group_sum<LPR>(gh.x * zh.x + gh.y * zh.y + gh.z * zh.z + gh.w * zh.w) * inv_c;
        if (live) {
            const float4 gz = f4(rstd * (gh.x - m1 - zh.x * m2), rstd * (gh.y - m1 - zh.y * m2), rstd * (gh.z - m1 - zh.z * m2),
                                 rstd * (gh.w - m1 - zh.w * m2));
            st4(gx + r * a.C + c4 * 4, gz);
            if (gsub) st4(gsub + r * a.C + c4 * 4, mul4(gz, ds));
        }
    }
    // column partials of this CTA: threads with equal c4 hold the same columns
    red[0][threadIdx.x] = agm;
    red[1][threadIdx.x] = abt;
    __syncthreads();
    if (threadIdx.x < LPR) {
        float4 sg = f4(0.f, 0.f, 0.f, 0.f), sb = sg;
        for (int k = threadIdx.x; k < 256; k += LPR) { sg = sg + red[0][k]; sb = sb + red[1][k]; }
        red_add4(ggamma + threadIdx.x * 4, sg);
        red_add4(gbeta + threadIdx.x * 4, sb);
    }
}


// y = dropout(relu(x)) (the feed-forward's inner activation, tabgnn.py:199-209) and its backward g_x = g * scale [y > 0]:
// element-wise over quads, the mask is the same counter hash keyed by the flat quad index
__global__ void __launch_bounds__(256) relu_dropout_kernel(const float* __restrict__ x, const float* __restrict__ g, int64_t quads,
                                                           float p, uint64_t seed, float* __restrict__ out) {
    const float inv_keep = p > 0.f ? 1.f / (1.f - p) : 1.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = ldg4(x + i * 4);                        // forward: x; backward: the saved output y
        const float4 ds = p > 0.f ? drop4(seed, i, 0, p, inv_keep) : f4(1.f, 1.f, 1.f, 1.f);
        if (!g) {
            st4(out + i * 4, mul4(relu4(v), ds));
        } else {
            const float4 gv = ldg4(g + i * 4);
            st4(out + i * 4, f4(v.x > 0.f ? gv.x * ds.x : 0.f, v.y > 0.f ? gv.y * ds.y : 0.f, v.z > 0.f ? gv.z * ds.z : 0.f,
                                v.w > 0.f ? gv.w * ds.w : 0.f));
        }
    }
}

// out = sum_i s_i t_i over up to 6 tensors (quads): a gradient carried as a sum of terms, written once where it leaves the column
// stack (functional._ColumnStack) instead of being added up pair by pair
struct ScaledSumArgs { const float* t[6]; float s[6]; int n; };
__global__ void __launch_bounds__(256) scaled_sum_kernel(ScaledSumArgs a, int64_t quads, float* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (int64_t)gridDim.x * blockDim.x) {
        float4 v = a.s[0] * ldg4(a.t[0] + i * 4);
#pragma unroll
        for (int k = 1; k < 6; ++k)
            if (k < a.n) v = v + a.s[k] * ldg4(a.t[k] + i * 4);
        st4(out + i * 4, v);
    }
}

static int row_norm_grid(int64_t rows, int lpr) {
    const int64_t iters = (rows + 256 / lpr - 1) / (256 / lpr);
    const int64_t cap = (int64_t)device_info().sms * 8;
    return (int)(iters < cap ? iters : cap);
}

static bool row_norm_ok(int64_t rows, int C, float p) {
    return rows >= 0 && (C == 16 || C == 32 || C == 64 || C == 128) && p >= 0.f && p < 1.f;
}

}  // namespace b200mp

extern "C" int b200mp_add_dropout_layernorm_fwd(const float* x, const float* sub, int64_t rows, int C, float p_drop,
                                                uint64_t seed, const float* gamma, const float* beta, float eps, float* y,
                                                float* stats, void* stream) {
    return b200mp_add_dropout_layernorm_fwd_ex(x, sub, rows, C, p_drop, seed, gamma, beta, eps, 1.f, nullptr, 0.f, nullptr, 0.f, y, stats,
                                               stream);
}

extern "C" int b200mp_add_dropout_layernorm_fwd_ex(const float* x, const float* sub, int64_t rows, int C, float p_drop,
                                                   uint64_t seed, const float* gamma, const float* beta, float eps, float yscale,
                                                   const float* b0, float bs0, const float* b1, float bs1, float* y, float* stats,
                                                   void* stream) {
    using namespace b200mp;
    if (!row_norm_ok(rows, C, p_drop)) return B200MP_EINVAL;
    if (rows == 0) return 0;
    if (!x || !gamma || !beta || !y || !stats || (b1 && !b0)) return B200MP_EINVAL;
    RowNormArgs a{rows, C, x, sub, sub ? p_drop : 0.f, seed, gamma, beta, eps, yscale, b0, bs0, b1, bs1, nullptr, nullptr, 1.f, 0.f, 0.f};
    cudaStream_t st = (cudaStream_t)stream;
    const int lpr = C / 4, grid = row_norm_grid(rows, lpr);
    switch (lpr) {
        case 4: row_norm_fwd_kernel<4><<<grid, 256, 0, st>>>(a, y, stats); break;
        case 8: row_norm_fwd_kernel<8><<<grid, 256, 0, st>>>(a, y, stats); break;
        case 16: row_norm_fwd_kernel<16><<<grid, 256, 0, st>>>(a, y, stats); break;
        default: row_norm_fwd_kernel<32><<<grid, 256, 0, st>>>(a, y, stats); break;
    }
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_add_dropout_layernorm_bwd(const float* g, const float* x, const float* sub, const float* stats,
                                                int64_t rows, int C, float p_drop, uint64_t seed, const float* gamma, float* gx,
                                                float* gsub, float* ggamma, float* gbeta, void* stream) {
    return b200mp_add_dropout_layernorm_bwd_ex(g, 1.f, nullptr, 0.f, nullptr, 0.f, x, sub, stats, rows, C, p_drop, seed, gamma, gx, gsub,
                                               ggamma, gbeta, stream);
}

extern "C" int b200mp_add_dropout_layernorm_bwd_ex(const float* g0, float gs0, const float* g1, float gs1, const float* g2, float gs2,
                                                   const float* x, const float* sub, const float* stats, int64_t rows, int C,
                                                   float p_drop, uint64_t seed, const float* gamma, float* gx, float* gsub,
                                                   float* ggamma, float* gbeta, void* stream) {
    using namespace b200mp;
    const float* g = g0;
    if (!row_norm_ok(rows, C, p_drop)) return B200MP_EINVAL;
    if (rows == 0) return 0;
    if (!g || !x || !stats || !gamma || !gx || !ggamma || !gbeta || (sub && p_drop > 0.f && !gsub) || (g2 && !g1)) return B200MP_EINVAL;
    RowNormArgs a{rows, C, x, sub, sub ? p_drop : 0.f, seed, gamma, nullptr, 0.f, 1.f, nullptr, 0.f, nullptr, 0.f, g1, g2, gs0, gs1, gs2};
    cudaStream_t st = (cudaStream_t)stream;
    const int lpr = C / 4, grid = row_norm_grid(rows, lpr);
    switch (lpr) {
        case 4: row_norm_bwd_kernel<4><<<grid, 256, 0, st>>>(a, g, stats, gx, gsub, ggamma, gbeta); break;
        case 8: row_norm_bwd_kernel<8><<<grid, 256, 0, st>>>(a, g, stats, gx, gsub, ggamma, gbeta); break;
        case 16: row_norm_bwd_kernel<16><<<grid, 256, 0, st>>>(a, g, stats, gx, gsub, ggamma, gbeta); break;
        default: row_norm_bwd_kernel<32><<<grid, 256, 0, st>>>(a, g, stats, gx, gsub, ggamma, gbeta); break;
    }
    B200MP_LAUNCH_CHECK();
    return 0;
}

/* x (forward) or the saved output y (backward, g != NULL): numel must be a multiple of 4 */
extern "C" int b200mp_relu_dropout(const float* x_or_y, const float* g, int64_t numel, float p_drop, uint64_t seed, float* out,
                                   void* stream) {
    using namespace b200mp;
    if (numel < 0 || (numel & 3) || p_drop < 0.f || p_drop >= 1.f) return B200MP_EINVAL;
    if (numel == 0) return 0;
    if (!x_or_y || !out) return B200MP_EINVAL;
    const int64_t quads = numel / 4, want = (quads + 255) / 256, cap = (int64_t)device_info().sms * 16;
    relu_dropout_kernel<<<(int)(want < cap ? want : cap), 256, 0, (cudaStream_t)stream>>>(x_or_y, g, quads, p_drop, seed, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

/* out = sum_{i < n} scales[i] * terms[i]  (n <= 6 tensors of numel floats, numel % 4 == 0; host arrays of pointers / scales) */
extern "C" int b200mp_scaled_sum(const float* const* terms, const float* scales, int n, int64_t numel, float* out, void* stream) {
    using namespace b200mp;
    if (n < 1 || n > 6 || numel < 0 || (numel & 3) || !terms || !scales) return B200MP_EINVAL;
    if (numel == 0) return 0;
    if (!out) return B200MP_EINVAL;
    ScaledSumArgs a{};
    a.n = n;
    for (int i = 0; i < n; ++i) {
        if (!terms[i]) return B200MP_EINVAL;
        a.t[i] = terms[i];
        a.s[i] = scales[i];
    }
    const int64_t quads = numel / 4, want = (quads + 255) / 256, cap = (int64_t)device_info().sms * 16;
    scaled_sum_kernel<<<(int)(want < cap ? want : cap), 256, 0, (cudaStream_t)stream>>>(a, quads, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

// Column (stype-wise) feature encoders, SURVEY.md section 8 row a5: what the reference's
// StypeWiseFeatureEncoder call computes (/root/reference/src/utils.py:139-140,357-359, fused.py:71-72; built at
// src/datasets/ibm_transactions_for_aml.py:283-294 from EmbeddingEncoder / LinearEncoder / TimestampEncoder of
// the pytorch-frame fork, which is absent from the reference tree -- semantics restated in SURVEY.md App. A.4 and
// oracle/stype_encoders.py, parity unpinned):
//   numerical   x[r, c] = ((v_r - mean) / std) * w[c] + b[c]
//   categorical x[r, :] = table[id_r + 1, :]                     (id -1 = missing -> row 0, the padding row)
//   timestamp   x[r, :] = enc(r)[56] . W[56, C] + b,  enc = 7 calendar fields x 8: year -> sin/cos((y - min_year) 10^-q),
//               the six cyclic fields f / max_f -> sin(x (q+1) pi), cos(x (q+1) 2 pi), q = 0..3
// One launch writes the whole [rows, S, C] token tensor once -- all columns of all stypes in their output order and,
// when the caller passes one, the CLS token in slot 0 (the concat of tabgnn.py:214 / fused.py:149 folded in).
// Sixteen lanes own one (row, token) and write its 4 C bytes as 16-byte stores; the 56 timestamp encodings of a row
// are computed once by 14 of the lanes and shared through shared memory.
//
// The inputs are data, so backward is parameter gradients only: per column a persistent grid reduces over the rows
// (numerical: sum z g and sum g; categorical: per-category sums in a shared-memory table when it fits; timestamp: the
// [56 x rows] . [rows x C] contraction in 32-row tiles with 16 accumulators per thread) and adds its partial to the
// caller's gradient buffers with one atomic per element per CTA.
#include "common.cuh"

namespace b200mp {

constexpr int ENC_MAXCOL = 16;
constexpr int TS_K = 56;             // 7 fields x 8 encodings
constexpr float PI_F = 3.14159265358979323846f;

struct EncArgs {
    int64_t rows; int ncol; int C; int S; int tok0;
    b200mp_column col[ENC_MAXCOL];
    const float* cls;
};

// encodings p = field * 4 + q of a timestamp row: e[field * 8 + q] = sin, e[field * 8 + 4 + q] = cos
__device__ __forceinline__ void ts_encode(const int64_t* __restrict__ f, float min_year, int p, float* __restrict__ e) {
    const int fld = p >> 2, q = p & 3;
    const float v = (float)__ldg(f + fld);
    float sn, cs;
    if (fld == 0) {
        const float mult = q == 0 ? 1.f : q == 1 ? 0.1f : q == 2 ? 0.01f : 0.001f;   // (1/10000)^(2q/8)
        sincosf((v - min_year) * mult, &sn, &cs);
    } else {
        const float maxv = fld == 1 ? 12.f : fld == 2 ? 31.f : fld == 3 ? 7.f : fld == 4 ? 24.f : 60.f;
        const float m = (v / maxv) * (float)(q + 1);
        sn = sinf(m * PI_F);
        cs = cosf(m * 2.f * PI_F);
    }
    e[fld * 8 + q] = sn;
    e[fld * 8 + 4 + q] = cs;
}

__global__ void __launch_bounds__(256) col_encode_fwd_kernel(const __grid_constant__ EncArgs a, float* __restrict__ out) {
    __shared__ float enc_s[16][TS_K];
    const int lane = threadIdx.x & 15, grp = threadIdx.x >> 4;
    const int C4 = a.C >> 2;
    // a CTA iteration = 16 rows (one per lane group), tokens one after the other: every warp of the CTA works on the same
    // column kind at the same time (mixing kinds in a warp serialised the 56-term timestamp loop with the one-line cases)
    const int64_t chunks = (a.rows + 15) / 16;
    for (int64_t chunk = blockIdx.x; chunk < chunks; chunk += gridDim.x)
    for (int s = 0; s < a.S; ++s) {
        const int64_t r = chunk * 16 + grp;
        const bool valid = r < a.rows;
        const int64_t item = r * a.S + s;
        const int ci = s - a.tok0;
        const int kind = ci < 0 ? -1 : a.col[ci].kind;
        if (valid && kind == B200MP_COL_TIMESTAMP && lane < 14) {
            const int64_t* f = (const int64_t*)a.col[ci].feat + r * a.col[ci].stride + a.col[ci].offset;
            ts_encode(f, a.col[ci].min_year, lane * 2, enc_s[grp]);
            ts_encode(f, a.col[ci].min_year, lane * 2 + 1, enc_s[grp]);
        }
        __syncwarp();
        if (valid) {
            float* o = out + item * a.C;
            if (kind < 0) {
                for (int c4 = lane; c4 < C4; c4 += 16) st4(o + c4 * 4, ldg4(a.cls + c4 * 4));
            } else if (kind == B200MP_COL_NUMERICAL) {
                const b200mp_column& c = a.col[ci];
                const float z = (__ldg((const float*)c.feat + r * c.stride + c.offset) - c.mean) / c.std;
                for (int c4 = lane; c4 < C4; c4 += 16) {
                    const float4 w = ldg4(c.weight + c4 * 4), b = ldg4(c.bias + c4 * 4);
                    st4(o + c4 * 4, f4(z * w.x + b.x, z * w.y + b.y, z * w.z + b.z, z * w.w + b.w));
                }
            } else if (kind == B200MP_COL_CATEGORICAL) {
                const b200mp_column& c = a.col[ci];
                const int64_t id = __ldg((const int64_t*)c.feat + r * c.stride + c.offset) + 1;
                const float* w = c.weight + id * a.C;
                for (int c4 = lane; c4 < C4; c4 += 16) st4(o + c4 * 4, ldg4(w + c4 * 4));
            } else {
                const b200mp_column& c = a.col[ci];
                for (int c4 = lane; c4 < C4; c4 += 16) {
                    float4 acc = f4(0.f, 0.f, 0.f, 0.f);
                    const float* w = c.weight + c4 * 4;
#pragma unroll 8
                    for (int k = 0; k < TS_K; ++k) {
                        const float e = enc_s[grp][k];
                        const float4 wk = ldg4(w + (int64_t)k * a.C);
                        acc.x = fmaf(e, wk.x, acc.x); acc.y = fmaf(e, wk.y, acc.y);
                        acc.z = fmaf(e, wk.z, acc.z); acc.w = fmaf(e, wk.w, acc.w);
                    }
                    st4(o + c4 * 4, acc + ldg4(c.bias + c4 * 4));
                }
            }
        }
        __syncwarp();
    }
}

struct EncBwdArgs {
    int64_t rows; int C; int S; int tok;      // tok: token slot of this column in g_out [rows, S, C]
    b200mp_column col;
    const float* g_out; float* g_weight; float* g_bias;
};

// numerical column (and, with g_weight == nullptr, the CLS token): sum_r z_r g_r and sum_r g_r
__global__ void __launch_bounds__(256) col_encode_bwd_linear_kernel(const __grid_constant__ EncBwdArgs a) {
    __shared__ float sw[512], sb[512];
    const int lane = threadIdx.x & 15, grp = threadIdx.x >> 4;
    for (int i = threadIdx.x; i < a.C; i += 256) { sw[i] = 0.f; sb[i] = 0.f; }
    __syncthreads();
    const bool has_w = a.g_weight != nullptr;
    for (int c4 = lane; c4 < (a.C >> 2); c4 += 16) {
        float4 aw = f4(0.f, 0.f, 0.f, 0.f), ab = aw;
        for (int64_t r = (int64_t)blockIdx.x * 16 + grp; r < a.rows; r += (int64_t)gridDim.x * 16) {
            const float4 g = ldg4(a.g_out + (r * a.S + a.tok) * a.C + c4 * 4);
            ab = ab + g;
            if (has_w) {
                const float z = (__ldg((const float*)a.col.feat + r * a.col.stride + a.col.offset) - a.col.mean) / a.col.std;
                aw = aw + z * g;
            }
        }
        const float w4[4] = {aw.x, aw.y, aw.z, aw.w}, b4[4] = {ab.x, ab.y, ab.z, ab.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) { atomicAdd(&sb[c4 * 4 + i], b4[i]); if (has_w) atomicAdd(&sw[c4 * 4 + i], w4[i]); }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < a.C; i += 256) {
        atomicAdd(a.g_bias + i, sb[i]);
        if (has_w) atomicAdd(a.g_weight + i, sw[i]);
    }
}

// categorical column: g_table[id + 1] += g_r (row 0, the padding row, gets none).  The table is accumulated in
// shared memory when (ncat + 1) * C floats fit (categories are few: currencies, payment formats), else with vector
// reductions straight into global memory.
constexpr int CAT_SMEM_FLOATS = 10240;
__global__ void __launch_bounds__(256) col_encode_bwd_cat_kernel(const __grid_constant__ EncBwdArgs a) {
    __shared__ float tab[CAT_SMEM_FLOATS];
    const int lane = threadIdx.x & 15, grp = threadIdx.x >> 4;
    const int n = (a.col.ncat + 1) * a.C;
    const bool in_smem = n <= CAT_SMEM_FLOATS;
    if (in_smem) {
        for (int i = threadIdx.x; i < n; i += 256) tab[i] = 0.f;
        __syncthreads();
    }
    for (int64_t r = (int64_t)blockIdx.x * 16 + grp; r < a.rows; r += (int64_t)gridDim.x * 16) {
        const int64_t id = __ldg((const int64_t*)a.col.feat + r * a.col.stride + a.col.offset) + 1;
        if (id == 0) continue;
        for (int c4 = lane; c4 < (a.C >> 2); c4 += 16) {
            const float4 g = ldg4(a.g_out + (r * a.S + a.tok) * a.C + c4 * 4);
            if (in_smem) {
                float* t = tab + id * a.C + c4 * 4;
                atomicAdd(t, g.x); atomicAdd(t + 1, g.y); atomicAdd(t + 2, g.z); atomicAdd(t + 3, g.w);
            } else {
                red_add4(a.g_weight + id * a.C + c4 * 4, g);
            }
        }
    }
    if (in_smem) {
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += 256) if (tab[i] != 0.f) atomicAdd(a.g_weight + i, tab[i]);
    }
}

// timestamp column: g_W[56, C] += enc^T g, g_b += sum_r g.  Thread (lane, kg) owns encodings kg, kg + 16, kg + 32,
// kg + 48 and channels 4 lane .. 4 lane + 3 of each 64-channel chunk.
constexpr int TS_TR = 32, TS_LDE = TS_K + 1;
template <int NCH>
__global__ void __launch_bounds__(256) col_encode_bwd_ts_kernel(const __grid_constant__ EncBwdArgs a) {
    __shared__ float enc_s[TS_TR][TS_LDE];
    __shared__ __align__(16) float g_s[TS_TR * NCH * 64];
    const int lane = threadIdx.x & 15, kg = threadIdx.x >> 4;
    const int C = a.C, C4 = C >> 2;
    float4 acc[NCH][4], accb[NCH];
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) {
        accb[ch] = f4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 4; ++u) acc[ch][u] = f4(0.f, 0.f, 0.f, 0.f);
    }
    for (int64_t r0 = (int64_t)blockIdx.x * TS_TR; r0 < a.rows; r0 += (int64_t)gridDim.x * TS_TR) {
        __syncthreads();
        for (int t = threadIdx.x; t < TS_TR * 28; t += 256) {
            const int rr = t / 28, p = t - rr * 28;
            if (r0 + rr < a.rows)
                ts_encode((const int64_t*)a.col.feat + (r0 + rr) * a.col.stride + a.col.offset, a.col.min_year, p, enc_s[rr]);
            else { enc_s[rr][(p >> 2) * 8 + (p & 3)] = 0.f; enc_s[rr][(p >> 2) * 8 + 4 + (p & 3)] = 0.f; }
        }
        for (int t = threadIdx.x; t < TS_TR * C4; t += 256) {
            const int rr = t / C4, c4 = t - rr * C4;
            float4 g = f4(0.f, 0.f, 0.f, 0.f);
            if (r0 + rr < a.rows) g = ldg4(a.g_out + ((r0 + rr) * a.S + a.tok) * C + c4 * 4);
            st4(g_s + rr * C + c4 * 4, g);
        }
        __syncthreads();
#pragma unroll
        for (int ch = 0; ch < NCH; ++ch) {
            const int c4 = ch * 16 + lane;
            if (c4 < C4) {
#pragma unroll 4
                for (int rr = 0; rr < TS_TR; ++rr) {
                    const float4 g = lds4(g_s + rr * C + c4 * 4);
                    if (kg == 0) accb[ch] = accb[ch] + g;
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int k = kg + 16 * u;
                        if (k < TS_K) acc[ch][u] = acc[ch][u] + enc_s[rr][k] * g;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int ch = 0; ch < NCH; ++ch) {
        const int c4 = ch * 16 + lane;
        if (c4 < C4) {
            if (kg == 0) red_add4(a.g_bias + c4 * 4, accb[ch]);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int k = kg + 16 * u;
                if (k < TS_K) red_add4(a.g_weight + (int64_t)k * C + c4 * 4, acc[ch][u]);
            }
        }
    }
}

static int enc_grid(int64_t work_per_cta_units) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t cap = (int64_t)sms * 4;
    return (int)(work_per_cta_units < 1 ? 1 : work_per_cta_units < cap ? work_per_cta_units : cap);
}

static bool column_ok(const b200mp_column& c) {
    if (!c.feat || !c.weight || c.stride < 1 || c.offset < 0) return false;
    switch (c.kind) {
        case B200MP_COL_NUMERICAL: return c.bias != nullptr && c.std != 0.f;
        case B200MP_COL_CATEGORICAL: return c.ncat >= 0;
        case B200MP_COL_TIMESTAMP: return c.bias != nullptr;
        default: return false;
    }
}

}  // namespace b200mp

extern "C" int b200mp_col_encode_fwd(const b200mp_column* cols, int ncol, const float* cls, int64_t rows, int C, float* out,
                                     void* stream) {
    using namespace b200mp;
    if (rows < 0 || ncol < 0 || ncol > ENC_MAXCOL || C < 4 || C % 4 || (ncol && !cols)) return B200MP_EINVAL;
    EncArgs a{};
    a.rows = rows; a.ncol = ncol; a.C = C; a.tok0 = cls ? 1 : 0; a.S = ncol + a.tok0; a.cls = cls;
    if (a.S == 0 || rows == 0) return 0;
    if (!out) return B200MP_EINVAL;
    for (int i = 0; i < ncol; ++i) {
        if (!column_ok(cols[i])) return B200MP_EINVAL;
        a.col[i] = cols[i];
    }
    col_encode_fwd_kernel<<<enc_grid((rows + 15) / 16), 256, 0, (cudaStream_t)stream>>>(a, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_col_encode_bwd(const b200mp_column* cols, int ncol, int has_cls, int64_t rows, int C, const float* g_out,
                                     float* const* g_weight, float* const* g_bias, float* g_cls, void* stream) {
    using namespace b200mp;
    if (rows < 0 || ncol < 0 || ncol > ENC_MAXCOL || C < 4 || C % 4 || C > 512 || (ncol && (!cols || !g_weight || !g_bias)))
        return B200MP_EINVAL;
    const int tok0 = has_cls ? 1 : 0, S = ncol + tok0;
    if (S == 0 || rows == 0) return 0;
    if (!g_out || (has_cls && !g_cls)) return B200MP_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (has_cls) {
        EncBwdArgs a{rows, C, S, 0, b200mp_column{}, g_out, nullptr, g_cls};
        col_encode_bwd_linear_kernel<<<enc_grid((rows + 15) / 16), 256, 0, st>>>(a);
        B200MP_LAUNCH_CHECK();
    }
    for (int i = 0; i < ncol; ++i) {
        if (!column_ok(cols[i]) || !g_weight[i]) return B200MP_EINVAL;
        EncBwdArgs a{rows, C, S, tok0 + i, cols[i], g_out, g_weight[i], g_bias[i]};
        switch (cols[i].kind) {
            case B200MP_COL_NUMERICAL:
                if (!g_bias[i]) return B200MP_EINVAL;
                col_encode_bwd_linear_kernel<<<enc_grid((rows + 15) / 16), 256, 0, st>>>(a);
                break;
            case B200MP_COL_CATEGORICAL:
                col_encode_bwd_cat_kernel<<<enc_grid((rows + 15) / 16), 256, 0, st>>>(a);
                break;
            default:
                if (!g_bias[i] || C > 128) return B200MP_EINVAL;
                if (C <= 64) col_encode_bwd_ts_kernel<1><<<enc_grid((rows + TS_TR - 1) / TS_TR), 256, 0, st>>>(a);
                else col_encode_bwd_ts_kernel<2><<<enc_grid((rows + TS_TR - 1) / TS_TR), 256, 0, st>>>(a);
        }
        B200MP_LAUNCH_CHECK();
    }
    return 0;
}

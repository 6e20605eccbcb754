// Node-level fused elementwise kernels around the PNA aggregation (HBM-bound, one pass each):
//   b200mp_pna_scale_fwd / _bwd : degree scalers applied AFTER the post GEMM,
//        out = base + sum_s f_s(deg) * Z[:, s]      (PyG DegreeScalerAggregation + post_nn, restated
//        with the scalers commuted past the Linear: SURVEY.md App. A.2; pna.py:71-79)
//   b200mp_pna_coef : per-node coefficient rows of the message backward (edge_bwd.cu) from the
//        aggregate gradients, in one pass instead of ~15 tensor ops.
#include "common.cuh"

namespace b200mp {

// scaler codes: 0 identity, 1 amplification, 2 attenuation, 3 linear, 4 inverse_linear
__device__ __forceinline__ float scaler_value(int code, float d, float avg_lin, float avg_log) {
    switch (code) {
        case 1: return logf(d + 1.f) / avg_log;
        case 2: return avg_log / logf(fmaxf(d, 1.f) + 1.f);
        case 3: return d / avg_lin;
        case 4: return avg_lin / fmaxf(d, 1.f);
    }
    return 1.f;
}

struct ScalerCodes { int n; int code[5]; };

template <bool BWD>
__global__ void pna_scale_kernel(int64_t N, int c4 /* F/4 */, const int64_t* __restrict__ ptr, const float* __restrict__ base_or_g,
                                 const float* __restrict__ Z, ScalerCodes sc, const float* __restrict__ avg_lin_p,
                                 const float* __restrict__ avg_log_p, float* __restrict__ out) {
    const float avg_lin = __ldg(avg_lin_p), avg_log = __ldg(avg_log_p);
    const int64_t total = N * c4;
    const int F = c4 * 4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = i / c4;
        const int c = (int)(i % c4) * 4;
        const float d = (float)(__ldg(ptr + v + 1) - __ldg(ptr + v));
        if (!BWD) {
            float4 o = ldg4(base_or_g + v * F + c);
            for (int s = 0; s < sc.n; ++s)
                o = o + scaler_value(sc.code[s], d, avg_lin, avg_log) * ldg4(Z + (v * sc.n + s) * F + c);
            st4(out + v * F + c, o);
        } else {
            const float4 g = ldg4(base_or_g + v * F + c);
            for (int s = 0; s < sc.n; ++s) st4(out + (v * sc.n + s) * F + c, scaler_value(sc.code[s], d, avg_lin, avg_log) * g);
        }
    }
}

// coef[N,7H] = [c0 | c1 | gmin' | min | gmax' | max | mean],  gA[N,H] = (g_mean + g_min + g_max) [deg > 0]
__global__ void pna_coef_kernel(int64_t N, int c4, const int64_t* __restrict__ ptr, const float* __restrict__ g_agg,
                                const float* __restrict__ agg, const float* __restrict__ tie, float* __restrict__ coef,
                                float* __restrict__ gA) {
    const int64_t total = N * c4;
    const int H = c4 * 4;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = i / c4;
        const int c = (int)(i % c4) * 4;
        const float deg = (float)(__ldg(ptr + v + 1) - __ldg(ptr + v));
        const float dn = fmaxf(deg, 1.f);
        const float* ga = g_agg + v * 4 * H + c;
        const float* ag = agg + v * 4 * H + c;
        const float4 g_mean = ldg4(ga), g_min = ldg4(ga + H), g_max = ldg4(ga + 2 * H), g_std = ldg4(ga + 3 * H);
        const float4 mean = ldg4(ag), mn = ldg4(ag + H), mx = ldg4(ag + 2 * H), sd = ldg4(ag + 3 * H);
        const float4 cn = ldg4(tie + v * 2 * H + c), cx = ldg4(tie + v * 2 * H + H + c);
        auto c1f = [dn](float gs, float s) { return s > 0.f ? gs / (dn * s) : 0.f; };
        float* o = coef + v * 7 * H + c;
        st4(o, f4(g_mean.x / dn, g_mean.y / dn, g_mean.z / dn, g_mean.w / dn));
        st4(o + H, f4(c1f(g_std.x, sd.x), c1f(g_std.y, sd.y), c1f(g_std.z, sd.z), c1f(g_std.w, sd.w)));
        st4(o + 2 * H, f4(g_min.x / fmaxf(cn.x, 1.f), g_min.y / fmaxf(cn.y, 1.f), g_min.z / fmaxf(cn.z, 1.f), g_min.w / fmaxf(cn.w, 1.f)));
        st4(o + 3 * H, mn);
        st4(o + 4 * H, f4(g_max.x / fmaxf(cx.x, 1.f), g_max.y / fmaxf(cx.y, 1.f), g_max.z / fmaxf(cx.z, 1.f), g_max.w / fmaxf(cx.w, 1.f)));
        st4(o + 5 * H, mx);
        st4(o + 6 * H, mean);
        if (gA) {
            const float has = deg > 0.f ? 1.f : 0.f;
            st4(gA + v * H + c, has * (g_mean + g_min + g_max));
        }
    }
}

static inline int node_grid(int64_t n) {
    int64_t g = (n + 255) / 256, cap = (int64_t)device_info().sms * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace b200mp

using namespace b200mp;

static int make_codes(int ns, const int* codes, ScalerCodes& sc) {
    if (ns < 1 || ns > 5 || !codes) return B200MP_EINVAL;
    sc.n = ns;
    for (int i = 0; i < ns; ++i) {
        if (codes[i] < 0 || codes[i] > 4) return B200MP_EINVAL;
        sc.code[i] = codes[i];
    }
    return 0;
}

extern "C" int b200mp_pna_scale_fwd(int F, int64_t N, const int64_t* ptr, const float* base, const float* Z, int ns,
                                    const int* host_codes, const float* avg_lin, const float* avg_log, float* out, void* stream) {
    ScalerCodes sc;
    if (make_codes(ns, host_codes, sc) || N < 0 || F <= 0 || (F & 3) || !ptr || !avg_lin || !avg_log) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!base || !Z || !out) return B200MP_EINVAL;
    pna_scale_kernel<false><<<node_grid(N * (F / 4)), 256, 0, (cudaStream_t)stream>>>(N, F / 4, ptr, base, Z, sc, avg_lin, avg_log, out);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_pna_scale_bwd(int F, int64_t N, const int64_t* ptr, const float* g, int ns, const int* host_codes,
                                    const float* avg_lin, const float* avg_log, float* gZ, void* stream) {
    ScalerCodes sc;
    if (make_codes(ns, host_codes, sc) || N < 0 || F <= 0 || (F & 3) || !ptr || !avg_lin || !avg_log) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!g || !gZ) return B200MP_EINVAL;
    pna_scale_kernel<true><<<node_grid(N * (F / 4)), 256, 0, (cudaStream_t)stream>>>(N, F / 4, ptr, g, nullptr, sc, avg_lin, avg_log, gZ);
    B200MP_LAUNCH_CHECK();
    return 0;
}

extern "C" int b200mp_pna_coef(int H, int64_t N, const int64_t* ptr, const float* g_agg, const float* agg, const float* tie,
                               float* coef, float* gA, void* stream) {
    if (N < 0 || H <= 0 || (H & 3) || !ptr) return B200MP_EINVAL;
    if (N == 0) return 0;
    if (!g_agg || !agg || !tie || !coef) return B200MP_EINVAL;
    pna_coef_kernel<<<node_grid(N * (H / 4)), 256, 0, (cudaStream_t)stream>>>(N, H / 4, ptr, g_agg, agg, tie, coef, gA);
    B200MP_LAUNCH_CHECK();
    return 0;
}

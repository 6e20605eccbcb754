// Segmented reductions over messages stored in segment order (b200mp_segment_reduce_fwd).
// No atomics: a warp owns a whole light segment (H/4 lanes per row, several rows per warp
// iteration, shuffle merge); heavy segments (hubs) are cut into chunks reduced by one CTA each
// into partial statistics that a third kernel merges in chunk order (deterministic).
//
// Restates PyG 2.5.3 DegreeScalerAggregation's aggregators for 3-D messages:
//   mean = sum / max(cnt,1);  min/max = scatter_reduce(amin/amax, include_self=False) -> 0 on
//   empty segments;  std = sqrt(clamp(E[m^2]-E[m]^2, 1e-5)) masked to 0 at <= sqrt(1e-5).
// and GINEConv's 'add' aggregation.  Called for /root/reference/src/nn/gnn/pna.py:71-74,96 and
// gine.py:66-70,87.
#include "umma_edge.cuh"      // ldg_e4: float | bf16 message rows
#include <math_constants.h>
#include <cstdlib>

namespace b200mp {
using ue::bf16;
using ue::ldg_e4;

#ifndef SEG_MINB
#define SEG_MINB 4
#endif
constexpr float STD_EPS = 1e-5f;
constexpr float STD_THR = 0.0031622776601683794f;   // (float)math.sqrt(1e-5)

template <bool STATS>
struct Acc {
    float4 s, q, mn, cn, mx, cx;
    __device__ __forceinline__ void init() {
        s = q = cn = cx = f4(0.f, 0.f, 0.f, 0.f);
        mn = f4(CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F);
        mx = f4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
    }
    static __device__ __forceinline__ void add1(float v, float& s, float& q, float& mn, float& cn, float& mx, float& cx) {
        s += v;
        if (STATS) {                       // branch free: the lanes of a warp hold different columns and would diverge
            q = fmaf(v, v, q);
            const float nmn = fminf(mn, v), nmx = fmaxf(mx, v);
            cn = (v == nmn) ? ((nmn == mn) ? cn + 1.f : 1.f) : cn;
            cx = (v == nmx) ? ((nmx == mx) ? cx + 1.f : 1.f) : cx;
            mn = nmn;
            mx = nmx;
        }
    }
    __device__ __forceinline__ void add(float4 v) {
        add1(v.x, s.x, q.x, mn.x, cn.x, mx.x, cx.x);
        add1(v.y, s.y, q.y, mn.y, cn.y, mx.y, cx.y);
        add1(v.z, s.z, q.z, mn.z, cn.z, mx.z, cx.z);
        add1(v.w, s.w, q.w, mn.w, cn.w, mx.w, cx.w);
    }
    static __device__ __forceinline__ void merge1(float& s, float& q, float& mn, float& cn, float& mx, float& cx,
                                                  float os, float oq, float omn, float ocn, float omx, float ocx) {
        s += os;
        if (STATS) {
            q += oq;
            const float nmn = fminf(mn, omn), nmx = fmaxf(mx, omx);
            cn = (omn == nmn ? ocn : 0.f) + (mn == nmn ? cn : 0.f);
            cx = (omx == nmx ? ocx : 0.f) + (mx == nmx ? cx : 0.f);
            mn = nmn;
            mx = nmx;
        }
    }
    __device__ __forceinline__ void merge(const Acc& o) {
        merge1(s.x, q.x, mn.x, cn.x, mx.x, cx.x, o.s.x, o.q.x, o.mn.x, o.cn.x, o.mx.x, o.cx.x);
        merge1(s.y, q.y, mn.y, cn.y, mx.y, cx.y, o.s.y, o.q.y, o.mn.y, o.cn.y, o.mx.y, o.cx.y);
        merge1(s.z, q.z, mn.z, cn.z, mx.z, cx.z, o.s.z, o.q.z, o.mn.z, o.cn.z, o.mx.z, o.cx.z);
        merge1(s.w, q.w, mn.w, cn.w, mx.w, cx.w, o.s.w, o.q.w, o.mn.w, o.cn.w, o.mx.w, o.cx.w);
    }
};

__device__ __forceinline__ float4 shfl_xor4(float4 v, int off) {
    return f4(__shfl_xor_sync(0xffffffffu, v.x, off), __shfl_xor_sync(0xffffffffu, v.y, off),
              __shfl_xor_sync(0xffffffffu, v.z, off), __shfl_xor_sync(0xffffffffu, v.w, off));
}

// Merge the row groups of a warp (lanes with equal lane % LPR hold the same columns).
template <bool STATS, int LPR>
__device__ __forceinline__ void warp_merge(Acc<STATS>& a) {
#pragma unroll
    for (int off = LPR; off < 32; off <<= 1) {
        Acc<STATS> o;
        o.s = shfl_xor4(a.s, off);
        if (STATS) {
            o.q = shfl_xor4(a.q, off); o.mn = shfl_xor4(a.mn, off); o.cn = shfl_xor4(a.cn, off);
            o.mx = shfl_xor4(a.mx, off); o.cx = shfl_xor4(a.cx, off);
        }
        a.merge(o);
    }
}

// rows [begin, end) of m, this lane's 4 columns starting at c; rows are strided by `stride`
template <bool STATS, int H, typename T>
__device__ __forceinline__ void accumulate_rows(Acc<STATS>& a, const T* __restrict__ m, int64_t first, int64_t end,
                                                int stride, int c) {
    int64_t r = first;
    for (; r + 3 * (int64_t)stride < end; r += 4 * (int64_t)stride) {      // 4 independent loads in flight
        float4 v0 = ldg_e4(m + r * H + c);
        float4 v1 = ldg_e4(m + (r + stride) * H + c);
        float4 v2 = ldg_e4(m + (r + 2 * (int64_t)stride) * H + c);
        float4 v3 = ldg_e4(m + (r + 3 * (int64_t)stride) * H + c);
        a.add(v0); a.add(v1); a.add(v2); a.add(v3);
    }
    for (; r < end; r += stride) a.add(ldg_e4(m + r * H + c));
}

template <bool STATS, int H>
__device__ __forceinline__ void finalize_store(const Acc<STATS>& a, int64_t deg, int64_t node, int c,
                                               float* __restrict__ agg, float* __restrict__ tie) {
    if (!STATS) {
        st4(agg + node * H + c, a.s);      // empty segment: s == 0
        return;
    }
    float* o = agg + node * 4 * H + c;
    float* t = tie + node * 2 * H + c;
    if (deg == 0) {
        float4 z = f4(0.f, 0.f, 0.f, 0.f);
        st4(o, z); st4(o + H, z); st4(o + 2 * H, z); st4(o + 3 * H, z); st4(t, z); st4(t + H, z);
        return;
    }
    const float d = (float)deg;
    float4 mean = f4(a.s.x / d, a.s.y / d, a.s.z / d, a.s.w / d);
    float4 m2 = f4(a.q.x / d, a.q.y / d, a.q.z / d, a.q.w / d);
    auto sd = [](float mean2, float mu) {
        float var = mean2 - mu * mu;
        float s = sqrtf(fmaxf(var, STD_EPS));
        return s <= STD_THR ? 0.f : s;
    };
    st4(o, mean);
    st4(o + H, a.mn);
    st4(o + 2 * H, a.mx);
    st4(o + 3 * H, f4(sd(m2.x, mean.x), sd(m2.y, mean.y), sd(m2.z, mean.z), sd(m2.w, mean.w)));
    st4(t, a.cn);
    st4(t + H, a.cx);
}

// Two-pass statistics of a light segment (TWOPASS): sums and extrema first (4 instructions per element instead of the 14 of the
// branch-free tie bookkeeping in Acc::add1), then the tie counts against the FINAL extrema over the same rows -- which the first
// pass has just pulled through L1 (a 15-row segment is 3.8 KB).  Same summation order, same extrema, integer-valued counts: the
// results are bit-identical to the one-pass form.
__device__ __forceinline__ float4 min4(float4 a, float4 b) { return f4(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z), fminf(a.w, b.w)); }
__device__ __forceinline__ float4 max4(float4 a, float4 b) { return f4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w)); }
__device__ __forceinline__ float4 fma4(float4 a, float4 b, float4 c) { return f4(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z), fmaf(a.w, b.w, c.w)); }
__device__ __forceinline__ float4 eq4(float4 v, float4 r, float4 c) {
    return f4(c.x + (v.x == r.x ? 1.f : 0.f), c.y + (v.y == r.y ? 1.f : 0.f), c.z + (v.z == r.z ? 1.f : 0.f), c.w + (v.w == r.w ? 1.f : 0.f));
}
template <int H, int LPR, typename T>
__device__ __forceinline__ void two_pass_stats(Acc<true>& a, const T* __restrict__ m, int64_t first, int64_t end, int stride, int c) {
    float4 s = f4(0.f, 0.f, 0.f, 0.f), q = s;
    float4 mn = f4(CUDART_INF_F, CUDART_INF_F, CUDART_INF_F, CUDART_INF_F), mx = f4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
    int64_t r = first;
    for (; r + 3 * (int64_t)stride < end; r += 4 * (int64_t)stride) {
        const float4 v0 = ldg_e4(m + r * H + c), v1 = ldg_e4(m + (r + stride) * H + c);
        const float4 v2 = ldg_e4(m + (r + 2 * (int64_t)stride) * H + c), v3 = ldg_e4(m + (r + 3 * (int64_t)stride) * H + c);
        s = s + v0; q = fma4(v0, v0, q); mn = min4(mn, v0); mx = max4(mx, v0);
        s = s + v1; q = fma4(v1, v1, q); mn = min4(mn, v1); mx = max4(mx, v1);
        s = s + v2; q = fma4(v2, v2, q); mn = min4(mn, v2); mx = max4(mx, v2);
        s = s + v3; q = fma4(v3, v3, q); mn = min4(mn, v3); mx = max4(mx, v3);
    }
    for (; r < end; r += stride) {
        const float4 v = ldg_e4(m + r * H + c);
        s = s + v; q = fma4(v, v, q); mn = min4(mn, v); mx = max4(mx, v);
    }
#pragma unroll
    for (int off = LPR; off < 32; off <<= 1) {
        s = s + shfl_xor4(s, off); q = q + shfl_xor4(q, off);
        mn = min4(mn, shfl_xor4(mn, off)); mx = max4(mx, shfl_xor4(mx, off));
    }
    float4 cn = f4(0.f, 0.f, 0.f, 0.f), cx = cn;
    for (r = first; r + (int64_t)stride < end; r += 2 * (int64_t)stride) {
        const float4 v0 = ldg_e4(m + r * H + c), v1 = ldg_e4(m + (r + stride) * H + c);
        cn = eq4(v0, mn, cn); cx = eq4(v0, mx, cx);
        cn = eq4(v1, mn, cn); cx = eq4(v1, mx, cx);
    }
    for (; r < end; r += stride) {
        const float4 v = ldg_e4(m + r * H + c);
        cn = eq4(v, mn, cn); cx = eq4(v, mx, cx);
    }
#pragma unroll
    for (int off = LPR; off < 32; off <<= 1) { cn = cn + shfl_xor4(cn, off); cx = cx + shfl_xor4(cx, off); }
    a.s = s; a.q = q; a.mn = mn; a.mx = mx; a.cn = cn; a.cx = cx;
}

// ---- light segments: one warp per node
template <bool STATS, int H, typename T, bool TWOPASS = false>
__global__ void __launch_bounds__(NT, SEG_MINB) seg_light_kernel(int64_t N, const int64_t* __restrict__ ptr, const T* __restrict__ m,
                                                       float* __restrict__ agg, float* __restrict__ tie, int64_t light_max) {
    constexpr int LPR = H / 4, RPW = 32 / LPR;
    const int lane = threadIdx.x & 31;
    const int g = lane / LPR, c = (lane % LPR) * 4;
    const int64_t warps = (int64_t)gridDim.x * (NT / 32);
    int64_t node = (int64_t)blockIdx.x * (NT / 32) + (threadIdx.x >> 5);
    int64_t nb = 0, ne = 0;
    if (node < N) { nb = __ldg(ptr + node); ne = __ldg(ptr + node + 1); }
    for (; node < N; node += warps) {
        const int64_t begin = nb, end = ne;
        if (node + warps < N) { nb = __ldg(ptr + node + warps); ne = __ldg(ptr + node + warps + 1); }   // next segment's bounds in flight
        if (end - begin > light_max) continue;          // handled by the heavy path
        Acc<STATS> a;
        if constexpr (STATS && TWOPASS) {
            two_pass_stats<H, LPR>(a, m, begin + g, end, RPW, c);
        } else {
            a.init();
            accumulate_rows<STATS, H>(a, m, begin + g, end, RPW, c);
            warp_merge<STATS, LPR>(a);
        }
        if (g == 0) finalize_store<STATS, H>(a, end - begin, node, c, agg, tie);
    }
}

// ---- heavy segments: one CTA per chunk -> partial [6][H] (sum, sumsq, min, cmin, max, cmax)
template <bool STATS, int H, typename T>
__global__ void __launch_bounds__(NT) seg_heavy_partial_kernel(const int64_t* __restrict__ ptr, const T* __restrict__ m,
                                                               const int32_t* __restrict__ heavy_node,
                                                               const int32_t* __restrict__ chunk_heavy,
                                                               const int64_t* __restrict__ chunk_begin, int64_t chunk,
                                                               float* __restrict__ partial) {
    constexpr int LPR = H / 4, RPW = 32 / LPR, NW = NT / 32;
    __shared__ float4 sm[NW][LPR][6];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int g = lane / LPR, cl = lane % LPR, c = cl * 4;
    const int64_t ch = blockIdx.x;
    const int64_t node = __ldg(heavy_node + __ldg(chunk_heavy + ch));
    const int64_t begin = __ldg(chunk_begin + ch);
    const int64_t end = min64(begin + chunk, __ldg(ptr + node + 1));
    Acc<STATS> a;
    a.init();
    accumulate_rows<STATS, H>(a, m, begin + w * RPW + g, end, RPW * NW, c);
    warp_merge<STATS, LPR>(a);
    if (g == 0) {
        sm[w][cl][0] = a.s; sm[w][cl][1] = a.q; sm[w][cl][2] = a.mn;
        sm[w][cl][3] = a.cn; sm[w][cl][4] = a.mx; sm[w][cl][5] = a.cx;
    }
    __syncthreads();
    if (threadIdx.x < LPR) {
        Acc<STATS> t;
        t.init();
        for (int i = 0; i < NW; ++i) {
            Acc<STATS> o;
            o.s = sm[i][threadIdx.x][0]; o.q = sm[i][threadIdx.x][1]; o.mn = sm[i][threadIdx.x][2];
            o.cn = sm[i][threadIdx.x][3]; o.mx = sm[i][threadIdx.x][4]; o.cx = sm[i][threadIdx.x][5];
            t.merge(o);
        }
        float* p = partial + ch * 6 * H + threadIdx.x * 4;
        st4(p, t.s); st4(p + H, t.q); st4(p + 2 * H, t.mn); st4(p + 3 * H, t.cn); st4(p + 4 * H, t.mx); st4(p + 5 * H, t.cx);
    }
}

// ---- heavy segments: merge the chunk partials of one node in chunk order
template <bool STATS, int H>
__global__ void seg_heavy_combine_kernel(const int64_t* __restrict__ ptr, const int32_t* __restrict__ heavy_node,
                                         const int32_t* __restrict__ heavy_chunk_ptr, const float* __restrict__ partial,
                                         float* __restrict__ agg, float* __restrict__ tie) {
    constexpr int LPR = H / 4;
    if (threadIdx.x >= LPR) return;
    const int64_t hn = blockIdx.x;
    const int64_t node = __ldg(heavy_node + hn);
    const int c = threadIdx.x * 4;
    Acc<STATS> t;
    t.init();
    for (int64_t ch = __ldg(heavy_chunk_ptr + hn); ch < __ldg(heavy_chunk_ptr + hn + 1); ++ch) {
        const float* p = partial + ch * 6 * H + c;
        Acc<STATS> o;
        o.s = ldg4(p); o.q = ldg4(p + H); o.mn = ldg4(p + 2 * H); o.cn = ldg4(p + 3 * H); o.mx = ldg4(p + 4 * H); o.cx = ldg4(p + 5 * H);
        t.merge(o);
    }
    finalize_store<STATS, H>(t, __ldg(ptr + node + 1) - __ldg(ptr + node), node, c, agg, tie);
}

template <bool STATS, int H, typename T = float>
static int launch_reduce(int64_t N, const int64_t* ptr, const T* m, float* agg, float* tie, int64_t light_max,
                         int64_t chunk, int64_t n_heavy, const int32_t* heavy_node, const int32_t* heavy_chunk_ptr,
                         int64_t n_chunks, const int32_t* chunk_heavy, const int64_t* chunk_begin, float* partial,
                         cudaStream_t st) {
    if (N == 0) return 0;
    const int64_t want = (N + NT / 32 - 1) / (NT / 32);
    const int grid = (int)min64(want, (int64_t)device_info().sms * 32);
    // two-pass statistics by default (same-box A/B on hi_medium, two pairs: 3.53 -> 2.93 ms per launch, bit-identical results);
    // B200MP_SEG_TWOPASS=0 selects the one-pass bookkeeping
    const char* tp = getenv("B200MP_SEG_TWOPASS");
    const bool two_pass = !(tp && tp[0] == '0');
    if (STATS && two_pass) seg_light_kernel<STATS, H, T, true><<<grid, NT, 0, st>>>(N, ptr, m, agg, tie, light_max);
    else seg_light_kernel<STATS, H, T><<<grid, NT, 0, st>>>(N, ptr, m, agg, tie, light_max);
    B200MP_LAUNCH_CHECK();
    if (n_heavy > 0) {
        if (!heavy_node || !heavy_chunk_ptr || !chunk_heavy || !chunk_begin || !partial || n_chunks <= 0) return B200MP_EINVAL;
        seg_heavy_partial_kernel<STATS, H, T><<<(unsigned)n_chunks, NT, 0, st>>>(ptr, m, heavy_node, chunk_heavy, chunk_begin, chunk, partial);
        B200MP_LAUNCH_CHECK();
        seg_heavy_combine_kernel<STATS, H><<<(unsigned)n_heavy, 32, 0, st>>>(ptr, heavy_node, heavy_chunk_ptr, partial, agg, tie);
        B200MP_LAUNCH_CHECK();
    }
    return 0;
}

}  // namespace b200mp

extern "C" int b200mp_segment_reduce_fwd(int kind, int H, int64_t N, const int64_t* ptr, const float* m,
                                         float* agg, float* tie, int64_t light_max, int64_t chunk,
                                         int64_t n_heavy, const int32_t* heavy_node, const int32_t* heavy_chunk_ptr,
                                         int64_t n_chunks, const int32_t* chunk_heavy, const int64_t* chunk_begin,
                                         float* partial, void* stream) {
    using namespace b200mp;
    if (N < 0 || !ptr || !agg || (kind == 0 && !tie) || light_max < 0 || chunk <= 0) return B200MP_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
#define B200MP_RED(STATS, HH) \
    return launch_reduce<STATS, HH>(N, ptr, m, agg, tie, light_max, chunk, n_heavy, heavy_node, heavy_chunk_ptr, n_chunks, chunk_heavy, chunk_begin, partial, st)
    if (kind == 0) {
        switch (H) { case 16: B200MP_RED(true, 16); case 32: B200MP_RED(true, 32); case 64: B200MP_RED(true, 64); case 128: B200MP_RED(true, 128); }
    } else if (kind == 1) {
        switch (H) { case 16: B200MP_RED(false, 16); case 32: B200MP_RED(false, 32); case 64: B200MP_RED(false, 64); case 128: B200MP_RED(false, 128); }
    }
#undef B200MP_RED
    return B200MP_EINVAL;
}

/* the same reductions over bf16-stored messages m [E,64] (fp32 statistics and outputs) */
extern "C" int b200mp_segment_reduce_fwd_bf16(int kind, int64_t N, const int64_t* ptr, const void* m, float* agg, float* tie,
                                              int64_t light_max, int64_t chunk, int64_t n_heavy, const int32_t* heavy_node,
                                              const int32_t* heavy_chunk_ptr, int64_t n_chunks, const int32_t* chunk_heavy,
                                              const int64_t* chunk_begin, float* partial, void* stream) {
    using namespace b200mp;
    if (N < 0 || !ptr || !agg || (kind == 0 && !tie) || light_max < 0 || chunk <= 0 || (kind != 0 && kind != 1)) return B200MP_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const bf16* mb = (const bf16*)m;
    if (kind == 0)
        return launch_reduce<true, 64, bf16>(N, ptr, mb, agg, tie, light_max, chunk, n_heavy, heavy_node, heavy_chunk_ptr, n_chunks, chunk_heavy,
                                             chunk_begin, partial, st);
    return launch_reduce<false, 64, bf16>(N, ptr, mb, agg, tie, light_max, chunk, n_heavy, heavy_node, heavy_chunk_ptr, n_chunks, chunk_heavy,
                                          chunk_begin, partial, st);
}

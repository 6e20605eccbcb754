// Tensor-core (tcgen05, 3xTF32) backward kernels for H = 64: edge-update backward and message
// backward.  Same math and outputs as the SIMT kernels in edge_bwd.cu (which remain the path for
// H = 16 / 32 / 128); see b200mp.h for the contract and edge_bwd.cu for the gradient rules.
//
// Per tile of 128 edges the per-edge GEMMs run on the tensor cores with accumulators in tensor
// memory; the two weight gradients accumulate in tensor memory across ALL tiles of the CTA and
// are written once as a per-CTA partial.
#include "umma_edge.cuh"

namespace b200mp {
using namespace ue;

// ------------------------------------------------------------------ edge-update backward
//   Gh  = (beta g W2) * [h > 0]            U    = g W2            (data GEMM, tmem cols   0.. 63)
//   g_e = alpha g + Gh W1                  ACC2 = alpha g + Gh W1 (data GEMM, tmem cols  64..127)
//   gW2 = beta g^T h                       GW2                    (wgrad,     tmem cols 128..191)
//   gW1 = Gh^T e                           GW1                    (wgrad,     tmem cols 192..255)
struct UpdBwdUArgs {
    int64_t E;
    const float* g_out; const float* h; const float* e;
    const int32_t* src; const int32_t* dst;
    const float* W1; const float* W2; float alpha, beta;
    float* g_e; float* gP; float* gQ;
    float* gW1_part; float* gb1_part; float* gW2_part; float* gb2_part;
};

constexpr int NTB = 512;                  // threads per CTA of the backward kernels (one CTA per SM)
constexpr int CWB = HH / (NTB / TE);      // 16 columns per epilogue thread

// g_e rows of a finished tile: ACC2 (tensor memory) -> global.  Runs one tile late, while the next
// tile's phase A MMAs execute.
__device__ __forceinline__ void upd_bwd_store_ge(float* __restrict__ g_e, uint32_t T_ACC2, int64_t row0, int nrows, int t, int ch) {
    float v[16];
    tmem_ld16(T_ACC2, v);
    if (t < nrows) {
        float* o = g_e + (row0 + t) * HH + ch * CWB;
#pragma unroll
        for (int c = 0; c < 16; c += 4) st4(o + c, f4(v[c], v[c + 1], v[c + 2], v[c + 3]));
    }
}

__global__ void __launch_bounds__(NTB, 1) edge_update_bwd_umma_kernel(UpdBwdUArgs a) {
    constexpr int NTH = NTB, CW = CWB;
    static_assert(CW == 16, "epilogues below read one 16-column tensor-memory group per thread");
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;   // keeps the shared address space
    float* G_hi = smem;                       // g tile, later the e tile
    float* G_lo = G_hi + RT::FLOATS;
    float* H_hi = G_lo + RT::FLOATS;          // h tile, later the Gh tile
    float* H_lo = H_hi + RT::FLOATS;
    float* W2_hi = H_lo + RT::FLOATS;         // B[n=k][K=j] = W2[j][k]   (U = g W2)
    float* W2_lo = W2_hi + WT::FLOATS;
    float* W1_hi = W2_lo + WT::FLOATS;        // B[n=k][K=j] = W1[j][k]   (g_e = Gh W1)
    float* W1_lo = W1_hi + WT::FLOATS;
    float* bsum = W1_lo + WT::FLOATS;         // [2][NTH / 32 warps][32] bias-gradient partials (kernel end)
    __shared__ TileSync sync;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int t = tid & (TE - 1), ch = tid >> 7;
    if (warp == 0) tmem_alloc(&sync.tmem_base, 256);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_init(&sync.bar[1], 1); mbar_fence_init(); }
    load_weight_tiles<true, NTH>(W2_hi, W2_lo, a.W2, tid);
    load_weight_tiles<true, NTH>(W1_hi, W1_lo, a.W1, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t T_GW2 = tm + 128, T_GW1 = tm + 192;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t T_U = tm + lane_off + ch * CW, T_ACC2 = tm + 64 + lane_off + ch * CW;

    float accb1 = 0.f, accb2 = 0.f;           // lane L: column CW ch + L % CW summed over this warp's rows
    const int64_t ntiles = (a.E + TE - 1) / TE;
    uint32_t it = 0;
    RowRegs<NTH> R, R2;
    int nrows_pf = 0, drow_pf = 0, srow_pf = 0;
    if ((int64_t)blockIdx.x < ntiles) {
        nrows_pf = (int)min64(TE, a.E - (int64_t)blockIdx.x * TE);
        load_rows(R, a.g_out, nullptr, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        load_rows(R2, a.h, nullptr, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        drow_pf = __ldg(a.dst + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
        srow_pf = __ldg(a.src + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
    }
    int64_t prev_row0 = 0;
    int prev_nrows = 0;

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE;
        const int nrows = nrows_pf;
        const uint32_t par = it & 1;
        const bool valid = t < nrows;

        // ---- phase A: g and h tiles (prefetched) -> shared memory; U = g W2; GW2 += g^T h
        // (the previous tile's phase B MMAs, which read these buffers, were waited for below)
        store_rows_split(R, G_hi, G_lo, nrows, tid);
        store_rows_split(R2, H_hi, H_lo, nrows, tid);
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, G_hi, G_lo, W2_hi, W2_lo, 0);
            gemm3_wgrad(T_GW2, G_hi, G_lo, H_hi, H_lo, it > 0);
            mma_commit(&sync.bar[0]);
        }
        load_rows(R, a.e, nullptr, row0, nrows, tid);                 // e tile in flight during the MMAs
        const int drow = valid ? drow_pf : -1;
        const int srow = srow_pf;
        // the previous tile's g_e rows leave ACC2 while this tile's phase A runs on the tensor core
        if (it > 0) upd_bwd_store_ge(a.g_e, T_ACC2, prev_row0, prev_nrows, t, ch);
        mbar_wait(&sync.bar[0], par);
        fence_after();

        // ---- epilogue 1 (thread = CW columns of an edge row): Gh, gP[src], gQ[dst], alpha g -> ACC2
        {
            float gcol[CW], ghcol[CW];
            float u[16], ag[16];
            tmem_ld16(T_U, u);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int o = RT::off(t, ch * (CW / 4) + j);
                const float4 hv = lds4(H_hi + o) + lds4(H_lo + o);
                const float4 gv = lds4(G_hi + o) + lds4(G_lo + o);
                const int c = 4 * j;
                gcol[c] = gv.x; gcol[c + 1] = gv.y; gcol[c + 2] = gv.z; gcol[c + 3] = gv.w;
                ag[c] = a.alpha * gv.x; ag[c + 1] = a.alpha * gv.y; ag[c + 2] = a.alpha * gv.z; ag[c + 3] = a.alpha * gv.w;
                const float4 G = f4(hv.x > 0.f ? a.beta * u[c] : 0.f, hv.y > 0.f ? a.beta * u[c + 1] : 0.f,
                                    hv.z > 0.f ? a.beta * u[c + 2] : 0.f, hv.w > 0.f ? a.beta * u[c + 3] : 0.f);
                ghcol[c] = G.x; ghcol[c + 1] = G.y; ghcol[c + 2] = G.z; ghcol[c + 3] = G.w;
                if (valid && a.gP) red_add4(a.gP + (int64_t)srow * HH + ch * CW + c, G);
                float4 gh, gl;
                split4(G, gh, gl);
                st4(H_hi + o, gh);
                st4(H_lo + o, gl);
            }
            tmem_st16(T_ACC2, ag);
            tmem_st_wait();
            // grad Q[dst]: a warp whose 32 rows share one destination reduces first (hubs), else per-row REDs
            const int d0 = __shfl_sync(0xffffffffu, drow, 0);
            const bool uniform = __all_sync(0xffffffffu, drow == d0) && d0 >= 0;
            if (!uniform && valid && a.gQ) {
#pragma unroll
                for (int c = 0; c < CW; c += 4)
                    red_add4(a.gQ + (int64_t)drow * HH + ch * CW + c, f4(ghcol[c], ghcol[c + 1], ghcol[c + 2], ghcol[c + 3]));
            }
            accb2 += warp_column_sums(gcol, lane);
            const float s1 = warp_column_sums(ghcol, lane);
            accb1 += s1;
            if (uniform && a.gQ && lane < CW) atomicAdd(a.gQ + (int64_t)d0 * HH + ch * CW + lane, s1);
        }
        __syncthreads();                                 // every thread is done reading the g tile
        store_rows_split(R, G_hi, G_lo, nrows, tid);     // e tile replaces it
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();

        // ---- phase B: ACC2 += Gh W1; GW1 += Gh^T e
        if (tid == 0) {
            gemm3_data(tm + 64, H_hi, H_lo, W1_hi, W1_lo, 1);
            gemm3_wgrad(T_GW1, H_hi, H_lo, G_hi, G_lo, it > 0);
            mma_commit(&sync.bar[1]);
        }
        if (tile + gridDim.x < ntiles) {                 // next tile's g and h in flight
            nrows_pf = (int)min64(TE, a.E - (tile + gridDim.x) * TE);
            load_rows(R, a.g_out, nullptr, (tile + gridDim.x) * TE, nrows_pf, tid);
            load_rows(R2, a.h, nullptr, (tile + gridDim.x) * TE, nrows_pf, tid);
            drow_pf = __ldg(a.dst + (tile + gridDim.x) * TE + min(t, nrows_pf - 1));
            srow_pf = __ldg(a.src + (tile + gridDim.x) * TE + min(t, nrows_pf - 1));
        }
        prev_row0 = row0;
        prev_nrows = nrows;
        mbar_wait(&sync.bar[1], par);                    // both tile buffers are free again; ACC2 is final
        fence_after();
    }
    if (it > 0) upd_bwd_store_ge(a.g_e, T_ACC2, prev_row0, prev_nrows, t, ch);

    const bool any = it > 0;
    store_wgrad_partial(T_GW1, a.gW1_part + (int64_t)blockIdx.x * HH * HH, 1.f, any, tid);
    store_wgrad_partial(T_GW2, a.gW2_part + (int64_t)blockIdx.x * HH * HH, a.beta, any, tid);
    bsum[warp * 32 + lane] = accb1;
    bsum[NTH + warp * 32 + lane] = accb2;
    fence_before();
    __syncthreads();
    if (tid < HH) {
        a.gb1_part[(int64_t)blockIdx.x * HH + tid] = column_total<NTH>(bsum, tid);
        a.gb2_part[(int64_t)blockIdx.x * HH + tid] = a.beta * column_total<NTH>(bsum + NTH, tid);
    }
    if (warp == 0) tmem_dealloc(tm, 256);
}

int launch_upd_bwd_umma(int64_t E, const float* g_out, const float* h, const float* e, const int32_t* src, const int32_t* dst,
                        const float* W1, const float* W2, float alpha, float beta, float* g_e, float* gP, float* gQ,
                        float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part, int n_part, cudaStream_t st) {
    UpdBwdUArgs a{E, g_out, h, e, src, dst, W1, W2, alpha, beta, g_e, gP, gQ, gW1_part, gb1_part, gW2_part, gb2_part};
    const size_t smem = 4 * RT::BYTES + 4 * WT::BYTES + 2 * NTB * sizeof(float) + 1024;
    B200MP_CUDA(cudaFuncSetAttribute(edge_update_bwd_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    edge_update_bwd_umma_kernel<<<n_part, NTB, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

// ------------------------------------------------------------------ message backward
//   G (per edge, from the saved message m and per-node coefficient rows; edge_bwd.cu has the rules)
//   g_e = G W        ACC (data GEMM, tmem cols  0.. 63)        gT1[nbr] += G   (vector REDs)
//   gW  = G^T e      GW  (wgrad,     tmem cols 64..127)        gb = sum_k G
struct MsgBwdUArgs {
    int64_t E;
    const float* m; const float* e; const int32_t* eid;
    const float* node_in; const int32_t* seg; const int32_t* nbr;
    const float* W; float* g_e; int accumulate; float* gT1;
    float* gW_part; float* gb_part;
};

template <int KIND>
__global__ void __launch_bounds__(NTB, 1) message_bwd_umma_kernel(MsgBwdUArgs a) {
    constexpr int NTH = NTB, CW = CWB;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;   // keeps the shared address space
    float* G_hi = smem;                       // first the raw m tile, then G (hi)
    float* G_lo = G_hi + RT::FLOATS;
    float* E_hi = G_lo + RT::FLOATS;
    float* E_lo = E_hi + RT::FLOATS;
    float* W_hi = E_lo + RT::FLOATS;          // B[n=k][K=j] = W[j][k]   (g_e = G W)
    float* W_lo = W_hi + WT::FLOATS;
    float* bsum = W_lo + WT::FLOATS;          // [NTH / 32 warps][32]
    __shared__ TileSync sync;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int t = tid & (TE - 1), ch = tid >> 7;
    if (warp == 0) tmem_alloc(&sync.tmem_base, 128);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_fence_init(); }
    load_weight_tiles<true, NTH>(W_hi, W_lo, a.W, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t T_GW = tm + 64;
    const uint32_t T_ACC = tm + ((uint32_t)((warp & 3) * 32) << 16) + ch * CW;

    float accb = 0.f;
    const int64_t ntiles = (a.E + TE - 1) / TE;
    uint32_t it = 0;
    RowRegs<NTH> R, R2;
    int nrows_pf = 0, seg_pf = 0, nbr_pf = 0;
    if ((int64_t)blockIdx.x < ntiles) {
        nrows_pf = (int)min64(TE, a.E - (int64_t)blockIdx.x * TE);
        load_rows(R, a.m, nullptr, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        load_rows(R2, a.e, a.eid, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        seg_pf = __ldg(a.seg + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
        if (a.gT1) nbr_pf = __ldg(a.nbr + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
    }

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE;
        const int nrows = nrows_pf;
        const bool valid = t < nrows;

        store_rows_raw(R, G_hi, nrows, tid);                           // m tile (prefetched)
        store_rows_split(R2, E_hi, E_lo, nrows, tid);
        const int64_t d = seg_pf;
        const int64_t nrow = nbr_pf;
        __syncthreads();

        // ---- per-edge gradient G (thread = CW columns of an edge row), in place over the m tile
        {
            float gcol[CW];
#pragma unroll
            for (int j = 0; j < CW / 4; ++j) {
                const int o = RT::off(t, ch * (CW / 4) + j);
                const int c = ch * CW + j * 4;
                const float4 mv = lds4(G_hi + o);
                float4 G;
                if (KIND == 0) {
                    const float* q = a.node_in + d * 7 * HH + c;
                    const float4 c0 = ldg4(q), c1 = ldg4(q + HH), gmn = ldg4(q + 2 * HH), mn = ldg4(q + 3 * HH),
                                 gmx = ldg4(q + 4 * HH), mx = ldg4(q + 5 * HH), mu = ldg4(q + 6 * HH);
                    G.x = fmaf(c1.x, mv.x - mu.x, c0.x) + (mv.x == mn.x ? gmn.x : 0.f) + (mv.x == mx.x ? gmx.x : 0.f);
                    G.y = fmaf(c1.y, mv.y - mu.y, c0.y) + (mv.y == mn.y ? gmn.y : 0.f) + (mv.y == mx.y ? gmx.y : 0.f);
                    G.z = fmaf(c1.z, mv.z - mu.z, c0.z) + (mv.z == mn.z ? gmn.z : 0.f) + (mv.z == mx.z ? gmx.z : 0.f);
                    G.w = fmaf(c1.w, mv.w - mu.w, c0.w) + (mv.w == mn.w ? gmn.w : 0.f) + (mv.w == mx.w ? gmx.w : 0.f);
                } else {
                    const float4 ga = ldg4(a.node_in + d * HH + c);
                    G = f4(mv.x > 0.f ? ga.x : 0.f, mv.y > 0.f ? ga.y : 0.f, mv.z > 0.f ? ga.z : 0.f, mv.w > 0.f ? ga.w : 0.f);
                }
                if (!valid) G = f4(0.f, 0.f, 0.f, 0.f);
                if (valid && a.gT1) red_add4(a.gT1 + nrow * HH + c, G);
                gcol[4 * j] = G.x; gcol[4 * j + 1] = G.y; gcol[4 * j + 2] = G.z; gcol[4 * j + 3] = G.w;
                float4 gh, gl;
                split4(G, gh, gl);
                st4(G_hi + o, gh);
                st4(G_lo + o, gl);
            }
            accb += warp_column_sums(gcol, lane);
        }
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, G_hi, G_lo, W_hi, W_lo, 0);
            gemm3_wgrad(T_GW, G_hi, G_lo, E_hi, E_lo, it > 0);
            mma_commit(&sync.bar[0]);
        }
        int64_t row = row0 + t;
        if (a.eid && valid) row = __ldg(a.eid + row0 + t);
        float* o = a.g_e + row * HH + ch * CW;
        float4 prev[CW / 4];
        if (a.accumulate && valid) {
#pragma unroll
            for (int c = 0; c < CW / 4; ++c) prev[c] = *reinterpret_cast<const float4*>(o + c * 4);
        }
        if (tile + gridDim.x < ntiles) {                 // next tile's m and e in flight
            nrows_pf = (int)min64(TE, a.E - (tile + gridDim.x) * TE);
            load_rows(R, a.m, nullptr, (tile + gridDim.x) * TE, nrows_pf, tid);
            load_rows(R2, a.e, a.eid, (tile + gridDim.x) * TE, nrows_pf, tid);
            seg_pf = __ldg(a.seg + (tile + gridDim.x) * TE + min(t, nrows_pf - 1));
            if (a.gT1) nbr_pf = __ldg(a.nbr + (tile + gridDim.x) * TE + min(t, nrows_pf - 1));
        }
        mbar_wait(&sync.bar[0], it & 1);
        fence_after();

        // ---- epilogue: g_e rows (row(k) = eid ? eid[k] : k)
#pragma unroll
        for (int q = 0; q < CW / 16; ++q) {
            float v[16];
            tmem_ld16(T_ACC + q * 16, v);
            if (valid) {
#pragma unroll
                for (int c = 0; c < 16; c += 4) {
                    float4 r4 = f4(v[c], v[c + 1], v[c + 2], v[c + 3]);
                    if (a.accumulate) r4 = r4 + prev[q * 4 + c / 4];
                    st4(o + q * 16 + c, r4);
                }
            }
        }
        fence_before();
        __syncthreads();
        fence_after();
    }

    store_wgrad_partial(T_GW, a.gW_part + (int64_t)blockIdx.x * HH * HH, 1.f, it > 0, tid);
    bsum[warp * 32 + lane] = accb;
    fence_before();
    __syncthreads();
    if (tid < HH) a.gb_part[(int64_t)blockIdx.x * HH + tid] = column_total<NTH>(bsum, tid);
    if (warp == 0) tmem_dealloc(tm, 128);
}

int launch_msg_bwd_umma(int kind, int64_t E, const float* m, const float* e, const int32_t* eid, const float* node_in,
                        const int32_t* seg, const int32_t* nbr, const float* W, float* g_e, int accumulate, float* gT1,
                        float* gW_part, float* gb_part, int n_part, cudaStream_t st) {
    MsgBwdUArgs a{E, m, e, eid, node_in, seg, nbr, W, g_e, accumulate, gT1, gW_part, gb_part};
    const size_t smem = 4 * RT::BYTES + 2 * WT::BYTES + NTB * sizeof(float) + 1024;
    auto kern = kind == 0 ? message_bwd_umma_kernel<0> : message_bwd_umma_kernel<1>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<n_part, NTB, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200mp

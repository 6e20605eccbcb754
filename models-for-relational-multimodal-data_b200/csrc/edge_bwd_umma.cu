// Tensor-core (tcgen05, 3xTF32) backward kernels for H = 64: edge-update backward and message
// backward.  Same math and outputs as the SIMT kernels in edge_bwd.cu (which remain the path for
// H = 16 / 32 / 128); see b200mp.h for the contract and edge_bwd.cu for the gradient rules.
//
// Per tile of 128 edges the per-edge GEMMs run on the tensor cores with accumulators in tensor
// memory; the two weight gradients accumulate in tensor memory across ALL tiles of the CTA and
// are written once as a per-CTA partial.
#include <cstdlib>
#include "umma_edge.cuh"

namespace b200mp {
using namespace ue;

// ------------------------------------------------------------------ edge-update backward
//   Gh  = (beta g W2) * [h > 0]            U    = g W2            (data GEMM, tmem cols   0.. 63)
//   g_e = alpha g + Gh W1                  ACC2 = alpha g + Gh W1 (data GEMM, tmem cols  64..127)
//   gW2 = beta g^T h                       GW2                    (wgrad,     tmem cols 128..191)
//   gW1 = Gh^T e                           GW1                    (wgrad,     tmem cols 192..255)
template <typename T>                   // T: storage type of the [E,H] tensors (float | bf16)
struct UpdBwdUArgs {
    int64_t E;
    const T* g_out; const T* h; const T* e;
    const int32_t* src; const int32_t* dst;
    const float* W1; const float* W2; float alpha, beta;
    T* g_e; float* gP; float* gQ;
    float* gW1_part; float* gb1_part; float* gW2_part; float* gb2_part;
    const int32_t* g_eid; const int32_t* e_eid;       // storage rows of g_out / e for segment position k (null = k)
    float* gh_out;                                    // optional [E,H] (segment order): Gh rows, for the atomic-free node gradients
};

constexpr int NTB = 512;                  // threads per CTA of the backward kernels (one CTA per SM)
constexpr int CWB = HH / (NTB / TE);      // 16 accumulator columns per tensor-memory thread
constexpr int NPB = RowRegs<NTB>::NP;     // 4 adjacent rows per thread in the coalesced mapping
// edges between drains of the weight-gradient accumulators (B200MP_WGRAD_FLUSH overrides; 0 = never: round-1 behaviour)
static uint32_t wgrad_flush_edges() {
    if (const char* v = getenv("B200MP_WGRAD_FLUSH")) {
        const long n = atol(v);
        return n <= 0 ? 0x40000000u : (uint32_t)n;
    }
    return WGRAD_FLUSH_EDGES;
}

// One RED per run of equal node ids among a thread's adjacent rows (edges are sorted by
// destination, so runs are the rule on that side); when the whole warp (8 adjacent rows) is one
// run the two half warps are combined first.  v[i] must be zero for rows past the tile's end.
__device__ __forceinline__ void red_runs(float* __restrict__ table, const int (&node)[NPB], const float4 (&v)[NPB], int j, int lane) {
    bool one = true;
#pragma unroll
    for (int i = 1; i < NPB; ++i) one = one && node[i] == node[0];
    const int n0 = __shfl_sync(0xffffffffu, node[0], 0);
    if (__all_sync(0xffffffffu, one && node[0] == n0)) {
        float4 sum = v[0];
#pragma unroll
        for (int i = 1; i < NPB; ++i) sum = sum + v[i];
        sum.x += __shfl_xor_sync(0xffffffffu, sum.x, 16);
        sum.y += __shfl_xor_sync(0xffffffffu, sum.y, 16);
        sum.z += __shfl_xor_sync(0xffffffffu, sum.z, 16);
        sum.w += __shfl_xor_sync(0xffffffffu, sum.w, 16);
        if (lane < 16) red_add4(table + (int64_t)n0 * HH + j * 4, sum);
        return;
    }
    float4 run = v[0];
    int cur = node[0];
#pragma unroll
    for (int i = 1; i < NPB; ++i) {
        if (node[i] == cur) {
            run = run + v[i];
        } else {
            red_add4(table + (int64_t)cur * HH + j * 4, run);
            run = v[i];
            cur = node[i];
        }
    }
    red_add4(table + (int64_t)cur * HH + j * 4, run);
}

// Per-thread column partials (thread (g, j): 4 columns of chunk j summed over its rows of all
// tiles) -> per-CTA column sums; red = [NTB / 16][64] floats of shared memory.
__device__ __forceinline__ void cta_column_sums(float* red, float4 acc, float* __restrict__ out, float scale, int tid) {
    __syncthreads();
    st4(red + (tid >> 4) * HH + (tid & 15) * 4, acc);
    __syncthreads();
    if (tid < HH) {
        float sum = 0.f;
#pragma unroll 8
        for (int g = 0; g < NTB / 16; ++g) sum += red[g * HH + tid];
        out[tid] = scale * sum;
    }
}

// (namespace dual -- the two-pipelines-per-CTA helpers -- lives in umma_edge.cuh: the update forward uses it too)

// SPLIT: the data product and the weight-gradient product of a phase are committed to separate mbarriers and the
// accumulator rows are staged in a tile of their own (2 x 16 KB more shared memory), so that the epilogue of a phase only waits
// for the product it reads: the weight-gradient MMAs (half of the tensor-pipe time of this kernel, 42 % active under ncu) run
// behind the epilogue instead of in front of it, and are only awaited where their operand tiles are overwritten.
template <typename T, bool ROWS, bool SPLIT>   // ROWS: write the Gh rows (atomic-free node gradients) instead of scattering them
__global__ void __launch_bounds__(NTB, 1) edge_update_bwd_umma_kernel(UpdBwdUArgs<T> a, int n_virtual, uint32_t FLUSH_U) {
    using namespace dual;
    constexpr int NP = NPG;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;   // keeps the shared address space
    float* W2_hi = smem;                      // B[n=k][K=j] = W2[j][k]   (U = g W2)
    float* W2_lo = W2_hi + WT::FLOATS;
    float* W1_hi = W2_lo + WT::FLOATS;        // B[n=k][K=j] = W1[j][k]   (g_e = Gh W1)
    float* W1_lo = W1_hi + WT::FLOATS;
    __shared__ uint64_t bars[2][4];           // [group][U | ACC2 | GW2 | GW1] (the last two only with SPLIT)
    __shared__ uint32_t tmem_base_s;
    __shared__ int sidx_s[2][2][RG];          // [group][src / dst][row]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int grp = tid >> 8, lt = tid & (NTG - 1), lwarp = lt >> 5;
    const int j = lt & 15, g = lt >> 4;                    // global-memory side
    float* G_hi = W1_lo + WT::FLOATS + grp * 4 * RG_T::FLOATS;   // g tile; staging tile once its MMAs are done; then the e tile
    float* G_lo = G_hi + RG_T::FLOATS;
    float* H_hi = G_lo + RG_T::FLOATS;        // h tile, later the Gh tile
    float* H_lo = H_hi + RG_T::FLOATS;
    uint64_t* bar = bars[grp];
    int (*sidx)[RG] = sidx_s[grp];
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    if (lt == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_init(&bar[2], 1); mbar_init(&bar[3], 1); mbar_fence_init(); }
    load_weight_tiles<true, NTB>(W2_hi, W2_lo, a.W2, tid);
    load_weight_tiles<true, NTB>(W1_hi, W1_lo, a.W1, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = tmem_base_s + grp * 256;
    const uint32_t T_U = tm, T_ACC2 = tm + 64, T_GW2 = tm + 128, T_GW1 = tm + 192;
    float* stage = SPLIT ? W1_lo + WT::FLOATS + (8 + grp) * RG_T::FLOATS : G_hi;

    float4 accb1 = f4(0.f, 0.f, 0.f, 0.f), accb2 = f4(0.f, 0.f, 0.f, 0.f);   // columns 4 j .. 4 j + 3 over this thread's rows
    const int64_t ntiles = (a.E + RG - 1) / RG;
    const int64_t v0 = (int64_t)blockIdx.x * 2 + grp;      // this group's virtual CTA
    const int64_t stride = n_virtual;
    const bool active = v0 < n_virtual;
    uint32_t it = 0;
    Rows Rg, Rh, Re;
    int nrows_pf = 0, idx_pf = 0;
    const int t = lt & (RG - 1), which = lt >> 6;          // index staging: thread lt < 128 -> (src | dst) of row t
    const int32_t* idx_src = which == 0 ? a.src : which == 1 ? a.dst : nullptr;
    if (active && v0 < ntiles) {
        nrows_pf = (int)min64(RG, a.E - v0 * RG);
        load_rows(Rg, a.g_out, a.g_eid, v0 * RG, nrows_pf, lt);
        load_rows(Rh, a.h, nullptr, v0 * RG, nrows_pf, lt);
        if (idx_src) idx_pf = __ldg(idx_src + v0 * RG + min(t, nrows_pf - 1));
    }

    for (int64_t tile = v0; active && tile < ntiles; tile += stride, ++it) {
        const int64_t row0 = tile * RG;
        const int nrows = nrows_pf;
        const uint32_t par = it & 1;

        // ---- phase A: g and h tiles (prefetched, kept in registers too) -> shared memory; U = g W2; GW2 += g^T h
        store_rows_split(Rg, G_hi, G_lo, nrows, lt);
        store_rows_split(Rh, H_hi, H_lo, nrows, lt);
        if (which < 2) sidx[which][t] = idx_pf;
        fence_async_smem();
        fence_before();
        group_sync(grp);
        fence_after();
        if (lt == 0) {
            gemm3_data64(T_U, G_hi, G_lo, W2_hi, W2_lo, 0);
            if (SPLIT) mma_commit(&bar[0]);
            gemm3_wgrad64(T_GW2, G_hi, G_lo, H_hi, H_lo, (it % FLUSH_U) != 0);
            mma_commit(&bar[SPLIT ? 2 : 0]);
        }
        load_rows(Re, a.e, a.e_eid, row0, nrows, lt);                 // e tile in flight during the MMAs
        int srow[NP], drow[NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) { srow[i] = sidx[0][g * NP + i]; drow[i] = sidx[1][g * NP + i]; }
        mbar_wait(&bar[0], par);
        fence_after();
        stage_from_tmem64(T_U, stage, lwarp, lane);                   // the g tile is dead: U rows take its place
        group_sync(grp);

        // ---- Gh = beta U [h > 0] in the coalesced mapping; gP[src], gQ[dst], bias partials; Gh -> MMA operand
        {
            float4 Gh[NP];
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int r = g * NP + i;
                const float4 u = lds4(stage + stage_off(r, j));
                const float4 hv = Rh.v[i];
                const bool valid = r < nrows;
                Gh[i] = f4(valid && hv.x > 0.f ? a.beta * u.x : 0.f, valid && hv.y > 0.f ? a.beta * u.y : 0.f,
                           valid && hv.z > 0.f ? a.beta * u.z : 0.f, valid && hv.w > 0.f ? a.beta * u.w : 0.f);
                accb1 = accb1 + Gh[i];
                if (valid) accb2 = accb2 + Rg.v[i];
                if (!ROWS && valid && a.gP) red_add4(a.gP + (int64_t)srow[i] * HH + j * 4, Gh[i]);
                if (ROWS && valid) st4(a.gh_out + (row0 + r) * HH + j * 4, Gh[i]);
                if (!SPLIT) {
                    float4 gh, gl;
                    split4(Gh[i], gh, gl);
                    const int o = RG_T::off(r, j);
                    st4(H_hi + o, gh);
                    st4(H_lo + o, gl);
                }
            }
            if (!ROWS && a.gQ) red_runs(a.gQ, drow, Gh, j, lane);
            if (SPLIT) {
                mbar_wait(&bar[2], par);                 // GW2 has read the g and h tiles: both may be overwritten
                fence_after();
#pragma unroll
                for (int i = 0; i < NP; ++i) {
                    float4 gh, gl;
                    split4(Gh[i], gh, gl);
                    const int o = RG_T::off(g * NP + i, j);
                    st4(H_hi + o, gh);
                    st4(H_lo + o, gl);
                }
            }
        }
        if (!SPLIT) group_sync(grp);                     // every thread is done reading the staged U rows
        store_rows_split(Re, G_hi, G_lo, nrows, lt);     // e tile replaces them (SPLIT: replaces the g tile; the stage is separate)
        fence_async_smem();
        fence_before();
        group_sync(grp);
        fence_after();

        // ---- phase B: ACC2 = Gh W1; GW1 += Gh^T e
        if (lt == 0) {
            gemm3_data64(T_ACC2, H_hi, H_lo, W1_hi, W1_lo, 0);
            if (SPLIT) mma_commit(&bar[1]);
            gemm3_wgrad64(T_GW1, H_hi, H_lo, G_hi, G_lo, (it % FLUSH_U) != 0);
            mma_commit(&bar[SPLIT ? 3 : 1]);
        }
        int nrows_next = 0;
        if (tile + stride < ntiles) {                    // next tile's h and g in flight
            nrows_next = (int)min64(RG, a.E - (tile + stride) * RG);
            load_rows(Rh, a.h, nullptr, (tile + stride) * RG, nrows_next, lt);
            load_rows(Re, a.g_out, a.g_eid, (tile + stride) * RG, nrows_next, lt);
            if (idx_src) idx_pf = __ldg(idx_src + (tile + stride) * RG + min(t, nrows_next - 1));
        }
        mbar_wait(&bar[1], par);
        fence_after();
        stage_from_tmem64(T_ACC2, stage, lwarp, lane);   // the e tile is dead too
        group_sync(grp);

        // ---- g_e = alpha g + Gh W1, coalesced rows
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            const float4 v = a.alpha * Rg.v[i] + lds4(stage + stage_off(r, j));
            if (r < nrows) st_e4(a.g_e + (row0 + r) * HH + j * 4, v);
        }
#pragma unroll
        for (int i = 0; i < NP; ++i) Rg.v[i] = Re.v[i];
        nrows_pf = nrows_next;
        if (SPLIT) { mbar_wait(&bar[3], par); fence_after(); }    // GW1 done: its operand tiles are free for the next tile, and the drain may read it
        if ((it % FLUSH_U) == FLUSH_U - 1) {             // drain the two weight-gradient accumulators (both products are complete)
            store_wgrad_partial(T_GW1, a.gW1_part + v0 * HH * HH, 1.f, true, lt, it >= FLUSH_U);
            store_wgrad_partial(T_GW2, a.gW2_part + v0 * HH * HH, a.beta, true, lt, it >= FLUSH_U);
        }
        fence_before();
        group_sync(grp);
        fence_after();
    }

    if (active) {
        const bool any = (it % FLUSH_U) != 0, flushed = it >= FLUSH_U;        // tiles since the last drain | drained before
        if (any || !flushed) {
            store_wgrad_partial(T_GW1, a.gW1_part + v0 * HH * HH, 1.f, any, lt, flushed);
            store_wgrad_partial(T_GW2, a.gW2_part + v0 * HH * HH, a.beta, any, lt, flushed);
        }
        group_column_sums(G_hi, accb1, a.gb1_part + v0 * HH, 1.f, lt, grp);
        group_column_sums(G_hi, accb2, a.gb2_part + v0 * HH, a.beta, lt, grp);
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base_s, 512);
}

template <typename T>
static int launch_upd_bwd_umma_t(int64_t E, const T* g_out, const T* h, const T* e, const int32_t* src, const int32_t* dst,
                                 const float* W1, const float* W2, float alpha, float beta, T* g_e, float* gP, float* gQ,
                                 float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part, int n_part,
                                 const int32_t* g_eid, const int32_t* e_eid, cudaStream_t st, float* gh_out) {
    UpdBwdUArgs<T> a{E, g_out, h, e, src, dst, W1, W2, alpha, beta, g_e, gP, gQ, gW1_part, gb1_part, gW2_part, gb2_part, g_eid, e_eid, gh_out};
    // B200MP_UPD_BWD_SPLIT=1 selects the split-barrier phases.  OPT-IN: same-box A/B on hi_medium 16.2 ms against 13.8 ms -- the
    // tensor-memory reads of the epilogue queue behind the weight-gradient MMAs they were meant to overlap, and every thread
    // polls four mbarriers per tile instead of two
    const char* sv = getenv("B200MP_UPD_BWD_SPLIT");
    const bool split = sv && sv[0] == '1';
    const size_t smem = (split ? 10 : 8) * dual::RG_T::BYTES + 4 * WT::BYTES + 1024;
    auto kern = split ? (gh_out ? edge_update_bwd_umma_kernel<T, true, true> : edge_update_bwd_umma_kernel<T, false, true>)
                      : (gh_out ? edge_update_bwd_umma_kernel<T, true, false> : edge_update_bwd_umma_kernel<T, false, false>);
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const uint32_t fl = wgrad_flush_edges() / 64;          // tiles (64 edges) between drains
    kern<<<(n_part + 1) / 2, NTB, smem, st>>>(a, n_part, fl < 1 ? 1 : fl);   // two 64-edge pipelines (partial rows) per CTA
    B200MP_LAUNCH_CHECK();
    return 0;
}

int launch_upd_bwd_umma(int64_t E, const float* g_out, const float* h, const float* e, const int32_t* src, const int32_t* dst,
                        const float* W1, const float* W2, float alpha, float beta, float* g_e, float* gP, float* gQ,
                        float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part, int n_part,
                        const int32_t* g_eid, const int32_t* e_eid, cudaStream_t st, float* gh_out) {
    return launch_upd_bwd_umma_t<float>(E, g_out, h, e, src, dst, W1, W2, alpha, beta, g_e, gP, gQ, gW1_part, gb1_part, gW2_part, gb2_part,
                                        n_part, g_eid, e_eid, st, gh_out);
}

// ------------------------------------------------------------------ message backward
//   G (per edge, from the saved message m and per-node coefficient rows; edge_bwd.cu has the rules)
//   g_e = G W        ACC (data GEMM, tmem cols  0.. 63)        gT1[nbr] += G   (vector REDs)
//   gW  = G^T e      GW  (wgrad,     tmem cols 64..127)        gb = sum_k G
template <typename T>
struct MsgBwdUArgs {
    int64_t E;
    const T* m; const T* e; const int32_t* eid;
    const float* node_in; const int32_t* seg; const int32_t* nbr;
    const float* W; T* g_e; int accumulate; float* gT1;
    float* gW_part; float* gb_part;
    const T* g_acc;                                   // optional [E,H] in segment order added to g_e
    float* G_out;                                     // optional [E,H] (segment order): per-edge gradient rows G
    int l2pf;                                         // 64-edge kernel: announce the next tile's node_in rows to L2 one tile ahead
};

// PNA per-edge gradient from the saved message and the destination's coefficient rows
// (c0, c1, g_min, min, g_max, max, mean; node_ops.cu / edge_bwd.cu have the rules)
struct PnaCoef { float4 c0, c1, gmn, mn, gmx, mx, mu; };
__device__ __forceinline__ void load_coef(PnaCoef& c, const float* __restrict__ q) {
    c.c0 = ldg4(q); c.c1 = ldg4(q + HH); c.gmn = ldg4(q + 2 * HH); c.mn = ldg4(q + 3 * HH);
    c.gmx = ldg4(q + 4 * HH); c.mx = ldg4(q + 5 * HH); c.mu = ldg4(q + 6 * HH);
}
__device__ __forceinline__ float pna_g(float m, float c0, float c1, float gmn, float mn, float gmx, float mx, float mu) {
    return fmaf(c1, m - mu, c0) + (m == mn ? gmn : 0.f) + (m == mx ? gmx : 0.f);
}

template <int KIND, typename T, bool ROWS>
__global__ void __launch_bounds__(NTB, 1) message_bwd_umma_kernel(MsgBwdUArgs<T> a, uint32_t FLUSH_M) {
    constexpr int NTH = NTB, CW = CWB, NP = NPB;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;   // keeps the shared address space
    float* G_hi = smem;                       // per-edge gradient G
    float* G_lo = G_hi + RT::FLOATS;
    float* E_hi = G_lo + RT::FLOATS;          // e tile; staging tile once the MMAs are done
    float* E_lo = E_hi + RT::FLOATS;
    float* W_hi = E_lo + RT::FLOATS;          // B[n=k][K=j] = W[j][k]   (g_e = G W)
    float* W_lo = W_hi + WT::FLOATS;
    __shared__ TileSync sync;

    const int tid = threadIdx.x, warp = tid >> 5;
    const int t = tid & (TE - 1), ch = tid >> 7;           // tensor-memory side
    const int j = tid & 15, g = tid >> 4;                  // global-memory side
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
    float* stage = E_hi;

    float4 accb = f4(0.f, 0.f, 0.f, 0.f);     // columns 4 j .. 4 j + 3 over this thread's rows
    const int64_t ntiles = (a.E + TE - 1) / TE;
    uint32_t it = 0;
    RowRegs<NTH> Rm, Re;
    int nrows_pf = 0, seg_pf[NP], nbr_pf[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) { seg_pf[i] = 0; nbr_pf[i] = 0; }
    auto prefetch = [&](int64_t tile_) {
        const int64_t r0_ = tile_ * TE;
        nrows_pf = (int)min64(TE, a.E - r0_);
        load_rows(Rm, a.m, nullptr, r0_, nrows_pf, tid);
        load_rows(Re, a.e, a.eid, r0_, nrows_pf, tid);
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int rc = min(g * NP + i, nrows_pf - 1);
            seg_pf[i] = __ldg(a.seg + r0_ + rc);
            if (a.gT1) nbr_pf[i] = __ldg(a.nbr + r0_ + rc);
        }
    };
    if ((int64_t)blockIdx.x < ntiles) prefetch(blockIdx.x);

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE;
        const int nrows = nrows_pf;

        store_rows_split(Re, E_hi, E_lo, nrows, tid);
        // ---- per-edge gradient G in the coalesced mapping (m rows and ids are in registers)
        {
            float4 Gv[NP];
            PnaCoef cf;
            float4 ga = f4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int r = g * NP + i;
                const float4 mv = Rm.v[i];
                const int64_t d = seg_pf[i];
                const bool fresh = i == 0 || seg_pf[i] != seg_pf[i - 1];      // adjacent rows mostly share their segment
                float4 G;
                if (KIND == 0) {
                    if (fresh) load_coef(cf, a.node_in + d * 7 * HH + j * 4);
                    G.x = pna_g(mv.x, cf.c0.x, cf.c1.x, cf.gmn.x, cf.mn.x, cf.gmx.x, cf.mx.x, cf.mu.x);
                    G.y = pna_g(mv.y, cf.c0.y, cf.c1.y, cf.gmn.y, cf.mn.y, cf.gmx.y, cf.mx.y, cf.mu.y);
                    G.z = pna_g(mv.z, cf.c0.z, cf.c1.z, cf.gmn.z, cf.mn.z, cf.gmx.z, cf.mx.z, cf.mu.z);
                    G.w = pna_g(mv.w, cf.c0.w, cf.c1.w, cf.gmn.w, cf.mn.w, cf.gmx.w, cf.mx.w, cf.mu.w);
                } else {
                    if (fresh) ga = ldg4(a.node_in + d * HH + j * 4);
                    G = f4(mv.x > 0.f ? ga.x : 0.f, mv.y > 0.f ? ga.y : 0.f, mv.z > 0.f ? ga.z : 0.f, mv.w > 0.f ? ga.w : 0.f);
                }
                if (r >= nrows) G = f4(0.f, 0.f, 0.f, 0.f);
                if (!ROWS && r < nrows && a.gT1) red_add4(a.gT1 + (int64_t)nbr_pf[i] * HH + j * 4, G);
                if (ROWS && r < nrows) st4(a.G_out + (row0 + r) * HH + j * 4, G);
                accb = accb + G;
                Gv[i] = G;
            }
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                float4 gh, gl;
                split4(Gv[i], gh, gl);
                const int o = RT::off(g * NP + i, j);
                st4(G_hi + o, gh);
                st4(G_lo + o, gl);
            }
        }
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, G_hi, G_lo, W_hi, W_lo, 0);
            gemm3_wgrad(T_GW, G_hi, G_lo, E_hi, E_lo, (it % FLUSH_M) != 0);
            mma_commit(&sync.bar[0]);
        }
        int64_t orow[NP];
        float4 prev[NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            orow[i] = (a.eid && r < nrows) ? (int64_t)__ldg(a.eid + row0 + r) : row0 + r;
        }
#pragma unroll
        for (int i = 0; i < NP; ++i) prev[i] = f4(0.f, 0.f, 0.f, 0.f);
        if (a.accumulate) {
#pragma unroll
            for (int i = 0; i < NP; ++i)
                if (g * NP + i < nrows) prev[i] = ld_e4(a.g_e + orow[i] * HH + j * 4);
        }
        if (a.g_acc) {
#pragma unroll
            for (int i = 0; i < NP; ++i)
                if (g * NP + i < nrows) prev[i] = prev[i] + ldg_e4(a.g_acc + (row0 + g * NP + i) * HH + j * 4);
        }
        if (tile + gridDim.x < ntiles) prefetch(tile + gridDim.x);    // next tile's m, e and ids in flight
        mbar_wait(&sync.bar[0], it & 1);
        fence_after();
        stage_from_tmem<NTH>(T_ACC, stage, t, ch);        // the e tile is dead: g_e rows take its place
        __syncthreads();

        // ---- g_e rows, coalesced (row(k) = eid ? eid[k] : k)
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            float4 v = lds4(stage + stage_off(r, j));
            v = v + prev[i];
            if (r < nrows) st_e4(a.g_e + orow[i] * HH + j * 4, v);
        }
        if ((it % FLUSH_M) == FLUSH_M - 1)                // drain the weight-gradient accumulator
            store_wgrad_partial(T_GW, a.gW_part + (int64_t)blockIdx.x * HH * HH, 1.f, true, tid, it >= FLUSH_M);
        fence_before();
        __syncthreads();
        fence_after();
    }

    if ((it % FLUSH_M) != 0 || it < FLUSH_M)
        store_wgrad_partial(T_GW, a.gW_part + (int64_t)blockIdx.x * HH * HH, 1.f, (it % FLUSH_M) != 0, tid, it >= FLUSH_M);
    cta_column_sums(G_hi, accb, a.gb_part + (int64_t)blockIdx.x * HH, 1.f, tid);
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 128);
}

// ---- message backward on 64-edge tiles, two CTAs per SM ---------------------------------------------------------------
// Round 2 measured that two resident CTAs are what hides the per-tile phase chain (message forward: -16 %); the 128-edge
// kernel above needs 160 KB of shared memory, so it runs alone on its SM.  This variant is the same computation on 64-edge
// tiles with 256 threads: 96 KB (G and e hi / lo tiles of 16 KB, the hi / lo weight tile), two CTAs per SM, each with its own
// tensor-memory columns and weight-gradient partial.  The data product runs at M = 64 (half the tensor-pipe efficiency of
// M = 128, as in the two-pipeline edge-update backward); the weight-gradient product is M = 64 in both.
template <int KIND, typename T, bool ROWS>
__global__ void __launch_bounds__(256, 2) message_bwd_umma64_kernel(MsgBwdUArgs<T> a, uint32_t FLUSH_T) {
    using namespace dual;
    constexpr int NP = NPG;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;
    float* G_hi = smem;
    float* G_lo = G_hi + RG_T::FLOATS;
    float* E_hi = G_lo + RG_T::FLOATS;        // e tile; staging tile once the MMAs are done
    float* E_lo = E_hi + RG_T::FLOATS;
    float* W_hi = E_lo + RG_T::FLOATS;
    float* W_lo = W_hi + WT::FLOATS;
    __shared__ TileSync sync;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int j = tid & 15, g = tid >> 4;                  // global-memory side: chunk j of rows 4 g .. 4 g + 3
    if (warp == 0) tmem_alloc(&sync.tmem_base, 128);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_fence_init(); }
    load_weight_tiles<true, NTG>(W_hi, W_lo, a.W, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t T_ACC = tm, T_GW = tm + 64;
    float* stage = E_hi;

    float4 accb = f4(0.f, 0.f, 0.f, 0.f);
    const int64_t ntiles = (a.E + RG - 1) / RG;
    uint32_t it = 0;
    Rows Rm, Re;
    int nrows_pf = 0, seg_pf[NP], nbr_pf[NP];
#pragma unroll
    for (int i = 0; i < NP; ++i) { seg_pf[i] = 0; nbr_pf[i] = 0; }
    auto prefetch = [&](int64_t tile_) {
        const int64_t r0_ = tile_ * RG;
        nrows_pf = (int)min64(RG, a.E - r0_);
        load_rows(Rm, a.m, nullptr, r0_, nrows_pf, tid);
        load_rows(Re, a.e, a.eid, r0_, nrows_pf, tid);
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int rc = min(g * NP + i, nrows_pf - 1);
            seg_pf[i] = __ldg(a.seg + r0_ + rc);
            if (a.gT1) nbr_pf[i] = __ldg(a.nbr + r0_ + rc);
        }
    };
    if ((int64_t)blockIdx.x < ntiles) prefetch(blockIdx.x);
    // node_in rows (PNA: 7 x 256 B of coefficients per destination, GINE: 256 B) are read by the one or two tiles that hold the node's
    // edges: first touches in DRAM, and the per-edge gradient cannot start without them (ncu: 17 % of this kernel's samples wait at
    // their first uses, the block barrier behind them another 24 %).  Threads 0..63 carry the destination of row tid of the NEXT tile
    // (and of row tid - 1) and announce each distinct node's rows to L2 at the top of the iteration before.
    int pf_id = -1, pf_prev = -1;
    auto load_pf_ids = [&](int64_t tile_) {
        pf_id = -1;
        if (a.l2pf && tid < RG && tile_ < ntiles) {
            const int64_t r0_ = tile_ * RG;
            if (r0_ + tid < a.E) {
                pf_id = __ldg(a.seg + r0_ + tid);
                pf_prev = tid > 0 ? __ldg(a.seg + r0_ + tid - 1) : -1;
            }
        }
    };
    load_pf_ids((int64_t)blockIdx.x + gridDim.x);

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * RG;
        const int nrows = nrows_pf;

        if (pf_id >= 0 && pf_id != pf_prev) {
            const float* q = a.node_in + (int64_t)pf_id * (KIND == 0 ? 7 * HH : HH);
#pragma unroll
            for (int k = 0; k < (KIND == 0 ? 14 : 2); ++k) prefetch_l2(q + k * 32);
        }
        load_pf_ids(tile + 2 * (int64_t)gridDim.x);
        store_rows_split(Re, E_hi, E_lo, nrows, tid);
        {
            float4 Gv[NP];
            PnaCoef cf;
            float4 ga = f4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int r = g * NP + i;
                const float4 mv = Rm.v[i];
                const int64_t d = seg_pf[i];
                const bool fresh = i == 0 || seg_pf[i] != seg_pf[i - 1];
                float4 G;
                if (KIND == 0) {
                    if (fresh) load_coef(cf, a.node_in + d * 7 * HH + j * 4);
                    G.x = pna_g(mv.x, cf.c0.x, cf.c1.x, cf.gmn.x, cf.mn.x, cf.gmx.x, cf.mx.x, cf.mu.x);
                    G.y = pna_g(mv.y, cf.c0.y, cf.c1.y, cf.gmn.y, cf.mn.y, cf.gmx.y, cf.mx.y, cf.mu.y);
                    G.z = pna_g(mv.z, cf.c0.z, cf.c1.z, cf.gmn.z, cf.mn.z, cf.gmx.z, cf.mx.z, cf.mu.z);
                    G.w = pna_g(mv.w, cf.c0.w, cf.c1.w, cf.gmn.w, cf.mn.w, cf.gmx.w, cf.mx.w, cf.mu.w);
                } else {
                    if (fresh) ga = ldg4(a.node_in + d * HH + j * 4);
                    G = f4(mv.x > 0.f ? ga.x : 0.f, mv.y > 0.f ? ga.y : 0.f, mv.z > 0.f ? ga.z : 0.f, mv.w > 0.f ? ga.w : 0.f);
                }
                if (r >= nrows) G = f4(0.f, 0.f, 0.f, 0.f);
                if (!ROWS && r < nrows && a.gT1) red_add4(a.gT1 + (int64_t)nbr_pf[i] * HH + j * 4, G);
                if (ROWS && r < nrows) st4(a.G_out + (row0 + r) * HH + j * 4, G);
                accb = accb + G;
                Gv[i] = G;
            }
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                float4 gh, gl;
                split4(Gv[i], gh, gl);
                const int o = RG_T::off(g * NP + i, j);
                st4(G_hi + o, gh);
                st4(G_lo + o, gl);
            }
        }
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data64(T_ACC, G_hi, G_lo, W_hi, W_lo, 0);
            gemm3_wgrad64(T_GW, G_hi, G_lo, E_hi, E_lo, (it % FLUSH_T) != 0);
            mma_commit(&sync.bar[0]);
        }
        int64_t orow[NP];
        float4 prev[NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            orow[i] = (a.eid && r < nrows) ? (int64_t)__ldg(a.eid + row0 + r) : row0 + r;
        }
#pragma unroll
        for (int i = 0; i < NP; ++i) prev[i] = f4(0.f, 0.f, 0.f, 0.f);
        if (a.accumulate) {
#pragma unroll
            for (int i = 0; i < NP; ++i)
                if (g * NP + i < nrows) prev[i] = ld_e4(a.g_e + orow[i] * HH + j * 4);
        }
        if (a.g_acc) {
#pragma unroll
            for (int i = 0; i < NP; ++i)
                if (g * NP + i < nrows) prev[i] = prev[i] + ldg_e4(a.g_acc + (row0 + g * NP + i) * HH + j * 4);
        }
        if (tile + gridDim.x < ntiles) prefetch(tile + gridDim.x);
        mbar_wait(&sync.bar[0], it & 1);
        fence_after();
        stage_from_tmem64(T_ACC, stage, warp, lane);      // the e tile is dead: g_e rows take its place
        __syncthreads();
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            float4 v = lds4(stage + stage_off(r, j));
            v = v + prev[i];
            if (r < nrows) st_e4(a.g_e + orow[i] * HH + j * 4, v);
        }
        if ((it % FLUSH_T) == FLUSH_T - 1)
            store_wgrad_partial(T_GW, a.gW_part + (int64_t)blockIdx.x * HH * HH, 1.f, true, tid, it >= FLUSH_T);
        fence_before();
        __syncthreads();
        fence_after();
    }

    if ((it % FLUSH_T) != 0 || it < FLUSH_T)
        store_wgrad_partial(T_GW, a.gW_part + (int64_t)blockIdx.x * HH * HH, 1.f, (it % FLUSH_T) != 0, tid, it >= FLUSH_T);
    group_column_sums(G_hi, accb, a.gb_part + (int64_t)blockIdx.x * HH, 1.f, tid, 0);
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 128);
}

template <typename T>
static int launch_msg_bwd_umma_t(int kind, int64_t E, const T* m, const T* e, const int32_t* eid, const float* node_in,
                                 const int32_t* seg, const int32_t* nbr, const float* W, T* g_e, int accumulate, float* gT1,
                                 float* gW_part, float* gb_part, int n_part, const T* g_acc, cudaStream_t st, float* G_out) {
    const char* lv = getenv("B200MP_L2PF");                 // A/B switch of the L2 announcements (default on)
    MsgBwdUArgs<T> a{E, m, e, eid, node_in, seg, nbr, W, g_e, accumulate, gT1, gW_part, gb_part, g_acc, G_out, (lv && lv[0] == '0') ? 0 : 1};
    {   // the 64-edge / two-CTAs-per-SM variant is the default (same-box A/B on hi_medium: PNA 11.77 -> 11.36 ms, GINE 10.85 -> 10.61 ms);
        // B200MP_MSG_BWD64=0 selects the 128-edge kernel
        const char* v = getenv("B200MP_MSG_BWD64");
        if (!(v && v[0] == '0')) {
            const size_t smem64 = 4 * dual::RG_T::BYTES + 2 * WT::BYTES + 1024;
            auto k64 = G_out ? (kind == 0 ? message_bwd_umma64_kernel<0, T, true> : message_bwd_umma64_kernel<1, T, true>)
                             : (kind == 0 ? message_bwd_umma64_kernel<0, T, false> : message_bwd_umma64_kernel<1, T, false>);
            B200MP_CUDA(cudaFuncSetAttribute(k64, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem64));
            (void)cudaFuncSetAttribute(k64, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
            const uint32_t fl64 = wgrad_flush_edges() / 64;
            k64<<<n_part, 256, smem64, st>>>(a, fl64 < 1 ? 1 : fl64);
            B200MP_LAUNCH_CHECK();
            return 0;
        }
    }
    const size_t smem = 4 * RT::BYTES + 2 * WT::BYTES + 1024;
    auto kern = G_out ? (kind == 0 ? message_bwd_umma_kernel<0, T, true> : message_bwd_umma_kernel<1, T, true>)
                      : (kind == 0 ? message_bwd_umma_kernel<0, T, false> : message_bwd_umma_kernel<1, T, false>);
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const uint32_t fl = wgrad_flush_edges() / TE;
    kern<<<n_part, NTB, smem, st>>>(a, fl < 1 ? 1 : fl);
    B200MP_LAUNCH_CHECK();
    return 0;
}

int launch_msg_bwd_umma(int kind, int64_t E, const float* m, const float* e, const int32_t* eid, const float* node_in,
                        const int32_t* seg, const int32_t* nbr, const float* W, float* g_e, int accumulate, float* gT1,
                        float* gW_part, float* gb_part, int n_part, const float* g_acc, cudaStream_t st, float* G_out) {
    return launch_msg_bwd_umma_t<float>(kind, E, m, e, eid, node_in, seg, nbr, W, g_e, accumulate, gT1, gW_part, gb_part, n_part, g_acc, st, G_out);
}

}  // namespace b200mp

/* bf16 STORAGE of the [E,64] tensors (m, e, g_e, g_acc / g_out, h), fp32 node tables / weights / partials; tensor-core path only */
extern "C" int b200mp_message_bwd_bf16(int kind, int64_t E, const void* m, const void* e, const int32_t* eid, const float* node_in,
                                       const int32_t* seg, const int32_t* nbr, const float* W, void* g_e, int accumulate, float* gT1,
                                       float* gW_part, float* gb_part, int n_part, const void* g_acc, void* stream) {
    using namespace b200mp;
    if (E < 0 || n_part <= 0 || !W || !gW_part || !gb_part || (kind != 0 && kind != 1)) return B200MP_EINVAL;
    if (E > 0 && (!g_e || !m || !e || !node_in || !seg || (gT1 && !nbr))) return B200MP_EINVAL;
    return launch_msg_bwd_umma_t<ue::bf16>(kind, E, (const ue::bf16*)m, (const ue::bf16*)e, eid, node_in, seg, nbr, W, (ue::bf16*)g_e, accumulate,
                                           gT1, gW_part, gb_part, n_part, (const ue::bf16*)g_acc, (cudaStream_t)stream, nullptr);
}

extern "C" int b200mp_edge_update_bwd_bf16(int64_t E, const void* g_out, const void* h, const void* e, const int32_t* src,
                                           const int32_t* dst, const float* W1, const float* W2, float alpha, float beta, void* g_e,
                                           float* gP, float* gQ, float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part,
                                           int n_part, const int32_t* g_eid, const int32_t* e_eid, void* stream) {
    using namespace b200mp;
    if (E < 0 || n_part <= 0 || !W1 || !W2 || !gW1_part || !gb1_part || !gW2_part || !gb2_part) return B200MP_EINVAL;
    if (E > 0 && (!g_e || !g_out || !h || !e || !dst || (gP && !src))) return B200MP_EINVAL;
    return launch_upd_bwd_umma_t<ue::bf16>(E, (const ue::bf16*)g_out, (const ue::bf16*)h, (const ue::bf16*)e, src, dst, W1, W2, alpha, beta,
                                           (ue::bf16*)g_e, gP, gQ, gW1_part, gb1_part, gW2_part, gb2_part, n_part, g_eid, e_eid,
                                           (cudaStream_t)stream, nullptr);
}

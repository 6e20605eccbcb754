// Tensor-core (tcgen05, 3xTF32) per-edge linear forward for H = 64 (b200mp_edge_linear_fwd):
// PNA / GINE message (modes 0 / 1) and the edge update (mode 2).  Same contract as the SIMT
// kernels in edge_fwd.cu, which remain the path for H = 16 / 32 / 128.
//
//   z = e W1^T + b1 + T1[idx1] + T2[idx2]     Z (data GEMM, tmem cols 0..63); the gathered node rows
//                                             are fetched coalesced while the MMAs run
//   mode 0: out1 = z     mode 1: out1 = relu(z)
//   mode 2: out1 = h = relu(z);  out2[row] = alpha e + beta (h W2^T + b2)     U (tmem cols 64..127)
#include <cstdlib>
#include "umma_edge.cuh"

namespace b200mp {
using namespace ue;

template <typename T>                   // T: storage type of the [E,H] tensors (float | bf16)
struct EdgeFwdUArgs {
    int64_t E;
    const T* e; const int32_t* eid;
    const float* W1; const float* b1;
    const float* T1; const int32_t* idx1;
    const float* T2; const int32_t* idx2;
    const float* W2; const float* b2;
    float alpha, beta;
    T* out1; T* out2; const int32_t* out_eid;
    int l2pf;                                 // modes 0 / 1: announce the next tile's T1 rows (random source rows) to L2 at the end of a tile
};

// threads per CTA: the message modes fit two CTAs of 256 per SM; the update's tiles fill shared
// memory (one CTA per SM), so it runs 512 threads
template <int MODE> struct FwdCfg { static constexpr int NTH = MODE == 2 ? 512 : 256; };

// Two thread mappings per tile (umma_edge.cuh): the tensor-memory side is row-per-thread
// (thread (t, ch) reads CW accumulator columns of edge row t and drops them into a staging tile);
// everything that touches global memory -- e rows, gathered node rows, h / e' / m rows -- runs in
// the coalesced mapping (thread (g, j): 16-byte chunk j of NP adjacent rows), with the gathered
// rows and the residual e rows kept in registers instead of shared memory.
// GPF (mode 2 only): the node rows P[src] / Q[dst] of the NEXT tile are gathered behind this tile's second product -- the
// registers that held this tile's gathered rows are free from the first epilogue on -- so that they are in flight across the
// tile boundary instead of between the first MMA and its epilogue, where a random-row DRAM latency (~1 us) outlasts the 0.4 us
// of tensor work it was meant to hide behind.  The next tile's indices go into the other half of a double-buffered index stage.
template <int MODE, typename T, bool GPF = false>
__global__ void __launch_bounds__(FwdCfg<MODE>::NTH, MODE == 2 ? 1 : 2) edge_linear_fwd_umma_kernel(EdgeFwdUArgs<T> a) {
    static_assert(!GPF || MODE == 2, "gather prefetch is an edge-update variant");
    constexpr int NTH = FwdCfg<MODE>::NTH;
    constexpr int CW = HH / (NTH / TE);       // accumulator columns per tensor-memory thread
    constexpr int NP = RowRegs<NTH>::NP;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;   // keeps the shared address space
    float* E_hi = smem;                       // e tile; mode 2: later h.  Modes 0/1: staging tile once the MMAs are done
    float* E_lo = E_hi + RT::FLOATS;
    float* ST = E_lo + RT::FLOATS;            // mode 2 only: staging tile
    float* W1_hi = MODE == 2 ? ST + RT::FLOATS : ST;
    float* W1_lo = W1_hi + WT::FLOATS;
    float* W2_hi = W1_lo + WT::FLOATS;        // mode 2 only
    float* W2_lo = W2_hi + WT::FLOATS;
    __shared__ TileSync sync;
    __shared__ int sidx_s[GPF ? 2 : 1][2][TE];   // gather indices of the tile (clamped rows); GPF: this tile's | the next one's

    const int tid = threadIdx.x, warp = tid >> 5;
    const int t = tid & (TE - 1), ch = tid >> 7;           // tensor-memory side: row t, columns CW ch ..
    const int j = tid & 15, g = tid >> 4;                  // global-memory side: chunk j of rows g NP ..
    constexpr uint32_t NCOLS = MODE == 2 ? 128 : 64;
    if (warp == 0) tmem_alloc(&sync.tmem_base, NCOLS);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_init(&sync.bar[1], 1); mbar_fence_init(); }
    load_weight_tiles<false, NTH>(W1_hi, W1_lo, a.W1, tid);
    if (MODE == 2) load_weight_tiles<false, NTH>(W2_hi, W2_lo, a.W2, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t T_Z = tm + lane_off + ch * CW, T_U = tm + 64 + lane_off + ch * CW;
    float* stage = MODE == 2 ? ST : E_hi;
    const float4 bias1 = ldg4(a.b1 + j * 4);
    const float4 bias2 = MODE == 2 ? ldg4(a.b2 + j * 4) : f4(0.f, 0.f, 0.f, 0.f);

    const int64_t ntiles = (a.E + TE - 1) / TE;
    uint32_t it = 0;
    RowRegs<NTH> R, G;
    int nrows_pf = 0, idx_pf = 0;             // thread tid < 256: index (tid / 128) of row tid % 128, one tile ahead
    RowIds<NTH> eids;                         // a.eid: storage rows of the tile AFTER the next one (two tiles ahead)
    const int32_t* idx_src = ch == 0 ? a.idx1 : ch == 1 ? a.idx2 : nullptr;
    if ((int64_t)blockIdx.x < ntiles) {
        nrows_pf = (int)min64(TE, a.E - (int64_t)blockIdx.x * TE);
        load_rows(R, a.e, a.eid, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        if (idx_src) idx_pf = __ldg(idx_src + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
        if (a.eid && (int64_t)blockIdx.x + gridDim.x < ntiles)
            load_row_ids(eids, a.eid, ((int64_t)blockIdx.x + gridDim.x) * TE,
                         (int)min64(TE, a.E - ((int64_t)blockIdx.x + gridDim.x) * TE), tid);
    }

    RowRegs<NTH> G2;                         // mode 2 (registers to spare): the add waits until the prefetch is issued
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE;
        const int nrows = nrows_pf;
        const uint32_t par = it & 1;
        int (*sidx)[TE] = sidx_s[GPF ? (it & 1) : 0];
        const bool own_gather = !GPF || it == 0;          // GPF: later tiles had theirs issued during the previous tile

        store_rows_split(R, E_hi, E_lo, nrows, tid);
        if (own_gather && ch < 2) sidx[ch][t] = idx_pf;
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, E_hi, E_lo, W1_hi, W1_lo, 0);
            mma_commit(&sync.bar[0]);
        }
        if (MODE == 2) { if (own_gather) gather_rows_issue(G, G2, a.T1, sidx[0], a.T2, sidx[1], tid); }
        else gather_rows(G, a.T1, sidx[0], a.T2, sidx[1], tid);   // coalesced, in flight while the MMAs run
        RowRegs<NTH> Rn;                     // mode 2 keeps this tile's e rows for the residual
        int nrows_next = 0;
        if (tile + gridDim.x < ntiles) {     // next tile's e rows in flight
            const int64_t nt = tile + gridDim.x;
            nrows_next = (int)min64(TE, a.E - nt * TE);
            if (a.eid) {
                load_rows_by_id(MODE == 2 ? Rn : R, a.e, eids, tid);       // ids were fetched a tile ago
                if (nt + gridDim.x < ntiles) load_row_ids(eids, a.eid, (nt + gridDim.x) * TE, (int)min64(TE, a.E - (nt + gridDim.x) * TE), tid);
            } else {
                load_rows(MODE == 2 ? Rn : R, a.e, nullptr, nt * TE, nrows_next, tid);
            }
            if (idx_src) idx_pf = __ldg(idx_src + nt * TE + min(t, nrows_next - 1));
        }
        if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < NP; ++i) G.v[i] = G.v[i] + G2.v[i];
        }
        mbar_wait(&sync.bar[0], par);
        fence_after();
        stage_from_tmem<NTH>(T_Z, stage, t, ch);
        __syncthreads();

        // ---- z rows in the coalesced mapping: + bias + gathered node rows (registers)
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            float4 v = lds4(stage + stage_off(r, j)) + bias1 + G.v[i];
            if (MODE >= 1) v = relu4(v);
            if (r < nrows) st_e4(a.out1 + (row0 + r) * HH + j * 4, v);
            if (MODE == 2) {                 // h -> MMA operand (the e tile is dead: its MMAs are done)
                float4 hh, hl;
                split4(r < nrows ? v : f4(0.f, 0.f, 0.f, 0.f), hh, hl);
                const int o = RT::off(r, j);
                st4(E_hi + o, hh);
                st4(E_lo + o, hl);
            }
        }
        if (MODE == 2) {
            const bool has_next = GPF && nrows_next > 0;
            if (has_next && ch < 2) sidx_s[GPF ? ((it + 1) & 1) : 0][ch][t] = idx_pf;      // the next tile's indices (fetched above)
            fence_async_smem();
            fence_before();
            __syncthreads();
            fence_after();
            if (tid == 0) {
                gemm3_data(tm + 64, E_hi, E_lo, W2_hi, W2_lo, 0);
                mma_commit(&sync.bar[1]);
            }
            if (has_next) {                  // G / G2 are dead since the first epilogue: the next tile's node rows take their place
                int (*nidx)[TE] = sidx_s[GPF ? ((it + 1) & 1) : 0];
                gather_rows_issue(G, G2, a.T1, nidx[0], a.T2, nidx[1], tid);
            }
            int64_t orow[NP];
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int r = g * NP + i;
                orow[i] = (a.out_eid && r < nrows) ? (int64_t)__ldg(a.out_eid + row0 + r) : row0 + r;
            }
            mbar_wait(&sync.bar[1], par);
            fence_after();
            stage_from_tmem<NTH>(T_U, stage, t, ch);
            __syncthreads();
#pragma unroll
            for (int i = 0; i < NP; ++i) {
                const int r = g * NP + i;
                const float4 v = a.alpha * R.v[i] + a.beta * (lds4(stage + stage_off(r, j)) + bias2);
                if (r < nrows) st_e4(a.out2 + orow[i] * HH + j * 4, v);
            }
#pragma unroll
            for (int i = 0; i < NP; ++i) R.v[i] = Rn.v[i];
        }
        if (MODE < 2 && a.l2pf && nrows_next > 0 && ch == 0 && a.T1) {      // idx_pf = idx1 of row t of the next tile (arrived by now)
            const float* q = a.T1 + (int64_t)idx_pf * HH;
            prefetch_l2(q);
            prefetch_l2(q + 32);
        }
        nrows_pf = nrows_next;
        fence_before();
        __syncthreads();
        fence_after();
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, NCOLS);
}


// (Measured and dropped: the update forward as two 64-edge pipelines per CTA, the structure that pays in the update
// backward.  10.9 ms vs 9.2 ms on hi_medium -- the forward has no weight-gradient MMAs to hide, and M = 64 MMAs and
// half-filled tensor-memory loads double the per-edge cost of the phases it does have.)


// ---- edge update, software-pipelined over two operand sets -----------------------------------------------------------
// The kernel above runs one tile as a serial chain (split e -> MMA1 -> epilogue 1 -> MMA2 -> epilogue 2): ncu put its
// loads in flight for ~1.5 us of the ~5.6 us a tile takes.  Here the NEXT tile's e rows are split into a second operand
// set while this tile is in its first epilogue, and its MMA1 is issued right behind this tile's MMA2, so that at the top of
// an iteration the first product is already in tensor memory: split / fence / barrier / MMA1 leave the critical path.
// Shared memory: two e / h operand sets (4 x 32 KB), staging tile (32 KB), weights (64 KB) = 224 KB, one CTA per SM.
// The residual rows e[row(k)] are re-read (L2) behind MMA2 instead of being held in registers for a whole iteration:
// the register budget (512 threads x 128) has room for four row sets -- prefetch, residual, two gathered tables.
template <typename T>
__global__ void __launch_bounds__(512, 1) edge_update_fwd_pipe_kernel(EdgeFwdUArgs<T> a) {
    constexpr int NTH = 512;
    constexpr int CW = HH / (NTH / TE);
    constexpr int NP = RowRegs<NTH>::NP;
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;
    auto EBhi = [&](uint32_t set) { return smem + set * 2 * RT::FLOATS; };                 // operand set: hi tile, lo tile
    auto EBlo = [&](uint32_t set) { return smem + set * 2 * RT::FLOATS + RT::FLOATS; };
    float* ST = smem + 4 * RT::FLOATS;
    float* W1_hi = ST + RT::FLOATS;
    float* W1_lo = W1_hi + WT::FLOATS;
    float* W2_hi = W1_lo + WT::FLOATS;
    float* W2_lo = W2_hi + WT::FLOATS;
    __shared__ TileSync sync;
    __shared__ int sidx[2][TE];               // [src | dst][row] of the tile whose gathers are issued next

    const int tid = threadIdx.x, warp = tid >> 5;
    const int t = tid & (TE - 1), ch = tid >> 7;
    const int j = tid & 15, g = tid >> 4;
    if (warp == 0) tmem_alloc(&sync.tmem_base, 128);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_init(&sync.bar[1], 1); mbar_fence_init(); }
    load_weight_tiles<false, NTH>(W1_hi, W1_lo, a.W1, tid);
    load_weight_tiles<false, NTH>(W2_hi, W2_lo, a.W2, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t T_Z = tm + lane_off + ch * CW, T_U = tm + 64 + lane_off + ch * CW;
    const float4 bias1 = ldg4(a.b1 + j * 4), bias2 = ldg4(a.b2 + j * 4);
    const int64_t ntiles = (a.E + TE - 1) / TE;
    const int32_t* idx_src = ch == 0 ? a.idx1 : ch == 1 ? a.idx2 : nullptr;

    RowRegs<NTH> Rp, Rr, G, G2;               // prefetched e rows of a later tile | residual rows of this tile | gathered tables
    int idx_pf = 0;
    int64_t tile = blockIdx.x;
    if (tile < ntiles) {                      // prologue: first tile's e -> set 0, its MMA1, its gathers; second tile's rows in flight
        const int nrows = (int)min64(TE, a.E - tile * TE);
        load_rows(Rp, a.e, a.eid, tile * TE, nrows, tid);
        if (idx_src) idx_pf = __ldg(idx_src + tile * TE + min(t, nrows - 1));
        store_rows_split(Rp, EBhi(0), EBlo(0), nrows, tid);
        if (ch < 2) sidx[ch][t] = idx_pf;
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, EBhi(0), EBlo(0), W1_hi, W1_lo, 0);
            mma_commit(&sync.bar[0]);
        }
        gather_rows_issue(G, G2, a.T1, sidx[0], a.T2, sidx[1], tid);
        const int64_t nt = tile + gridDim.x;
        if (nt < ntiles) {
            const int nn = (int)min64(TE, a.E - nt * TE);
            load_rows(Rp, a.e, a.eid, nt * TE, nn, tid);
            if (idx_src) idx_pf = __ldg(idx_src + nt * TE + min(t, nn - 1));
        }
    }

    uint32_t it = 0;
    for (; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE, nt = tile + gridDim.x;
        const int nrows = (int)min64(TE, a.E - row0);
        const bool has_next = nt < ntiles;
        const int nrows_next = has_next ? (int)min64(TE, a.E - nt * TE) : 0;
        const uint32_t s = it & 1, par = it & 1;
        float* E_hi = EBhi(s);
        float* E_lo = EBlo(s);

        // ---- Z = e W1^T is (long) done: accumulator rows -> staging tile
        mbar_wait(&sync.bar[0], par);
        fence_after();
        stage_from_tmem<NTH>(T_Z, ST, t, ch);
        __syncthreads();

        // ---- epilogue 1: h = relu(z + b1 + P[src] + Q[dst]) -> global and -> this set (MMA2 operand); next tile's e -> other set
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            float4 v = relu4(lds4(ST + stage_off(r, j)) + bias1 + (G.v[i] + G2.v[i]));
            if (r < nrows) st_e4(a.out1 + (row0 + r) * HH + j * 4, v);
            float4 hh, hl;
            split4(r < nrows ? v : f4(0.f, 0.f, 0.f, 0.f), hh, hl);
            const int o = RT::off(r, j);
            st4(E_hi + o, hh);
            st4(E_lo + o, hl);
        }
        if (has_next) {
            store_rows_split(Rp, EBhi(s ^ 1), EBlo(s ^ 1), nrows_next, tid);
            if (ch < 2) sidx[ch][t] = idx_pf;
        }
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm + 64, E_hi, E_lo, W2_hi, W2_lo, 0);
            mma_commit(&sync.bar[1]);
            if (has_next) {                   // the next tile's first product queues right behind
                gemm3_data(tm, EBhi(s ^ 1), EBlo(s ^ 1), W1_hi, W1_lo, 0);
                mma_commit(&sync.bar[0]);
            }
        }
        // ---- behind the MMAs: residual rows of this tile (L2), next tile's gathered rows, the tile after's e rows
        load_rows(Rr, a.e, a.eid, row0, nrows, tid);
        int64_t orow[NP];
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            orow[i] = (a.out_eid && r < nrows) ? (int64_t)__ldg(a.out_eid + row0 + r) : row0 + r;
        }
        if (has_next) {
            gather_rows_issue(G, G2, a.T1, sidx[0], a.T2, sidx[1], tid);
            const int64_t n2 = nt + gridDim.x;
            if (n2 < ntiles) {
                const int nn = (int)min64(TE, a.E - n2 * TE);
                load_rows(Rp, a.e, a.eid, n2 * TE, nn, tid);
                if (idx_src) idx_pf = __ldg(idx_src + n2 * TE + min(t, nn - 1));
            }
        }
        mbar_wait(&sync.bar[1], par);
        fence_after();
        stage_from_tmem<NTH>(T_U, ST, t, ch);
        __syncthreads();
        // ---- epilogue 2: e' = alpha e + beta (h W2^T + b2)
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            const float4 v = a.alpha * Rr.v[i] + a.beta * (lds4(ST + stage_off(r, j)) + bias2);
            if (r < nrows) st_e4(a.out2 + orow[i] * HH + j * 4, v);
        }
        __syncthreads();                      // the staging tile is free for the next iteration's Z rows
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 128);
}

template <typename T>
static int launch_update_fwd_pipe(const EdgeFwdUArgs<T>& a, cudaStream_t st) {
    const size_t smem = 5 * RT::BYTES + 4 * WT::BYTES + 1024;
    auto kern = edge_update_fwd_pipe_kernel<T>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.E + TE - 1) / TE;
    int grid = persistent_grid((const void*)kern, smem, ntiles, 512);
    kern<<<grid, 512, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

template <int MODE, typename T, bool GPF = false>
static int launch_fwd_umma_mode(const EdgeFwdUArgs<T>& a, cudaStream_t st) {
    const size_t smem = (MODE == 2 ? 3 * RT::BYTES + 4 * WT::BYTES : 2 * RT::BYTES + 2 * WT::BYTES) + 1024;
    auto kern = edge_linear_fwd_umma_kernel<MODE, T, GPF>;
    constexpr int NTH = FwdCfg<MODE>::NTH;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.E + TE - 1) / TE;
    int grid = persistent_grid((const void*)kern, smem, ntiles, NTH);
    kern<<<grid, NTH, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    return 0;
}

template <typename T>
static int launch_edge_fwd_umma_t(int mode, int64_t E, const T* e, const int32_t* eid, const float* W1, const float* b1,
                                  const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                                  const float* W2, const float* b2, float alpha, float beta, T* out1, T* out2,
                                  const int32_t* out_eid, cudaStream_t st) {
    const char* lv = getenv("B200MP_L2PF");                 // A/B switch of the L2 announcements (default on)
    EdgeFwdUArgs<T> a{E, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out1, out2, out_eid, (lv && lv[0] == '0') ? 0 : 1};
    switch (mode) {
        case 0: return launch_fwd_umma_mode<0, T>(a, st);
        case 1: return launch_fwd_umma_mode<1, T>(a, st);
        case 2: {
            // B200MP_UPD_PIPE=1: the software-pipelined kernel.  Opt-in: same-box A/B on B200 -- hi_small 1.63 -> 1.54 ms, but
            // hi_medium 9.80 -> 9.94 ms (the residual rows re-read behind the MMAs stop hitting L2 once 148 SMs stream 8 GB tensors)
            const char* v = getenv("B200MP_UPD_PIPE");
            const bool pipe = v && v[0] == '1';
            if (pipe) return launch_update_fwd_pipe<T>(a, st);
            // default: the next tile's node rows gathered behind this tile's second product (same-box A/B on hi_medium, two pairs:
            // 9.86 / 9.92 -> 9.38 / 9.41 ms; hi_small 1.54 -> 1.44 ms; bit-identical outputs); B200MP_UPD_GPF=0 gathers per tile
            const char* gv = getenv("B200MP_UPD_GPF");
            const bool gpf = !(gv && gv[0] == '0');
            return gpf ? launch_fwd_umma_mode<2, T, true>(a, st) : launch_fwd_umma_mode<2, T, false>(a, st);
        }
    }
    return B200MP_EINVAL;
}

int launch_edge_fwd_umma(int mode, int64_t E, const float* e, const int32_t* eid, const float* W1, const float* b1,
                         const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                         const float* W2, const float* b2, float alpha, float beta, float* out1, float* out2,
                         const int32_t* out_eid, cudaStream_t st) {
    return launch_edge_fwd_umma_t<float>(mode, E, e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta, out1, out2, out_eid, st);
}

}  // namespace b200mp

/* bf16 STORAGE of the [E,64] tensors (e, out1, out2), fp32 node tables / weights / arithmetic; tensor-core path only */
extern "C" int b200mp_edge_linear_fwd_bf16(int mode, int64_t E, const void* e, const int32_t* eid, const float* W1, const float* b1,
                                           const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                                           const float* W2, const float* b2, float alpha, float beta, void* out1, void* out2,
                                           const int32_t* out_eid, void* stream) {
    using namespace b200mp;
    if (E < 0 || mode < 0 || mode > 2) return B200MP_EINVAL;
    if (E == 0) return 0;
    if (!W1 || !b1 || !out1 || !e || (T1 && !idx1) || (T2 && !idx2) || (mode == 2 && (!W2 || !b2 || !out2))) return B200MP_EINVAL;
    return launch_edge_fwd_umma_t<ue::bf16>(mode, E, (const ue::bf16*)e, eid, W1, b1, T1, idx1, T2, idx2, W2, b2, alpha, beta,
                                            (ue::bf16*)out1, (ue::bf16*)out2, out_eid, (cudaStream_t)stream);
}

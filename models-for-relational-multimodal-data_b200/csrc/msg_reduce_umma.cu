// Fused gather -> per-edge linear -> segmented reduction for H = 64 (b200mp_message_fwd_reduce): the PNA / GINE message
// forward of edge_fwd_umma.cu (modes 0 / 1) with the destination-wise aggregation of seg_reduce.cu done in the tile
// epilogue, so the messages are written once (backward needs them) and never read back.
//
//   m_k   = e[row(k)] W^T + b + T1[idx1[k]] (+ T2[seg[k]])          KIND 0 (PNA; PyG PNAConv.message, pna.py:71-74)
//   m_k   = relu(e[row(k)] W^T + b + T1[idx1[k]])                   KIND 1 (GINE; gine.py:66-70)
//   KIND 0: agg[n] = [mean | min | max | std] over the messages of segment n, tie[n] = multiplicity of min | max
//   KIND 1: agg[n] = sum                                            (same definitions as seg_reduce.cu)
//
// Edges arrive in segment (CSC) order, so a 128-edge tile holds a run of whole segments plus at most one segment cut at
// each end.  Per tile, after the messages are final in the shared-memory staging tile:
//   - "walkers": thread (c, q) = (column, quarter) scans rows 32 q .. 32 q + 31 of column c top to bottom keeping the
//     running sum / sum of squares / min / max / tie counts of the current run; runs that begin and end inside the
//     quarter are finished on the spot (coalesced 128-byte stores per statistic), the first and last run of a quarter
//     become records in shared memory;
//   - 64 threads stitch the <= 8 quarter records of the tile in row order: runs closed inside the tile are finished,
//     the run touching row 0 ("head") and the run touching the last row ("tail") go to a per-tile record in global memory;
//   - a second kernel stitches the tile records: every segment that crosses a tile boundary (hubs cross thousands) is
//     merged in tile order by the block of the tile it starts in.  No atomics: results are run-to-run identical.
// A third, trivial kernel zeroes the rows of empty segments (no tile ever sees them).
#include <math_constants.h>
#include "umma_edge.cuh"

namespace b200mp {
using namespace ue;

constexpr float MR_STD_EPS = 1e-5f;
constexpr float MR_STD_THR = 0.0031622776601683794f;   // (float)sqrt(1e-5), as seg_reduce.cu

struct MsgRedArgs {
    int64_t E;
    const float* e; const int32_t* eid;
    const float* W; const float* b;
    const float* T1; const int32_t* idx1;     // source-side node table and ids
    const float* T2;                          // destination-side node table (KIND 0) or null, indexed by seg
    const int32_t* seg;                       // aggregation node of sorted edge k
    float* m;
    float* agg; float* tie;
    float* part;                              // [ntiles][2][NST][64] head / tail records
    int32_t* meta;                            // [ntiles][4] = head_seg, head_cnt, tail_seg (-1: none), tail_cnt
};

// running statistics of one column
template <bool STATS>
struct Run {
    float s, q, mn, cn, mx, cx;
    __device__ __forceinline__ void init() { s = q = cn = cx = 0.f; mn = CUDART_INF_F; mx = -CUDART_INF_F; }
    __device__ __forceinline__ void add(float v) {
        s += v;
        if (STATS) {
            q = fmaf(v, v, q);
            const float nmn = fminf(mn, v), nmx = fmaxf(mx, v);
            cn = (v == nmn) ? ((nmn == mn) ? cn + 1.f : 1.f) : cn;
            cx = (v == nmx) ? ((nmx == mx) ? cx + 1.f : 1.f) : cx;
            mn = nmn; mx = nmx;
        }
    }
    __device__ __forceinline__ void merge(const Run& o) {
        s += o.s;
        if (STATS) {
            q += o.q;
            const float nmn = fminf(mn, o.mn), nmx = fmaxf(mx, o.mx);
            cn = (o.mn == nmn ? o.cn : 0.f) + (mn == nmn ? cn : 0.f);
            cx = (o.mx == nmx ? o.cx : 0.f) + (mx == nmx ? cx : 0.f);
            mn = nmn; mx = nmx;
        }
    }
    // record = NST x 64 floats, this thread's column c
    __device__ __forceinline__ void store(float* rec, int c) const {
        rec[c] = s;
        if (STATS) { rec[64 + c] = q; rec[128 + c] = mn; rec[192 + c] = cn; rec[256 + c] = mx; rec[320 + c] = cx; }
    }
    __device__ __forceinline__ void load(const float* rec, int c) {
        s = rec[c];
        if (STATS) { q = rec[64 + c]; mn = rec[128 + c]; cn = rec[192 + c]; mx = rec[256 + c]; cx = rec[320 + c]; }
        else { q = cn = cx = 0.f; mn = mx = 0.f; }
    }
    // the finished segment -> agg / tie (same arithmetic as seg_reduce.cu: finalize_store)
    __device__ __forceinline__ void finish(int64_t node, int cnt, int c, float* __restrict__ agg, float* __restrict__ tie) const {
        if (!STATS) { agg[node * HH + c] = s; return; }
        const float d = (float)cnt, mean = s / d, m2 = q / d;
        const float var = m2 - mean * mean;
        float sd = sqrtf(fmaxf(var, MR_STD_EPS));
        sd = sd <= MR_STD_THR ? 0.f : sd;
        float* o = agg + node * 4 * HH + c;
        o[0] = mean; o[HH] = mn; o[2 * HH] = mx; o[3 * HH] = sd;
        float* t = tie + node * 2 * HH + c;
        t[0] = cn; t[HH] = cx;
    }
};

template <int KIND>
__global__ void __launch_bounds__(256, 2) message_fwd_reduce_umma_kernel(MsgRedArgs a) {
    constexpr int NTH = 256;
    constexpr bool STATS = KIND == 0;
    constexpr int NST = STATS ? 6 : 1;
    constexpr int CW = HH / (NTH / TE);       // 32 accumulator columns per tensor-memory thread
    constexpr int NP = RowRegs<NTH>::NP;      // 8 adjacent rows per thread in the coalesced mapping
    constexpr int REC = NST * HH;             // floats per record
    extern __shared__ __align__(128) float smem_raw[];
    float* smem = smem_raw + (((smem_u32(smem_raw) + 1023u) & ~1023u) - smem_u32(smem_raw)) / 4;
    float* E_hi = smem;                       // e tile; staging tile (final messages) once the MMAs are done
    float* E_lo = E_hi + RT::FLOATS;          // e tile; quarter records once the MMAs are done
    float* W_hi = E_lo + RT::FLOATS;
    float* W_lo = W_hi + WT::FLOATS;
    __shared__ TileSync sync;
    __shared__ int sidx[2][TE];               // [0]: source ids, [1]: segment ids of the tile (clamped rows)
    __shared__ int qmeta[4][4];               // per quarter: head_seg (-1: empty), head_cnt, tail_seg (-1: none), tail_cnt

    const int tid = threadIdx.x, warp = tid >> 5;
    const int t = tid & (TE - 1), ch = tid >> 7;
    const int j = tid & 15, g = tid >> 4;
    if (warp == 0) tmem_alloc(&sync.tmem_base, 64);
    if (tid == 0) { mbar_init(&sync.bar[0], 1); mbar_fence_init(); }
    load_weight_tiles<false, NTH>(W_hi, W_lo, a.W, tid);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tm = sync.tmem_base;
    const uint32_t T_Z = tm + ((uint32_t)((warp & 3) * 32) << 16) + ch * CW;
    float* stage = E_hi;
    float* qrec = E_lo;                       // [4 quarters][2][REC]
    const float4 bias = ldg4(a.b + j * 4);

    const int64_t ntiles = (a.E + TE - 1) / TE;
    uint32_t it = 0;
    RowRegs<NTH> R, G;
    int nrows_pf = 0, idx_pf = 0;
    RowIds<NTH> eids;
    const int32_t* idx_src = ch == 0 ? a.idx1 : a.seg;
    if ((int64_t)blockIdx.x < ntiles) {
        nrows_pf = (int)min64(TE, a.E - (int64_t)blockIdx.x * TE);
        load_rows(R, a.e, a.eid, (int64_t)blockIdx.x * TE, nrows_pf, tid);
        idx_pf = __ldg(idx_src + (int64_t)blockIdx.x * TE + min(t, nrows_pf - 1));
        if (a.eid && (int64_t)blockIdx.x + gridDim.x < ntiles)
            load_row_ids(eids, a.eid, ((int64_t)blockIdx.x + gridDim.x) * TE,
                         (int)min64(TE, a.E - ((int64_t)blockIdx.x + gridDim.x) * TE), tid);
    }

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int64_t row0 = tile * TE;
        const int nrows = nrows_pf;

        store_rows_split(R, E_hi, E_lo, nrows, tid);
        sidx[ch][t] = idx_pf;
        fence_async_smem();
        fence_before();
        __syncthreads();
        fence_after();
        if (tid == 0) {
            gemm3_data(tm, E_hi, E_lo, W_hi, W_lo, 0);
            mma_commit(&sync.bar[0]);
        }
        gather_rows(G, a.T1, sidx[0], a.T2, sidx[1], tid);         // coalesced, in flight while the MMAs run
        int nrows_next = 0;
        if (tile + gridDim.x < ntiles) {                           // next tile's e rows and ids in flight
            const int64_t nt = tile + gridDim.x;
            nrows_next = (int)min64(TE, a.E - nt * TE);
            if (a.eid) {
                load_rows_by_id(R, a.e, eids, tid);
                if (nt + gridDim.x < ntiles) load_row_ids(eids, a.eid, (nt + gridDim.x) * TE, (int)min64(TE, a.E - (nt + gridDim.x) * TE), tid);
            } else {
                load_rows(R, a.e, nullptr, nt * TE, nrows_next, tid);
            }
            idx_pf = __ldg(idx_src + nt * TE + min(t, nrows_next - 1));
        }
        mbar_wait(&sync.bar[0], it & 1);
        fence_after();
        stage_from_tmem<NTH>(T_Z, stage, t, ch);
        __syncthreads();

        // ---- messages in the coalesced mapping: + bias + gathered node rows; to global (for backward) and back to the stage
#pragma unroll
        for (int i = 0; i < NP; ++i) {
            const int r = g * NP + i;
            float* sp = stage + stage_off(r, j);
            float4 v = lds4(sp) + bias + G.v[i];
            if (KIND == 1) v = relu4(v);
            if (r < nrows) st4(a.m + (row0 + r) * HH + j * 4, v);
            st4(sp, v);
        }
        __syncthreads();

        // ---- walkers: column c, rows 32 q .. 32 q + 31
        {
            const int c = tid & 63, q = tid >> 6;
            const int r0 = q * 32, r1 = min(r0 + 32, nrows);
            const int so = ((c >> 2) & 8) << 2, sj = c >> 2, sl = c & 3;      // stage_off pieces of this column
            Run<STATS> run;
            run.init();
            int cur = -1, cnt = 0, hs = -1, hc = 0;
            bool first = true;
            if (r0 < r1) cur = sidx[1][r0];
            for (int r = r0; r < r1; ++r) {
                const int s = sidx[1][r];
                if (s != cur) {
                    if (first) { run.store(qrec + (q * 2) * REC, c); hs = cur; hc = cnt; first = false; }
                    else run.finish(cur, cnt, c, a.agg, a.tie);
                    run.init();
                    cnt = 0;
                    cur = s;
                }
                run.add(stage[r * HH + (((sj ^ r) & 7) << 2) + so + sl]);
                ++cnt;
            }
            int ts = -1, tc = 0;
            if (r0 < r1) {
                if (first) { run.store(qrec + (q * 2) * REC, c); hs = cur; hc = cnt; }        // one run covers the quarter
                else { run.store(qrec + (q * 2 + 1) * REC, c); ts = cur; tc = cnt; }
            }
            if (c == 0) { qmeta[q][0] = hs; qmeta[q][1] = hc; qmeta[q][2] = ts; qmeta[q][3] = tc; }
        }
        __syncthreads();

        // ---- stitch the quarter records in row order (64 threads, one per column)
        if (tid < HH) {
            const int c = tid;
            Run<STATS> carry, o;
            carry.init();
            int cseg = -1, ccnt = 0;
            bool at_start = true;             // the carried run touches row 0 of the tile
            float* ph = a.part + tile * 2 * REC;
            int head_seg = -1, head_cnt = 0;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int hs = qmeta[q][0];
                if (hs >= 0) {
                    o.load(qrec + (q * 2) * REC, c);
                    if (hs == cseg) { carry.merge(o); ccnt += qmeta[q][1]; }
                    else {
                        if (cseg >= 0) {
                            if (at_start) { carry.store(ph, c); head_seg = cseg; head_cnt = ccnt; at_start = false; }
                            else carry.finish(cseg, ccnt, c, a.agg, a.tie);
                        }
                        carry = o; cseg = hs; ccnt = qmeta[q][1];
                    }
                    const int ts = qmeta[q][2];
                    if (ts >= 0) {
                        if (at_start) { carry.store(ph, c); head_seg = cseg; head_cnt = ccnt; at_start = false; }
                        else carry.finish(cseg, ccnt, c, a.agg, a.tie);
                        carry.load(qrec + (q * 2 + 1) * REC, c);
                        cseg = ts; ccnt = qmeta[q][3];
                    }
                }
            }
            int tail_seg = -1, tail_cnt = 0;
            if (at_start) { carry.store(ph, c); head_seg = cseg; head_cnt = ccnt; }            // one run covers the tile
            else { carry.store(ph + REC, c); tail_seg = cseg; tail_cnt = ccnt; }
            if (c == 0) *reinterpret_cast<int4*>(a.meta + tile * 4) = make_int4(head_seg, head_cnt, tail_seg, tail_cnt);
        }
        nrows_pf = nrows_next;
        fence_before();
        __syncthreads();
        fence_after();
    }
    fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tm, 64);
}

// Tile records -> segments that touch a tile boundary.  Block = tile t, thread = column.  A group of records (same segment,
// consecutive in the order head_0, tail_0, head_1, tail_1, ...) is merged, in that order, by the block of the tile it
// starts in: the head record of t starts a group iff the last run of tile t - 1 belongs to another segment; the tail
// record of t always starts one (it follows a different run of the same tile) and continues through the head records of
// the following tiles while they carry the same segment and reach their tile's end.
template <bool STATS>
__global__ void __launch_bounds__(64) tile_records_stitch_kernel(const float* __restrict__ part, const int32_t* __restrict__ meta,
                                                                 int64_t ntiles, float* __restrict__ agg, float* __restrict__ tie) {
    constexpr int NST = STATS ? 6 : 1, REC = NST * HH;
    const int64_t t = blockIdx.x;
    const int c = threadIdx.x;
    const int4 mt = __ldg(reinterpret_cast<const int4*>(meta) + t);
    bool head_starts = true;
    if (t > 0) {
        const int4 mp = __ldg(reinterpret_cast<const int4*>(meta) + t - 1);
        head_starts = (mp.z >= 0 ? mp.z : mp.x) != mt.x;
    }
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
        if (which == 0 ? !head_starts : mt.z < 0) continue;
        Run<STATS> acc, o;
        acc.load(part + (t * 2 + which) * REC, c);
        const int seg = which == 0 ? mt.x : mt.z;
        int64_t cnt = which == 0 ? mt.y : mt.w;
        bool open = which == 1 || mt.z < 0;           // the run reaches the end of tile t
        int64_t u = t + 1;
        while (open && u < ntiles) {
            // up to 8 following tiles at a time: their head records are loaded together (independent addresses), then
            // applied in order until the segment ends
            int4 mu[8];
            float rv[8][NST];
            const int nb = (int)min64(8, ntiles - u);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (k < nb) mu[k] = __ldg(reinterpret_cast<const int4*>(meta) + u + k);
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (k < nb && mu[k].x == seg) {
#pragma unroll
                    for (int s = 0; s < NST; ++s) rv[k][s] = __ldg(part + ((u + k) * 2) * REC + s * HH + c);
                }
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (!open || k >= nb) break;
                if (mu[k].x != seg) { open = false; break; }
                o.s = rv[k][0];
                if (STATS) { o.q = rv[k][1]; o.mn = rv[k][2]; o.cn = rv[k][3]; o.mx = rv[k][4]; o.cx = rv[k][5]; }
                acc.merge(o);
                cnt += mu[k].y;
                if (mu[k].z >= 0) open = false;        // that head run ended inside its tile
            }
            u += 8;
        }
        acc.finish(seg, (int)cnt, c, agg, tie);
    }
}

// rows of empty segments: zero (PyG: scatter into zeros; App. A.2)
__global__ void __launch_bounds__(256) zero_empty_segments_kernel(const int64_t* __restrict__ ptr, int64_t N, int w4a, int w4t,
                                                                  float* __restrict__ agg, float* __restrict__ tie) {
    const int lane = threadIdx.x & 31;
    for (int64_t n = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; n < N; n += ((int64_t)gridDim.x * blockDim.x) >> 5) {
        if (__ldg(ptr + n + 1) != __ldg(ptr + n)) continue;
        const float4 z = f4(0.f, 0.f, 0.f, 0.f);
        for (int k = lane; k < w4a; k += 32) st4(agg + (n * w4a + k) * 4, z);
        for (int k = lane; k < w4t; k += 32) st4(tie + (n * w4t + k) * 4, z);
    }
}

template <int KIND>
static int launch_msg_reduce(const MsgRedArgs& a, const int64_t* ptr, int64_t N, cudaStream_t st) {
    constexpr bool STATS = KIND == 0;
    const size_t smem = 2 * RT::BYTES + 2 * WT::BYTES + 1024;
    auto kern = message_fwd_reduce_umma_kernel<KIND>;
    B200MP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t ntiles = (a.E + TE - 1) / TE;
    if (N > 0) {
        const int64_t want = (N * 32 + 255) / 256, cap = (int64_t)device_info().sms * 16;
        zero_empty_segments_kernel<<<(int)(want < cap ? want : cap), 256, 0, st>>>(ptr, N, STATS ? 64 : 16, STATS ? 32 : 0, a.agg, a.tie);
        B200MP_LAUNCH_CHECK();
    }
    if (ntiles == 0) return 0;
    const int grid = persistent_grid((const void*)kern, smem, ntiles, 256);
    kern<<<grid, 256, smem, st>>>(a);
    B200MP_LAUNCH_CHECK();
    tile_records_stitch_kernel<STATS><<<(unsigned)ntiles, 64, 0, st>>>(a.part, a.meta, ntiles, a.agg, a.tie);
    B200MP_LAUNCH_CHECK();
    return 0;
}

}  // namespace b200mp

extern "C" int b200mp_message_reduce_workspace(int kind, int64_t E, size_t* part_floats, size_t* meta_ints) {
    using namespace b200mp;
    if (!part_floats || !meta_ints || E < 0 || kind < 0 || kind > 1) return B200MP_EINVAL;
    const size_t ntiles = (size_t)((E + TE - 1) / TE);
    *part_floats = ntiles * 2 * (kind == 0 ? 6 : 1) * HH;
    *meta_ints = ntiles * 4;
    return 0;
}

extern "C" int b200mp_message_fwd_reduce(int kind, int H, int64_t E, int64_t N, const int64_t* ptr, const float* e,
                                         const int32_t* eid, const float* W, const float* b, const float* T1,
                                         const int32_t* idx1, const float* T2, const int32_t* seg, float* m, float* agg,
                                         float* tie, float* part, int32_t* meta, void* stream) {
    using namespace b200mp;
    if (H != HH || E < 0 || N < 0 || kind < 0 || kind > 1) return B200MP_EINVAL;
    if (!ptr || !agg || (kind == 0 && !tie)) return B200MP_EINVAL;
    if (E > 0 && (!e || !W || !b || !T1 || !idx1 || !seg || !m || !part || !meta)) return B200MP_EINVAL;
    MsgRedArgs a{E, e, eid, W, b, T1, idx1, kind == 0 ? T2 : nullptr, seg, m, agg, tie, part, meta};
    cudaStream_t st = (cudaStream_t)stream;
    return kind == 0 ? launch_msg_reduce<0>(a, ptr, N, st) : launch_msg_reduce<1>(a, ptr, N, st);
}

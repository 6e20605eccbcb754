/*
 * b200mp.h -- C ABI of the B200-native (sm_100a) message-passing kernels.
 *
 * This is the drop-in boundary one level below the reference's torch.nn.Module API
 * (src/nn/gnn/gine.py:80-91, src/nn/gnn/pna.py:90-101 in the reference).  The reference
 * itself binds no native code on this path: everything below replaces *third-party*
 * operator calls (torch_geometric 2.5.3 / ATen) made from the cited reference lines.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless named host_*; tensors are dense row-major;
 *   - fp32 activations ("float"), int64 indices at the reference-facing surface
 *     (edge_index, perm, colptr), int32 indices inside the kernels (E, N < 2^31);
 *   - `stream` is a cudaStream_t passed as void*; calls enqueue work and return, they never
 *     synchronise, never allocate device memory (the caller passes workspaces) and never
 *     throw;  return value 0 = ok, >0 = cudaError_t of the failed launch, <0 = B200MP_E*.
 *   - "segment order": edges sorted (stably) by the node they are aggregated at.  For the
 *     forward direction that is CSC order (sorted by destination); e-storage is kept in that
 *     order for the whole backbone, so edge tensors stream sequentially from HBM.
 */
#ifndef B200MP_H
#define B200MP_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200MP_EINVAL   (-1)   /* bad argument (unsupported H, null pointer, ...) */
#define B200MP_EWORKSP  (-2)   /* workspace too small */
#define B200MP_ELIMIT   (-3)   /* size beyond the fixed maximum of an entry point that owns its scratch memory */

/* Library / device introspection. */
int b200mp_version(void);
/* Number of CTAs the persistent edge kernels are launched with (= rows of the per-CTA
 * weight-gradient partial buffers the caller must allocate). */
int b200mp_partial_rows(int* rows);
/* The same for hidden width H (the tensor-core kernels for H = 64 run one CTA per SM). */
int b200mp_partial_rows_for(int H, int* rows);
/* Edges per tile of the per-edge kernels for hidden width H (callers size n_part with it). */
int b200mp_tile_edges(int H, int* te);

/* ------------------------------------------------------------------ graph structure
 * Replaces torch_geometric `to_csc` reached from create_graph (src/datasets/util/graph.py:31-57)
 * and `sampler.edge_permutation` (src/datasets/ibm_transactions_for_aml.py:91):
 *   perm   = stable argsort(key)          int64 [E]
 *   ptr    = index2ptr(key[perm], N)      int64 [N+1]
 *   other  = other_end[perm]              int64 [E]   (CSC `row` when key = dst)
 *   other32/key32 = int32 copies in sorted order (kernel-side arrays)
 * Bit-exact. */
int b200mp_csc_workspace_bytes(int64_t E, int64_t N, size_t* bytes);
int b200mp_csc_build(const int64_t* key, const int64_t* other_end, int64_t E, int64_t N,
                     int64_t* perm, int64_t* ptr, int64_t* other,
                     int32_t* other32, int32_t* key32,
                     void* workspace, size_t workspace_bytes, void* stream);

/* In-degree and its histogram (main.py:284-286, utils.py:179-182):
 *   deg[v] = ptr[v+1]-ptr[v];  *max_deg (device scalar) = max(deg);  hist = bincount(deg). */
int b200mp_degree(const int64_t* ptr, int64_t N, int64_t* deg, int64_t* max_deg, void* stream);
int b200mp_bincount(const int64_t* vals, int64_t n, int64_t* hist, int64_t nbins, void* stream);

/* Batch relabel (src/utils/batch_processing.py:65-72, ibm_transactions_for_aml.py:166-173):
 *   nodes = unique(ids) ascending (first *n_unique entries of `nodes` valid),
 *   local[i] = rank of ids[i] in nodes. */
int b200mp_relabel_workspace_bytes(int64_t M, size_t* bytes);
int b200mp_relabel(const int64_t* ids, int64_t M, int64_t* nodes, int64_t* n_unique,
                   int64_t* local, void* workspace, size_t workspace_bytes, void* stream);

/* b200mp_relabel without any allocation on either side of the call.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results (bit-exact) as b200mp_relabel; the workspace is a file-scope
 * `static __device__` array of B200MP_RELABEL_STATIC_BYTES inside the library (held by every GPU that loads it until exit),
 * resolved once per device with cudaGetSymbolAddress.  Limit: M <= B200MP_RELABEL_STATIC_MAX_IDS, beyond it -> B200MP_ELIMIT
 * (nothing is launched).  The entry point calls no cudaMalloc / cudaFree / allocating library routine.  All calls share the one
 * array (it keeps the previous call's contents; nothing in it needs to start at zero): calls that overlap on the GPU --
 * different streams or host threads on one device -- race; serialise them on one stream. */
#define B200MP_RELABEL_STATIC_MAX_IDS 262144
#define B200MP_RELABEL_STATIC_BYTES   (8u << 20)
int b200mp_relabel_static(const int64_t* ids, int64_t M, int64_t* nodes, int64_t* n_unique, int64_t* local, void* stream);

/* Row permutation of an [n,H] fp32 matrix: out[i] = in[idx[i]] / out[idx[i]] = in[i]. */
int b200mp_gather_rows(const float* in, const int64_t* idx, float* out, int64_t n, int H, void* stream);
int b200mp_scatter_rows(const float* in, const int64_t* idx, float* out, int64_t n, int H, void* stream);

/* x = hi + lo, hi = x rounded to tf32, lo = x - hi (exact); n a multiple of 4.  Operand split of
 * the node-level 3xTF32 linears (functional.linear3x): three tensor-core library GEMMs
 * lo*hi + hi*lo + hi*hi reproduce the fp32 product to ~1e-7 (PyG Linear / torch addmm call sites:
 * pna.py:71-79, gine.py:66-70 through PNAConv.update / GINEConv.nn). */
int b200mp_split_tf32(const float* x, float* hi, float* lo, int64_t n, void* stream);

/* Node-level linear on the tensor cores (tcgen05, 3xTF32; fp32 accuracy ~1e-7), one pass over HBM:
 *   Y[N, NO] = X[N, K] W^T (+ bias)         X rows at stride ldx floats; K % 32 == 0; NO in {64,128,192,256}
 * Wimg is the weight pre-split and pre-tiled by the caller (tiny): [K/32][2 (hi, lo)][8][NO][4] floats,
 * Wimg[c][p][q][n][i] = part_p(W[n][32 c + 4 q + i]) with hi = W rounded to tf32, lo = W - hi.
 * Replaces torch addmm at PyG PNAConv.update's post_nn (pna.py:71-79) and, with W^T, its data gradient. */
int b200mp_node_linear(const float* X, int64_t N, int K, int ldx, const float* Wimg, int NO, const float* bias,
                       float* Y, void* stream);
/* The same with the rows of Y at stride ldy >= NO floats: a call fills a column block of a wider output, so linears with more than
 * 256 outputs (PNAConv post_nn at H = 128: 3 scalers x 128 = 384, its data gradient 512) are column blocks of <= 256. */
int b200mp_node_linear_ld(const float* X, int64_t N, int K, int ldx, const float* Wimg, int NO, const float* bias,
                          float* Y, int ldy, void* stream);
/* Weight gradient of the above: part[p][k][n] = sum over the rows handled by CTA p of X[r][k] Y[r][n]
 * (K % 128 == 0, NO in {64,128,192,256}); the caller sums the n_part partials (deterministic). */
int b200mp_node_linear_wgrad(const float* X, int ldx, int K, const float* Y, int ldy, int NO, int64_t N,
                             float* part, int n_part, void* stream);
/* The same with row maps: reduction row r reads X[xrow[r]] and Y[yrow[r]] (NULL = r) -- the weight gradients of the H = 128 per-edge
 * products pair gradient rows in segment order with edge features in the caller's order (b200mp_row_linear). */
int b200mp_node_linear_wgrad_rows(const float* X, const int32_t* xrow, int ldx, int K, const float* Y, const int32_t* yrow,
                                  int ldy, int NO, int64_t N, float* part, int n_part, void* stream);
/* Several H = 64 node linears in one pass: X = [X0 | X1 | X2] (nx blocks of [N,64], e.g. the gradients of
 * A | B | out0), Y = X W^T + bias, NO in {64,128,192,256}; ysplit != 0 writes Y as NO/64 separate [N,64]
 * blocks (A, B, out0 of PNAConv's hoisted pre_nn / post_nn self term; P, Q of the edge update). */
int b200mp_node_linear64(const float* X0, const float* X1, const float* X2, int nx, int64_t N, const float* Wimg,
                         int NO, const float* bias, float* Y, int ysplit, void* stream);
/* b200mp_node_linear64 (NO = 64, Y [N,64]) that also returns the column sums of X = [X0 | X1 | X2] as per-CTA partials
 * colsum_part [n_part][64 nx] (zeroed by the caller; summed by the caller, fixed order): with X = the output gradients of the Q | K | V
 * projections (or of one linear) of nn.TransformerEncoderLayer (tabgnn.py:116-128 through torch MultiheadAttention / Linear backward),
 * one pass yields the data gradient and the bias gradients, which torch computes with a separate sum over every gradient tensor. */
int b200mp_node_linear64_colsum(const float* X0, const float* X1, const float* X2, int nx, int64_t N, const float* Wimg,
                                const float* bias, float* Y, float* colsum_part, int n_part, void* stream);
/* Weight gradients of two H = 64 node linears sharing their input Y (PNAConv pre_nn halves A | B, edge
 * update halves P | Q, pna.py:71-84): part[p][0:64] = G0^T Y, part[p][64:128] = G1^T Y; all [N,64].  G1 == NULL: one gradient
 * (G1 is not read, part[p][64:128] = 0). */
int b200mp_node_pair_wgrad(const float* G0, const float* G1, const float* Y, int64_t N, float* part, int n_part,
                           void* stream);

/* ------------------------------------------------------------------ per-edge linear (forward)
 * One pass over E edges in segment order k = 0..E-1; H in {16,32,64,128}.
 *   z_k   = e[row(k)] W1^T + b1 + T1[idx1[k]] + T2[idx2[k]]        (T1/T2 nullable)
 *   mode 0: out1[k] = z_k                      PNAConv message with hoisted node terms
 *                                              (pna.py:71-74 -> PyG PNAConv.message)
 *   mode 1: out1[k] = relu(z_k)                GINEConv message (gine.py:66-70)
 *   mode 2: out1[k] = h_k = relu(z_k);  out2[orow(k)] = alpha e[row(k)] + beta (h_k W2^T + b2)
 *                                              edge update (gine.py:88-89, pna.py:97-99,
 *                                              models/fused.py:254)
 * row(k) = eid ? eid[k] : k  (eid maps segment order -> e-storage row; null = identity);
 * orow(k) = out_eid ? out_eid[k] : k.  The two maps let the first layer read the caller's edge
 * order and the last one write it, with no separate permutation pass
 * (the backward row maps g_eid / e_eid / g_acc exist on the tensor-core path, H = 64, only). */
int b200mp_edge_linear_fwd(int mode, int H, int64_t E,
                           const float* e, const int32_t* eid,
                           const float* W1, const float* b1,
                           const float* T1, const int32_t* idx1,
                           const float* T2, const int32_t* idx2,
                           const float* W2, const float* b2, float alpha, float beta,
                           float* out1, float* out2, const int32_t* out_eid, void* stream);

/* ------------------------------------------------------------------ segmented reductions
 * Replaces PyG DegreeScalerAggregation's scatter kernels (mean/min/max/std; App. A.2 of
 * SURVEY.md) and GINEConv's scatter_add, without atomics.
 * Segments with more than `light_max` rows are split into chunks of `chunk` rows listed in
 * hchunk_* (built once per graph), reduced to partials and combined in order.
 *   kind 0 (PNA): agg[N,4H] = [mean | min | max | std], tie[N,2H] = multiplicity of min | max
 *   kind 1 (sum): agg[N,H]  = sum                              (tie unused) */
int b200mp_segment_reduce_fwd(int kind, int H, int64_t N, const int64_t* ptr, const float* m,
                              float* agg, float* tie,
                              int64_t light_max, int64_t chunk,
                              int64_t n_heavy, const int32_t* heavy_node, const int32_t* heavy_chunk_ptr,
                              int64_t n_chunks, const int32_t* chunk_heavy, const int64_t* chunk_begin,
                              float* partial /* [n_chunks, 6H] */, void* stream);

/* Fused message forward + reduction for H = 64 on the tensor-core path (csrc/msg_reduce_umma.cu): what
 * b200mp_edge_linear_fwd modes 0 / 1 followed by b200mp_segment_reduce_fwd kinds 0 / 1 compute, in one pass over the
 * edges -- the messages m are written once (the backward reads them) and never read back; no atomics.
 *   kind 0 (PNA):  m_k = e[row(k)] W^T + b + T1[idx1[k]] + T2[seg[k]];  agg [N,256] = mean|min|max|std, tie [N,128]
 *   kind 1 (GINE): m_k = relu(e[row(k)] W^T + b + T1[idx1[k]]);          agg [N,64] = sum (tie, T2 unused)
 * seg[k] = aggregation node of sorted edge k (non-decreasing), ptr = the segment boundaries [N+1].
 * part / meta: scratch of the sizes b200mp_message_reduce_workspace reports (records of the segments cut by tile
 * boundaries, stitched by a second kernel in tile order). */
int b200mp_message_reduce_workspace(int kind, int64_t E, size_t* part_floats, size_t* meta_ints);
int b200mp_message_fwd_reduce(int kind, int H, int64_t E, int64_t N, const int64_t* ptr, const float* e,
                              const int32_t* eid, const float* W, const float* b, const float* T1,
                              const int32_t* idx1, const float* T2, const int32_t* seg, float* m, float* agg,
                              float* tie, float* part, int32_t* meta, void* stream);

/* ------------------------------------------------------------------ node-level PNA glue
 * Degree scalers of PyG DegreeScalerAggregation (constructed at pna.py:71-74)
 * commuted past post_nn's Linear:  out[N,F] = base + sum_s f_s(deg) Z[:, s*F:(s+1)*F],
 * deg = ptr[v+1]-ptr[v]; host_codes[s]: 0 identity, 1 amplification log(d+1)/avg_log,
 * 2 attenuation avg_log/log(max(d,1)+1), 3 linear d/avg_lin, 4 inverse_linear avg_lin/max(d,1).
 * avg_lin / avg_log are DEVICE scalars (the module's buffers).  _bwd: gZ[:, s] = f_s(deg) g. */
int b200mp_pna_scale_fwd(int F, int64_t N, const int64_t* ptr, const float* base, const float* Z, int ns,
                         const int* host_codes, const float* avg_lin, const float* avg_log, float* out, void* stream);
int b200mp_pna_scale_bwd(int F, int64_t N, const int64_t* ptr, const float* g, int ns, const int* host_codes,
                         const float* avg_lin, const float* avg_log, float* gZ, void* stream);
/* Coefficient rows of b200mp_message_bwd kind 0 from the aggregate gradient g_agg[N,4H]
 * (mean|min|max|std), agg[N,4H] and tie[N,2H]:  coef[N,7H], gA[N,H] = (g_mean+g_min+g_max)[deg>0]
 * (gA nullable). */
int b200mp_pna_coef(int H, int64_t N, const int64_t* ptr, const float* g_agg, const float* agg, const float* tie,
                    float* coef, float* gA, void* stream);

/* ------------------------------------------------------------------ per-edge backward
 * Message backward (autograd of the mode 0 / mode 1 forward + its reduction):
 *   kind 0 (PNA):  G_k = c0[d] + c1[d] (m_k - mean[d]) + gmin'[d] [m_k==min[d]] + gmax'[d] [m_k==max[d]]
 *                  with node_in = coef[N,7H] = [c0 | c1 | gmin' | min | gmax' | max | mean]
 *                  (c0 = g_mean/deg, c1 = g_std/(deg std) where std is not clamped/masked, else 0,
 *                   gmin' = g_min / #ties, gmax' likewise)
 *   kind 1 (GINE): G_k = node_in[d] [m_k > 0],  node_in = grad_agg[N,H]
 *   d = seg[k].  Then
 *     gT1[nbr[k]] += G_k                  (vector atomics)
 *     g_e[row(k)] (+)= G_k W (+ g_acc[k])  (accumulate != 0 adds to the existing value; g_acc:
 *                                          optional [E,H] in segment order, H = 64 only -- the
 *                                          gradient the same layer's edge update left for e)
 *     gW_part[cta]  = sum_k G_k^T e_k ,  gb_part[cta] = sum_k G_k     (per-CTA partials) */
int b200mp_message_bwd(int kind, int H, int64_t E,
                       const float* m, const float* e, const int32_t* eid,
                       const float* node_in, const int32_t* seg, const int32_t* nbr,
                       const float* W, float* g_e, int accumulate, float* gT1,
                       float* gW_part, float* gb_part, int n_part, const float* g_acc, void* stream);

/* Edge-update backward (mode 2).  g_out = dL/d out2 [E,H]; row k of the segment order is read at
 * g_eid ? g_eid[k] : k, e at e_eid ? e_eid[k] : k (maps: H = 64 only); h and g_e are in segment order.
 *   g_e = alpha g_out + Gh W1 ,  Gh = (beta g_out W2) * [h>0]
 *   gP[src[k]] += Gh_k ,  gQ[dst[k]] += Gh_k  (run-length reduced over equal consecutive dst)
 *   partials of grad W1, b1, W2, b2 per CTA. */
int b200mp_edge_update_bwd(int H, int64_t E,
                           const float* g_out, const float* h, const float* e,
                           const int32_t* src, const int32_t* dst,
                           const float* W1, const float* W2, float alpha, float beta,
                           float* g_e, float* gP, float* gQ,
                           float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part,
                           int n_part, const int32_t* g_eid, const int32_t* e_eid, void* stream);

/* ------------------------------------------------------------------ column attention core
 * softmax(Q K^T / sqrt(d)) V of nn.MultiheadAttention inside the reference's TransformerEncoderLayer
 * (src/nn/models/tabgnn.py:199-209, fused.py:83-93,187-197, inteleaved.py:155) over column tokens:
 * q, k, v, out, gradients: [R, S, C] fp32 contiguous, S <= 8, head_dim = C / nhead in {4, 8, 16}.
 * p_drop: dropout on the attention probabilities (training), mask = hash(seed, row, head, i, j), the
 * backward recomputes it from the same seed. */
int b200mp_col_attention_fwd(const float* q, const float* k, const float* v, int64_t R, int S, int C, int nhead,
                             float p_drop, uint64_t seed, float* out, void* stream);
int b200mp_col_attention_bwd(const float* q, const float* k, const float* v, const float* g_out, int64_t R, int S,
                             int C, int nhead, float p_drop, uint64_t seed, float* gq, float* gk, float* gv,
                             void* stream);

/* ------------------------------------------------------------------ BatchNorm + ReLU + residual average
 * x' = (x + relu(BatchNorm(c))) / 2 over node rows [N, H], H in {16, 32, 64, 128}, training-mode statistics: the tail of
 * every backbone layer (src/nn/gnn/gine.py:87, pna.py:96; models/tabgnn.py:185-188).  sums: fp64 [2, H], zeroed by the
 * caller, accumulated into.  mean_rstd: fp32 [2, H] = (mean, 1 / sqrt(biased var + eps)) computed by the caller from sums.
 *   b200mp_bn_stats                 sums = (sum c, sum c^2) per column
 *   b200mp_bn_relu_res_fwd          out = (x + relu((c - mean) rstd gamma + beta)) / 2
 *   b200mp_bn_relu_res_bwd_reduce   sums = (sum gr, sum gr chat), gr = g / 2 where BatchNorm(c) > 0  (= g_beta, g_gamma)
 *   b200mp_bn_relu_res_bwd_apply    gc = gamma rstd (gr - gmeans[0] - chat gmeans[1]),  gx = g / 2;  gmeans = sums / N */
int b200mp_bn_stats(const float* c, int64_t N, int H, double* sums, void* stream);
int b200mp_bn_relu_res_fwd(const float* c, const float* x, int64_t N, int H, const float* gamma, const float* beta,
                           const float* mean_rstd, float* out, void* stream);
int b200mp_bn_relu_res_bwd_reduce(const float* g, const float* c, int64_t N, int H, const float* gamma, const float* beta,
                                  const float* mean_rstd, double* sums, void* stream);
int b200mp_bn_relu_res_bwd_apply(const float* g, const float* c, int64_t N, int H, const float* gamma, const float* beta,
                                 const float* mean_rstd, const float* gmeans, float* gc, float* gx, void* stream);

/* ------------------------------------------------------------------ residual + dropout + LayerNorm
 * y = LayerNorm(x + dropout(sub)) over rows of C in {16, 32, 64, 128} floats (sub == NULL: y = LayerNorm(x)): norm1 /
 * norm2 with dropout1 / dropout2 of the post-norm nn.TransformerEncoderLayer and the LayerNorm after it
 * (src/nn/models/tabgnn.py:199-209,219, fused.py:83-93,249).  stats [rows, 2] = (mean, rstd) for backward.
 * Backward writes g_x, g_sub (NULL allowed when p_drop == 0: it equals g_x) and ACCUMULATES into ggamma / gbeta [C]. */
int b200mp_add_dropout_layernorm_fwd(const float* x, const float* sub, int64_t rows, int C, float p_drop, uint64_t seed,
                                     const float* gamma, const float* beta, float eps, float* y, float* stats, void* stream);
int b200mp_add_dropout_layernorm_bwd(const float* g, const float* x, const float* sub, const float* stats, int64_t rows,
                                     int C, float p_drop, uint64_t seed, const float* gamma, float* gx, float* gsub,
                                     float* ggamma, float* gbeta, void* stream);

/* Extended forms for a STACK of encoder layers evaluated as one unit (tabgnn.py:116-131: `x = (x + layer(x)) / 2` per layer, then
 * `(x + x_tab) / 2` into the embeddings): the residual blends are written by the layer's last LayerNorm instead of separate add / div
 * passes, and a gradient that is still a sum of terms is consumed as such:
 *   forward   y = yscale * LayerNorm(x + dropout(sub)) + bs0 * b0 + bs1 * b1       (b0, b1 nullable; b1 needs b0)
 *   backward  as above with the incoming gradient g = gs0 * g0 + gs1 * g1 + gs2 * g2  (g1, g2 nullable; g2 needs g1)
 * b200mp_scaled_sum: out = sum_{i<n} scales[i] * terms[i], n <= 6 tensors of numel floats (numel % 4 == 0; `terms` / `scales` are
 * HOST arrays): writes such a gradient once where it leaves the stack. */
int b200mp_add_dropout_layernorm_fwd_ex(const float* x, const float* sub, int64_t rows, int C, float p_drop, uint64_t seed,
                                        const float* gamma, const float* beta, float eps, float yscale, const float* b0, float bs0,
                                        const float* b1, float bs1, float* y, float* stats, void* stream);
int b200mp_add_dropout_layernorm_bwd_ex(const float* g0, float gs0, const float* g1, float gs1, const float* g2, float gs2,
                                        const float* x, const float* sub, const float* stats, int64_t rows, int C, float p_drop,
                                        uint64_t seed, const float* gamma, float* gx, float* gsub, float* ggamma, float* gbeta,
                                        void* stream);
int b200mp_scaled_sum(const float* const* terms, const float* scales, int n, int64_t numel, float* out, void* stream);

/* out = dropout(relu(x)) (g == NULL, first argument x) or its backward out = g * keep_scale where y > 0 (g != NULL, first
 * argument the saved forward output y): the feed-forward's inner activation of the encoder layer.  numel % 4 == 0. */
int b200mp_relu_dropout(const float* x_or_y, const float* g, int64_t numel, float p_drop, uint64_t seed, float* out,
                        void* stream);

/* ------------------------------------------------------------------ column (stype-wise) encoders
 * What the reference's StypeWiseFeatureEncoder call computes (src/utils.py:139-140,357-359, fused.py:71-72;
 * encoders chosen at src/datasets/ibm_transactions_for_aml.py:283-294): every column of every stype -> one
 * [C] token, all tokens of a row written once as out [rows, S, C], S = ncol (+ 1 when cls != NULL: the CLS
 * token of src/nn/models/tabgnn.py:214 / fused.py:149 goes to slot 0).  Columns appear in output order. */
enum { B200MP_COL_NUMERICAL = 0, B200MP_COL_CATEGORICAL = 1, B200MP_COL_TIMESTAMP = 2 };
typedef struct b200mp_column {
    int32_t kind;        /* B200MP_COL_* */
    int32_t offset;      /* element offset of this column inside a row of feat */
    int64_t stride;      /* elements per row of feat */
    const void* feat;    /* numerical: float32 [rows, stride]; categorical: int64 ids, -1 = missing;
                            timestamp: int64, 7 consecutive fields (year, month, day, weekday, hour, minute, second) */
    const float* weight; /* numerical [C]; categorical [ncat + 1, C], row 0 = missing; timestamp [7 * 8, C] */
    const float* bias;   /* numerical, timestamp: [C]; categorical: NULL */
    float mean, std;     /* numerical: x = (v - mean) / std * weight + bias */
    float min_year;      /* timestamp: the positional encoding takes year - min_year */
    int32_t ncat;        /* categorical: number of categories */
} b200mp_column;
int b200mp_col_encode_fwd(const b200mp_column* cols, int ncol, const float* cls, int64_t rows, int C, float* out,
                          void* stream);
/* parameter gradients (the inputs are data): g_weight[i] / g_bias[i] have the shape of cols[i].weight / bias and
 * are ACCUMULATED into (zero them first); g_bias[i] is NULL for categorical columns; g_cls [C] when has_cls. */
int b200mp_col_encode_bwd(const b200mp_column* cols, int ncol, int has_cls, int64_t rows, int C, const float* g_out,
                          float* const* g_weight, float* const* g_bias, float* g_cls, void* stream);

/* ------------------------------------------------------------------ batch construction
 * Negative sampler (replaces src/primitives/negative_sampling/negative_sampling.cpp:10-76, the
 * reference's only native component, called at src/utils/batch_processing.py:145).  edge_index [2,E]
 * and pos [2,B] are int64 device arrays of batch-local node ids (< 2^31).  Per positive (src, dst):
 * num_neg/2 edges (src, n) then num_neg/2 edges (n, dst); n uniform on [0, num_nodes) excluding src, dst
 * and the undirected neighbours of src and dst in edge_index.  num_nodes <= 0: the number of distinct
 * ids in edge_index (what the reference's nodeset.size() is).  neg: [2, B * 2 * (num_neg/2)].  The draw
 * is a hash of (seed, output slot, attempt): reproducible (the reference's std::random_device is not).
 * num_neg <= 0 -> B200MP_EINVAL (reference: std::invalid_argument). */
int b200mp_negative_sample_workspace_bytes(int64_t E, size_t* bytes);
int b200mp_negative_sample(const int64_t* edge_index, int64_t E, int64_t num_nodes, const int64_t* pos, int64_t B,
                           int num_neg, uint64_t seed, int64_t* neg, void* workspace, size_t workspace_bytes,
                           void* stream);

/* CSC neighbour sampler, one hop (PyG NeighborSampler.sample_from_edges / pyg-lib neighbor_sample as
 * reached from src/datasets/ibm_transactions_for_aml.py:88-102; fan-out [100,100], non-disjoint, no
 * replacement).  For frontier nodes f[0..F) in order: take[i] = min(in-degree, fanout) (fanout < 0: all);
 * after the caller's exclusive scan `offs`, out_pos[offs[i] + j] = CSC positions of the taken in-edges:
 * the whole slice in slice order when deg <= fanout (bit-exact), else `fanout` distinct positions drawn
 * with Floyd's algorithm from a hash of (seed, node).
 * b200mp_first_occurrence: flags[k] = 1 iff cand[k] is an unvisited node and k is the first position
 * holding it (discovery order of the next frontier); marks those nodes visited.  visited: [N] bytes,
 * first: [N] scratch preset to 0xFF..FF (left in that state). */
int b200mp_sample_hop_count(const int64_t* colptr, const int64_t* frontier, int64_t F, int64_t fanout, int64_t* take,
                            void* stream);
int b200mp_sample_hop_fill(const int64_t* colptr, const int64_t* frontier, int64_t F, int64_t fanout,
                           const int64_t* offs, uint64_t seed, int64_t* out_pos, void* stream);
int b200mp_first_occurrence(const int64_t* cand, int64_t K, uint8_t* visited, uint64_t* first, uint8_t* flags,
                            void* stream);

/* Atomic-free variants of the two backward kernels (H = 64, tensor-core path): the per-edge gradient rows that
 * b200mp_message_bwd / b200mp_edge_update_bwd scatter into node tables with vector REDs are written once as [E,H] in
 * segment order instead (G_rows: d loss / d m_k; Gh_rows: the gradient at the edge update's hidden layer), and the caller sums
 * them per node over the CSC / CSR orders with b200mp_segment_reduce_fwd kind 1.  Everything else (g_e, weight / bias partials,
 * row maps, g_acc) as in the scattering entry points.  Results are bit-identical from run to run. */
int b200mp_message_bwd_rows(int kind, int64_t E, const float* m, const float* e, const int32_t* eid,
                            const float* node_in, const int32_t* seg, const float* W, float* g_e, int accumulate,
                            float* G_rows, float* gW_part, float* gb_part, int n_part, const float* g_acc, void* stream);
int b200mp_edge_update_bwd_rows(int64_t E, const float* g_out, const float* h, const float* e, const float* W1,
                                const float* W2, float alpha, float beta, float* g_e, float* Gh_rows, float* gW1_part,
                                float* gb1_part, float* gW2_part, float* gb2_part, int n_part, const int32_t* g_eid,
                                const int32_t* e_eid, const int32_t* src, const int32_t* dst, void* stream);

/* ------------------------------------------------------------------ bf16 storage (BASELINE.json config 3)
 * The per-edge kernels and the segmented reduction with every [E,64] tensor (e, m, h, e', the gradients) STORED as bf16 -- half
 * the bytes of the streams that dominate the step -- while node tables, weights, partials and all arithmetic stay fp32 (bf16
 * values are split exactly into the tf32 operands, accumulation in fp32 tensor memory).  Same contracts as the fp32 entry
 * points (H = 64, tensor-core path); `void*` arguments are bf16 device arrays. */
int b200mp_edge_linear_fwd_bf16(int mode, int64_t E, const void* e, const int32_t* eid, const float* W1, const float* b1,
                                const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                                const float* W2, const float* b2, float alpha, float beta, void* out1, void* out2,
                                const int32_t* out_eid, void* stream);
int b200mp_segment_reduce_fwd_bf16(int kind, int64_t N, const int64_t* ptr, const void* m, float* agg, float* tie,
                                   int64_t light_max, int64_t chunk, int64_t n_heavy, const int32_t* heavy_node,
                                   const int32_t* heavy_chunk_ptr, int64_t n_chunks, const int32_t* chunk_heavy,
                                   const int64_t* chunk_begin, float* partial, void* stream);
int b200mp_message_bwd_bf16(int kind, int64_t E, const void* m, const void* e, const int32_t* eid, const float* node_in,
                            const int32_t* seg, const int32_t* nbr, const float* W, void* g_e, int accumulate, float* gT1,
                            float* gW_part, float* gb_part, int n_part, const void* g_acc, void* stream);
int b200mp_edge_update_bwd_bf16(int64_t E, const void* g_out, const void* h, const void* e, const int32_t* src,
                                const int32_t* dst, const float* W1, const float* W2, float alpha, float beta, void* g_e,
                                float* gP, float* gQ, float* gW1_part, float* gb1_part, float* gW2_part, float* gb2_part,
                                int n_part, const int32_t* g_eid, const int32_t* e_eid, void* stream);

/* ------------------------------------------------------------------ read-out heads, ranking metric, pooled write-back
 * (kernels in csrc/readout.cu)
 *
 * b200mp_pair_gather_act: first layer of the reference's edge read-out heads with its two node gathers
 *   (src/nn/gnn/decoder.py:17-21 ClassifierHead.forward, :62-71 LinkPredHead.forward):
 *     out[e] = act(Rs[src[e]] + Rd[dst[e]] + T[e]),  act = relu (relu != 0) or identity
 *   with Rs = relu(x) Ws^T, Rd = relu(x) Wd^T ([N, W]) and T = edge_attr We^T + b ([E, W]) computed by the caller: relu is
 *   element-wise, so this equals Linear([relu(x[src]) | relu(x[dst]) | edge_attr]) without ever building the [E, 3H]
 *   concatenation.  Any of Rs / Rd / T may be NULL.  W % 4 == 0; src / dst int64 (edge_index rows).
 * b200mp_rank_positive: SSMetric.mrr's ranking (src/utils/metric.py:30-43): rank[b] = position (1-based) of pos[b] in the
 *   descending order of [neg[b, 0..K), pos[b]], the positive last among equal scores (a stable argsort of -combined).
 * b200mp_segment_mean_blend_fwd / _bwd: pooled write-back of the fused layer (src/nn/models/fused.py:256-268):
 *   index = target_edge_index.flatten() sorted by node with b200mp_csc_build (ptr [N+1], perm [R]);
 *     out[n] = deg_n > 0 ? (x[n] + mean_{k in segment n} emb[perm[k]]) / 2 : x[n]
 *   backward: gx[n] = g[n] * (deg_n > 0 ? 1/2 : 1),  gemb[r] = g[index[r]] / (2 deg_index[r])  (gx / gemb nullable).
 * No atomics: all four are run-to-run identical. */
int b200mp_pair_gather_act(const float* Rs, const float* Rd, const float* T, const int64_t* src, const int64_t* dst,
                           int64_t E, int W, int relu, float* out, void* stream);
int b200mp_rank_positive(const float* pos, const float* neg, int64_t B, int K, int32_t* rank, void* stream);
int b200mp_segment_mean_blend_fwd(const float* x, const float* emb, const int64_t* ptr, const int64_t* perm, int64_t N,
                                  int W, float* out, void* stream);
int b200mp_segment_mean_blend_bwd(const float* g, const int64_t* ptr, const int64_t* index, int64_t N, int64_t R, int W,
                                  float* gx, float* gemb, void* stream);

/* ------------------------------------------------------------------ per-edge linears at H = 128 on the tensor cores
 * `n_hidden = 128` is the reference's PNAS default (src/nn/gnn/pna.py:50) and the only width its notebook timed
 * (benchmark.ipynb:90,100).  The fused H = 64 kernels keep a whole weight image (and, in the backward, two weight-gradient
 * accumulators) resident per CTA; at H = 128 that no longer fits next to the operand tiles, so the same products are composed from
 * the node-level tcgen05 kernels over the E rows:
 * b200mp_row_linear: for tile row k (N rows, K % 32 == 0, NO in {64,128,192,256}; T1 / T2 / mask / R / R2 / Y rows are NO wide)
 *     z = X[xrow[k]] W^T + bias + T1[idx1[k]] + T2[idx2[k]];   relu != 0: z = max(z, 0);   mask: z = z * [mask[k] > 0]
 *     Y[orow[k]] = alpha R[rrow[k]] + beta z + R2[k]
 *   every pointer after Wimg except Y is nullable (row maps NULL = k).  Wimg as b200mp_node_linear.  R may alias Y when
 *   rrow == orow.  Covers PNAConv / GINEConv messages (pna.py:71-74, gine.py:66-70: T1 = B[src], T2 = A[dst]), both halves of the
 *   edge update (pna.py:93, gine.py:90: h = relu(..), e' = alpha e + beta (h W2^T + b2)) and the data products of their backward
 *   (Gh = beta (g W2) * [h > 0];  g_e = alpha g + Gh W1;  g_e += G W).
 * b200mp_message_grad_rows: G[k] = d loss / d m_k from the saved message and the per-node coefficient rows (kind 0: PNA, node_in
 *   [N,7H] of b200mp_pna_coef; kind 1: GINE, node_in [N,H] = gradient at the aggregate), same rules as b200mp_message_bwd;
 *   gT1 (nullable): gT1[nbr[k]] += G[k] with vector REDs.  H % 4 == 0.
 * b200mp_scatter_add_rows: out[idx[k]] += rows[k] (vector REDs), H % 4 == 0.
 * The weight gradients of the composition are b200mp_node_linear_wgrad over the E rows. */
int b200mp_row_linear(const float* X, const int32_t* xrow, int64_t N, int K, int ldx, const float* Wimg, int NO,
                      const float* bias, const float* T1, const int32_t* idx1, const float* T2, const int32_t* idx2,
                      int relu, const float* mask, const float* R, const int32_t* rrow, float alpha, float beta,
                      const float* R2, float* Y, const int32_t* orow, void* stream);
int b200mp_message_grad_rows(int kind, int H, int64_t E, const float* m, const float* node_in, const int32_t* seg,
                             const int32_t* nbr, float* G, float* gT1, void* stream);
int b200mp_scatter_add_rows(const float* rows, const int32_t* idx, int64_t n, int H, float* out, void* stream);

/* ------------------------------------------------------------------ tensor-core self test
 * One-CTA check of the tcgen05 building blocks (csrc/umma.cuh) used by the per-edge kernels:
 *   mode 0: Y[128,64] = X[128,64] W[64,64]^T     mode 1: Y = X W      (3xTF32)
 *   mode 2: Y[64,64]  = X[128,64]^T W[128,64]    (weight-gradient form, 3xTF32)
 *   mode 3: mode 0 with the low-order A operand read from tensor memory
 *   mode 4: mode 0 in a single TF32 pass (no split)
 * Test infrastructure only (tests/test_gpu_umma.py); no reference counterpart. */
int b200mp_umma_selftest(int mode, const float* X, const float* W, float* Y, void* stream);
/* Raw tensor-core probe (scripts/umma_probe.py): host-built shared-memory images (<= 32 KB each)
 * and descriptor fields, one TF32 pass of `ksteps` MMAs; D = the 128 x 64 accumulator lanes. */
int b200mp_umma_probe(const float* a_img, int a_bytes, const float* b_img, int b_bytes,
                      uint64_t adesc, uint64_t bdesc, int a_kstep, int b_kstep, int ksteps,
                      uint32_t idesc, float* D, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200MP_H */

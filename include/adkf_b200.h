/* adkf_b200.h -- C-ABI of the B200-native ADKF-IFT hot path (libadkf_b200.so).
 *
 * The reference (Wenlin-Chen/ADKF-IFT) is pure Python and has NO native interface: these entry points
 * are what a ctypes binding on the reference side would bind to replace the Python/gpytorch/torch_scatter
 * code cited beside each function (paths relative to the reference root).  See INTEGRATION.md for the
 * reference-side ctypes stub.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned, contiguous memory unless marked "host";
 *   - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, no host sync inside;
 *   - return value: 0 = launched, <0 = argument error (ADKF_ERR_*), >0 = cudaError_t of the launch;
 *   - numerical failures (non-PD after the jitter retries, fit not converged) never raise: they are
 *     reported per task in an int32 `status[T]` array (bit flags ADKF_ST_*);
 *   - tasks are batched: T tasks per call, ragged sizes through int32/int64 offset arrays;
 *   - fp32 arithmetic (the reference dtype), fp64 only inside block-wide reductions.
 */
#ifndef ADKF_B200_H
#define ADKF_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADKF_ERR_BADARG   (-1)
#define ADKF_ERR_TOOLARGE (-2)

/* per-task status bit flags */
#define ADKF_ST_OK            0
#define ADKF_ST_JITTER        1   /* Cholesky needed diagonal jitter (psd_safe_cholesky fallback) */
#define ADKF_ST_NOT_PD        2   /* not positive definite after 3 jitter retries                 */
#define ADKF_ST_NOT_CONVERGED 4   /* inner fit hit max_iter before |grad|_inf <= gtol             */
#define ADKF_ST_BADWS        16   /* workspace smaller than adkf_gp_ws_floats() for this batch: nothing was computed   */
#define ADKF_ST_HESS_DAMPED   8   /* inner Hessian singular in adkf_gp_ift: no hypergradient for the task */
/* A task flagged NOT_PD or HESS_DAMPED by adkf_gp_ift gets Wm = 0 (zero gradient) and aux[ADKF_AUX_FOUTER] = NaN: the
 * reference raises in these cases (gpytorch NotPSDError / torch.linalg.solve); the caller drops the task from the meta-batch
 * mean and renormalises (adaptive_dkt_utils.ADKTModelTrainer.train_step). */

/* GP kernel families (fs_mol/utils/gp_utils.py:22-43, :106-134) */
#define ADKF_KERNEL_MATERN52 0   /* pair statistic u = centred squared distance, third parameter = lengthscale */
#define ADKF_KERNEL_RBF      1
#define ADKF_KERNEL_LINEAR   2   /* u = x.x',  K = s v u,       third parameter = variance v (gpytorch LinearKernel)   */
#define ADKF_KERNEL_POLY1    3   /* u = x.x',  K = s (u + c),   third parameter = offset c (PolynomialKernel power 1)  */
#define ADKF_KERNEL_POLY2    4   /* u = x.x',  K = s (u + c)^2                                                        */
#define ADKF_KERNEL_COSSIM   5   /* u = x.x' of L2-normalised features, K = s u (variance fixed to 1), 2 parameters   */
#define ADKF_KERNEL_TANIMOTO 6   /* u = Tanimoto similarity (gp_utils.py:106-117), K = s u, 2 parameters              */
/* OR-ed into the `kernel` argument of adkf_gp_fit / adkf_gp_ift: keep that raw parameter fixed in the inner problem */
#define ADKF_KERNEL_FREEZE_NOISE        0x100   /* ExactTanimotoGP for classification (gp_utils.py:143) */
#define ADKF_KERNEL_FREEZE_OUTPUTSCALE  0x200
#define ADKF_KERNEL_FREEZE_THIRD        0x400
/* pair statistics produced by adkf_gram_pairstat* */
#define ADKF_STAT_SQDIST   0
#define ADKF_STAT_DOT      1
#define ADKF_STAT_TANIMOTO 2

/* layout of the per-task float32 `aux` record written by adkf_gp_fit / adkf_gp_ift (stride ADKF_AUX) */
#define ADKF_AUX            32
#define ADKF_AUX_F           0   /* f = -MLL at phi*                       */
#define ADKF_AUX_G           1   /* d f / d raw phi            [3]          */
#define ADKF_AUX_H           4   /* d2 f / d raw phi2          [9] row-major */
#define ADKF_AUX_L0         13   /* median-heuristic lengthscale init       */
#define ADKF_AUX_LS_LOC     14   /* LogNormal prior loc on the lengthscale  */
#define ADKF_AUX_ITERS      15   /* Newton iterations used                  */
#define ADKF_AUX_FOUTER     16   /* F = -log p(y_q | y_s) (sum, joint)      */
#define ADKF_AUX_DF_DPHI    17   /* d F / d raw phi            [3]          */
#define ADKF_AUX_V          20   /* v = H^-1 dF/dphi           [3]          */
#define ADKF_AUX_JITTER     23   /* jitter finally added to diag(A)         */
#define ADKF_AUX_NFEV       24   /* function evaluations used by the fit    */

int adkf_version(void);
/* kernels launched through this library since it was loaded (bench.py's gpu_launches) */
int64_t adkf_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * Graph indexing (integer path, bit-exact).  Replaces GNN.forward's bidirectional expansion
 * (fs_mol/modules/gnn.py:540-544) + the implicit scatter index of torch_scatter (gnn.py:197-265).
 * Input : adjacency lists as the reference batcher produces them (fs_mol/data/fsmol_batcher.py:197),
 *         one [E_t, 2] int64 array per edge type, each bond stored once.
 * Output: CSR over TARGET nodes of the bidirectional edge set: rowptr[V+1]; for every incoming edge its
 *         source node, edge type and its index in the reference's concatenated message order
 *         (type-major; within a type forward edges then flipped edges) -- sorted by that index inside
 *         a row, so the layout is a deterministic, permutation-stable transform of the input.
 * ------------------------------------------------------------------------------------------- */
int adkf_graph_csr_build(const int64_t* adj0, int64_t e0, const int64_t* adj1, int64_t e1,
                         const int64_t* adj2, int64_t e2, int32_t num_nodes,
                         int32_t* rowptr /*[V+1]*/, int32_t* col_src /*[2E]*/, uint8_t* col_type /*[2E]*/,
                         int32_t* col_tgt /*[2E]*/, int32_t* col_eid /*[2E]*/,
                         int32_t* rev_pos /*[2E] CSR position of the reversed copy of each edge*/,
                         int32_t* degree /*[V]*/, int32_t* scratch /*[2E]*/, void* stream);

/* ---------------------------------------------------------------------------------------------
 * PNA relational message passing, all 4 towers of one GNN block in one launch.
 * Replaces RelationalMP.forward + RelationalMultiAggrMP._aggregate_messages (gnn.py:129-150, 197-265):
 *   m_e   = relu(W[k][type_e] [h_src[k]; h_tgt[k]] + b[k][type_e])                 (192 per tower)
 *   agg_t = [sum | mean | sqrt(sum relu(m^2 - mean^2) + 1e-7) | max] x {1, log(d+1)/1.1515, 1.1515/(log(d+1)+1e-7)}
 * h [V,128] -> out [V, 4*768]; weights W [4][3][192][64], b [4][3][192].
 * aux_cnt / aux_arg [V,4,64] uint8: per channel the number of edges with m^2 > mean^2 and the arg-max edge
 * (index inside the CSR row), kept for the backward.
 * ------------------------------------------------------------------------------------------- */
int adkf_mp_pna_fwd(const float* h, const float* W, const float* b, const int32_t* rowptr,
                    const int32_t* col_src, const uint8_t* col_type, int32_t num_nodes,
                    float* out, uint8_t* aux_cnt, uint8_t* aux_arg, void* stream);
/* backward: d_out [V,3072] -> d_h [V,128], d_W, d_b (deterministic: no float atomics).
 * scratch: d_pre [2E,4,192], d_xs [2E,128], d_xt [V,128], partials [num_partials, 12, 192*64+192]. */
int adkf_mp_pna_bwd(const float* h, const float* W, const float* b, const int32_t* rowptr,
                    const int32_t* col_src, const int32_t* col_tgt, const uint8_t* col_type,
                    const int32_t* rev_pos, int32_t num_nodes, int32_t num_edges, const float* fwd_out,
                    const uint8_t* aux_cnt, const uint8_t* aux_arg, const float* d_out, float* d_pre,
                    float* d_xs, float* d_xt, float* d_h, float* d_W, float* d_b, float* partials,
                    int32_t num_partials, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Elementwise half of a dense layer's backward (autograd of nn.Linear + ReLU / LeakyReLU(0.01):
 * fs_mol/modules/gnn.py:79-95, fs_mol/modules/mlp.py:6-22) in one pass over dY [M,N]:
 *   dZ = dY * act'(Y)   (act 1 relu, 2 leaky_relu; Y = the saved activation OUTPUT; act 0: no dZ is written)
 *   db[c] = sum_r dZ[r,c]   (db may be NULL; deterministic two-pass column sum, ws >= adkf_act_bwd_colsum_ws_floats)
 * N % 4 == 0, 16-byte aligned rows.
 * ------------------------------------------------------------------------------------------- */
int64_t adkf_act_bwd_colsum_ws_floats(int64_t M, int32_t N);
int adkf_act_bwd_colsum(const float* dY, int64_t ldy, const float* Y, int64_t ldyy, int64_t M, int32_t N, int32_t act,
                        float* dZ, int64_t ldz, float* db, float* ws, int64_t ws_floats, void* stream);
/* adkf_act_bwd_colsum for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * NO COUNTERPART IN THE REFERENCE.  Same kernels and the same result bit for bit;
 * the column-sum partials live in a file-scope `static __device__` array of adkf_act_bwd_colsum_noalloc_max_floats() floats
 * (1184 x 1024 = 4.6 MiB, held by every GPU that loads the library until exit) instead of a caller workspace; the path calls no
 * cudaMalloc / cudaFree / allocating routine.  Returns ADKF_ERR_TOOLARGE (nothing launched) when
 * adkf_act_bwd_colsum_ws_floats(M, N) exceeds that maximum.  All calls share the array: calls that overlap on the GPU (streams
 * without an ordering between them) race; calls on one stream are safe.  Nothing needs clearing between calls (every partial
 * that is read was written by the same call). */
int64_t adkf_act_bwd_colsum_noalloc_max_floats(void);
int adkf_act_bwd_colsum_noalloc(const float* dY, int64_t ldy, const float* Y, int64_t ldyy, int64_t M, int32_t N, int32_t act,
                                float* dZ, int64_t ldz, float* db, void* stream);

/* ---------------------------------------------------------------------------------------------
 * ReZero residual of a GNN block (fs_mol/modules/gnn.py:497-498, 512-513): out = h + alpha * x with a learned scalar alpha read
 * from the device; backward in one pass over g: dx = alpha * g (dx may be NULL) and dalpha = <g, x> (dalpha may be NULL;
 * deterministic, fp64 partials in ws[adkf_rezero_bwd_ws_doubles()]).  n % 4 == 0, 16-byte aligned arrays.
 * ------------------------------------------------------------------------------------------- */
int64_t adkf_rezero_bwd_ws_doubles(void);
int adkf_rezero_fwd(const float* h, const float* x, const float* alpha, int64_t n, float* out, void* stream);
int adkf_rezero_bwd(const float* g, const float* x, const float* alpha, int64_t n, float* dx, float* dalpha, double* ws, void* stream);

/* ---------------------------------------------------------------------------------------------
 * PNA degree scalers folded out of the aggregate (gnn.py:242-263 + msg_out_projection, gnn.py:492-495; pna_combine.cu):
 * out[v] = P[v,0:128] + amp(v) P[v,128:256] + att(v) P[v,256:384] + bias, where P = agg_unscaled [V,1024] @ [W_1;W_amp;W_att]^T.
 *   adkf_pna_degree_scalers: amp = log(deg+1)/1.1515, att = 1.1515/(log(deg+1)+1e-7) from the type-sorted tptr [3V+1]
 *   adkf_pna_combine_bwd   : dP [V,384] = [g | amp g | att g]; dbias [128] (may be NULL; ws >= adkf_pna_combine_bwd_ws_floats(V))
 * ------------------------------------------------------------------------------------------- */
int adkf_pna_degree_scalers(const int32_t* tptr, int32_t V, float* amp, float* att, void* stream);
int adkf_pna_combine_fwd(const float* P, const float* amp, const float* att, const float* bias, int64_t V, float* out,
                         void* stream);
int64_t adkf_pna_combine_bwd_ws_floats(int64_t V);
int adkf_pna_combine_bwd(const float* g, const float* amp, const float* att, int64_t V, float* dP, float* dbias, float* ws,
                         int64_t ws_floats, void* stream);

/* ---------------------------------------------------------------------------------------------
 * nn.LayerNorm(128) of the GNN block (`boom_norm_layer`, fs_mol/modules/gnn.py:462, applied at :504; torch semantics:
 * biased variance, eps inside the square root) on x [M,128] fp32, 16-byte aligned: one warp per row.
 *   fwd: y = (x - mean) * rstd * gamma + beta; mean / rstd [M] are kept for the backward pass
 *   bwd: dx [M,128], dgamma [128], dbeta [128] (deterministic two-pass column sums; ws >= adkf_layernorm128_bwd_ws_floats(M))
 * ------------------------------------------------------------------------------------------- */
int adkf_layernorm128_fwd(const float* x, int64_t M, const float* gamma, const float* beta, float eps, float* y,
                          float* mean, float* rstd, void* stream);
int64_t adkf_layernorm128_bwd_ws_floats(int64_t M);
int adkf_layernorm128_bwd(const float* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                          int64_t M, float* dx, float* dgamma, float* dbeta, float* ws, int64_t ws_floats, void* stream);

/* ---------------------------------------------------------------------------------------------
 * The same PNA block with its three dense products on the tensor cores (pna_gemm.cu; default path of the extractor):
 *   XG [E,256]   = adkf_pna_gather_inputs(h)            [h_src | h_tgt] per tower, edges sorted by (type, target, message index)
 *   pre [E,768]  = adkf_gemm_tf32x3_batched(XG, W, b)   message MLP pre-activations, 4 towers per launch, one launch per type
 *   out [V,3072] = adkf_pna_aggregate_fwd(pre)          relu + sum | mean | pseudo-std | max x 3 degree scalers
 * backward: d_pre = adkf_pna_dpre(d out); dXG, dW through the batched GEMM; d_h = adkf_pna_dh_gather(dXG).
 * tptr [3V+1]: tptr[type*V + v] .. tptr[type*V + v + 1] = sorted positions of the type-`type` edges arriving at v;
 * rev [E]: sorted position of the reversed copy of every edge.
 * ------------------------------------------------------------------------------------------- */
int adkf_pna_gather_inputs(const float* h, const int32_t* src, const int32_t* tgt, int64_t E, float* XG, void* stream);
int adkf_pna_aggregate_fwd(const float* pre, const int32_t* tptr, int32_t V, float* out, uint8_t* aux_cnt, uint8_t* aux_arg,
                           void* stream);
int adkf_pna_dpre(const float* pre, const int32_t* tptr, int32_t V, const float* fwd_out, const uint8_t* aux_cnt,
                  const uint8_t* aux_arg, const float* d_out, float* d_pre, void* stream);
int adkf_pna_dh_gather(const float* dXG, const int32_t* tptr, const int32_t* rev, int32_t V, float* d_h, void* stream);
/* ---------------------------------------------------------------------------------------------
 * The tower-style message passing of a PNA / MultiAggr block as ONE kernel per direction (pna_fused.cu; default when the
 * largest in-degree is <= adkf_pna_fused_max_degree()): gather [h_src | h_tgt] -> message MLP on tcgen05 (3xTF32) -> bias + ReLU ->
 * sum | mean | pseudo-std | max per target node.  Replaces RelationalMP.forward + RelationalMultiAggrMP._aggregate_messages
 * (fs_mol/modules/gnn.py:129-150, 197-240) without writing the gathered inputs [E,256] or the pre-activations [E,768] to HBM.
 *   fwd : h [V,128], W [4][3][192][64], b [4][3][192] -> out [V,1024] (unscaled aggregate, ADKF_MP_MULTIAGGR layout),
 *         aux_cnt / aux_arg [V,4,64]; bit-identical to gather + adkf_gemm_tf32x3_batched + adkf_mp_aggregate_fwd
 *   dpre: recomputes the pre-activations and emits d pre [E,768] in (type, target)-sorted row order (csr2ts [E] maps a CSR
 *         row to its sorted position), ready for the weight / input gradient GEMMs
 * CSR arrays as built by adkf_graph_csr_build; tile_node0 [adkf_pna_fused_num_tiles(E) + 1]: first node v with
 * rowptr[v] >= 112 j (last entry V).
 * ------------------------------------------------------------------------------------------- */
int32_t adkf_pna_fused_num_tiles(int64_t E);
int32_t adkf_pna_fused_max_degree(void);
/* development probe: per-CTA cycle counters [148][16] of the next fused launches (waits / work per warp role); NULL = off */
void adkf_pna_fused_set_debug(void* buf);
int adkf_pna_fused_fwd(const float* h, const float* W, const float* b, const int32_t* rowptr, const int32_t* col_src,
                       const int32_t* col_tgt, const uint8_t* col_type, const int32_t* tile_node0, int32_t V, int64_t E,
                       float* out, uint8_t* aux_cnt, uint8_t* aux_arg, void* stream);
int adkf_pna_fused_dpre(const float* h, const float* W, const float* b, const int32_t* rowptr, const int32_t* col_src,
                        const int32_t* col_tgt, const uint8_t* col_type, const int32_t* tile_node0, const int32_t* csr2ts,
                        int32_t V, int64_t E, const float* fwd_out, const uint8_t* aux_cnt, const uint8_t* aux_arg,
                        const float* d_out, float* d_pre, void* stream);
/* The other two tower-style message-passing types of GNNBlock (fs_mol/modules/gnn.py:421-447), same edge layout:
 *   ADKF_MP_MULTIAGGR  RelationalMultiAggrMP without PNA scalers: pre [E,768] -> out [V,1024]  (gnn.py:197-240)
 *   ADKF_MP_PLAIN      RelationalMP: relu + scatter_sum of 64-wide messages, pre [E,256] -> out [V,256]  (gnn.py:152-165;
 *                      fwd_out / aux arrays unused, may be NULL)
 * ADKF_MP_PNA is what adkf_pna_aggregate_fwd / adkf_pna_dpre run. */
#define ADKF_MP_PNA 0
#define ADKF_MP_MULTIAGGR 1
#define ADKF_MP_PLAIN 2
int adkf_mp_aggregate_fwd(const float* pre, const int32_t* tptr, int32_t V, int32_t mode, float* out, uint8_t* aux_cnt,
                          uint8_t* aux_arg, void* stream);
int adkf_mp_dpre(const float* pre, const int32_t* tptr, int32_t V, int32_t mode, const float* fwd_out, const uint8_t* aux_cnt,
                 const uint8_t* aux_arg, const float* d_out, float* d_pre, void* stream);

/* ---------------------------------------------------------------------------------------------
 * `--gnn_type MultiHeadAttention`: RelationalMultiHeadAttentionMP (fs_mol/modules/gnn.py:268-387), hidden 128, 4 heads x 64,
 * 3 edge types, on the same type-sorted edge arrays (src / tgt / rev [E], tptr [3V+1]) as the tower kernels.
 *   XG [E,256]  = adkf_mha_gather_inputs(x)        [x_src | x_tgt]  (message MLP input; the MLP itself is adkf_gemm_tf32x3 per type)
 *   QK [V,1536] = x [Wq_0 Wq_1 Wq_2 Wk_0 Wk_1 Wk_2]^T  (adkf_gemm_tf32x3; target-query / neighbour-key projections of every node)
 *   out [V,256] = adkf_mha_attention_fwd(pre [E,256], QK): per (target, head) softmax over ALL arriving edges of
 *                 64^-1/2 <Q_type[tgt], K_type[src]>, weighted sum of relu(pre); probs [E,4] kept for the backward
 *   adkf_mha_attention_bwd: d_pre [E,256], dQK [V,1536] (ds_ws [E,4] scratch); adkf_mha_dx_gather: d_x from dXG.
 * ------------------------------------------------------------------------------------------- */
int adkf_mha_gather_inputs(const float* x, const int32_t* src, const int32_t* tgt, int64_t E, float* XG, void* stream);
int adkf_mha_dx_gather(const float* dXG, const int32_t* tptr, const int32_t* rev, int32_t V, float* d_x, void* stream);
int adkf_mha_attention_fwd(const float* pre, const float* QK, const int32_t* src, const int32_t* tptr, int32_t V, float* out,
                           float* probs, void* stream);
int adkf_mha_attention_bwd(const float* pre, const float* QK, const int32_t* src, const int32_t* rev, const int32_t* tptr,
                           int32_t V, const float* probs, const float* d_out, float* d_pre, float* ds_ws, float* dQK,
                           void* stream);

/* ---------------------------------------------------------------------------------------------
 * Graph readout segment kernels (fs_mol/modules/graph_readout.py:224-259, 286-296).  Nodes of a graph
 * are contiguous: graph_ptr[G+1].
 *   weighted pool : w = softmax over the graph's nodes (mode 0) or sigmoid (mode 1) of scores [V,heads];
 *                   out[g, h*hd + c] = sum_v w[v,h] * values[v, h*hd + c]
 *   max pool      : out[g, c] = max_v x[v, c]  (empty graph -> 0), argmax kept for the backward
 * ------------------------------------------------------------------------------------------- */
int adkf_readout_weighted_fwd(const float* scores, const float* values, const int32_t* graph_ptr,
                              int32_t num_graphs, int32_t heads, int32_t head_dim, int32_t mode,
                              float* weights /*[V,heads]*/, float* out /*[G,heads*head_dim]*/, void* stream);
int adkf_readout_weighted_bwd(const float* weights, const float* values, const int32_t* graph_ptr,
                              int32_t num_graphs, int32_t heads, int32_t head_dim, int32_t mode,
                              const float* d_out, float* d_scores, float* d_values, void* stream);
int adkf_readout_max_fwd(const float* x, const int32_t* graph_ptr, int32_t num_graphs, int32_t dim,
                         float* out, int32_t* argmax, void* stream);
int adkf_readout_max_bwd(const float* d_out, const int32_t* argmax, const int32_t* graph_ptr,
                         int32_t num_graphs, int32_t dim, float* d_x, void* stream);

/* ---------------------------------------------------------------------------------------------
 * ECFP branch of the first FC layer (fs_mol/models/adaptive_dkt.py:145-146,158-160): the fingerprint
 * arrives as int32 counts with ~2% non-zeros, so  fc0(cat(gnn, ecfp.float()))  is computed as
 *   out[g,:] += sum_{j : fp[g,j] != 0} fp[g,j] * Wt[j,:]       (Wt = fc.0.weight[:, 512:].T, [2048,2048])
 * instead of a dense [G,2048]x[2048,2048] GEMM.  `out` must already hold the dense gnn part + bias.
 * ------------------------------------------------------------------------------------------- */
int adkf_ecfp_linear_fwd(const int32_t* fp, int32_t num_graphs, int32_t fp_dim, const float* Wt,
                         int32_t out_dim, float* out, void* stream);
/* d_Wt[j,:] += sum_g fp[g,j] * d_out[g,:]  (deterministic: one CTA owns a weight row) */
int adkf_ecfp_linear_bwd(const int32_t* fp, int32_t num_graphs, int32_t fp_dim, const float* d_out,
                         int32_t out_dim, float* d_Wt, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Batched GP tail.  Task t owns rows [z_off[t], z_off[t+1]) of Z [rows, d] (support rows first:
 * n_sup[t] of them, then the query rows) and the N_t x N_t matrix at D2 + d2_off[t] (ld = N_t).
 * ------------------------------------------------------------------------------------------- */

/* Joint squared-distance matrix per task from the centred quadratic expansion (gpytorch sq_dist; call
 * sites fs_mol/utils/gp_utils.py:28,31).  qq_block = 0 skips the query x query block (meta-test needs
 * only diag(K_qq) = outputscale). */
int adkf_gram_sqdist(const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int32_t* n_sup,
                     const int64_t* d2_off, int32_t max_n /*max rows of any task (grid sizing)*/,
                     int32_t qq_block, float* mu_ws /*[T,d]*/, float* nrm_ws /*[rows]*/, float* D2, void* stream);

/* Same result as adkf_gram_sqdist, computed on the 5th-generation tensor cores: tcgen05.mma kind::tf32 with the 3xTF32
 * error-compensated split (fp32-class accuracy), TMA-staged operands, accumulator in TMEM (gemm_tf32x3.cu).
 * Needs d % 4 == 0 and a 16-byte aligned Z; total_rows = number of rows of Z (bounds of the TMA tensor map). */
int adkf_gram_sqdist_tc(const float* Z, int32_t d, int32_t T, int64_t total_rows, const int32_t* z_off,
                        const int32_t* n_sup, const int64_t* d2_off, int32_t max_n, int32_t qq_block,
                        float* mu_ws /*[T,d]*/, float* nrm_ws /*[rows]*/, float* D2, void* stream);

/* Generalisation of the two entry points above to the other kernel families' pair statistics (ADKF_STAT_*):
 * dot products for linear / polynomial / cossim (gpytorch LinearKernel / PolynomialKernel behind
 * fs_mol/utils/gp_utils.py:22-43) and the Tanimoto similarity (batch_tanimoto_sim, gp_utils.py:106-117).  For these the
 * inputs are NOT centred; the diagonal (|z|^2 resp. 1) is always written because the predictive variance needs k(x*,x*).
 * The backward exists for SQDIST and DOT (Tanimoto inputs are fixed fingerprints). */
int adkf_gram_pairstat(int32_t stat, const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int32_t* n_sup,
                       const int64_t* d2_off, int32_t max_n, int32_t qq_block, float* mu_ws, float* nrm_ws, float* D2,
                       void* stream);
int adkf_gram_pairstat_tc(int32_t stat, const float* Z, int32_t d, int32_t T, int64_t total_rows, const int32_t* z_off,
                          const int32_t* n_sup, const int64_t* d2_off, int32_t max_n, int32_t qq_block, float* mu_ws,
                          float* nrm_ws, float* D2, void* stream);
int adkf_gram_pairstat_bwd(int32_t stat, const float* Z, int32_t d, int32_t T, const int32_t* z_off,
                           const int64_t* d2_off, int32_t max_n, const float* Wm, const float* mu_ws, float scale,
                           float* dZ, void* stream);

/* Dense layer forward on the same tensor-core kernel:  Y[M,N] = act(X[M,K] W[N,K]^T + bias[N])  (torch.nn.Linear of the
 * extractor: fs_mol/modules/gnn.py:79-95,477-515; graph_readout.py:224-259; fs_mol/models/adaptive_dkt.py:61-65).
 * act: 0 none, 1 relu, 2 leaky_relu(0.01).  ld* in floats (multiples of 4), pointers 16-byte aligned, bias may be NULL. */
int adkf_linear_tf32x3(const float* X, int64_t M, int32_t K, int64_t ldx, const float* W, int32_t N, int64_t ldw,
                       const float* bias, int32_t act, float* Y, int64_t ldy, void* stream);

/* General form of the same kernel, used by the backward of every dense layer (what autograd's mm backward does in the
 * reference):   D[M,N] = act( opA(A)[M,K] . opB(B)[N,K]^T + bias[N] )
 *   transA = 0: A stored [M,K] (lda floats per row);  transA = 1: A stored [K,M]  (contraction index is the slow one)
 *   transB = 0: B stored [N,K];                       transB = 1: B stored [K,N]
 * Supported combinations: (0,0) forward y = x W^T;  (0,1) input gradient dX = dY W;  (1,1) weight gradient dW = dY^T X.
 * When the output has fewer than 148 tiles (weight gradients: K = number of nodes) and a workspace of
 * adkf_gemm_splitk_ws_floats(M,N,K) floats is passed (bias NULL, act 0, ldd == N), the contraction is split over CTAs and
 * reduced by a second, deterministic pass.  act = 3 (ADKF_EPI_SUB_FROM_D): D <- D - opA(A) opB(B)^T, the trailing update
 * of the blocked Cholesky below. */
#define ADKF_EPI_SUB_FROM_D 3
int adkf_gemm_tf32x3(const float* A, int32_t transA, int64_t lda, const float* B, int32_t transB, int64_t ldb,
                     int64_t M, int32_t N, int32_t K, const float* bias, int32_t act, float* D, int64_t ldd,
                     float* splitk_ws, int64_t splitk_ws_floats, void* stream);

/* D[M,N] = (opA(A) . opB(B)^T) * act'(mask[M,N]): the input-gradient GEMM of a dense layer fused with the activation derivative
 * of the layer below (autograd of Linear -> ReLU / LeakyReLU(0.01) -> Linear: BOOMLayer gnn.py:79-95, MLP mlp.py:6-22).
 * mask = the saved activation OUTPUT of the layer below (ldm floats per row), mask_act 1 relu / 2 leaky_relu. */
int adkf_gemm_tf32x3_masked(const float* A, int32_t transA, int64_t lda, const float* B, int32_t transB, int64_t ldb,
                            int64_t M, int32_t N, int32_t K, const float* mask, int64_t ldm, int32_t mask_act, float* D,
                            int64_t ldd, void* stream);
int64_t adkf_gemm_splitk_ws_floats(int64_t M, int32_t N, int32_t K);

/* ---- --use-ard: one lengthscale per feature (ExactGPLayer(..., ard_num_dims=d), fs_mol/utils/gp_utils.py:14,27,30) -------------
 * The kernel acts on u_ab = sum_k w_k (z_ak - z_bk)^2, w_k = 1 / l_k^2; the n x n work runs in adkf_gp_fit / _predict / _ift on u
 * with the scalar lengthscale fixed to 1.  Task layout as in the batched GP tail (z_off, n_sup, d2_off; leading dimension N_t).
 *   adkf_ard_wsqdist : D2[t][a][b] = sum_k c[t][k] (z_ak - z_bk)^2 (c of either sign); support_only != 0: rows/cols < n_sup[t] only
 *   adkf_ard_pairquad: q[t][k] = sum_ab G[t][a][b] (z_ak - z_bk)^2 over the same index range
 * adkf_gp_ift mode 4 (aux[ADKF_AUX_V..+1] = direction of the raw noise / outputscale, support block of Wm = direction dD2 on
 * entry) returns Wm_ss = d/deps df/dD2 and aux[ADKF_AUX_DF_DPHI..+1] = d/deps of the raw (noise, outputscale) gradient: with
 * the two contractions above that is one Hessian-vector product of the (d+2)-dimensional inner problem, no Hessian formed. */
int adkf_ard_wsqdist(const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                     int32_t max_n, const float* c, int32_t support_only, float* D2, void* stream);
int adkf_ard_pairquad(const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                      const float* G, int32_t support_only, float* q, void* stream);

/* Pre-split weight images for the forward / input-gradient GEMMs (gemm_pair.cu, PRE variant).  The B operand of those products is
 * a weight matrix (nn.Linear: y = x W^T, fs_mol/modules/gnn.py:79-95, graph_readout.py:224-259; its autograd dX = dY W) that every
 * one of the ~1000 M-tiles would split into (hi, lo) again.  adkf_gemm_presplit writes, once per optimizer step, for W [rows, K]
 * (K % 32 == 0, K-major): hi [rows, K] = tf32-rounded values and co [rows, K] 32-bit words = per 32-wide k-block the 16-bit pairs of
 * hi then of lo (corr 1: bf16; corr 2: fp16 with lo x 2^11 -- see adkf_gemm_configure).  adkf_gemm_tf32x3_pre computes
 *   D [M,N] = act(A [M,K] . W^T + bias)       or, with mask != NULL,   D = (A . W^T) * act'(mask)   (mask_act 1 relu | 2 leaky_relu)
 * from the images of W [N,K] (for dX pass the images of the TRANSPOSED weight): TMA loads the image tiles straight into the MMA's
 * B stages, the B-splitter warps split half of the A rows instead.  M > 128; `corr` must be the flavour the images were built with. */
int adkf_gemm_presplit(const float* W, int64_t rows, int32_t K, int64_t ldw, int32_t corr, float* hi, float* co, void* stream);
int adkf_gemm_tf32x3_pre(const float* A, int64_t lda, const float* B_hi, const float* B_co, int32_t corr, int64_t M, int32_t N, int32_t K,
                         const float* bias, int32_t act, const float* mask, int64_t ldm, int32_t mask_act, float* D, int64_t ldd,
                         void* stream);

/* Process-wide GEMM variant (development / A-B measurements; initialised from ADKF_GEMM_IMPL / ADKF_GEMM_CORR):
 *   impl   1 = one CTA per 128 x 128 tile (gemm_tf32x3.cu), 2 = CTA pairs on 256 x 128 tiles (gemm_pair.cu, default); other: keep
 *   corr16 correction products of the pair kernel: 0 = tf32, 1 = bf16, 2 = tf32 forward / bf16 for the backward operand forms
 *          (K,T) / (T,T), 3 = scaled fp16 forward / bf16 backward (default), 4 = scaled fp16 everywhere; other: keep.
 *          Scaled fp16: lo parts x 2^11 as fp16, accumulator x 2^-11 -- as exact as tf32 corrections for operand magnitudes in
 *          [6.1e-5, 65504]; elements outside that range lose their correction (TF32-class accuracy for them).
 * Returns the previous setting as impl | corr16 << 8.  Not thread-safe against concurrent launches.
 * NOTE: adkf_gemm_splitk_ws_floats depends on impl -- size workspaces after configuring. */
int adkf_gemm_configure(int32_t impl, int32_t corr16);
/* Batched form: `batch` products in one launch whose operands are column / row blocks of the same 2-D arrays (one TMA tensor
 * map per operand): batch b uses opA at (row + b*a_mn_stride, k + b*a_k_stride), opB likewise, writes D + b*d_stride and adds
 * bias + b*bias_stride.  a_slow / a_fast (b_slow / b_fast) are the extents of the whole underlying arrays (slow = rows).
 * Used by the PNA message passing: 4 towers x {message MLP, its input gradient, its weight gradient} per edge type
 * (fs_mol/modules/gnn.py:129-150).  K must be a multiple of 32 when a k stride is used. */
int adkf_gemm_tf32x3_batched(const float* A, int32_t transA, int64_t lda, int64_t a_slow, int64_t a_fast, int32_t a_mn_stride,
                             int32_t a_k_stride, const float* B, int32_t transB, int64_t ldb, int64_t b_slow, int64_t b_fast,
                             int32_t b_mn_stride, int32_t b_k_stride, int64_t M, int32_t N, int32_t K, int32_t batch,
                             const float* bias, int64_t bias_stride, int32_t act, float* D, int64_t ldd, int64_t d_stride,
                             float* splitk_ws, int64_t splitk_ws_floats, void* stream);
int64_t adkf_gemm_batched_splitk_ws_floats(int64_t M, int32_t N, int32_t K, int32_t batch);

/* dZ[a,:] = scale * 4 * sum_b Wm[a,b] (z_a - z_b)  for a symmetric per-task weight matrix Wm (same layout
 * as D2): the chain rule of anything that depends on Z only through D2. */
int adkf_gram_sqdist_bwd(const float* Z, int32_t d, int32_t T, const int32_t* z_off, const int64_t* d2_off,
                         int32_t max_n, const float* Wm, const float* mu_ws, float scale, float* dZ, void* stream);

/* Per-task inner fit of (raw_noise, raw_outputscale, raw_lengthscale) by minimising f = -MLL (with the
 * LogNormal priors) on the support set.  Replaces reinit_gp_params + compute_median_lengthscale_init
 * (fs_mol/models/adaptive_dkt.py:88-131) and botorch fit_gpytorch_scipy / SciPy L-BFGS-B
 * (fs_mol/utils/adaptive_dkt_utils.py:91) with an on-device damped Newton iteration using the analytic
 * 3x3 Hessian.  max_iter = 0 evaluates f, grad, Hessian at the given phi (fixed-phi mode, phi_io is input).
 * init_mode: 1 = reference initialisation (noise_init, outputscale softplus(0), median lengthscale),
 *            0 = start from phi_io.  ls_prior: 1 = LogNormal(lengthscale) prior with loc log(l0)+0.0625,
 *            2 = use the loc already stored in aux[ADKF_AUX_LS_LOC], 0 = no lengthscale prior. */
int adkf_gp_fit(int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                const float* D2, const float* y, int32_t kernel, float noise_init, float noise_prior_loc,
                int32_t ls_prior, int32_t init_mode, int32_t max_iter, float gtol,
                float* phi_io /*[T,3]*/, float* aux /*[T,ADKF_AUX]*/, int32_t* status /*[T]*/,
                float* ws, int64_t ws_floats, void* stream);

/* Posterior predictive of the query rows: mean = K_qs A^-1 y, variance = diag(K_qq - K_qs A^-1 K_sq) + noise
 * (gpytorch ExactGP eval + GaussianLikelihood; call sites fs_mol/models/adaptive_dkt.py:198-203,
 * bayes_opt/bo_utils.py:367-373).  mean/var are aligned with the Z rows (support rows are left untouched). */
int adkf_gp_predict(int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                    const float* D2, const float* y, int32_t kernel, const float* phi,
                    float* mean, float* var, int32_t* status, float* ws, int64_t ws_floats, void* stream);

/* IFT hypergradient pieces for a meta-training task (fs_mol/utils/cauchy_hypergradient.py:5-164 with
 * f_inner / f_outer of fs_mol/utils/adaptive_dkt_utils.py:383-395):
 *   F = -log N(y_q | posterior(y_s) + noise)  (summed joint log-prob, adaptive_dkt.py:183-191),
 *   H = d2f/dphi2 (3x3), v = H^-1 dF/dphi, and the symmetric weight matrix
 *   Wm = dF/dD2 - d/deps [ df/dD2 (phi + eps v) ]   (SURVEY App. B), to be fed to adkf_gram_sqdist_bwd.
 * mode: 0 = full hypergradient; 1 = drop the second term (the reference's ignore_grad_correction flag);
 *       2 = inner loss only: Wm = df/dD2 on the support block (first-order gradient of f = -MLL w.r.t. Z_s,
 *           the functional train-loss branch of adaptive_dkt.py:173-176), aux F/G/H refreshed. */
int adkf_gp_ift(int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                const float* D2, const float* y, int32_t kernel, float noise_prior_loc, int32_t ls_prior,
                int32_t mode, const float* phi, float* aux /*[T,ADKF_AUX] in/out*/,
                float* Wm, int32_t* status, float* ws, int64_t ws_floats, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Large-pool path (bayes_opt / predictive_performance: fit on <= 64 molecules, evaluate on ~2.25k):
 * joint predictive NLL of ONE task's query rows through a blocked Cholesky of the (support + query) covariance, trailing
 * updates on the tensor-core GEMM (chol_blocked.cu).  Replaces likelihood(model(x_test)).log_prob(y_test)
 * (bayes_opt/predictive_performance/adkt.py:93-96; gpytorch would switch to CG/Lanczos above 800 rows).
 *   U [N,N] pair statistics (adkf_gram_pairstat* with qq_block = 1), support rows first; phi [3] raw parameters (device).
 *   out4 = { NLL of all N rows, NLL of the n_sup support rows, their difference F = -log p(y_q | y_s), log det J }.
 *   status[0] gets ADKF_ST_NOT_PD if a pivot fails (retry with a larger `jitter`, psd_safe_cholesky policy, on the host).
 * ------------------------------------------------------------------------------------------- */
int64_t adkf_chol_blocked_ws_floats(int32_t N);
int adkf_gp_joint_nll_blocked(const float* U, int32_t N, int32_t n_sup, const float* y, int32_t kernel, const float* phi,
                              float jitter, float* ws, int64_t ws_floats, float* out4, int32_t* status, void* stream);
/* The same with gpytorch's psd_safe_cholesky jitter ladder (0, 1e-6, 1e-5, 1e-4) run ON THE DEVICE: the jittered attempts are
 * enqueued behind the first and skip themselves unless the previous attempt flagged ADKF_ST_NOT_PD; no host synchronisation.
 * status[0]: 0, ADKF_ST_JITTER (a jittered attempt succeeded) or ADKF_ST_NOT_PD. */
int adkf_gp_joint_nll_blocked_ladder(const float* U, int32_t N, int32_t n_sup, const float* y, int32_t kernel, const float* phi,
                                     float* ws, int64_t ws_floats, float* out4, int32_t* status, void* stream);

/* Analytic expected improvement for MINIMISATION over a candidate pool, on the latent posterior (variance minus noise):
 * botorch ExpectedImprovement(model, best_f, maximize=False) evaluated per molecule in bayes_opt/bo_utils.py:368-378.
 * exclude[i] != 0 (already queried) -> -inf. */
int adkf_ei_pool(const float* mean, const float* var, int32_t m, float noise, float best_f, const uint8_t* exclude,
                 float* ei, void* stream);

/* ---------------------------------------------------------------------------------------------
 * MoleculeNet branch (config 4): GIN molecule encoder, MoleculeNet/chem_lib/models/encoder.py (gin.cu).
 * Graphs arrive in the PyG layout of the reference: x [V,2] int64 (atom type, chirality), edge_index [2,E] int64
 * (row 0 source j, row 1 target i; both directions present), edge_attr [E,2] int64 (bond type, direction).
 * ------------------------------------------------------------------------------------------- */
/* CSR of edge_index by target (by_source = 0) or by source (1): rowptr [V+1], col_other [E] = the other endpoint,
 * col_eid [E] = edge id; rows sorted by edge id (deterministic).  scratch: V int32. */
int adkf_csr_from_edge_index(const int64_t* edge_index, int64_t E, int32_t V, int32_t by_source, int32_t* rowptr,
                             int32_t* col_other, int32_t* col_eid, int32_t* scratch, void* stream);
/* out[r] = T1[idx[r,0]] + T2[idx[r,1]]  (x_embedding1 + x_embedding2, encoder.py:269) */
int adkf_embed_sum2(const int64_t* idx, int64_t rows, int32_t D, const float* T1, const float* T2, float* out, void* stream);
/* GINConv propagate (encoder.py:40-58): out[i] = x[i] + E1[4] + E2[0] + sum_{p in row i} (x[col[p]] + E1[attr[eid[p],0]] +
 * E2[attr[eid[p],1]]).  With edge_attr == NULL the edge terms are dropped: out[i] = x[i] + sum x[col[p]] -- the transpose
 * used by the backward on the by-source CSR. */
int adkf_gin_gather(const float* x, int32_t V, int32_t D, const int32_t* rowptr, const int32_t* col, const int32_t* col_eid,
                    const int64_t* edge_attr, const float* E1, const float* E2, float* out, void* stream);
/* BatchNorm1d (+ReLU when act = 1, + inverted dropout when keep < 1) with the statistics taken PER SEGMENT of rows
 * seg_ptr[S+1]: one segment = one encoder call of the reference (support or query molecules of one task), so that merging
 * T tasks into one launch leaves the train-mode statistics of encoder.py:273-281 unchanged.  mode 0 = batch statistics
 * (saved: mean, rstd, unbiased variance, each [S,D]); mode 1 = running statistics. */
int adkf_seg_bn_fwd(const float* x, int32_t D, const int32_t* seg_ptr, int32_t S, const float* gamma, const float* beta,
                    const float* run_mean, const float* run_var, float eps, int32_t mode, int32_t act, float keep,
                    uint64_t seed, float* y, float* save_mean, float* save_rstd, float* save_uvar, void* stream);
int adkf_seg_bn_bwd(const float* x, const float* dy, int32_t D, const int32_t* seg_ptr, int32_t S, const float* gamma,
                    const float* beta, const float* save_mean, const float* save_rstd, int32_t act, float keep, uint64_t seed,
                    float* dx, float* dgamma_part /*[S,D]*/, float* dbeta_part /*[S,D]*/, void* stream);
/* global_mean_pool (encoder.py:325): nodes of a graph are contiguous, graph_ptr[G+1] */
int adkf_seg_mean_fwd(const float* x, int32_t D, const int32_t* graph_ptr, int32_t G, float* out, void* stream);
int adkf_seg_mean_bwd(const float* dout, int32_t D, const int32_t* graph_ptr, int32_t G, float* dx, void* stream);

/* workspace sizes (floats) for the three GP entry points, host-side helper */
int64_t adkf_gp_ws_floats(int32_t T, const int32_t* z_off_host, const int32_t* n_sup_host, int32_t which);

#ifdef __cplusplus
}
#endif
#endif /* ADKF_B200_H */

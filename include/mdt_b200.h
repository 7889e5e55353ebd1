/*
 * mdt_b200.h -- C ABI of libmdt_b200.so, the sm_100a (B200) implementation of mDT's
 * graph-transformer hot path.
 *
 * Conventions:
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - the caller owns every buffer (inputs, outputs, workspaces); the library keeps no
 *     device memory of its own;
 *   - `stream` is a cudaStream_t passed as void*; all work is stream-ordered, nothing
 *     synchronises the device;
 *   - every entry returns 0 on success and a negative MDT_E* code on failure; the text
 *     of the last failure of the calling thread is mdt_last_error_string();
 *   - nothing throws across this boundary; there is no CPU fallback behind it.
 *
 * Each entry cites the reference interface (file:line under /root/reference) it replaces.
 */
#ifndef MDT_B200_H
#define MDT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MDT_ABI_VERSION 1

#define MDT_OK 0
#define MDT_EINVAL (-1)   /* bad argument (shape, alignment, dtype, range) */
#define MDT_ECUDA (-2)    /* CUDA runtime / driver error, see mdt_last_error_string() */
#define MDT_ENOSUP (-3)   /* valid request outside what the kernels cover */

/* dtype codes for float buffers */
#define MDT_F32 0
#define MDT_BF16 1

/* spatial_pos flavours produced by mdt_collate_graphs */
#define MDT_SPATIAL_CANTOR 0 /* mDT: Cantor-paired (hops up, hops down), pre_processing.py:31-53 */
#define MDT_SPATIAL_SPD 1    /* upstream Graphormer: shortest-path distance itself */

/* bias-code bytes consumed by the fused attention (bias synthesised in-kernel).
 * Legal bytes: spatial codes 0..MDT_CODE_SPATIAL_MAX, MDT_CODE_VIRTUAL, MDT_CODE_MASKED.  Bytes 62..253 are
 * RESERVED: the persistent forward indexes a 64-entry table by (code & 63), so they would alias; callers must not
 * pass them (mdt_collate_graphs never emits them: Cantor codes are <= 21, SPD codes <= spatial_pos_max, which it
 * limits to MDT_CODE_GRAD_MAX when codes are requested).  grad_lut is exact for codes 0..MDT_CODE_GRAD_MAX and
 * MDT_CODE_VIRTUAL; gradients of spatial codes above MDT_CODE_GRAD_MAX are accumulated into entry MDT_CODE_GRAD_MAX
 * (use the materialised bias for wider tables). */
#define MDT_CODE_SPATIAL_MAX 61
#define MDT_CODE_GRAD_MAX 22
#define MDT_CODE_VIRTUAL 254 /* graph-token virtual distance t[h], graphormer_layers.py:105-106 */
#define MDT_CODE_MASKED 255  /* -inf */

typedef void *mdt_stream_t;

int mdt_abi_version(void);
const char *mdt_last_error_string(void);
/* 1 when the current device is sm_100 (B200); the library refuses to launch otherwise. */
int mdt_device_ok(void);

/* ------------------------------------------------------------------------------------
 * A1  floyd_warshall                       mDT/src/data/algos.pyx:7-52
 * Batched all-pairs shortest paths on dense adjacency matrices, bit-exact against the
 * Cython routine: M = min(distance, 510), diagonal 0, zero entries mean "no edge".
 *   adj        [B, n, n], elem_bytes 1 (uint8) | 4 (int32) | 8 (int64), values >= 0
 *   M          [B, n, n] int64 out (may be NULL)
 *   path       [B, n, n] int64 out, last strictly-improving k (algos.pyx:41-43), -1 / 510;
 *              NULL skips the predecessor matrix (and the slower kernel that orders k
 *              strictly); needs workspace of mdt_apsp_workspace_bytes(B, n)
 *   n <= 512.
 * ------------------------------------------------------------------------------------ */
int64_t mdt_apsp_workspace_bytes(int B, int n);
int mdt_apsp_batched(const void *adj, int elem_bytes, int B, int n, int64_t *M, int64_t *path,
                     void *workspace, mdt_stream_t stream);

/* A1b gen_edge_input                         mDT/src/data/algos.pyx:55-89
 *   path [B,n,n] int64 (from mdt_apsp_batched), edge_feat [B,n,n,F] int64
 *   out  [B,n,n,max_dist,F] int64, -1 filled
 *   status [B] int32: 0, or 1 where a path was longer than max_dist (the reference raises). */
int mdt_gen_edge_input(const int64_t *path, const int64_t *edge_feat, int B, int n, int F, int max_dist,
                       int64_t *out, int32_t *status, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * A2 + A3  preprocess_item + collator (graph part)
 *          mDT/src/data/pyg_datasets/pre_processing.py:23-69, mDT/src/data/collator.py:69-179
 * One launch builds every padded graph tensor of a batch from raw edge lists.
 *   edge_src/edge_dst [E_total] int32 node ids local to their graph; graph b owns
 *                     edges [edge_offsets[b], edge_offsets[b+1])
 *   n_nodes           [B] int32 (1 <= n_b <= N)
 *   N                 padded node count (max over the batch), N <= 512;  T = N + 1
 *   spatial_pos       [B,N,N] int32 out: index+1, 0 = padding          (collator.py:59-66)
 *   attn_bias         [B,T,T] out, 0 / -inf: -inf where hop distance >= spatial_pos_max
 *                     (collator.py:122-126) and on padded key columns (collator.py:38-47)
 *   in_degree         [B,N] int64 out: degree+1, 0 = padding            (collator.py:8-20)
 *   bias_code         [B,T,code_pitch] uint8 out or NULL: one byte per (query,key) for the
 *                     fused attention: spatial_pos value, MDT_CODE_VIRTUAL on the graph-token
 *                     row/column, MDT_CODE_MASKED where attn_bias is -inf
 *   adj_workspace     [B,N8,N8] uint8 scratch, N8 = N rounded up to 8
 *   The hop pairs of the reference's distance_matrix are recovered from the tree rooted at
 *   node 0: depth = M[0,:], up = (M + d_i - d_j)/2, down = (M - d_i + d_j)/2.  This equals the reference's
 *   distance_matrix for undirected trees rooted at node 0 only (MDT_SPATIAL_CANTOR); the host side
 *   (data/collator.py: check_discussion_tree) refuses anything else.  Edge endpoints must lie inside their graph.
 * ------------------------------------------------------------------------------------ */
int mdt_collate_graphs(const int32_t *edge_src, const int32_t *edge_dst, const int32_t *edge_offsets,
                       const int32_t *n_nodes, int B, int N, int spatial_pos_max, int flavour,
                       int32_t *spatial_pos, void *attn_bias, int bias_dtype, int64_t *in_degree,
                       uint8_t *bias_code, int code_pitch, uint8_t *adj_workspace, mdt_stream_t stream);

/* mdt_collate_graphs without a caller-provided workspace, for callers that cannot allocate device memory (CUDA graph
 * capture, serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the reference.
 * Same arguments, results and errors as mdt_collate_graphs minus adj_workspace: the
 * [B,N8,N8] adjacency scratch is a file-scope device array of MDT_COLLATE_STATIC_MAX_BYTES inside the library (held by
 * every GPU that loads it until exit).  B * N8 * N8 beyond that limit is rejected with MDT_EINVAL (e.g. 32 graphs of 512 nodes,
 * 128 of 256, 2048 of 64 fit).  The call allocates nothing (no cudaMalloc / cudaFree).  All calls share the one array:
 * the part a call uses is cleared on `stream` first, and calls that overlap on the same GPU (different streams) race --
 * serialise them. */
#define MDT_COLLATE_STATIC_MAX_BYTES (8u << 20)
int mdt_collate_graphs_static(const int32_t *edge_src, const int32_t *edge_dst, const int32_t *edge_offsets,
                              const int32_t *n_nodes, int B, int N, int spatial_pos_max, int flavour,
                              int32_t *spatial_pos, void *attn_bias, int bias_dtype, int64_t *in_degree,
                              uint8_t *bias_code, int code_pitch, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * A5  GraphAttnBias.forward / backward      mDT/src/modules/graphormer_layers.py:86-110
 *   out[b,h,i,j] = 2*attn_bias[b,i,j] + (i>=1 && j>=1 ? table[spatial_pos[b,i-1,j-1], h]
 *                                                      : virt[h])
 *   attn_bias [B,T,T], out / grad_out [B,H,T,T] in `dtype`; table [num_spatial,H] and
 *   virt [H] in float32 (the parameters' master copies).
 *   backward: grad_table [num_spatial,H] f32 (row 0 = padding_idx gets no gradient),
 *   grad_virt [H] f32; both are accumulated into (caller zeroes them).
 * ------------------------------------------------------------------------------------ */
int mdt_attn_bias_fwd(const void *attn_bias, const int32_t *spatial_pos, const float *table, const float *virt,
                      int B, int H, int N, int num_spatial, int dtype, void *out, mdt_stream_t stream);
int mdt_attn_bias_bwd(const void *grad_out, const int32_t *spatial_pos, int B, int H, int N, int num_spatial,
                      int dtype, float *grad_table, float *grad_virt, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * A6  MultiheadAttention core                mDT/src/modules/multihead_attention.py:168-199
 *   S = q k^T (+bias) (+key padding -> -inf); P = softmax_fp32(S); O = P v
 * tcgen05 / TMEM / TMA kernel, head_dim 64, bf16 operands, fp32 accumulation and softmax.
 *   q,k,v     bf16, element (b,t,h,c) at b*stride_b + t*stride_t + h*stride_h + c
 *             (strides in elements, multiples of 8; q is pre-scaled by d^-0.5 as at
 *             multihead_attention.py:137)
 *   bias      materialised [B,H,T,T] (`bias_dtype` MDT_F32|MDT_BF16) or NULL
 *   code      [B,T,code_pitch] uint8 + lut [H,256] f32 (bias = lut[h][code]) or NULL
 *             (in-kernel synthesis: the [B,H,T,T] tensor never exists)
 *   key_pad   [B,T] uint8 (non-zero = padded key) or NULL
 *   out       bf16, same addressing as q with its own strides
 *   lse       [B,H,T] f32 out: log-sum-exp of each row (saved for backward)
 * backward: grad_out like out; dq/dk/dv like q/k/v with their own strides (dq must be
 *   zero-filled f32 scratch `dq_accum` [B,H,T,64] -- it is reduced across key tiles -- and is
 *   converted into dq); grad_bias [B,H,T,T] (`bias_dtype`) or NULL; grad_lut [H,256] f32
 *   accumulated or NULL.  code_t [B,T,code_pitch] (optional, backward only) is the transpose of code inside each
 *   graph, code_t[b][k][q] = code[b][q][k] (columns q >= T: MDT_CODE_MASKED); build it once per batch -- with it the
 *   backward runs its warp-specialised kernel (a thread reads the codes of one key against 64 queries as 4 vector
 *   loads); without it (NULL) the byte-gather kernel runs.  delta [B,H,T] f32 is scratch (may stay untouched).
 * ------------------------------------------------------------------------------------ */
typedef struct mdt_attn_dims {
    int32_t B, H, T;
    int64_t q_stride_b, q_stride_t, q_stride_h;
    int64_t k_stride_b, k_stride_t, k_stride_h;
    int64_t v_stride_b, v_stride_t, v_stride_h;
    int64_t o_stride_b, o_stride_t, o_stride_h;
} mdt_attn_dims;

/* dropout_p in [0,1) drops attention probabilities as multihead_attention.py:196 (after the softmax, survivors scaled by
 * 1/(1-p)); the keep mask is a counter hash of (dropout_seed, b, h, q, k) that the backward regenerates from the same
 * seed -- nothing is stored.  dropout_p = 0 switches it off. */
int mdt_biased_attn_fwd(const void *q, const void *k, const void *v, const mdt_attn_dims *dims,
                        const void *bias, int bias_dtype, const uint8_t *code, int code_pitch,
                        const float *lut, const uint8_t *key_pad, void *out, float *lse, float dropout_p,
                        uint64_t dropout_seed, mdt_stream_t stream);

/* Diagnostics (no reference counterpart): when dev_buffer is non-null the persistent forward writes 64
 * %globaltimer stamps per CTA into it (u64 [n_ctas,64], device memory); null switches tracing off. */
int mdt_debug_attn_trace(unsigned long long *dev_buffer);

int mdt_biased_attn_bwd(const void *q, const void *k, const void *v, const void *out, const void *grad_out,
                        const float *lse, const mdt_attn_dims *dims, const void *bias, int bias_dtype,
                        const uint8_t *code, const uint8_t *code_t, int code_pitch, const float *lut,
                        const uint8_t *key_pad, void *dq, void *dk, void *dv, const mdt_attn_dims *grad_dims, float *dq_accum,
                        float *delta, void *grad_bias, float *grad_lut, float dropout_p, uint64_t dropout_seed,
                        mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * N1  self-attention over PACKED (unpadded) sequences: the BERT side of the fusion layers
 *     mDT/src/modules/multi_graphormer_fusion_layer.py:105-149 (Hugging Face BertSelfAttention called at :138-148:
 *     softmax(q k^T / sqrt(d) + key mask) v with dropout on the probabilities), mDT/src/modules/multigraphormer_graph_encoder.py:341-362
 *   Tokens of all comments are laid end to end: sequence s = rows [cu_seqlens[s], cu_seqlens[s+1]) of q / k / v
 *   (bf16, row r / head h / channel c at r*stride + h*64 + c -- e.g. the three slices of one [total,3,H,64]
 *   projection); scores are q k^T * softmax_scale (the scale is folded into the exponent: q is NOT pre-scaled).
 *   Every sequence must be <= 128 rows.  A padded token never influences
 *   a real one in the reference (additive key mask), so attention restricted to each sequence is the same function.
 *   mdt_varlen_attn_plan groups consecutive sequences into <= 128-row tiles (group != 0; group == 0: one sequence
 *   per tile) and precomputes every packed row's key window.  plan is int32 [mdt_varlen_attn_plan_words(n_seq, total)]:
 *     [0] n_items  [1] error flag (a sequence > 128 rows: the attention calls then do nothing)  [2] n_seq  [3] total
 *     [4 + 2i] first row of item i, [5 + 2i] its first sequence (i = 0..n_items; entry n_items = {total, n_seq})
 *     then n_seq scratch words and `total` row windows (lo | hi << 16, relative to the item's first row).
 *   Build it once per batch; every layer, forward and backward, reuses it (cu_seqlens is not read again).
 *   forward : out [total, H*64] (row stride o_stride), lse f32 [H, total]
 *   backward: dq / dk / dv rows like q / k / v with their own strides (each row written exactly once: no atomics,
 *             no scratch); out / grad_out rows with o_stride / g_stride.
 *   Dropout (BertSelfAttention.dropout) is a counter hash of (seed, item, head, row, column) regenerated in the
 *   backward, so forward and backward must be given the same plan.
 * ------------------------------------------------------------------------------------ */
int64_t mdt_varlen_attn_plan_words(int n_seq, int total);
int mdt_varlen_attn_plan(const int32_t *cu_seqlens, int n_seq, int total, int group, int32_t *plan, mdt_stream_t stream);
int mdt_varlen_attn_fwd(const void *q, const void *k, const void *v, int64_t q_stride, int64_t k_stride, int64_t v_stride,
                        int total, int H, int n_seq, const int32_t *plan, void *out, int64_t o_stride, float *lse,
                        float softmax_scale, float dropout_p, uint64_t dropout_seed, mdt_stream_t stream);
int mdt_varlen_attn_bwd(const void *q, const void *k, const void *v, int64_t q_stride, int64_t k_stride, int64_t v_stride,
                        const void *out, int64_t o_stride, const void *grad_out, int64_t g_stride, const float *lse,
                        int total, int H, int n_seq, const int32_t *plan, void *dq, void *dk, void *dv,
                        int64_t dq_stride, int64_t dk_stride, int64_t dv_stride, float softmax_scale, float dropout_p,
                        uint64_t dropout_seed, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * A4  GraphNodeFeature.forward               mDT/src/modules/graphormer_layers.py:39-50
 *   out[b,0,:] = graph_token; out[b,1+i,:] = x[b,i,:] + E_in[in_degree[b,i]] + E_out[out_degree[b,i]]
 *   x [B,N,C], tables [num_degree,C], out [B,N+1,C], all `dtype`; degrees int64.
 * backward: grad_x = grad_out[:,1:,:]; grad tables / token accumulated in f32
 *   (index 0 = padding_idx gets no gradient).
 * ------------------------------------------------------------------------------------ */
int mdt_node_feature_fwd(const void *x, const int64_t *in_degree, const int64_t *out_degree, const void *w_in,
                         const void *w_out, const void *graph_token, int B, int N, int C, int num_in,
                         int num_out, int dtype, void *out, mdt_stream_t stream);
int mdt_node_feature_bwd(const void *grad_out, const int64_t *in_degree, const int64_t *out_degree, int B, int N,
                         int C, int num_in, int num_out, int dtype, void *grad_x, float *grad_w_in,
                         float *grad_w_out, float *grad_token, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * K12  node-major <-> padded row moves       mDT/src/modules/multigraphormer_graph_encoder.py:363-371,425,435
 * Replaces boolean-mask indexing (which costs a device->host sync per call) by a
 * precomputed int32 row list.
 *   gather : dst[r, :] = src[index[r], :]   for r < n_rows
 *   scatter: dst[index[r], :] = src[r, :]
 *   rows are `row_elems` elements of `dtype`, with independent row pitches (in elements).
 * ------------------------------------------------------------------------------------ */
int mdt_gather_rows(const void *src, int64_t src_pitch, const int32_t *index, int n_rows, int row_elems,
                    int dtype, void *dst, int64_t dst_pitch, mdt_stream_t stream);
int mdt_scatter_rows(const void *src, int64_t src_pitch, const int32_t *index, int n_rows, int row_elems,
                     int dtype, void *dst, int64_t dst_pitch, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * N2  dense -> dropout -> + residual -> LayerNorm block tail, one pass each way
 *     (Hugging Face BertSelfOutput / BertOutput inside the fusion layers, mDT/src/modules/multi_graphormer_fusion_layer.py:105-149;
 *      mDT/src/modules/graphormer_graph_encoder_layer.py:127-130,139-141)
 *   forward : z = dropout_p(x) + residual (residual may be null);  y = LayerNorm(z) * gamma + beta
 *             x, residual, gamma, beta, y, z: bf16 [rows, C] / [C];  mean, rstd: f32 [rows];  C in {256,512,768,1024}
 *             z (the rounded pre-norm sum) is saved for the backward; it may be null only without residual and
 *             dropout (then the backward is given x as z).  The keep mask is a counter hash of (seed, element
 *             index): it is never stored, the backward regenerates it from the same seed.
 *   backward: dres = d z (null to skip), dx = keep * d z / (1 - p) (null to skip), dgamma / dbeta f32 [C] are
 *             ACCUMULATED into (the caller zeroes them).
 * ------------------------------------------------------------------------------------ */
int mdt_dropout_add_ln_fwd(const void *x, const void *residual, const void *gamma, const void *beta, void *y, void *z,
                           float *mean, float *rstd, int64_t rows, int C, float p, uint64_t seed, float eps,
                           mdt_stream_t stream);
int mdt_dropout_add_ln_bwd(const void *dy, const void *z, const float *mean, const float *rstd, const void *gamma,
                           void *dx, void *dres, float *dgamma, float *dbeta, int64_t rows, int C, float p,
                           uint64_t seed, mdt_stream_t stream);
/* the same backward with one more accumulated output: dxsum f32 [C] += column sums of the (rounded) tensor written to dx
 * (to dres when dx is null) -- the bias gradient of the dense layer whose output was x (BertSelfOutput.dense /
 * BertOutput.dense, graphormer_graph_encoder_layer.py fc2 / out_proj), so that gradient needs no pass of its own.
 * dadd (bf16 [rows, C], may be null) is added to d z before it is written to dres and masked into dx: the gradient of a
 * residual connection that joins at z (the pre-LN blocks of the ViT: d x = d(residual path) + LayerNorm-path gradient) needs
 * no add pass of its own.  dxsum and dadd null: this is mdt_dropout_add_ln_bwd. */
int mdt_dropout_add_ln_bwd_colsum(const void *dy, const void *z, const float *mean, const float *rstd, const void *gamma,
                                  const void *dadd, void *dx, void *dres, float *dgamma, float *dbeta, float *dxsum,
                                  int64_t rows, int C, float p, uint64_t seed, mdt_stream_t stream);
/* y = dropout_p(x) + residual (residual may be null) without a LayerNorm behind it -- the second residual of a pre-LN block
 * (Hugging Face ViTOutput, called from mDT/src/modules/multi_graphormer_fusion_layer.py:73-100) -- and its backward
 * dx = keep * dy / (1 - p) (dx may be null) with dxsum f32 [C] += column sums of the rounded dx (the bias gradient of the
 * dense layer that produced x; may be null).  bf16 [rows, C], C a multiple of 8; same counter-hash mask as the LayerNorm
 * entries (a function of seed and element index, never stored). */
int mdt_dropout_add_fwd(const void *x, const void *residual, void *y, int64_t rows, int C, float p, uint64_t seed,
                        mdt_stream_t stream);
int mdt_dropout_bwd(const void *dy, void *dx, float *dxsum, int64_t rows, int C, float p, uint64_t seed, mdt_stream_t stream);

/* ------------------------------------------------------------------------------------
 * N2  GELU of the feed-forward blocks (exact erf form, fp32 arithmetic on bf16 tensors)
 *     mDT/src/modules/graphormer_graph_encoder_layer.py:132-136 (fc1 -> gelu(x.float()).type_as(x)); Hugging Face
 *     BertIntermediate / ViTIntermediate called from mDT/src/modules/multi_graphormer_fusion_layer.py:94-98,138-148
 *   forward : g = u * Phi(u)                                   u, g bf16 [rows, C], C a multiple of 8
 *   backward: du = dg * (Phi(u) + u phi(u));  dbias[c] += sum_r du[r,c]   (f32 [C], accumulated; NULL to skip) --
 *             the bias gradient of the dense layer in front of the activation, without a further pass over du.
 *             du may be dg (the gradient is overwritten in place).
 * ------------------------------------------------------------------------------------ */
int mdt_gelu_fwd(const void *u, void *g, int64_t rows, int C, mdt_stream_t stream);
int mdt_gelu_bwd(const void *dg, const void *u, void *du, float *dbias, int64_t rows, int C, mdt_stream_t stream);

/* N2  dense (+ GELU) as one tcgen05 GEMM with a fused epilogue     (same reference lines as mdt_gelu_fwd)
 *   out[M,N] = a[M,K] * w[N,K]^T + bias[N];  act = 1: aux = that (the pre-activation, saved for the backward) and
 *   out = gelu(aux) -- the activation then costs one extra write instead of a read + write pass over [M,N].
 *   bf16 row-major; lda / ldw = leading dimensions of a / w in elements; out / aux dense; N % 256 == 0, K % 64 == 0. */
int mdt_gemm_bias_act(const void *a, int64_t lda, const void *w, int64_t ldw, const void *bias, void *out, void *aux,
                      int M, int N, int K, int act, mdt_stream_t stream);
/* dU[M, N] = (dY[M, K] W[K, N]) * gelu'(U[M, N]): the input-gradient GEMM of the dense layer behind an exact GELU (fc2 of the
 * feed-forward blocks, same reference lines) with the activation's derivative applied in its epilogue; W row-major [K, N] (the
 * layer's weight as stored), bf16, N % 256 == 0, K % 64 == 0.  The bias gradient of fc1 (column sums of dU) comes out of
 * mdt_wgrad_gemm's colsum on the same dU. */
int mdt_gemm_dgrad_gelu(const void *dy, int64_t ldy, const void *w, int64_t ldw, const void *u, void *du, int M, int N, int K,
                        mdt_stream_t stream);

/* N2  weight gradient of a dense layer: out[N, C] += dy[M, N]^T * x[M, C]   (torch: dy.t() @ x; Hugging Face dense layers called
 *     from mDT/src/modules/multi_graphormer_fusion_layer.py:94-98,138-148, backward pass)
 *   dy, x bf16 row-major with leading dimensions ldy, ldx (elements, multiples of 8); out fp32 [N, ldo], ACCUMULATED into with
 *   vector atomics (split-K: the caller zeroes it, or passes a running fp32 gradient); N, C multiples of 256; M < 2^31 rows.
 *   colsum fp32 [N] (may be NULL) += the column sums of dy -- the bias gradient of the same layer -- added up from the tiles
 *   the kernel has staged anyway. */
int mdt_wgrad_gemm(const void *dy, int64_t ldy, const void *x, int64_t ldx, float *out, int64_t ldo, float *colsum, int64_t M,
                   int N, int C, mdt_stream_t stream);

/* Self-test of the tcgen05 building blocks (shared-memory descriptor layouts, TMA 128-byte
 * swizzle, TMEM read-back); called by tests/ only.  All operands bf16, row major:
 *   d1[128,128] = a[128,64] * b[128,64]^T          (A and B K-major, both via TMA)
 *   d2[128,64]  = p[128,128] * v[128,64]           (A K-major written by threads, B MN-major)
 *   d3[128,64]  = at[128(K),128(M)]^T * bk[64(N),128(K)]^T   (A MN-major, B K-major)
 *   d4[128,64]  = p[128,128] * v[128,64]           (A read from TENSOR MEMORY, written there by tcgen05.st; B MN-major)
 * d1..d4 are f32 device buffers. */
int mdt_umma_selftest(const void *a, const void *b, const void *p, const void *v, const void *at, const void *bk,
                      float *d1, float *d2, float *d3, float *d4, mdt_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MDT_B200_H */

/* stemgnn_b200.h -- C-ABI of the B200-native STEM-GNN encoder/VQ hot path.
 *
 * The reference (GXG-CS/STEM-GNN) is pure Python; the "FFI" it binds for this path is the
 * ATen / torch_scatter operator layer underneath torch_geometric.  Each entry point below
 * replaces one such operator sequence, cited as reference file:line (paths relative to
 * /root/reference/STEM-GNN).  Conventions:
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - the caller owns all memory (outputs and workspace); kernels are stream-ordered on
 *     `stream`, never synchronise, never allocate;
 *   - return value 0 on success, a negative STEMGNN_E* code otherwise;
 *     stemgnn_last_error() returns a thread-local message;
 *   - row-major contiguous fp32 features; int64 edge_index [2,E] as the reference passes it;
 *     the internal CSR uses int32 (E, N < 2^31).
 */
#ifndef STEMGNN_B200_H
#define STEMGNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* stemgnn_stream_t; /* cudaStream_t */

#define STEMGNN_OK 0
#define STEMGNN_EINVAL (-1)
#define STEMGNN_ENOMEM_WS (-2)
#define STEMGNN_ECUDA (-3)
#define STEMGNN_EUNSUPPORTED (-4)

const char* stemgnn_last_error(void);
/* Layout version of the argument structs below; a binding checks it at load time.
 * 2: out_ld / row_x_ld in stemgnn_aggregate_args; 3: stemgnn_bn_act_backward_apply and the loop-side
 * entry points (augmentations, sampler, l2norm, negative sampling); 4: hub_ws / hub_ws_bytes appended to
 * stemgnn_aggregate_args; 5: accumulate appended to stemgnn_aggregate_args; 6: row_div appended to
 * stemgnn_linear_args, row_add / row_add_ld appended to stemgnn_aggregate_args. */
int stemgnn_abi_version(void);
/* SMs left free by the persistent one-CTA-per-SM kernels (stemgnn_linear, stemgnn_linear_wgrad) from now on;
 * returns the previous value.  A NCCL collective running beside them (the halo exchange of a partitioned graph,
 * the all-gather the reference's propagate at model/encoder.py:82 needs once the nodes are split over GPUs) holds a
 * few dozen SMs whose remaining shared memory cannot take a 200 KB CTA: with a grid of all SMs those CTAs -- and
 * their statically dealt share of the tiles -- would wait for the collective to END.  Process-wide, 0 by default. */
int stemgnn_set_sm_reserve(int sms);
/* number of CUDA kernels this library has launched in this process (all entry points) */
uint64_t stemgnn_launch_count(void);

/* ------------------------------------------------------------------ K0: graph structure
 * Replaces the implicit COO scatter order of MessagePassing.propagate (model/encoder.py:82)
 * and scatter_mean (model/encoder.py:124) with an explicit CSR grouped by
 * edge_index[group_row] using a STABLE sort (ties keep original edge order), which is the
 * order scatter_add_ sums in.  Outputs: rowptr[N+1], col[E] (the other endpoint), perm[E]
 * (original edge id per slot), cnt[N] = float(max(deg,1)) (MeanAggregation's clamped count). */
int stemgnn_csr_build_workspace(int64_t num_edges, int64_t num_nodes, size_t* bytes_host);
int stemgnn_csr_build(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes,
                      int group_row, int32_t* rowptr, int32_t* col, int32_t* perm, float* cnt,
                      void* workspace, size_t workspace_bytes, stemgnn_stream_t stream);

/* Both groupings of ONE edge_index in the same launches (index 0: grouped by edge_index[0], the transposed
 * structure the backward pass walks; index 1: grouped by edge_index[1], the forward structure): what a training
 * step needs of every fresh mini-batch edge list (STEM-GNN/pretrain.py:41-45,151-153 hand the encoder a new
 * edge_index per batch; model/encoder.py:82 and its autograd consume both orders).  Same outputs, bit for bit,
 * as two stemgnn_csr_build calls; 4 kernel launches instead of 8 for 18-bit node ids. */
int stemgnn_csr_build_pair_workspace(int64_t num_edges, int64_t num_nodes, size_t* bytes_host);
int stemgnn_csr_build_pair(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes,
                           int32_t* const rowptr[2], int32_t* const col[2], int32_t* const perm[2],
                           float* const cnt[2], void* workspace, size_t workspace_bytes, stemgnn_stream_t stream);

/* GCNConv.gcn_norm (torch_geometric 2.3.0, called from model/encoder.py:197): given the
 * edge list WITH remaining self loops already appended (edge_index_sl; the host side builds
 * it with integer mask/concat ops) and the CSR rowptr grouped by destination,
 * deg[i] = float(in-degree incl. self loop), dis = deg^-1/2 (inf -> 0),
 * w[e] = dis[src] * 1.0f * dis[dst] in original edge order (two separate multiplies). */
int stemgnn_gcn_edge_weight(const int64_t* edge_index_sl, int64_t num_edges, int64_t num_nodes,
                            const int32_t* rowptr_by_dst, float* w_out, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K1/K2: aggregation
 * Forward: replaces index_select + message + scatter_add_ + count/clamp/div
 * (model/encoder.py:82,94-97; PyG MeanAggregation; torch_scatter.scatter_mean at :124; GCN/GIN
 * sums at :197/:199).  For every row r of the CSR:
 *     acc = 0;  for slots p in [rowptr[r], rowptr[r+1]) in order:
 *         m = x[col[p]] (+ xe[perm[p]] | + table[etype[perm[p]]]);  if act==RELU m = max(m,0);
 *         if w: m = w[perm[p]] * m (separate rounding);  acc = acc + m
 *     out[r] = reduce==MEAN ? acc / cnt[r] : acc
 * Backward (same entry point on the TRANSPOSED structure, mode flags below):
 *     g_x[r] = sum_p mask(r,p) * ( g_out[col[p]] / cnt_col[col[p]] ) [* w[perm[p]]]
 *     mask = (x[r] (+ xe[perm[p]]) > 0) when act==RELU
 * which is autograd's index_add_ of the masked, pre-divided upstream gradient in edge order. */
#define STEMGNN_REDUCE_SUM 0
#define STEMGNN_REDUCE_MEAN 1
#define STEMGNN_ACT_NONE 0
#define STEMGNN_ACT_RELU 1

typedef struct {
    const int32_t* rowptr;   /* [n_rows+1] */
    const int32_t* col;      /* [E] gathered-row index per slot */
    const int32_t* perm;     /* [E] original edge id per slot (needed for xe/etype/w) or NULL */
    int64_t n_rows;          /* rows produced */
    int64_t n_gather_rows;   /* rows of the gathered matrix (bounds) */
    int64_t num_edges;
    int64_t feat;            /* F */
    const float* gathered;   /* fwd: x [n_gather_rows,F]; bwd: g_out */
    const float* row_cnt;    /* fwd MEAN: cnt[n_rows]; else NULL */
    const float* gather_cnt; /* bwd MEAN: cnt over gathered rows (divide each gathered row); else NULL */
    const float* edge_feat;  /* dense xe [E,F] in ORIGINAL edge order, or NULL */
    const float* type_table; /* [T,F] or NULL (used with edge_type) */
    const int32_t* edge_type;/* [E] in original edge order or NULL */
    const float* edge_weight;/* [E] in original edge order or NULL */
    const float* row_x;      /* bwd RELU: x [n_rows,F] whose sign (plus xe) gives the mask; else NULL */
    int act;                 /* STEMGNN_ACT_* */
    int backward;            /* 0 fwd, 1 bwd */
    float* out;              /* [n_rows,F] */
    int64_t out_ld;          /* row stride of out in floats; 0 = feat (contiguous).  ABI 2 */
    int64_t row_x_ld;        /* row stride of row_x in floats; 0 = feat */
    void* hub_ws;            /* device scratch of stemgnn_aggregate_hub_workspace(num_edges) bytes, or NULL.  ABI 4 */
    size_t hub_ws_bytes;
    int accumulate;          /* ABI 5.  1: every row sum STARTS from the value already in out[r] (the additions of this
                              * call follow it, in edge order) instead of 0 -- one aggregation cut into several calls
                              * over disjoint edge subsets (source-owner blocks of a partitioned graph, aggregated
                              * as their halo blocks land).  Pass row_cnt / the backward relu mask (act with
                              * backward=1) only in the LAST call of such a chain: both apply to the finished sum. */
    int reserved;
    const float* row_add;    /* ABI 6.  [n_rows, F] (row stride row_add_ld floats, 0 = feat) or NULL: added to the
                              * FINISHED row -- after the count division / backward relu mask -- with one rounding:
                              * out[r] = finish(sum) + row_add[r].  The backward of a SAGE layer adds the lin_r term
                              * g @ W_r to the aggregated one (autograd's accumulation into x.grad, reference
                              * model/encoder.py:84-88) here instead of in a pass of its own; may alias out. */
    int64_t row_add_ld;
} stemgnn_aggregate_args;

int stemgnn_aggregate(const stemgnn_aggregate_args* args_host, stemgnn_stream_t stream);

/* Hub rows.  Real citation / co-purchase graphs have rows with 10^4 edges (ogbn-arxiv: largest
 * in-degree 13,155; ogbn-products: 17,481): walked by one group of lanes such a row alone outlasts
 * the rest of the launch.  With hub_ws set, rows of >= 256 edges are put on a device work list by
 * the first launch and summed by a second one that cuts each of them into column slices (32 columns,
 * 8 from 4,096 edges on) over many CTAs, longest rows first, and stages 256 / 1,024 edges at a time in
 * shared memory -- the additions of every column stay in edge order, so the result is the same bits.
 * The scratch holds four counters and the list.  The caller zero-fills it ONCE (cudaMemset) before its
 * first use; every call leaves it zeroed again.  It must not be shared by calls that may run
 * concurrently.  hub_ws == NULL keeps every row in the first launch. */
int stemgnn_aggregate_hub_workspace(int64_t num_edges, size_t* bytes);

/* K2a. out[r, :] = x[r, :] / cnt[r] (IEEE division, rows of `feat` floats; out may alias x).
 * MeanAggregation's backward divides the upstream gradient of a destination by its clamped count
 * BEFORE index_add_ scatters it over the edges (torch_geometric MeanAggregation -> scatter 'mean'
 * autograd; reference call site STEM-GNN/model/encoder.py:82).  The quotient is the same for every
 * edge of that destination, so the host op applies it once per row with this call and then runs
 * stemgnn_aggregate(backward=1, gather_cnt=NULL): same bits as dividing per edge, E*F fewer
 * divisions, and in the partitioned path no count exchange. */
int stemgnn_row_scale(int64_t n_rows, int64_t feat, const float* x, const float* cnt, float* out,
                      stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K3: VQ assignment
 * Replaces dist = -cdist / einsum, argmax, one_hot, one_hot @ embed (or the eval gather) and
 * the MSE commitment partial (model/vq.py:472-482, 650-659, 79-80, 1007), and fuses the head
 * rearranges 'b n (h d) -> h b n d' / 'h b n d -> b n (h d)' / 'h b n -> b n h'
 * (model/vq.py:885-887, 971, 1034) through strides:
 *   element (h, m, d) of x        is x       [m*x_row_stride + h*x_head_stride + d]
 *   element (h, m, d) of quantize is quantize[m*q_row_stride + h*q_head_stride + d]
 *   element (h, m)    of ind      is ind     [m*ind_row_stride + h*ind_head_stride]
 * embed is [embed_heads, K, D] contiguous with embed_heads in {1, heads}.
 * metric 0: euclidean, argmax of -sqrt(clamp((x2 + e2) + (-2 x.e), 0)), first index on ties;
 * metric 1: dot / cosine, argmax of x.e (the caller l2-normalises, model/vq.py:891).
 * sq_err[h] (fp64) = sum over (m,d) of (quantize - x)^2, reduced in a fixed order.
 * path 0 = auto, 1 = exact fp32 CUDA-core scan, 2 = tcgen05 fp16 GEMM with a rigorous
 * per-row error band: rows whose band holds one code are final, the others are re-scored in
 * fp32 over the codes inside the band (same formula, first index on ties), so path 2 returns
 * the fp32 answer up to the summation order of the dot product.  stats (optional, int64[4],
 * device): [0] rows decided by fp32 re-scoring of a candidate list, [1] rows that needed an
 * exact scan over all codes (candidate overflow / |x| outside the fp16 range), [2..3] = 0. */
typedef struct {
    const float* x; int64_t x_row_stride, x_head_stride;
    const float* embed;
    int64_t heads, embed_heads, rows, codes, dim;
    int metric, path;
    int64_t* ind; int64_t ind_row_stride, ind_head_stride;
    float* quantize; int64_t q_row_stride, q_head_stride;
    double* sq_err;   /* [heads] or NULL */
    int64_t* stats;   /* [4] or NULL */
} stemgnn_vq_assign_args;

int stemgnn_vq_assign_workspace(const stemgnn_vq_assign_args* args_host, size_t* bytes_host);
int stemgnn_vq_assign(const stemgnn_vq_assign_args* args_host, void* workspace,
                      size_t workspace_bytes, stemgnn_stream_t stream);

/* backward of straight-through + commitment (model/vq.py:937,1007-1009), elementwise over
 * `numel` contiguous floats: g_x = g_q + g_loss[0] * coef * (x - q), coef = 2*w/numel.
 * g_q may be NULL (treated as 0); g_loss may be NULL (treated as 0). */
int stemgnn_vq_backward(const float* g_q, const float* x, const float* q, const float* g_loss,
                        float coef, int64_t numel, float* g_x, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K5: dense layers (next row f-1)
 * out[m, n] = sum_{s < n_src} sum_k A_s[m, k] * W_s[n, k]  (+ bias[n]),  fp32 in / fp32 out.
 * Replaces `self.lin_l(out) + self.lin_r(x_r)` of MySAGEConv.forward (reference
 * STEM-GNN/model/encoder.py:84-88) as ONE call with two sources, the single Linear of the GCN / GIN
 * layers, and -- with w_transposed = 1, W_s given as [k, n] -- any A @ W product such as the
 * backward-data term g_out @ lin.weight.  Runs on the tcgen05 tensor cores as 3xTF32 (operands
 * split hi + lo, three MMAs, fp32 accumulate): within 1e-5 of torch's fp32 result, full fp32 range.
 * The tensor-core accumulator truncates (about 7.5e-9 of the output scale per accumulated column), so
 * a caller that needs 1e-5 beyond sum(k) ~ 1024 splits K into slabs and chains calls with accumulate = 1
 * (the drop-in host op does: stem-gnn_b200/dense.py).
 * k[s], lda[s] must be multiples of 4 and a[s] 16-byte aligned (row segments are read as float4). */
typedef struct {
    int64_t m;            /* rows */
    int64_t n;            /* output features */
    int32_t n_src;        /* 1 or 2 */
    int32_t w_transposed; /* 0: w[s] is [n, k[s]] (nn.Linear.weight); 1: w[s] is [k[s], n] */
    const float* a[2];    /* [m, k[s]] row-major, row stride lda[s] floats */
    int64_t lda[2];
    int64_t k[2];
    const float* w[2];    /* row stride ldw[s] floats */
    int64_t ldw[2];
    const float* bias;    /* [n] or NULL */
    float* out;           /* [m, n], row stride ldo floats */
    int64_t ldo;
    int32_t accumulate;   /* 1: out += product (round-to-nearest add in the epilogue) instead of out = */
    int32_t reserved;
    const float* row_div; /* ABI 6.  [m] or NULL: out[m, :] = (product + bias) / row_div[m] (IEEE division) -- the
                           * `g_out / count[dst]` of MeanAggregation's backward (stemgnn_row_scale) applied where the
                           * backward-data product g @ lin_l.weight is written, same bits, one pass over the matrix
                           * less.  Not with accumulate = 1. */
} stemgnn_linear_args;

int stemgnn_linear_workspace(const stemgnn_linear_args* args_host, size_t* bytes_host);
int stemgnn_linear(const stemgnn_linear_args* args_host, void* workspace, size_t workspace_bytes,
                   stemgnn_stream_t stream);

/* K6. Weight (and bias) gradients of the same layers:
 *     gw[s][n, k] = sum_m g[m, n] * a[s][m, k],      gbias[n] = sum_m g[m, n]
 * (autograd of nn.Linear inside MySAGEConv / GCN / GIN, reference model/encoder.py:84-88).  3xTF32 on
 * tcgen05 with the row reduction cut into slabs of 512 rows whose fp32 tiles are added round-to-
 * nearest: the tensor-core accumulator truncates, so a single 169k-row accumulation would not hold
 * 1e-5.  Deterministic (fixed summation order).  n, k[s], ldg, lda[s] multiples of 4; g, a[s]
 * 16-byte aligned; at most 64 tiles of 256 x 256 weights per call. */
typedef struct {
    int64_t m;
    int64_t n;
    int32_t n_src;
    int32_t reserved;
    const float* g;       /* [m, n] upstream gradient, row stride ldg */
    int64_t ldg;
    const float* a[2];    /* [m, k[s]] layer inputs */
    int64_t lda[2];
    int64_t k[2];
    float* gw[2];         /* [n, k[s]] outputs, row stride ldgw[s] */
    int64_t ldgw[2];
    float* gbias;         /* [n] or NULL */
} stemgnn_linear_wgrad_args;

int stemgnn_linear_wgrad_workspace(const stemgnn_linear_wgrad_args* args_host, size_t* bytes_host);
int stemgnn_linear_wgrad(const stemgnn_linear_wgrad_args* args_host, void* workspace,
                         size_t workspace_bytes, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K7: BatchNorm + ReLU + dropout
 * Replaces `z = self.norms[i](z)` and `z = self.dropout(self.activation(z))` (reference
 * STEM-GNN/model/encoder.py:313-317) and their autograd with three passes over [rows, feat] fp32
 * (feat % 4 == 0, 16-byte aligned rows):
 *   stemgnn_bn_stats        mean[c], biased var[c] of column c (what nn.BatchNorm1d uses in training);
 *                           running-statistics updates and invstd = rsqrt(var + eps) are the caller's
 *                           ([feat]-sized torch ops in the drop-in module)
 *   stemgnn_bn_act_forward  out = dropout(relu((x - mean) * invstd * gamma + beta), p)
 *   stemgnn_bn_act_backward sum_g[c] = sum_r g, sum_gx[c] = sum_r g * xhat (= grad of beta, gamma) with
 *                           g = g_y * mask/(1-p) * [pre > 0];  gx = gamma * invstd * (g - sum_g/N -
 *                           xhat * sum_gx/N) when batch_stats, gamma * invstd * g otherwise
 * The dropout mask is Philox4x32-7 keyed by *seed (a uint64 in DEVICE memory, read at kernel time,
 * so a captured CUDA graph sees a fresh seed on every replay) and the element index: it is recomputed
 * in the backward, never stored.  relu / p = 0 switch the activation / dropout off. */
int stemgnn_bn_workspace(int64_t rows, int64_t feat, size_t* bytes_host);
int stemgnn_bn_stats(int64_t rows, int64_t feat, const float* x, int64_t ldx, float* mean, float* var,
                     void* workspace, size_t workspace_bytes, stemgnn_stream_t stream);
/* stemgnn_bn_stats plus what nn.BatchNorm1d does with the statistics of a training batch (F.batch_norm under
 * model/encoder.py:313), in the tail of the same two launches instead of ten [feat]-sized tensor ops per layer:
 * invstd = rsqrt(var + eps); when running_mean / running_var are given, running = running * (1 - m) + stat * m with
 * the unbiased variance var * rows / (rows - 1), m = momentum, or 1 / *num_batches_tracked when momentum < 0 (the
 * cumulative average of momentum=None; the caller increments the module's device counter first).  Each product and
 * sum is rounded separately, like the tensor ops. */
int stemgnn_bn_train_stats(int64_t rows, int64_t feat, const float* x, int64_t ldx, float eps, double momentum,
                           const int64_t* num_batches_tracked, float* running_mean, float* running_var, float* mean,
                           float* var, float* invstd, void* workspace, size_t workspace_bytes, stemgnn_stream_t stream);
/* stemgnn_bn_stats without a caller workspace and without any allocation.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, arithmetic
 * and results (bit for bit) as stemgnn_bn_stats; the per-block partial sums live in a 4 MiB
 * `static __device__` array inside the library (held by every GPU that loads it until the process
 * exits), whose device address is looked up once per device with cudaGetSymbolAddress.  Limit: the
 * partials, 2 * feat floats for each of at most 3 * SM-count row blocks, must fit the 4 MiB -- any row
 * count with feat <= 1,180 on a 148-SM B200 (wider only while rows <= 64 * 4 MiB / (8 * feat));
 * beyond it the call returns STEMGNN_EINVAL and says so.  The call makes no cudaMalloc / cudaFree /
 * allocating library call.  Every call on a device shares the one array: it carries the last call's
 * partials (nothing in it needs to start at zero -- each call overwrites every word it reads), and two
 * calls that may overlap on the GPU (different streams without an ordering between them) race on it. */
int stemgnn_bn_stats_static(int64_t rows, int64_t feat, const float* x, int64_t ldx, float* mean, float* var,
                            stemgnn_stream_t stream);
int stemgnn_bn_act_forward(int64_t rows, int64_t feat, const float* x, int64_t ldx, const float* mean,
                           const float* invstd, const float* gamma, const float* beta, int relu, float p,
                           const uint64_t* seed, float* out, int64_t ldo, stemgnn_stream_t stream);
int stemgnn_bn_act_backward(int64_t rows, int64_t feat, const float* x, int64_t ldx, const float* gy,
                            int64_t ldg, const float* mean, const float* invstd, const float* gamma,
                            const float* beta, int relu, float p, const uint64_t* seed, int batch_stats,
                            float* sum_g, float* sum_gx, float* gx, int64_t ldgx, void* workspace,
                            size_t workspace_bytes, stemgnn_stream_t stream);
/* The same backward with the batch split over several processes (1-D partition, SURVEY.md 8e "BatchNorm
 * batch statistics all-reduced"): call stemgnn_bn_act_backward with gx = NULL for this rank's sum_g /
 * sum_gx, all-reduce them, then this entry point applies gx with the GLOBAL sums and row count
 * (mean / invstd are the global statistics). */
int stemgnn_bn_act_backward_apply(int64_t rows, int64_t feat, const float* x, int64_t ldx, const float* gy,
                                  int64_t ldg, const float* mean, const float* invstd, const float* gamma,
                                  const float* beta, int relu, float p, const uint64_t* seed, const float* sum_g,
                                  const float* sum_gx, int64_t batch_rows, float* gx, int64_t ldgx,
                                  stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K4: EMA statistics
 * Replaces onehot.sum(1) and einsum('h n d, h n c -> h c d') (model/vq.py:492,497 / 665,670):
 * counts [embed_heads,K] fp32 and sums [embed_heads,K,D] fp32.  Deterministic: rows are
 * stable-sorted by code, each code's rows are added in row order inside fixed-size chunks and
 * the chunk partials are combined in a fixed order.  With embed_heads==1 and heads>1 all heads
 * feed the shared codebook in (row, head) order like '1 (b h) n d'.  x / ind use the same
 * strides as stemgnn_vq_assign.  The all-reduce of [counts|sums] (model/vq.py:494,499) sits
 * between this call and stemgnn_vq_ema_apply. */
typedef struct {
    const float* x; int64_t x_row_stride, x_head_stride;
    const int64_t* ind; int64_t ind_row_stride, ind_head_stride;
    int64_t heads, embed_heads, rows, codes, dim;
    float* counts; float* sums;
} stemgnn_vq_code_stats_args;

int stemgnn_vq_code_stats_workspace(const stemgnn_vq_code_stats_args* args_host, size_t* bytes_host);
int stemgnn_vq_code_stats(const stemgnn_vq_code_stats_args* args_host, void* workspace,
                          size_t workspace_bytes, stemgnn_stream_t stream);

/* lerp_ + laplace smoothing + divide (+ double l2norm for cosine), model/vq.py:495-506 /
 * 668-681, 40-46, 102-104.  Updates cluster_size [Hc,K], embed_avg [Hc,K,D], embed in place. */
int stemgnn_vq_ema_apply(const float* counts, const float* sums, int64_t embed_heads,
                         int64_t codes, int64_t dim, double decay, double eps, int cosine,
                         float* cluster_size, float* embed_avg, float* embed,
                         float* totals_scratch /* [embed_heads] */, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K8: edge-wise inner product
 * Replaces InnerProductDecoder.forward (STEM-GNN/model/encoder.py:353-366; called on the positive
 * and the negative edges by model/pt_model.py:62-63):
 *     out[e] = sum_d z[edge_index[0][e], d] * z[edge_index[1][e], d]      (then sigmoid if asked)
 * edge_index is the reference's int64 [2, E] COO tensor (row 0 at edge_index, row 1 at
 * edge_index + E); z has `feat` columns and row stride `ldz` floats.  The sum over d runs as an
 * fma chain per lane and a fixed shuffle tree: deterministic, within fp32 rounding of torch's sum.
 * An index outside [0, num_nodes) gives NaN for that edge (torch raises a device assert).
 * The backward needs no entry point of its own: it is two stemgnn_aggregate(SUM) calls with
 * edge_weight = upstream gradient (see stemgnn_b200/decoder.py). */
int stemgnn_edge_dot(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes, int64_t feat,
                     const float* z, int64_t ldz, int apply_sigmoid, float* out, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K9: graph augmentations
 * The pretrain step augments the sampled subgraph before it reaches the encoder (reference
 * STEM-GNN/pretrain.py:41-44, torch_geometric.utils 2.3.0 `mask_feature` / `dropout_adj`).
 *
 * dropout_adj: edge e survives iff u[e] >= p and (force_undirected ? row[e] <= col[e] : 1), with
 * u = rand[e] when `rand` is given (the caller's torch.rand(E): same stream as the reference on the
 * same device and seed) or a Philox draw keyed by (*seed, e) otherwise.  Survivors keep their order;
 * with force_undirected the output is [row_kept ++ col_kept ; col_kept ++ row_kept] (2*kept columns),
 * else [row_kept ; col_kept].  edge_id_out (optional, same length as an output row) holds the original
 * edge id of every output edge, which is how edge_attr / edge types follow (`edge_attr[mask]`,
 * `cat([edge_attr, edge_attr])`).  The survivor count is data-dependent, so the call is split like the
 * boolean-mask indexing it replaces: _count leaves the count in device memory (and the tile offsets in
 * the workspace), the caller reads it, allocates, and calls _write with the same arguments and
 * workspace contents. */
int stemgnn_dropout_adj_workspace(int64_t num_edges, size_t* bytes_host);
int stemgnn_dropout_adj_count(const int64_t* edge_index, int64_t num_edges, const float* rand,
                              const uint64_t* seed, float p, int force_undirected, int64_t* kept_out,
                              void* workspace, size_t workspace_bytes, stemgnn_stream_t stream);
int stemgnn_dropout_adj_write(const int64_t* edge_index, int64_t num_edges, const float* rand,
                              const uint64_t* seed, float p, int force_undirected, int64_t kept,
                              int64_t* edge_index_out, int64_t* edge_id_out, const void* workspace,
                              size_t workspace_bytes, stemgnn_stream_t stream);

/* mask_feature: out[r, c] = keep ? x[r, c] : fill (masked_fill semantics: a masked NaN becomes fill)
 * with keep = row_keep[r] & col_keep[c] & (u(r, c) >= p_elem); row_keep / col_keep are torch.bool
 * vectors (1 byte each) or NULL, p_elem = 0 switches the element-wise Philox mask off.  mask_out
 * (optional, [rows, feat] bytes) receives the mask (mode 'all' returns it).  The same call with
 * fill = 0 on the upstream gradient is the backward.  feat % 4 == 0, 16-byte aligned rows. */
int stemgnn_mask_fill(int64_t rows, int64_t feat, const float* x, int64_t ldx, const uint8_t* row_keep,
                      const uint8_t* col_keep, float p_elem, const uint64_t* seed, float fill, float* out,
                      int64_t ldo, uint8_t* mask_out, stemgnn_stream_t stream);

/* out[i, :] = src[ids[i], :] -- `node_text_feat[data.x]`, `edge_text_feat[data.xe]`
 * (STEM-GNN/pretrain.py:33-38) and `edge_attr[mask]` of dropout_adj.  An id outside [0, n_src) gives
 * a NaN row.  feat % 4 == 0, 16-byte aligned rows. */
int stemgnn_gather_rows(const float* src, int64_t lds, int64_t n_src, const int64_t* ids, int64_t n_out,
                        int64_t feat, float* out, int64_t ldo, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K10: neighbour sampling
 * Replaces the sampling step of torch_geometric.loader.NeighborLoader (2.3.0; torch_sparse 0.6.15
 * `neighbor_sample`, replace=False, directed=True) as driven by STEM-GNN/pretrain.py:151-153 and
 * utils/loader.py:11-24.  The graph is the K0 CSR grouped by destination (in-edges per node).
 * nodes = seeds in order (a duplicated seed keeps both positions, the map points at the first);
 * per hop every frontier node takes min(deg, fanout) distinct in-edges (all, in edge order, when
 * fanout < 0 or deg <= fanout; a uniform random subset otherwise, Philox keyed by *seed); unseen
 * sources are appended to n_id in order of first appearance and form the next frontier.
 * Outputs (caller-allocated at the capacities stemgnn_neighbor_sample_bounds returns):
 *   n_id [node_cap] global ids; edge_index_out [2, edge_cap] LOCAL ids, row 0 = source, row 1 =
 *   target (row 1 starts at edge_index_out + edge_cap); e_id [edge_cap] original edge ids;
 *   counts [3 + 2*num_hops] int64 (device): [0] nodes, [1] edges, [2+2h], [3+2h] cumulative nodes /
 *   edges after hop h, last word = error bits (1 seed out of range, 2 / 4 capacity exceeded).
 * No host synchronisation inside the call: stages are launched over host-side bounds, true sizes stay
 * on the device.  node_map [num_nodes] int32 is scratch that persists across calls: initialise it once
 * with stemgnn_node_map_init; every call leaves it initialised again.  fan-outs above 64 (other than
 * -1) return STEMGNN_EUNSUPPORTED. */
typedef struct {
    const int32_t* rowptr; const int32_t* col; const int32_t* perm;   /* CSR grouped by destination */
    int64_t num_nodes, num_edges;
    const int64_t* seeds; int64_t num_seeds;
    int32_t num_hops; int32_t reserved;
    const int32_t* fanout_host;     /* [num_hops] HOST array */
    const uint64_t* seed;           /* device; may be NULL when every fan-out is < 0 */
    int32_t* node_map;
    int64_t node_cap, edge_cap;
    int64_t* n_id; int64_t* edge_index_out; int64_t* e_id; int64_t* counts;
} stemgnn_neighbor_sample_args;

int stemgnn_neighbor_sample_bounds(int64_t num_seeds, int64_t num_nodes, int64_t num_edges,
                                   int32_t num_hops, const int32_t* fanout_host, int64_t* node_cap_host,
                                   int64_t* edge_cap_host, size_t* workspace_bytes_host);
int stemgnn_node_map_init(int32_t* node_map, int64_t num_nodes, stemgnn_stream_t stream);
int stemgnn_neighbor_sample(const stemgnn_neighbor_sample_args* args_host, void* workspace,
                            size_t workspace_bytes, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K11: row-wise l2 normalisation
 * y[r, :] = x[r, :] / max(||x[r, :]||_2, eps)  (F.normalize(p=2, dim=-1), the `l2norm` applied per
 * head in front of the cosine codebook, STEM-GNN/model/vq.py:48-49,891) over `rows` contiguous rows
 * of `dim` floats (dim % 4 == 0); norm_out [rows] (optional) keeps ||x|| for the backward:
 * gx = (g - y * <g, y>) / max(||x||, eps), or g / eps where the norm was clamped. */
int stemgnn_l2norm_forward(int64_t rows, int64_t dim, const float* x, float eps, float* y, float* norm_out,
                           stemgnn_stream_t stream);
int stemgnn_l2norm_backward(int64_t rows, int64_t dim, const float* g, const float* y, const float* norm,
                            float eps, float* gx, stemgnn_stream_t stream);

/* ------------------------------------------------------------------ K12: negative edge sampling
 * Replaces torch_geometric.utils.negative_sampling(pos_edge_index, num_nodes) (2.3.0, sparse method)
 * called at STEM-GNN/model/pt_model.py:60.  `num_candidates` Philox draws (keyed by *seed, a device
 * uint64) over the N(N-1) ordered pairs without self loops; a candidate survives if it is not an edge
 * of edge_index and is the first of its value; survivors are written in candidate order into
 * neg_out [2, capacity] (row 1 at neg_out + capacity), count_out (device int64) = number of survivors
 * (may exceed capacity: only the first `capacity` are stored).  The caller sizes num_candidates like
 * PyG (1.1 * wanted / (1 - E / N(N-1))) and retries with more if the count falls short. */
int stemgnn_negative_sampling_workspace(int64_t num_edges, int64_t num_candidates, size_t* bytes_host);
int stemgnn_negative_sampling(const int64_t* edge_index, int64_t num_edges, int64_t num_nodes,
                              int64_t num_candidates, int64_t capacity, const uint64_t* seed, int64_t* neg_out,
                              int64_t* count_out, void* workspace, size_t workspace_bytes, stemgnn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* STEMGNN_B200_H */

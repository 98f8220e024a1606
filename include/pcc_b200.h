/* pcc_b200.h — C-ABI of the B200-native encode/decode hot path.
 *
 * Drop-in boundary.  The reference has no FFI layer of its own: its hot path
 * calls the Python class surface of MinkowskiEngine / CompressAI, whose native back ends are
 * pybind11 modules (MinkowskiEngineBackend._C, compressai.ans, compressai._CXX).  Every entry
 * point below names the reference call sites (file:line under /root/reference) whose native
 * work it replaces.  INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *   - all functions return 0 on success or a negative PCC_ERR_*; pcc_last_error() gives text
 *   - no exceptions, no torch types; pointers marked [dev] are device pointers owned by the
 *     caller, [host] are host pointers; `stream` is a cudaStream_t passed as void*
 *   - work is enqueued on `stream`; a function synchronises the stream only where it must hand
 *     a count back to the host (documented per function)
 *   - coordinates are int32 rows (b, x, y, z), batch first, |x|,|y|,|z| < 32768, 0 <= b < 65536
 *   - a coordinate key is the uint64  (b << 48) | (x+32768) << 32 | (y+32768) << 16 | (z+32768):
 *     unsigned key order == lexicographic (b,x,y,z) == the reference's canonical order
 *     (utils.py:155-204 sort_tensor / sort_points)
 *   - features are fp32 row-major [N, C]
 */
#ifndef PCC_B200_H
#define PCC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PCC_OK 0
#define PCC_ERR_INVALID_ARG (-1)
#define PCC_ERR_OOM (-2)
#define PCC_ERR_CUDA (-3)
#define PCC_ERR_CAPACITY (-4)

#define PCC_TILE_M 128 /* output rows per kernel-map / convolution tile */

int pcc_version(void);
const char* pcc_last_error(void);
/* SM count, compute capability major*10+minor of the current device. */
int pcc_device_info(int* sm_count, int* cc);

/* ------------------------------------------------------------------ coordinates
 * Replaces ME's coordinate manager (coordinate_map_gpu.cu) behind every ME.SparseTensor(...)
 * construction: model/model.py:122,188; model/blocks.py:30,43,49; model/transforms.py:94,123,260;
 * model/entropy_models.py:356,394,410; utils.py:174,438-445. */

/* coords [dev, n x 4] -> keys [dev, n].  status [dev, 1 int32, may be NULL] is OR-ed with 1 when
 * a coordinate is out of the packable range. */
int pcc_coords_pack(const int32_t* coords, int64_t n, uint64_t* keys, int32_t* status, void* stream);
int pcc_coords_unpack(const uint64_t* keys, int64_t n, int32_t* coords, void* stream);

/* Canonical order: utils.sort_tensor / sort_points (utils.py:155-204), the Sorted* shims
 * (model/entropy_models.py:12-102).  Sorts keys ascending; perm_out[i] = original index of the
 * i-th smallest key (may be NULL).  ws: scratch of pcc_sort_ws_bytes(n) bytes. */
size_t pcc_sort_ws_bytes(int64_t n);
int pcc_sort_keys(const uint64_t* keys_in, int64_t n, uint64_t* keys_out, int32_t* perm_out,
                  void* ws, size_t ws_bytes, void* stream);

/* Open-addressing hash table over unique keys: tkeys [dev, cap] / tvals [dev, cap],
 * cap = pcc_hash_capacity(n) (power of two >= 2n).  value = index of the key in `keys`. */
int64_t pcc_hash_capacity(int64_t n);
int pcc_hash_build(const uint64_t* keys, int64_t n, uint64_t* tkeys, int32_t* tvals, int64_t cap,
                   void* stream);
/* rows[i] = index of q_keys[i] or -1.  Replaces the coordinate lookup inside
 * SparseTensor.features_at_coordinates (model/blocks.py:37,50; model/transforms.py:96,124,262;
 * model/entropy_models.py:364,401) and torch.isin in _prune_coords (model/blocks.py:118-125). */
int pcc_hash_probe(const uint64_t* tkeys, const int32_t* tvals, int64_t cap, const uint64_t* q_keys,
                   int64_t nq, int32_t* rows, void* stream);

/* Stride-2 output map of ME.MinkowskiConvolution(stride=2) (model/transforms.py:49-51,238;
 * model/blocks.py:203; model/entropy_models.py:276,280): unique(floor(c / ts_out) * ts_out),
 * returned sorted.  out_keys capacity >= n.  Synchronises `stream` to return *n_out [host]. */
size_t pcc_stride_ws_bytes(int64_t n);
int pcc_coords_stride(const uint64_t* keys, int64_t n, int32_t ts_out, uint64_t* out_keys,
                      int64_t* n_out, void* ws, size_t ws_bytes, void* stream);

/* Generative output map of ME.MinkowskiGenerativeConvolutionTranspose (model/blocks.py:84;
 * model/entropy_models.py:286,290) and of ME.MinkowskiConvolutionTranspose on a fresh manager
 * (model/entropy_models.py:298,302): unique{ p + d * ts_out : d in kernel region }, returned
 * sorted.  `keys` must be the sorted unique input map hashed in (tkeys, tvals, cap); ts_in is the
 * input tensor stride (ts_out = ts_in / 2).  Each candidate is emitted once by its
 * lexicographically smallest parent, so no deduplication pass is needed.
 * out_keys capacity >= kernel_size^3 * n.  Synchronises `stream` to return *n_out [host]. */
size_t pcc_expand_ws_bytes(int64_t n, int kernel_size);
int pcc_coords_expand(const uint64_t* tkeys, const int32_t* tvals, int64_t cap, const uint64_t* keys,
                      int64_t n, int kernel_size, int32_t ts_in, uint64_t* out_keys, int64_t* n_out,
                      void* ws, size_t ws_bytes, void* stream);

/* Kernel map (ME kernel_map.cuh) of every convolution on the path.  For tile row j (output row
 * o = order ? order[j] : j) and kernel offset k (ME order: first spatial axis fastest; size 3 ->
 * {-1,0,1}^3, size 2 -> {0,1}^3):
 *     nbr[k * nbr_stride + j] = row of (c_o + d_k * step) in the hashed input map, or -1
 * step = +input tensor stride for (strided) convolution, -output tensor stride for transposed
 * convolution.  in_align = input tensor stride: targets not aligned to it are not probed.
 * tapmask[j / PCC_TILE_M] bit k is set iff some row of that tile has a neighbour at offset k.
 * order [dev, n_out] (may be NULL = identity) RECEIVES the tile-row permutation: output rows
 * sorted by (row / seg_rows, neighbour mask), so that rows with the same neighbour pattern share
 * a tile and the convolution skips the offsets a tile does not have (the reference's ME kernel
 * map keeps per-offset pair lists for the same purpose).  seg_rows: power of two >= PCC_TILE_M.
 * n_groups (1 <= n_groups <= K; > 1 needs order): the K offsets are cut into n_groups consecutive
 * ranges of ceil(K / n_groups) offsets and EACH range gets its own tile-row permutation, sorted by
 * that range's bits of the neighbour mask — ragged maps (rows with unrelated neighbour patterns)
 * then still form tiles whose rows agree on a 9-offset range.  order is [n_groups x n_out], tapmask
 * [n_groups x ceil(n_out / PCC_TILE_M)] (group g's masks carry only its own offsets' bits), and
 * nbr[k] is indexed by the tile rows of k's group.  pcc_conv_fwd then runs one pass per group,
 * continuing every output element's FMA chain from the previous pass, so results do not change.
 * n_groups == 0 chooses: the rows are sorted by the whole mask, the share of useful MACs of such
 * tiles is measured on the device, and one group is kept when it is >= 0.90 or the map has fewer
 * than 2 x SM-count tiles; two groups are built down to 0.70 and three below; order / tapmask /
 * tile_order must be sized for 3 groups and the stream is synchronised to read the
 * measurement.  *n_groups_used [host, may be NULL unless n_groups == 0] receives the group count.
 * tile_order [dev, n_groups x ceil(n_out / PCC_TILE_M), may be NULL; needs tapmask] RECEIVES, per
 * group, the tiles sorted by descending number of active offsets: the convolution's CTA b works on
 * tile tile_order[b], so the long tiles start first and the short ones fill the tail of the launch.
 * ws: pcc_kernel_map_ws_bytes(n_out) bytes of scratch, needed only with order. */
size_t pcc_kernel_map_ws_bytes(int64_t n_out);
int pcc_kernel_map(const uint64_t* tkeys, const int32_t* tvals, int64_t cap, const uint64_t* out_keys,
                   int64_t n_out, int kernel_size, int32_t step, int32_t in_align, int32_t* nbr,
                   int64_t nbr_stride, uint32_t* tapmask, int32_t* order, int32_t* tile_order,
                   int64_t seg_rows, int32_t n_groups, int32_t* n_groups_used, void* ws, size_t ws_bytes,
                   void* stream);

/* ------------------------------------------------------------------ sparse convolution
 * Replaces ME convolution_kernel.cu / convolution_transpose_gpu.cu behind ME.MinkowskiConvolution
 * (57 sites), ...GenerativeConvolutionTranspose (model/blocks.py:84; model/entropy_models.py:286,290)
 * and ...ConvolutionTranspose (model/entropy_models.py:298,302), with the element-wise work that
 * follows them on the path fused into the epilogue, applied in this order:
 *   + bias                                   (ME bias add)
 *   activation                               (ME.MinkowskiReLU / MinkowskiLeakyReLU, 32 sites)
 *   * beta + gamma                           (ScaledBlock FiLM, model/blocks.py:37-40)
 *   + residual                               (ScaledBlock skip, model/blocks.py:49-52)
 * Output-stationary: out[order[j]] = sum_k in[nbr[k][j]] @ W[k]; no atomics, fixed summation order.
 * film_rows / res_rows are indexed by OUTPUT row. */
#define PCC_ACT_NONE 0
#define PCC_ACT_RELU 1
#define PCC_ACT_LEAKY 2

#define PCC_MODE_EXACT 0 /* fp32 FFMA, k-major / ci-minor order: bit-identical to the oracle */
#define PCC_MODE_TF32 1  /* tcgen05.mma kind::tf32, fp32 accumulate in TMEM (cin, cout % 32 == 0) */

/* Store the outputs rounded to tf32 (round-to-nearest-away, cvt.rna.tf32.f32).  The tensor core
 * truncates fp32 operands to tf32, which biases every product towards zero; when the producer has
 * already rounded to nearest, the truncation is a no-op and the error stays unbiased. */
#define PCC_CONV_ROUND_OUT_TF32 1

typedef struct pcc_conv_args {
  const float* in;          /* [dev, n_in x cin] */
  int64_t n_in;
  int32_t cin;
  int32_t cout;
  const float* weight;      /* [dev, K x cin x cout]  (ME `kernel` layout) */
  const float* bias;        /* [dev, cout] or NULL */
  const int32_t* nbr;       /* [dev, K x nbr_stride] */
  int64_t nbr_stride;
  const uint32_t* tapmask;  /* [dev, n_groups x ceil(n_out / PCC_TILE_M)] or NULL (= all taps) */
  int32_t K;
  int32_t act;              /* PCC_ACT_* */
  float slope;              /* leaky slope */
  int32_t film_c;           /* 0: no FiLM; cout: per-channel; 1: broadcast single channel */
  const float* film;        /* [dev, n_film x 2*film_c]: beta | gamma */
  const int32_t* film_rows; /* [dev, n_out] row into film, -1 -> zeros; NULL -> identity */
  const float* residual;    /* [dev, n_res x cout] or NULL */
  const int32_t* res_rows;  /* [dev, n_out] row into residual, -1 -> zeros; NULL -> identity */
  float* out;               /* [dev, n_out x out_stride] */
  int64_t n_out;
  int64_t out_stride;       /* elements between output rows (>= cout) */
  int32_t mode;             /* PCC_MODE_* */
  int32_t flags;            /* PCC_CONV_* bit set */
  const float* weight_t;    /* [dev, K x cout x cin] (pcc_weight_transpose); required by PCC_MODE_TF32 */
  const int32_t* order;     /* [dev, n_groups x n_out] tile row -> output row (pcc_kernel_map), NULL = identity */
  int32_t n_groups;         /* offset groups of the kernel map (0 or 1: one); > 1: PCC_MODE_EXACT only */
  int32_t reserved;         /* 0 */
  const int32_t* tile_order; /* [dev, n_groups x tiles] launch order of the tiles (pcc_kernel_map), NULL = identity */
} pcc_conv_args;

int pcc_conv_fwd(const pcc_conv_args* args, void* stream);
/* One-off re-layout of a kernel for the tensor-core path: wt[k][co][ci] = tf32_round(w[k][ci][co]). */
int pcc_weight_transpose(const float* w, int32_t K, int32_t cin, int32_t cout, float* wt, void* stream);

/* Channel 0 of a convolution whose other output channels are never read — the occupancy head
 * (model/blocks.py:142 reads F[:, 0] of occ_predict only) — in factorised form: P[i][k] =
 * <in[i], weight[k][:, 0]> in one streaming pass, then out[order[j]] = bias[0] + sum_k P[nbr[k][j]][k],
 * a 4-byte read per kernel-map pair instead of a cin-channel row.  fp32, but summed over ci before k:
 * not the exact path's FMA order, so callers use it in the tolerance mode only.  Single-group kernel
 * maps.  out [dev, n_out]; ws: pcc_conv_head0_ws_bytes(n_in). */
size_t pcc_conv_head0_ws_bytes(int64_t n_in);
int pcc_conv_head0(const float* in, int64_t n_in, int32_t cin, const float* weight, int32_t cout, const float* bias,
                   const int32_t* nbr, int64_t nbr_stride, const int32_t* order, int64_t n_out, int32_t K,
                   float* out, void* ws, size_t ws_bytes, void* stream);

/* SparseTensor.features_at_coordinates for on-grid queries == gather-or-zero
 * (sites listed at pcc_hash_probe).  out[i] = rows[i] >= 0 ? in[rows[i]] : 0, optionally
 * out[i] = add[i] + that  (Q.F + Q_plus.features_at_coordinates(Q.C), model/transforms.py:96,262). */
int pcc_gather_rows(const float* in, int32_t c, const int32_t* rows, int64_t n, const float* add,
                    float* out, void* stream);

/* ------------------------------------------------------------------ top-k pruning
 * GenerativeUpBlock._topk_prediction + _prune_coords (model/blocks.py:101-150): torch.topk over
 * channel 0 of the occupancy logits, torch.isin, ME.MinkowskiPruning.
 * Radix select on the order-preserving uint32 image of the logit; ties at the threshold are
 * broken towards the smaller row index.  sel_rows [dev, k] receives the selected rows in
 * ascending order.  logits are read at logits[i * stride].  ws: pcc_topk_ws_bytes(m). */
size_t pcc_topk_ws_bytes(int64_t m);
int pcc_topk_select(const float* logits, int64_t stride, int64_t m, int64_t k, int32_t* sel_rows,
                    void* ws, size_t ws_bytes, void* stream);
/* Row gather used by pruning and sorting: out[i] = in[rows[i]] for fp32 rows / uint64 keys. */
int pcc_take_keys(const uint64_t* keys, const int32_t* rows, int64_t n, uint64_t* out, void* stream);

/* ------------------------------------------------------------------ entropy stage (device)
 * CompressAI EntropyBottleneck / GaussianConditional tensor math (model/entropy_models.py:
 * 313,330,352-353,371-372,393,407-408).  Latents come row-major [N, C]; the coder consumes
 * channel-major [C, N] (z.F.t().unsqueeze(0)), so these kernels transpose through shared memory.
 * params [dev, n_params x 2C] = scales | means (h_s output); rows [dev, N] maps latent row ->
 * params row (-1 -> zeros), i.e. gaussian_params.features_at_coordinates(y.C). */
int pcc_gc_quantize_index(const float* y, const float* params, const int32_t* rows, int64_t n,
                          int32_t c, const float* scale_table, int32_t n_scales, float scale_bound,
                          int32_t* symbols, int32_t* indexes, void* stream);
int pcc_gc_build_indexes(const float* params, const int32_t* rows, int64_t n, int32_t c,
                         const float* scale_table, int32_t n_scales, float scale_bound,
                         int32_t* indexes, void* stream);
int pcc_gc_dequantize(const int32_t* symbols, const float* params, const int32_t* rows, int64_t n,
                      int32_t c, float* y_hat, void* stream);
int pcc_gc_likelihood(const float* y, const float* params, const int32_t* rows, int64_t n, int32_t c,
                      float scale_bound, float likelihood_bound, float* y_hat, float* likelihood,
                      void* stream);
int pcc_eb_quantize(const float* z, const float* medians, int64_t n, int32_t c, int32_t* symbols,
                    float* z_hat, void* stream);
int pcc_eb_dequantize(const int32_t* symbols, const float* medians, int64_t n, int32_t c,
                      float* z_hat, void* stream);
/* Factorised prior with filters (3,3,3,3).  eb_params [dev]: per channel 58 floats:
 * softplus(matrix0)[3] bias0[3] tanh(factor0)[3] | m1[9] b1[3] f1[3] | m2 | m3 | m4[3] b4[1]. */
#define PCC_EB_PARAMS_PER_CHANNEL 58
int pcc_eb_likelihood(const float* z, const float* medians, const float* eb_params, int64_t n,
                      int32_t c, float likelihood_bound, float* z_hat, float* likelihood, void* stream);
/* Final colour quantisation (model/model.py:206): clamp(round(F * 255), 0, 255) / 255 and the
 * [x, y, z, r, g, b] float rows of ColorModel.decompress's return value (model/model.py:207). */
int pcc_color_finalize(const uint64_t* keys, const float* feats, int64_t n, int32_t c, float* out,
                       void* stream);

/* ------------------------------------------------------------------ training path
 * ColorModel.forward (model/model.py:51-93) under autograd, loss.backward() (train.py:189-203).
 * Replaces ME's ConvolutionBackward / ConvolutionTransposeBackward and the autograd graph of
 * CompressAI's GaussianConditional.forward(training=True).
 *
 * Data gradient: d in = conv(d out) over the REVERSED kernel map (pcc_kernel_map from the output
 * side: transposed probe for a forward convolution and vice versa) with pcc_weight_transpose'd
 * weights — pcc_conv_fwd, no extra entry point.
 * Weight gradient: dweight[k][ci][co] = sum_j in[nbr[k][j]][ci] * gout[order[j]][co] over the
 * forward kernel map (same nbr / tapmask / order / n_groups as pcc_conv_fwd took).  Deterministic:
 * partial sums of fixed row ranges are added in a fixed order.  ws: pcc_conv_wgrad_ws_bytes(). */
typedef struct pcc_wgrad_args {
  const float* in;          /* [dev, n_in x cin] forward input */
  const float* gout;        /* [dev, n_out x cout] gradient of the convolution's output (before bias) */
  int32_t cin;
  int32_t cout;
  const int32_t* nbr;       /* [dev, K x nbr_stride] */
  int64_t nbr_stride;
  const uint32_t* tapmask;  /* [dev, n_groups x ceil(n_out / PCC_TILE_M)] or NULL */
  const int32_t* order;     /* [dev, n_groups x n_out] or NULL */
  int64_t n_out;
  int32_t K;
  int32_t n_groups;
  float* dweight;           /* [dev, K x cin x cout] receives the gradient (overwritten) */
  void* ws;
  size_t ws_bytes;
  int32_t mode;             /* PCC_MODE_EXACT: fp32 CUDA cores; PCC_MODE_TF32: tcgen05 (tf32 operands, fp32
                             * accumulation in TMEM) when cin, cout % 32 == 0 and cout <= 256, else fp32 */
  int32_t reserved;         /* 0 */
} pcc_wgrad_args;
size_t pcc_conv_wgrad_ws_bytes(int32_t K, int32_t cin, int32_t cout);
int pcc_conv_wgrad(const pcc_wgrad_args* args, void* stream);
/* Backward of pcc_gather_rows: out[rows[i]] += g[i] (rows[i] < 0 skipped; NULL = identity). */
int pcc_scatter_rows_add(const float* g, int32_t c, const int32_t* rows, int64_t n, float* out, void* stream);
/* Bias gradient of every convolution (grad_bias of ME's ConvolutionBackward): out[j] = sum_i in[i][j] over a
 * row-major [n x c] matrix, c <= 1024; fixed summation order (deterministic).  ws: pcc_col_sum_ws_bytes(c). */
size_t pcc_col_sum_ws_bytes(int32_t c);
int pcc_col_sum(const float* in, int64_t n, int32_t c, float* out, void* ws, size_t ws_bytes, void* stream);
/* pcc_col_sum for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels and the same bits as pcc_col_sum, but the block partials live in a file-scope
 * `static __device__` array inside the library instead of a caller workspace: every GPU that loads the
 * library holds its 592 x PCC_COL_SUM_STATIC_MAX_C x 4 = 606,208 bytes until the process exits.
 * The call itself allocates nothing (no cudaMalloc / cudaFree / allocating library routine); the array's
 * device address is looked up once per device (cudaGetSymbolAddress, with that device current).
 * c > PCC_COL_SUM_STATIC_MAX_C is rejected with PCC_ERR_CAPACITY and a message naming the limit.
 * Every call on a device shares the one array: calls whose kernels may overlap on the GPU (different
 * streams without an ordering between them) RACE on it — order them, or use pcc_col_sum with separate
 * workspaces.  Nothing in the array has to start at zero (each call overwrites what it reads). */
#define PCC_COL_SUM_STATIC_MAX_C 256
int pcc_col_sum_static(const float* in, int64_t n, int32_t c, float* out, void* stream);
/* GaussianConditional.forward(training=True) (model/entropy_models.py:330 in train mode): y_noisy =
 * y + U(-0.5, 0.5) is made by the caller; likelihood [n x c] row-major = max(Phi((0.5 - v) / s) -
 * Phi((-0.5 - v) / s), likelihood_bound), v = |y_noisy - mean|, s = max(scale, scale_bound).  The
 * backward recomputes it and applies CompressAI LowerBound's gradient rule to both bounds;
 * g_param_rows [n x 2c] is the gradient of the GATHERED parameter rows (scatter it with
 * pcc_scatter_rows_add). */
int pcc_gc_likelihood_train_fwd(const float* y_noisy, const float* params, const int32_t* rows, int64_t n,
                                int32_t c, float scale_bound, float likelihood_bound, float* likelihood,
                                void* stream);
int pcc_gc_likelihood_train_bwd(const float* y_noisy, const float* params, const int32_t* rows, int64_t n,
                                int32_t c, float scale_bound, float likelihood_bound, const float* g_likelihood,
                                float* g_y, float* g_param_rows, void* stream);

/* ------------------------------------------------------------------ coordinate side channel
 * Replaces ColorModel.gpcc_encode / gpcc_decode (model/model.py:318-358, 360-395): the reference
 * writes the stride-8 latent coordinates to an ASCII PLY, runs the G-PCC `tmc3` binary as a
 * subprocess (lossless geometry, positionQuantizationScale = 1) and reads the file back.  tmc3 is
 * an un-vendored dependency, so the stream here is this library's own lossless code, NOT G-PCC
 * compatible: breadth-first octree over (x, y, z) / stride, Morton child order, one occupancy byte
 * per internal node, `depth` levels (2^depth = side of the cube in stride units), no header.
 * keys [dev]: unique packed keys of batch 0, coordinates >= 0 and multiples of `stride`.
 * out [dev, 4-byte aligned, capacity >= n * depth rounded up to 4]; *nbytes [host] = stream length.
 * Decoding writes the keys in Morton order (callers re-sort with pcc_sort_keys, as the reference
 * re-sorts what tmc3 returns, model/model.py:262-264); cap = capacity of keys_out, *n_out [host].
 * Both synchronise the stream.  ws: pcc_octree_ws_bytes(n) (encode) / (cap) (decode). */
size_t pcc_octree_ws_bytes(int64_t n);
int pcc_octree_encode(const uint64_t* keys, int64_t n, int32_t stride, int32_t depth, uint8_t* out,
                      int64_t out_cap, int64_t* nbytes, void* ws, size_t ws_bytes, void* stream);
int pcc_octree_decode(const uint8_t* bytes, int64_t nbytes, int32_t stride, int32_t depth,
                      uint64_t* keys_out, int64_t cap, int64_t* n_out, void* ws, size_t ws_bytes,
                      void* stream);

/* ------------------------------------------------------------------ entropy coder (host)
 * CompressAI cpp_exts/rans/rans_interface.cpp + ryg_rans rans64.h and cpp_exts/ops/ops.cpp,
 * reached through EntropyBottleneck/GaussianConditional .compress/.decompress/.update
 * (model/entropy_models.py:352-353,372,393,408; model/model.py:30-36).  The coder is the
 * reference's sequential CPU algorithm, kept on the CPU and timed separately (north_star).
 * cdfs [host, n_tables x cdf_stride] int32. */
int64_t pcc_rans_encode_with_indexes(const int32_t* symbols, const int32_t* indexes, int64_t n,
                                     const int32_t* cdfs, int32_t cdf_stride, const int32_t* cdf_sizes,
                                     const int32_t* offsets, int32_t n_tables, uint8_t* out,
                                     int64_t out_cap); /* returns bytes written, <0 on error */
int pcc_rans_decode_with_indexes(const uint8_t* enc, int64_t nbytes, const int32_t* indexes, int64_t n,
                                 const int32_t* cdfs, int32_t cdf_stride, const int32_t* cdf_sizes,
                                 const int32_t* offsets, int32_t n_tables, int32_t* symbols);
int pcc_pmf_to_quantized_cdf(const float* pmf, int32_t n, int32_t precision, uint32_t* cdf);
/* GaussianConditional.update(): tables for `n` scales.  Call with cdf == NULL to get the row
 * stride (*cdf_stride = max pmf length + 2). */
int pcc_gaussian_cdf_tables(const float* scale_table, int32_t n, double tail_mass, int32_t* cdf,
                            int32_t* cdf_stride, int32_t* cdf_length, int32_t* offset);
/* EntropyBottleneck.update(): raw parameters [host] in CompressAI layout; filters (3,3,3,3).
 * matrices[i] / biases[i] / factors[i] are the _matrix{i} / _bias{i} / _factor{i} tensors
 * flattened row-major; quantiles [c x 3]. Call with cdf == NULL to get the row stride. */
int pcc_factorized_cdf_tables(const float* const* matrices, const float* const* biases,
                              const float* const* factors, const float* quantiles, int32_t c,
                              int32_t* cdf, int32_t* cdf_stride, int32_t* cdf_length, int32_t* offset);

#ifdef __cplusplus
}
#endif
#endif /* PCC_B200_H */

/*
 * kgcnn_b200.h -- C-ABI of libkgcnn_b200.so: the B200 (sm_100a) kernels behind kgcnn's ragged
 * message-passing hot path (gather -> per-edge math -> segment aggregation).
 *
 * The reference (Tacitus523/gcnn_keras, pure Python on TensorFlow) has no FFI of its own.  Each entry
 * point below names the reference code (file:line under the reference tree) whose arithmetic it
 * replaces; a host framework (PyTorch binding in gcnn_keras_b200/, TensorFlow custom-op shim in
 * tf_shim/) wraps them behind the unchanged kgcnn layer API.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller; outputs and workspaces are pre-allocated
 *    by the host framework (torch.empty / TF allocate_output);
 *  - every call is asynchronous on `stream`, never synchronises, never allocates, and is re-entrant
 *    (the only process-global state is a monotonically increasing launch counter; the one exception is
 *    kgcnn_colsum_static, which owns a scratch array shared by all of its calls -- see its comment);
 *  - float tensors are fp32 row-major contiguous; indices are int32 after kgcnn_shift_index (the
 *    int64 -> int32 conversion happens once per batch, the reference redoes it in every layer);
 *  - return value: 0 = KGCNN_OK, negative = KGCNN_E_*; no C++ exception crosses the ABI;
 *  - reductions are deterministic: CSR segments, fixed in-segment order, no floating-point atomics.
 */
#ifndef KGCNN_B200_H_
#define KGCNN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* kgcnn_stream_t; /* cudaStream_t */

enum {
  KGCNN_OK = 0,
  KGCNN_E_INVALID = -1,     /* bad argument (null pointer, negative size, unknown enum) */
  KGCNN_E_UNSUPPORTED = -2, /* shape outside what the kernels are instantiated for */
  KGCNN_E_CUDA = -3,        /* a CUDA launch error was pending after the launch */
  KGCNN_E_WORKSPACE = -4,   /* workspace smaller than *_workspace_bytes() */
  KGCNN_E_STATIC_LIMIT = -5 /* problem needs more scratch than a *_static entry point owns */
};

/* segment_ops_by_name names (kgcnn/ops/segment.py:39-46) */
enum { KGCNN_OP_SUM = 0, KGCNN_OP_MEAN = 1, KGCNN_OP_MAX = 2, KGCNN_OP_MIN = 3 };

/* radial basis selector for the fused edge-geometry kernels */
enum { KGCNN_BASIS_NONE = 0, KGCNN_BASIS_GAUSS = 1, KGCNN_BASIS_BESSEL = 2 };

/* activations with f, f', f'' (kgcnn/ops/activ.py:5-107, keras 'tanh', 'linear') */
enum {
  KGCNN_ACT_LINEAR = 0,
  KGCNN_ACT_SHIFTED_SOFTPLUS = 1,
  KGCNN_ACT_SWISH = 2,
  KGCNN_ACT_TANH = 3,
  KGCNN_ACT_RELU = 4,
  KGCNN_ACT_SOFTPLUS2 = 5,
  KGCNN_ACT_SIGMOID = 6,
  KGCNN_ACT_LEAKY_RELU = 7 /* alpha = 0.05 */
};

int kgcnn_version(void);
const char* kgcnn_error_string(int code);
/* number of kernel launches issued through this library since load / since the last reset */
long long kgcnn_launch_count(void);
void kgcnn_launch_count_reset(void);

/* ---------------------------------------------------------------------------------------------
 * a1  partition_row_indexing(sample -> batch)      kgcnn/ops/partition.py:97-161
 *     idx_batch[e,k] = idx[e,k] + node_row_splits[graph_of(e)], graph_of from edge_row_splits
 *     (the reference: repeat(range(B), row_lengths) :29,45 -> gather :146 -> broadcast add :150-155).
 *     Writes int64 [E,K] (same layout as the input; the bit-exact comparand) and/or int32 [K,E]
 *     (column-major: each index column is contiguous, which is what every other kernel consumes).
 * ------------------------------------------------------------------------------------------- */
int kgcnn_shift_index(const int64_t* idx, int64_t E, int K, const int64_t* node_row_splits,
                      const int64_t* edge_row_splits, int64_t B, int64_t* out_i64 /*nullable*/,
                      int32_t* out_i32 /*nullable*/, kgcnn_stream_t stream);

/* value_rowids from row_splits (kgcnn/ops/partition.py:41-45): ids[n] = graph of element n */
int kgcnn_value_rowids(const int64_t* row_splits, int64_t B, int64_t n_values, int32_t* ids,
                       kgcnn_stream_t stream);

/* int64 row_splits -> int32 CSR row pointer (PoolingEmbedding, kgcnn/layers/pooling.py:215-218) */
int kgcnn_cast_i64_i32(const int64_t* src, int64_t n, int32_t* dst, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a20  ragged <-> padded (+ mask)     kgcnn/layers/casting.py:78-84 (RaggedTensor.to_tensor and the ones_like
 *      mask of ChangeTensorType 'mask'), kgcnn/model/force.py:204-211 (boolean-mask gather back to ragged).
 *      Elements are moved as 32-bit words: words_per_row = trailing elements per value row x elem_words
 *      (float32 / int32: elem_words 1, int64: 2).  out[b, j, :] = values[row_splits[b] + j, :] for
 *      j < len(b), else the fill element (fill_word0 [, fill_word1]).  mask (nullable) has the shape of out
 *      (1.0f / integer 1 where a value exists, 0 elsewhere) or, with mask_is_bool, one byte per 1-word element.
 *      Rows longer than max_len are truncated (to_tensor(shape=...) semantics).  kgcnn_unpad is the inverse /
 *      adjoint: values[n, :] = padded[graph_of(n), n - row_splits[graph_of(n)], :].
 * ------------------------------------------------------------------------------------------- */
int kgcnn_pad(const void* values, const int64_t* row_splits, int64_t B, int64_t max_len, int64_t words_per_row,
              uint32_t fill_word0, uint32_t fill_word1, int elem_words, void* out, void* mask /*nullable*/,
              int mask_is_bool, kgcnn_stream_t stream);
int kgcnn_unpad(const void* padded, const int64_t* row_splits, int64_t B, int64_t max_len, int64_t words_per_row,
                int64_t n_values, void* values, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * plan: replaces tf.argsort(stable) + permutation gathers of PoolingLocalEdges
 *       kgcnn/layers/pooling.py:63-68 (and :156-159), once per batch instead of once per layer.
 *  kgcnn_check_sorted : flags[0] = 1 if key is NOT ascending, flags[1] = 1 if any key outside [0,N)
 *                       (flags must be zeroed by the caller; OR-accumulated).  clamp_in_place != 0 also
 *                       rewrites out-of-range keys to the nearest valid id, so that an already enqueued
 *                       (CUDA-graph) step can never read out of bounds; the flags still report it and
 *                       kgcnn_adam_clipnorm's guard skips the update.
 *  kgcnn_plan_build   : rowptr[N+1] = CSR row pointer of key (histogram + scan);
 *                       perm[E] = stable ascending argsort of key (identity when assume_sorted!=0).
 *                       key is read with stride `key_stride` (column of an [E,K] int32 tensor).
 * ------------------------------------------------------------------------------------------- */
int kgcnn_check_sorted(int32_t* key, int key_stride, int64_t E, int64_t N, int32_t* flags, int clamp_in_place,
                       kgcnn_stream_t stream);
size_t kgcnn_plan_workspace_bytes(int64_t E, int64_t N);
int kgcnn_plan_build(const int32_t* key, int key_stride, int64_t E, int64_t N, int assume_sorted,
                     int32_t* rowptr, int32_t* perm /*nullable if assume_sorted*/, void* workspace,
                     size_t workspace_bytes, kgcnn_stream_t stream);
/* out[p] = src[perm[p]*src_stride]  (pre-permuted gather index so the CSR kernels read one array) */
int kgcnn_permute_i32(const int32_t* src, int src_stride, const int32_t* perm, int64_t E,
                      int32_t* out, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a2/a3  GatherEmbedding / GatherEmbeddingSelection   kgcnn/layers/gather.py:83, :228
 *     y[e, 0:F] = x[idx[e*idx_stride], 0:F]; y rows are `y_stride` floats apart so K selections can
 *     be written side by side (concat_axis=2, gather.py:88-90) without a second pass.
 * ------------------------------------------------------------------------------------------- */
int kgcnn_gather_rows(const float* x, int64_t N, int F, const int32_t* idx, int idx_stride, int64_t E,
                      float* y, int64_t y_stride, kgcnn_stream_t stream);
/* same for int64 payloads (atomic numbers: acsf_conv.py:204, :488) */
int kgcnn_gather_rows_i64(const int64_t* x, int64_t N, int F, const int32_t* idx, int idx_stride,
                          int64_t E, int64_t* y, int64_t y_stride, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a4/a5/a7/a13/a14/a18  fused gather * weight -> segment reduce (CSR, deterministic)
 *     y[i, f] = OP_{p in [rowptr[i], rowptr[i+1])}  x[src(p), f] * w[e(p), f or 0]
 *         e(p)   = perm ? perm[p] : p                      (original edge id, stable order)
 *         src(p) = gidx ? gidx[e(p)*gidx_stride] : e(p)    (gidx==NULL: x is the edge tensor itself)
 *         gidx_permuted != 0: gidx is already in CSR order, src(p) = gidx[p*gidx_stride]
 *     n_elements = rowptr[n_out] (number of edges); only used to pick the mapping: short segments
 *     (n_elements <= 64 n_out) get one lane group per row, long ones one warp per row.
 *     w_cols in {0 (no weight), 1 (per-edge scalar), F (per-edge row)}; op MAX/MIN require w_cols==0.
 *     Empty segment -> 0 for every op (kgcnn/layers/pooling.py:71-76: sorted segment op + scatter_nd
 *     zero padding).  MEAN divides by the segment length (kgcnn/ops/segment.py:39-40).
 *     Replaces: tf.gather (gather.py:228) + LazyMultiply (schnet_conv.py:77) + argsort, permutation gathers,
 *     sorted segment op and scatter_nd (pooling.py:63-76); PoolingWeightedLocalEdges (pooling.py:152-172);
 *     PoolingEmbedding (pooling.py:215-218) with rowptr = row_splits.
 * ------------------------------------------------------------------------------------------- */
int kgcnn_gather_mul_segment_reduce(const float* x, int F, const float* w, int w_cols,
                                    const int32_t* rowptr, const int32_t* perm, const int32_t* gidx,
                                    int gidx_stride, int gidx_permuted, int64_t n_out, int64_t n_elements,
                                    int op, float* y, const int32_t* long_rows /*nullable*/,
                                    const int32_t* n_long /*nullable*/, int long_cap, int long_thresh,
                                    kgcnn_stream_t stream);
/* hub rows: stats[0] = max segment length (atomicMax into a zeroed word); kgcnn_long_rows_build appends every
 * row longer than `thresh` to rows[cap] (count zeroed by the caller).  Passing the list to the call above makes
 * the row-parallel kernel skip those rows and reduces each of them with a whole CTA instead. */
int kgcnn_row_stats(const int32_t* rowptr, int64_t n_rows, int32_t* stats, kgcnn_stream_t stream);
int kgcnn_long_rows_build(const int32_t* rowptr, int64_t n_rows, int thresh, int cap, int32_t* rows,
                          int32_t* count, kgcnn_stream_t stream);

/* z[e, f] = a[ia[e], f] * b[ib[e], f]   (reduce_f==0)   or   z[e] = sum_f a[ia[e],f]*b[ib[e],f]
 * ia/ib == NULL means "row e".  VJP of the fused op above with respect to w. */
int kgcnn_gather2_mul(const float* a, const int32_t* ia, int ia_stride, const float* b,
                      const int32_t* ib, int ib_stride, int F, int64_t E, int reduce_f, float* z,
                      kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a19  bias + activation on flat values (kgcnn/layers/modules.py:85; kgcnn/ops/activ.py:5-107)
 *     out[m,u] = g1[m,u] * g2[m,u] * f^(order)(pre[m,u] + bias[u]);  bias, g1, g2 nullable; order 0..3.
 *     order k+1 with one more multiplier is the VJP of order k: closed under differentiation.
 *  kgcnn_colsum: out[u] = sum_m a[m,u] (bias gradient), deterministic two-stage tree.
 * ------------------------------------------------------------------------------------------- */
int kgcnn_bias_act(const float* pre, const float* bias, const float* g1, const float* g2, int act, int order,
                   int64_t M, int U, float* out, kgcnn_stream_t stream);
/* both VJPs of out = g1 * f^(order)(pre + bias) for the cotangent gy in one pass (order 0..2):
 *     d_pre = gy * g1 * f^(order+1)(pre + bias),  d_g1 = gy * f^(order)(pre + bias) */
int kgcnn_bias_act_pair(const float* pre, const float* bias, const float* gy, const float* g1, int act, int order,
                        int64_t M, int U, float* d_pre, float* d_g1, kgcnn_stream_t stream);
size_t kgcnn_colsum_workspace_bytes(int64_t M, int U);
int kgcnn_colsum(const float* a, int64_t M, int U, float* out, void* workspace, size_t workspace_bytes,
                 kgcnn_stream_t stream);
/* kgcnn_colsum_static -- no counterpart in the reference; it exists for the host side of this library.
 *     Same result as kgcnn_colsum (same kernels, same summation order, bit for bit), but the caller brings only the
 *     input and the output: the scratch between the two stages is a 1 MiB array inside the library image.  Meant for
 *     hosts that cannot produce a device workspace -- an op running under a framework that has already claimed all of
 *     HBM for its own arena (cudaMalloc then fails), a C program with a device budget fixed at start-up, a serving
 *     thread that may not call cudaMalloc / cudaFree because they synchronise the device.  It calls no allocator.
 *   Limit: ceil(M / 512) * U <= KGCNN_COLSUM_STATIC_MAX_FLOATS when M > 512 (e.g. M <= 1 048 576 rows at U = 128);
 *     anything larger returns KGCNN_E_STATIC_LIMIT before touching the device -- use kgcnn_colsum there.
 *   Cost: every GPU that loads the library holds the 1 MiB until the process exits.
 *   First call on each device (with that device current) looks up the array's device address; make that call before
 *     any stream capture begins.  Later calls only launch.
 *   The array is shared by ALL calls on a device and keeps the previous call's partial sums.  Nothing needs clearing
 *     (stage 1 overwrites every element stage 2 reads), but two calls that can overlap on the GPU -- different
 *     streams, or different host threads on different streams -- race on it: keep them on one stream or order them
 *     with events.  This is the one entry point of the library that is not re-entrant in that sense. */
#define KGCNN_COLSUM_STATIC_MAX_FLOATS (1 << 18)
int kgcnn_colsum_static(const float* a, int64_t M, int U, float* out, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a19  Dense contractions on the tcgen05 tensor cores (three-piece bf16 split, fp32 TMEM accumulators):
 *   op 0 (NN): out[M,U] = A[M,K] B[K,U] + bias[U] (bias nullable)      forward, kgcnn/layers/modules.py:85
 *   op 1 (NT): out[M,K] = A[M,U] B[K,U]^T                              input gradient
 *   op 2 (TN): out[K,U] = A[M,K]^T B[M,U]                              weight gradient (needs the workspace)
 * ------------------------------------------------------------------------------------------- */
size_t kgcnn_dense_tc_workspace_bytes(int op, int64_t M, int K, int U);
int kgcnn_dense_tc(int op, const float* A, const float* B, int64_t M, int K, int U, const float* bias, float* out,
                   void* workspace, size_t workspace_bytes, kgcnn_stream_t stream);

/* a19  Dense with its activation in the GEMM epilogue (kgcnn/layers/modules.py:61-85, `activation(x W + b)`):
 *   pre[M,U] = X[M,K] W[K,U] + bias[U]  (kept: every derivative of the activation is evaluated at pre),
 *   y[M,U]   = f(pre),  f = enum kgcnn_activation.  One launch instead of GEMM + bias/activation pass. */
int kgcnn_dense_act(const float* X, const float* W, int64_t M, int K, int U, const float* bias, int act, float* pre,
                    float* y, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a19  grouped weight gradients: every  dW_i[K_i,U_i] += X_i[M_i,K_i]^T G_i[M_i,U_i],  db_i[U_i] += colsum(G_i)
 *   of one backward sweep in two launches (tcgen05 bf16x3 + fixed-order reduce).  The products are leaves of the
 *   backward graph (backward of kgcnn/layers/modules.py:85 for every Dense, once through the energy and once
 *   through the force), so the host defers them and submits the list.  X, G, M, K, U, dW, db are HOST arrays of
 *   length n (device pointers inside); db[i] nullable.  Results are ACCUMULATED into dW / db (zero them first);
 *   problems with the same dW must be adjacent in the list.  Deterministic for a fixed list.
 * ------------------------------------------------------------------------------------------- */
size_t kgcnn_dense_tn_grouped_workspace_bytes(int n, const int64_t* M, const int32_t* K, const int32_t* U);
int kgcnn_dense_tn_grouped(int n, const float* const* X, const float* const* G, const int64_t* M, const int32_t* K,
                           const int32_t* U, float* const* dW, float* const* db, void* workspace,
                           size_t workspace_bytes, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * f1  RelationalDense (kgcnn/layers/relational.py:232-234: kernel_per_feature = gather(kernel, relation);
 *   out = batch_dot(x, kernel_per_feature)) as R grouped Dense problems in one launch.  The M rows are SORTED by
 *   relation; offs (DEVICE, int32 [R + 1], offs[0] = 0, offs[R] = M) gives each relation's row range and is read on
 *   the device, so the launch shape depends on (M, R) only.  kernels = [R, K, U] row-major.
 *     op 0 (NN): out[M,U] rows of relation r = A[M,K] rows of r  @ kernels[r]   (+ bias[U], nullable)
 *     op 1 (NT): out[M,K] rows of relation r = A[M,U] rows of r  @ kernels[r]^T         (input gradient)
 *     op 2 (TN): out[R,K,U]; out[r] = A[M,K] rows of r ^T  @ B[M,U] rows of r           (kernel gradient, B = G;
 *                relations without rows give zeros; fixed-order split reduce, needs the workspace)
 * ------------------------------------------------------------------------------------------- */
size_t kgcnn_dense_rel_workspace_bytes(int op, int R, int K, int U);
int kgcnn_dense_rel(int op, const float* A, const float* B, const int32_t* offs, int R, int64_t M, int K, int U,
                    const float* bias, float* out, void* workspace, size_t workspace_bytes, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a9-a12  fused edge geometry: coordinates + index pair -> distance, unit direction, radial basis, cosine cutoff
 *   v = x_i - x_j; d = sqrt(relu(sum v^2) + eps)        NodePosition + NodeDistanceEuclidean  geom.py:14-73,181-187
 *   u = v * divide_no_nan(1, d)                          EdgeDirectionNormalized               geom.py:369-371
 *   basis 1 (gauss):  rbf_k = exp(-gamma (d - offset - k step)^2)      GaussBasisLayer         geom.py:567-570
 *         basis_params (HOST, 3 floats) = (gamma, offset, distance / bins)
 *   basis 2 (bessel): rbf_n = env_p(d / c) sin(f_n d / c)              BesselBasisLayer        geom.py:772-785
 *         basis_params = (1 / c, p = envelope_exponent + 1, unused); frequencies[R] on the device
 *   cos_cutoff > 0:   fc = 0.5 (cos(pi clip(d, -c, c) / c) + 1)        CosCutOffEnvelope       geom.py:834-835
 *   mode 0: values into out_d[E], out_u[E,3], out_rbf[E,R], out_fc[E] (each nullable = not wanted)
 *   mode 1: VJP; cot_d / cot_u / cot_rbf / cot_fc (each nullable) -> out_gv[E,3] = cotangent of v; the caller adds
 *           it to x_i and subtracts it from x_j with the deterministic segment kernel
 *   mode 2: JVP; tangent_xyz[N,3] -> tangents of (d, u, rbf, fc) into the mode-0 outputs
 *   mode 3: VJP of mode 1 with respect to the coordinates (_vjp_vjp, the coordinate-Hessian term of
 *           kgcnn/layers/geom.py:181-190 under two tapes): the cotangents cot_* that entered mode 1 and the cotangent
 *           tangent_xyz[N,3] of its result -> out_gv[E,3] = d/dv sum <cot, D out(v)[h_i - h_j]>  (scattered +i / -j
 *           by the caller like mode 1)
 *   idx_i / idx_j: batch-indexed int32 columns in the ORIGINAL edge order.  Frequencies are constants here (a
 *   trainable Bessel basis keeps its own differentiable path on top of out_d).
 * ------------------------------------------------------------------------------------------- */
int kgcnn_edge_geometry(const float* xyz, const int32_t* idx_i, const int32_t* idx_j, int64_t E, int mode,
                        int basis, int R, const float* basis_params, const float* frequencies, float eps,
                        int no_nan, float cos_cutoff, const float* cot_d, const float* cot_u, const float* cot_rbf,
                        const float* cot_fc, const float* tangent_xyz, float* out_d, float* out_u, float* out_rbf,
                        float* out_fc, float* out_gv, kgcnn_stream_t stream);

/* a10  gradient with respect to the TRAINABLE Bessel frequencies f_n (kgcnn/layers/geom.py:766-770, tf.Variable):
 *   tangent_xyz null : out[n] = sum_e cot_rbf[e,n] * d rbf_n / d f_n                    (loss -> energy -> basis)
 *   tangent_xyz given: out[n] = sum_e cot_rbf[e,n] * t_d[e] * d2 rbf_n / (d d  d f_n),  t_d = (d d_e / d x) . tangent
 *                      (loss -> force = coordinate VJP of the basis; cot_rbf is the cotangent that entered that VJP)
 *   basis_params as for basis 2 of kgcnn_edge_geometry.  Block partials + fixed-order sum (deterministic). */
size_t kgcnn_bessel_frequency_grad_workspace_bytes(int R);
int kgcnn_bessel_frequency_grad(const float* xyz, const int32_t* idx_i, const int32_t* idx_j, int64_t E, int R,
                                const float* basis_params, const float* frequencies, float eps, const float* cot_rbf,
                                const float* tangent_xyz /*nullable*/, float* out, void* workspace,
                                size_t workspace_bytes, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a15/a16  fused ACSF symmetry functions (kgcnn/layers/conv/acsf_conv.py:199-211 and :473-494)
 *   mode 0: G[N, R*m]                  (out)
 *   mode 1: VJP, cot = dG[N, R*m]   -> out = d/dx of the CENTRE atoms [N,3]; the neighbour parts are written per
 *           pair (edge_out [M,3], for the sender) / per triple (triple_out [2,A,3], for j and k) in the ORIGINAL
 *           pair / triple order and are summed by the caller with the deterministic segment kernel
 *   mode 2: JVP, cot = u[N,3]       -> out = J u  [N, R*m]   (what force training needs: parameters are constants)
 *   rel_atom[n]  : relation of atom n as a NEIGHBOUR (reverse_mapping[Z_n], < 0 = element not mapped -> skipped)
 *   tset_atom[n] : nullable; index of atom n as a CENTRE into a (N_el, R, m, .) table (use_target_set)
 *   z_atom[n]    : atomic number; rev_pair[zmax*zmax]: pair relation of (Z_j, Z_k), < 0 = not mapped
 *   rowptr/perm  : CSR of the pairs / triples by centre atom; table: [T][R][m][3] (eta, mu, Rc) / [T][R][m][4]
 *   (eta, zeta, lambda, Rc); eps: added under the square root when add_eps (else 0); uniform_rc != 0 promises
 *   that every table entry has the same Rc (the three cutoffs are then evaluated once per triple).
 * a17  weighted ACSF (kgcnn/layers/conv/wacsf_conv.py:195-203 wACSFRad, :325-356 wACSFAng) runs through the same
 *   kernels with R = 1, tset_atom = Z_i (parameters looked up by the CENTRE atom) and
 *   weight[M | A]  : nullable per-pair / per-triple factor in the ORIGINAL order (g(Z_j) = Z_j, h = Z_j Z_k, or the
 *                    layer's external weights);
 *   mu_table       : nullable [T][R][m]; shifts the three Gaussians of the angular function,
 *                    exp(-eta ((r_ij-mu)^2 + (r_ik-mu)^2 + (r_jk-mu)^2))  (wacsf_conv.py:277-281); null = ACSF G4.
 *   factor_table / factor_index : nullable pair (uniform_rc only).  The m parameter sets of one (T, R) row are
 *                    products of few distinct Gaussians and few distinct angular terms; factor_table [T][R][32][2]
 *                    holds the distinct (eta, mu) in slots 0-15 and the distinct (zeta, lambda) in slots 16-31
 *                    (unused slots 0), factor_index [T][R][m] = gaussian_slot | angular_slot << 8.  With it every
 *                    distinct term is evaluated once per triple (5 + 10 instead of 3 x 50 transcendental chains for
 *                    force_hdnnp2nd.py:49-55); results agree with the unfactorised path to fp32 rounding.
 *   n_triples      : A, the stride between the j and k halves of triple_out (mode 1).
 * ------------------------------------------------------------------------------------------- */
int kgcnn_acsf_g2(const float* xyz, const int32_t* rel_atom, const int32_t* tset_atom, const int32_t* rowptr,
                  const int32_t* perm, const int32_t* send, int send_permuted, const float* table, int R, int m,
                  float eps, int64_t N, const float* weight, int mode, const float* cot, float* out,
                  float* edge_out, kgcnn_stream_t stream);
int kgcnn_acsf_g4(const float* xyz, const int32_t* z_atom, const int32_t* tset_atom, const int32_t* rev_pair,
                  int zmax, const int32_t* rowptr, const int32_t* perm, const int32_t* idx_j, const int32_t* idx_k,
                  const float* table, int R, int m, float eps, float multiplicity, int uniform_rc, int64_t N,
                  const float* mu_table, const float* weight, const float* factor_table,
                  const int32_t* factor_index, int64_t n_triples, int mode, const float* cot, float* out,
                  float* triple_out, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a14  fused PaiNN message (kgcnn/layers/conv/painn_conv.py:97-115) and its derivatives of any order.
 *   M(Gs,Gv,Phi,V,W,R) = sum_e sum_f  Gs[i,f] P1[j,f] W1[e,f]                         term 1 (mask bit 0)
 *                                    + sum_c Gv[i,c,f] P2[j,f] V[j,c,f] W2[e,f]        term 2 (mask bit 1)
 *                                    + sum_c Gv[i,c,f] P3[j,f] W3[e,f] R[e,c]          term 3 (mask bit 2)
 *   i = receiver, j = sender of edge e; Phi = (P1|P2|P3) [N,3F]; W = (W1|W2|W3) [E,3F]; V, Gv [N,3,F]; Gs [N,F];
 *   R [E,3].  F % 4 == 0, all float pointers 16 B aligned.  Operands of inactive terms may be NULL; output blocks
 *   of inactive terms are written as zeros; a NULL output is skipped.
 *   _nodes   : (ds, dv)   = dM/d(Gs, Gv)   the layer's forward; rowptr/perm = CSR of the RECEIVER relation,
 *                                          nbr = sender ids (in CSR order when nbr_permuted)
 *   _senders : (gphi, gvv) = dM/d(Phi, V)  rowptr/perm = CSR of the SENDER relation, nbr = receiver ids
 *   _edges   : (gw, gr)   = dM/d(W, R)     recv/send = the two id columns in the original edge order
 * ------------------------------------------------------------------------------------------- */
int kgcnn_painn_msg_nodes(const float* phi, const float* v, const float* w, const float* r, const int32_t* rowptr,
                          const int32_t* perm /*nullable*/, const int32_t* nbr, int nbr_permuted, int64_t N,
                          int64_t E, int F, int mask, float* ds /*nullable*/, float* dv /*nullable*/,
                          kgcnn_stream_t stream);
int kgcnn_painn_msg_senders(const float* gs, const float* gv, const float* phi, const float* v, const float* w,
                            const float* r, const int32_t* rowptr, const int32_t* perm /*nullable*/,
                            const int32_t* nbr, int nbr_permuted, int64_t N, int64_t E, int F, int mask,
                            float* gphi /*nullable*/, float* gvv /*nullable*/, kgcnn_stream_t stream);
int kgcnn_painn_msg_edges(const float* gs, const float* gv, const float* phi, const float* v, const float* w,
                          const float* r, const int32_t* recv, const int32_t* send, int64_t E, int F, int mask,
                          const float* gw_acc /*nullable: gw = gw_acc + dM/dW, may not alias gw*/,
                          float* gw /*nullable*/, float* gr /*nullable*/, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * optimizer of the force_*.py drivers (force_schnet.py:196-203): Keras Adam with per-variable clipnorm on one
 * flat parameter buffer.  sumsq[var] = sum of squared gradients of variable var (from the segment kernel with
 * rowptr = variable offsets), var_of[i] = variable of element i, *step = device-resident step counter.
 * guard (nullable): n_guard device words; if (guard[i] & guard_mask[i]) != 0 for any i the whole update is
 * skipped (parameters, moments and the step counter stay untouched) -- the plan-validation flags of a replayed
 * CUDA graph, so that a batch that breaks the recorded assumptions cannot corrupt the weights.
 * ------------------------------------------------------------------------------------------- */
int kgcnn_adam_clipnorm(float* params, const float* grads, float* m, float* v, const int32_t* var_of,
                        const float* sumsq, int64_t n, float lr,
                        const float* lr_device /*nullable; DEVICE scalar that overrides lr (schedules under a graph)*/,
                        float beta1, float beta2, float eps, float clipnorm, int32_t* step,
                        const int32_t* guard /*nullable*/,
                        const int32_t* guard_mask /*nullable*/, int n_guard, kgcnn_stream_t stream);
/* kgcnn_guard_fold: *word = 1.0f if (guard[i] & guard_mask[i]) != 0 for any i, else 0.0f.  Data-parallel training:
 * the word sits behind the last variable of the flat gradient buffer and travels through the step's ONE all-reduce
 * (SURVEY 8e), so that after it every rank holds the same non-zero word when ANY rank's batch broke its captured plan;
 * kgcnn_adam_clipnorm then takes that reduced word as its guard (mask = all bits) and all replicas skip the update
 * together instead of one rank skipping while the others apply its garbage gradients.  No reference counterpart
 * (kgcnn/utils/devices.py:11-64 is all the distributed code the reference has). */
int kgcnn_guard_fold(const int32_t* guard, const int32_t* guard_mask, int n_guard, float* word, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Batch graph construction (the step in front of the path).  Two-pass: count -> offsets ->
 * emit; the host reads offsets[n] (one scalar) to size the output.  Integer outputs are bit-exact against the
 * reference's NumPy code.
 *
 * kgcnn_offsets_from_counts: offsets[0] = 0, offsets[i+1] = sum(count[0..i]) in int64.
 *
 * Angle triples -- kgcnn/graph/adj.py:300-385 (get_angle_indices, check_sorted=True ordering) for every graph
 * of the batch at once.  idx_cm = batch-indexed int32 index, column-major [2,M] (kgcnn_shift_index);
 * (rowptr, perm) = CSR of column pos_fix (kgcnn_plan_build; perm NULL when that column is ascending);
 * pos_fix / pos_ij follow adj.py:333-335 (pos_k = 1 - pos_fix).  allow_self with pos_fix != pos_ij is
 * KGCNN_E_UNSUPPORTED.  Outputs in SAMPLE indexing like the reference: ijk[A,3] node triples, ij_k[A,2] edge
 * labels (nullable), angle_row_splits[B+1] (nullable).
 *
 * Range neighbours -- kgcnn/graph/adj.py:466-481 (coordinates_to_distancematrix) + :537-593
 * (define_adjacency_from_distance) as called by SetRange (kgcnn/graph/preprocessor.py:302-304).  xyz is
 * [N,3] fp32 or fp64 (is_f64); distances are evaluated in that dtype in NumPy's operation order.
 * use_max_distance = 0 <=> max_distance=None; max_neighbours < 0 <=> None.  max_nodes >= the largest graph
 * (one distance row per warp is staged in shared memory; KGCNN_E_UNSUPPORTED beyond ~3000 atoms per graph).
 * Ties in the neighbour ranking are broken by index (np.argsort leaves them unspecified).  Outputs: edges[M,2]
 * int64 in sample indexing, row-major order of the adjacency matrix; dist[M] (nullable, coordinate dtype);
 * edge_row_splits[B+1] (nullable).
 * ------------------------------------------------------------------------------------------- */
size_t kgcnn_offsets_workspace_bytes(int64_t n);
int kgcnn_offsets_from_counts(const int32_t* count, int64_t n, int64_t* offsets /*[n+1]*/, void* workspace,
                              size_t workspace_bytes, kgcnn_stream_t stream);
int kgcnn_angle_count(const int32_t* idx_cm, int64_t M, int64_t N, const int32_t* rowptr, const int32_t* perm,
                      int pos_fix, int pos_ij, int allow_multi, int allow_reverse, int allow_self, int32_t* count,
                      kgcnn_stream_t stream);
int kgcnn_angle_emit(const int64_t* idx_sample, const int32_t* idx_cm, int64_t M, int64_t N, const int32_t* rowptr,
                     const int32_t* perm, int pos_fix, int pos_ij, int allow_multi, int allow_reverse,
                     int allow_self, const int64_t* offsets, const int64_t* edge_row_splits, int64_t B,
                     int64_t* ijk, int64_t* ij_k, int64_t* angle_row_splits, kgcnn_stream_t stream);
int kgcnn_range_count(const void* xyz, int is_f64, const int64_t* node_row_splits, int64_t B, int64_t N,
                      int max_nodes, int use_max_distance, double max_distance, int max_neighbours, int exclusive,
                      int self_loops, int32_t* count, kgcnn_stream_t stream);
int kgcnn_range_emit(const void* xyz, int is_f64, const int64_t* node_row_splits, int64_t B, int64_t N,
                     int max_nodes, int use_max_distance, double max_distance, int max_neighbours, int exclusive,
                     int self_loops, const int64_t* offsets, int64_t* edges, void* dist, int64_t* edge_row_splits,
                     kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Segment softmax -- kgcnn/ops/segment.py:4-24 (segment_softmax) behind
 * PoolingLocalEdgesAttention (kgcnn/layers/pooling.py:521-533).  a, out: [E,H] in the ORIGINAL edge order;
 * (rowptr, perm) = CSR of the pooling column (perm NULL when ascending).  out[e,h] =
 * exp(a[e,h] - max_seg) / sum_seg exp(a - max_seg)  (normalize = 0 skips the max shift, segment.py:16).
 * Elements whose id is outside [0, n_out) are not written.
 * ------------------------------------------------------------------------------------------- */
int kgcnn_segment_softmax(const float* a, const int32_t* rowptr, const int32_t* perm /*nullable*/, int64_t n_out,
                          int64_t n_elements, int H, int normalize, float* out, kgcnn_stream_t stream);

/* ---------------------------------------------------------------------------------------------
 * a13  SchNet continuous-filter convolution with the filter network inside the gather-scale-scatter
 *   (kgcnn/layers/conv/schnet_conv.py:73-79 fed by GaussBasisLayer, kgcnn/layers/geom.py:567-570):
 *       rbf_k(d) = exp(-gamma (d - offset - k step)^2),  k < R <= 32
 *       phi(d)   = ssp(rbf(d) W1 + b1) W2 + b2           W1[R,128], W2[128,128]   (units = 128)
 *   No [E, units] tensor is read or written: the basis, both Dense layers (tcgen05, three-piece bf16 split) and the
 *   multiply / segment sum run per tile of 128 edges.  The filter depends on one scalar per edge, so the derivatives
 *   of the layer are forward-mode chains phi', phi'' through the same network; the four modes are closed under the
 *   two differentiations of force training (i = output row of an edge, o = its other end; a, b, c: [*,128]):
 *     mode 0  out[i] = sum a[o] phi                                                   value (CSR of the receivers)
 *     mode 1  out[i] = sum a[o] phi;            edge_out[e] = <a[o] b[i], phi'>          VJP   (CSR of the senders)
 *     mode 2  out[i] = sum a[o] phi + t b[o] phi'                                       JVP   (CSR of the receivers)
 *     mode 3  out[i] = sum t a[o] phi';  edge_out[e] = <a[o] c[i], phi'> + t <a[o] b[i], phi''>   VJP of mode 1
 *   (rowptr, rows, other, eperm): CSR of the output rows; rows[p] / other[p] = output / gathered row of CSR position
 *   p, eperm[p] = original edge id (NULL: identity).  dist, tang (t), edge_out are in the ORIGINAL edge order.
 *   Rows without edges are written as 0.  Summation inside a row is in CSR order (deterministic, no atomics).
 *   KGCNN_E_UNSUPPORTED: units != 128, R > 32, activation other than shifted softplus.
 * ------------------------------------------------------------------------------------------- */
size_t kgcnn_cfconv_fused_workspace_bytes(int64_t E);
/* out[e,k] = scale[e] * (d/dd)^order exp(-gamma (dist[e] - offset - k step)^2), order 0..2, scale nullable: the
 * basis operands of the filter network's weight gradients (kgcnn/layers/geom.py:567-570 and its derivatives). */
int kgcnn_gauss_basis(const float* dist, int64_t E, int R, float gamma, float offset, float step, int order,
                      const float* scale /*nullable*/, float* out, kgcnn_stream_t stream);
int kgcnn_cfconv_fused(int mode, const float* dist, int64_t E, int R, float gamma, float offset, float step,
                       const float* W1, const float* b1 /*nullable*/, const float* W2, const float* b2 /*nullable*/,
                       int units, int act, int static_weights, const int32_t* rowptr, const int32_t* rows,
                       const int32_t* other,
                       const int32_t* eperm /*nullable*/, int64_t n_out, const float* a, const float* b,
                       const float* c, const float* tang, float* out, float* edge_out,
                       float* const* wg /*nullable; HOST array of 8 device pointers*/, void* workspace,
                       size_t workspace_bytes, kgcnn_stream_t stream);
/* static_weights != 0: W1, b1, W2, b2 are model variables that no kernel launched since the last optimizer update
 * writes; the kernel then stages them before it waits for its programmatic launch dependency (csrc/common.cuh).
 * wg (modes 1 and 3 only): the operands of the filter network's weight gradients, produced in the same pass, rows in
 * CSR order of this call:   dW2 += h^T A (+ hd^T Bt),  db2 += colsum A,  dW1 += rbf^T gp (+ rbf1^T gpd),  db1 += colsum gp
 *   wg[0] rbf [E,R]   wg[1] h [E,128]   wg[2] A [E,128] = a[o] b[i] (mode 1) | a[o] c[i] (mode 3)   wg[3] gp [E,128]
 *   mode 3 also:  wg[4] rbf1 [E,R] = t rbf'   wg[5] hd [E,128]   wg[6] Bt [E,128] = a[o] b[i]   wg[7] gpd [E,128]
 * (reverse sweep of kgcnn/layers/conv/schnet_conv.py:74-75 on the edge rows; the pairs go to kgcnn_dense_tn_grouped). */

/* ---------------------------------------------------------------------------------------------
 * 8e  weight-gradient all-reduce over peer memory (the one exchange of the data-parallel step; the reference has no
 *   distributed code, SURVEY F-list): out[i] = scale * sum_r bufs[r][i] with the sum in RANK order (bitwise identical
 *   on every rank).  bufs = HOST array of `world` (<= 16) device pointers to the ranks' flat gradient buffers, all mapped
 *   into this process (CUDA IPC / symmetric memory); the caller orders the launch after a cross-rank barrier and holds
 *   the buffers unchanged until a second barrier.  out must not alias any of the buffers.
 * ------------------------------------------------------------------------------------------- */
int kgcnn_allreduce_peer(const float* const* bufs, int world, int64_t n, float scale, float* out,
                         kgcnn_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* KGCNN_B200_H_ */

/*
 * minpy_cuda.h -- C ABI of libminpy_cuda.so, the sm_100a compute library behind the
 * `cuda` array variant of MinPy (minpy/array_variants/cuda).
 *
 * The reference (dmlc/minpy @2e44927) has no native code of its own: its GPU variant
 * calls into MXNet's NDArray library (minpy/array_variants/mxnet/mxnet_core.py:100-170).
 * Every entry point below therefore replaces one *Python-level* call site of the
 * reference; the file:line each one stands in for is cited beside it.  A maintainer
 * binds these with ctypes; no C++ or torch types cross the ABI.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; mc_last_error() gives
 *     the message of the last failure on the calling thread;
 *   - all array pointers are DEVICE pointers owned by the caller, dense row-major,
 *     float32 unless the parameter name says otherwise (i32 / i64 / u8 / f64);
 *   - work is enqueued on the library's current stream (mc_set_stream); nothing
 *     synchronises unless the name says so (mc_sync, mc_memcpy_d2h);
 *   - `n`, `rows`, `cols`, `outer`, `red`, `inner` are element counts (int64_t).
 */
#ifndef MINPY_CUDA_H_
#define MINPY_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ runtime --- */
/* Device selection: replaces minpy/context.py:55-56 (mxnet.Context construction).  Each device gets
 * its own compute stream on first selection; later calls run on the CURRENT device's stream. */
int mc_init(int device);
int mc_device_count(int* count);
int mc_set_device(int device);
int mc_get_device(int* device);
int mc_sm_count(int* count);
/* Stream used by every later call; NULL selects the legacy default stream. */
int mc_set_stream(void* cuda_stream);
int mc_get_stream(void** cuda_stream);
/* Blocks until the current stream drains: replaces NDArray.wait_to_read (minpy/array.py:480-487). */
int mc_sync(void);
const char* mc_last_error(void);
/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
int mc_launch_count(uint64_t* count);

/* Side stream: library calls between mc_side_begin and mc_side_end run on a second stream forked from the
 * compute stream (and, inside mc_graph_begin..end, become a parallel branch of the graph); mc_side_join makes
 * the compute stream wait for them (mc_sync and mc_graph_end join implicitly).  Used for the weight-gradient
 * contractions of the backward pass, which only the update rule reads (nn/solver.py:190-195) while the
 * data-gradient chain continues (minpy/tape.py:223-234 replays VJPs strictly in sequence).  The caller keeps
 * the operands of side work alive until the join. */
int mc_side_begin(void);
int mc_side_end(void);
int mc_side_join(void);

/* Stream-ordered device allocator (cudaMallocAsync pool): replaces NDArray allocation
 * inside mxnet.ndarray.array / mx.nd.* results (minpy/array.py:388-389). */
int mc_alloc(void** ptr, size_t nbytes);
int mc_free(void* ptr);
int mc_host_alloc(void** ptr, size_t nbytes);   /* pinned host memory */
int mc_host_free(void* ptr);
int mc_mem_info(size_t* free_bytes, size_t* total_bytes);
/* Copies: replace Array._synchronize_data (minpy/array.py:375-390). d2h synchronises. */
int mc_memcpy_h2d(void* dst_dev, const void* src_host, size_t nbytes);
int mc_memcpy_d2h(void* dst_host, const void* src_dev, size_t nbytes);
int mc_memcpy_d2h_async(void* dst_host, const void* src_dev, size_t nbytes);
int mc_memcpy_d2d(void* dst_dev, const void* src_dev, size_t nbytes);
/* Copy between two devices of this process: replaces NDArray.copyto(ctx) / as_in_context for operands
 * that live on another device than the `with gpu(i):` scope (minpy/context.py:92-98, array.py:375-390).
 * Drains the source device's stream, then copies on the current device's stream. */
int mc_memcpy_peer(void* dst_dev, int dst_device, const void* src_dev, int src_device, size_t nbytes);
int mc_memset(void* dst_dev, int byte, size_t nbytes);
/* Input prefetch beside the compute stream (minpy/nn/io.py:238-258 sliced on the host and copied
 * every step; here batch i+1 is copied on a second stream while step i runs).
 * acquire: the compute stream waits for all prefetch copies enqueued so far;
 * release: later prefetch copies wait for the compute work enqueued so far. */
int mc_prefetch_h2d(void* dst_dev, const void* src_host, size_t nbytes);
int mc_prefetch_acquire(void);
int mc_prefetch_release(void);
/* Deferred read of a small device value (<= 64 bytes) -- replaces the blocking loss read of
 * minpy/nn/solver.py:187 (`loss.asnumpy()` every step): `begin` enqueues a device->host copy into a
 * pinned ring slot on a third stream, ordered after the compute work enqueued so far; `wait` blocks
 * on that copy alone (not on later compute work) and copies the value out.  A ticket is valid for
 * the next 256 begins; `wait` on an older ticket returns status 4. */
int mc_readback_begin(const void* src_dev, size_t nbytes, uint64_t* ticket);
int mc_readback_wait(uint64_t ticket, void* dst_host, size_t nbytes);
/* CUDA events on the current stream (bench timing). */
int mc_event_create(void** ev);
int mc_event_record(void* ev);
int mc_event_sync(void* ev);
int mc_event_elapsed_ms(void* start, void* stop, float* ms);
int mc_event_destroy(void* ev);
/* CUDA-graph capture of a step on the current stream (must be a non-default stream). */
int mc_graph_begin(void);
int mc_graph_end(void** graph_exec);
int mc_graph_launch(void* graph_exec);
int mc_graph_destroy(void* graph_exec);
int mc_graph_abort(void);
int mc_graph_capturing(int* flag);

/* ------------------------------------------------------------- elementwise --- */
/* Unary ufuncs: numpy_core.py:228-230,267-276,303 / mxnet_core.py:116-118,150-159. */
enum mc_unary_op {
  MC_NEG = 0, MC_ABS, MC_SIGN, MC_ROUND, MC_CEIL, MC_FLOOR, MC_SQRT, MC_EXP, MC_LOG,
  MC_SIN, MC_COS, MC_TANH, MC_SIGMOID, MC_SQUARE, MC_RECIP, MC_RELU, MC_COPY, MC_COSH,
  MC_LOGICAL_NOT, MC_UNARY_COUNT
};
int mc_unary(int op, const float* x, float* y, int64_t n);

/* Binary ufuncs and comparisons (0/1 float results, as the MXNet variant gave):
 * numpy_core.py:240-291; comparisons minpy/array.py:73-89. */
enum mc_binary_op {
  MC_ADD = 0, MC_SUB, MC_MUL, MC_DIV, MC_POW, MC_MAX, MC_MIN, MC_MOD,
  MC_EQ, MC_NE, MC_LT, MC_GT, MC_LE, MC_GE, MC_BINARY_COUNT
};
int mc_binary(int op, const float* a, const float* b, float* out, int64_t n);
/* scalar_left != 0 computes op(s, a[i]) instead of op(a[i], s). */
int mc_binary_scalar(int op, const float* a, float s, int scalar_left, float* out, int64_t n);
/* out(rows, cols) = op(a(rows, cols), v) with v a row vector (cols,) [vec_is_col == 0] or a
 * column vector (rows,) [vec_is_col != 0]; vec_left puts v on the left of op.
 * The (N,D) o (D,) bias / batchnorm-scale pattern of nn/layers.py:27,107. */
int mc_binary_vec(int op, const float* a, const float* v, int vec_is_col, int vec_left,
                  float* out, int64_t rows, int64_t cols);
/* General NumPy broadcasting, up to MC_MAX_DIMS dims; strides in elements, 0 on a
 * broadcast axis. */
#define MC_MAX_DIMS 6
int mc_binary_bcast(int op, const float* a, const float* b, float* out, int ndim,
                    const int64_t* out_shape, const int64_t* a_strides, const int64_t* b_strides);
int mc_fill(float* dst, float value, int64_t n);
int mc_arange(float* dst, float start, float step, int64_t n);
/* dtype casts; codes: 0 f32, 1 i32, 2 i64, 3 f64, 4 u8. */
enum mc_dtype { MC_F32 = 0, MC_I32, MC_I64, MC_F64, MC_U8 };
int mc_cast(const void* src, int src_dtype, void* dst, int dst_dtype, int64_t n);
/* Multi-tensor float32 copy in one launch per 48 tensors: packs the tape's gradients into the
 * flat all-reduce bucket inserted between nn/solver.py:184 and :190 (srcs/dsts/ns are HOST
 * arrays of `count` device pointers / element counts). */
int mc_multi_copy(const void* const* srcs, void* const* dsts, const int64_t* ns, int count);

/* Fused chain of elementwise VJPs -- the rewrite of the per-record loop of
 * minpy/tape.py:223-234: out[i] = (acc ? acc[i] : 0) + g[i] * prod_k f_k(a_k, b_k, s_k).
 * Each step multiplies the running gradient by one primitive's local derivative. */
enum mc_vjp_op {
  MC_VJP_MUL_A = 0,     /* g * a                    multiply, exp (a = ans)            */
  MC_VJP_MUL_S,         /* g * s                    scalar multiply                    */
  MC_VJP_NEG,           /* -g                       negative / subtract rhs            */
  MC_VJP_DIV_A,         /* g / a                    divide lhs, log (a = x)            */
  MC_VJP_DIV_S,         /* g / s                                                        */
  MC_VJP_TANH_ANS,      /* g * (1 - a^2)            tanh (a = ans)                     */
  MC_VJP_SQRT_ANS,      /* g * 0.5 / a              sqrt (a = ans)                     */
  MC_VJP_SIN_X,         /* g * cos(a)                                                   */
  MC_VJP_COS_X,         /* -g * sin(a)                                                  */
  MC_VJP_ABS_X,         /* g * sign(a)                                                  */
  MC_VJP_SIGMOID_ANS,   /* g * a * (1 - a)                                              */
  MC_VJP_EQ_AB,         /* g * (a == b)             maximum / minimum / relu           */
  MC_VJP_EQ_AS,         /* g * (a == s)                                                 */
  MC_VJP_DIV_Y,         /* -g * a / (b * b)         divide rhs (a = x, b = y)          */
  MC_VJP_S_DIV_Y,       /* -g * s / (a * a)         scalar / array rhs (a = y)         */
  MC_VJP_POW_XS,        /* g * s * a^(s-1)          power, scalar exponent (a = x)     */
  MC_VJP_POW_SY,        /* g * log(s) * a           power, scalar base (a = ans)       */
  MC_VJP_POW_X,         /* g * b * a^(b-1)          power lhs (a = x, b = y)           */
  MC_VJP_POW_Y,         /* g * log(a) * b           power rhs (a = x, b = ans)         */
  MC_VJP_RECIP_ANS,     /* -g * a * a               1/x (a = ans)                      */
  MC_VJP_GT0_A,         /* g * (a > 0)                                                  */
  MC_VJP_GE0_A,         /* g * (a >= 0)             layers.relu: maximum(0,x) passes g at x == 0 */
  MC_VJP_SIGNPOS_A,     /* g * !signbit(a)          relu from its signed-zero OUTPUT (mc_gemm relu mode 2) */
  MC_VJP_COUNT
};
/* Operand addressing inside a step: how element i of the output maps to a / b. */
enum mc_vjp_mode { MC_OPD_FULL = 0, MC_OPD_ROW = 1 /* a[i % inner] */, MC_OPD_COL = 2 /* a[i / inner] */,
                   MC_OPD_ONE = 3 /* a[0] */ };
typedef struct mc_vjp_step {
  int32_t op;        /* mc_vjp_op */
  int32_t a_mode;    /* mc_vjp_mode */
  int32_t b_mode;
  float scalar;
  const float* a;
  const float* b;
  int64_t inner;     /* for ROW / COL modes */
} mc_vjp_step;
#define MC_VJP_MAX_STEPS 12
/* g may be NULL (then g_scalar is the incoming gradient); g_mode addresses g like an operand. */
int mc_vjp_chain(const float* g, int g_mode, int64_t g_inner, float g_scalar, const float* acc,
                 const mc_vjp_step* steps, int nsteps, float* out, int64_t n);

/* -------------------------------------------------------------- reductions --- */
/* x viewed as (outer, red, inner); the middle axis is reduced: out is (outer, inner).
 * np.sum/max/min with axis/keepdims: numpy_core.py:233-237, mxnet_core.py:120-122;
 * `_unbroadcast` sums: numpy_core.py:119-146. */
enum mc_reduce_op { MC_RED_SUM = 0, MC_RED_MAX, MC_RED_MIN, MC_RED_COUNT };
int mc_reduce(int op, const float* x, float* out, int64_t outer, int64_t red, int64_t inner);
/* np.argmax (first maximal index, int64, bit-exact): NumPy-only in the reference
 * (docs/feature/op_stat.rst:113-119); used by nn/solver.py:231. */
int mc_argmax(const float* x, int64_t* out_i64, int64_t outer, int64_t red, int64_t inner);
/* out(outer, red, inner) = g(outer, inner) broadcast over red: VJP of sum
 * (numpy_core.py:150-162 `_sum_grad`). scale multiplies. */
int mc_bcast_mid(const float* g, float* out, float scale, int64_t outer, int64_t red, int64_t inner);
/* VJP of max/min: out = g_b * (x == ans_b), ties all receive g
 * (numpy_core.py:66-70 `_chooser_grad`; tests/unittest/test_autograd.py:107-130). */
int mc_select_grad(const float* x, const float* ans, const float* g, float* out,
                   int64_t outer, int64_t red, int64_t inner);

/* ----------------------------------------------------- softmax / losses ------ */
/* Row softmax: `_softmax_output` numpy_core.py:164-168 (4 NDArray ops in mxnet_core.py:85-90). */
int mc_softmax_rows(const float* x, float* out, int64_t rows, int64_t cols);
int mc_logsumexp_rows(const float* x, float* out, int64_t rows, int64_t cols);
/* One-hot in place: `_onehot_encode` numpy_core.py:47-63 (out[i, idx[i]] = 1). */
int mc_onehot(const int32_t* idx_i32, float* out, int64_t rows, int64_t cols);
/* Fused softmax_loss of nn/layers.py:176-198: probs = softmax(x);
 * loss[0] = -sum_i log(probs[i, y_i]) / rows (labels) or -sum(log(p) * onehot)/rows.
 * Exactly one of labels_i32 / onehot is non-NULL. */
int mc_softmax_ce_fwd(const float* x, const int32_t* labels_i32, const float* onehot,
                      float* probs, float* loss, int64_t rows, int64_t cols);
/* softmax_loss forward AND its gradient in one call: probs (may be NULL), dx = (probs - y) / rows (may be
 * NULL; the VJP ignores the upstream gradient, numpy_core.py:170-175, so it is known at forward time) and
 * loss[0].  A classifier head (integer labels, cols <= 32) is ONE launch (warp per row, last block sums). */
int mc_softmax_ce_fused(const float* x, const int32_t* labels_i32, const float* onehot,
                        float* probs, float* dx, float* loss, int64_t rows, int64_t cols);
/* mc_softmax_ce_fused for callers that cannot allocate.  NO counterpart in the reference: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc.
 * Same results as mc_softmax_ce_fused with integer labels (probs or dx may be NULL, not both), but the row losses and
 * the ticket counter live in a file-scope `static __device__` array of the library (257 KB per GPU, held until
 * exit) instead of the scratch arena: the call makes no cudaMalloc / cudaFree / allocating library call.
 * Limits: cols <= 32, rows <= MC_SOFTMAX_CE_STATIC_MAX_ROWS; anything larger is rejected (status 2, mc_last_error
 * says which limit).  All calls share the one array: calls that overlap on the GPU (different streams, devices are
 * separate) race; calls on the library's stream are ordered and safe. */
#define MC_SOFTMAX_CE_STATIC_MAX_ROWS 65536
int mc_softmax_ce_fused_static(const float* x, const int32_t* labels_i32, float* probs, float* dx, float* loss,
                               int64_t rows, int64_t cols);
/* The classifier head as one launch: logits = h (rows, K) . W (K, C) + bias, C <= 16; loss[0] = softmax_loss(logits,
 * labels) (nn/layers.py:27, 201-221); dlogits = (softmax - onehot) / rows (numpy_core.py:170-175); dh = dlogits . W^T
 * (the VJP of the affine with respect to its input; dh / logits may be NULL).  FP32 arithmetic. */
int mc_head_softmax_loss(const float* h, const float* W, const float* bias, const int32_t* labels_i32, float* logits,
                         float* dlogits, float* dh, float* loss, int64_t rows, int64_t K, int64_t C);
/* temporal_softmax_loss of nn/layers.py:444-480 over rows = N*T score vectors of `cols` classes:
 * loss[0] = -sum_r mask[r] * log softmax(x[r])[y[r]] / batch   and   dx = mask * (softmax(x) - onehot(y)) / batch
 * (the derivative of that loss; the VJP multiplies it by the upstream scalar).  mask may be NULL (all ones). */
int mc_temporal_softmax_loss(const float* x, const int32_t* y_i32, const float* mask, float* dx, float* loss,
                             int64_t rows, int64_t cols, int64_t batch);
/* dx = (probs - y) / rows: `_softmax_output_grad` numpy_core.py:170-175 (ignores upstream g). */
int mc_softmax_ce_bwd(const float* probs, const int32_t* labels_i32, const float* onehot,
                      float* dx, int64_t rows, int64_t cols);

/* ---------------------------------------------------------------- indexing --- */
/* dst (dense, shape[]) = src[offset + sum_k idx_k * strides[k]]: basic slices, strided
 * slices and axis permutations of `_minpy_getitem` / np.transpose
 * (numpy_core.py:19-21; minpy/array.py:432-443). */
int mc_strided_copy(const float* src, float* dst, int ndim, const int64_t* shape,
                    const int64_t* src_strides, int64_t src_offset);
/* dst[offset + sum idx*strides] += src (dense): VJP of a basic slice
 * (`_minpy_getitem_grad`, numpy_core.py:24-28, np.add.at semantics). */
int mc_strided_add(float* dst, const float* src, int ndim, const int64_t* shape,
                   const int64_t* dst_strides, int64_t dst_offset);
/* out[i] = x[i, idx[i]]  (probs[np.arange(N), y], nn/layers.py:478) and its scatter-add VJP. */
int mc_pick_cols(const float* x, const int32_t* idx_i32, float* out, int64_t rows, int64_t cols);
int mc_pick_cols_grad(const float* g, const int32_t* idx_i32, float* dx_zeroed, int64_t rows, int64_t cols);
/* out[i, :] = x[idx[i], :] integer-array gather on axis 0, and np.add.at scatter VJP
 * (duplicates accumulate). */
int mc_take_rows(const float* x, const int32_t* idx_i32, float* out, int64_t nidx, int64_t row_elems);
int mc_take_rows_grad(const float* g, const int32_t* idx_i32, float* dx_zeroed, int64_t nidx, int64_t row_elems);
/* concatenation along the middle axis of (outer, na + nb, inner): np.append / concatenate
 * (numpy_core.py:296-302). */
int mc_concat_mid(const float* a, const float* b, float* out, int64_t outer, int64_t na, int64_t nb, int64_t inner);

/* -------------------------------------------------------------------- GEMM --- */
/* C(M,N) = op(A) . op(B) [+ bias(N,)] [relu] [+ C if accumulate]; row-major, leading
 * dimensions in elements.  np.dot and its two VJPs: numpy_core.py:224-225,
 * mxnet_core.py:112-114; bias/relu epilogue = nn/layers.py:27,41.
 * relu: 0 none, 1 max(0, v), 2 the same values with the sign of the pre-activation kept in the zero's sign
 * bit (-0.0f where v < 0): the VJP of layers.relu passes the gradient where x >= 0 INCLUDING x == 0
 * (numpy_core.py:284-286), and a fused affine+relu never stores x.
 * precision: 0 = FP32 CUDA-core (exact-accumulate cross-check), 1 = TF32 tcgen05/TMEM
 * fed by TMA (FP32 accumulate). Falls back from 1 to 0 only for shapes TMA cannot
 * describe (row pitch not a multiple of 16 bytes). */
enum mc_gemm_precision { MC_GEMM_FP32 = 0, MC_GEMM_TF32 = 1 };
int mc_gemm(int trans_a, int trans_b, int64_t M, int64_t N, int64_t K,
            const float* A, int64_t lda, const float* B, int64_t ldb,
            float* C, int64_t ldc, const float* bias, int relu, int accumulate, int precision);
/* Which path the last mc_gemm call on this thread took (0 fp32, 1 tf32 tcgen05). */
int mc_gemm_last_path(int* path);

/* ------------------------------------------------- data-parallel exchange ---- */
/* The one collective of the path (no reference counterpart: the reference has no
 * multi-device training loop, minpy/context.py:92-98): in-place all-reduce of the flat fp32
 * gradient buffer between the tape's backward pass and the update rules (the gap between
 * minpy/nn/solver.py:184 and :190), NCCL over NVLink/NVSwitch on the LIBRARY stream, so that it
 * is captured into the step graph.  One process per GPU: rank 0 calls mc_comm_unique_id, the
 * 128 bytes travel through the launcher's rendezvous, every rank calls mc_comm_init. */
int mc_comm_unique_id(void* id128);
int mc_comm_init(int rank, int world, const void* id128);
int mc_comm_info(int* rank, int* world, int* nccl_version);
int mc_allreduce_avg(float* buf, int64_t n);
int mc_allreduce_sum(float* buf, int64_t n);
int mc_broadcast(float* buf, int64_t n, int root);
int mc_comm_destroy(void);
/* Peer-memory window for the gradient bucket (one node, 2..8 ranks).  Every rank calls mc_p2p_create
 * (cudaMalloc + two IPC handles, 128 bytes), the handles of all ranks travel through the rendezvous
 * (world x 128 bytes, rank-major) into mc_p2p_connect.  From then on mc_allreduce_avg on a range inside
 * the window runs the library's own two-shot all-reduce over NVLink peer loads (csrc/comm.cu) instead
 * of ncclAllReduce; ranges elsewhere still go to NCCL.  mc_p2p_info: the window (NULL while inactive)
 * and whether a cross-GPU barrier ever timed out (20 s; the results are then invalid). */
int mc_p2p_create(int rank, int world, size_t bytes, void* handles128);
int mc_p2p_connect(const void* all_handles);
int mc_p2p_info(void** ptr, size_t* bytes, int* timed_out);

/* ------------------------------------------------------------ nn.layers ------ */
/* layers.batchnorm train mode, nn/layers.py:93-110: biased variance, returns updated
 * running stats (momentum*r + (1-momentum)*stat). save_mean/save_invstd feed the VJP. */
int mc_batchnorm_train_fwd(const float* x, const float* gamma, const float* beta,
                           const float* running_mean, const float* running_var,
                           float* out, float* new_running_mean, float* new_running_var,
                           float* save_mean, float* save_invstd,
                           int64_t N, int64_t D, float eps, float momentum);
int mc_batchnorm_train_bwd(const float* x, const float* dy, const float* gamma,
                           const float* save_mean, const float* save_invstd,
                           float* dx, float* dgamma, float* dbeta, int64_t N, int64_t D);
/* The hidden-layer block of examples/nn/mlp_bn_dropout.py:41-54 -- affine, relu, batchnorm, dropout -- in
 * TWO forward and ONE elementwise backward launch (plus the two VJP GEMMs):
 *   forward : mc_gemm(..., relu = 2)  then  mc_bn_dropout_fwd  = batchnorm (train) with the dropout mask
 *             applied at the store (mask of element i = lane i & 3 of Philox(seed)(offset + i / 4), as mc_dropout);
 *   backward: mc_bn_dropout_relu_bwd  = dropout VJP (mask regenerated while dy is loaded, `dropout` != 0),
 *             batchnorm VJP (nn/layers.py:93-110 differentiated), and -- `relu_signed` != 0, x produced by
 *             mc_gemm relu mode 2 -- the relu VJP (numpy_core.py:284-286) plus the column sums of the result,
 *             i.e. the bias gradient of the affine layer (nn/layers.py:27), into dbias.
 * Small batches only (N <= 1024, N*D <= 2^20, D % 8 == 0: mc_bn_fused_supported); otherwise the callers
 * compose mc_batchnorm_train_* / mc_dropout / mc_vjp_chain / mc_reduce. */
int mc_bn_fused_supported(int64_t N, int64_t D, int* ok);
int mc_bn_dropout_fwd(const float* x, const float* gamma, const float* beta,
                      const float* running_mean, const float* running_var,
                      float* out, float* new_running_mean, float* new_running_var,
                      float* save_mean, float* save_invstd,
                      int64_t N, int64_t D, float eps, float momentum, float p, uint64_t seed, uint64_t offset);
int mc_bn_dropout_relu_bwd(const float* x, const float* dy, const float* gamma,
                           const float* save_mean, const float* save_invstd,
                           float* dx, float* dgamma, float* dbeta, float* dbias, int64_t N, int64_t D,
                           int dropout, float p, uint64_t seed, uint64_t offset, int relu_signed);
/* test mode, nn/layers.py:111-113. */
int mc_batchnorm_test_fwd(const float* x, const float* gamma, const float* beta,
                          const float* running_mean, const float* running_var,
                          float* out, int64_t N, int64_t D, float eps);
/* layers.dropout train mode, nn/layers.py:141-144: keep where u > p, NOT rescaled.
 * The Philox stream is keyed by (seed, offset); the same pair regenerates the mask in bwd. */
int mc_dropout(const float* x, float* out, int64_t n, float p, uint64_t seed, uint64_t offset);
/* lstm_step of nn/layers.py:353-380 after the two GEMMs: a is (N, 4H) pre-activations. */
int mc_lstm_cell_fwd(const float* a, const float* prev_c, float* next_h, float* next_c,
                     float* gates /* (N,4H) saved i,f,o,g */, int64_t N, int64_t H);
int mc_lstm_cell_bwd(const float* gates, const float* prev_c, const float* next_c,
                     const float* dnext_h, const float* dnext_c /* may be NULL */,
                     float* da /* (N,4H) */, float* dprev_c, int64_t N, int64_t H);
/* lstm_temporal of nn/layers.py:383-420 on TIME-MAJOR buffers, the whole loop over t in one call
 * (the reference issues 26 primitive dispatches + a slice + an append per step from Python).
 *   fwd: A (T,N,4H) holds x_t.Wx + b on entry; hbuf / cbuf (T+1,N,H) hold h0 / c0 in slot 0 and receive
 *        h_t / c_t in slot t+1; gates (T,N,4H) receives i,f,o,g.  Per step: the K-sliced GEMM
 *        h_{t-1}.Wh leaves raw partials, the gate kernel adds them to A[t] while it reads it.
 *   bwd: dH (T+1,N,H) holds the upstream gradient of h_t in slot t+1 (slot 0 zero) on entry; on exit
 *        slot 0 is dL/dh0 and dA (T,N,4H) the pre-activation gradients (slots 1..T of dH are scratch).
 * precision as mc_gemm (MC_GEMM_FP32 runs the exact-accumulate GEMM per step). */
int mc_lstm_seq_fwd(float* A, const float* Wh, float* hbuf, float* cbuf, float* gates,
                    int64_t T, int64_t N, int64_t H, int precision);
int mc_lstm_seq_bwd(const float* gates, const float* cbuf, const float* Wh, float* dH, float* dA,
                    int64_t T, int64_t N, int64_t H, int precision);

/* --------------------------------------------------------------- nn.optim ---- */
/* Update rules of nn/optim.py:33-141; outputs may alias inputs. */
int mc_sgd(const float* w, const float* dw, float* w_out, float lr, int64_t n);
int mc_sgd_momentum(const float* w, const float* dw, const float* v, float* w_out, float* v_out,
                    float lr, float momentum, int64_t n);
int mc_rmsprop(const float* w, const float* dw, const float* cache, float* w_out, float* cache_out,
               float lr, float decay, float eps, int64_t n);
int mc_adam(const float* w, const float* dw, const float* m, const float* v,
            float* w_out, float* m_out, float* v_out,
            float lr, float beta1, float beta2, float eps, int64_t t, int64_t n);

/* The per-parameter update loop of nn/solver.py:190-195 as ONE launch per 32 tensors: rule 0 sgd,
 * 1 sgd_momentum, 2 rmsprop, 3 adam, same formulas as above.  w / dw / s1 / s2 / *_out / ns are HOST
 * arrays of `count` device pointers / element counts; s1 = velocity | cache | m, s2 = v (entries may be
 * NULL = zero state; unused arrays may be NULL).  hyper (host, 4 floats): sgd {lr}; sgd_momentum
 * {lr, momentum}; rmsprop {lr, decay, eps}; adam {lr, beta1, beta2, eps} with step count t. */
int mc_multi_update(int rule, int count, const void* const* w, const void* const* dw, const void* const* s1,
                    const void* const* s2, void* const* w_out, void* const* s1_out, void* const* s2_out,
                    const int64_t* ns, const float* hyper, int64_t t);

/* -------------------------------------------------------------------- RNG ---- */
/* Philox4x32-10 counter RNG: replaces numpy.random.rand/randn (array_variants/numpy/random.py)
 * and mx.random.uniform/normal (array_variants/mxnet/random.py:11-16). */
int mc_rand_uniform(float* out, int64_t n, uint64_t seed, uint64_t offset);
int mc_rand_normal(float* out, int64_t n, uint64_t seed, uint64_t offset);

/* ---------------------------------------------------- convolution / pooling -- */
/* NCHW cross-correlation, the MXNet `Convolution` / `Pooling(max)` / `Activation(relu)`
 * symbols reached through core.Function (minpy/core.py:148-190; examples/nn/cnn.py:32-40). */
/* precision (mc_gemm_precision): 1 = implicit GEMM on tcgen05 (kind::tf32, TMA-gathered windows, FP32
 * accumulate in TMEM) when stride == 1 and rows are 16-byte multiples; 0 = direct FP32 CUDA-core kernels
 * (exact accumulate; also the fallback for shapes the tensor-core path does not cover). */
int mc_conv2d_fwd(const float* x, const float* w, const float* bias, float* y,
                  int64_t N, int64_t C, int64_t H, int64_t W, int64_t F, int64_t KH, int64_t KW,
                  int64_t stride, int64_t pad, int relu, int precision);
int mc_conv2d_dgrad(const float* dy, const float* w, float* dx,
                    int64_t N, int64_t C, int64_t H, int64_t W, int64_t F, int64_t KH, int64_t KW,
                    int64_t stride, int64_t pad, int precision);
int mc_conv2d_wgrad(const float* x, const float* dy, float* dw, float* dbias,
                    int64_t N, int64_t C, int64_t H, int64_t W, int64_t F, int64_t KH, int64_t KW,
                    int64_t stride, int64_t pad, int precision);
/* Convolution + bias (+ relu) + 2x2 / stride-2 max pooling in one kernel (examples/nn/cnn.py:32-40: Convolution ->
 * Activation('relu') -> Pooling(max, 2x2 / 2)): only the pooled tensor y_pooled (N, OH/2, OW/2, Fp) and the window
 * index (idx_u8, same shape, 0..3 = kh*2 + kw of the first maximum; may be NULL) are written -- the format
 * mc_maxpool2d_nhwc_fwd produces, so mc_maxpool2d_nhwc_bwd is its VJP.  Shapes: OH, OW even, OW <= 32, F % 32 == 0,
 * Fp == F (mc_conv2d_nhwc_pool_supported). */
int mc_conv2d_nhwc_pool_supported(int64_t OH, int64_t OW, int64_t F, int64_t Fp, int* ok);
int mc_conv2d_nhwc_fwd_pool(const float* x, const float* w, const float* bias, float* y_pooled, uint8_t* idx_u8,
                            int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp, int64_t wpad,
                            int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad, int relu);
/* Channels-last (NHWC) convolution path -- what the tensor-core kernels natively consume.  TMA needs
 * the innermost box coordinate 16-byte aligned, so the filter-tap shifts must fall on outer
 * dimensions: activations are (N, H, W, Cp) with Cp = channel pitch, a multiple of 4.  The symbol
 * executor (minpy_b200/symbol_exec.py, the stand-in for the MXNet executor of minpy/core.py:148-190)
 * keeps a whole conv trunk in this layout and converts at its boundaries.
 *   x  : (N, H, W + 2*wpad, Cp), physically zero-padded along W by wpad pixels per side; wpad is
 *        either 0 or == pad with KW*Cp <= 32 ("merged" few-channel mode: one k-block per filter row);
 *   w  : (F, C, KH, KW), the reference layout (examples/nn/cnn.py conv_weight);   stride is 1;
 *   y / dy : (N, OH, OW, Fp);  pitch channels (>= C / >= F) are written as zeros by the producers. */
int mc_nchw_to_nhwc(const float* x, float* y, int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp, int64_t wpad);
int mc_nhwc_to_nchw(const float* x, float* y, int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp);
int mc_conv2d_nhwc_fwd(const float* x, const float* w, const float* bias, float* y,
                       int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp, int64_t wpad,
                       int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad, int relu);
int mc_conv2d_nhwc_dgrad(const float* dy, const float* w, float* dx,
                         int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp,
                         int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad);
int mc_conv2d_nhwc_wgrad(const float* x, const float* dy, float* dw, float* dbias,
                         int64_t N, int64_t C, int64_t H, int64_t W, int64_t Cp, int64_t wpad,
                         int64_t F, int64_t Fp, int64_t KH, int64_t KW, int64_t pad);
/* NHWC max pooling; idx_u8 = kh*KW + kw of the first maximal element of each window.  bwd needs
 * non-overlapping windows (stride == K): a gather, no atomics; relu_y (optional, the pooled output of
 * a relu'd tensor) folds the producer's relu mask `out > 0` into the same pass. */
int mc_maxpool2d_nhwc_fwd(const float* x, float* y, uint8_t* idx_u8, int64_t N, int64_t H, int64_t W, int64_t Cp,
                          int64_t KH, int64_t KW, int64_t stride);
int mc_maxpool2d_nhwc_bwd(const float* dy, const uint8_t* idx_u8, const float* relu_y, float* dx,
                          int64_t N, int64_t H, int64_t W, int64_t Cp, int64_t K, int64_t OH, int64_t OW);
/* Which path the last mc_conv2d_* call on this thread took (0 direct fp32, 1 tcgen05 implicit GEMM). */
int mc_conv_last_path(int* path);
int mc_maxpool2d_fwd(const float* x, float* y, int32_t* argmax_i32, int64_t NC, int64_t H, int64_t W,
                     int64_t KH, int64_t KW, int64_t stride);
int mc_maxpool2d_bwd(const float* dy, const int32_t* argmax_i32, float* dx_zeroed, int64_t NC,
                     int64_t H, int64_t W, int64_t OH, int64_t OW);
/* MXNet Pooling(pool_type='avg'), 'valid' windows, NCHW planes (nn/modules.py:273-290, resnet.py:39):
 * y = mean over the KH x KW window; dx gathers dy over the windows containing each element. */
int mc_avgpool2d_fwd(const float* x, float* y, int64_t NC, int64_t H, int64_t W, int64_t KH, int64_t KW, int64_t stride);
int mc_avgpool2d_bwd(const float* dy, float* dx, int64_t NC, int64_t H, int64_t W, int64_t KH, int64_t KW, int64_t stride);

#ifdef __cplusplus
}
#endif
#endif  /* MINPY_CUDA_H_ */

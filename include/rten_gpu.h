/* rten_gpu.h — C ABI of librten_gpu.so, the B200 (sm_100a) replacement for RTen's dense-compute hot path.
 *
 * This is exactly the surface a Rust `rten-gpu-sys` FFI crate would bind.  Every entry point
 * cites the reference interface it replaces as  path:line  relative to the RTen 0.12.0 tree.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ / torch types cross this boundary.
 *   - every function returns rtg_status.  0 = ok; 1..6 mirror the six OpError variants of the reference
 *     (src/ops/mod.rs:665-705) and rtg_last_error() returns the same static message string the reference uses;
 *     RTG_ERR_PANIC marks conditions where the reference panics (src/gemm.rs:743-750,773).
 *   - tensors are (pointer, dtype, shape, element strides) views like the reference's `Input` views
 *     (src/ops/mod.rs:177-180); `device` says where `data` lives: RTG_HOST (-1) or a CUDA device ordinal.  Host
 *     tensors are staged through pinned memory inside the call; device tensors are used in place.
 *   - there is NO CPU fallback: a call fails with RTG_ERR_CUDA if no sm_100 device is usable.
 */
#ifndef RTEN_GPU_H_
#define RTEN_GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t rtg_status;
enum {
  RTG_OK = 0,
  RTG_ERR_INCORRECT_INPUT_TYPE = 1,      /* OpError::IncorrectInputType      src/ops/mod.rs:667 */
  RTG_ERR_INCORRECT_OUTPUT_TYPE = 2,     /* OpError::IncorrectOutputType     src/ops/mod.rs:671 */
  RTG_ERR_INCOMPATIBLE_INPUT_SHAPES = 3, /* OpError::IncompatibleInputShapes src/ops/mod.rs:675 */
  RTG_ERR_MISSING_INPUTS = 4,            /* OpError::MissingInputs           src/ops/mod.rs:678 */
  RTG_ERR_INVALID_VALUE = 5,             /* OpError::InvalidValue            src/ops/mod.rs:682 */
  RTG_ERR_UNSUPPORTED_VALUE = 6,         /* OpError::UnsupportedValue        src/ops/mod.rs:685 */
  RTG_ERR_PANIC = 50,                    /* the reference would panic (assert!) here */
  RTG_ERR_CUDA = 100,
  RTG_ERR_NCCL = 101,
  RTG_ERR_INTERNAL = 102,
  RTG_ERR_FORMAT = 103                   /* malformed .rten file: ModelLoadError, src/model.rs:60-110 */
};

/* Thread-local message of the last failing call on this thread (never NULL). */
const char* rtg_last_error(void);

/* ------------------------------------------------------------------------------------------------ tensors */
enum { RTG_F32 = 0, RTG_I32 = 1 };       /* the reference's only two element types, src/ops/mod.rs:168-180 */
enum { RTG_HOST = -1 };
#define RTG_MAX_DIMS 8

typedef struct rtg_tensor {
  void* data;
  int32_t dtype;
  int32_t ndim;
  int64_t shape[RTG_MAX_DIMS];
  int64_t strides[RTG_MAX_DIMS]; /* in elements; may describe transposed / NHWC views */
  int32_t device;                /* RTG_HOST or CUDA ordinal */
} rtg_tensor;

/* ------------------------------------------------------------------------------------------------ context */
/* One context per (GPU, host thread): owns a stream, the device TensorPool (src/tensor_pool.rs:103-199) and the
 * cache of device-resident, pre-arranged weights (the role of prepack_a/prepack_b, src/gemm.rs:337-438). */
typedef struct rtg_ctx rtg_ctx;

enum { RTG_MATH_TF32 = 0, RTG_MATH_3XTF32 = 1, RTG_MATH_TF32_BF16X2 = 2 };

rtg_status rtg_ctx_create(int32_t device, rtg_ctx** out);
rtg_status rtg_ctx_destroy(rtg_ctx* ctx);
/* Use an externally owned cudaStream_t (e.g. torch's current stream) instead of the context's own. */
rtg_status rtg_ctx_set_stream(rtg_ctx* ctx, void* cuda_stream);
/* RTG_MATH_3XTF32 (default): results within rtol 1e-5 of the reference.  RTG_MATH_TF32: rtol 2e-3.
 * RTG_MATH_TF32_BF16X2: the same split as 3xTF32 (a = hi + lo) with hi*hi on the TF32 pipe and the two correction
 * products lo(A)*hi(B) + hi(A)*lo(B) as ONE bf16 MMA of doubled depth: 2 tensor-core passes per product instead of 3.
 * hi is the operand ROUNDED to nearest TF32 (|lo| <= 2^-12 |a|), so the error per product a*b is at most
 * 2^-18 + 2^-19 (about 5.7e-6) of |a*b|: lo(A)*lo(B) is dropped (2^-24) and the correction operands are rounded to bf16
 * (2 x 2^-12 x 2^-8 = 2^-19 ... 2^-18 with the hi parts' bf16 rounding); held to rtol 1e-5 like 3xTF32
 * (tests/test_fullsize_gpu.py asserts the bound on adversarial operands).
 * Environment: RTEN_GPU_MATH = tf32 | 3xtf32 | tf32_bf16x2 sets the initial mode of every new context; RTEN_GPU_SMS,
 * RTEN_USE_POOL, RTEN_TIMING and RTEN_GPU_GRAPH are described in DESIGN.md section 8. */
rtg_status rtg_ctx_set_math(rtg_ctx* ctx, int32_t mode);
rtg_status rtg_ctx_get_math(rtg_ctx* ctx, int32_t* mode);
rtg_status rtg_ctx_sync(rtg_ctx* ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
rtg_status rtg_ctx_launch_count(rtg_ctx* ctx, uint64_t* count);
/* Name of the GEMM kernel variant chosen by the last contraction call (diagnostics / tests). */
const char* rtg_ctx_last_kernel(rtg_ctx* ctx);
/* Measurement aid (no reference counterpart; the reference's RunOptions::timing works per operator, src/graph.rs
 * :467-483): when enabled, every launch of the tcgen05 contraction kernel is bracketed by CUDA events on the context
 * stream.  rtg_ctx_kernel_timing synchronises, returns the number of launches recorded since the last call, their summed
 * device time, and their algorithmic flops (2*M*N*K with the true K) and bytes (operands once + output), then clears. */
rtg_status rtg_ctx_set_kernel_timing(rtg_ctx* ctx, int32_t enabled);
rtg_status rtg_ctx_kernel_timing(rtg_ctx* ctx, uint64_t* launches, double* total_ms, double* flops, double* bytes);

/* TensorPool::alloc / add / alloc_count / hit_count / len      src/tensor_pool.rs:129-192 */
rtg_status rtg_pool_alloc(rtg_ctx* ctx, size_t bytes, void** dptr);
rtg_status rtg_pool_free(rtg_ctx* ctx, void* dptr);
rtg_status rtg_pool_stats(rtg_ctx* ctx, uint64_t* alloc_count, uint64_t* hit_count, uint64_t* len);
rtg_status rtg_upload(rtg_ctx* ctx, void* dst_device, const void* src_host, size_t bytes);
rtg_status rtg_download(rtg_ctx* ctx, void* dst_host, const void* src_device, size_t bytes);

/* ------------------------------------------------------------------------------------------------ GemmExecutor */
/* GemmExecutor::gemm / gemm_bias / gemm_uninit_bias                src/gemm.rs:448-542
 *   out[m*out_row_stride + n] = alpha * sum_k a[m,k]*b[k,n] + beta * out[..] + bias[m]
 * a is [m,k] with element strides (a_rs, a_cs); b is [k,n] with (b_rs, b_cs).  bias has length m or is NULL.
 * beta == 0 never reads `out` (it may hold NaN; src/gemm.rs:1349-1387).  k == 0 gives out = beta*out (:757-765).
 * `device` is RTG_HOST or the ordinal all pointers live on.  out_len is the length of `out` in elements; a short
 * output is RTG_ERR_PANIC "Output buffer should be large enough" (src/gemm.rs:1167-1183). */
rtg_status rtg_gemm(rtg_ctx* ctx, int32_t device, float* out, size_t out_len, size_t out_row_stride, const float* a,
                    size_t m, size_t k, ptrdiff_t a_rs, ptrdiff_t a_cs, const float* b, size_t b_rows, size_t n,
                    ptrdiff_t b_rs, ptrdiff_t b_cs, float alpha, float beta, const float* bias, size_t bias_len);

/* Split-K GEMM that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arithmetic as rtg_gemm's split-K path (partial sums of `splits` K ranges in scratch planes, folded in split
 * order by the reduction pass together with alpha / beta / bias[m]); the scratch is a file-scope device array of
 * RTG_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit, instead of the context pool.
 *   - all pointers are DEVICE pointers on the context's device; a is a dense row-major [m, k], b a dense row-major
 *     [k, n]; a, b and out 16-byte aligned, k, n and out_row_stride multiples of 4 (so that no operand needs a gathered
 *     copy); k must be long enough to split (two 32-deep blocks in TF32 mode, more than 128 otherwise).
 *   - splits: 0 = the library's choice (at least 2), else the requested number of K ranges.
 *   - the call needs  splits * m * n * 4  bytes (+ n * ceil32(k) * 4 in RTG_MATH_TF32_BF16X2): anything beyond
 *     RTG_STATIC_SCRATCH_BYTES fails with RTG_ERR_UNSUPPORTED_VALUE and a message naming both numbers.
 *   - no cudaMalloc / cudaFree / pool call; fails if per-launch kernel timing (rtg_ctx_set_kernel_timing) is on.
 *   - ALL calls on a GPU share the one array (it keeps the last call's contents; nothing in it has to start at zero,
 *     every element is written before it is read): calls that overlap on the GPU — two streams, two contexts of one
 *     device — race.  Calls on one stream are ordered and safe. */
#define RTG_STATIC_SCRATCH_BYTES (16u << 20)
rtg_status rtg_gemm_static_scratch(rtg_ctx* ctx, float* out, size_t out_len, size_t out_row_stride, const float* a,
                                   size_t m, size_t k, const float* b, size_t n, float alpha, float beta,
                                   const float* bias, int32_t splits);

/* N-split GEMM across GPUs (BASELINE.json configs[4]).  The reference has no multi-device path: this is
 * the exchange step the split adds after GemmExecutor::gemm (src/gemm.rs:448-542), fused into the GEMM epilogue.
 * Every rank owns a peer-visible row-major [m, n_total] result buffer (rtg_symm_alloc; the 64-byte handle is sent to the
 * other ranks, which map it with rtg_symm_open).  rtg_gemm_scatter computes a[m,k] . b[k,n_block] and writes the
 * block into columns [col_offset, col_offset + n_block) of dst[0] (this rank's buffer) AND of dst[1..n_dst) (the peers'
 * buffers): each finished tile leaves shared memory by one TMA store per destination, over NVLink for the peers, so
 * the gather overlaps the contraction tile by tile.  All pointers are device pointers on `device`; a and b may have
 * any element strides.  After the call returns and the stream has drained, a cross-rank barrier makes every buffer
 * complete.  n_dst <= 8; col_offset and n_total must be multiples of 4 (16-byte aligned windows). */
#define RTG_SYMM_HANDLE_BYTES 64
rtg_status rtg_symm_alloc(rtg_ctx* ctx, size_t bytes, void** dptr, uint8_t handle[RTG_SYMM_HANDLE_BYTES]);
rtg_status rtg_symm_free(rtg_ctx* ctx, void* dptr);
rtg_status rtg_symm_open(rtg_ctx* ctx, const uint8_t handle[RTG_SYMM_HANDLE_BYTES], void** peer_dptr);
rtg_status rtg_symm_close(rtg_ctx* ctx, void* peer_dptr);
rtg_status rtg_gemm_scatter(rtg_ctx* ctx, const float* a, size_t m, size_t k, ptrdiff_t a_rs, ptrdiff_t a_cs,
                            const float* b, size_t n_block, ptrdiff_t b_rs, ptrdiff_t b_cs, size_t col_offset,
                            size_t n_total, float* const* dst, int32_t n_dst);

/* ------------------------------------------------------------------------------------------------ operators */
enum { RTG_PAD_FIXED = 0, RTG_PAD_SAME = 1 }; /* Padding::{Fixed,Same}  src/ops/mod.rs:127-166 */
enum {
  RTG_ACT_NONE = 0, RTG_ACT_RELU = 1, RTG_ACT_CLIP = 2, RTG_ACT_GELU = 3, RTG_ACT_SIGMOID = 4, RTG_ACT_SILU = 5,
  RTG_ACT_ERF = 6, RTG_ACT_TANH = 7, RTG_ACT_SQRT = 8
};

typedef struct rtg_conv_attrs {   /* struct Conv {groups, dilations, padding, strides}  src/ops/conv.rs:282-293 */
  int32_t groups;
  int32_t pad_mode;               /* RTG_PAD_FIXED | RTG_PAD_SAME */
  int64_t pads[4];                /* [top, left, bottom, right]  (1-D conv: [left, right]) */
  int32_t n_pads;
  int64_t strides[2];
  int32_t n_strides;
  int64_t dilations[2];
  int32_t n_dilations;
} rtg_conv_attrs;

/* calc_output_size_and_padding + conv() shape checks             src/ops/pooling.rs:27-89, src/ops/conv.rs:86-197
 * Writes the output shape ([N,O,Oh,Ow], or [N,O,Ow] for 3-D input). */
rtg_status rtg_conv_out_shape(const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias_or_null,
                              const rtg_conv_attrs* attrs, int64_t* out_shape, int32_t* out_ndim);
/* Conv::run / conv()                                             src/ops/conv.rs:86-311
 * input NCHW (any strides; NHWC-strided views are consumed without a copy), weight [O, C/groups, Kh, Kw], bias [O].
 * `residual` (same shape as output, or NULL) and `act` are the fused epilogue: out = act(conv + bias + residual).
 * With residual == NULL and act == RTG_ACT_NONE this is exactly the reference operator. */
rtg_status rtg_conv(rtg_ctx* ctx, const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias_or_null,
                    const rtg_conv_attrs* attrs, const rtg_tensor* residual_or_null, int32_t act, float act_lo,
                    float act_hi, rtg_tensor* output);

/* VirtualIm2Col and the layouts around it                          src/ops/conv/im2col.rs:75-260, src/ops/conv.rs:212-271
 * The reference gathers the patch matrix on demand with i32 offset tables and a zero-fill mask; here TMA's im2col mode
 * does it in hardware, with three software layout kernels around it.  This entry point writes what those kernels build
 * to a HOST buffer so that it can be compared bit for bit with the reference's tables (every element is a copy or 0):
 *   kind 0  patch matrix of `group`: [n][oh*ow][kh*kw*kpt], kpt = ceil32(C/groups), column (ky*kw + kx)*kpt + c
 *           (the software gather the convolution falls back to outside TMA's im2col limits)
 *   kind 1  zero-padded NHWC image of the few-channel (row-window) path: [n][hp][wp][ceil4(C)]
 *   kind 2  NHWC copy [n][h*w][ceil4(C)]        kind 3  kind 2 converted back to dense NCHW [n][c][h][w]
 * out_dims receives 4 extents.  RTG_DEBUG_FORCE_IM2COL_FALLBACK (rtg_ctx_set_debug_flags) makes rtg_conv itself take the
 * software gather, whose contraction result must equal the TMA path's bit for bit. */
rtg_status rtg_conv_patches(rtg_ctx* ctx, int32_t kind, const rtg_tensor* input, const rtg_tensor* weight,
                            const rtg_conv_attrs* attrs, int32_t group, float* out_host, size_t out_floats,
                            int64_t* out_dims);
#define RTG_DEBUG_FORCE_IM2COL_FALLBACK (1 << 16)
/* Developer / test switches of a context (no reference counterpart). */
rtg_status rtg_ctx_set_debug_flags(rtg_ctx* ctx, int32_t flags);

/* conv_transpose() / ConvTranspose::run                           src/ops/conv.rs:377-577
 * input NCHW (or NCW), kernel [C, O, Kh, Kw] (or [C, O, Kw]), bias [O] or NULL.  Fixed padding is [top, left, bottom,
 * right] ([left, right] for 1-D); RTG_PAD_SAME gives out = in * stride per axis.  Errors carry the reference's strings
 * ("Input is too small", "Strides must be > 0", "Wrong number of pad values", ...). */
typedef struct rtg_conv_transpose_attrs {   /* struct ConvTranspose {padding, strides}  src/ops/conv.rs:547-551 */
  int32_t pad_mode;               /* RTG_PAD_FIXED | RTG_PAD_SAME */
  int64_t pads[4];
  int32_t n_pads;
  int64_t strides[2];
  int32_t n_strides;
} rtg_conv_transpose_attrs;
rtg_status rtg_conv_transpose_out_shape(const rtg_tensor* input, const rtg_tensor* weight, const rtg_tensor* bias_or_null,
                                        const rtg_conv_transpose_attrs* attrs, int64_t* out_shape, int32_t* out_ndim);
rtg_status rtg_conv_transpose(rtg_ctx* ctx, const rtg_tensor* input, const rtg_tensor* weight,
                              const rtg_tensor* bias_or_null, const rtg_conv_transpose_attrs* attrs, rtg_tensor* output);

/* matmul() / MatMul::run                                          src/ops/matmul.rs:119-254 */
rtg_status rtg_matmul_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* out_shape, int32_t* out_ndim);
rtg_status rtg_matmul(rtg_ctx* ctx, const rtg_tensor* a, const rtg_tensor* b, rtg_tensor* output);
/* gemm_op() / Gemm::run  (ONNX Gemm: alpha*A'B' + beta*C)          src/ops/matmul.rs:27-81 */
rtg_status rtg_gemm_op(rtg_ctx* ctx, const rtg_tensor* a, const rtg_tensor* b, const rtg_tensor* c_or_null,
                       float alpha, float beta, int32_t transpose_a, int32_t transpose_b, rtg_tensor* output);

/* UnaryFloatOp::{run,run_in_place} for Relu/Clip/Gelu/Sigmoid/Silu/Erf   src/ops/unary_elementwise.rs:22-63
 * In place when output->data == input->data. */
rtg_status rtg_unary(rtg_ctx* ctx, int32_t act, const rtg_tensor* input, float lo, float hi, rtg_tensor* output);
enum { RTG_BIN_ADD = 0, RTG_BIN_SUB = 1, RTG_BIN_MUL = 2, RTG_BIN_DIV = 3, RTG_BIN_POW = 4 };
/* binary_op / binary_op_in_place with numpy broadcasting          src/ops/binary_elementwise.rs:65-157,261-322 */
rtg_status rtg_binary_out_shape(const rtg_tensor* a, const rtg_tensor* b, int64_t* out_shape, int32_t* out_ndim);
rtg_status rtg_binary(rtg_ctx* ctx, int32_t op, const rtg_tensor* a, const rtg_tensor* b, rtg_tensor* output);

/* MaxPool / AveragePool / GlobalAveragePool                        src/ops/pooling.rs:91-420 */
typedef struct rtg_pool_attrs {
  int64_t kernel[2];
  int64_t strides[2];
  int32_t pad_mode;
  int64_t pads[4];
  int32_t count_include_pad;
} rtg_pool_attrs;
enum { RTG_POOL_MAX = 0, RTG_POOL_AVG = 1 };
rtg_status rtg_pool2d_out_shape(const rtg_tensor* input, const rtg_pool_attrs* attrs, int64_t* out_shape);
rtg_status rtg_pool2d(rtg_ctx* ctx, int32_t kind, const rtg_tensor* input, const rtg_pool_attrs* attrs,
                      rtg_tensor* output);
rtg_status rtg_global_average_pool(rtg_ctx* ctx, const rtg_tensor* input, rtg_tensor* output);

/* ------------------------------------------------------------------------------------------------ Model / Graph */
/* Model::load / load_file, Model::run, Graph::run                  src/model.rs:214-309,580-646; src/graph.rs:733-1073
 * A `.rten` file (V1 or V2 container, src/header.rs:60-80, src/schema.fbs) is parsed by a hand-written FlatBuffers
 * reader; operators are instantiated through the op registry (src/op_registry.rs:36-62) and executed on the GPU. */
typedef struct rtg_model rtg_model;
rtg_status rtg_model_load_file(rtg_ctx* ctx, const char* path, int32_t optimize, rtg_model** out);
rtg_status rtg_model_load(rtg_ctx* ctx, const void* data, size_t len, int32_t optimize, rtg_model** out);
rtg_status rtg_model_free(rtg_model* model);
rtg_status rtg_model_num_inputs(const rtg_model* model, int32_t* n);
rtg_status rtg_model_num_outputs(const rtg_model* model, int32_t* n);
/* Name and (possibly symbolic: -1) shape of the i-th graph input / output:  Model::input_ids / node_info */
rtg_status rtg_model_input_info(const rtg_model* model, int32_t i, const char** name, int64_t* shape, int32_t* ndim,
                                int32_t* dtype);
rtg_status rtg_model_output_info(const rtg_model* model, int32_t i, const char** name, int64_t* shape, int32_t* ndim,
                                 int32_t* dtype);
rtg_status rtg_model_num_ops(const rtg_model* model, int32_t* n);

/* Run handle: the outputs of one Model::run.  Shapes are known after the run; copy out with rtg_run_output_copy.
 * Lifetime of HOST inputs: rtg_model_run* enqueue the uploads on the context's stream and return without waiting (with
 * pinned memory the copy really is asynchronous), so the caller must leave the input buffers unchanged until it has
 * copied an output to the host (rtg_run_output_copy synchronises) or called rtg_ctx_sync. */
typedef struct rtg_run rtg_run;
rtg_status rtg_model_run(rtg_model* model, const rtg_tensor* inputs, int32_t n_inputs, rtg_run** out);
rtg_status rtg_run_num_outputs(const rtg_run* run, int32_t* n);
rtg_status rtg_run_output_info(const rtg_run* run, int32_t i, int64_t* shape, int32_t* ndim, int32_t* dtype);
/* Copies output i (contiguous, reference layout) to `dst` on `device` (RTG_HOST or ordinal). */
rtg_status rtg_run_output_copy(rtg_run* run, int32_t i, void* dst, size_t dst_bytes, int32_t device);
rtg_status rtg_run_free(rtg_run* run);
/* Model::run proper: inputs and outputs addressed by node id                    src/model.rs:580-592
 * (rtg_model_run above is run with the graph's own input / output lists).  Any value node may be requested, e.g. an
 * intermediate activation; ids must be unique ("Inputs are not unique" / "Outputs are not unique"). */
rtg_status rtg_model_run_ids(rtg_model* model, const int32_t* input_ids, const rtg_tensor* inputs, int32_t n_inputs,
                             const int32_t* output_ids, int32_t n_outputs, rtg_run** out);
/* Model::partial_run / Graph::partial_run                                      src/model.rs:639-646, src/graph.rs:1147-1235
 * Runs with an incomplete set of inputs: every operator of the plan whose inputs are available is evaluated; the run's
 * outputs are the leaves of that sub-plan (requested outputs, or values feeding operators that could not run).
 * rtg_run_output_node gives the node id of each; pass them to rtg_model_run_ids later with the remaining inputs. */
rtg_status rtg_model_partial_run(rtg_model* model, const int32_t* input_ids, const rtg_tensor* inputs, int32_t n_inputs,
                                 const int32_t* output_ids, int32_t n_outputs, rtg_run** out);
rtg_status rtg_run_output_node(const rtg_run* run, int32_t i, int32_t* node_id);
/* Model::find_node, input_ids, output_ids                                      src/model.rs:525-553 */
rtg_status rtg_model_find_node(const rtg_model* model, const char* name, int32_t* id);
rtg_status rtg_model_input_id(const rtg_model* model, int32_t i, int32_t* id);
rtg_status rtg_model_output_id(const rtg_model* model, int32_t i, int32_t* id);
/* Graph::get_node / get_source_node                                            src/graph.rs:560-600
 * kind: 0 operator, 1 constant, 2 value.  rtg_model_source_op: the operator producing value `value_id` after the
 * load-time optimiser (GraphOptimizer, src/optimize.rs:286-297) ran: its operator name ("Silu", "Gelu",
 * "LayerNormalization", "FusedMatMul", ...) and its node name. */
rtg_status rtg_model_node_kind(const rtg_model* model, int32_t id, int32_t* kind);
rtg_status rtg_model_source_op(const rtg_model* model, int32_t value_id, const char** op_name, const char** node_name);

/* Batch-sharded Model::run inside one process (the reference's advice for scaling is the same:
 * "use Rayon to execute the model on separate chunks in parallel", docs/performance.md:137-139).  The model is loaded
 * once per device of `devices` (weights replicated; a device may be named twice: two contexts on one GPU); a run splits
 * dim 0 of every (host, contiguous) input into contiguous chunks — the first batch % g shards take one item more —, runs
 * them concurrently (one host thread, stream and device pool per shard) and writes every shard's rows into the
 * caller's host output buffers ([batch, ...] contiguous).  There is no collective.  Loading fails with
 * RTG_ERR_UNSUPPORTED_VALUE when an operator of the graph is not known to keep batch items independent (a Reshape that
 * folds the batch, a Transpose that moves it, reductions over dim 0, activation x activation MatMul ...); a failed
 * shard fails the whole run. */
typedef struct rtg_sharded_model rtg_sharded_model;
rtg_status rtg_sharded_model_load(const int32_t* devices, int32_t n_devices, int32_t math, const void* data, size_t len,
                                  int32_t optimize, rtg_sharded_model** out);
rtg_status rtg_sharded_model_num_shards(const rtg_sharded_model* model, int32_t* n);
rtg_status rtg_sharded_model_run(rtg_sharded_model* model, const rtg_tensor* inputs, int32_t n_inputs,
                                 rtg_tensor* outputs, int32_t n_outputs);
rtg_status rtg_sharded_model_free(rtg_sharded_model* model);

/* Asynchronous Model::run over HOST buffers (no reference counterpart: the reference runs on the caller's thread;
 * this is what a serving loop binds instead of rtg_model_run to keep the PCIe copies off the critical path).
 * The inputs are uploaded on a second stream into one of two staging slots, so the upload of submit i + 1 overlaps
 * the graph of submit i; the outputs are written to `outputs[i].data` (host; `shape` gives the capacity on entry and
 * the actual shape on return).  All buffers must stay valid until rtg_ticket_wait, which also frees the ticket.
 * Use pinned host memory, or the copies serialise with the caller. */
typedef struct rtg_ticket rtg_ticket;
rtg_status rtg_model_submit(rtg_model* model, const rtg_tensor* inputs, int32_t n_inputs, rtg_tensor* outputs,
                            int32_t n_outputs, rtg_ticket** out);
rtg_status rtg_ticket_wait(rtg_ticket* ticket);
/* CUDA-graph replay of the cached plan (no reference counterpart: it replaces the per-operator dispatch loop of
 * run_plan, src/graph.rs:883-1051, whose cost on a GPU is launch latency).  On by default: the first run of a
 * (node ids, input shapes, math mode) key executes eagerly, the second is stream-captured, later runs replay one
 * cudaGraphLaunch.  Graphs that need a host round trip while running (shape tensors computed on the device and read
 * back) stay on the eager path.  Environment: RTEN_GPU_GRAPH=0 turns it off for every model.
 * rtg_model_graph_replays: how many runs of this model were replays (diagnostics / tests). */
rtg_status rtg_model_set_graph_capture(rtg_model* model, int32_t enabled);
rtg_status rtg_model_graph_replays(rtg_model* model, uint64_t* replays);
/* RunOptions{timing}: per-operator device time of the last run as text   src/graph.rs:467-483, src/timing.rs */
rtg_status rtg_model_set_timing(rtg_model* model, int32_t enabled);
const char* rtg_model_timing_report(rtg_model* model);

#ifdef __cplusplus
}
#endif
#endif /* RTEN_GPU_H_ */

/*
 * pdb200.h — C-ABI of the B200-native PHI hot path (libpdb200.so).
 *
 * One entry point per PHI kernel on the path named by BASELINE.json's north_star:
 * the broadcast/elementwise and reduce machinery of
 *   paddle/phi/kernels/funcs/broadcast_function.h, reduce_function.h
 * and the ops built directly on it.  Each function is what a PHI kernel body
 * (PD_REGISTER_KERNEL / PD_RegisterPhiKernel) forwards to after
 * it has pulled raw pointers, dims and the stream out of DenseTensor / GPUContext.
 *
 * Conventions (identical for every entry point)
 *   - plain pointers and sizes only; no C++/torch/Paddle types cross this boundary;
 *   - every tensor pointer is DEVICE memory of the current CUDA device, contiguous
 *     row-major (PHI's PrepareData guarantees this, paddle/phi/api/lib/data_transform.cc:394-428);
 *   - outputs are caller-allocated with the shape/dtype Paddle's InferMeta produces
 *     (the kernel's `dev_ctx.Alloc<T>(out)`, paddle/phi/core/device_context.h:156);
 *   - `stream` is a cudaStream_t passed as void* (dev_ctx.stream(),
 *     paddle/phi/backends/gpu/gpu_context.h:100); all work is enqueued on it and the
 *     call never synchronises the host;
 *   - scratch memory is caller-provided (`workspace`, `workspace_bytes`), sized by the
 *     matching *_workspace_bytes() query — the PHI wrapper allocates it from Paddle's
 *     stream-safe allocator the way reduce_function.h:287-295 does with phi::Empty;
 *   - return value: PDB200_OK or an error code; pdb200_last_error() returns the
 *     message of the last failure on the calling thread (the PHI wrapper turns it into
 *     PADDLE_THROW(InvalidArgument(...)) — same checks as the reference, cited below);
 *   - there is NO CPU fallback: without a CUDA device every compute call fails with
 *     PDB200_CUDA_ERROR.
 */
#ifndef PDB200_H_
#define PDB200_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define PDB200_API __attribute__((visibility("default")))
#else
#define PDB200_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* dtype codes (subset of phi::DataType, paddle/phi/common/data_type.h) */
typedef enum {
  PDB200_BOOL = 0,
  PDB200_UINT8 = 1,
  PDB200_INT8 = 2,
  PDB200_INT16 = 3,
  PDB200_INT32 = 4,
  PDB200_INT64 = 5,
  PDB200_FLOAT16 = 6,
  PDB200_BFLOAT16 = 7,
  PDB200_FLOAT32 = 8,
  PDB200_FLOAT64 = 9
} pdb200_dtype;

typedef enum {
  PDB200_OK = 0,
  PDB200_INVALID_ARGUMENT = 1, /* shape/axis/dtype rule violated (PADDLE_ENFORCE_* in the reference) */
  PDB200_UNSUPPORTED = 2,      /* dtype not registered for this op in the reference either */
  PDB200_CUDA_ERROR = 3,       /* launch/runtime failure, or no CUDA device */
  PDB200_WORKSPACE_TOO_SMALL = 4
} pdb200_status;

#define PDB200_MAX_RANK 9 /* common::DDim::kMaxRank, paddle/common/ddim.h */

/* Message of the last error raised on this thread ("" if none). */
PDB200_API const char* pdb200_last_error(void);
/* Library version string and the SM architecture it was compiled for ("sm_100a"). */
PDB200_API const char* pdb200_version(void);
/* Number of kernels this library has launched in this process (all threads). */
PDB200_API uint64_t pdb200_launch_count(void);

/* ---------------------------------------------------------------------------
 * Shape helpers (host only, no CUDA): what InferMeta computes for these ops.
 * ------------------------------------------------------------------------- */

/* NumPy right-aligned broadcast of two shapes.
 * Replaces: funcs::GetBroadcastDimsArrays (paddle/phi/kernels/funcs/common_shape.h:35-93) as used by
 * ElementwiseInferMeta (paddle/phi/infermeta/binary.cc:1652).  out_dims has room for PDB200_MAX_RANK. */
PDB200_API int pdb200_broadcast_shape(const int64_t* x_dims, int x_rank, const int64_t* y_dims, int y_rank,
                           int64_t* out_dims, int* out_rank);

/* ---------------------------------------------------------------------------
 * Elementwise with broadcasting.
 * Replaces: phi::AddKernel / phi::MultiplyKernel<T, GPUContext>
 *   (paddle/phi/kernels/kps/elementwise_kernel.cu:84-99, legacy/kps/elementwise_kernel.cu:25-33)
 *   -> funcs::BroadcastKernel (paddle/phi/kernels/funcs/broadcast_function.h:759-824).
 * out has the broadcast shape; x, y and out share `dtype`.  out may alias x or y when
 * that operand already has the output shape (add_ / multiply_ in-place variants).
 * dtypes: all of pdb200_dtype; multiply<bool> is logical AND (elementwise_functor.h:72-77).
 * ------------------------------------------------------------------------- */
PDB200_API int pdb200_add(const void* x, const int64_t* x_dims, int x_rank, const void* y,
               const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
PDB200_API int pdb200_multiply(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                    const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);

/* The other binary ops on the same broadcast engine.
 * Replace phi::SubtractKernel / DivideKernel / MaximumKernel / MinimumKernel<T, GPUContext>
 *   (paddle/phi/kernels/kps/elementwise_kernel.cu:30-53,183-202,338-381) with the functors of
 *   paddle/phi/kernels/funcs/elementwise_functor.h:80-120,471-511.
 * dtypes: subtract {f32,f64,i16,i32,i64,f16,bf16}; divide {f32,f64,i8,u8,i16,i32,i64,f16,bf16};
 * maximum / minimum {f32,f64,i32,i64,f16,bf16} (the GPU registrations; complex is out of scope).
 * maximum is `a > b ? a : b` and minimum `a < b ? a : b` (a NaN operand returns b), not fmax/fmin.
 * Integer divide truncates toward zero; a zero divisor yields 0 (the reference raises
 * "Integer division by zero" through a device-side assert, which a C-ABI kernel cannot do). */
PDB200_API int pdb200_subtract(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                    const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
PDB200_API int pdb200_divide(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                  const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
PDB200_API int pdb200_maximum(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                   const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
PDB200_API int pdb200_minimum(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                   const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);

/* ---------------------------------------------------------------------------
 * Compare / logical / bitwise with broadcasting — the remaining binary callers of
 * funcs::BroadcastKernel.  Shapes broadcast exactly as for pdb200_add; `dtype` is the dtype of x and y.
 *
 * pdb200_compare replaces LessThan/LessEqual/GreaterThan/GreaterEqual/Equal/NotEqual Kernel
 *   paddle/phi/kernels/kps/compare_kernel.cu:49-66,143-173 (functors funcs/compare_functors.h:21-62).
 *   out is BOOL (one byte per element).  All ten dtypes.  equal / not_equal on the floating types (bf16
 *   and fp16 included) follow the reference's EqualFunctor: |a - b| < 1e-8 counts as equal (a - b taken in
 *   the tensor's type), infinities compare exactly, NaN equals nothing; integers and bool use ==.
 * pdb200_logical replaces LogicalAnd/Or/Xor Kernel, kps/logical_kernel.cu:60-69,112-134: out (BOOL) =
 *   bool(x) op bool(y); every dtype except uint8.
 * pdb200_bitwise replaces BitwiseAnd/Or/Xor Kernel, kps/bitwise_kernel.cu:55-67,94-125: out has the input
 *   dtype; bool and the integer dtypes (bool uses &&, ||, !=). */
enum pdb200_compare_op {
  PDB200_LESS_THAN = 0, PDB200_LESS_EQUAL = 1, PDB200_GREATER_THAN = 2, PDB200_GREATER_EQUAL = 3,
  PDB200_EQUAL = 4, PDB200_NOT_EQUAL = 5
};
enum pdb200_logical_op { PDB200_LOGICAL_AND = 0, PDB200_LOGICAL_OR = 1, PDB200_LOGICAL_XOR = 2 };
enum pdb200_bitwise_op { PDB200_BITWISE_AND = 0, PDB200_BITWISE_OR = 1, PDB200_BITWISE_XOR = 2 };
PDB200_API int pdb200_compare(int op, const void* x, const int64_t* x_dims, int x_rank, const void* y,
                   const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
PDB200_API int pdb200_logical(int op, const void* x, const int64_t* x_dims, int x_rank, const void* y,
                   const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
PDB200_API int pdb200_bitwise(int op, const void* x, const int64_t* x_dims, int x_rank, const void* y,
                   const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);
/* bitwise_left_shift / bitwise_right_shift (kps/bitwise_kernel.cu:138-156, funcs/bitwise_functors.h:51-200): out =
 * x << y or x >> y elementwise with broadcasting; right != 0 selects the right shift, is_arithmetic != 0 the
 * sign-filling one (a logical right shift of a signed value shifts its unsigned representation).  A count outside
 * [0, bits) gives 0, or the sign fill for an arithmetic right shift.  dtypes: uint8, int8, int16, int32, int64. */
PDB200_API int pdb200_bitwise_shift(int right, int is_arithmetic, const void* x, const int64_t* x_dims, int x_rank,
                         const void* y, const int64_t* y_dims, int y_rank, void* out, int dtype, void* stream);

/* ---------------------------------------------------------------------------
 * One-input maps and `where` — the callers of funcs::ElementwiseKernel.
 * `numel` elements, contiguous.  out may be x itself when both have the same element size.
 *
 * pdb200_logical_not  LogicalNotKernel, kps/logical_kernel.cu:71-91: out (BOOL) = !bool(x); no uint8.
 * pdb200_bitwise_not  BitwiseNotKernel, kps/bitwise_kernel.cu:69-80: ~x (bool: !x); bool and integers.
 * pdb200_scale        ScaleKernel, gpu/scale_kernel.cu:24-65: out = scale*x + bias (bias_after_scale != 0)
 *                     or scale*(x + bias); computed in float for bf16/fp16 and in the tensor's own type
 *                     otherwise, scale and bias converted to that type first (Scalar::to<>).
 *                     Integer overflow wraps.  All ten dtypes.
 * pdb200_cast         CastKernel, gpu/cast_kernel.cu + gpu/cast_impl.h: static_cast per element between any
 *                     two of the ten dtypes (bf16/fp16 through float; to bool = "!= 0"; float -> integer
 *                     truncates toward zero; out-of-range float -> integer is undefined in the reference
 *                     and is not reproduced).
 * pdb200_where        WhereKernel, gpu/where_kernel.cu:24-51: out = condition ? x : y, three same-shape
 *                     tensors (broadcasting is done by the Python layer in the reference); condition is
 *                     BOOL; any of the ten dtypes (a bit select, built per element size). */
PDB200_API int pdb200_logical_not(const void* x, int64_t numel, int dtype, void* out, void* stream);
PDB200_API int pdb200_bitwise_not(const void* x, int64_t numel, int dtype, void* out, void* stream);
PDB200_API int pdb200_scale(const void* x, int64_t numel, int dtype, double scale, double bias,
                 int bias_after_scale, void* out, void* stream);
PDB200_API int pdb200_cast(const void* x, int64_t numel, int x_dtype, int out_dtype, void* out, void* stream);
PDB200_API int pdb200_where(const void* condition, const void* x, const void* y, int64_t numel, int dtype,
                 void* out, void* stream);

/* ---------------------------------------------------------------------------
 * Backward of add / multiply with broadcasting: the gradient of an operand that was
 * broadcast is dout summed over the broadcast dimensions (funcs::GetReduceDim, axis = -1).
 * Replaces phi::AddGradKernel / MultiplyGradKernel on GPU:
 *   paddle/phi/kernels/gpu/elementwise_grad_kernel.cu:361-372,431-443 (registrations)
 *   paddle/phi/kernels/gpu/elementwise_grad.h:148-232 (add), :57-107,371-399 (multiply)
 * dout has the broadcast shape of x_dims and y_dims; dx has x's shape, dy has y's; either may be NULL (not
 * wanted).  dx or dy may be the dout buffer itself when their shapes agree (the in-place case: nothing is
 * copied).  add_grad: dx = sum(dout), dy = sum(dout).  multiply_grad: dx = sum(dout * y), dy = sum(dout * x),
 * each product rounded to the tensor dtype first (as the reference's temporaries are), sums in float for
 * bf16/fp16 (double for float64, wrapping integers) and rounded once.
 * dtypes: float32, float64, int32, int64, float16, bfloat16 (complex, and bool multiply_grad, are out of scope).
 * Workspace: ask pdb200_elementwise_grad_workspace_bytes (op = PDB200_GRAD_ADD / PDB200_GRAD_MULTIPLY); 0 when
 * nothing is reduced; 16-byte aligned; PDB200_WORKSPACE_TOO_SMALL if the buffer is smaller. */
enum pdb200_grad_op { PDB200_GRAD_ADD = 0, PDB200_GRAD_MULTIPLY = 1 };
PDB200_API size_t pdb200_elementwise_grad_workspace_bytes(int op, const int64_t* x_dims, int x_rank,
                                               const int64_t* y_dims, int y_rank, int dtype, int want_dx,
                                               int want_dy);
PDB200_API int pdb200_add_grad(const int64_t* x_dims, int x_rank, const int64_t* y_dims, int y_rank, const void* dout,
                    int dtype, void* dx, void* dy, void* workspace, size_t workspace_bytes, void* stream);
PDB200_API int pdb200_multiply_grad(const void* x, const int64_t* x_dims, int x_rank, const void* y,
                         const int64_t* y_dims, int y_rank, const void* dout, int dtype, void* dx, void* dy,
                         void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * Reductions.  `axes` may hold negative values; n_axes == 0 or reduce_all != 0 or a full axis
 * set means "reduce everything" (phi::recompute_reduce_all, paddle/phi/core/kernel_utils.h:407-417).
 * keep_dim only changes the output SHAPE (InferMeta), never the data, so it is not a parameter.
 * x_dtype != out_dtype means "cast each element to out_dtype, then reduce in it"
 * (paddle/phi/kernels/gpu/reduce.h:48-69); supported pairs: equal dtypes, any integer/bool -> INT64,
 * FLOAT16/BFLOAT16 -> FLOAT32 (kps/reduce_kernel.cu:251-265), FLOAT32 -> FLOAT64.
 * Accumulation type follows phi::dtype::MPTypeTrait (float for fp16/bf16, else out dtype;
 * paddle/phi/common/amp_type_traits.h:24-54).  x must not be empty
 * (reduce_function.h:1052-1055).  64-bit indexing throughout.
 * ------------------------------------------------------------------------- */
typedef enum {
  PDB200_REDUCE_SUM = 0,
  PDB200_REDUCE_MEAN = 1,
  PDB200_REDUCE_MAX = 2,
  PDB200_REDUCE_MIN = 3,
  PDB200_REDUCE_PROD = 4
} pdb200_reduce_op;

/* Bytes of device scratch the call below needs for these arguments (0 when none). */
PDB200_API size_t pdb200_reduce_workspace_bytes(int op, const int64_t* x_dims, int x_rank, const int64_t* axes,
                                     int n_axes, int reduce_all, int x_dtype, int out_dtype);

/* Replaces phi::SumRawKernel<T, GPUContext> (paddle/phi/kernels/kps/reduce_kernel.cu:184-271). */
PDB200_API int pdb200_sum_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                   int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                   size_t workspace_bytes, void* stream);
/* Replaces phi::MeanRawKernel (kps/reduce_kernel.cu:109-120). Integer mean truncates toward zero. */
PDB200_API int pdb200_mean_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                    int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                    size_t workspace_bytes, void* stream);
/* Replaces phi::MaxRawKernel (paddle/phi/kernels/legacy/kps/reduce_max_kernel.cu:21-32). */
PDB200_API int pdb200_max_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                   int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                   size_t workspace_bytes, void* stream);
/* Replaces phi::MinRawKernel (kps/reduce_kernel.cu:122-133). */
PDB200_API int pdb200_min_raw(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                   int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                   size_t workspace_bytes, void* stream);
/* Replaces phi::ProdKernel (kps/reduce_kernel.cu:34-45). */
PDB200_API int pdb200_prod(const void* x, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                int reduce_all, int x_dtype, int out_dtype, void* out, void* workspace,
                size_t workspace_bytes, void* stream);

/* Column sums without any allocation: out[c] = sum over r of x[r][c] for a contiguous [rows, cols]
 * matrix, the same result as pdb200_sum_raw(x, {rows, cols}, axes = {0}).
 * No counterpart in the reference: every PHI kernel gets its scratch from the device context's
 * allocator.  This variant is for callers that have no allocator to offer and may not call cudaMalloc
 * themselves (memory planned and handed out at start-up, real-time loops, plugin hosts without an
 * allocator): it takes no workspace argument and keeps its partial sums in a fixed array of
 * PDB200_STATIC_SCRATCH_BYTES (1.5 MiB) inside the library's device image — present on every GPU that
 * loads the library, for the life of the process.  Limits and rules:
 *   - shapes whose partial sums need more than PDB200_STATIC_SCRATCH_BYTES are refused with
 *     PDB200_WORKSPACE_TOO_SMALL (the message says how much was needed); shapes wide enough to need
 *     no partials at all are unaffected by the limit;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *     that device current); make that first call before starting any stream capture;
 *   - the array is shared by all calls on a device and keeps the previous call's contents (nothing in
 *     it needs to start at zero); calls that can overlap on the GPU — different streams without an
 *     ordering between them — race on it, so keep them on one stream or order them with events;
 *   - it calls no cudaMalloc / cudaFree and no allocating library routine.
 * dtype pairs: FLOAT32->FLOAT32, BFLOAT16->BFLOAT16, BFLOAT16->FLOAT32, INT64->INT64. */
#define PDB200_STATIC_SCRATCH_BYTES (3u << 19)
PDB200_API int pdb200_sum_axis0_static(const void* x, int64_t rows, int64_t cols, int x_dtype, int out_dtype,
                                       void* out, void* stream);

/* ---------------------------------------------------------------------------
 * Backward of sum / mean: x_grad = out_grad broadcast back over the reduced dimensions
 * (mean: divided by n = the number of reduced elements, in float for bf16/fp16 — the CPU kernel's g / n; the
 * GPU kernel's g * (1.0 / n) differs by at most half an fp32 ulp, and for integers by being zero).
 * Replaces phi::ReduceSumGradKernel / ReduceMeanGradKernel, paddle/phi/kernels/gpu/reduce_kernel.cu:40-76,122-155.
 * x_dims is the shape of the forward INPUT (= of x_grad); out_grad holds one element per kept position (its own
 * rank — keep_dim or not — does not matter: the data layout is the same); axes / reduce_all as for the forward.
 * dtypes: float32, float64, float16, bfloat16, int32, int64.  No workspace. */
PDB200_API int pdb200_sum_grad(const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                    int reduce_all, int dtype, void* x_grad, void* stream);
PDB200_API int pdb200_mean_grad(const void* out_grad, const int64_t* x_dims, int x_rank, const int64_t* axes, int n_axes,
                     int reduce_all, int dtype, void* x_grad, void* stream);

/* Backward of max / min / prod.  x is the forward input, out the forward result and out_grad its gradient (one element
 * per kept position each, as above); x_grad has x's shape.
 *   max_grad, min_grad: x_grad = out_grad * (x equals out ? 1 : 0) — "equals" is the reference's EqualFunctor
 *     (compare_functors.h:35-48: |a - b| < 1e-8 for floating types, infinities exact, NaN equals nothing), every tied
 *     position receives the whole out_grad, and out_grad * 0 keeps inf * 0 = NaN.  Replaces
 *     phi::ReduceMaxGradKernel / ReduceMinGradKernel, paddle/phi/kernels/gpu/reduce_kernel.cu:76-120,159-203 (two
 *     broadcast kernels and an x-sized temporary there, one pass here).  The two entry points run the same kernel.
 *   prod_grad: x_grad = (out_grad * out) * (1 / x), each step rounded to the tensor type (integers: 1 / x truncates to
 *     1, -1 or 0; x == 0, a division by zero in the reference, gives 0).  Replaces phi::ProdGradKernel,
 *     paddle/phi/kernels/gpu/prod_grad_kernel.cu:21-30 -> funcs::ProdGradFunctor (funcs/reduce_functor.h:143-159).
 * dtypes: float32, float64, float16, bfloat16, int32, int64.  No workspace. */
PDB200_API int pdb200_max_grad(const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                    const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream);
PDB200_API int pdb200_min_grad(const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                    const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream);
PDB200_API int pdb200_prod_grad(const void* x, const int64_t* x_dims, int x_rank, const void* out, const void* out_grad,
                     const int64_t* axes, int n_axes, int reduce_all, int dtype, void* x_grad, void* stream);

/* ---------------------------------------------------------------------------
 * argmax.  Replaces phi::ArgMaxKernel<T, GPUContext> (paddle/phi/kernels/gpu/arg_min_max_kernel.cu:59-204).
 * Lowest index wins ties; out_dtype is PDB200_INT64 or PDB200_INT32; flatten != 0 reduces over the
 * flattened tensor (axis ignored); 0-D input yields 0.  x must not be empty (:214-219).
 * ------------------------------------------------------------------------- */
PDB200_API size_t pdb200_argmax_workspace_bytes(const int64_t* x_dims, int x_rank, int64_t axis, int flatten,
                                     int x_dtype);
PDB200_API int pdb200_argmax(const void* x, const int64_t* x_dims, int x_rank, int64_t axis, int flatten,
                  int x_dtype, int out_dtype, void* out, void* workspace, size_t workspace_bytes,
                  void* stream);
/* argmin.  Replaces phi::ArgMinKernel<T, GPUContext> (paddle/phi/kernels/gpu/arg_min_max_kernel.cu:233-246, the same
 * ArgCUDAKernel with cub::ArgMin).  Same arguments, rules and workspace query as pdb200_argmax; lowest index wins ties. */
PDB200_API int pdb200_argmin(const void* x, const int64_t* x_dims, int x_rank, int64_t axis, int flatten,
                  int x_dtype, int out_dtype, void* out, void* workspace, size_t workspace_bytes,
                  void* stream);

/* ---------------------------------------------------------------------------
 * softmax forward along `axis`.  Replaces phi::SoftmaxGPUDNNKernel / SoftmaxKernel<T, GPUContext>
 * (paddle/phi/kernels/gpudnn/softmax_kernel.cu:25-39 -> softmax_gpudnn.h:1239-1329).
 * dtypes FLOAT32, FLOAT64, FLOAT16, BFLOAT16 (half types accumulate in float, one rounding at the end).
 * numel == 0 is a no-op, 0-D yields 1.  x - max is clamped at -64 as the CPU kernel does
 * (paddle/phi/kernels/funcs/softmax_impl.h:36-42).  out may alias x.
 * ------------------------------------------------------------------------- */
PDB200_API int pdb200_softmax(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out,
                   void* stream);

/* log_softmax forward along `axis`.  Replaces phi::LogSoftmaxKernel<T, GPUContext>
 * (paddle/phi/kernels/gpu/log_softmax_kernel.cu; numerics of the CPU kernel,
 * paddle/phi/kernels/cpu/log_softmax_kernel.cc:40-127): y = d - log(sum(exp(d))) with d = max(x - max, -64).
 * Same routes, dtypes and edge cases as pdb200_softmax (0-D yields 0). */
PDB200_API int pdb200_log_softmax(const void* x, const int64_t* dims, int rank, int axis, int dtype, void* out,
                       void* stream);

/* cross_entropy_with_softmax forward.  Replaces phi::CrossEntropyWithSoftmaxKernel<T, GPUContext>
 * (paddle/phi/kernels/gpu/cross_entropy_kernel.cu:1452-1476) with the numerics of the CPU kernel
 * (paddle/phi/kernels/cpu/cross_entropy_kernel.cc:27-104, funcs/cross_entropy.cc:29-134).
 * logits dims/rank, softmax along `axis` -> `softmax` (same shape; may be the logits buffer, as the op is declared
 * in-place) and `loss` (the logits shape with the axis dimension set to 1).
 *   soft_label == 0: label holds one class index per position (label_dtype INT64 or INT32, the logits shape with the
 *     axis dimension 1); loss = -log(softmax[label]), 0 where label == ignore_index, NaN for an index out of range
 *     (the reference raises there).
 *   soft_label != 0: label has the logits shape and dtype; loss = -sum(label * log(softmax)) along the axis.
 *   use_softmax == 0: the input already is a softmax: it is copied to `softmax` and only the loss is computed.
 * The loss is computed in float (double for FLOAT64) from the logits and the two statistics of each row (max, log of the
 * sum) that the softmax kernel hands over through `workspace` (pdb200_cross_entropy_workspace_bytes; 16-byte aligned):
 * log softmax[k] = max(x[k] - max, -64) - log(sum).  Without use_softmax, or when `softmax` is the logits buffer, the
 * loss is -log of the stored probability (+-inf become +-1e20, TolerableValue) and no workspace is needed.
 * dtypes FLOAT32, FLOAT64, FLOAT16 (the reference's GPU registrations) and BFLOAT16. */
PDB200_API size_t pdb200_cross_entropy_workspace_bytes(const int64_t* dims, int rank, int axis, int dtype);
PDB200_API int pdb200_cross_entropy_with_softmax(const void* logits, const int64_t* dims, int rank, const void* label,
                                      int label_dtype, int soft_label, int use_softmax, int ignore_index, int axis,
                                      int dtype, void* softmax, void* loss, void* workspace, size_t workspace_bytes,
                                      void* stream);
/* Backward of cross_entropy_with_softmax with use_softmax != 0: logits_grad from the saved softmax, the labels and loss_grad
 * (one value per position, the logits' shape with a 1 along `axis`).  Replaces phi::CrossEntropyWithSoftmaxGradKernel<T,
 * GPUContext> (paddle/phi/kernels/gpu/cross_entropy_grad_kernel.cu:281-306); arithmetic as the CPU kernel
 * (cpu/cross_entropy_grad_kernel.cc:143-179): soft labels g * (p - y); hard labels p * g, minus g at the label's class (two
 * roundings), zero for a position whose label is ignore_index.  dims / rank describe softmax and logits_grad.  label:
 * INT64 / INT32 class indices or, with soft_label, distributions of the tensor dtype.  dtypes FLOAT32, FLOAT64, FLOAT16,
 * BFLOAT16 (16-bit: float arithmetic, one rounding).  use_softmax == 0 returns PDB200_UNSUPPORTED (the PHI binding
 * forwards it).  logits_grad may alias softmax (the reference computes in place). */
PDB200_API int pdb200_cross_entropy_with_softmax_grad(const void* label, int label_dtype, const void* softmax, const void* loss_grad,
                                           const int64_t* dims, int rank, int soft_label, int use_softmax, int ignore_index,
                                           int axis, int dtype, void* logits_grad, void* stream);

/* ---------------------------------------------------------------------------
 * layer_norm forward.  Replaces phi::LayerNormKernel<T, GPUContext>
 * (paddle/phi/kernels/gpu/layer_norm_kernel.cu:490-665).
 * rows = prod(dims[:begin_norm_axis]), D = prod(dims[begin_norm_axis:]).  scale / bias may be NULL;
 * param_dtype is `dtype` or PDB200_FLOAT32 and applies to both (:516-538).  y has x's dtype; mean and
 * variance are FLOAT32 for fp16/bf16/fp32 inputs, FLOAT64 for FLOAT64 (LayerNormParamType,
 * funcs/layer_norm_impl.cu.h:40-43) and hold `rows` values (biased variance, not rstd).
 * mean / variance may be NULL when the caller does not need them.
 * ------------------------------------------------------------------------- */
PDB200_API int pdb200_layer_norm(const void* x, const int64_t* dims, int rank, const void* scale,
                      const void* bias, int param_dtype, float epsilon, int begin_norm_axis,
                      int dtype, void* y, void* mean, void* variance, void* stream);

/* ---------------------------------------------------------------------------
 * Backward kernels.
 * softmax_grad: x_grad = (out_grad - sum(out * out_grad, axis)) * out.  Replaces phi::SoftmaxGradGPUDNNKernel /
 *   SoftmaxGradKernel<T, GPUContext> (paddle/phi/kernels/gpudnn/softmax_grad_kernel.cu ->
 *   softmax_gpudnn.h:1331-1407); numerics of paddle/phi/kernels/funcs/softmax_impl.h:275-308,406-445.
 *   dtypes FLOAT32/FLOAT64/FLOAT16/BFLOAT16; 0-D yields 0; numel == 0 is a no-op; x_grad may alias out_grad.
 * layer_norm_grad: Replaces phi::LayerNormGradKernel<T, GPUContext> (paddle/phi/kernels/gpu/layer_norm_grad_kernel.cu);
 *   formulas of paddle/phi/kernels/cpu/layer_norm_grad_kernel.cc:28-155.  mean / variance are the forward
 *   kernel's outputs (FLOAT32, or FLOAT64 for FLOAT64 x).  scale may be NULL; x_grad, scale_grad, bias_grad
 *   may each be NULL (at least one is required); scale_grad / bias_grad have param_dtype (= dtype or FLOAT32).
 *   Unlike the CPU kernel, scale is applied whenever it is given, not only when scale_grad is requested.
 *   Scratch: pdb200_layer_norm_grad_workspace_bytes() bytes when scale_grad or bias_grad is requested.
 * ------------------------------------------------------------------------- */
PDB200_API int pdb200_softmax_grad(const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                        int dtype, void* x_grad, void* stream);
/* Backward of log_softmax on the same kernels: s = sum over `axis` of out_grad, x_grad = out_grad - exp(out) * s.
 * Replaces phi::LogSoftmaxGradKernel<T, GPUContext> (paddle/phi/kernels/gpu/log_softmax_grad_kernel.cu:25-60; the CPU
 * functor the oracle is built from: cpu/log_softmax_grad_kernel.cc:30-62).  Same arguments, dtypes, routes and 0-D rule
 * (x_grad = 0) as pdb200_softmax_grad. */
PDB200_API int pdb200_log_softmax_grad(const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                            int dtype, void* x_grad, void* stream);
PDB200_API size_t pdb200_layer_norm_grad_workspace_bytes(const int64_t* dims, int rank, int begin_norm_axis, int dtype);
PDB200_API int pdb200_layer_norm_grad(const void* x, const int64_t* dims, int rank, const void* scale, int param_dtype,
                           const void* mean, const void* variance, const void* out_grad, float epsilon,
                           int begin_norm_axis, int dtype, void* x_grad, void* scale_grad, void* bias_grad,
                           void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------
 * rms_norm forward with the optional fused bias + residual add.  Replaces phi::RmsNormKernel<T,
 * GPUContext> without quantisation (paddle/phi/kernels/gpu/rms_norm_kernel.cu:1080-1260; python:
 * paddle.incubate.nn.functional.fused_rms_norm):
 *   v = x (+ residual + bias); residual_out = v (rounded to the dtype) when a residual is given;
 *   inv_var[row] = rsqrt(mean(v^2) + epsilon) (float, optional); out = (v * inv_var) * norm_weight (+ norm_bias).
 * rows = prod(dims[:begin_norm_axis]), D = prod(dims[begin_norm_axis:]); bias, norm_weight, norm_bias have D elements of
 * the tensor dtype; bias, residual, norm_bias, inv_var may be NULL (bias only together with residual).  residual_out may
 * be the residual buffer.  dtypes FLOAT32, FLOAT16, BFLOAT16 (computed in float).  The int8 / fp8 quantised outputs of
 * the reference stay with the in-tree kernel. */
PDB200_API int pdb200_rms_norm(const void* x, const int64_t* dims, int rank, const void* bias, const void* residual,
                    const void* norm_weight, const void* norm_bias, float epsilon, int begin_norm_axis, int dtype,
                    void* out, void* residual_out, void* inv_var, void* stream);

/* ---------------------------------------------------------------------------
 * Masked softmax of attention scores.
 * pdb200_fused_softmax_mask replaces phi::fusion::FusedSoftmaxMaskKernel
 * (paddle/phi/kernels/fusion/gpu/fused_softmax_mask_kernel.cu:149-316; python paddle.incubate.softmax_mask_fuse):
 *   out[b,h,q,:] = softmax(float(x[b,h,q,:]) + float(mask[b,0,q,:]))   x_dims = [B,H,Sq,Sk], mask_dims = [B,1,Sq,Sk]
 * computed in float with plain exp(d - max) (no clip, as the reference kernel), rounded once.  dtype FLOAT32 / FLOAT16 /
 * BFLOAT16 (the reference registers float and float16); mask_dtype = dtype or FLOAT32.  Checks reproduced: mask's
 * second dim is 1, the other dims equal x's (:177-202).  The reference's 32 <= Sk < 8192 and Sq-divisibility limits
 * belong to its warp-per-row layout and are not imposed.
 * pdb200_fused_softmax_mask_upper_triangle replaces phi::fusion::FusedSoftmaxMaskFuseUpperTriangleKernel
 * (fusion/gpu/fused_softmax_mask_upper_triangle_kernel.cu:255-400; python paddle.incubate.softmax_mask_fuse_upper_triangle):
 *   out[b,h,q,k] = softmax over k <= q of x[b,h,q,k], 0 for k > q      x_dims = [B,H,S,S] (Sq == Sk is checked, :271-277)
 * Only the part of x on and below the diagonal is read.  out must not alias x. */
PDB200_API int pdb200_fused_softmax_mask(const void* x, const int64_t* x_dims, const void* mask, const int64_t* mask_dims,
                              int dtype, int mask_dtype, void* out, void* stream);
PDB200_API int pdb200_fused_softmax_mask_upper_triangle(const void* x, const int64_t* x_dims, int dtype, void* out,
                                             void* stream);

/* ---------------------------------------------------------------------------
 * layer_norm(residual + dropout(x + bias)) in one pass.
 * Replaces phi::fusion::FusedBiasDropoutResidualLnKernel<T, GPUContext>
 * (paddle/phi/kernels/fusion/gpu/fused_bias_dropout_residual_layer_norm_kernel.cu:24-98; element rule
 * fusion/gpu/fused_residual_dropout_bias.h:40-156, statistics and output fused_layernorm_residual_dropout_bias.h:117-221;
 * python paddle.incubate.nn.functional.fused_bias_dropout_residual_layer_norm):
 *   d = (T(x) + T(bias)) * factor + residual, each step rounded to T;  bias_dropout_residual_out = d;
 *   ln_mean / ln_variance = mean and biased variance of float(d) over the last axis (float; double for FLOAT64);
 *   y = T(ln_scale * (d - mean) * rsqrt(variance + ln_epsilon) + ln_bias).
 * dims / rank describe x (and residual, y, bias_dropout_residual_out); D = dims[rank-1]; bias, ln_scale, ln_bias have D
 * elements and may be NULL; param_dtype = dtype of ln_scale / ln_bias: the tensor dtype or FLOAT32 (the reference uses
 * LayerNormParamType<T> = float for float16).  The deterministic cases run here: dropout_rate == 0, or is_test != 0
 * (factor = 1 resp. 1 - dropout_rate for downscale_in_infer, fused_dropout_common.h:171-181); no dropout mask is written
 * in those cases (neither does the reference).  Training with dropout_rate > 0 returns PDB200_UNSUPPORTED: the random
 * mask comes from the framework's Philox generator state (the PHI binding forwards that case to the in-tree kernel).
 * dtypes FLOAT32, FLOAT64, FLOAT16 (the reference's list) and BFLOAT16.  Outputs must not alias inputs or each other. */
PDB200_API int pdb200_fused_bias_dropout_residual_layer_norm(const void* x, const void* residual, const int64_t* dims, int rank,
                                                  const void* bias, const void* ln_scale, const void* ln_bias,
                                                  int param_dtype, float dropout_rate, int is_test, int upscale_in_train,
                                                  float ln_epsilon, int dtype, void* y, void* bias_dropout_residual_out,
                                                  void* ln_mean, void* ln_variance, void* stream);

/* layer_norm(x + bias + residual_alpha * residual) in one pass — paddle.incubate.nn.functional.fused_layer_norm.
 * Replaces phi::fusion::FusedLayerNormKernel<T, GPUContext> (paddle/phi/kernels/fusion/gpu/fused_layernorm_kernel.cu:996-1216,
 * registered as fused_bias_residual_layernorm :1222-1260), its three unquantised forms:
 *   residual != NULL, norm_weight == norm_bias == NULL:  out = (x + bias) + residual * T(residual_alpha), every step rounded
 *       to T (ResidualDropoutBias, fused_residual_dropout_bias.h:40-156, dropout off); no other output is written;
 *   residual != NULL, a norm parameter given:  residual_out = the same sum; mean / variance (float, one per row) = mean and
 *       biased variance of float(residual_out); out = T(norm_weight * (residual_out - mean) * rsqrt(variance + epsilon) +
 *       norm_bias)  (LayernormResidualDropoutBias, fused_dropout_helper.h:428-489);
 *   residual == NULL:  out, mean, variance = layer_norm(x) with the float parameters; bias is not read (:1204-1213).
 * dims / rank describe x; rows = prod(dims[:begin_norm_axis]), the rest is normalised.  bias has the row length and x's
 * dtype; norm_weight / norm_bias are FLOAT32 (LayerNormParamType<T>) and may each be NULL.  dtypes FLOAT32, FLOAT16,
 * BFLOAT16.  The quantised outputs (int8 / float8 out with quant_scale > 0) are not on this path: the PHI binding forwards
 * them to the in-tree kernel.  Outputs must not alias inputs or each other. */
PDB200_API int pdb200_fused_bias_residual_layernorm(const void* x, const int64_t* dims, int rank, const void* bias,
                                         const void* residual, const void* norm_weight, const void* norm_bias,
                                         float epsilon, float residual_alpha, int begin_norm_axis, int dtype, void* out,
                                         void* residual_out, void* mean, void* variance, void* stream);

/* ---------------------------------------------------------------------------
 * The exchange step of sharded reductions over NVLink peer memory — one process per GPU.
 * Replaces paddle.distributed.all_reduce -> ProcessGroupNCCL::AllReduce -> ncclAllReduce
 * (paddle/fluid/distributed/collective/process_group_nccl.cc:267-297; ReduceScatter :467-500; the bucketed gradient
 * reduce of paddle/fluid/distributed/collective/reducer.cc:1089-1111) for what this path exchanges: one scalar per rank
 * for a full reduce, a vector of column sums for a row-sharded sum(axis=0).
 * Each rank creates a window (device memory: control block + two slots + two result regions of `slot_bytes` each),
 * sends its 64-byte IPC handle to the other ranks (any transport; the Python layer uses torch.distributed.all_gather
 * once, at set-up) and opens theirs.
 * Per call: write the local partial result into pdb200_peer_slot(window, slot_bytes, epoch) — e.g. pass it as the `out`
 * of a reduction, in `slot_dtype` (FLOAT32 partials for a BFLOAT16 / FLOAT16 sum: the cross-rank sum is rounded once) —
 * then pdb200_peer_all_reduce with the same epoch on every rank (1, 2, 3, ... per window).  ONE kernel publishes the
 * epoch to all peers, waits for theirs and combines all slots in rank order (bit-identical on every rank): up to 64 KB
 * every rank reads every slot (one NVLink round trip); beyond that rank r reduces only its 1/world of the columns and
 * stores the result chunk into every rank's result region (reduce-scatter + all-gather: each byte crosses NVLink once
 * per direction), and `out` is filled from the local result region.  op codes: 0 sum, 1 max, 2 min, 3 prod.
 * (slot_dtype, out_dtype): equal, or FLOAT32 -> BFLOAT16 / FLOAT16.  `out` must be aligned to its element size
 * (16 bytes for the fast paths).
 * The wait is bounded in time (30 s; PDB200_PEER_TIMEOUT_MS): a peer that never arrives makes the kernel fill `out`
 * with NaN (all-ones for integers) and raise the error word; every later call on that window then fails with
 * PDB200_CUDA_ERROR without touching the GPU — the group must be rebuilt (pdb200_peer_error reads the word at once,
 * synchronising).  Launch-order contract: every rank enqueues its exchanges in the same order, one process per GPU,
 * and nothing enqueued before rank B's exchange may depend on work rank A enqueues after its own.
 * windows[r] = this process's mapping of rank r's window (its own window for r == rank).
 * These entry points allocate (window_create) and synchronise (window_create, peer_error): set-up and checks only. */
PDB200_API size_t pdb200_peer_window_bytes(size_t slot_bytes);
PDB200_API int pdb200_peer_window_create(size_t slot_bytes, void** window, void* ipc_handle_64);
PDB200_API int pdb200_peer_window_open(const void* ipc_handle_64, void** mapped);
PDB200_API int pdb200_peer_window_close(void* mapped);
PDB200_API int pdb200_peer_window_destroy(void* window);
PDB200_API void* pdb200_peer_slot(void* window, size_t slot_bytes, unsigned long long epoch);
PDB200_API int pdb200_peer_error(void* window);
PDB200_API int pdb200_peer_all_reduce(int op, void* const* windows, int world, int rank, size_t slot_bytes,
                           unsigned long long epoch, int64_t n, int slot_dtype, int out_dtype, void* out, void* stream);
/* Self-test of the exchange on ONE GPU: all `world` windows were created by this process on the current device; the same
 * device code runs for every rank in one cooperative grid (blockIdx.y = rank), rank r's result goes to outs[r].  No
 * counterpart in the reference; it exists because ranks that wait for one another must never be separate launches on
 * one GPU.  op 0 (sum) / 1 (max); (slot, out) dtypes (FLOAT32, FLOAT32), (FLOAT32, BFLOAT16), (INT64, INT64). */
PDB200_API int pdb200_peer_all_reduce_emulated(int op, void* const* windows, int world, size_t slot_bytes,
                                    unsigned long long epoch, int64_t n, int slot_dtype, int out_dtype,
                                    void* const* outs, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PDB200_H_ */

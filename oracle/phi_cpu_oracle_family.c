/*
 * phi_cpu_oracle_family.c — TEST INFRASTRUCTURE ONLY (see phi_cpu_oracle.c for the rules: only tests/,
 * __graft_entry__.smoke() and bench.py's CPU legs may load the library this is compiled into).
 *
 * Plain-C restatement of the rest of the elementwise family of PaddlePaddle's CPU PHI kernels
 * (SURVEY.md §8f rank 1): compare, logical, bitwise, logical_not / bitwise_not, scale, cast, where.
 *
 * Parity status: PINNED — tests/test_oracle_pin.py checks every function against
 * oracle/_ref/libphi_cpu_ref.so (the reference's own CPU kernels, reached through oracle/ref_shim_family.cc)
 * and against tests/golden/phi_family_golden.npz recorded from it.  Where the CPU build registers no kernel
 * for a dtype the GPU build has (bf16 / fp16 logical ops; where on 1- and 2-byte types) the same functor is
 * applied to that dtype — an extension by analogy that the reference CPU cannot pin; it is marked below.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

enum { DT_BOOL = 0, DT_UINT8, DT_INT8, DT_INT16, DT_INT32, DT_INT64, DT_FLOAT16, DT_BFLOAT16, DT_FLOAT32, DT_FLOAT64 };
#define MAXR 9

int oracle_broadcast_shape(const int64_t* xd, int xr, const int64_t* yd, int yr, int64_t* od, int* orank);

static float bf2f(uint16_t b) { uint32_t u = (uint32_t)b << 16; float f; memcpy(&f, &u, 4); return f; }
static float h2f(uint16_t b) { _Float16 h; memcpy(&h, &b, 2); return (float)h; }
/* float -> bf16, paddle/phi/common/bfloat16.h:61-88: NaN -> 0x7FFF, else round to nearest even */
static uint16_t f2bf(float f) {
  uint32_t u; memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7FFF;
  u += 0x7FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
/* float -> fp16: round to nearest even (the GPU build's __float2half; the non-F16C CPU routine of
 * phi/common/float16.h:121-138 truncates instead — tests allow one unit in the last place for fp16 outputs) */
static uint16_t f2h(float f) { _Float16 h = (_Float16)f; uint16_t b; memcpy(&b, &h, 2); return b; }

/* ---- the broadcast walk of CommonForwardBroadcastCPU (funcs/elementwise_base.h:274-307) ------------------------ */
typedef struct { int R; int64_t od[MAXR], xs[MAXR], ys[MAXR], numel; } Walk;
static int walk_init(Walk* w, const int64_t* xd, int xr, const int64_t* yd, int yr) {
  if (oracle_broadcast_shape(xd, xr, yd, yr, w->od, &w->R)) return 1;
  int64_t sx = 1, sy = 1;
  w->numel = 1;
  for (int i = w->R - 1; i >= 0; --i) {
    int xi = i - (w->R - xr), yi = i - (w->R - yr);
    int64_t a = xi >= 0 ? xd[xi] : 1, b = yi >= 0 ? yd[yi] : 1;
    w->xs[i] = a == 1 ? 0 : sx;
    w->ys[i] = b == 1 ? 0 : sy;
    sx *= a; sy *= b;
    w->numel *= w->od[i];
  }
  return 0;
}
#define WALK(BODY)                                                                       \
  {                                                                                      \
    int64_t idx[MAXR] = {0};                                                             \
    for (int64_t i = 0; i < w.numel; ++i) {                                              \
      int64_t xo = 0, yo = 0;                                                            \
      for (int d = 0; d < w.R; ++d) { xo += idx[d] * w.xs[d]; yo += idx[d] * w.ys[d]; }  \
      BODY;                                                                              \
      for (int d = w.R - 1; d >= 0; --d) { if (++idx[d] < w.od[d]) break; idx[d] = 0; }  \
    }                                                                                    \
  }

/* ------------------------------------------------------------------------------------------
 * compare.  funcs/compare_functors.h:21-62; cpu/compare_kernel.cc:26-44 (when x has the lower rank the
 * inverse functor is applied to swapped operands — the value is the same).
 * op: 0 <, 1 <=, 2 >, 3 >=, 4 equal, 5 not_equal.  Equal on float/double: infinities exact, NaN never,
 * else fabs((double)(a - b)) < 1e-8 with a - b evaluated in T; isinf / isnan are tested on the value cast to
 * float, so a finite double beyond float range is treated as infinite.  bf16 / fp16 take the same route
 * (std::is_floating_point is specialised for them, bfloat16.h:373-379, float16.h:1019): a - b is one float
 * subtraction rounded back to the 16-bit type.
 * ---------------------------------------------------------------------------------------- */
static int eq_f32(float a, float b) {
  if (isinf(a) || isinf(b)) return a == b;
  if (isnan(a) || isnan(b)) return 0;
  volatile float d = a - b;
  return fabs((double)d) < 1e-8;
}
static int eq_f64(double a, double b) {
  float fa = (float)a, fb = (float)b;
  if (isinf(fa) || isinf(fb)) return a == b;
  if (isnan(fa) || isnan(fb)) return 0;
  return fabs(a - b) < 1e-8;
}
static int eq_bf16(float a, float b) {
  if (isinf(a) || isinf(b)) return a == b;
  if (isnan(a) || isnan(b)) return 0;
  return fabs((double)bf2f(f2bf(a - b))) < 1e-8;
}
static int eq_fp16(float a, float b) {
  if (isinf(a) || isinf(b)) return a == b;
  if (isnan(a) || isnan(b)) return 0;
  /* fp16 cannot hold a non-zero magnitude below 1e-8 (smallest subnormal 6e-8): equal iff the rounded difference is 0,
   * whichever way the conversion rounds (the truncating CPU routine flushes what it cannot represent to zero too) */
  return fabs((double)h2f(f2h(a - b))) < 1e-8;
}
#define CMP_CASES(T, LOAD, EQ)                                                         \
  {                                                                                    \
    const T* X = (const T*)x; const T* Y = (const T*)y;                                \
    switch (op) {                                                                      \
      case 0: WALK(O[i] = LOAD(X[xo]) < LOAD(Y[yo])) break;                            \
      case 1: WALK(O[i] = LOAD(X[xo]) <= LOAD(Y[yo])) break;                           \
      case 2: WALK(O[i] = LOAD(X[xo]) > LOAD(Y[yo])) break;                            \
      case 3: WALK(O[i] = LOAD(X[xo]) >= LOAD(Y[yo])) break;                           \
      case 4: WALK(O[i] = EQ(LOAD(X[xo]), LOAD(Y[yo])) ? 1 : 0) break;                 \
      case 5: WALK(O[i] = EQ(LOAD(X[xo]), LOAD(Y[yo])) ? 0 : 1) break;                 \
      default: return 2;                                                               \
    }                                                                                  \
  }
#define ID(v) (v)
#define EQ_PLAIN(a, b) ((a) == (b))
int oracle_compare(int op, int dtype, const void* x, const int64_t* xd, int xr, const void* y, const int64_t* yd,
                   int yr, uint8_t* O) {
  Walk w;
  if (walk_init(&w, xd, xr, yd, yr)) return 1;
  switch (dtype) {
    case DT_BOOL: case DT_UINT8: CMP_CASES(uint8_t, ID, EQ_PLAIN) break;
    case DT_INT8: CMP_CASES(int8_t, ID, EQ_PLAIN) break;
    case DT_INT16: CMP_CASES(int16_t, ID, EQ_PLAIN) break;
    case DT_INT32: CMP_CASES(int32_t, ID, EQ_PLAIN) break;
    case DT_INT64: CMP_CASES(int64_t, ID, EQ_PLAIN) break;
    case DT_FLOAT16: CMP_CASES(uint16_t, h2f, eq_fp16) break;
    case DT_BFLOAT16: CMP_CASES(uint16_t, bf2f, eq_bf16) break;
    case DT_FLOAT32: CMP_CASES(float, ID, eq_f32) break;
    case DT_FLOAT64: CMP_CASES(double, ID, eq_f64) break;
    default: return 2;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * logical and / or / xor: static_cast<bool>(a) op static_cast<bool>(b), funcs/logical_functor.h:20-32.
 * CPU registers float, double, bool, int64, int, int8, int16 (cpu/logical_kernel.cc:93-111); fp16 / bf16
 * (GPU only, kps/logical_kernel.cu:117-119) follow by analogy: bool(v) = (float(v) != 0).
 * ---------------------------------------------------------------------------------------- */
#define LOG_CASES(T, TRUTH)                                                            \
  {                                                                                    \
    const T* X = (const T*)x; const T* Y = (const T*)y;                                \
    switch (op) {                                                                      \
      case 0: WALK(O[i] = (TRUTH(X[xo]) && TRUTH(Y[yo])) ? 1 : 0) break;               \
      case 1: WALK(O[i] = (TRUTH(X[xo]) || TRUTH(Y[yo])) ? 1 : 0) break;               \
      case 2: WALK(O[i] = (TRUTH(X[xo]) != TRUTH(Y[yo])) ? 1 : 0) break;               \
      default: return 2;                                                               \
    }                                                                                  \
  }
#define T_PLAIN(v) ((v) != 0)
#define T_BF(v) (bf2f(v) != 0.0f)
#define T_H(v) (h2f(v) != 0.0f)
int oracle_logical(int op, int dtype, const void* x, const int64_t* xd, int xr, const void* y, const int64_t* yd,
                   int yr, uint8_t* O) {
  Walk w;
  if (walk_init(&w, xd, xr, yd, yr)) return 1;
  switch (dtype) {
    case DT_BOOL: LOG_CASES(uint8_t, T_PLAIN) break;
    case DT_INT8: LOG_CASES(int8_t, T_PLAIN) break;
    case DT_INT16: LOG_CASES(int16_t, T_PLAIN) break;
    case DT_INT32: LOG_CASES(int32_t, T_PLAIN) break;
    case DT_INT64: LOG_CASES(int64_t, T_PLAIN) break;
    case DT_FLOAT16: LOG_CASES(uint16_t, T_H) break;
    case DT_BFLOAT16: LOG_CASES(uint16_t, T_BF) break;
    case DT_FLOAT32: LOG_CASES(float, T_PLAIN) break;
    case DT_FLOAT64: LOG_CASES(double, T_PLAIN) break;
    default: return 2;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * bitwise and / or / xor: a & b, a | b, a ^ b; bool: &&, ||, != (funcs/bitwise_functors.h:20-35).
 * ---------------------------------------------------------------------------------------- */
#define BIT_CASES(T)                                                                   \
  {                                                                                    \
    const T* X = (const T*)x; const T* Y = (const T*)y; T* Q = (T*)out;                \
    switch (op) {                                                                      \
      case 0: WALK(Q[i] = (T)(X[xo] & Y[yo])) break;                                   \
      case 1: WALK(Q[i] = (T)(X[xo] | Y[yo])) break;                                   \
      case 2: WALK(Q[i] = (T)(X[xo] ^ Y[yo])) break;                                   \
      default: return 2;                                                               \
    }                                                                                  \
  }
int oracle_bitwise(int op, int dtype, const void* x, const int64_t* xd, int xr, const void* y, const int64_t* yd,
                   int yr, void* out) {
  Walk w;
  if (walk_init(&w, xd, xr, yd, yr)) return 1;
  switch (dtype) {
    case DT_BOOL: {
      const uint8_t* X = (const uint8_t*)x; const uint8_t* Y = (const uint8_t*)y; uint8_t* Q = (uint8_t*)out;
      switch (op) {
        case 0: WALK(Q[i] = (X[xo] && Y[yo]) ? 1 : 0) break;
        case 1: WALK(Q[i] = (X[xo] || Y[yo]) ? 1 : 0) break;
        case 2: WALK(Q[i] = ((X[xo] != 0) != (Y[yo] != 0)) ? 1 : 0) break;
        default: return 2;
      }
      break;
    }
    case DT_UINT8: BIT_CASES(uint8_t) break;
    case DT_INT8: BIT_CASES(int8_t) break;
    case DT_INT16: BIT_CASES(int16_t) break;
    case DT_INT32: BIT_CASES(int32_t) break;
    case DT_INT64: BIT_CASES(int64_t) break;
    default: return 2;
  }
  return 0;
}

/* logical_not: !a (funcs/logical_functor.h:34-38) -> bool; bitwise_not: ~a, bool: !a (bitwise_functors.h:38-49).
 * which: 0 logical_not, 1 bitwise_not. */
int oracle_not(int which, int dtype, const void* x, int64_t n, void* out) {
  uint8_t* O = (uint8_t*)out;
#define LNOT(T, TRUTH) { const T* X = (const T*)x; for (int64_t i = 0; i < n; ++i) O[i] = TRUTH(X[i]) ? 0 : 1; }
#define BNOT(T) { const T* X = (const T*)x; T* Q = (T*)out; for (int64_t i = 0; i < n; ++i) Q[i] = (T)~X[i]; }
  if (which == 0) {
    switch (dtype) {
      case DT_BOOL: LNOT(uint8_t, T_PLAIN) break;
      case DT_INT8: LNOT(int8_t, T_PLAIN) break;
      case DT_INT16: LNOT(int16_t, T_PLAIN) break;
      case DT_INT32: LNOT(int32_t, T_PLAIN) break;
      case DT_INT64: LNOT(int64_t, T_PLAIN) break;
      case DT_FLOAT16: LNOT(uint16_t, T_H) break;
      case DT_BFLOAT16: LNOT(uint16_t, T_BF) break;
      case DT_FLOAT32: LNOT(float, T_PLAIN) break;
      case DT_FLOAT64: LNOT(double, T_PLAIN) break;
      default: return 2;
    }
  } else {
    switch (dtype) {
      case DT_BOOL: LNOT(uint8_t, T_PLAIN) break;
      case DT_UINT8: BNOT(uint8_t) break;
      case DT_INT8: BNOT(int8_t) break;
      case DT_INT16: BNOT(int16_t) break;
      case DT_INT32: BNOT(int32_t) break;
      case DT_INT64: BNOT(int64_t) break;
      default: return 2;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * scale.  phi::ScaleKernel<T,CPUContext> (cpu/scale_kernel.cc:27-48) -> EigenScale (funcs/eigen/scale.cc:22-40):
 *   out = scale * x + bias  or  scale * (x + bias), every operation in T, with scale.to<T>() / bias.to<T>()
 *   (= static_cast<T> of the Python float, which arrives as a double).
 *   bf16 / fp16: T arithmetic = float operation + rounding per operation (bfloat16.h:264-292).
 *   Integers wrap.  bool: (s && x) || b, resp. s && (x || b).
 *   float / double: whether the multiply-add is fused is the host compiler's choice in the reference build
 *   (-mfma, GNU C++ contracts by default); this restatement does not fuse — tests compare to one unit in
 *   the last place.
 * ---------------------------------------------------------------------------------------- */
int oracle_scale(int dtype, const void* x, int64_t n, double scale, double bias, int after, void* out) {
#define SC_INT(T, U) { const T* X = (const T*)x; T* Q = (T*)out; const U s = (U)(T)scale, b = (U)(T)bias;           \
    for (int64_t i = 0; i < n; ++i) Q[i] = after ? (T)(U)(s * (U)X[i] + b) : (T)(U)(s * (U)((U)X[i] + b)); }
  switch (dtype) {
    case DT_BOOL: {
      const uint8_t* X = (const uint8_t*)x; uint8_t* Q = (uint8_t*)out;
      const int s = scale != 0.0, b = bias != 0.0;
      for (int64_t i = 0; i < n; ++i) Q[i] = after ? ((s && X[i]) || b) : (s && (X[i] || b));
      break;
    }
    case DT_UINT8: SC_INT(uint8_t, uint8_t) break;
    case DT_INT8: SC_INT(int8_t, uint8_t) break;
    case DT_INT16: SC_INT(int16_t, uint16_t) break;
    case DT_INT32: SC_INT(int32_t, uint32_t) break;
    case DT_INT64: SC_INT(int64_t, uint64_t) break;
    case DT_BFLOAT16: {
      const uint16_t* X = (const uint16_t*)x; uint16_t* Q = (uint16_t*)out;
      const float s = bf2f(f2bf((float)scale)), b = bf2f(f2bf((float)bias));
      for (int64_t i = 0; i < n; ++i)
        Q[i] = after ? f2bf(bf2f(f2bf(s * bf2f(X[i]))) + b) : f2bf(s * bf2f(f2bf(bf2f(X[i]) + b)));
      break;
    }
    case DT_FLOAT16: {
      const uint16_t* X = (const uint16_t*)x; uint16_t* Q = (uint16_t*)out;
      const float s = h2f(f2h((float)scale)), b = h2f(f2h((float)bias));
      for (int64_t i = 0; i < n; ++i)
        Q[i] = after ? f2h(h2f(f2h(s * h2f(X[i]))) + b) : f2h(s * h2f(f2h(h2f(X[i]) + b)));
      break;
    }
    case DT_FLOAT32: {
      const float* X = (const float*)x; float* Q = (float*)out;
      const float s = (float)scale, b = (float)bias;
      for (int64_t i = 0; i < n; ++i) Q[i] = after ? s * X[i] + b : s * (X[i] + b);
      break;
    }
    case DT_FLOAT64: {
      const double* X = (const double*)x; double* Q = (double*)out;
      for (int64_t i = 0; i < n; ++i) Q[i] = after ? scale * X[i] + bias : scale * (X[i] + bias);
      break;
    }
    default: return 2;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * cast.  phi::CastKernel -> CastKernelImpl -> CastOpTransformFunctor: static_cast<OutT>(in) per element
 * (cpu/cast_impl.h:22-40).  bf16 / fp16 sources convert through float, every source reaches bf16 / fp16
 * through float (the constructors of phi::dtype::bfloat16 / float16).  To bool: != 0.
 * Float -> integer truncates toward zero; a value outside the target range is undefined behaviour in the
 * reference and is kept out of the test inputs.  Narrow integer targets keep the low bits.
 * ---------------------------------------------------------------------------------------- */
typedef struct { int is_float; double f; int64_t i; uint64_t u; int is_unsigned; } Num;
static Num load_num(const void* x, int dt, int64_t k) {
  Num v = {0, 0.0, 0, 0, 0};
  switch (dt) {
    case DT_BOOL: v.i = ((const uint8_t*)x)[k] != 0; break;
    case DT_UINT8: v.i = ((const uint8_t*)x)[k]; break;
    case DT_INT8: v.i = ((const int8_t*)x)[k]; break;
    case DT_INT16: v.i = ((const int16_t*)x)[k]; break;
    case DT_INT32: v.i = ((const int32_t*)x)[k]; break;
    case DT_INT64: v.i = ((const int64_t*)x)[k]; break;
    case DT_FLOAT16: v.is_float = 1; v.f = h2f(((const uint16_t*)x)[k]); break;
    case DT_BFLOAT16: v.is_float = 1; v.f = bf2f(((const uint16_t*)x)[k]); break;
    case DT_FLOAT32: v.is_float = 1; v.f = ((const float*)x)[k]; break;
    case DT_FLOAT64: v.is_float = 2; v.f = ((const double*)x)[k]; break;
  }
  return v;
}
int oracle_cast(int x_dtype, int out_dtype, const void* x, int64_t n, void* out) {
  if (x_dtype < 0 || x_dtype > DT_FLOAT64 || out_dtype < 0 || out_dtype > DT_FLOAT64) return 2;
  for (int64_t k = 0; k < n; ++k) {
    const Num v = load_num(x, x_dtype, k);
    /* the value as the source type would convert it */
    const float as_f = v.is_float ? (float)v.f : (float)v.i;     /* int64 -> float: round to nearest */
    const double as_d = v.is_float ? v.f : (double)v.i;
    const int64_t as_i = v.is_float ? (int64_t)v.f : v.i;         /* truncation toward zero */
    switch (out_dtype) {
      case DT_BOOL: ((uint8_t*)out)[k] = v.is_float ? (v.f != 0.0) : (v.i != 0); break;
      case DT_UINT8: ((uint8_t*)out)[k] = (uint8_t)(v.is_float ? (int64_t)(int32_t)v.f : as_i); break;
      case DT_INT8: ((int8_t*)out)[k] = (int8_t)(v.is_float ? (int64_t)(int32_t)v.f : as_i); break;
      case DT_INT16: ((int16_t*)out)[k] = (int16_t)(v.is_float ? (int64_t)(int32_t)v.f : as_i); break;
      case DT_INT32: ((int32_t*)out)[k] = v.is_float ? (int32_t)v.f : (int32_t)as_i; break;
      case DT_INT64: ((int64_t*)out)[k] = as_i; break;
      case DT_FLOAT16: ((uint16_t*)out)[k] = x_dtype == DT_FLOAT16 ? ((const uint16_t*)x)[k] : f2h(as_f); break;
      case DT_BFLOAT16: ((uint16_t*)out)[k] = x_dtype == DT_BFLOAT16 ? ((const uint16_t*)x)[k] : f2bf(as_f); break;
      case DT_FLOAT32: ((float*)out)[k] = as_f; break;
      case DT_FLOAT64: ((double*)out)[k] = as_d; break;
    }
  }
  return 0;
}

/* where: out[i] = cond[i] ? x[i] : y[i] (cpu/where_kernel.cc:20-36).  CPU registers float, double, int, int64;
 * the select copies bits, so the other element sizes follow without any arithmetic. */
int oracle_where(int dtype, const uint8_t* cond, const void* x, const void* y, int64_t n, void* out) {
  size_t es;
  switch (dtype) {
    case DT_BOOL: case DT_UINT8: case DT_INT8: es = 1; break;
    case DT_INT16: case DT_FLOAT16: case DT_BFLOAT16: es = 2; break;
    case DT_INT32: case DT_FLOAT32: es = 4; break;
    case DT_INT64: case DT_FLOAT64: es = 8; break;
    default: return 2;
  }
  for (int64_t i = 0; i < n; ++i)
    memcpy((char*)out + i * es, (cond[i] ? (const char*)x : (const char*)y) + i * es, es);
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * add_grad / multiply_grad with broadcasting.
 * phi::AddGradKernel / MultiplyGradKernel<T,CPUContext> (cpu/elementwise_add_grad_kernel.cc:25-62,
 * cpu/elementwise_multiply_grad_kernel.cc:26-39) -> ElemwiseGradCompute / ElemwiseExplicitGradCompute
 * (funcs/elementwise_grad_base.h:1802-1821).  add: dx_op = dy_op = dout (IdentityGrad); multiply:
 * dx_op = dout * y, dy_op = dout * x (MulGradDX / MulGradDY, cpu/elementwise_grad.h), each a T-typed product.
 * Route (elementwise_grad_base.h:266-355, funcs/elementwise_utils.h:40-75 GetMidDims):
 *   - the lower-rank operand, trailing 1s trimmed, equals a contiguous run of the other's dims starting at
 *     axis = |rank_x - rank_y|  ->  (pre, n, post) loops (:103-211): the full operand's gradient is the plain
 *     product, the broadcast operand's is accumulated per element in MPType (float for bf16) over increasing
 *     dout index and rounded once;
 *   - anything else -> CommonGradBroadcastCPU (:52-98): both gradients start at 0 and take "+=" in T, in
 *     increasing dout index (bf16: rounded after every addition).
 * float / double: the reference build (-O3 -mfma, GNU contraction default) fuses "sum += a * b" into one fma in
 *   every loop except the (pre, n, post > 1) loop compiled for "full operand's gradient not wanted", which GCC
 *   vectorises as multiply-then-ordered-add.  This restatement always fuses: bit-identical to oracle/_ref
 *   everywhere else, within a few units in the last place there (tests/test_oracle_pin_family.py states both).
 * acc_mode 1 (not a CPU route): the GPU kernels' arithmetic for every shape — products rounded to T, sums in
 *   MPType, rounded once (gpu/elementwise_grad.h:57-107 + ReduceKernel).  It differs from acc_mode 0 only for
 *   16-bit types on the common route.  fp16 exists only in this mode (no CPU registration).
 * dx / dy may be NULL.  op: 0 add_grad, 1 multiply_grad.
 * ---------------------------------------------------------------------------------------- */
#include <stdlib.h>
static int grad_common_route(const int64_t* xd, int xr, const int64_t* yd, int yr) {
  const int x_larger = xr >= yr;
  const int64_t* bd = x_larger ? xd : yd; const int br = x_larger ? xr : yr;
  const int64_t* sd = x_larger ? yd : xd; int sr = x_larger ? yr : xr;
  const int axis0 = br - sr;
  while (sr > 0 && sd[sr - 1] == 1) --sr;            /* TrimTrailingSingularDims */
  const int axis = sr == 0 ? br : axis0;
  for (int i = 0; i < sr; ++i) if (bd[i + axis] != sd[i]) return 1;
  return 0;
}
int oracle_elementwise_grad(int op, int dtype, int acc_mode, const void* x, const int64_t* xd, int xr, const void* y,
                            const int64_t* yd, int yr, const void* dout, void* dx, void* dy) {
  Walk w;
  if (walk_init(&w, xd, xr, yd, yr)) return 1;
  int64_t xn = 1, yn = 1;
  for (int i = 0; i < xr; ++i) xn *= xd[i];
  for (int i = 0; i < yr; ++i) yn *= yd[i];
#define GRAD_NATIVE(T, ACC_ADD, ACC_MULADD)                                                    \
  {                                                                                            \
    const T* X = (const T*)x; const T* Y = (const T*)y; const T* G = (const T*)dout;           \
    T* DX = (T*)dx; T* DY = (T*)dy;                                                            \
    if (DX) memset(DX, 0, (size_t)xn * sizeof(T));                                             \
    if (DY) memset(DY, 0, (size_t)yn * sizeof(T));                                             \
    WALK({                                                                                     \
      if (DX) DX[xo] = op == 0 ? ACC_ADD(DX[xo], G[i]) : ACC_MULADD(G[i], Y[yo], DX[xo]);      \
      if (DY) DY[yo] = op == 0 ? ACC_ADD(DY[yo], G[i]) : ACC_MULADD(G[i], X[xo], DY[yo]);      \
    })                                                                                         \
  }
#define F_ADD(a, b) ((a) + (b))
#define I32_ADD(a, b) ((int32_t)((uint32_t)(a) + (uint32_t)(b)))
#define I32_MULADD(a, b, c) ((int32_t)((uint32_t)(a) * (uint32_t)(b) + (uint32_t)(c)))
#define I64_ADD(a, b) ((int64_t)((uint64_t)(a) + (uint64_t)(b)))
#define I64_MULADD(a, b, c) ((int64_t)((uint64_t)(a) * (uint64_t)(b) + (uint64_t)(c)))
  switch (dtype) {
    case DT_FLOAT32: GRAD_NATIVE(float, F_ADD, fmaf) break;
    case DT_FLOAT64: GRAD_NATIVE(double, F_ADD, fma) break;
    case DT_INT32: GRAD_NATIVE(int32_t, I32_ADD, I32_MULADD) break;
    case DT_INT64: GRAD_NATIVE(int64_t, I64_ADD, I64_MULADD) break;
    case DT_BFLOAT16: case DT_FLOAT16: {
      const int bf = dtype == DT_BFLOAT16;
      if (!bf && !acc_mode) return 2;
      const int per_add = !acc_mode && grad_common_route(xd, xr, yd, yr);
      const uint16_t* X = (const uint16_t*)x; const uint16_t* Y = (const uint16_t*)y; const uint16_t* G = (const uint16_t*)dout;
      uint16_t* DX = (uint16_t*)dx; uint16_t* DY = (uint16_t*)dy;
      float* ax = DX ? (float*)calloc((size_t)xn, sizeof(float)) : NULL;
      float* ay = DY ? (float*)calloc((size_t)yn, sizeof(float)) : NULL;
#define H2F(v) (bf ? bf2f(v) : h2f(v))
#define F2H(v) (bf ? f2bf(v) : f2h(v))
      WALK({
        if (DX) { float v = op == 0 ? H2F(G[i]) : H2F(F2H(H2F(G[i]) * H2F(Y[yo]))); ax[xo] = per_add ? H2F(F2H(ax[xo] + v)) : ax[xo] + v; }
        if (DY) { float v = op == 0 ? H2F(G[i]) : H2F(F2H(H2F(G[i]) * H2F(X[xo]))); ay[yo] = per_add ? H2F(F2H(ay[yo] + v)) : ay[yo] + v; }
      })
      if (DX) { for (int64_t k = 0; k < xn; ++k) DX[k] = F2H(ax[k]); free(ax); }
      if (DY) { for (int64_t k = 0; k < yn; ++k) DY[k] = F2H(ay[k]); free(ay); }
      break;
    }
    default: return 2;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * sum_grad / mean_grad.  phi::ReduceSumGradKernel / ReduceMeanGradKernel<T,CPUContext>
 * (cpu/reduce_sum_grad_kernel.cc:26-43, cpu/reduce_mean_grad_kernel.cc) -> ReduceGradKernel (impl/reduce_grad.h) ->
 * funcs::SumGradFunctor / MeanGradFunctor (funcs/reduce_functor.h:91-123):
 *     dx = dy.broadcast(dim)                 dx = dy.broadcast(dim) / dx.constant(size)
 * i.e. every x_grad element is its out_grad element (mean: divided by the number of reduced elements, in T;
 * integers divide truncating).  bf16 / fp16 mean_grad (no CPU kernel): float division, one rounding.
 * red[i] != 0 marks a reduced dimension of x.  op: 0 sum_grad, 1 mean_grad.
 * ---------------------------------------------------------------------------------------- */
int oracle_reduce_grad(int op, int dtype, const int64_t* xd, int xr, const int* red, const void* out_grad, void* x_grad) {
  int64_t gs[MAXR], idx[MAXR] = {0}, numel = 1, n = 1, s = 1;
  for (int i = xr - 1; i >= 0; --i) {
    gs[i] = red[i] ? 0 : s;
    if (!red[i]) s *= xd[i]; else n *= xd[i];
    numel *= xd[i];
  }
  for (int64_t i = 0; i < numel; ++i) {
    int64_t go = 0;
    for (int d = 0; d < xr; ++d) go += idx[d] * gs[d];
    switch (dtype) {
      case DT_FLOAT32: { float v = ((const float*)out_grad)[go]; ((float*)x_grad)[i] = op ? v / (float)n : v; break; }
      case DT_FLOAT64: { double v = ((const double*)out_grad)[go]; ((double*)x_grad)[i] = op ? v / (double)n : v; break; }
      case DT_INT32: { int32_t v = ((const int32_t*)out_grad)[go]; ((int32_t*)x_grad)[i] = op ? (int32_t)(v / (int32_t)n) : v; break; }
      case DT_INT64: { int64_t v = ((const int64_t*)out_grad)[go]; ((int64_t*)x_grad)[i] = op ? v / n : v; break; }
      case DT_BFLOAT16: { uint16_t v = ((const uint16_t*)out_grad)[go]; ((uint16_t*)x_grad)[i] = op ? f2bf(bf2f(v) / (float)n) : v; break; }
      case DT_FLOAT16: { uint16_t v = ((const uint16_t*)out_grad)[go]; ((uint16_t*)x_grad)[i] = op ? f2h(h2f(v) / (float)n) : v; break; }
      default: return 2;
    }
    for (int d = xr - 1; d >= 0; --d) { if (++idx[d] < xd[d]) break; idx[d] = 0; }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * cross_entropy_with_softmax_grad, use_softmax = true.  phi::CrossEntropyWithSoftmaxGradKernel<T,CPUContext>
 * (cpu/cross_entropy_grad_kernel.cc:143-179) on [n][axis_dim][remain], p = softmax, g = loss_grad[i][j]:
 *   soft: dx = g * (p - y)
 *   hard: dx = p * g; dx[label] -= g (a second rounding); label == ignore_index: the whole position 0.   (label: int64 here)
 * T = float or double.
 * ---------------------------------------------------------------------------------------- */
int oracle_cross_entropy_with_softmax_grad(int dtype, const void* label, const void* softmax, const void* loss_grad,
                                           const int64_t* dims, int rank, int soft_label, int ignore_index, int axis,
                                           void* logits_grad) {
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  if (rank < 1 || axis < -rank || axis >= rank) return 1;
  if (axis < 0) axis += rank;
  int64_t n = 1, ad = dims[axis], rem = 1;
  for (int i = 0; i < axis; ++i) n *= dims[i];
  for (int i = axis + 1; i < rank; ++i) rem *= dims[i];
  for (int64_t b = 0; b < n; ++b)
    for (int64_t j = 0; j < rem; ++j) {
      const int64_t pos = b * rem + j;
      const int64_t lbl = soft_label ? -1 : ((const int64_t*)label)[pos];
      for (int64_t k = 0; k < ad; ++k) {
        const int64_t i = (b * ad + k) * rem + j;
        if (dtype == DT_FLOAT32) {
          const float p = ((const float*)softmax)[i], g = ((const float*)loss_grad)[pos];
          volatile float r;
          if (soft_label) { r = p - ((const float*)label)[i]; r = g * r; }
          else { r = p * g; if (k == lbl) r = r - g; if (lbl == ignore_index) r = 0.f; }
          ((float*)logits_grad)[i] = r;
        } else {
          const double p = ((const double*)softmax)[i], g = ((const double*)loss_grad)[pos];
          volatile double r;
          if (soft_label) { r = p - ((const double*)label)[i]; r = g * r; }
          else { r = p * g; if (k == lbl) r = r - g; if (lbl == ignore_index) r = 0.0; }
          ((double*)logits_grad)[i] = r;
        }
      }
    }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * log_softmax_grad.  phi::LogSoftmaxGradKernel<T,CPUContext> -> LogSoftmaxGradFunctor
 * (cpu/log_softmax_grad_kernel.cc:30-62): x viewed as [n][axis_dim][remain];
 *     dx = dy - exp(y) * sum_along_axis(dy)
 * 0-D -> 0; 0-size -> no-op.  The sum runs along the axis in increasing index (Eigen's reducer order only moves the last
 * bits: tests compare to 1e-5 of the row scale).  T = float or double.
 * ---------------------------------------------------------------------------------------- */
int oracle_log_softmax_grad(int dtype, const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                            void* x_grad) {
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  if (rank == 0) { if (dtype == DT_FLOAT32) *(float*)x_grad = 0.f; else *(double*)x_grad = 0.0; return 0; }
  if (axis < -rank || axis >= rank) return 1;
  if (axis < 0) axis += rank;
  int64_t n = 1, ad = dims[axis], rem = 1;
  for (int i = 0; i < axis; ++i) n *= dims[i];
  for (int i = axis + 1; i < rank; ++i) rem *= dims[i];
  if (n * ad * rem == 0) return 0;
  for (int64_t b = 0; b < n; ++b)
    for (int64_t r = 0; r < rem; ++r) {
      const int64_t base = b * ad * rem + r;
      if (dtype == DT_FLOAT32) {
        const float* Y = (const float*)out; const float* G = (const float*)out_grad; float* DX = (float*)x_grad;
        float s = 0.f;
        for (int64_t k = 0; k < ad; ++k) s += G[base + k * rem];
        for (int64_t k = 0; k < ad; ++k) DX[base + k * rem] = G[base + k * rem] - expf(Y[base + k * rem]) * s;
      } else {
        const double* Y = (const double*)out; const double* G = (const double*)out_grad; double* DX = (double*)x_grad;
        double s = 0.0;
        for (int64_t k = 0; k < ad; ++k) s += G[base + k * rem];
        for (int64_t k = 0; k < ad; ++k) DX[base + k * rem] = G[base + k * rem] - exp(Y[base + k * rem]) * s;
      }
    }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * max_grad / min_grad / prod_grad, as the GPU kernels compute them.
 *   max_grad, min_grad: phi::ReduceMaxGradKernel / ReduceMinGradKernel<T,GPUContext> (gpu/reduce_kernel.cu:76-120,159-203):
 *       equal_out = Equal(out, x) with funcs::EqualFunctor<T> (funcs/compare_functors.h:35-48), stored as T;
 *       x_grad = out_grad * equal_out.
 *     (The CPU kernel, funcs::MaxOrMinGradFunctor reduce_functor.h:161-186, uses an exact ==; the two agree unless two
 *     different values lie within 1e-8 of each other.  tol != 0 selects the GPU rule, tol == 0 the CPU rule.)
 *   prod_grad: funcs::ProdGradFunctor (reduce_functor.h:143-159): dx = dy.broadcast * y.broadcast * x.inverse(), every
 *     step in T (16-bit types: float arithmetic rounded back after each step); integer 1 / 0 -> 0 here (undefined there).
 * op: 0 max/min, 1 prod.  red[i] != 0 marks a reduced dimension of x.
 * ---------------------------------------------------------------------------------------- */
static int eq_tol_f(float a, float b, float d, int tol) {  /* d = a - b rounded to T, widened */
  if (!tol) return a == b;
  if (isinf(a) || isinf(b)) return a == b;
  if (isnan(a) || isnan(b)) return 0;
  return fabs((double)d) < 1e-8;
}
int oracle_select_grad(int op, int tol, int dtype, const int64_t* xd, int xr, const int* red, const void* x, const void* out,
                       const void* out_grad, void* x_grad) {
  int64_t gs[MAXR], idx[MAXR] = {0}, numel = 1, s = 1;
  for (int i = xr - 1; i >= 0; --i) {
    gs[i] = red[i] ? 0 : s;
    if (!red[i]) s *= xd[i];
    numel *= xd[i];
  }
  for (int64_t i = 0; i < numel; ++i) {
    int64_t go = 0;
    for (int d = 0; d < xr; ++d) go += idx[d] * gs[d];
    switch (dtype) {
      case DT_FLOAT32: {
        const float xv = ((const float*)x)[i], y = ((const float*)out)[go], dy = ((const float*)out_grad)[go];
        volatile float t;
        if (op == 0) { t = y - xv; t = dy * (eq_tol_f(y, xv, t, tol) ? 1.f : 0.f); }
        else { volatile float inv = 1.f / xv; t = dy * y; t = t * inv; }
        ((float*)x_grad)[i] = t; break; }
      case DT_FLOAT64: {
        const double xv = ((const double*)x)[i], y = ((const double*)out)[go], dy = ((const double*)out_grad)[go];
        volatile double t;
        if (op == 0) {
          int e;
          if (!tol) e = (y == xv);
          else if (isinf((float)y) || isinf((float)xv)) e = (y == xv);
          else if (isnan((float)y) || isnan((float)xv)) e = 0;
          else { t = y - xv; e = fabs(t) < 1e-8; }
          t = dy * (e ? 1.0 : 0.0);
        } else { volatile double inv = 1.0 / xv; t = dy * y; t = t * inv; }
        ((double*)x_grad)[i] = t; break; }
      case DT_INT32: {
        const int32_t xv = ((const int32_t*)x)[i], y = ((const int32_t*)out)[go], dy = ((const int32_t*)out_grad)[go];
        if (op == 0) ((int32_t*)x_grad)[i] = y == xv ? dy : 0;
        else ((int32_t*)x_grad)[i] = (int32_t)((uint32_t)dy * (uint32_t)y * (uint32_t)(xv == 1 ? 1 : xv == -1 ? -1 : 0));
        break; }
      case DT_INT64: {
        const int64_t xv = ((const int64_t*)x)[i], y = ((const int64_t*)out)[go], dy = ((const int64_t*)out_grad)[go];
        if (op == 0) ((int64_t*)x_grad)[i] = y == xv ? dy : 0;
        else ((int64_t*)x_grad)[i] = (int64_t)((uint64_t)dy * (uint64_t)y * (uint64_t)(xv == 1 ? 1 : xv == -1 ? -1 : 0));
        break; }
      case DT_BFLOAT16: case DT_FLOAT16: {
        const int b = dtype == DT_BFLOAT16;
        const uint16_t xb = ((const uint16_t*)x)[i], yb = ((const uint16_t*)out)[go], gb = ((const uint16_t*)out_grad)[go];
        const float xv = b ? bf2f(xb) : h2f(xb), y = b ? bf2f(yb) : h2f(yb), dy = b ? bf2f(gb) : h2f(gb);
#define RT(v) (b ? bf2f(f2bf(v)) : h2f(f2h(v)))
        float t;
        if (op == 0) { const float d = RT(y - xv); t = RT(dy * (eq_tol_f(y, xv, d, tol) ? 1.f : 0.f)); }
        else { const float inv = RT(1.f / xv); t = RT(dy * y); t = RT(t * inv); }
#undef RT
        ((uint16_t*)x_grad)[i] = b ? f2bf(t) : f2h(t); break; }
      default: return 2;
    }
    for (int d = xr - 1; d >= 0; --d) { if (++idx[d] < xd[d]) break; idx[d] = 0; }
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * log_softmax.  phi::LogSoftmaxKernel<T,CPUContext> -> LogSoftmaxFunctor
 * (cpu/log_softmax_kernel.cc:40-127): x viewed as [n][axis_dim][remain];
 *     d = max(x - max_along_axis, -64)   (ValueClip :32-38)
 *     y = d - log(sum_along_axis(exp(d)))
 * 0-D -> 0; 0-size -> no-op.  Sums run along the axis in increasing index (Eigen's reducer order only moves the
 * last bits: tests compare to 1e-6).  T = float or double on the CPU; bf16 / fp16 (GPU registration) follow by
 * computing in float and rounding once.
 * ---------------------------------------------------------------------------------------- */
int oracle_log_softmax(int dtype, const void* x, const int64_t* dims, int rank, int axis, void* out) {
  if (rank == 0) {
    switch (dtype) {
      case DT_FLOAT32: *(float*)out = 0.f; break;
      case DT_FLOAT64: *(double*)out = 0.0; break;
      case DT_BFLOAT16: case DT_FLOAT16: *(uint16_t*)out = 0; break;
      default: return 2;
    }
    return 0;
  }
  if (axis < -rank || axis >= rank) return 1;
  if (axis < 0) axis += rank;
  int64_t n = 1, ad = dims[axis], rem = 1;
  for (int i = 0; i < axis; ++i) n *= dims[i];
  for (int i = axis + 1; i < rank; ++i) rem *= dims[i];
  if (n * ad * rem == 0) return 0;
  for (int64_t b = 0; b < n; ++b)
    for (int64_t r = 0; r < rem; ++r) {
      const int64_t base = b * ad * rem + r;
#define LSM_AT(k) (base + (k) * rem)
      if (dtype == DT_FLOAT64) {
        const double* X = (const double*)x; double* Y = (double*)out;
        double m = X[LSM_AT(0)], s = 0.0;
        for (int64_t k = 1; k < ad; ++k) if (X[LSM_AT(k)] > m) m = X[LSM_AT(k)];
        for (int64_t k = 0; k < ad; ++k) { double d = X[LSM_AT(k)] - m; if (d < -64.0) d = -64.0; Y[LSM_AT(k)] = d; s += exp(d); }
        const double l = log(s);
        for (int64_t k = 0; k < ad; ++k) Y[LSM_AT(k)] -= l;
      } else {
        float m = -INFINITY, s = 0.f;
#define LSM_LOAD(k) (dtype == DT_FLOAT32 ? ((const float*)x)[LSM_AT(k)] : dtype == DT_BFLOAT16 ? bf2f(((const uint16_t*)x)[LSM_AT(k)]) : h2f(((const uint16_t*)x)[LSM_AT(k)]))
        if (dtype != DT_FLOAT32 && dtype != DT_BFLOAT16 && dtype != DT_FLOAT16) return 2;
        for (int64_t k = 0; k < ad; ++k) { const float v = LSM_LOAD(k); if (v > m) m = v; }
        for (int64_t k = 0; k < ad; ++k) { float d = LSM_LOAD(k) - m; if (d < -64.f) d = -64.f; s += expf(d); }
        const float l = logf(s);
        for (int64_t k = 0; k < ad; ++k) {
          float d = LSM_LOAD(k) - m; if (d < -64.f) d = -64.f;
          const float y = d - l;
          if (dtype == DT_FLOAT32) ((float*)out)[LSM_AT(k)] = y;
          else ((uint16_t*)out)[LSM_AT(k)] = dtype == DT_BFLOAT16 ? f2bf(y) : f2h(y);
        }
      }
    }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * cross_entropy_with_softmax.  phi::CrossEntropyWithSoftmaxKernel<T,CPUContext>
 * (cpu/cross_entropy_kernel.cc:73-104): softmax = SoftmaxKernel(logits, axis) (oracle_softmax in phi_cpu_oracle.c), or
 * the input itself when use_softmax is false; then CrossEntropyFunctor (funcs/cross_entropy.cc:29-134) on
 * [n][axis_dim][remain]:
 *   hard: loss[i][j] = label == ignore_index ? 0 : -TolerableValue(log(prob[i][label][j]))   (label: int64 here)
 *   soft: loss[i][j] = -sum_k label[i][k][j] * TolerableValue(log(prob[i][k][j]))   (summed along k in increasing order)
 * TolerableValue: +-inf -> +-1e20 (funcs/cross_entropy.h:28-38).  A label outside [0, axis_dim) raises in the
 * reference: rc 3 here.  T = float or double.
 * ---------------------------------------------------------------------------------------- */
int oracle_softmax(int dtype, const void* x, const int64_t* dims, int rank, int axis, void* out);
static double tol_d(double v) { return v == INFINITY ? 1e20 : (v == -INFINITY ? -1e20 : v); }
static float tol_f(float v) { return v == INFINITY ? 1e20f : (v == -INFINITY ? -1e20f : v); }
int oracle_cross_entropy_with_softmax(int dtype, const void* logits, const int64_t* dims, int rank, const void* label,
                                      int soft_label, int use_softmax, int ignore_index, int axis, void* softmax,
                                      void* loss) {
  if (rank < 1 || axis < -rank || axis >= rank) return 1;
  if (axis < 0) axis += rank;
  int64_t n = 1, ad = dims[axis], rem = 1;
  for (int i = 0; i < axis; ++i) n *= dims[i];
  for (int i = axis + 1; i < rank; ++i) rem *= dims[i];
  if (ad <= 0 || n <= 0) return 1;
  const size_t es = dtype == DT_FLOAT64 ? 8 : 4;
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  if (use_softmax) { int rc = oracle_softmax(dtype, logits, dims, rank, axis, softmax); if (rc) return rc; }
  else if (softmax != logits) memcpy(softmax, logits, (size_t)(n * ad * rem) * es);
  for (int64_t i = 0; i < n; ++i)
    for (int64_t j = 0; j < rem; ++j) {
      const int64_t pos = i * rem + j;
      if (soft_label) {
        if (dtype == DT_FLOAT64) {
          double s = 0.0;
          for (int64_t k = 0; k < ad; ++k) s += ((const double*)label)[(i * ad + k) * rem + j] * tol_d(log(((const double*)softmax)[(i * ad + k) * rem + j]));
          ((double*)loss)[pos] = -s;
        } else {
          float s = 0.f;
          for (int64_t k = 0; k < ad; ++k) s += ((const float*)label)[(i * ad + k) * rem + j] * tol_f(logf(((const float*)softmax)[(i * ad + k) * rem + j]));
          ((float*)loss)[pos] = -s;
        }
      } else {
        const int64_t lbl = ((const int64_t*)label)[pos];
        if (lbl != ignore_index && (lbl < 0 || lbl >= ad)) return 3;
        if (dtype == DT_FLOAT64) ((double*)loss)[pos] = lbl == ignore_index ? 0.0 : -tol_d(log(((const double*)softmax)[(i * ad + lbl) * rem + j]));
        else ((float*)loss)[pos] = lbl == ignore_index ? 0.f : -tol_f(logf(((const float*)softmax)[(i * ad + lbl) * rem + j]));
      }
    }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * bitwise_left_shift / bitwise_right_shift (funcs/bitwise_functors.h:51-200; cpu/bitwise_kernel.cc:40-78 applies the
 * inverse functor to swapped operands when x has the lower rank — the value is x shifted by y either way).
 *   left (arithmetic and logic alike): count outside [0, bits) -> 0, else x << y
 *   right arithmetic: count outside -> the sign fill (-1 for negative x, 0 otherwise; uint8: 0), else x >> y
 *   right logic: count outside -> 0, else the unsigned representation shifted (logic_shift_func for signed types)
 * ---------------------------------------------------------------------------------------- */
#define SHIFT_CASES(T, U, SIGNED)                                                                          \
  {                                                                                                        \
    const T* X = (const T*)x; const T* Y = (const T*)y; T* Q = (T*)out;                                    \
    const int bits = (int)sizeof(T) * 8;                                                                   \
    WALK({                                                                                                 \
      const T a = X[xo]; const T b = Y[yo];                                                                \
      const int bad = (SIGNED && b < 0) || (int64_t)b >= bits;                                             \
      if (!right) Q[i] = bad ? (T)0 : (T)(U)((U)a << b);                                                   \
      else if (is_arithmetic) Q[i] = bad ? (T)((SIGNED && a < 0) ? -1 : 0) : (T)(a >> b);                  \
      else Q[i] = bad ? (T)0 : (T)(U)((U)a >> b);                                                          \
    })                                                                                                     \
  }
int oracle_bitwise_shift(int right, int is_arithmetic, int dtype, const void* x, const int64_t* xd, int xr, const void* y,
                         const int64_t* yd, int yr, void* out) {
  Walk w;
  if (walk_init(&w, xd, xr, yd, yr)) return 1;
  switch (dtype) {
    case DT_UINT8: SHIFT_CASES(uint8_t, uint8_t, 0) break;
    case DT_INT8: SHIFT_CASES(int8_t, uint8_t, 1) break;
    case DT_INT16: SHIFT_CASES(int16_t, uint16_t, 1) break;
    case DT_INT32: SHIFT_CASES(int32_t, uint32_t, 1) break;
    case DT_INT64: SHIFT_CASES(int64_t, uint64_t, 1) break;
    default: return 2;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * rms_norm.  PARITY UNPINNED for this function: the reference has no CPU kernel for rms_norm
 * (GPU only, paddle/phi/kernels/gpu/rms_norm_kernel.cu), so it cannot be compiled into oracle/_ref.  This restates the
 * formula the reference's own test checks its kernel against (test/legacy_test/test_rms_norm_op.py:37-47,74-79):
 *     v = x + residual + bias;  variance = mean(v^2, last dims);  out = rsqrt(variance + eps) * v * gamma + beta
 * in double precision; callers round to the tensor dtype.  residual / bias / beta may be NULL.
 * ---------------------------------------------------------------------------------------- */
int oracle_rms_norm(const double* x, int64_t rows, int64_t D, const double* bias, const double* residual, const double* gamma,
                    const double* beta, double eps, double* out, double* residual_out, double* inv_var) {
  for (int64_t r = 0; r < rows; ++r) {
    double ss = 0.0;
    for (int64_t i = 0; i < D; ++i) {
      double v = x[r * D + i];
      if (residual) { v += residual[r * D + i]; if (bias) v += bias[i]; }
      if (residual_out) residual_out[r * D + i] = v;
      ss += v * v;
    }
    const double inv = 1.0 / sqrt(ss / (double)D + eps);
    if (inv_var) inv_var[r] = inv;
    for (int64_t i = 0; i < D; ++i) {
      double v = x[r * D + i];
      if (residual) { v += residual[r * D + i]; if (bias) v += bias[i]; }
      out[r * D + i] = inv * v * gamma[i] + (beta ? beta[i] : 0.0);
    }
  }
  return 0;
}

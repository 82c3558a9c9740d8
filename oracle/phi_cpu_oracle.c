/*
 * phi_cpu_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the algorithms of PaddlePaddle's CPU PHI kernels for the hot
 * path (elementwise add/multiply with broadcasting, sum/mean/max/min/prod, argmax, softmax,
 * layer_norm).  It exists to CHECK the CUDA kernels; it is never the thing measured or
 * shipped.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  Nothing under paddlepaddle_b200/ imports, links or
 * executes it.
 *
 * Parity status: PINNED.  tests/test_oracle_pin.py checks every function below against
 *   (a) oracle/_ref/libphi_cpu_ref.so = the reference's own CPU kernels compiled from
 *       /root/reference (oracle/build_ref.sh), when that library is present, and
 *   (b) tests/golden/*.npz = outputs of (a) recorded by tests/golden/make_golden.py,
 *   (c) the docstring known answers of python/paddle/tensor/math.py:1530-1585.
 * The reference keeps no golden-vector files for these ops.
 *
 * Third-party arithmetic restated here: Eigen (vendored at third_party/eigen3, commit
 * f612df27 per .SUBMODULES.json) — the reduction order of TensorReduction.h.
 *
 * Every function cites the reference lines it follows.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* dtype codes shared with include/pdb200.h */
enum { DT_BOOL = 0, DT_UINT8, DT_INT8, DT_INT16, DT_INT32, DT_INT64, DT_FLOAT16, DT_BFLOAT16, DT_FLOAT32, DT_FLOAT64 };
enum { OP_SUM = 0, OP_MEAN, OP_MAX, OP_MIN, OP_PROD };
#define MAXR 9

static size_t dt_size(int dt) {
  switch (dt) {
    case DT_BOOL: case DT_UINT8: case DT_INT8: return 1;
    case DT_INT16: case DT_FLOAT16: case DT_BFLOAT16: return 2;
    case DT_INT32: case DT_FLOAT32: return 4;
    default: return 8;
  }
}

/* ------------------------------------------------------------------------------------------
 * bf16 / fp16 scalars
 * float -> bf16: paddle/phi/common/bfloat16.h:61-88 (cpu_float_to_bfloat16): zero keeps sign,
 * inf passes, NaN -> 0x7FFF, else add 0x7FFF + lsb and truncate (round-to-nearest-even).
 * bf16 -> float: :233-249 (bits << 16).
 * fp16: paddle/phi/common/float16.h converts with F16C/IEEE RNE; gcc's _Float16 is the same rule.
 * ---------------------------------------------------------------------------------------- */
static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) == 0) return (uint16_t)((u >> 16) & 0x8000u);
  if ((u & 0x7f800000u) == 0x7f800000u) {
    if (u & 0x007fffffu) return 0x7FFF;
    return (uint16_t)(u >> 16);
  }
  u += 0x00007FFFu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static float bf2f(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static uint16_t f2h(float f) {
  _Float16 h = (_Float16)f;
  uint16_t b;
  memcpy(&b, &h, 2);
  return b;
}
static float h2f(uint16_t b) {
  _Float16 h;
  memcpy(&h, &b, 2);
  return (float)h;
}

void oracle_float_to_bf16(const float* in, uint16_t* out, int64_t n) { for (int64_t i = 0; i < n; ++i) out[i] = f2bf(in[i]); }
void oracle_bf16_to_float(const uint16_t* in, float* out, int64_t n) { for (int64_t i = 0; i < n; ++i) out[i] = bf2f(in[i]); }
void oracle_float_to_fp16(const float* in, uint16_t* out, int64_t n) { for (int64_t i = 0; i < n; ++i) out[i] = f2h(in[i]); }
void oracle_fp16_to_float(const uint16_t* in, float* out, int64_t n) { for (int64_t i = 0; i < n; ++i) out[i] = h2f(in[i]); }

/* ------------------------------------------------------------------------------------------
 * Elementwise add / multiply with broadcasting.
 * Shape rule: funcs::GetBroadcastDimsArrays, paddle/phi/kernels/funcs/common_shape.h:35-93
 *   (right-aligned, axis = |rank_x - rank_y|, dims equal or 1).
 * Walk: the general odometer of CommonForwardBroadcastCPU,
 *   paddle/phi/kernels/funcs/elementwise_base.h:274-307 (index array incremented from the last
 *   dim, per-operand offset = sum(idx[d] * stride[d]) with stride 0 on broadcast dims).  The
 *   RowWise / MidWise fast paths (:54-151, 200-271) compute the same values.
 * Scalar arithmetic: funcs/elementwise_functor.h:38-94 — a + b, a * b in T; bool multiply is AND;
 *   bf16/fp16 round once from the float result (bfloat16.h:264-276).  Integer overflow wraps.
 * ---------------------------------------------------------------------------------------- */
int oracle_broadcast_shape(const int64_t* xd, int xr, const int64_t* yd, int yr, int64_t* od, int* orank) {
  int R = xr > yr ? xr : yr;
  for (int i = 0; i < R; ++i) {
    int xi = i - (R - xr), yi = i - (R - yr);
    int64_t a = xi >= 0 ? xd[xi] : 1, b = yi >= 0 ? yd[yi] : 1;
    if (a != b && a != 1 && b != 1) return 1;
    od[i] = a == 1 ? b : a;
  }
  *orank = R;
  return 0;
}

#define EW_LOOP(T, EXPR)                                                     \
  {                                                                          \
    const T* X = (const T*)x; const T* Y = (const T*)y; T* O = (T*)out;      \
    for (int64_t i = 0; i < numel; ++i) {                                    \
      int64_t xo = 0, yo = 0;                                                \
      for (int d = 0; d < R; ++d) { xo += idx[d] * xs[d]; yo += idx[d] * ys[d]; } \
      T a = X[xo], b = Y[yo];                                                \
      O[i] = (EXPR);                                                         \
      for (int d = R - 1; d >= 0; --d) { if (++idx[d] < od[d]) break; idx[d] = 0; } \
    }                                                                        \
  }

int oracle_elementwise(int op, int dtype, const void* x, const int64_t* xd, int xr, const void* y,
                       const int64_t* yd, int yr, void* out) {
  int64_t od[MAXR], xs[MAXR], ys[MAXR], idx[MAXR] = {0};
  int R;
  if (oracle_broadcast_shape(xd, xr, yd, yr, od, &R)) return 1;
  int64_t numel = 1, sx = 1, sy = 1;
  for (int i = R - 1; i >= 0; --i) {
    int xi = i - (R - xr), yi = i - (R - yr);
    int64_t a = xi >= 0 ? xd[xi] : 1, b = yi >= 0 ? yd[yi] : 1;
    xs[i] = (a == 1 && od[i] != 1) ? 0 : sx;
    ys[i] = (b == 1 && od[i] != 1) ? 0 : sy;
    if (a == 1) xs[i] = 0;
    if (b == 1) ys[i] = 0;
    sx *= a; sy *= b;
    numel *= od[i];
  }
  if (op >= 2) {
    /* subtract / divide / maximum / minimum: funcs/elementwise_functor.h:70-76 (a - b), :102-121 (a / b;
     * integer divide raises on b == 0 — here rc 3), :471-475 (a > b ? a : b), :507-511 (a < b ? a : b).
     * When x has the lower rank the CPU kernels swap the operands: subtract/divide through the Inverse
     * functors (same value), maximum/minimum by evaluating f(y, x) (same value unless NaN / signed zero). */
#define EW_OP2(T, SUB, DIV)                                                        \
  switch (op) {                                                                    \
    case 2: EW_LOOP(T, SUB) break;                                                 \
    case 3: EW_LOOP(T, DIV) break;                                                 \
    case 4: EW_LOOP(T, a > b ? a : b) break;                                       \
    case 5: EW_LOOP(T, a < b ? a : b) break;                                       \
    default: return 2;                                                             \
  }
#define EW_INT_DIVZERO(T)                                                          \
  if (op == 3) {                                                                   \
    int64_t ny = 1; for (int i = 0; i < yr; ++i) ny *= yd[i];                      \
    for (int64_t i = 0; i < ny; ++i) if (((const T*)y)[i] == 0) return 3;          \
  }
    switch (dtype) {
      case DT_FLOAT32: EW_OP2(float, a - b, a / b) break;
      case DT_FLOAT64: EW_OP2(double, a - b, a / b) break;
      case DT_INT32: EW_INT_DIVZERO(int32_t) EW_OP2(int32_t, (int32_t)((uint32_t)a - (uint32_t)b), (b == -1 ? (int32_t)(0u - (uint32_t)a) : a / b)) break;
      case DT_INT64: EW_INT_DIVZERO(int64_t) EW_OP2(int64_t, (int64_t)((uint64_t)a - (uint64_t)b), (b == -1 ? (int64_t)(0ull - (uint64_t)a) : a / b)) break;
      case DT_INT16: EW_INT_DIVZERO(int16_t) EW_OP2(int16_t, (int16_t)(a - b), (int16_t)(a / b)) break;
      case DT_INT8: EW_INT_DIVZERO(int8_t) EW_OP2(int8_t, (int8_t)(a - b), (int8_t)(a / b)) break;
      case DT_UINT8: EW_INT_DIVZERO(uint8_t) EW_OP2(uint8_t, (uint8_t)(a - b), (uint8_t)(a / b)) break;
      case DT_BFLOAT16:
        switch (op) {
          case 2: EW_LOOP(uint16_t, f2bf(bf2f(a) - bf2f(b))) break;
          case 3: EW_LOOP(uint16_t, f2bf(bf2f(a) / bf2f(b))) break;
          case 4: EW_LOOP(uint16_t, bf2f(a) > bf2f(b) ? a : b) break;
          case 5: EW_LOOP(uint16_t, bf2f(a) < bf2f(b) ? a : b) break;
          default: return 2;
        }
        break;
      case DT_FLOAT16:
        switch (op) {
          case 2: EW_LOOP(uint16_t, f2h(h2f(a) - h2f(b))) break;
          case 3: EW_LOOP(uint16_t, f2h(h2f(a) / h2f(b))) break;
          case 4: EW_LOOP(uint16_t, h2f(a) > h2f(b) ? a : b) break;
          case 5: EW_LOOP(uint16_t, h2f(a) < h2f(b) ? a : b) break;
          default: return 2;
        }
        break;
      default: return 2;
    }
    return 0;
  }
  const int add = (op == 0);
  switch (dtype) {
    case DT_FLOAT32: EW_LOOP(float, add ? a + b : a * b) break;
    case DT_FLOAT64: EW_LOOP(double, add ? a + b : a * b) break;
    case DT_INT32: EW_LOOP(int32_t, add ? (int32_t)((uint32_t)a + (uint32_t)b) : (int32_t)((uint32_t)a * (uint32_t)b)) break;
    case DT_INT64: EW_LOOP(int64_t, add ? (int64_t)((uint64_t)a + (uint64_t)b) : (int64_t)((uint64_t)a * (uint64_t)b)) break;
    case DT_INT16: EW_LOOP(int16_t, add ? (int16_t)(a + b) : (int16_t)(a * b)) break;
    case DT_INT8: EW_LOOP(int8_t, add ? (int8_t)(a + b) : (int8_t)(a * b)) break;
    case DT_UINT8: EW_LOOP(uint8_t, add ? (uint8_t)(a + b) : (uint8_t)(a * b)) break;
    case DT_BOOL: EW_LOOP(uint8_t, add ? (uint8_t)((a + b) != 0) : (uint8_t)(a && b)) break;
    case DT_BFLOAT16: EW_LOOP(uint16_t, add ? f2bf(bf2f(a) + bf2f(b)) : f2bf(bf2f(a) * bf2f(b))) break;
    case DT_FLOAT16: EW_LOOP(uint16_t, add ? f2h(h2f(a) + h2f(b)) : f2h(h2f(a) * h2f(b))) break;
    default: return 2;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * Reductions.
 * Dispatch: phi::Reduce<CPUContext> (paddle/phi/kernels/cpu/reduce.h:26-67) ->
 *   funcs::ReduceKernelImpl (paddle/phi/kernels/funcs/reduce_function.h:1354-1402) ->
 *   Eigen functors y = x.sum(dim) / mean / maximum / minimum / prod
 *   (paddle/phi/kernels/funcs/reduce_functor.h:52-89), single thread.
 * Axis rules: negative axes += rank; empty axes / full set = reduce all
 *   (core/kernel_utils.h:407-417, cpu/reduce.h:35-45).  keep_dim is shape-only.
 * Reduction ORDER restated from Eigen (third_party/eigen3/unsupported/Eigen/CXX11/src/Tensor/
 *   TensorReduction.h).  Paddle passes the reduced dims as a run-time Eigen::array, so Eigen's
 *   static ReducingInnerMostDims / PreservingInnerMostDims fast paths (:100-139) are OFF:
 *   - reduce_all (Paddle flattens to 1-D, reduce_function.h:1368-1377) runs FullReducer ->
 *     InnerMostDimReducer::reduce over the whole buffer (:169-260): for float/double sum and
 *     prod a binary tree with 1024-packet leaves, two packet accumulators inside a leaf,
 *     horizontal add of the packet at the end; for mean one packet accumulator, no tree;
 *     integers and max/min are order-free;
 *   - every partial reduction evaluates each output with coeff() (:783-800) ->
 *     GenericDimReducer (:142-160): a plain sequential accumulation in T whose OUTERMOST loop
 *     is the last reduced dim and whose innermost loop is the first reduced dim.
 *   Packet = 8 floats / 4 doubles (AVX, Paddle's x86 release flags).
 * mean = sum / count (MeanReducer::finalize, TensorFunctors.h); integer mean truncates.
 * max/min identities: lowest()/max() — irrelevant to the result for non-empty input.
 * ---------------------------------------------------------------------------------------- */
#define DEF_TREE_SUM(NAME, T, PK)                                                         \
  static T NAME##_hadd(const T* p) {                                                      \
    /* predux: fold the upper half onto the lower half until one lane is left */          \
    T t[PK];                                                                              \
    for (int i = 0; i < PK; ++i) t[i] = p[i];                                             \
    if (PK == 8) { for (int i = 0; i < 4; ++i) t[i] = t[i] + t[i + 4]; }                  \
    /* Packet4: (t0+t2) + (t1+t3) */                                                      \
    if (PK >= 4) return (t[0] + t[2]) + (t[1] + t[3]);                                    \
    return t[0] + t[1];                                                                   \
  }                                                                                       \
  static T NAME##_hmul(const T* p) {                                                      \
    T t[PK];                                                                              \
    for (int i = 0; i < PK; ++i) t[i] = p[i];                                             \
    if (PK == 8) { for (int i = 0; i < 4; ++i) t[i] = t[i] * t[i + 4]; }                  \
    if (PK >= 4) return (t[0] * t[2]) * (t[1] * t[3]);                                    \
    return t[0] * t[1];                                                                   \
  }                                                                                       \
  static T NAME(const T* x, int64_t first, int64_t n, int is_prod) {                      \
    const T ident = is_prod ? (T)1 : (T)0;                                                \
    T accum = ident;                                                                      \
    if (n > (int64_t)PK * 1024) {                                                         \
      int64_t half_up = (n + 1) / 2;                                                      \
      int64_t split = PK * ((first + half_up + PK - 1) / PK);                             \
      int64_t num_left = split - first < n ? split - first : n;                           \
      T l = NAME(x, first, num_left, is_prod);                                            \
      accum = is_prod ? accum * l : accum + l;                                            \
      if (num_left < n) {                                                                 \
        T r = NAME(x, split, n - num_left, is_prod);                                      \
        accum = is_prod ? accum * r : accum + r;                                          \
      }                                                                                   \
      return accum;                                                                       \
    }                                                                                     \
    int64_t unroll = (n / (2 * PK)) * 2 * PK, vec = (n / PK) * PK;                        \
    T p1[PK], p2[PK];                                                                     \
    for (int i = 0; i < PK; ++i) { p1[i] = ident; p2[i] = ident; }                        \
    int64_t j = 0;                                                                        \
    for (; j < unroll; j += 2 * PK)                                                       \
      for (int i = 0; i < PK; ++i) {                                                      \
        if (is_prod) { p1[i] *= x[first + j + i]; p2[i] *= x[first + j + PK + i]; }       \
        else { p1[i] += x[first + j + i]; p2[i] += x[first + j + PK + i]; }               \
      }                                                                                   \
    for (; j < vec; j += PK)                                                              \
      for (int i = 0; i < PK; ++i) { if (is_prod) p1[i] *= x[first + j + i]; else p1[i] += x[first + j + i]; } \
    for (int i = 0; i < PK; ++i) { if (is_prod) p1[i] *= p2[i]; else p1[i] += p2[i]; }    \
    for (j = vec; j < n; ++j) { if (is_prod) accum *= x[first + j]; else accum += x[first + j]; } \
    return is_prod ? accum * NAME##_hmul(p1) : accum + NAME##_hadd(p1);                   \
  }                                                                                       \
  /* MeanReducer: stateful => no tree, ONE packet accumulator (TensorReduction.h:183-198) */ \
  static T NAME##_mean(const T* x, int64_t first, int64_t n) {                            \
    int64_t vec = (n / PK) * PK;                                                          \
    T p1[PK];                                                                             \
    for (int i = 0; i < PK; ++i) p1[i] = (T)0;                                            \
    for (int64_t j = 0; j < vec; j += PK)                                                 \
      for (int i = 0; i < PK; ++i) p1[i] += x[first + j + i];                             \
    T accum = (T)0;                                                                       \
    for (int64_t j = vec; j < n; ++j) accum += x[first + j];                              \
    return (accum + NAME##_hadd(p1)) / (T)n;                                              \
  }
DEF_TREE_SUM(tree_f32, float, 8)
DEF_TREE_SUM(tree_f64, double, 4)

/* normalise axes -> red[] flags; returns 0 ok */
static int norm_axes(int rank, const int64_t* axes, int n_axes, int reduce_all, int* red) {
  for (int i = 0; i < rank; ++i) red[i] = 0;
  if (rank == 0 || n_axes == 0) reduce_all = 1;
  for (int i = 0; i < n_axes; ++i) {
    int64_t a = axes[i];
    if (rank > 0) {
      if (a < -rank || a >= rank) return 1;
      red[a < 0 ? a + rank : a] = 1;
    }
  }
  if (reduce_all) for (int i = 0; i < rank; ++i) red[i] = 1;
  return 0;
}

#define RED_GENERIC(T, INIT, STEP, FINAL)                                                       \
  {                                                                                             \
    const T* X = (const T*)x; T* O = (T*)out;                                                   \
    for (int64_t o = 0; o < out_numel; ++o) {                                                   \
      /* decode the kept coordinates of output o */                                             \
      int64_t base = 0, rem = o;                                                                \
      for (int d = nk - 1; d >= 0; --d) { base += (rem % kdim[d]) * kstr[d]; rem /= kdim[d]; }  \
      T acc = (INIT);                                                                           \
      int64_t ridx[MAXR] = {0};                                                                 \
      for (int64_t r = 0; r < red_numel; ++r) {                                                 \
        int64_t off = base;                                                                     \
        for (int d = 0; d < nr; ++d) off += ridx[d] * rstr[d];                                  \
        T v = X[off];                                                                           \
        STEP;                                                                                   \
        /* GenericDimReducer: the FIRST reduced dim is the innermost loop */                    \
        for (int d = 0; d < nr; ++d) { if (++ridx[d] < rdim[d]) break; ridx[d] = 0; }           \
      }                                                                                         \
      O[o] = (FINAL);                                                                           \
    }                                                                                           \
  }

int oracle_reduce(int op, int dtype, const void* x, const int64_t* dims, int rank, const int64_t* axes,
                  int n_axes, int reduce_all, void* out) {
  int red[MAXR];
  if (rank > MAXR || norm_axes(rank, axes, n_axes, reduce_all, red)) return 1;
  int64_t stride[MAXR], kdim[MAXR], kstr[MAXR], rdim[MAXR], rstr[MAXR];
  int nk = 0, nr = 0;
  int64_t s = 1, numel = 1;
  for (int i = rank - 1; i >= 0; --i) { stride[i] = s; s *= dims[i]; numel *= dims[i]; }
  if (numel <= 0) return 1; /* "Tensor need be reduced must not empty." reduce_function.h:1360-1363 */
  int64_t out_numel = 1, red_numel = 1;
  for (int i = 0; i < rank; ++i) {
    if (red[i]) { rdim[nr] = dims[i]; rstr[nr] = stride[i]; ++nr; red_numel *= dims[i]; }
    else { kdim[nk] = dims[i]; kstr[nk] = stride[i]; ++nk; out_numel *= dims[i]; }
  }
  /* full reduction -> FullReducer -> InnerMostDimReducer order for float sum/prod/mean */
  const int full = (nk == 0);
  if (full && (dtype == DT_FLOAT32 || dtype == DT_FLOAT64) && (op == OP_SUM || op == OP_PROD || op == OP_MEAN)) {
    for (int64_t o = 0; o < out_numel; ++o) {
      if (dtype == DT_FLOAT32) {
        const float* X = (const float*)x;
        ((float*)out)[o] = op == OP_MEAN ? tree_f32_mean(X, o * red_numel, red_numel)
                                         : tree_f32(X, o * red_numel, red_numel, op == OP_PROD);
      } else {
        const double* X = (const double*)x;
        ((double*)out)[o] = op == OP_MEAN ? tree_f64_mean(X, o * red_numel, red_numel)
                                          : tree_f64(X, o * red_numel, red_numel, op == OP_PROD);
      }
    }
    return 0;
  }
#define RED_BY_OP(T, UT, IS_INT)                                                                 \
  switch (op) {                                                                                  \
    case OP_SUM: RED_GENERIC(T, (T)0, acc = IS_INT ? (T)((UT)acc + (UT)v) : (T)(acc + v), acc) break; \
    case OP_MEAN: RED_GENERIC(T, (T)0, acc = IS_INT ? (T)((UT)acc + (UT)v) : (T)(acc + v), (T)(acc / (T)red_numel)) break; \
    case OP_MAX: RED_GENERIC(T, X[base], acc = (acc < v) ? v : acc, acc) break;                  \
    case OP_MIN: RED_GENERIC(T, X[base], acc = (v < acc) ? v : acc, acc) break;                  \
    case OP_PROD: RED_GENERIC(T, (T)1, acc = IS_INT ? (T)((UT)acc * (UT)v) : (T)(acc * v), acc) break; \
    default: return 2;                                                                           \
  }
  switch (dtype) {
    case DT_FLOAT32: RED_BY_OP(float, float, 0) break;
    case DT_FLOAT64: RED_BY_OP(double, double, 0) break;
    case DT_INT32: RED_BY_OP(int32_t, uint32_t, 1) break;
    case DT_INT64: RED_BY_OP(int64_t, uint64_t, 1) break;
    default: return 2; /* CPU max/min/prod: f32,f64,i32,i64 only (cpu/reduce_max_kernel.cc etc.) */
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * argmax.  phi::ArgMaxKernel<T,CPUContext> -> ArgMinMaxFunctor -> Eigen argmax(axis)
 *   paddle/phi/kernels/cpu/arg_min_max_kernel.cc:35-66,150-196.
 * Eigen's ArgMaxTupleReducer keeps the accumulator unless t.second > accum.second (strict), so the
 *   FIRST maximal index wins (TensorFunctors.h:370-381).  flatten -> index into the flat tensor.
 * Output int64 (default) or int32.  Values are compared in T.  The CPU kernel is not registered for
 *   bf16/fp16 (arg_min_max_kernel.cc:212-224); for those the same rule is applied to their float values
 *   with float's lowest() as the start — an extension by analogy, not pinned by the reference.
 * ---------------------------------------------------------------------------------------- */
static double load_as_double(const void* x, int dtype, int64_t i) {
  switch (dtype) {
    case DT_FLOAT32: return ((const float*)x)[i];
    case DT_FLOAT64: return ((const double*)x)[i];
    case DT_INT32: return ((const int32_t*)x)[i];
    case DT_INT16: return ((const int16_t*)x)[i];
    case DT_UINT8: return ((const uint8_t*)x)[i];
    case DT_BFLOAT16: return bf2f(((const uint16_t*)x)[i]);
    case DT_FLOAT16: return h2f(((const uint16_t*)x)[i]);
    default: return 0;
  }
}

int oracle_argmax(int dtype, const void* x, const int64_t* dims, int rank, int64_t axis, int flatten,
                  int out_dtype, void* out) {
  int64_t pre = 1, n = 1, post = 1, numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (numel <= 0) return 1;
  if (flatten || rank == 0) { n = numel; }
  else {
    if (axis < -rank || axis >= rank) return 1;
    if (axis < 0) axis += rank;
    for (int i = 0; i < rank; ++i) { if (i < axis) pre *= dims[i]; else if (i == axis) n = dims[i]; else post *= dims[i]; }
  }
  for (int64_t p = 0; p < pre; ++p)
    for (int64_t q = 0; q < post; ++q) {
      int64_t best = 0;
      if (dtype == DT_INT64) {  /* int64 does not fit a double: compare natively */
        const int64_t* X = (const int64_t*)x;
        int64_t bv = X[p * n * post + q];
        for (int64_t k = 1; k < n; ++k) { int64_t v = X[(p * n + k) * post + q]; if (v > bv) { bv = v; best = k; } }
      } else {
        /* the accumulator starts as (index 0, lowest()) — NOT as element 0 — so a NaN in front is skipped
         * like any other NaN, and a row that never beats lowest() (all -inf / NaN) answers 0 */
        double bv = dtype == DT_FLOAT64 ? -1.7976931348623157e308 : (dtype == DT_INT32 ? -2147483648.0 : (dtype == DT_INT16 ? -32768.0 : (dtype == DT_UINT8 ? 0.0 : -(double)3.40282346638528859812e38f)));
        for (int64_t k = 0; k < n; ++k) { double v = load_as_double(x, dtype, (p * n + k) * post + q); if (v > bv) { bv = v; best = k; } }
      }
      if (out_dtype == DT_INT32) ((int32_t*)out)[p * post + q] = (int32_t)best;
      else ((int64_t*)out)[p * post + q] = best;
    }
  return 0;
}

/* argmin.  phi::ArgMinKernel<T,CPUContext> -> ArgMinMaxFunctor<..., kArgMin> -> Eigen argmin(axis)
 *   paddle/phi/kernels/cpu/arg_min_max_kernel.cc:68-98,175-196.  Eigen's ArgMinTupleReducer keeps the accumulator
 *   unless t.second < accum.second (strict) and starts at (index 0, highest()): the FIRST minimal index wins, NaN
 *   never wins, a row that never goes below highest() answers 0 (TensorFunctors.h:383-394). */
int oracle_argmin(int dtype, const void* x, const int64_t* dims, int rank, int64_t axis, int flatten,
                  int out_dtype, void* out) {
  int64_t pre = 1, n = 1, post = 1, numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (numel <= 0) return 1;
  if (flatten || rank == 0) { n = numel; }
  else {
    if (axis < -rank || axis >= rank) return 1;
    if (axis < 0) axis += rank;
    for (int i = 0; i < rank; ++i) { if (i < axis) pre *= dims[i]; else if (i == axis) n = dims[i]; else post *= dims[i]; }
  }
  for (int64_t p = 0; p < pre; ++p)
    for (int64_t q = 0; q < post; ++q) {
      int64_t best = 0;
      if (dtype == DT_INT64) {
        const int64_t* X = (const int64_t*)x;
        int64_t bv = INT64_MAX;
        for (int64_t k = 0; k < n; ++k) { int64_t v = X[(p * n + k) * post + q]; if (v < bv) { bv = v; best = k; } }
      } else {
        double bv = dtype == DT_FLOAT64 ? 1.7976931348623157e308 : (dtype == DT_INT32 ? 2147483647.0 : (dtype == DT_INT16 ? 32767.0 : (dtype == DT_UINT8 ? 255.0 : (double)3.40282346638528859812e38f)));
        for (int64_t k = 0; k < n; ++k) { double v = load_as_double(x, dtype, (p * n + k) * post + q); if (v < bv) { bv = v; best = k; } }
      }
      if (out_dtype == DT_INT32) ((int32_t*)out)[p * post + q] = (int32_t)best;
      else ((int64_t*)out)[p * post + q] = best;
    }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * softmax.  phi::SoftmaxKernel<T,CPUContext> (paddle/phi/kernels/impl/softmax_kernel_impl.h:25-53):
 *   0-size -> no-op; 0-D -> 1; else x viewed as [n][d] with d = prod(dims[axis:]), axis_dim = dims[axis],
 *   -> SoftmaxFunctor<CPUContext> (paddle/phi/kernels/funcs/softmax_impl.h:234-273):
 *   last axis + AVX: per row max_element; y = x - max; clip at -64 (vec_clip / ValueClip :36-42);
 *   exp (std::exp when MKLML is absent, funcs/cpu_vec.h:40-67); sum = vec_sum (AVX, 8 lanes then
 *   horizontal add, cpu_vec.h:152-186); y *= 1/sum.
 *   other axes: SoftmaxEigen (:44-102) — same formula along the axis; sums are sequential over the axis.
 * T = float or double only (cpu/softmax_kernel.cc:21-22).
 * ---------------------------------------------------------------------------------------- */
static float vec_sum_avx_f32(const float* x, int64_t n) {
  /* cpu_vec.h:152-186: 8-lane accumulate, hadd twice, add the two 128-bit halves, then the tail */
  const int64_t rest = n % 8, end = n - rest;
  if (end == 0) { float s = x[0]; for (int64_t i = 1; i < n; ++i) s += x[i]; return s; }
  float l[8];
  for (int i = 0; i < 8; ++i) l[i] = x[i];
  for (int64_t i = 8; i < end; i += 8) for (int k = 0; k < 8; ++k) l[k] += x[i + k];
  /* _mm256_hadd_ps(h,h) twice: lanes 0 and 4 hold (l0+l1)+(l2+l3) and (l4+l5)+(l6+l7) */
  float lo = (l[0] + l[1]) + (l[2] + l[3]);
  float hi = (l[4] + l[5]) + (l[6] + l[7]);
  float s = lo + hi;
  for (int64_t i = end; i < n; ++i) s += x[i];
  return s;
}

int oracle_softmax(int dtype, const void* x, const int64_t* dims, int rank, int axis, void* out) {
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (rank > 0 && numel == 0) return 0;
  if (rank == 0) { if (dtype == DT_FLOAT32) *(float*)out = 1.0f; else *(double*)out = 1.0; return 0; }
  if (axis < -rank || axis >= rank) return 1;
  if (axis < 0) axis += rank;
  int64_t outer = 1, dim = dims[axis], inner = 1;
  for (int i = 0; i < axis; ++i) outer *= dims[i];
  for (int i = axis + 1; i < rank; ++i) inner *= dims[i];
#define SOFTMAX_BODY(T, EXP, LAST_SUM)                                                  \
  {                                                                                     \
    const T* X = (const T*)x; T* Y = (T*)out;                                           \
    for (int64_t o = 0; o < outer; ++o)                                                 \
      for (int64_t q = 0; q < inner; ++q) {                                             \
        const T* src = X + o * dim * inner + q; T* dst = Y + o * dim * inner + q;       \
        T m = src[0];                                                                   \
        for (int64_t k = 1; k < dim; ++k) if (m < src[k * inner]) m = src[k * inner];   \
        for (int64_t k = 0; k < dim; ++k) {                                             \
          T d = src[k * inner] - m;                                                     \
          if (d < (T)-64) d = (T)-64;                                                   \
          dst[k * inner] = EXP(d);                                                      \
        }                                                                               \
        T sum;                                                                          \
        if (inner == 1) { sum = LAST_SUM; }                                             \
        else { sum = 0; for (int64_t k = 0; k < dim; ++k) sum += dst[k * inner]; }      \
        T inv = (T)1 / sum;                                                             \
        for (int64_t k = 0; k < dim; ++k) dst[k * inner] = dst[k * inner] * inv;        \
      }                                                                                 \
  }
  if (dtype == DT_FLOAT32) SOFTMAX_BODY(float, expf, vec_sum_avx_f32(dst, dim))
  else SOFTMAX_BODY(double, exp, tree_f64(dst, 0, dim, 0))
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * layer_norm.  phi::LayerNormKernel<T,CPUContext>, CUDA-build branch
 *   paddle/phi/kernels/cpu/layer_norm_kernel.cc:32-102:
 *   x -> [left][right]; mean = RowwiseMean (sequential T sum, times double 1/right, stored as T;
 *   funcs/math_function_impl.h:158-191); out = (x-mean)^2 (SubAndSquareFunctor,
 *   funcs/layer_norm_util.h:145-148); var = RowwiseMean(out); out = x - mean;
 *   out = out / sqrt(var + eps) (DivAndSqrtFunctor :150-161, eps cast to T);
 *   out *= scale; out += bias (each a separate pass, so separately rounded).
 * Outputs y, mean, variance (biased; NOT rstd).  T = float or double (:151-155).
 * ---------------------------------------------------------------------------------------- */
int oracle_layer_norm(int dtype, const void* x, const int64_t* dims, int rank, const void* scale,
                      const void* bias, float epsilon, int begin_norm_axis, void* y, void* mean, void* var) {
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  if (begin_norm_axis < 0 || begin_norm_axis > rank) return 1;
  int64_t left = 1, right = 1;
  for (int i = 0; i < rank; ++i) { if (i < begin_norm_axis) left *= dims[i]; else right *= dims[i]; }
#define LN_BODY(T, SQRT)                                                                 \
  {                                                                                      \
    const T* X = (const T*)x; T* Y = (T*)y; T* M = (T*)mean; T* V = (T*)var;             \
    const T* S = (const T*)scale; const T* B = (const T*)bias;                           \
    const double inv = 1.0 / (double)right;                                              \
    const T eps = (T)epsilon;                                                            \
    for (int64_t r = 0; r < left; ++r) {                                                 \
      const T* row = X + r * right; T* o = Y + r * right;                                \
      T sum = 0;                                                                         \
      for (int64_t j = 0; j < right; ++j) sum += row[j];                                 \
      const T mu = (T)(sum * inv);                                                       \
      for (int64_t j = 0; j < right; ++j) { T d = row[j] - mu; o[j] = d * d; }           \
      sum = 0;                                                                           \
      for (int64_t j = 0; j < right; ++j) sum += o[j];                                   \
      const T va = (T)(sum * inv);                                                       \
      const T den = SQRT(va + eps);                                                      \
      for (int64_t j = 0; j < right; ++j) {                                              \
        T t = row[j] - mu;                                                               \
        t = t / den;                                                                     \
        if (S) t = t * S[j];                                                             \
        if (B) t = t + B[j];                                                             \
        o[j] = t;                                                                        \
      }                                                                                  \
      if (M) M[r] = mu;                                                                  \
      if (V) V[r] = va;                                                                  \
    }                                                                                    \
  }
  if (dtype == DT_FLOAT32) LN_BODY(float, sqrtf)
  else LN_BODY(double, sqrt)
  (void)dt_size;
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * softmax_grad.  phi::SoftmaxGradKernel<T,CPUContext>
 *   (paddle/phi/kernels/impl/softmax_grad_kernel_impl.h:25-59): 0-D -> 0; 0-size -> no-op; else
 *   SoftmaxGradFunctor<CPUContext> (paddle/phi/kernels/funcs/softmax_impl.h:406-445):
 *   last axis + AVX: per row dot = vec_mul_reduce(dy, y) (float: 8 AVX lanes, then
 *   ((l0+l1)+(l4+l5)) + ((l2+l3)+(l6+l7)), then the tail, funcs/cpu_vec.h:229-263; double: sequential);
 *   dx = y * (dy + (-dot)).  Other axes: SoftmaxGradEigen (:275-308): dot = sum over the axis
 *   (sequential), dx = (dy - dot) * y.
 * ---------------------------------------------------------------------------------------- */
static float vec_mul_reduce_avx_f32(const float* x, const float* y, int64_t n) {
  if (n < 8) { float z = x[0] * y[0]; for (int64_t i = 1; i < n; ++i) z += x[i] * y[i]; return z; }
  const int64_t end = n & ~(int64_t)7;
  float l[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int64_t i = 0; i < end; i += 8) for (int k = 0; k < 8; ++k) l[k] = l[k] + x[i + k] * y[i + k];
  float z = ((l[0] + l[1]) + (l[4] + l[5])) + ((l[2] + l[3]) + (l[6] + l[7]));
  for (int64_t i = end; i < n; ++i) z += x[i] * y[i];
  return z;
}

int oracle_softmax_grad(int dtype, const void* out, const void* out_grad, const int64_t* dims, int rank, int axis,
                        void* x_grad) {
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) numel *= dims[i];
  if (rank == 0) { if (dtype == DT_FLOAT32) *(float*)x_grad = 0.f; else *(double*)x_grad = 0.0; return 0; }
  if (numel == 0) return 0;
  if (axis < -rank || axis >= rank) return 1;
  if (axis < 0) axis += rank;
  int64_t outer = 1, dim = dims[axis], inner = 1;
  for (int i = 0; i < axis; ++i) outer *= dims[i];
  for (int i = axis + 1; i < rank; ++i) inner *= dims[i];
#define SMG_BODY(T, LAST_DOT)                                                              \
  {                                                                                        \
    const T* Y = (const T*)out; const T* G = (const T*)out_grad; T* DX = (T*)x_grad;       \
    for (int64_t o = 0; o < outer; ++o)                                                    \
      for (int64_t q = 0; q < inner; ++q) {                                                \
        const T* y = Y + o * dim * inner + q; const T* g = G + o * dim * inner + q;        \
        T* dx = DX + o * dim * inner + q;                                                  \
        if (inner == 1) {                                                                  \
          T dot = LAST_DOT;                                                                \
          dot = dot * (T)-1;                                                               \
          for (int64_t k = 0; k < dim; ++k) dx[k] = y[k] * (g[k] + dot);                   \
        } else {                                                                           \
          T dot = 0;                                                                       \
          for (int64_t k = 0; k < dim; ++k) dot += y[k * inner] * g[k * inner];            \
          for (int64_t k = 0; k < dim; ++k) dx[k * inner] = (g[k * inner] - dot) * y[k * inner]; \
        }                                                                                  \
      }                                                                                    \
  }
  if (dtype == DT_FLOAT32) SMG_BODY(float, vec_mul_reduce_avx_f32(g, y, dim))
  else {
#define SEQ_DOT_F64 ({ double z_ = g[0] * y[0]; for (int64_t i_ = 1; i_ < dim; ++i_) z_ += g[i_] * y[i_]; z_; })
    SMG_BODY(double, SEQ_DOT_F64)
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------
 * layer_norm_grad.  phi::LayerNormGradKernel<T,CPUContext> (paddle/phi/kernels/cpu/layer_norm_grad_kernel.cc:28-155):
 *   xhat = (x - mean) / sqrt(var + eps)                 (SubtractFunctor, DivAndSqrtFunctor)
 *   d_bias[j]  = sum over rows of dy[r][j]               (ColwiseSum: sequential over rows, math_function_impl.h)
 *   d_scale[j] = sum over rows of xhat[r][j] * dy[r][j]
 *   g = dy * scale  (when scale is given)                else g = dy
 *   dx = ((g - rowmean(g)) - xhat * rowmean(g * xhat)) / sqrt(var + eps)
 *   rowmean = sequential T sum times double 1/right (RowwiseMean).  Every step is a separate pass over
 *   memory, hence separately rounded.  T = float or double.
 *   (The CPU kernel multiplies by *scale exactly when d_scale is requested, :110-125; callers of this
 *   oracle request d_scale whenever scale is given, which is also what the GPU kernel's contract is.)
 * ---------------------------------------------------------------------------------------- */
int oracle_layer_norm_grad(int dtype, const void* x, const int64_t* dims, int rank, const void* scale,
                           const void* mean, const void* variance, const void* out_grad, float epsilon,
                           int begin_norm_axis, void* x_grad, void* scale_grad, void* bias_grad) {
  if (dtype != DT_FLOAT32 && dtype != DT_FLOAT64) return 2;
  if (begin_norm_axis < 0 || begin_norm_axis > rank) return 1;
  int64_t left = 1, right = 1;
  for (int i = 0; i < rank; ++i) { if (i < begin_norm_axis) left *= dims[i]; else right *= dims[i]; }
#define LNG_BODY(T, SQRT)                                                                          \
  {                                                                                                \
    const T* X = (const T*)x; const T* S = (const T*)scale; const T* M = (const T*)mean;           \
    const T* V = (const T*)variance; const T* G = (const T*)out_grad;                              \
    T* DX = (T*)x_grad; T* DS = (T*)scale_grad; T* DB = (T*)bias_grad;                             \
    const double inv = 1.0 / (double)right; const T eps = (T)epsilon;                              \
    T* xh = (T*)malloc(sizeof(T) * (size_t)(left * right)); T* tmp = (T*)malloc(sizeof(T) * (size_t)right); \
    for (int64_t r = 0; r < left; ++r) {                                                           \
      const T den = SQRT(V[r] + eps);                                                              \
      for (int64_t j = 0; j < right; ++j) { T d = X[r * right + j] - M[r]; xh[r * right + j] = d / den; } \
    }                                                                                              \
    if (DB) for (int64_t j = 0; j < right; ++j) { T s = 0; for (int64_t r = 0; r < left; ++r) s += G[r * right + j]; DB[j] = s; } \
    if (DS) for (int64_t j = 0; j < right; ++j) { T s = 0; for (int64_t r = 0; r < left; ++r) { T p = xh[r * right + j] * G[r * right + j]; s += p; } DS[j] = s; } \
    if (DX) for (int64_t r = 0; r < left; ++r) {                                                   \
      const T* g = G + r * right; const T* h = xh + r * right; T* dx = DX + r * right;             \
      T sum = 0;                                                                                   \
      for (int64_t j = 0; j < right; ++j) { tmp[j] = S ? (T)(g[j] * S[j]) : g[j]; sum += tmp[j]; } \
      const T m1 = (T)(sum * inv);                                                                 \
      sum = 0;                                                                                     \
      for (int64_t j = 0; j < right; ++j) { dx[j] = tmp[j] - m1; T p = tmp[j] * h[j]; tmp[j] = p; sum += p; } \
      const T m2 = (T)(sum * inv);                                                                 \
      const T den = SQRT(V[r] + eps);                                                              \
      for (int64_t j = 0; j < right; ++j) { T t = h[j] * m2; t = dx[j] - t; dx[j] = t / den; }     \
    }                                                                                              \
    free(xh); free(tmp);                                                                           \
  }
  if (dtype == DT_FLOAT32) LNG_BODY(float, sqrtf)
  else LNG_BODY(double, sqrt)
  return 0;
}

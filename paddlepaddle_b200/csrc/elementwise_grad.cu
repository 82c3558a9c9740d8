// elementwise_grad.cu — backward of add and multiply with broadcasting: a gradient that
// was broadcast in the forward pass is the SUM of dout over the broadcast dimensions.
//
// Replaces (fresh design): phi::AddGradKernel / MultiplyGradKernel on GPU —
//   paddle/phi/kernels/gpu/elementwise_grad.h:57-107   GetGradXAndYOut / GetGradXOrYOut (+ ReduceWrapper)
//   paddle/phi/kernels/gpu/elementwise_grad.h:148-232  DefaultElementwiseAddGrad / ElementwiseAddGrad
//   paddle/phi/kernels/gpu/elementwise_grad.h:371-399  ElementwiseMulGrad
//   paddle/phi/kernels/funcs/common_shape.h (GetReduceDim): which dimensions of dout an operand's gradient sums over
// The reference materialises dout*y and dout*x in dout-shaped temporaries (rounded to T) and reduces each with a
// second kernel.  Here:
//   * add_grad: same-shape gradients are one copy that reads dout once and writes both; a broadcast operand's
//     gradient is one run of the reduction engine over dout (reduce_impl.cuh).
//   * multiply_grad, the shape of config 2 (x[M,N] * y[N]): ONE kernel reads dout and x once, tile by tile through
//     the TMA tensor-map ring, writes dx = dout*y and accumulates dy = sum over rows of round_T(dout*x) in fp32 —
//     3 tensor passes of traffic instead of the reference's 7 (two products written and re-read).
//   * multiply_grad, any other broadcast: product into a workspace temporary (same rounding to T as the
//     reference), then the reduction engine.
// Sums run in MPType (fp32 for bf16/fp16) and are rounded once, like the reference's ReduceKernel.
#include "elementwise_impl.cuh"
#include "reduce_dispatch.cuh"

namespace pdb200 {

struct GradShape {
  int R = 0;
  int64_t od[PDB200_MAX_RANK];
  int64_t numel = 1, xnumel = 1, ynumel = 1;
  std::vector<int64_t> x_axes, y_axes;  // dims of dout that dx / dy sum over
};

// funcs::GetReduceDim with axis = -1: leading dims the operand lacks, and dims where it has 1 but dout has more.
inline std::string grad_shape(const int64_t* xd, int xr, const int64_t* yd, int yr, GradShape* g) {
  if (xr < 0 || yr < 0 || xr > PDB200_MAX_RANK || yr > PDB200_MAX_RANK) return "tensor rank must be in [0, 9]";
  g->R = std::max(xr, yr);
  for (int i = 0; i < g->R; ++i) {
    const int xi = i - (g->R - xr), yi = i - (g->R - yr);
    const int64_t a = xi >= 0 ? xd[xi] : 1, b = yi >= 0 ? yd[yi] : 1;
    if (a < 0 || b < 0) return "negative dimension";
    if (a != b && a != 1 && b != 1)
      return "Broadcast dimension mismatch. Operands could not be broadcast together with the shape of X and Y: " +
             std::to_string(a) + " != " + std::to_string(b) + " at output axis " + std::to_string(i);
    g->od[i] = a == 1 ? b : a;
    g->numel *= g->od[i];
    g->xnumel *= a;
    g->ynumel *= b;
    if (g->od[i] != 1) {
      if (a == 1) g->x_axes.push_back(i);
      if (b == 1) g->y_axes.push_back(i);
    }
  }
  return "";
}

// ---- dout -> dx and dy at once (same shapes): SimpleElemwiseAddGradCUDAKernel, elementwise_grad.h:115-145 ----
template <typename T>
__global__ void __launch_bounds__(kThreads) copy2_kernel(const T* __restrict__ src, T* __restrict__ a, T* __restrict__ b,
                                                         int64_t nvec, int64_t numel) {
  constexpr int V = Vec<T>::N;
  const int64_t base = (int64_t)blockIdx.x * (kThreads * kUnroll) + threadIdx.x;
  Vec<T> v[kUnroll];
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const int64_t i = base + u * kThreads;
    if (i < nvec) v[u] = ld_stream(src + i * V);
  }
#pragma unroll
  for (int u = 0; u < kUnroll; ++u) {
    const int64_t i = base + u * kThreads;
    if (i < nvec) {
      if (a) st_stream(a + i * V, v[u]);
      if (b) st_stream(b + i * V, v[u]);
    }
  }
  if (blockIdx.x == gridDim.x - 1) {
    const int64_t t = nvec * V + threadIdx.x;
    if (t < numel) {
      const T s = src[t];
      if (a) a[t] = s;
      if (b) b[t] = s;
    }
  }
}
template <typename T>
__global__ void __launch_bounds__(kThreads) copy2_scalar_kernel(const T* src, T* a, T* b, int64_t numel) {
  int64_t i = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * kThreads;
  for (; i < numel; i += stride) {
    const T s = src[i];
    if (a) a[i] = s;
    if (b) b[i] = s;
  }
}
// a and/or b may be NULL or equal to src (then that store is skipped: the gradient already lives there)
template <typename T>
int launch_copy2(const T* src, T* a, T* b, int64_t numel, cudaStream_t st) {
  if (a == src) a = nullptr;
  if (b == src) b = nullptr;
  if ((!a && !b) || numel == 0) return PDB200_OK;
  constexpr int V = Vec<T>::N;
  if (is_aligned16(src) && (!a || is_aligned16(a)) && (!b || is_aligned16(b))) {
    const int64_t nvec = numel / V;
    const int64_t tiles = std::max<int64_t>(1, (nvec + kThreads * kUnroll - 1) / (kThreads * kUnroll));
    copy2_kernel<T><<<(unsigned)tiles, kThreads, 0, st>>>(src, a, b, nvec, numel);
  } else {
    const int64_t blocks = std::min<int64_t>((numel + kThreads - 1) / kThreads, (int64_t)sm_count() * 32);
    copy2_scalar_kernel<T><<<(unsigned)blocks, kThreads, 0, st>>>(src, a, b, numel);
  }
  return after_launch("grad_copy");
}

// Sum `src` (dout-shaped) over `axes` into `dst`.
template <typename T>
int reduce_to(const GradShape& g, const std::vector<int64_t>& axes, const T* src, T* dst, void* ws, size_t ws_bytes,
              cudaStream_t st) {
  ReducePlan plan;
  std::string err = plan_reduce(g.od, g.R, axes.data(), (int)axes.size(), false, &plan);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
  using R = ReduceRunner<T, T, SumOp, kFinCast>;
  const size_t need = R::workspace(plan);
  if (need > ws_bytes)
    return fail(PDB200_WORKSPACE_TOO_SMALL, "elementwise grad: reduction needs " + std::to_string(need) +
                                                " workspace bytes, got " + std::to_string(ws_bytes));
  return R::run(plan, src, dst, ws, ws_bytes, st);
}
template <typename T>
size_t reduce_ws(const GradShape& g, const std::vector<int64_t>& axes) {
  ReducePlan plan;
  if (!plan_reduce(g.od, g.R, axes.data(), (int)axes.size(), false, &plan).empty()) return 0;
  return ReduceRunner<T, T, SumOp, kFinCast>::workspace(plan);
}

inline size_t align256(size_t v) { return (v + 255) & ~size_t(255); }

}  // namespace pdb200

namespace pdb200 {

// ---- multiply_grad for big[rows][cols] * row[cols] in one pass ---------------------------------------------------
// Grid = column tiles (512 bytes wide) x row splits, as in reduce_cols_tma_kernel.  A ring stage holds the [32 rows]
// x [512 B] box of dout and the matching box of `big`, both fetched by cp.async.bulk.tensor.2d on one mbarrier.
// Thread (vc, rg) owns vector-column vc of the tile and rows rg, rg+8, ... of each box: it multiplies dout by its
// (register-resident) slice of `row` and streams dbig out, and folds round_T(dout * big) into fp32 accumulators
// that meet in shared memory at the end.  part[split][cols] (accumulator type) is folded by a second tiny launch.
constexpr int kMgStages = 3;

template <typename T, bool WANT_DBIG>
__global__ void __launch_bounds__(kTmaThreads) mulgrad_inner_tma_kernel(const __grid_constant__ CUtensorMap map_dout,
                                                                        const __grid_constant__ CUtensorMap map_big,
                                                                        const T* __restrict__ rowv, T* __restrict__ dbig,
                                                                        typename AccType<T>::type* __restrict__ part,
                                                                        int64_t rows, int64_t cols, int64_t seg,
                                                                        int splits, int64_t col_tiles, int nstages) {
  using A = typename AccType<T>::type;
  using Sum = SumOp<A>;
  constexpr int V = Vec<T>::N;
  constexpr int TI = 32 * V;
  constexpr uint32_t kTileBytes = (uint32_t)kTmaTileRows * TI * sizeof(T);
  extern __shared__ __align__(128) unsigned char mg_ring_raw[];
  T* ring = reinterpret_cast<T*>(mg_ring_raw);  // [kMgStages][2][kTmaTileRows][TI]
  A* fold = reinterpret_cast<A*>(mg_ring_raw + (size_t)kMgStages * 2 * kTileBytes);  // [8][TI]
  __shared__ uint64_t full[kMgStages];
  const int tid = threadIdx.x;
  const int vc = tid & 31, rg = tid >> 5;
  int64_t b = blockIdx.x;
  const int64_t ct = b % col_tiles;
  const int split = (int)(b / col_tiles);
  const int64_t begin = (int64_t)split * seg;
  const int64_t end = begin + seg < rows ? begin + seg : rows;
  const int32_t c0 = (int32_t)(ct * TI);
  const int64_t col = (int64_t)c0 + vc * V;
  const bool col_ok = col < cols;  // cols % V == 0: a vector is entirely inside or outside
  const int64_t ntiles = (end - begin + kTmaTileRows - 1) / kTmaTileRows;
  if (tid == 0) {
    for (int s = 0; s < kMgStages; ++s) mbar_init(&full[s], 1);
    mbar_fence_init();
  }
  __syncthreads();
  auto issue = [&](int s, int64_t t) {
    T* dst = ring + (size_t)s * 2 * kTmaTileRows * TI;
    mbar_expect_tx(&full[s], 2 * kTileBytes);
    tma_load_2d(dst, &map_dout, c0, (int32_t)(begin + t * kTmaTileRows), &full[s]);
    tma_load_2d(dst + (size_t)kTmaTileRows * TI, &map_big, c0, (int32_t)(begin + t * kTmaTileRows), &full[s]);
  };
  if (tid == 0)
    for (int s = 0; s < nstages && s < ntiles; ++s) issue(s, s);
  Vec<T> yv;
  if (WANT_DBIG && col_ok) yv = ld_cached(rowv + col);
  A acc[V];
#pragma unroll
  for (int k = 0; k < V; ++k) acc[k] = Sum::identity();
  int s = 0;
  uint32_t phase = 0;
  for (int64_t t = 0; t < ntiles; ++t) {
    const int64_t r0 = begin + t * kTmaTileRows;
    const int nvalid = (int)(end - r0 < kTmaTileRows ? end - r0 : kTmaTileRows);
    mbar_wait(&full[s], phase);
    const T* tile_d = ring + (size_t)s * 2 * kTmaTileRows * TI;
    const T* tile_b = tile_d + (size_t)kTmaTileRows * TI;
#pragma unroll
    for (int i = 0; i < kTmaTileRows / 8; ++i) {
      const int r = rg + i * 8;
      if (r < nvalid) {
        const Vec<T> d = lds_vec(tile_d + (size_t)r * TI + vc * V);
        const Vec<T> g = lds_vec(tile_b + (size_t)r * TI + vc * V);
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = Sum::combine(acc[k], cvt<A>(MulOp<T>::apply(d.v[k], g.v[k])));
        if (WANT_DBIG && col_ok) {
          Vec<T> o;
#pragma unroll
          for (int k = 0; k < V; ++k) o.v[k] = MulOp<T>::apply(d.v[k], yv.v[k]);
          st_stream(dbig + (r0 + r) * cols + col, o);
        }
      }
    }
    __syncthreads();
    if (tid == 0 && t + nstages < ntiles) {
      fence_proxy_async_smem();
      issue(s, t + nstages);
    }
    if (++s == nstages) {
      s = 0;
      phase ^= 1;
    }
  }
#pragma unroll
  for (int k = 0; k < V; ++k) fold[(size_t)rg * TI + vc * V + k] = acc[k];
  __syncthreads();
  if (rg == 0 && col_ok) {
    A* dst = part + (int64_t)split * cols + col;
#pragma unroll
    for (int k = 0; k < V; ++k) {
      A a = fold[vc * V + k];
#pragma unroll
      for (int g = 1; g < 8; ++g) a = Sum::combine(a, fold[(size_t)g * TI + vc * V + k]);
      dst[k] = a;
    }
  }
}

template <typename T>
struct MulGradInner {
  using A = typename AccType<T>::type;
  // can the one-pass kernel take this shape?  (the tensor map needs 16-byte rows and 2/4/8-byte elements)
  static bool usable(int64_t rows, int64_t cols, const void* dout, const void* big, const void* rowv, const void* dbig) {
    constexpr int V = Vec<T>::N;
    if (!(sizeof(T) == 2 || sizeof(T) == 4 || sizeof(T) == 8)) return false;
    if (cols % V != 0 || cols / V < 32 || rows < 64) return false;
    if (rows >= (int64_t(1) << 31) || cols >= (int64_t(1) << 31)) return false;
    if (!is_aligned16(dout) || !is_aligned16(big) || !is_aligned16(rowv) || (dbig && !is_aligned16(dbig))) return false;
    return !knobs().no_fused_mulgrad && tma_encode_fn() != nullptr;
  }
  static PassGeom geometry(int64_t rows, int64_t cols) {
    PassGeom g = plan_pass(ReducePass{1, rows, cols}, sizeof(T), Vec<T>::N);
    // More, shorter CTAs than the read-only reduction wants: this kernel also WRITES dx, and write streams run best from
    // short-lived CTAs — 128 rows (4 boxes) per CTA, at most 1024 splits: bf16 [131072,4096] 0.503 -> 0.479 ms
    // (74 -> 1024 splits; 4096 splits of one box each: 0.673 ms).  PDB200_MULGRAD_SPLITS overrides.
    int64_t sp = std::min<int64_t>(1024, rows / 128);
    if (knobs().mulgrad_splits > 0) sp = knobs().mulgrad_splits;
    {
      if (sp > g.splits) {
        sp = std::min<int64_t>(sp, std::max<int64_t>(1, rows / kTmaTileRows));
        g.seg = (rows + sp - 1) / sp;
        g.seg = (g.seg + kTmaTileRows - 1) / kTmaTileRows * kTmaTileRows;
        g.splits = (int)((rows + g.seg - 1) / g.seg);
      }
    }
    return g;  // kind 3 by construction of usable(): TX = 32, TY = 8, splits / seg / col_tiles set
  }
  static ReducePlan fold_plan(int64_t rows, int64_t cols) {
    const int64_t dims[2] = {geometry(rows, cols).splits, cols};
    const int64_t axis0 = 0;
    ReducePlan p;
    plan_reduce(dims, 2, &axis0, 1, false, &p);
    if (p.passes.empty()) p.passes.push_back(ReducePass{p.numel, 1, 1});  // one split: the "fold" is just the rounding to T
    return p;
  }
  static size_t part_bytes(int64_t rows, int64_t cols) { return align256((size_t)geometry(rows, cols).splits * cols * sizeof(A)); }
  // workspace = the partials + whatever the fold of the partials asks for
  static size_t ws_bytes(int64_t rows, int64_t cols) {
    return part_bytes(rows, cols) + align256(reduce_scratch_bytes<A, A>(fold_plan(rows, cols)));
  }
  // dbig may be NULL (only the broadcast operand's gradient is wanted).
  static int run(int64_t rows, int64_t cols, const T* dout, const T* big, const T* rowv, T* dbig, T* drow, void* ws,
                 size_t ws_size, cudaStream_t st) {
    constexpr int TI = 32 * Vec<T>::N;
    const PassGeom g = geometry(rows, cols);
    const size_t need = ws_bytes(rows, cols);
    if (need > ws_size || (need > 0 && ws == nullptr))
      return fail(PDB200_WORKSPACE_TOO_SMALL, "multiply_grad: needs " + std::to_string(need) + " workspace bytes, got " +
                                                  std::to_string(ws_size));
    CUtensorMap md, mb;
    if (!make_cols_tensor_map(&md, dout, rows, cols, sizeof(T)) || !make_cols_tensor_map(&mb, big, rows, cols, sizeof(T)))
      return fail(PDB200_CUDA_ERROR, "multiply_grad: cuTensorMapEncodeTiled failed");
    const size_t smem = (size_t)kMgStages * 2 * kTmaTileRows * TI * sizeof(T) + (size_t)8 * TI * sizeof(A);
    A* part = static_cast<A*>(ws);
    const int64_t blocks = g.col_tiles * g.splits;
    const int nstages = knobs().mulgrad_stages >= 2 && knobs().mulgrad_stages <= kMgStages ? knobs().mulgrad_stages
                                                                                          : kMgStages;  // tuning knob
    if (dbig) {
      auto kern = mulgrad_inner_tma_kernel<T, true>;
      FuncAttrHold hold;  // one size per instantiation: set once per device, a map lookup afterwards
      func_attr_apply(reinterpret_cast<const void*>(kern), (int)smem, -1, &hold);
      kern<<<(unsigned)blocks, kTmaThreads, smem, st>>>(md, mb, rowv, dbig, part, rows, cols, g.seg, g.splits, g.col_tiles, nstages);
    } else {
      auto kern = mulgrad_inner_tma_kernel<T, false>;
      FuncAttrHold hold;
      func_attr_apply(reinterpret_cast<const void*>(kern), (int)smem, -1, &hold);
      kern<<<(unsigned)blocks, kTmaThreads, smem, st>>>(md, mb, rowv, dbig, part, rows, cols, g.seg, g.splits, g.col_tiles, nstages);
    }
    const int rc = after_launch("mulgrad_inner_tma");
    if (rc != PDB200_OK) return rc;
    // part[splits][cols] -> drow[cols] through the reduction engine (sum over axis 0, rounded once)
    const size_t pb = part_bytes(rows, cols);
    return run_reduce<A, T, SumOp<A>>(fold_plan(rows, cols), part, drow, CastFin<A, T>(), static_cast<char*>(ws) + pb,
                                      ws_size - pb, st);
  }
};

}  // namespace pdb200
namespace pdb200 {

// Is `small` (dims sd, rank sr) exactly the trailing dims of the output, and `big` the whole output?  Then
// out = big[rows][cols] (op) small[cols] with rows = product of the leading output dims.
inline bool inner_pattern(const GradShape& g, int64_t big_numel, const int64_t* sd, int sr, int64_t* rows, int64_t* cols) {
  if (big_numel != g.numel) return false;
  int k = 0;
  while (k < sr && sd[k] == 1) ++k;  // leading 1s of the small operand are broadcast dims too
  int64_t c = 1;
  for (int i = k; i < sr; ++i) {
    if (sd[i] != g.od[g.R - sr + i]) return false;
    c *= sd[i];
  }
  if (c <= 1 || g.numel % c != 0) return false;
  *cols = c;
  *rows = g.numel / c;
  return *rows > 1;
}

template <typename T>
struct GradRunner {
  // ---- add_grad ----
  static size_t add_ws(const GradShape& g, bool want_dx, bool want_dy) {
    size_t w = 0;
    if (want_dx && g.xnumel != g.numel) w = std::max(w, reduce_ws<T>(g, g.x_axes));
    if (want_dy && g.ynumel != g.numel) w = std::max(w, reduce_ws<T>(g, g.y_axes));
    return w;
  }
  static int add(const GradShape& g, const T* dout, T* dx, T* dy, void* ws, size_t ws_bytes, cudaStream_t st) {
    T* cx = (dx && g.xnumel == g.numel) ? dx : nullptr;
    T* cy = (dy && g.ynumel == g.numel) ? dy : nullptr;
    int rc = launch_copy2<T>(dout, cx, cy, g.numel, st);
    if (rc != PDB200_OK) return rc;
    if (dx && !cx) {
      rc = reduce_to<T>(g, g.x_axes, dout, dx, ws, ws_bytes, st);
      if (rc != PDB200_OK) return rc;
    }
    if (dy && !cy) rc = reduce_to<T>(g, g.y_axes, dout, dy, ws, ws_bytes, st);
    return rc;
  }

  // ---- multiply_grad ----
  // layout of the general route's workspace: [ product temporary (dout-shaped, T) | reduction scratch ]
  static size_t mul_ws(const GradShape& g, const int64_t* xd, int xr, const int64_t* yd, int yr, bool want_dx,
                       bool want_dy) {
    int64_t rows, cols;
    size_t fused = 0;
    if (want_dy && inner_pattern(g, g.xnumel, yd, yr, &rows, &cols)) fused = std::max(fused, MulGradInner<T>::ws_bytes(rows, cols));
    if (want_dx && inner_pattern(g, g.ynumel, xd, xr, &rows, &cols)) fused = std::max(fused, MulGradInner<T>::ws_bytes(rows, cols));
    size_t red = 0;
    if (want_dx && g.xnumel != g.numel) red = std::max(red, reduce_ws<T>(g, g.x_axes));
    if (want_dy && g.ynumel != g.numel) red = std::max(red, reduce_ws<T>(g, g.y_axes));
    const bool any_reduce = (want_dx && g.xnumel != g.numel) || (want_dy && g.ynumel != g.numel);
    const size_t general = any_reduce ? align256((size_t)g.numel * sizeof(T)) + align256(red) : 0;
    return std::max(fused, general);  // the fused kernel may be refused at run time (alignment): size for both
  }
  // grad of one operand: d_self = sum over self_axes of (dout * other)
  static int mul_one(const GradShape& g, const T* dout, const int64_t* od, const T* other, const int64_t* other_d,
                     int other_r, int64_t self_numel, const std::vector<int64_t>& self_axes, T* d_self, void* ws,
                     size_t ws_bytes, cudaStream_t st) {
    BroadcastPlan bp;
    std::string err = plan_broadcast(od, g.R, other_d, other_r, &bp);
    if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
    if (self_numel == g.numel) return launch_elementwise<T, MulOp<T>>(bp, dout, other, d_self, st);
    const size_t tmp_bytes = align256((size_t)g.numel * sizeof(T));
    if (ws == nullptr || ws_bytes < tmp_bytes)
      return fail(PDB200_WORKSPACE_TOO_SMALL, "multiply_grad: needs " + std::to_string(tmp_bytes) +
                                                  "+ workspace bytes, got " + std::to_string(ws_bytes) +
                                                  " (ask pdb200_elementwise_grad_workspace_bytes)");
    T* tmp = static_cast<T*>(ws);
    int rc = launch_elementwise<T, MulOp<T>>(bp, dout, other, tmp, st);
    if (rc != PDB200_OK) return rc;
    return reduce_to<T>(g, self_axes, tmp, d_self, static_cast<char*>(ws) + tmp_bytes, ws_bytes - tmp_bytes, st);
  }
  static int mul(const GradShape& g, const T* x, const int64_t* xd, int xr, const T* y, const int64_t* yd, int yr,
                 const T* dout, T* dx, T* dy, void* ws, size_t ws_bytes, cudaStream_t st) {
    int64_t rows, cols;
    // x[rows][cols] * y[cols]: dy (and dx, if wanted) in one pass over dout and x
    if (dy && inner_pattern(g, g.xnumel, yd, yr, &rows, &cols) && MulGradInner<T>::usable(rows, cols, dout, x, y, dx) &&
        ws_bytes >= MulGradInner<T>::ws_bytes(rows, cols))
      return MulGradInner<T>::run(rows, cols, dout, x, y, dx, dy, ws, ws_bytes, st);
    // x[cols] * y[rows][cols]: the mirror image
    if (dx && inner_pattern(g, g.ynumel, xd, xr, &rows, &cols) && MulGradInner<T>::usable(rows, cols, dout, y, x, dy) &&
        ws_bytes >= MulGradInner<T>::ws_bytes(rows, cols))
      return MulGradInner<T>::run(rows, cols, dout, y, x, dy, dx, ws, ws_bytes, st);
    int rc = PDB200_OK;
    if (dx) rc = mul_one(g, dout, g.od, y, yd, yr, g.xnumel, g.x_axes, dx, ws, ws_bytes, st);
    if (rc != PDB200_OK) return rc;
    if (dy) rc = mul_one(g, dout, g.od, x, xd, xr, g.ynumel, g.y_axes, dy, ws, ws_bytes, st);
    return rc;
  }
};

// gpu/elementwise_grad_kernel.cu:361-372 (add_grad), :431-443 (multiply_grad); complex is out of scope, and bool
// multiply_grad (a sum of bools kept as bool) stays with the in-tree kernel.
#define PDB_GRAD_DTYPES(M) M(PDB200_FLOAT32) M(PDB200_FLOAT64) M(PDB200_INT32) M(PDB200_INT64) M(PDB200_FLOAT16) M(PDB200_BFLOAT16)

}  // namespace pdb200

extern "C" {

size_t pdb200_elementwise_grad_workspace_bytes(int op, const int64_t* x_dims, int x_rank, const int64_t* y_dims,
                                               int y_rank, int dtype, int want_dx, int want_dy) {
  using namespace pdb200;
  GradShape g;
  if ((x_rank > 0 && !x_dims) || (y_rank > 0 && !y_dims) || !grad_shape(x_dims, x_rank, y_dims, y_rank, &g).empty()) return 0;
  if (g.numel == 0) return 0;
  switch (dtype) {
#define M(DT)                                                                                                   \
  case DT: {                                                                                                    \
    using T = typename CType<DT>::type;                                                                         \
    return op == PDB200_GRAD_ADD ? GradRunner<T>::add_ws(g, want_dx != 0, want_dy != 0)                         \
                                 : GradRunner<T>::mul_ws(g, x_dims, x_rank, y_dims, y_rank, want_dx != 0, want_dy != 0); \
  }
    PDB_GRAD_DTYPES(M)
#undef M
  }
  return 0;
}

int pdb200_add_grad(const int64_t* x_dims, int x_rank, const int64_t* y_dims, int y_rank, const void* dout, int dtype,
                    void* dx, void* dy, void* workspace, size_t workspace_bytes, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(x_rank == 0 || x_dims != nullptr, "x_dims is NULL");
  PDB_CHECK_ARG(y_rank == 0 || y_dims != nullptr, "y_dims is NULL");
  GradShape g;
  std::string err = grad_shape(x_dims, x_rank, y_dims, y_rank, &g);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                  \
  case DT: {                                                                                                   \
    using T = typename CType<DT>::type;                                                                        \
    if (g.numel == 0 || (!dx && !dy)) return PDB200_OK;                                                        \
    PDB_CHECK_ARG(dout != nullptr, "NULL tensor pointer");                                                     \
    return GradRunner<T>::add(g, static_cast<const T*>(dout), static_cast<T*>(dx), static_cast<T*>(dy),        \
                              workspace, workspace_bytes, st);                                                 \
  }
    PDB_GRAD_DTYPES(M)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "add_grad: dtype code " + std::to_string(dtype) + " is not registered for this op");
}

int pdb200_multiply_grad(const void* x, const int64_t* x_dims, int x_rank, const void* y, const int64_t* y_dims,
                         int y_rank, const void* dout, int dtype, void* dx, void* dy, void* workspace,
                         size_t workspace_bytes, void* stream) {
  using namespace pdb200;
  PDB_CHECK_ARG(x_rank == 0 || x_dims != nullptr, "x_dims is NULL");
  PDB_CHECK_ARG(y_rank == 0 || y_dims != nullptr, "y_dims is NULL");
  GradShape g;
  std::string err = grad_shape(x_dims, x_rank, y_dims, y_rank, &g);
  if (!err.empty()) return fail(PDB200_INVALID_ARGUMENT, err);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  switch (dtype) {
#define M(DT)                                                                                                  \
  case DT: {                                                                                                   \
    using T = typename CType<DT>::type;                                                                        \
    if (g.numel == 0 || (!dx && !dy)) return PDB200_OK;                                                        \
    PDB_CHECK_ARG(dout != nullptr && (x != nullptr || !dy) && (y != nullptr || !dx), "NULL tensor pointer");   \
    return GradRunner<T>::mul(g, static_cast<const T*>(x), x_dims, x_rank, static_cast<const T*>(y), y_dims,   \
                              y_rank, static_cast<const T*>(dout), static_cast<T*>(dx), static_cast<T*>(dy),   \
                              workspace, workspace_bytes, st);                                                 \
  }
    PDB_GRAD_DTYPES(M)
#undef M
  }
  return fail(PDB200_UNSUPPORTED, "multiply_grad: dtype code " + std::to_string(dtype) + " is not registered for this op");
}

}  // extern "C"

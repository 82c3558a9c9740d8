// shape.h — host-side launch planning: everything index-related is resolved here, once
// per call, so that the device inner loops are pure streaming (no div/mod per element).
//
// Replaces, as a fresh design (not a translation):
//   BroadcastDimsSimplifier   paddle/phi/kernels/funcs/dims_simplifier.h:25-229
//   BroadcastTypeClassifier   paddle/phi/kernels/funcs/broadcast_function.h:35-113
//   ReduceConfig::SetReduceDim/SetReduceType  paddle/phi/kernels/funcs/reduce_function.h:301-431
//   GetReduceDim / recompute_reduce_all       reduce_function.h:100-123, core/kernel_utils.h:407-417
#pragma once

#include <stdint.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/pdb200.h"

namespace pdb200 {

inline std::string dims_str(const int64_t* d, int r) {
  std::string s = "[";
  for (int i = 0; i < r; ++i) {
    if (i) s += ", ";
    s += std::to_string(d[i]);
  }
  return s + "]";
}

// ------------------------------------------------------------------------------------
// Broadcast planning
// ------------------------------------------------------------------------------------
enum BroadcastKind {
  kBcastEmpty = 0,   // out numel == 0: nothing to launch
  kBcastSame = 1,    // both operands cover every output element: 1-D stream
  kBcastInner = 2,   // one operand is a row vector repeated over rows:  big[M,N] (op) small[N]
  kBcastOuter = 3,   // one operand is one value per row (column / scalar): big[M,N] (op) small[M]
  kBcastGeneral = 4  // anything else: merged N-D strided walk
};

struct BroadcastPlan {
  int kind = kBcastEmpty;
  bool small_is_x = false;  // for Inner/Outer: the broadcast ("small") operand is x, not y
  int64_t numel = 0;        // output elements
  int64_t rows = 1, cols = 1;  // Inner/Outer view of the output: [rows, cols]
  // General: merged dims, innermost first; element strides per operand (0 = broadcast)
  int rank = 0;
  int64_t dims[PDB200_MAX_RANK] = {0};
  int64_t xs[PDB200_MAX_RANK] = {0};
  int64_t ys[PDB200_MAX_RANK] = {0};
  // result of shape inference
  int out_rank = 0;
  int64_t out_dims[PDB200_MAX_RANK] = {0};
};

// Returns "" on success, else an error message (shapes not broadcastable).
inline std::string plan_broadcast(const int64_t* xd, int xr, const int64_t* yd, int yr,
                                  BroadcastPlan* p) {
  if (xr < 0 || yr < 0 || xr > PDB200_MAX_RANK || yr > PDB200_MAX_RANK)
    return "tensor rank must be in [0, " + std::to_string(PDB200_MAX_RANK) + "]";
  const int R = std::max(xr, yr);
  p->out_rank = R;
  int64_t X[PDB200_MAX_RANK], Y[PDB200_MAX_RANK];
  for (int i = 0; i < R; ++i) {  // right-aligned, i counts from the outermost output dim
    int xi = i - (R - xr), yi = i - (R - yr);
    X[i] = xi >= 0 ? xd[xi] : 1;
    Y[i] = yi >= 0 ? yd[yi] : 1;
    if (X[i] < 0 || Y[i] < 0) return "negative dimension";
    if (X[i] != Y[i] && X[i] != 1 && Y[i] != 1) {
      // same wording family as funcs::GetBroadcastDimsArrays (common_shape.h:35-93)
      return "Broadcast dimension mismatch. Operands could not be broadcast together with the shape of "
             "X = " + dims_str(xd, xr) + " and the shape of Y = " + dims_str(yd, yr) +
             ". Received [" + std::to_string(X[i]) + "] in X is not equal to [" + std::to_string(Y[i]) +
             "] in Y at i:" + std::to_string(i) + ".";
    }
    // a 0-sized dim wins over 1 (numpy rule: 1 broadcasts to 0)
    p->out_dims[i] = (X[i] == 1) ? Y[i] : ((Y[i] == 1) ? X[i] : X[i]);
  }
  int64_t numel = 1;
  for (int i = 0; i < R; ++i) numel *= p->out_dims[i];
  p->numel = numel;
  if (numel == 0) {
    p->kind = kBcastEmpty;
    return "";
  }
  // Merge, innermost first.  flag bit0: x broadcasts along this dim, bit1: y does.
  int n = 0;
  int64_t md[PDB200_MAX_RANK];
  int mf[PDB200_MAX_RANK];
  for (int i = R - 1; i >= 0; --i) {
    const int64_t o = p->out_dims[i];
    if (o == 1) continue;  // contributes nothing to any index
    const int f = (X[i] == 1 ? 1 : 0) | (Y[i] == 1 ? 2 : 0);
    if (n > 0 && mf[n - 1] == f) {
      md[n - 1] *= o;
    } else {
      md[n] = o;
      mf[n] = f;
      ++n;
    }
  }
  p->rank = n;
  int64_t sx = 1, sy = 1;
  for (int i = 0; i < n; ++i) {
    p->dims[i] = md[i];
    p->xs[i] = (mf[i] & 1) ? 0 : sx;
    p->ys[i] = (mf[i] & 2) ? 0 : sy;
    if (!(mf[i] & 1)) sx *= md[i];
    if (!(mf[i] & 2)) sy *= md[i];
  }
  if (n == 0) {  // every dim is 1: a single element
    p->kind = kBcastSame;
    return "";
  }
  if (n == 1) {
    if (mf[0] == 0) {
      p->kind = kBcastSame;
    } else {  // one side is a single value
      p->kind = kBcastOuter;
      p->small_is_x = (mf[0] == 1);
      p->rows = 1;
      p->cols = md[0];
    }
    return "";
  }
  if (n == 2) {
    // md[0] = cols (inner), md[1] = rows (outer)
    const int fi = mf[0], fo = mf[1];
    p->rows = md[1];
    p->cols = md[0];
    if (fi == 0 && (fo == 1 || fo == 2)) {  // small = row vector [cols]
      p->kind = kBcastInner;
      p->small_is_x = (fo == 1);
      return "";
    }
    if (fo == 0 && (fi == 1 || fi == 2)) {  // small = column [rows]
      p->kind = kBcastOuter;
      p->small_is_x = (fi == 1);
      return "";
    }
  }
  p->kind = kBcastGeneral;
  return "";
}

// ------------------------------------------------------------------------------------
// Reduce planning
// ------------------------------------------------------------------------------------
struct ReducePass {
  int64_t outer, red, inner;  // x viewed as [outer][red][inner] -> out [outer][inner]
};

struct ReducePlan {
  int64_t numel = 0;       // input elements
  int64_t reduce_num = 1;  // elements folded into each output
  int64_t out_numel = 1;
  bool reduce_all = false;
  // Passes are executed in order; each consumes the previous pass's output.  An empty list
  // means reduce_num == 1 (pure cast/copy).
  std::vector<ReducePass> passes;
};

inline std::string plan_reduce(const int64_t* dims, int rank, const int64_t* axes, int n_axes,
                               bool reduce_all, ReducePlan* p) {
  if (rank < 0 || rank > PDB200_MAX_RANK) return "tensor rank must be in [0, 9]";
  bool red[PDB200_MAX_RANK] = {false};
  // recompute_reduce_all (core/kernel_utils.h:407-417): empty axes, 0-D, or full axis set
  if (rank == 0 || n_axes == 0) reduce_all = true;
  for (int i = 0; i < n_axes; ++i) {
    int64_t a = axes[i];
    if (a < -rank || a >= std::max(rank, 1))
      return "The reduce dim index " + std::to_string(i) + " should be in the range [ -dimension(X), "
             "dimension(X) ) which dimension = " + std::to_string(rank) + ". But received dim index = " +
             std::to_string(a) + ".";
    if (rank > 0) red[a < 0 ? a + rank : a] = true;  // duplicates collapse, as the std::set does
  }
  if (reduce_all)
    for (int i = 0; i < rank; ++i) red[i] = true;
  int64_t numel = 1;
  for (int i = 0; i < rank; ++i) {
    if (dims[i] < 0) return "negative dimension";
    numel *= dims[i];
  }
  p->numel = numel;
  if (numel <= 0) return "Tensor need be reduced must not empty.";  // reduce_function.h:1052-1055
  // merge: drop size-1 dims, fuse neighbours with the same flag (outermost first)
  std::vector<int64_t> md;
  std::vector<bool> mr;
  for (int i = 0; i < rank; ++i) {
    if (dims[i] == 1) continue;
    if (!md.empty() && mr.back() == red[i]) {
      md.back() *= dims[i];
    } else {
      md.push_back(dims[i]);
      mr.push_back(red[i]);
    }
  }
  int64_t rn = 1, on = 1;
  for (size_t i = 0; i < md.size(); ++i) (mr[i] ? rn : on) *= md[i];
  p->reduce_num = rn;
  p->out_numel = on;
  p->reduce_all = (on == 1);
  p->passes.clear();
  // fold reduce runs from the innermost outwards; each fold is one [outer][red][inner] pass
  while (true) {
    int last = -1;
    for (int i = static_cast<int>(md.size()) - 1; i >= 0; --i)
      if (mr[i]) { last = i; break; }
    if (last < 0) break;
    ReducePass ps{1, md[last], 1};
    for (int i = 0; i < last; ++i) ps.outer *= md[i];
    for (size_t i = last + 1; i < md.size(); ++i) ps.inner *= md[i];
    p->passes.push_back(ps);
    md.erase(md.begin() + last);
    mr.erase(mr.begin() + last);
    // neighbours that are both kept can now fuse (purely cosmetic: passes use products)
  }
  return "";
}

}  // namespace pdb200

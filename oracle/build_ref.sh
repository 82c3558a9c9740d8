#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds oracle/_ref/libphi_cpu_ref.so: the reference's
# own CPU PHI kernels for the hot path, compiled IN PLACE from /root/reference
# (sources are never copied into this repo) plus oracle/ref_shim.cc (extern "C"
# doorway, no algorithm of its own).  Only g++ on the reference's own files — its
# cmake build system is not run.  Output goes to oracle/_ref/ only (git-ignored,
# NOT gpurun-ignored, so the .so travels to the GPU box where /root/reference
# does not exist).
#
#   usage: oracle/build_ref.sh [reference_root]     (default /root/reference)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
R="${1:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$R/paddle/phi" ]; then
  echo "build_ref.sh: reference tree not found at $R (expected on the GPU box) - keeping prebuilt $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
# -O3 -DNDEBUG + AVX2/FMA = Paddle release flags on x86 (cmake/simd.cmake, cmake/flags.cmake)
FLAGS="-std=c++17 -O3 -DNDEBUG -mavx2 -mfma -fPIC -w -DPADDLE_USE_REFERENCE_CBLAS -I$HERE/stub -I$R -I$R/third_party/eigen3"
TUS="
paddle/common/ddim.cc
paddle/common/enforce.cc
paddle/common/errors.cc
paddle/common/flags.cc
paddle/common/flags_native.cc
paddle/phi/common/place.cc
paddle/phi/common/memory_utils.cc
paddle/phi/core/dense_tensor.cc
paddle/phi/core/dense_tensor_impl.cc
paddle/phi/core/tensor_meta.cc
paddle/phi/core/tensor_base.cc
paddle/phi/core/storage_properties.cc
paddle/phi/core/utils/type_info.cc
paddle/phi/core/device_context.cc
paddle/phi/core/enforce.cc
paddle/phi/backends/cpu/cpu_context.cc
paddle/phi/backends/cpu/cpu_info.cc
paddle/phi/kernels/funcs/math_function.cc
paddle/phi/kernels/funcs/eigen/scale.cc
paddle/phi/kernels/funcs/cross_entropy.cc
"
pids=()
build_one() {  # $1 = source path, $2 = object
  if [ ! -f "$2" ] || [ "$1" -nt "$2" ]; then
    g++ $FLAGS -c "$1" -o "$2" || { rm -f "$2"; return 1; }  # no stale object left for the link step
  fi
}
JOBS="${JOBS:-$(nproc)}"
running=0
for f in $TUS; do
  o="$OUT/$(echo "$f" | tr '/' '_' | sed 's/\.cc$/.o/')"
  build_one "$R/$f" "$o" &
  running=$((running + 1))
  if [ "$running" -ge "$JOBS" ]; then wait -n; running=$((running - 1)); fi
done
build_one "$HERE/ref_shim.cc" "$OUT/ref_shim.o" &
build_one "$HERE/ref_shim_family.cc" "$OUT/ref_shim_family.o" &
# the reference's AVX-512 kernel (fused_bias_residual_layernorm on CPU): its own TU with the instruction set enabled; the
# shim checks the CPU before entering it
FLAGS="$FLAGS -mavx512f -mavx512dq -mavx512bw -mavx512vl" build_one "$HERE/ref_shim_avx512.cc" "$OUT/ref_shim_avx512.o" &
wait
g++ -shared -o "$OUT/libphi_cpu_ref.so" "$OUT"/*.o -Wl,--no-undefined -lpthread -ldl
echo "built $OUT/libphi_cpu_ref.so"

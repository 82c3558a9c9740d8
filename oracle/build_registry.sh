#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds, from the reference's own sources compiled in place (no cmake, no
# copied files), a minimal host for Paddle's kernel registry and the Paddle-side binding of this repo:
#   oracle/_ref/libphi_registry_min.so  KernelFactory / KernelRegistrar / DenseTensor / ... (+ oracle/phi_registry_shim.cc)
#   oracle/_ref/libpdb200_phi.so        paddlepaddle_b200/plugin/phi_kernels.cc compiled against the reference
#                                       headers exactly as a maintainer would (-DPADDLE_WITH_CUDA
#                                       -DPADDLE_WITH_CUSTOM_KERNEL), linked to libpdb200.so
# tests/test_phi_registration.py dlopens both and checks that PD_REGISTER_KERNEL put the B200 kernels under
# Paddle's own keys.  usage: oracle/build_registry.sh [reference_root]
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(dirname "$HERE")"
R="${1:-/root/reference}"
OUT="$HERE/_ref"
OBJ="$OUT/registry"
if [ ! -d "$R/paddle/phi" ]; then
  echo "build_registry.sh: reference tree not found at $R - keeping prebuilt libraries" >&2
  exit 0
fi
mkdir -p "$OBJ" "$OUT/plugin"
CUDA_INC="${CUDA_HOME:-/usr/local/cuda}/include"
CUDA_LIB="${CUDA_HOME:-/usr/local/cuda}/lib64"
FLAGS="-std=c++17 -O1 -DNDEBUG -fPIC -w -DPADDLE_WITH_CUDA -DPADDLE_USE_REFERENCE_CBLAS -I$ROOT/include -I$HERE/stub -I$R -I$R/third_party/eigen3 -I$CUDA_INC"
TUS="
paddle/common/ddim.cc paddle/common/enforce.cc paddle/common/errors.cc paddle/common/flags.cc paddle/common/flags_native.cc
paddle/phi/common/place.cc paddle/phi/common/memory_utils.cc paddle/phi/common/int_array.cc paddle/phi/common/scalar.cc
paddle/phi/core/dense_tensor.cc paddle/phi/core/dense_tensor_impl.cc paddle/phi/core/tensor_meta.cc paddle/phi/core/tensor_base.cc
paddle/phi/core/storage_properties.cc paddle/phi/core/utils/type_info.cc paddle/phi/core/device_context.cc
paddle/phi/core/kernel_factory.cc paddle/phi/core/kernel_context.cc paddle/phi/core/kernel_registry.cc paddle/phi/core/custom_kernel.cc
paddle/phi/core/compat/op_utils.cc paddle/phi/core/compat/convert_utils.cc
paddle/phi/core/extended_tensor.cc paddle/phi/core/sparse_coo_tensor.cc paddle/phi/core/sparse_csr_tensor.cc
paddle/phi/core/string_tensor.cc paddle/phi/core/tensor_array.cc paddle/phi/backends/cpu/cpu_context.cc
"
# a failed compile removes its (stale) object, so the link below fails instead of silently reusing it
build_one() { if [ ! -f "$2" ] || [ "$1" -nt "$2" ]; then g++ $FLAGS $3 -c "$1" -o "$2" || { rm -f "$2"; return 1; }; fi; }
JOBS="${JOBS:-$(nproc)}"
running=0
for f in $TUS; do
  o="$OBJ/$(echo "$f" | tr '/' '_' | sed 's/\.cc$/.o/')"
  build_one "$R/$f" "$o" "" &
  running=$((running + 1))
  if [ "$running" -ge "$JOBS" ]; then wait -n; running=$((running - 1)); fi
done
build_one "$HERE/phi_registry_shim.cc" "$OBJ/shim.o" "" &
# the binding itself is an out-of-tree build: PADDLE_WITH_CUSTOM_KERNEL as paddle.utils.cpp_extension passes it
build_one "$ROOT/paddlepaddle_b200/plugin/phi_kernels.cc" "$OUT/plugin/phi_kernels.o" "-DPADDLE_WITH_CUSTOM_KERNEL" &
wait
g++ -shared -o "$OUT/libphi_registry_min.so" "$OBJ"/*.o -Wl,--no-undefined -L"$CUDA_LIB" -lcudart -lpthread -ldl
g++ -shared -o "$OUT/libpdb200_phi.so" "$OUT/plugin/phi_kernels.o" -L"$ROOT/paddlepaddle_b200" -lpdb200 \
    -Wl,-rpath,'$ORIGIN/../../paddlepaddle_b200'
echo "built $OUT/libphi_registry_min.so and $OUT/libpdb200_phi.so"

#!/usr/bin/env bash
# Builds the UNMODIFIED reference backend (mast3r_slam_backends pybind module)
# from the sources where they lie under /root/reference into oracle/_ref/.
# Test / baseline infrastructure only -- nothing under oracle/ is product code.
# Usage: oracle/build_ref.sh [REFERENCE_ROOT]   (default /root/reference)
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${1:-/root/reference}"
SRC="$REF/mast3r_slam/backend"
OUT="$HERE/_ref"
[ -d "$SRC/src" ] || { echo "reference sources not found at $SRC" >&2; exit 3; }
mkdir -p "$OUT/obj"
PY=python
TORCH_DIR="$($PY -c 'import torch,os;print(os.path.dirname(torch.__file__))')"
PY_INC="$($PY -c 'import sysconfig;print(sysconfig.get_paths()["include"])')"
EXT="$($PY -c 'import sysconfig;print(sysconfig.get_config_var("EXT_SUFFIX"))')"
COMMON=(-O3 -std=c++17 -DTORCH_EXTENSION_NAME=mast3r_slam_backends -DTORCH_API_INCLUDE_EXTENSION_H
  -D_GLIBCXX_USE_CXX11_ABI=1
  -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include" -I"$PY_INC"
  -I"$SRC/include" -I"$REF/thirdparty/eigen")
NVFLAGS=(-gencode=arch=compute_100,code=sm_100 -lineinfo
  -D__CUDA_NO_HALF_OPERATORS__ -D__CUDA_NO_HALF_CONVERSIONS__
  -D__CUDA_NO_BFLOAT16_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__
  --expt-relaxed-constexpr -include "$HERE/ref_shim.h" -Xcompiler -fPIC)
build_cu() { # name
  if [ ! -f "$OUT/obj/$1.o" ] || [ "$SRC/src/$1.cu" -nt "$OUT/obj/$1.o" ]; then
    nvcc "${COMMON[@]}" "${NVFLAGS[@]}" -c "$SRC/src/$1.cu" -o "$OUT/obj/$1.o"
  fi
}
build_cu matching_kernels &
build_cu gn_kernels &
if [ ! -f "$OUT/obj/gn.o" ]; then
  g++ "${COMMON[@]}" -fPIC -include "$HERE/ref_shim.h" -c "$SRC/src/gn.cpp" -o "$OUT/obj/gn.o"
fi
wait
g++ -shared "$OUT/obj/gn.o" "$OUT/obj/gn_kernels.o" "$OUT/obj/matching_kernels.o" \
  -L"$TORCH_DIR/lib" -lc10 -ltorch -ltorch_cpu -ltorch_python -lc10_cuda -ltorch_cuda \
  -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,"$TORCH_DIR/lib" \
  -o "$OUT/mast3r_slam_backends$EXT"
echo "built $OUT/mast3r_slam_backends$EXT"

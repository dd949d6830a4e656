#!/usr/bin/env bash
# Build the reference's OWN CUDA correlation (models/correlation_package) for sm_100a into oracle/_ref/ -- TEST
# INFRASTRUCTURE: the second oracle of the cost-volume path and the GPU baseline of bench.py, never the product path.
# Runs in the build container only (/root/reference is absent on the GPU box; the outputs travel with the snapshot).
#   * correlation_cuda.cc and correlation_cuda_kernel.cuh are compiled from where they lie under /root/reference;
#   * correlation_cuda_kernel.cu needs the 7-token F8 fix (AT_DISPATCH_*(x.type(), ...) no longer
#     compiles under torch >= 2.x; `.type()` -> `.scalar_type()` at :386,393,403,495,507,524,541): the fix is applied
#     with sed to a scratch copy under oracle/_ref/build/ (a build artefact, git-ignored) -- no reference file is modified;
#   * models/PWCNet.py, models/correlation_package/correlation.py are copied UNMODIFIED to oracle/_ref/refpkg/ so that the
#     GPU box can run the reference's own forward (bench.py --impl reference, gpu_baseline, tests/test_gpu_reference.py).
# Outputs only into oracle/_ref/ (listed in .gitignore, not in .gpurunignore).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${PWC_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
PKG="$REF/models/correlation_package"
[ -f "$PKG/correlation_cuda_kernel.cu" ] || { echo "build_ref: reference tree not found at $REF" >&2; exit 0; }
mkdir -p "$OUT/build" "$OUT/refpkg/models/correlation_package"
PY="${PYTHON:-python}"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
read -r TORCH_INC1 TORCH_INC2 TORCH_LIB PY_INC EXT_SUFFIX < <("$PY" - <<'PYEOF'
import sysconfig, os, torch
r = os.path.dirname(torch.__file__)
print(os.path.join(r, "include"), os.path.join(r, "include/torch/csrc/api/include"), os.path.join(r, "lib"),
      sysconfig.get_paths()["include"], sysconfig.get_config_var("EXT_SUFFIX"))
PYEOF
)
SO="$OUT/correlation_cuda$EXT_SUFFIX"
STAMP="$OUT/build/stamp"
NEWEST=$(stat -c %Y "$PKG/correlation_cuda_kernel.cu" "$PKG/correlation_cuda.cc" "$PKG/correlation_cuda_kernel.cuh" "$0" | sort -n | tail -1)
if [ -f "$SO" ] && [ -f "$STAMP" ] && [ "$(cat "$STAMP")" = "$NEWEST" ]; then
  echo "build_ref: $SO is up to date"
else
  sed -E 's/AT_DISPATCH_FLOATING_TYPES_AND_HALF\(([A-Za-z0-9_]+)\.type\(\)/AT_DISPATCH_FLOATING_TYPES_AND_HALF(\1.scalar_type()/' \
      "$PKG/correlation_cuda_kernel.cu" > "$OUT/build/correlation_cuda_kernel.cu"
  [ "$(diff "$PKG/correlation_cuda_kernel.cu" "$OUT/build/correlation_cuda_kernel.cu" | grep -c '^>')" = 7 ] || \
      { echo "build_ref: expected exactly 7 patched lines" >&2; exit 1; }
  DEFS="-DTORCH_EXTENSION_NAME=correlation_cuda -DTORCH_API_INCLUDE_EXTENSION_H -D_GLIBCXX_USE_CXX11_ABI=1"
  INCS="-I$PKG -isystem $TORCH_INC1 -isystem $TORCH_INC2 -isystem /usr/local/cuda/include -isystem $PY_INC"
  "$NVCC" -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -w $DEFS $INCS \
      -c "$OUT/build/correlation_cuda_kernel.cu" -o "$OUT/build/correlation_cuda_kernel.o" &
  g++ -O2 -std=c++17 -fPIC -w $DEFS $INCS -c "$PKG/correlation_cuda.cc" -o "$OUT/build/correlation_cuda.o" &
  wait
  g++ -shared -o "$SO" "$OUT/build/correlation_cuda.o" "$OUT/build/correlation_cuda_kernel.o" \
      -L"$TORCH_LIB" -L/usr/local/cuda/lib64 -lc10 -lc10_cuda -ltorch_cpu -ltorch_cuda -ltorch -ltorch_python -lcudart \
      -Wl,-rpath,"$TORCH_LIB"
  echo "$NEWEST" > "$STAMP"
  echo "build_ref: built $SO"
fi
# the unmodified Python side of the reference's hot path, for the GPU box
chmod -R u+w "$OUT/refpkg" 2>/dev/null || true
cp "$REF/models/PWCNet.py" "$OUT/refpkg/models/PWCNet.py"
cp "$REF/models/__init__.py" "$OUT/refpkg/models/__init__.py" 2>/dev/null || : > "$OUT/refpkg/models/__init__.py"
cp "$PKG/correlation.py" "$OUT/refpkg/models/correlation_package/correlation.py"
cp "$PKG/__init__.py" "$OUT/refpkg/models/correlation_package/__init__.py" 2>/dev/null || : > "$OUT/refpkg/models/correlation_package/__init__.py"
( cd "$REF" && sha256sum models/PWCNet.py models/correlation_package/correlation.py ) > "$OUT/refpkg/SHA256SUMS"
chmod -R u+w "$OUT/refpkg"
echo "build_ref: reference python copied to $OUT/refpkg (sha256 recorded)"

#!/usr/bin/env bash
# TEST INFRASTRUCTURE — builds the REFERENCE's own backend sources (read where they lie under
# /root/reference) into oracle/_ref/.  Nothing from the reference is copied into the repo:
# the two one-token torch-2.11 API patches (SURVEY.md finding 4) are applied by `sed` into a
# scratch directory under /tmp, and only the built .so files land in oracle/_ref/ (gitignored,
# NOT gpurun-ignored, so they travel to the GPU box).
#
#   oracle/_ref/ref_backends.so        the reference CUDA extension, -gencode compute_100/sm_100
#                                      (GPU oracle; python module name `ref_backends`)
#   oracle/_ref/libref_matching_host.so  the reference's iter_proj / refine_matches KERNEL BODIES
#                                      (matching_kernels.cu:17-81,119-275) compiled as host C++
#                                      through oracle/host_shim.h (CPU "reference" baseline)
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${M3R_REFERENCE_ROOT:-/root/reference/MASt3R-SLAM}"
OUT="$HERE/_ref"
WHAT="${1:-all}"
if [ ! -d "$REF/mast3r_slam/backend/src" ]; then
  echo "[build_ref] reference not present at $REF — keeping prebuilt files in $OUT" >&2
  exit 0
fi
mkdir -p "$OUT"
TMP="$(mktemp -d /tmp/m3r_ref_build.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
SRC="$REF/mast3r_slam/backend/src"
PY=python
TORCH_INC=$($PY - <<'PY'
import torch.utils.cpp_extension as c, sysconfig
print(" ".join("-I"+p for p in c.include_paths("cuda")+[sysconfig.get_paths()["include"]]))
PY
)
TORCH_LIB=$($PY -c "import torch,os;print(os.path.join(os.path.dirname(torch.__file__),'lib'))")

if [ "$WHAT" = "all" ] || [ "$WHAT" = "gpu" ]; then
  if [ ! -f "$OUT/ref_backends.so" ] || [ "${M3R_FORCE:-0}" = "1" ]; then
    sed 's/D11\.type()/D11.scalar_type()/' "$SRC/matching_kernels.cu" > "$TMP/matching_kernels.cu"
    sed 's/torch::linalg::linalg_norm(dx, std::optional<c10::Scalar>(), {}, false, {})/at::linalg_norm(dx)/' \
        "$SRC/gn_kernels.cu" > "$TMP/gn_kernels.cu"
    NVFLAGS="-O3 -std=c++17 -gencode=arch=compute_100,code=sm_100 -D__CUDA_NO_HALF_OPERATORS__ \
      -D__CUDA_NO_HALF_CONVERSIONS__ -D__CUDA_NO_BFLOAT16_CONVERSIONS__ -D__CUDA_NO_HALF2_OPERATORS__ \
      --expt-relaxed-constexpr -Xcompiler -fPIC -DTORCH_EXTENSION_NAME=ref_backends \
      -DTORCH_API_INCLUDE_EXTENSION_H $TORCH_INC -I$REF/mast3r_slam/backend/include -I$REF/thirdparty/eigen"
    ( nvcc $NVFLAGS -c "$TMP/matching_kernels.cu" -o "$TMP/matching_kernels.o" ) &
    ( nvcc $NVFLAGS -c "$TMP/gn_kernels.cu" -o "$TMP/gn_kernels.o" ) &
    g++ -O3 -std=c++17 -fPIC -DTORCH_EXTENSION_NAME=ref_backends -DTORCH_API_INCLUDE_EXTENSION_H $TORCH_INC \
      -I/usr/local/cuda/include -I$REF/mast3r_slam/backend/include -c "$SRC/gn.cpp" -o "$TMP/gn.o"
    wait
    g++ -shared -o "$OUT/ref_backends.so" "$TMP/gn.o" "$TMP/gn_kernels.o" "$TMP/matching_kernels.o" \
      -L"$TORCH_LIB" -L/usr/local/cuda/lib64 -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -ltorch_python -lcudart \
      -Wl,-rpath,"$TORCH_LIB"
    echo "[build_ref] built $OUT/ref_backends.so"
  fi
fi

if [ "$WHAT" = "all" ] || [ "$WHAT" = "host" ]; then
  if [ ! -f "$OUT/libref_matching_host.so" ] || [ "${M3R_FORCE:-0}" = "1" ]; then
    # kernel bodies only (no launchers: `<<<>>>` is not host C++); line ranges per SURVEY §8a
    { sed -n '17,81p' "$SRC/matching_kernels.cu"; sed -n '119,275p' "$SRC/matching_kernels.cu"; } > "$TMP/ref_matching_kernels.inc"
    g++ -O3 -std=c++17 -fPIC -fopenmp -shared -DM3R_REF_KERNELS_INC="\"$TMP/ref_matching_kernels.inc\"" \
      $TORCH_INC -I/usr/local/cuda/include "$HERE/ref_matching_host.cpp" -o "$OUT/libref_matching_host.so" \
      -L"$TORCH_LIB" -ltorch -ltorch_cpu -lc10 -Wl,-rpath,"$TORCH_LIB"
    echo "[build_ref] built $OUT/libref_matching_host.so"
  fi
fi

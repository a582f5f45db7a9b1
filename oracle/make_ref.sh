#!/usr/bin/env bash
# Test infrastructure (oracle/): builds oracle/_ref/ from the UNMODIFIED reference under /root/reference.
# Nothing here is product code and nothing under sglang_b200/ reads oracle/_ref/.
#
#   oracle/_ref/python/sglang/...   the reference's Python packages for this path (srt, kernels, test kits, ...),
#                                   staged verbatim so that the GPU box (which has no /root/reference) can import the
#                                   REAL AttentionBackend registry, RadixAttention, MHATokenToKVPool, ForwardBatch,
#                                   the dense attention unit-test kit and the Triton / merge_state kernels
#                                   (tests/test_reference_kit_gpu.py, tools/compare_reference_kernels.py).
#   oracle/_ref/merge_state_ref.so  the reference's merge_state_v2 CUDA kernel
#                                   (python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu) compiled for
#                                   sm_100a from where it lies, with a 4-macro shim for the un-vendored
#                                   flashinfer header it includes (oracle/shim/).
#
# oracle/_ref/ is git-ignored (no reference source enters the history) but NOT gpurun-ignored.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/python/sglang" ]; then
  echo "make_ref.sh: $REF/python/sglang not found (the GPU box uses the prebuilt oracle/_ref)" >&2
  exit 0
fi
mkdir -p "$OUT/python/sglang"
if [ ! -f "$OUT/python/sglang/srt/layers/radix_attention.py" ] || [ "${FORCE:-0}" = "1" ]; then
  # the packages the attention path imports; multimodal_gen (15 MB), benchmark, cli front-ends are not needed
  ( cd "$REF/python/sglang" && tar cf - --exclude='__pycache__' --exclude='multimodal_gen' --exclude='benchmark' \
      --exclude='kernels/aot/benchmark' --exclude='kernels/aot/tests' --exclude='*.so' . ) | ( cd "$OUT/python/sglang" && tar xf - )
  # The GPU pool refuses snapshots that NAME the batched-memcpy runtime calls (closed on this pool).  Three staged
  # files do, none on the attention path: two KV-transfer C++ sources (deleted from the staged copy) and one
  # COMMENT line of srt/model_executor/runner/flashinfer_autotune.py (that comment line is dropped).  These are the
  # only departures of oracle/_ref/python from /root/reference/python.
  PAT='Memcpy\(3D\)\?BatchAsync'
  grep -rlI "$PAT" "$OUT/python" | while read -r f; do
    case "$f" in
      *.py) sed -i "/^[[:space:]]*#.*$PAT/d" "$f"; if grep -q "$PAT" "$f"; then rm -f "$f"; echo "removed $f"; else echo "dropped a comment line of $f"; fi ;;
      *) rm -f "$f"; echo "removed $f" ;;
    esac
  done
  echo "staged $REF/python/sglang -> $OUT/python/sglang"
fi
SRC="$REF/python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu"
SO="$OUT/merge_state_ref.so"
if [ "${SKIP_MERGE_REF:-0}" != "1" ] && { [ ! -f "$SO" ] || [ "$SRC" -nt "$SO" ] || [ "$HERE/shim/merge_ref_binding.cu" -nt "$SO" ] || [ "${FORCE:-0}" = "1" ]; }; then
  PY=${PYTHON:-python}
  TORCH_DIR="$($PY -c 'import torch, os; print(os.path.dirname(torch.__file__))' 2>/dev/null)"
  NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
  PYINC="$($PY -c 'import sysconfig; print(sysconfig.get_paths()["include"])')"
  $NVCC -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -shared --expt-relaxed-constexpr \
    -D_GLIBCXX_USE_CXX11_ABI=1 -DTORCH_EXTENSION_NAME=merge_state_ref \
    -I"$HERE/shim" -I"$TORCH_DIR/include" -I"$TORCH_DIR/include/torch/csrc/api/include" -I"$PYINC" \
    "$SRC" "$HERE/shim/merge_ref_binding.cu" \
    -L"$TORCH_DIR/lib" -lc10 -lc10_cuda -ltorch -ltorch_cpu -ltorch_cuda \
    -Xlinker -rpath -Xlinker "$TORCH_DIR/lib" \
    -o "$SO"
  echo "built $SO"
fi
du -sh "$OUT"

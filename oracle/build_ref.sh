#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY. Compiles the reference's own three hot-path .cu files, unchanged and
# where they lie under /root/reference, plus oracle/ref_shim.cpp, into oracle/_ref/libref_reward.so
# (git-ignored, NOT gpurun-ignored, so it travels to the GPU box). Needs ~4 min per .cu file.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REF:-/root/reference}"
OUT="$HERE/_ref"
[ -d "$REF/src" ] || { echo "no reference at $REF; keeping prebuilt $OUT" >&2; exit 0; }
mkdir -p "$OUT"
# the sidecar writer/reader: one plain C++ file, seconds (the stdexcept include is the only thing it lacks on this compiler)
if [ ! -f "$OUT/libref_metadata.so" ] || [ "$HERE/ref_metadata_shim.cpp" -nt "$OUT/libref_metadata.so" ]; then
  g++ -std=c++17 -O2 -fPIC -shared -include stdexcept -I"$REF/include" "$REF/src/metadata.cpp" "$HERE/ref_metadata_shim.cpp" \
    -o "$OUT/libref_metadata.so"
  echo "built $OUT/libref_metadata.so"
fi
TORCH="$(python -c 'import torch,os;print(os.path.dirname(torch.__file__))')"
INC="-I$REF/include -I$TORCH/include -I$TORCH/include/torch/csrc/api/include"
NVF="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -D_GLIBCXX_USE_CXX11_ABI=1 --expt-relaxed-constexpr -Xcompiler -fPIC"
pids=()
for f in segment_stats png_size_estimator compute_rewards; do
  if [ ! -f "$OUT/$f.o" ] || [ "$REF/src/$f.cu" -nt "$OUT/$f.o" ]; then
    nvcc $NVF $INC -c "$REF/src/$f.cu" -o "$OUT/$f.o" & pids+=($!)
  fi
done
if [ ! -f "$OUT/ref_shim.o" ] || [ "$HERE/ref_shim.cpp" -nt "$OUT/ref_shim.o" ]; then
  g++ -std=c++17 -O2 -fPIC -D_GLIBCXX_USE_CXX11_ABI=1 $INC -I/usr/local/cuda/include -c "$HERE/ref_shim.cpp" -o "$OUT/ref_shim.o" & pids+=($!)
fi
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
g++ -shared -o "$OUT/libref_reward.so" "$OUT"/segment_stats.o "$OUT"/png_size_estimator.o "$OUT"/compute_rewards.o "$OUT"/ref_shim.o \
  -L"$TORCH/lib" -ltorch -ltorch_cpu -ltorch_cuda -lc10 -lc10_cuda -L/usr/local/cuda/lib64 -lcudart \
  -Wl,-rpath,"$TORCH/lib"
rm -f "$OUT"/*.o.tmp
echo "built $OUT/libref_reward.so"

#!/bin/bash
# Builds experimental variants of libicr_b200.so (one translation unit compiled with different -D flags) into
# tools/_variants/ for A/B timing on the GPU box: ICR_LIB=tools/_variants/libicr_<name>.so python bench.py ...
# usage: tools/build_variants.sh unit:name:"-DX=1 -DY=2" ...      (unit = narrow | estimate | prep | big_segment | ... | all)
set -e
cd "$(dirname "$0")/../image-compression-cpp_b200/csrc"
make -s -j8 >/dev/null
mkdir -p ../../tools/_variants
NVF="-std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -Xptxas -v"
for spec in "$@"; do
  unit="${spec%%:*}"; rest="${spec#*:}"; name="${rest%%:*}"; flags="${rest#*:}"
  objs=""
  if [ "$unit" = "all" ]; then          # every translation unit with the flags (e.g. all:checked:-DICR_CHECKED)
    for u in api prep narrow estimate big_segment misc slices exchange; do
      nvcc $NVF $flags -c $u.cu -o ../../tools/_variants/${u}_$name.o 2> ../../tools/_variants/${u}_$name.log &
      objs="$objs ../../tools/_variants/${u}_$name.o"
    done
    wait
    unit=estimate
  else
    nvcc $NVF $flags -c $unit.cu -o ../../tools/_variants/${unit}_$name.o 2> ../../tools/_variants/${unit}_$name.log
    for u in api prep narrow estimate big_segment misc slices exchange; do
      if [ "$u" = "$unit" ]; then objs="$objs ../../tools/_variants/${unit}_$name.o"; else objs="$objs $u.o"; fi
    done
  fi
  nvcc -shared -cudart shared -o ../../tools/_variants/libicr_$name.so $objs -ldl 2>/dev/null
  echo "$name: $(grep -o 'Used [0-9]* registers' ../../tools/_variants/${unit}_$name.log | sort | uniq -c | tr '\n' ' ')"
done

#!/bin/bash
# Tuning variants of the CUDA library (atooms_b200/lib/variant_<name>.so) for tools/variants.py:
#   tools/build_variants.sh name1:"-DA=1 -DB=2" name2:"-DC=3" ...
set -e
cd "$(dirname "$0")/.."
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -shared --expt-relaxed-constexpr"
mkdir -p atooms_b200/lib
rm -f atooms_b200/lib/variant_*.so
for spec in "$@"; do
    name="${spec%%:*}"; defs="${spec#*:}"
    ( /usr/local/cuda/bin/nvcc $FLAGS $defs -o atooms_b200/lib/variant_$name.so atooms_b200/csrc/b2_engine.cu -ldl -Xlinker -Bsymbolic && echo built $name ) &
done
wait

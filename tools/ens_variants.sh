#!/bin/bash
# Developer probe (GPU box): throughput of the ensemble kernel for a list of -D variants.
# usage: tools/ens_variants.sh "<defs 1>" "<defs 2>" ...   ("" = the default build)
for v in "$@"; do
  SLAMCUDA_DEFS="$v" python slam.jl_b200/build.py -f >/dev/null 2>&1 || { echo "build failed: $v"; continue; }
  regs=$(cuobjdump --dump-resource-usage slam.jl_b200/lib/libslamcuda.so 2>/dev/null | grep -A1 "k_ensembleILi128" | grep -o "REG:[0-9]*")
  for f in 5 1; do
    echo "[$v] $regs first=$f $(python tools/ens_probe.py 8192 3 35 128 5 $f 2>&1 | tail -1 | cut -c1-110)"
  done
done

#!/bin/bash
# A/B timing of library variants built by tools/build_variant.sh: tools/gpu_ab.sh name1 name2 ...  ("default" = the in-tree library)
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O; : > $O/ab.txt
for v in "$@"; do
  if [ "$v" = default ]; then timeout 300 python tools/quick_bench.py 65536 20 >> $O/ab.txt 2>&1
  elif [ "$v" = t2 ]; then B2C_P108=t2 timeout 300 python tools/quick_bench.py 65536 20 >> $O/ab.txt 2>&1
  else B2C_LIB_PATH=build/variants/$v.so timeout 300 python tools/quick_bench.py 65536 20 >> $O/ab.txt 2>&1; fi
done
cat $O/ab.txt

#!/bin/bash
# build kernel variants (-D flags) and time the enumeration with each; usage: tools/variants.sh "flags1" "flags2" ...
cd "$(dirname "$0")/.."
i=0
for f in "$@"; do
  i=$((i+1))
  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared -Iinclude $f \
       -o /tmp/librca_v$i.so k8s-llm-rca_b200/csrc/rca_b200.cu -Xptxas -v 2>&1 | grep -E "Compiling|Used|spill" | grep -A1 "ILb1ELb1" | tail -2 | tr '\n' ' '
  echo
  echo "variant $i [$f]: $(RCA_B200_LIB=/tmp/librca_v$i.so timeout 300 python tools/profile_varlen.py 1000000 3 2>&1 | tail -1 | sed 's/.*"level_ms"/level_ms/; s/"level_bytes.*//')"
done

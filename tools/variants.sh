#!/bin/bash
# build kernel variants (-D flags) and time a driver with each; usage: tools/variants.sh "flags1" "flags2" ...
# DRIVER (env) defaults to tools/with_states_cost.py
cd "$(dirname "$0")/.."
DRIVER=${DRIVER:-tools/with_states_cost.py}
i=0
for f in "$@"; do
  i=$((i+1))
  nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared -Iinclude $f \
       -o /tmp/librca_v$i.so k8s-llm-rca_b200/csrc/rca_b200.cu 2>&1 | grep -E "error" | head -3
  echo "variant $i [$f]:"
  RCA_B200_LIB=/tmp/librca_v$i.so timeout 300 python $DRIVER 2>&1 | tail -3 | sed 's/^/    /'
done

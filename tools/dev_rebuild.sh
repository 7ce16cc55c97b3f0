#!/bin/bash
# Developer shortcut: recompile only the listed graph-kernel variants (e.g. "22:20 20:20") plus the two API units,
# relink with the other (cached) objects.  build.py's build_library(force=True) is the real build.
set -e
cd "$(dirname "$0")/../cgnn_b200"
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC"
pids=()
for v in "$@"; do
  dp=${v%%:*}; hp=${v##*:}
  nvcc $FLAGS -DCGNN_DP=$dp -DCGNN_HP=$hp -DCGNN_REGEN=0 $EXTRA -c csrc/graph_inst.cu -o build/graph_dp${dp}_hp${hp}.o & pids+=($!)
  nvcc $FLAGS -DCGNN_DP=$dp -DCGNN_HP=$hp -DCGNN_REGEN=1 $EXTRA -c csrc/graph_inst.cu -o build/graph_dp${dp}_hp${hp}_regen.o & pids+=($!)
done
nvcc $FLAGS -c csrc/capi.cu -o build/capi.o & pids+=($!)
if [ -n "$PW" ]; then for v in "512 2" "256 4" "128 8"; do set -- $v; nvcc $FLAGS -DCGNN_PW_NT=$1 -DCGNN_PW_CTAS=$2 -DCGNN_PW_NS=pw$1 -DCGNN_PW_TAG=$1 -c csrc/pairwise_inst.cu -o build/pairwise_$1.o & pids+=($!); done; fi
nvcc $FLAGS -c csrc/graph_api.cu -o build/graph_api.o & pids+=($!)
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -o libcgnn_b200.so build/*.o
echo relinked

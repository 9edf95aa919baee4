#!/bin/bash
# Development aid: build tools/libzkbatch_<name>.so with extra -D flags.
#   tools/build_variant.sh prof -DZK_PROFILE_PHASES
set -e
name=$1; shift
cd "$(dirname "$0")/.."
C=node-zkstream_b200/csrc
F="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden --cudart shared"
mkdir -p /tmp/zkv_$name
for s in zk_ctx zk_decode zk_encode zk_stats; do
  /usr/local/cuda/bin/nvcc $F "$@" -c $C/$s.cu -o /tmp/zkv_$name/$s.o &
done
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared --cudart shared -o tools/libzkbatch_$name.so /tmp/zkv_$name/*.o
echo tools/libzkbatch_$name.so

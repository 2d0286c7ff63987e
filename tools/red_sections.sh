#!/bin/bash
# Profiling aid: time reduce_tc_kernel with one of its sections compiled out (HEXGNN_RED_SKIP bits), on the GPU box.
cd "$(dirname "$0")/.."
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --compiler-options -fPIC"
cp gnn_hex_b200/libhexgnn.so /tmp/libhexgnn.keep
for k in 0 1 2 4 8 6 14; do
  nvcc $FLAGS -DHEXGNN_RED_SKIP=$k -c gnn_hex_b200/csrc/pack_tc.cu -o build/pack_tc.o || exit 1
  nvcc -shared -o gnn_hex_b200/libhexgnn.so build/*.o -gencode arch=compute_100a,code=sm_100a || exit 1
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:reduce_tc -c 12 --csv --log-file /tmp/red_$k.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  echo "skip=$k $(python tools/launch_summary.py /tmp/red_$k.csv | head -1)"
done
cp /tmp/libhexgnn.keep gnn_hex_b200/libhexgnn.so

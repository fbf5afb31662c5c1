#!/bin/bash
# Ablation variants of csrc/qgemm_stream.cu (tuning evidence only): one .so per -D flag under profiles/variants/.
#   profiles/build_variants.sh ABL_NO_MMA ABL_NO_X "ABL_NO_MATH -DQS_ABL_NO_STTM" UNIT_TRACE
# Run with B200_PAGED_OPS_LIB=profiles/variants/lib_<name>.so python profiles/bench_qstream.py ...
set -e
cd "$(dirname "$0")/.."
python -m vllm_metal_b200.metal.build >/dev/null
mkdir -p profiles/variants
OBJS=$(ls vllm_metal_b200/build/*.o | grep -v qgemm_stream.o)
for v in "$@"; do
  name=$(echo "$v" | tr -d ' ' | sed 's/-DQS_ABL_/_/g; s/-DQS_/_/g')
  ( nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DQS_$v \
      -c -o profiles/variants/qs_$name.o vllm_metal_b200/csrc/qgemm_stream.cu &&
    nvcc --shared -o profiles/variants/lib_$name.so $OBJS profiles/variants/qs_$name.o && echo built $name ) &
done
wait

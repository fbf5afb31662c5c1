#!/bin/bash
# usage: profiles/ablate_qstream.sh [bench_qstream args]; runs the shipped library then every profiles/variants/lib_*.so
cd "$(dirname "$0")/.."
unset B200_PAGED_OPS_LIB
echo "== base"; timeout 120 python profiles/bench_qstream.py "$@" 2>&1 | grep -v "^#"
for so in profiles/variants/lib_*.so; do
  export B200_PAGED_OPS_LIB=$PWD/$so
  echo "== $(basename $so .so | sed s/lib_//)"; timeout 120 python profiles/bench_qstream.py "$@" 2>&1 | grep -v "^#"
done

#!/bin/bash
# developer A/B run on the GPU box: tools/ab.sh <tag> <variant> [<variant> ...]   (variants = directories under cvxcompress_b200/lib_exp)
tag=$1; shift
O=gpurun_out; mkdir -p $O
for v in "$@"; do
  lib=cvxcompress_b200/lib_exp/$v/libcvxcompress.so
  CVXCOMPRESS_LIB=$PWD/$lib timeout 100 python tools/kbench.py 960 lifting 10 >> $O/${tag}_kbench.jsonl 2>> $O/${tag}_kbench.err
done
cat $O/${tag}_kbench.jsonl

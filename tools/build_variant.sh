#!/bin/bash
# developer tool: tools/build_variant.sh <name> [TIMING=1] [EXTRA="-D..."]  -- builds a snapshot of csrc/ into cvxcompress_b200/lib_exp/<name>
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
snap=/tmp/bv/$name
rm -rf $snap; mkdir -p $snap/cvxcompress_b200 $snap/include
cp -r $root/cvxcompress_b200/csrc $snap/cvxcompress_b200/csrc
cp $root/include/*.h $root/include/*.hxx $snap/include/
make -s -C $snap/cvxcompress_b200/csrc OUT=$root/cvxcompress_b200/lib_exp/$name "$@" > /tmp/b_$name.log 2>&1
echo "$name: rc=$? $(grep -c 'error' /tmp/b_$name.log) errors"

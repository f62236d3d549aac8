#!/bin/bash
# builds kernel-tuning variants of libhenjou_b200.so into build/variants/<name>.so :  tools/build_variants.sh name "-DFLAG=.." ...
set -e
cd "$(dirname "$0")/../henjou-renderer_b200/csrc"
mkdir -p ../../build/variants
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  rm -f hj_wavefront.o hj_bvh.o hj_api.o
  make -s -j4 EXTRA="$flags" OUT=../../build/variants/$name.so
  echo "built $name ($flags)"
done
rm -f hj_wavefront.o hj_bvh.o hj_api.o
make -s -j4

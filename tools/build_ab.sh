#!/bin/bash
# A/B builds of libpbe_b200.so: tools/build_ab.sh NAME "EXTRA_NVCC_FLAGS" [NAME2 "FLAGS2" ...]
# -> openpbe_b200/lib/ab/libNAME.so (travels to the GPU box; tools/*.py pick it up through PBE_LIB)
set -e
cd "$(dirname "$0")/../openpbe_b200/csrc"
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  ( make -j4 -s OUT=../lib/ab/lib$name.so OBJDIR=../lib/ab/obj$name EXTRA="$flags" 2>&1 | tail -3 ) &
done
wait

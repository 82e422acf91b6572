#!/bin/bash
# build library variants that differ in the macros given to eam.cu (timing experiments):
#   bash profiles/eam_variants_build.sh name1 "-DEAM_ABLATE=1" name2 "-DEAM_BLOCK=256" ...
# NOTE (round 2): the EAM_ABLATE branches this script used were removed from csrc/eam.cu once the experiments were
# recorded (profiles/r01g_eam_variants.txt); they are in the history up to commit 902509a.  EAM_BLOCK1/2, EAM_UNROLL,
# EAM_IDX_CS, EAM_GATHER_NC are still compile-time knobs.
# -> imd_b200/variants/libimdcu_<name>.so  (run on the GPU box by profiles/eam_variants_run.sh)
set -e
cd "$(dirname "$0")/../imd_b200/csrc"
make -j8 > /dev/null
mkdir -p ../variants
SRC=${SRC:-eam.cu}            # SRC=ttm.cu EXTRA=-fmad=false for the lattice kernel
BASE=${SRC%.cu}
FLAGS="$EXTRA -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-ffp-contract=off,-Wall"
while [ $# -ge 2 ]; do
  name=$1; defs=$2; shift 2
  nvcc $FLAGS $defs -c $SRC -o /tmp/${BASE}_$name.o
  objs=$(ls *.o | grep -v "^$BASE.o\$")
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libimdcu_$name.so $objs /tmp/${BASE}_$name.o -lcudart -ldl
  echo built $name "($defs)"
done

#!/bin/bash
# Build libpdgs_<name>.so from the CUDA sources of an earlier commit (A/B against HEAD with tools/ab.sh or PDGS_LIB).
# usage: tools/build_commit.sh <commit> <name>     (sources are unpacked under gpurun_out/src_<name>, which does not travel)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
commit=$1; name=$2
dst=$ROOT/gpurun_out/src_$name
rm -rf $dst && mkdir -p $dst
git -C $ROOT archive $commit priordepth-gaussiansplatting_b200/csrc include | tar -x -C $dst
objs=""
for f in $dst/priordepth-gaussiansplatting_b200/csrc/*.cu; do
  o=${f%.cu}.o
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
    --expt-relaxed-constexpr -c $f -o $o &
  objs="$objs $o"
done
wait
/usr/local/cuda/bin/nvcc -shared -o $ROOT/priordepth-gaussiansplatting_b200/lib/libpdgs_$name.so $objs -gencode arch=compute_100a,code=sm_100a -lcudart
echo built libpdgs_$name.so from $commit

#!/bin/bash
# Build a variant of libpdgs.so with extra -D flags for ONE translation unit (A/B runs of kernel options).
# usage: tools/build_variant.sh <name> <source.cu> [-DFOO=1 ...]   ->  priordepth-gaussiansplatting_b200/lib/libpdgs_<name>.so
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
PKG=$ROOT/priordepth-gaussiansplatting_b200
name=$1; src=$2; shift 2
python $PKG/build.py > /dev/null
base=${src%.cu}
mkdir -p $PKG/lib/var
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -ffp-contract=off \
  --expt-relaxed-constexpr "$@" -Xptxas -v -c $PKG/csrc/$src -o $PKG/lib/var/${base}_$name.o 2>&1 | grep -A2 "Compiling.*${KERNEL:-render_}" | grep -i "registers\|spill" || true
objs=""
for o in pdgs_preprocess pdgs_binning pdgs_render pdgs_loss pdgs_ssim pdgs_optim pdgs_capi; do
  if [ "$o" == "$base" ]; then objs="$objs $PKG/lib/var/${base}_$name.o"; else objs="$objs $PKG/lib/$o.o"; fi
done
/usr/local/cuda/bin/nvcc -shared -o $PKG/lib/libpdgs_$name.so $objs -gencode arch=compute_100a,code=sm_100a -lcudart
echo built $PKG/lib/libpdgs_$name.so

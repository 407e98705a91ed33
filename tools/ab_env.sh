#!/bin/bash
# usage: tools/ab_env.sh <lib variant> "ENV=val ENV2=val" ...  : one bench run per env setting
v=$1; shift
export PDGS_LIB=$PWD/priordepth-gaussiansplatting_b200/lib/libpdgs_$v.so
for e in "$@"; do
  tag=$(echo "$e" | tr ' =' '__')
  env $e timeout 300 python bench.py --steps 2 --warmup 2 --views 32 --no-cpu-baseline --no-e2e > gpurun_out/abe_$tag.json 2> gpurun_out/abe_$tag.err || { echo "$e FAILED"; tail -3 gpurun_out/abe_$tag.err; continue; }
  echo "== $e"; python profiles/show_stages.py gpurun_out/abe_$tag.json 2>/dev/null | sed -n "2,3p"
done

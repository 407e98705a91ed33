#!/bin/bash
# A/B: run the C4 bench (resident inputs only) once per library variant; prints the stage table of each.
# usage (on the GPU box): tools/ab.sh name1 name2 ...   (names as given to tools/build_variant.sh; "base" = libpdgs.so)
for v in "$@"; do
  if [ "$v" == "base" ]; then unset PDGS_LIB; else export PDGS_LIB=$PWD/priordepth-gaussiansplatting_b200/lib/libpdgs_$v.so; fi
  timeout 300 python bench.py --steps 2 --warmup 2 --views 32 --no-cpu-baseline --no-e2e > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err || { echo "$v FAILED"; tail -5 gpurun_out/ab_$v.err; continue; }
  echo "== $v"; python profiles/show_stages.py gpurun_out/ab_$v.json 2>/dev/null | sed -n "1,${AB_LINES:-3}p"
  python -c "import json;print('loss', json.loads(open('gpurun_out/ab_$v.json').read().strip().splitlines()[-1])['loss_mean'])"
done

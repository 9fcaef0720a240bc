#!/bin/bash
# A/B of environment knobs on the cfg2 bench line: tools/ab_bench.sh "VAR=a VAR2=b" "VAR=c" ...
OUT=gpurun_out/${TAG:-ab}; mkdir -p $OUT
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-v1 > $OUT/bench_$i.json 2> $OUT/bench_$i.err
  python - "$envs" $OUT/bench_$i.json <<'PY'
import json,sys
try:
    b=json.load(open(sys.argv[2]))
    print(f"{sys.argv[1]:50s} value {b['value']:.0f} ms {b['ms_per_step']:.2f} e2e {b['e2e']['value']:.0f} long {b['long_form']['value']:.0f} clk {b['clocks']['sm_mhz']}")
except Exception as e:
    print(sys.argv[1], "FAILED", e); print(open(sys.argv[2].replace('.json','.err')).read()[-800:])
PY
done

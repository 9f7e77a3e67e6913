#!/bin/bash
# One bench.py run condensed to a line: ms per step, value, e2e, SM MHz, max W, oracle check, roofline fraction.
# usage: [WL=workload] tools/bench_ab.sh LABEL [ENV=VAL ...]
label=$1; shift
env "$@" timeout 200 python bench.py --workload ${WL:-cfg3_b64} --steps 10 --warmup 3 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', round(d['ms_per_step'],3), round(d['value'],1), round(d['e2e']['value'],1), d['clocks']['sm_mhz'], d['clocks']['power_w_max'], d['result_check']['matches_oracle'], round(d['roofline']['frac'],3))"

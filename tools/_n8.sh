# The 8-GPU validation of the round (run from the repo root on an 8-GPU box; outputs in gpurun_out/n8): bench at N = 8 / 4 / 2,
# the int8, b = 64 and configs[3] workloads at N = 8, the CPU arm, the multi-GPU tests and the one-process 8-GPU bench.
set -x
O=gpurun_out/n8; mkdir -p $O
run() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $n "$@" 2> $O/err_$n.txt | tail -1; }
run 8 --steps 100 --warmup 5 > $O/r2_bench_n8.json
run 4 --steps 100 --warmup 5 > $O/r2_bench_n4.json
run 2 --steps 50 --warmup 5 > $O/r2_bench_n2.json
run 8 --steps 100 --warmup 5 --workload cfg3_b1_q8 > $O/r2_bench_cfg3_b1_q8_n8.json
run 8 --steps 30 --warmup 3 --workload cfg3_b64 > $O/r2_bench_cfg3_b64_n8.json
run 8 --steps 5 --warmup 1 --workload cfg4 > $O/r2_bench_cfg4_n8.json
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/ref_arm_8gpu_box.json 2>/dev/null
timeout 600 python -m pytest tests/test_gpu_multi_device.py tests/test_multi_rank.py -m gpu -x -q 2>&1 | tail -3 > $O/multi_tests.txt
timeout 300 python tools/inproc_bench.py --devices 8 > $O/inproc_n8.json 2> $O/err_inproc.txt
for f in $O/*.json; do echo $f; python - $f <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print({k: d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('e2e',{}).get('value'), d.get('result_check',{}).get('matches_oracle'))
except Exception as e: print('ERR', e)
PY
done
cat $O/multi_tests.txt

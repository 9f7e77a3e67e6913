# Short 8-GPU check of the final library: N = 8 and N = 4 bench lines and the multi-GPU tests (outputs in gpurun_out/n8b).
set -x
O=gpurun_out/n8b; mkdir -p $O
run() { n=$1; shift; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $n "$@" 2> $O/err_$n.txt | tail -1; }
run 8 --steps 100 --warmup 5 > $O/r2_bench_n8.json
run 4 --steps 100 --warmup 5 > $O/r2_bench_n4.json
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > $O/r2_bench_n1_same_box.json 2>/dev/null
timeout 600 python -m pytest tests/test_gpu_multi_device.py tests/test_multi_rank.py -m gpu -x -q 2>&1 | tail -3 > $O/multi_tests.txt
for f in $O/*.json; do python - $f <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], {k: d.get(k) for k in ('value','ms_per_step','n_gpus')}, d.get('e2e',{}).get('value'), d.get('result_check',{}).get('matches_oracle'))
PY
done
cat $O/multi_tests.txt

# How the round-2 records under profiles/ were regenerated on one GPU (run from the repo root on the GPU box; outputs in gpurun_out/regen).
set -x
O=gpurun_out/regen; mkdir -p $O
for w in cfg1 cfg2 cfg3_b3 cfg3_b8 cfg3_b64 cfg3_b1_q8 cfg3_b64_q8; do timeout 300 python bench.py --workload $w > $O/r2_bench_$w.json 2> $O/err_$w.txt; done
build/topk_latency 200 1 5 2000 16 > $O/r2_topk_latency_cfg0_cpp.json
build/topk_latency 1000000 16 10 2000 16 > $O/r2_topk_latency_cfg1_cpp.json
timeout 600 python tools/ivf_bench.py --rows 10000000 > $O/r2_ivf_10m_lists4096.json 2> $O/err_ivf.txt
timeout 300 python tools/ivf_probe_timing.py > $O/ivf_probe_timing.txt 2>&1
timeout 300 python bench.py --workload cfg3_b64 --steps 2 --warmup 1 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:tc_rows_candidates -c 1 -o $O/r2_tc_rows_b64 timeout 600 python bench.py --workload cfg3_b64 --steps 2 --warmup 1 > $O/ncu_b64.log 2>&1
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/r2_bench_n1_steps3.json 2>/dev/null && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2_launches_bench_steps3.csv timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
ls -la $O

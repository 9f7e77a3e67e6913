# Final validation of the round on one GPU (from the repo root on the GPU box; outputs in gpurun_out/final).
set -x
O=gpurun_out/final; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 > $O/gpu_tests.txt; cat $O/gpu_tests.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.txt 2>&1; tail -2 $O/smoke.txt
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; tail -c 600 $O/bench_default.json
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > $O/r2_bench_reference_arm.json 2> $O/ref.err; tail -c 800 $O/r2_bench_reference_arm.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:scan_topk_kernel -s 4 -c 2 -o $O/r2_scan_topk timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_scan.log 2>&1
ls -la $O

mkdir -p gpurun_out
timeout 900 python bench.py --trace-clocks gpurun_out/r2_clock_trace_200steps.json > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_p.err; echo "bench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_20steps.json 2>> gpurun_out/r2_p.err
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench_reference.json 2>> gpurun_out/r2_p.err
timeout 900 python tools/bench_distributions.py > gpurun_out/r2_distributions.jsonl 2>> gpurun_out/r2_p.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 1 --no-extra --no-cpu-baseline --clock-period-ms 0 > gpurun_out/r2_ncu_bench.log 2>&1
for e in 1000000 500000 250000 125000; do
timeout 600 ncu --set full --import-source on --clock-control none -k regex:qre_gemm_topk -s 4 -c 1 -o gpurun_out/r2_scan_full_$e -f python bench.py --steps 3 --warmup 2 --no-extra --no-cpu-baseline --oracle-queries 0 --clock-period-ms 0 --entries $e > /dev/null 2>&1
done
ls -la gpurun_out/*.ncu-rep

set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 600 gpurun_out/r2_bench_n1.json; tail -2 gpurun_out/r2_bench_n1.err
python bench.py --k 50 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu > gpurun_out/r2_bench_100M_k50_n1.json 2>&1; tail -c 300 gpurun_out/r2_bench_100M_k50_n1.json
python profiles/ref_gpu_table.py > gpurun_out/r2_ref_gpu_table.md 2>/dev/null; cat gpurun_out/r2_ref_gpu_table.md
python profiles/run_case.py --steps 2 > gpurun_out/plain.log 2>&1 && cat gpurun_out/plain.log &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 45 --csv --log-file gpurun_out/r2_launches_100M_k20.csv python profiles/run_case.py --steps 2 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kn_query_tile -c 1 -o gpurun_out/r2_tile_final python profiles/run_case.py --steps 1 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kn_query_tile -c 1 -o gpurun_out/r2_tile_final_k50 python profiles/run_case.py --steps 1 --k 50 > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log

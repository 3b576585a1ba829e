# Single-GPU evidence run of round 2:  gpurun --timeout 1800 -- 'bash tools/_run.sh'   (results land in gpurun_out/,
# the ones to keep are copied to profiles/ by hand; ncu only wraps commands that have just run clean without it)
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 400 gpurun_out/r2_bench_n1.json; tail -2 gpurun_out/r2_bench_n1.err
python bench.py --k 50 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu > gpurun_out/r2_bench_100M_k50_n1.json 2>&1; tail -c 200 gpurun_out/r2_bench_100M_k50_n1.json
python profiles/ref_gpu_table.py > gpurun_out/r2_ref_gpu_table.md 2>/dev/null; cat gpurun_out/r2_ref_gpu_table.md
python profiles/extra_cases.py > gpurun_out/r2_extra_cases.txt 2>&1; cat gpurun_out/r2_extra_cases.txt
python profiles/run_case.py --steps 2 > gpurun_out/plain.log 2>&1 && cat gpurun_out/plain.log &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 45 --csv --log-file gpurun_out/r2_launches_100M_k20.csv python profiles/run_case.py --steps 2 > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kn_query_tile -c 1 -o gpurun_out/r2_tile_final python profiles/run_case.py --steps 1 > gpurun_out/ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kn_query_select_kernel -s 1 -c 1 -o gpurun_out/r2_second_chance python profiles/run_case.py --steps 1 > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:kn_scatter_kernel -c 1 -o gpurun_out/r2_build_scatter python profiles/run_case.py --steps 1 --no-solve > gpurun_out/ncu4.log 2>&1
python profiles/run_case.py --steps 2 --k 50 > gpurun_out/plain50.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 45 --csv --log-file gpurun_out/r2_launches_100M_k50.csv python profiles/run_case.py --steps 2 --k 50 > gpurun_out/ncu5.log 2>&1

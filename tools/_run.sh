set -x
python -m pytest tests -m gpu -x -q --durations=6 2>&1 | tail -14
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 4500 gpurun_out/r2_bench_n1.json; tail -3 gpurun_out/r2_bench_n1.err
python bench.py --k 50 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-ref-gpu > gpurun_out/r2_bench_100M_k50_n1.json 2>&1; tail -c 1200 gpurun_out/r2_bench_100M_k50_n1.json
( time python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r2_bench_reference_arm.json ) 2> gpurun_out/ref_time.log; cat gpurun_out/r2_bench_reference_arm.json | cut -c1-1500; tail -4 gpurun_out/ref_time.log
python profiles/ref_gpu_table.py > gpurun_out/r2_ref_gpu_table.md 2>/dev/null; cat gpurun_out/r2_ref_gpu_table.md

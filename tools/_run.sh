set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 900 gpurun_out/r2_bench_n1.json; tail -2 gpurun_out/r2_bench_n1.err
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 45 --csv --log-file gpurun_out/r2_launches_100M_k20.csv python profiles/run_case.py --steps 2 > gpurun_out/ncu1.log 2>&1
python profiles/launch_bytes.py gpurun_out/r2_launches_100M_k20.csv 30 | tail -14

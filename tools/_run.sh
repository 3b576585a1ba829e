python -m pytest tests/test_multi_gpu.py tests/test_multi_gpu_nccl.py -x -q 2>&1 | tail -5
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-ref-gpu --no-e2e > gpurun_out/r2_bench_n1_mid.json 2> gpurun_out/err1.log; tail -c 600 gpurun_out/r2_bench_n1_mid.json; tail -3 gpurun_out/err1.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-e2e > gpurun_out/r2_bench_n2_mid.json 2> gpurun_out/err2.log; tail -c 2500 gpurun_out/r2_bench_n2_mid.json; tail -5 gpurun_out/err2.log

run() { N=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$((RANDOM%90+10)) bench.py --gpus $N "$@"; }
run 8 --steps 20 --warmup 3 --no-e2e > gpurun_out/r2_bench_n8b.json 2> gpurun_out/err8.log; tail -c 500 gpurun_out/r2_bench_n8b.json; grep -v "^\*\|OMP_NUM" gpurun_out/err8.log | tail -3

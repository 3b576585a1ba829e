python -m pytest tests/test_parity_gpu.py tests/test_random_differential_gpu.py -x -q 2>&1 | tail -3
for k in 20 8; do python profiles/run_case.py --steps 3 --k $k; done

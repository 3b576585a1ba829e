python -m pytest tests/test_parity_gpu.py tests/test_random_differential_gpu.py -m gpu -x -q 2>&1 | tail -3
echo "k 8: $(python profiles/run_case.py --steps 3 --k 8 2>&1 | tail -1 | cut -c1-200)"

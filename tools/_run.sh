python -m pytest tests/test_parity_gpu.py tests/test_random_differential_gpu.py -m gpu -x -q 2>&1 | tail -3
for k in 20 50 32 8; do
echo "k $k: $(python profiles/run_case.py --steps 3 --k $k 2>&1 | tail -1 | cut -c1-200)"
done

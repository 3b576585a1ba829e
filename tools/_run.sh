python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "50 or uniform" 2>&1 | tail -3
echo "k 50: $(python profiles/run_case.py --steps 3 --k 50 2>&1 | tail -1 | cut -c1-200)"

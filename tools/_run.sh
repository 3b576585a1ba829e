python -m pytest tests/test_parity_gpu.py tests/test_random_differential_gpu.py -m gpu -x -q 2>&1 | tail -3
python profiles/extra_cases.py > gpurun_out/r2_extra_cases.txt 2>&1; cat gpurun_out/r2_extra_cases.txt
echo "k 20: $(python profiles/run_case.py --steps 3 2>&1 | tail -1 | cut -c1-200)"

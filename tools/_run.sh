python tools/_case.py
python -m pytest tests/test_parity_gpu.py tests/test_random_differential_gpu.py tests/test_parity_reference_gpu.py tests/test_multi_gpu.py -x -q 2>&1 | tail -3
python profiles/small_case.py pts20k 50; python profiles/small_case.py u300k 50; python profiles/small_case.py u900k 50

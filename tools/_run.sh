python -m pytest tests/test_multi_gpu.py tests/test_dropin_cpp.py -x -q 2>&1 | tail -8

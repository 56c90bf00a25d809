timeout 600 python -m pytest tests/test_gpu_baseline_sizes.py tests/test_gpu_model.py -q -m gpu --timeout 500 -p no:cacheprovider 2>&1 | tail -2

timeout 900 python -m pytest tests/test_gpu_series_cache.py tests/test_gpu_range_client.py tests/test_gpu_multi_rank.py -x -q 2>&1 | tail -12
python bench_kernels.py 2>&1 | grep -E "series cache|fused align"

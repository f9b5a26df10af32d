timeout 300 python -m pytest tests/test_gpu_conv_tc.py -x -q 2>&1 | tail -4
python scripts/bench_conv.py 1 shuffle | cut -c1-40,60-100
timeout 600 python -m pytest tests/test_gpu_codec.py tests/test_gpu_kernels.py -x -q 2>&1 | tail -3

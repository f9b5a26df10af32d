timeout 300 python -m pytest tests/test_gpu_conv_tc.py -x -q 2>&1 | tail -6
python scripts/bench_conv.py 1 | cut -c1-40,60-100

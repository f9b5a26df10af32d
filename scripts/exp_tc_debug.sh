echo HALO
DCVC_TC_HALO=1 timeout 300 python -m pytest tests/test_gpu_conv_tc.py -x -q 2>&1 | tail -2
DCVC_TC_HALO=1 python scripts/bench_conv.py 1 | grep -E "3x3s1|7x7" | cut -c1-40,60-100
echo HALO ctas1
DCVC_TC_HALO_CTAS=1 DCVC_TC_HALO=1 python scripts/bench_conv.py 1 | grep -E "3x3s1|7x7" | cut -c1-40,60-100

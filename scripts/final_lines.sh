# final bench lines of the round for profiles/ (no profiler attached)
set -x
python bench.py --profile-out gpurun_out/r02_ptable_c.json > gpurun_out/r02_bench_c.log 2>&1; tail -1 gpurun_out/r02_bench_c.log | cut -c1-400
timeout 300 python bench.py --impl reference --steps 2 --warmup 3 > gpurun_out/r02_bench_ref_c.log 2>&1; tail -1 gpurun_out/r02_bench_ref_c.log | cut -c1-200
python bench.py --workload yuv420x8 > gpurun_out/r02_yuv_c.log 2>&1; tail -1 gpurun_out/r02_yuv_c.log | cut -c1-200
python bench.py --workload hem_sweep > gpurun_out/r02_hem_c.log 2>&1; tail -1 gpurun_out/r02_hem_c.log | cut -c1-200
python bench.py --workload entropy > gpurun_out/r02_entropy_c.log 2>&1; tail -1 gpurun_out/r02_entropy_c.log | cut -c1-200

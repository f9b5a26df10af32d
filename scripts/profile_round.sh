set -x
python bench.py --profile-out gpurun_out/r02_ptable.json > gpurun_out/r02_bench.log 2>&1; tail -1 gpurun_out/r02_bench.log | cut -c1-600
timeout 300 python bench.py --impl reference --steps 2 --warmup 3 > gpurun_out/r02_bench_ref.log 2>&1; tail -1 gpurun_out/r02_bench_ref.log | cut -c1-300
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --pipeline 0 --graphs 0 --no-gpu-reference --no-parity --no-extra-arms"
$CMD > gpurun_out/r02_ncu_plain.log 2>&1 && timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 9000 --csv --log-file gpurun_out/r02_launches_full.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1
tail -2 gpurun_out/r02_ncu_list.log | cut -c1-300
export BENCH_F16=io
python scripts/bench_conv.py 1 "conv3x3s1 48->48" > gpurun_out/r02_benchconv_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_tc -s 3 -c 1 -o gpurun_out/r02_conv_tc_3x3_48 python scripts/bench_conv.py 1 "conv3x3s1 48->48" > gpurun_out/r02_ncu_full.log 2>&1
tail -2 gpurun_out/r02_ncu_full.log
ls -la gpurun_out | tail -8

# ncu evidence of the final build: a launch list of the first frames (I + P frames, eager launches,
# sequential schedule) and one --set full capture of the most frequent heavy convolution
set -x
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --pipeline 0 --graphs 0 --no-gpu-reference --no-parity --no-extra-arms"
$CMD > gpurun_out/r02f_ncu_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2900 --csv --log-file gpurun_out/r02f_launches_full.csv $CMD > gpurun_out/r02f_ncu_list.log 2>&1
tail -2 gpurun_out/r02f_ncu_list.log | cut -c1-200
export BENCH_F16=io
python scripts/bench_conv.py 1 "conv3x3s1 48->48" > gpurun_out/r02f_benchconv_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:conv_tc -s 3 -c 1 -o gpurun_out/r02f_conv_tc_3x3_48 python scripts/bench_conv.py 1 "conv3x3s1 48->48" > gpurun_out/r02f_ncu_full.log 2>&1
tail -2 gpurun_out/r02f_ncu_full.log

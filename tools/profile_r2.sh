set -x
CMD="python bench.py --steps 2 --warmup 3 --no-configs --no-orderkey --no-cpu --e2e-rows 8000000"
$CMD > gpurun_out/r2_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:^bgpu_fused$ -s 3 -c 1 -o gpurun_out/r2_prof_fused $CMD > gpurun_out/r2_ncu2.log 2>&1
python tools/microbench_groupby.py 1e7 100 1e8 > gpurun_out/r2_groupby_plain.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"bgpu_intern|bgpu_acc_global|bgpu_sketch" -s 6 -c 3 -o gpurun_out/r2_prof_twopass python tools/microbench_groupby.py 1e7 100 1e8 > gpurun_out/r2_ncu3.log 2>&1
tail -3 gpurun_out/r2_groupby_plain.log
ls -la gpurun_out/*.ncu-rep

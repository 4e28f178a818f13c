# ncu captures of the repartition kernels (config 3 shape); reports land in gpurun_out/
set -x
ROWS=${ROWS:-1e8}
python tools/microbench_partition.py $ROWS 8 16 64 512 > gpurun_out/part_plain.log 2>&1 || exit 1
cat gpurun_out/part_plain.log
BGPU_PART_GEOMETRY=1024,4 python tools/microbench_partition.py $ROWS 64 512
for N in 8 64 512; do
  ncu --set full --clock-control none --import-source on -k regex:"scatter_kernel|count_kernel" -s 6 -c 2 -f -o gpurun_out/prof_scatter_n$N \
    python tools/microbench_partition.py $ROWS $N > gpurun_out/part_ncu_$N.log 2>&1
done
ls -la gpurun_out/*.ncu-rep

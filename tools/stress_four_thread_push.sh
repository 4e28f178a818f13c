# runs the four-thread push test repeatedly (a self-deadlock of the submit path showed up once in ~4 runs before the fix)
fails=0
for i in $(seq 1 ${1:-12}); do
  timeout 40 python -m pytest tests/test_ingest_gpu.py -x -q -k four_threads > gpurun_out/ingest_loop.out 2>&1 || { fails=$((fails+1)); cp gpurun_out/ingest_loop.out gpurun_out/ingest_loop_fail.out; }
done
echo "four-thread push test: $fails failures of ${1:-12}"

# accumulate pass of the two-pass path with and without the shared-memory hot-group cache (uniform keys: orderkey; Zipf: q5)
for H in default 0; do
  if [ $H = 0 ]; then export BGPU_HOT_CACHE=0; else unset BGPU_HOT_CACHE; fi
  echo "== BGPU_HOT_CACHE=$H"
  BGPU_AGG_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-configs --no-cpu --rows 100000000 --e2e-rows 2000000 2>gpurun_out/hc_trace.err | python tools/parse_bench_line.py hc | grep orderkey
  grep "intern\|accumulate" gpurun_out/hc_trace.err | tail -4 | tr "\n" " "; echo
  python tools/microbench_groupby.py 1e8 100 3e7 2>/dev/null | grep -E "q3|q5|orderkey"
done

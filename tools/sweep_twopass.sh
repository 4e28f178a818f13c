for G in 256,8,2 256,8,3 256,4,3 256,4,4 512,4,2 128,8,4; do
  echo "== geometry $G"
  BGPU_TWOPASS_GEOMETRY=$G BGPU_AGG_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 3 --no-configs --no-cpu --rows 100000000 --e2e-rows 2000000 2>gpurun_out/ok_trace_g.err | python tools/parse_bench_line.py orderkey_short
  grep "intern\|accumulate" gpurun_out/ok_trace_g.err | tail -4 | tr "\n" " "; echo
done

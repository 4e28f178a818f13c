# Q1 fused-kernel geometry sweep (GPU box): prints one summary line per variant.
ROWS=${ROWS:-300000000}
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 200 python bench.py --rows $ROWS --steps 5 --no-cpu --e2e-rows 1000000 2>&1 | python tools/parse_bench_line.py "$tag"
}
run "pipe1 row32" BGPU_X=0
run "pipe1 row64" BGPU_ROW64=1
run "pipe2 row32" BGPU_FUSED_PIPE=2
run "pipe2 row64" BGPU_FUSED_PIPE=2 BGPU_ROW64=1

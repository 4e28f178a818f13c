# Q1 fused-kernel geometry sweep (GPU box): prints one summary line per variant.
ROWS=${ROWS:-300000000}
run() { # tag, env...
  tag=$1; shift
  env "$@" timeout 200 python bench.py --rows $ROWS --steps 5 --no-cpu --e2e-rows 1000000 2>&1 | python tools/parse_bench_line.py "$tag"
}
run "default w4" BGPU_X=0
run "default w8" BGPU_FUSED_WAVES=8
run "704,1,1 w8" BGPU_FUSED_GEOMETRY=704,1,1 BGPU_FUSED_WAVES=8
run "768,1,1 w8" BGPU_FUSED_GEOMETRY=768,1,1 BGPU_FUSED_WAVES=8
run "608,1,1 w8" BGPU_FUSED_GEOMETRY=608,1,1 BGPU_FUSED_WAVES=8
run "512,1,1 w8" BGPU_FUSED_GEOMETRY=512,1,1 BGPU_FUSED_WAVES=8
run "640,2,1 w8" BGPU_FUSED_GEOMETRY=640,2,1 BGPU_FUSED_WAVES=8
run "640 gcap5" BGPU_FUSED_GCAP=5 BGPU_FUSED_WAVES=8

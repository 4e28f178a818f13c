# end-to-end leg: the one-pass packer against range + pack (BGPU_INGEST_TWO_PASS=1), AVX-512 against AVX2
nproc
for V in one two one two avx2; do
  unset BGPU_INGEST_TWO_PASS BGPU_INGEST_NO_AVX512
  if [ $V = two ]; then export BGPU_INGEST_TWO_PASS=1; fi
  if [ $V = avx2 ]; then export BGPU_INGEST_NO_AVX512=1; fi
  timeout 300 python bench.py --steps 4 --warmup 3 --no-configs --no-orderkey --no-cpu --rows 20000000 2>/dev/null | python -c "
import json,sys,os
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('$V', 'threads', e['feeder_threads'], 'e2e G rows/s', round(e['value']/1e9,3), 'h2d B/row', round(e['h2d_bytes_per_step']/e['rows_per_gpu'],2), 'verified', d['verified'])"
done

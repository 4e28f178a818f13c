# end-to-end leg: AVX-512 packer against the AVX2 one, and the feeder thread count
nproc
for V in 0 1 0 1; do
  if [ $V = 1 ]; then export BGPU_INGEST_NO_AVX512=1; else unset BGPU_INGEST_NO_AVX512; fi
  timeout 300 python bench.py --steps 4 --warmup 3 --no-configs --no-orderkey --no-cpu --rows 20000000 2>/dev/null | python -c "
import json,sys,os
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('no_avx512' if os.environ.get('BGPU_INGEST_NO_AVX512') else 'avx512   ', 'threads', e['feeder_threads'], 'e2e G rows/s', round(e['value']/1e9,3), 'verified', d['verified'])"
done

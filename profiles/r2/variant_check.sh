#!/bin/bash
# variant_check.sh NAME...: single-stream identity + per-frame device time of libthorcuda variants (build_variant.sh), then the
# multi-stream rate of bench.py's stream leg.  Output: gpurun_out/variants.log
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
for v in "$@"; do
  if [ "$v" = base ]; then export LD_LIBRARY_PATH=; else export LD_LIBRARY_PATH=$PWD/build/variants/$v; fi
  echo "=== $v" >> gpurun_out/variants.log
  [ -n "$SKIP_SINGLE" ] || timeout 600 python profiles/r2/enc_gpu_check.py LDB_medium_complexity:416:240:6:32 LDB_medium_complexity:1920:1080:4:32 2>&1 | python -c "
import sys, json
for ln in sys.stdin:
    try: d = json.loads(ln)
    except Exception: print(ln.strip()[:300]); continue
    print(d.get('size'), 'identical', d.get('identical_bit'), d.get('identical_rec'), [t for t in d.get('gpu_timing', []) if 'per frame' in t], d.get('gpu_error', '')[:300])
" >> gpurun_out/variants.log
  for sc in ${STREAMS:-16:9}; do
    s=${sc%%:*}; c=${sc##*:}
    timeout 900 python bench.py --no-cpu --no-decoder --no-segments --streams $s --ctas-per-frame $c --steps 6 --warmup 3 $BENCH_EXTRA 2>&1 | python -c "
import sys, json
for ln in sys.stdin:
    try: d = json.loads(ln)
    except Exception: print(ln.strip()[:300]); continue
    print('streams $s ctas $c value', round(d['value'], 2), 'e2e', round(d['e2e']['value'], 2))
" >> gpurun_out/variants.log
  done
done

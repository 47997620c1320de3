#!/bin/bash
# device-resident frames/s of bench.py for several batch sizes
for b in 8 16 32 64 128 256; do
  v=$(python bench.py --steps 10 --warmup 3 --batch $b --no-cpu-baseline --no-extra 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']))")
  echo "batch $b value/e2e $v"
done

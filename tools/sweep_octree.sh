#!/bin/bash
# quad-tree stage time (ms per batch) for each CTA size and batch size (ORBX_OCT_T is an experiment knob)
for b in 8 16 32 64 128 256; do
  for t in 128 256 512; do
    v=$(ORBX_OCT_T=$t python bench.py --steps 10 --warmup 3 --batch $b --no-cpu-baseline --no-extra 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['e2e']['value']), round(d['stages']['octree']['ms'],4))")
    echo "batch $b T $t value/e2e/octree_ms $v"
  done
done

#!/bin/bash
# e2e frames/s of bench.py for several host-pipeline chunk sizes (ORBX_PIPE_CHUNK); two runs each
for c in ${CHUNKS:-24 32 40 48 64}; do
  for r in 1 2; do
    v=$(ORBX_PIPE_CHUNK=$c python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['e2e']['value']))")
    echo "chunk $c e2e $v"
  done
done

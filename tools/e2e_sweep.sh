#!/bin/bash
# e2e throughput of bench.py as a function of the streaming chunk size (frames per upload step)
for c in "$@"; do
    python bench.py --steps 8 --warmup 3 --no-cpu-baseline --e2e-chunk $c 2>&1 | tail -1 | \
        python -c "import sys,json; d=json.loads(sys.stdin.read()); print('chunk', $c, 'e2e', round(d['e2e']['value'],1), 'bytes', d['e2e']['h2d_bytes_per_step'])"
done

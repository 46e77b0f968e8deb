#!/bin/bash
# bench.py headline over early-exit chunk sizes and sub-batch sizes (one ring render per run; ~40 s each)
for sb in 64 125; do for ec in 4 5 6 8; do
  python bench.py --no-cpu-baseline --trigger-steps 0 --e2e-steps 1 --e2e-events 16 --early-chunk $ec --sub-batch $sb --steps 12 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
k={x['kernel'].split()[0]:x['ms_per_step'] for x in d['roofline']['kernels']}
print('sub_batch $sb early_chunk $ec value %.1f ms %.2f K1 %.2f K2 %.2f pairs %.3f launches %d'%(d['value'], d['ms_per_step'], k.get('K1',0), k.get('K2',0), d['roofline']['frame_pairs_computed_frac'], d['gpu_launches']))"
done; done

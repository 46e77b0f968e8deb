#!/bin/bash
# bench.py headline over localizer scratch sizes (chunks per sub-batch), slot sets, sub-batch and early-exit chunk sizes
run() {
  python bench.py --no-cpu-baseline --trigger-steps 0 --e2e-steps 1 --e2e-events 16 --steps 12 "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
k={x['kernel'].split()[0]:x['ms_per_step'] for x in d['roofline']['kernels']}
loc=[x for x in d['roofline']['kernels'] if x['kernel'].startswith('localizer')][0]
print('$LABEL $* value %.1f ms %.2f K1 %.2f K2 %.2f loc %.2f (K5b %.2f K7b %.2f) host %.2f launches %d'%(d['value'], d['ms_per_step'], k.get('K1',0), k.get('K2',0), loc['ms_per_step'], k.get('K5b',0), k.get('K7b',0), 1e3*d['stage_seconds_per_step']['host_analysis_s'], d['gpu_launches']))"
}
LABEL="scratch 4GB(default)" run
for mb in 8000 16000; do LABEL="scratch ${mb}MB" ABUB_LOC_SCRATCH_MB=$mb run; done
LABEL="scratch 16000MB" ABUB_LOC_SCRATCH_MB=16000 run --early-chunk 6

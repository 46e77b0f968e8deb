#!/bin/bash
# After tools/round_profile.sh came back: summaries of gpurun_out/r2_* -> profiles/r2_* (run here, no GPU needed).
set -u
O=gpurun_out
P=profiles
for f in $O/r2_*.ncu-rep; do
    [ -e "$f" ] || continue
    t=$(basename "$f" .ncu-rep)
    python tools/ncu_summary.py "$f" > $P/${t}_full.txt
done
[ -e $O/r2_launches_bench.csv ] && python tools/launch_summary.py $O/r2_launches_bench.csv > $P/r2_launches_bench.txt
if [ -e $O/r2_step.csv ]; then
    if [ -e $O/r2_step_full.csv ]; then python tools/ncu_traffic.py $O/r2_step.csv $O/r2_step_full.csv > $P/r2_traffic.json
    else python tools/ncu_traffic.py $O/r2_step.csv > $P/r2_traffic.json; fi
    python tools/launch_summary.py $O/r2_step.csv > $P/r2_launches_step.txt 2>/dev/null || true
fi
for j in r2_bench_n1.json r2_stages.json; do [ -s $O/$j ] && cp $O/$j $P/$j; done
ls -la $P | grep r2_

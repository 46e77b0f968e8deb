#!/bin/bash
# One GPU-box pass that produces the evidence kept under profiles/ (round 2): the bench line, a launch list of the
# bench, DRAM traffic per kernel group of one bench step (-> profiles/r2_traffic.json, which bench.py reads), and one
# `ncu --set full` capture per kernel family.  Every profiled command first runs to exit code 0 without ncu.
#   gpurun --timeout 2400 -- 'bash tools/round_profile.sh'
# then here:  bash tools/round_profile_collect.sh   (summaries -> profiles/r2_*)
set -u
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --trigger-steps 0 --e2e-steps 1 --e2e-events 16"
python bench.py > $O/r2_bench_n1.json 2> $O/r2_bench_n1.err || echo "bench failed"
$B --profile-step > $O/r2_bench_short.json 2> $O/r2_bench_short.err || { echo "short bench failed"; tail -5 $O/r2_bench_short.err; }
K='regex:k[1-9]_|DeviceRadixSort|DeviceScan|DeviceSelect'
# launch list of a whole short bench run (own kernels + the CUB kernels they call)
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file $O/r2_launches_bench.csv \
    $B > $O/ncu_lb.log 2>&1
# one step, with DRAM bytes: early-exit form (the headline) and the full-table form
ncu --profile-from-start off --clock-control none --csv --log-file $O/r2_step.csv \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    $B --profile-step > $O/ncu_step.log 2>&1
ncu --profile-from-start off --clock-control none --csv --log-file $O/r2_step_full.csv \
    --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum \
    $B --profile-step --early-chunk 0 > $O/ncu_step_full.log 2>&1
# --set full, one launch per kernel family, inside the profiled step (ONLY="k5b k5a": just these tags)
SPECS="k1_trigger_tma:k1_early:0 k5_post_tma:k5b:0 k5_genesis_tma:k5a:0 k7_trace2:k7b:0"
for spec in $SPECS; do
    IFS=: read -r kern tag skip <<< "$spec"
    if [ -n "${ONLY:-}" ] && ! echo " $ONLY " | grep -q " $tag "; then continue; fi
    ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$kern" -s $skip -c 1 -o $O/r2_$tag -f \
        $B --profile-step > $O/ncu_$tag.log 2>&1
done
if [ -z "${ONLY:-}" ]; then
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:k1_trigger_tma -c 1 -o $O/r2_k1_full -f \
    $B --profile-step --early-chunk 0 > $O/ncu_k1_full.log 2>&1
fi
[ -n "${ONLY:-}" ] && { ls -la $O | tail -20; exit 0; }
# Trainer kernels (K3, K4) and the stage numbers of tools/bench_stages.py
python tools/bench_stages.py --events 32 --train-events 200 --reps 3 > $O/r2_stages.json 2> $O/r2_stages.err || echo "stages failed"
for spec in "k4_welford16:k4" "k3_gate_batch:k3"; do
    IFS=: read -r kern tag <<< "$spec"
    ncu --set full --clock-control none --import-source on -k regex:$kern -c 1 -o $O/r2_$tag -f \
        python tools/bench_stages.py --events 32 --train-events 64 --reps 1 > $O/ncu_$tag.log 2>&1
done
ls -la $O | tail -40

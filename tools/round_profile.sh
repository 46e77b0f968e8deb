#!/bin/bash
# One GPU-box pass that produces the evidence kept under profiles/: bench lines (own arm + reference arm), a launch
# list of the bench and of the per-stage tool (own kernels only), and one ncu --set full capture per kernel family.
# Run with:  gpurun --timeout 1500 -- 'bash tools/round_profile.sh'
set -u
O=gpurun_out
mkdir -p $O
python bench.py --steps 16 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err || echo "bench failed"
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err || echo "ref bench failed"
python tools/bench_stages.py --events 32 --train-events 200 --reps 3 > $O/stages.json 2> $O/stages.err || echo "stages failed"
K='regex:k[1-9]_|DeviceRadixSort|DeviceScan|DeviceSelect'
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file $O/launches_bench.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_lb.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file $O/launches_stages.csv \
    python tools/bench_stages.py --events 32 --train-events 32 --reps 1 > $O/ncu_ls.log 2>&1
python tools/bench_decode.py --events 12 --batch 12 > $O/decode_bmp.json 2> $O/decode_bmp.err || echo "decode bmp failed"
python tools/bench_decode.py --events 24 --batch 24 --format png > $O/decode_png.json 2> $O/decode_png.err || echo "decode png failed"
ncu --set full --clock-control none --import-source on -k regex:k9_inflate -c 1 -o $O/r1f_k9 -f \
    python tools/bench_stages.py --events 8 --png-events 8 --only-png --reps 1 > $O/ncu_k9.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k9_unfilter -c 1 -o $O/r1f_k9u -f \
    python tools/bench_stages.py --events 8 --png-events 8 --only-png --reps 1 > $O/ncu_k9u.log 2>&1
for spec in "k1_trigger_tma:k1_tma:1" "k5_post_stream:k5b:1" "k6_candidates:k6:1" "k7_trace2:k7:1" "k8_ncc:k8:1" "k4_welford16:k4:1" "k3_gate_batch:k3:1"; do
    IFS=: read -r kern tag cnt <<< "$spec"
    if [ "$kern" = "k1_trigger_tma" ]; then
        ncu --set full --clock-control none --import-source on -k regex:$kern -s 1 -c $cnt -o $O/r1f_$tag -f \
            python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/ncu_$tag.log 2>&1
    else
        ncu --set full --clock-control none --import-source on -k regex:$kern -c $cnt -o $O/r1f_$tag -f \
            python tools/bench_stages.py --events 32 --train-events 64 --reps 1 > $O/ncu_$tag.log 2>&1
    fi
done
ls -la $O | tail -30

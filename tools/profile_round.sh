#!/bin/bash
# One profiling pass on the GPU box: plain bench first (must exit 0), then the ncu launch list and --set full captures.
# usage: tools/profile_round.sh <tag>     outputs under gpurun_out/<tag>_*
set -u
tag=${1:-r02a}
O=gpurun_out
mkdir -p $O
python bench.py > $O/${tag}_bench_1gpu.json 2> $O/${tag}_bench.err || { echo "plain bench failed"; tail -5 $O/${tag}_bench.err; exit 1; }
SMALL="--steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-from-iq --no-configs --sustain-seconds 0"
python bench.py $SMALL > $O/${tag}_plain_small.json 2>> $O/${tag}_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${tag}_launches.csv \
    python bench.py $SMALL > $O/${tag}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -o $O/${tag}_main -f \
    -k regex:"eq_fast|snr_tags|ldpc_bp_ref|ldpc_screen|bp_gather|tb_finish|tb_gather|tb_plan|carry_save|stream_state|fft64" -c 22 \
    python bench.py $SMALL > $O/${tag}_ncu2.log 2>&1
python tools/ncu_metrics.py $O/${tag}_main.ncu-rep $O/${tag}_ncu_full_metrics.csv > $O/${tag}_ncu_full_metrics.txt 2>&1
for k in eq_fast ldpc_bp_ref ldpc_screen; do
  python tools/ncu_sass_hot.py $O/${tag}_main.ncu-rep $k 14 >> $O/${tag}_ncu_source_hotspots.txt 2>&1
done
python bench.py $SMALL --ldpc-mode layered > $O/${tag}_plain_lay.json 2>> $O/${tag}_bench.err && \
ncu --set full --clock-control none --import-source on -o $O/${tag}_lay -f -k regex:"ldpc_bp_lay" -c 2 \
    python bench.py $SMALL --ldpc-mode layered > $O/${tag}_ncu3.log 2>&1
python tools/ncu_metrics.py $O/${tag}_lay.ncu-rep $O/${tag}_ncu_lay_metrics.csv > $O/${tag}_ncu_lay_metrics.txt 2>&1
python tools/ncu_sass_hot.py $O/${tag}_lay.ncu-rep ldpc_bp_lay 14 >> $O/${tag}_ncu_source_hotspots.txt 2>&1
# configuration legs: config 1 (no-FEC byte kernel), config 5 (N = 1024), config 3 per constellation
python tools/leg.py config1 4 3 > $O/${tag}_leg_config1.json 2>> $O/${tag}_bench.err && \
ncu --set full --clock-control none -o $O/${tag}_c1 -f -k regex:"eq_fast|tb_|frame_gather|fft64" -c 8 python tools/leg.py config1 4 2 > $O/${tag}_ncu4.log 2>&1
python tools/ncu_metrics.py $O/${tag}_c1.ncu-rep $O/${tag}_ncu_config1_metrics.csv > $O/${tag}_ncu_config1_metrics.txt 2>&1
python tools/leg.py config5 6 3 > $O/${tag}_leg_config5.json 2>> $O/${tag}_bench.err && \
ncu --set full --clock-control none -o $O/${tag}_c5 -f -k regex:"eq_fast|fft1024|fft_fwd|ldpc_screen|tb_finish|tb_gather" -c 10 python tools/leg.py config5 6 2 > $O/${tag}_ncu5.log 2>&1
python tools/ncu_metrics.py $O/${tag}_c5.ncu-rep $O/${tag}_ncu_config5_metrics.csv > $O/${tag}_ncu_config5_metrics.txt 2>&1
for c in 1 2 3 5 6; do
  ncu --set full --clock-control none -o $O/${tag}_c3_$c -f -k regex:"eq_fast" -c 1 --launch-skip 2 python tools/leg.py config3 $c 2 > $O/${tag}_ncu6_$c.log 2>&1
  python tools/ncu_metrics.py $O/${tag}_c3_$c.ncu-rep $O/${tag}_ncu_config3_cnst$c.csv > $O/${tag}_ncu_config3_cnst$c.txt 2>&1
done
rm -f $O/${tag}_c3_*.ncu-rep $O/${tag}_c1.ncu-rep $O/${tag}_c5.ncu-rep $O/${tag}_main.ncu-rep $O/${tag}_lay.ncu-rep
python bench.py --impl reference --steps 3 --warmup 1 > $O/${tag}_bench_reference_arm.json 2>> $O/${tag}_bench.err
ls -la $O | grep ${tag} | head -40

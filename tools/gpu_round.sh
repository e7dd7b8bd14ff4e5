#!/bin/bash
# One GPU session: parity tests, bench, timing twin, ncu launch list, ncu --set full of the conv kernels.
# Usage (from the repo root on the GPU box): bash tools/gpu_round.sh <tag> [steps: t b f l n]
tag=${1:-x}; what=${2:-tbfln}
mkdir -p gpurun_out
if [[ $what == *t* ]]; then python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${tag}_pytest.log; fi
if [[ $what == *b* ]]; then python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; fi
if [[ $what == *f* ]]; then python tools/fused_timing.py 16 3000 > gpurun_out/${tag}_fused_timing.txt 2>&1; echo "timing rc=$?"; tail -8 gpurun_out/${tag}_fused_timing.txt; fi
if [[ $what == *l* ]]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches.csv \
  python bench.py --steps 1 --warmup 1 --chunks 16 --no-cpu > gpurun_out/${tag}_ncu_launch.log 2>&1; echo "launchlist rc=$?"; fi
if [[ $what == *n* ]]; then
timeout 900 ncu --set full --clock-control none -k regex:'mrf_fused|conv_mma' -c 48 -f -o /tmp/${tag}_full \
  python bench.py --steps 1 --warmup 1 --chunks 8 --no-cpu > gpurun_out/${tag}_ncu_full.log 2>&1; echo "full rc=$?"
python tools/ncu_summary.py /tmp/${tag}_full.ncu-rep > gpurun_out/${tag}_ncu_full_summary.csv 2> gpurun_out/${tag}_ncu_summary.err
sz=$(stat -c %s /tmp/${tag}_full.ncu-rep 2>/dev/null || echo 0); echo "rep bytes $sz"
if [ "$sz" -gt 0 ] && [ "$sz" -lt 40000000 ]; then cp /tmp/${tag}_full.ncu-rep gpurun_out/; fi
fi
du -sh gpurun_out
if [[ $what == *s* ]]; then
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_synth_launches.csv \
  python tools/synth_bench.py 4 3000 > gpurun_out/${tag}_ncu_synth.log 2>&1; echo "synth launchlist rc=$?"; fi

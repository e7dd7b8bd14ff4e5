#!/bin/bash
# ncu --set full with source correlation of the fused-MRF launches only; prints the hottest SASS per launch.
tag=${1:-x}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'mrf_fused' -c 6 -f -o /tmp/${tag}_hot \
  python bench.py --steps 1 --warmup 1 --chunks 8 --no-cpu > gpurun_out/${tag}_hot_ncu.log 2>&1; echo "ncu rc=$?"
for k in 1 3 4 6; do
  python tools/ncu_hot.py /tmp/${tag}_hot.ncu-rep $k 45 > gpurun_out/${tag}_hot_launch$k.txt 2>&1
done
python tools/ncu_summary.py /tmp/${tag}_hot.ncu-rep > gpurun_out/${tag}_hot_summary.csv 2>/dev/null; cp /tmp/${tag}_hot.ncu-rep gpurun_out/
ls -la /tmp/${tag}_hot.ncu-rep; du -sh gpurun_out

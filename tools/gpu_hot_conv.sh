#!/bin/bash
# ncu --set full with source correlation of ONE conv_mma launch of the bench's first pass (0-based index among the
# conv_mma launches: 0 conv_pre, 1 ups0, 2-19 stage 0, 20 ups1, ...); prints the hottest SASS and the summary row.
tag=${1:-x}; skip=${2:-20}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'conv_mma' --launch-skip $skip --launch-count 1 -f -o /tmp/${tag}_hotc \
  python bench.py --steps 1 --warmup 1 --chunks 8 --no-cpu > gpurun_out/${tag}_hotc_ncu.log 2>&1; echo "ncu rc=$?"
python tools/ncu_hot.py /tmp/${tag}_hotc.ncu-rep 1 70 > gpurun_out/${tag}_hotc_launch.txt 2>&1
python tools/ncu_summary.py /tmp/${tag}_hotc.ncu-rep > gpurun_out/${tag}_hotc_summary.csv 2>/dev/null
cp /tmp/${tag}_hotc.ncu-rep gpurun_out/
cat gpurun_out/${tag}_hotc_summary.csv

#!/bin/bash
# A/B of two builds of librvcdec.so on the bench workload in ONE call (same box): bash tools/ab_lib.sh <tag> <other.so> [reps]
tag=$1; other=$2; reps=${3:-2}
mkdir -p gpurun_out
for r in $(seq 1 $reps); do
  for which in cur other; do
    if [ $which = other ]; then export RVC_B200_LIB=$PWD/$other; else unset RVC_B200_LIB; fi
    python bench.py --chunks ${CHUNKS:-128} --steps 3 --warmup 2 --no-cpu --no-longform > gpurun_out/${tag}_${which}$r.json 2> gpurun_out/${tag}_${which}$r.err
    echo "[$which $r] rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_${which}$r.json"))
print(round(d["value"],1), round(d["ms_per_step"],2), d["clocks"]["sm_mhz"], {k: round(v["ms_per_step"],2) for k,v in d["layers"].items()})
PY
  done
done

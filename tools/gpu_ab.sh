#!/bin/bash
# A/B of tuning options on the bench workload (64 chunks to save GPU time): bash tools/gpu_ab.sh <tag> "<opts A>" "<opts B>" ...
tag=$1; shift
mkdir -p gpurun_out
i=0
for o in "$@"; do
  RVC_B200_OPTS="$o" python bench.py --chunks ${CHUNKS:-64} --steps 3 --warmup 2 --no-cpu --no-longform > gpurun_out/${tag}_ab$i.json 2> gpurun_out/${tag}_ab$i.err
  echo "[$o] rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_ab$i.json"))
print(round(d["value"],1), round(d["ms_per_step"],2), d["clocks"]["sm_mhz"], {k: round(v["ms_per_step"],2) for k,v in d["layers"].items() if k.startswith("mrf")}, {k: round(v["ms_per_step"],2) for k,v in d["layers"].items() if not k.startswith("mrf")})
PY
  i=$((i+1))
done

# A/B of the device-pointer path's scratch cap (number of passes per 128-chunk step): bash tools/ws_ab.sh
for ws in 28 44 56 120 28 56 120; do
RVC_B200_MAX_WORKSPACE=$((ws<<30)) python bench.py --steps 3 --warmup 2 --no-cpu > gpurun_out/ws_$ws.json 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/ws_$ws.json"))
print($ws, "GiB:", round(d["value"],1), round(d["ms_per_step"],2), d["clocks"]["sm_mhz"], d["gpu_launches"], {k: round(v["ms_per_step"],1) for k,v in d["layers"].items() if k.startswith("mrf")})
PY
done

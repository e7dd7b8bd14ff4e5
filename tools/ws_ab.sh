for ws in 34359738368 128849018880 34359738368 128849018880; do
RVC_B200_MAX_WORKSPACE=$ws python bench.py --steps 3 --warmup 2 --no-cpu > gpurun_out/ws_$ws.json 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/ws_$ws.json"))
print($ws>>30, "GiB:", round(d["value"],1), round(d["ms_per_step"],2), d["clocks"]["sm_mhz"], d["gpu_launches"], round(d["e2e"]["value"],1))
PY
done

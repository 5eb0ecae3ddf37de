for cfg in prodlda_10k neurallda_20k ctm_50k zeroshot_50k prodlda_200k; do
  timeout 300 python bench.py --config $cfg --steps 30 --warmup 5 2>/dev/null | tail -1 > gpurun_out/ours_$cfg.json
  timeout 600 python bench.py --impl reference --config $cfg --steps 5 --warmup 3 2>/dev/null | tail -1 > gpurun_out/ref_$cfg.json
  python - <<PY
import json
for arm in ("ours","ref"):
    try:
        d=json.load(open("gpurun_out/%s_$cfg.json"%arm))
        print("$cfg", arm, d.get("ms_per_step"), d.get("value"), (d.get("e2e") or {}).get("value"), d.get("gpu_launches"), d.get("unavailable"))
    except Exception as e:
        print("$cfg", arm, "ERR", e)
PY
done

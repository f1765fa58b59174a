# GPU check used through the round: GPU tests, then short bench lines of the configs given as arguments (default 1)
python -m pytest tests -m gpu -q -x 2>&1 | tail -6
for c in ${@:-1}; do
  python bench.py --config $c --no-extra > gpurun_out/b_c$c.json 2> gpurun_out/b_c$c.err
  python - <<PY
import json
d = json.loads(open("gpurun_out/b_c$c.json").read().strip().splitlines()[-1])
print("configs[$c]", round(d["value"], 1), round(d["ms_per_step"], 2), round(d["e2e"]["value"], 1), d.get("kernel_ms_per_step_single_stream"), d.get("decide_phase_gcycles"))
PY
done

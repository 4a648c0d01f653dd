mkdir -p gpurun_out
for O in 0 1; do
RCM_OVERLAP=$O python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/ov$O.json 2>gpurun_out/ov$O.err
python - <<PY
import json
d=json.load(open("gpurun_out/ov$O.json")); print("overlap=$O ms/step", d["ms_per_step"], d["value"]/1e9)
PY
done
python -m pytest tests -x -q -m gpu 2>&1 | tail -2

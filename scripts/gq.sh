python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu > gpurun_out/b.json 2>gpurun_out/b.err; tail -2 gpurun_out/b.err; python - <<PY
import json
d=json.load(open("gpurun_out/b.json"))
print(d["ms_per_step"], d["value"]/1e9)
for k,v in d["roofline"]["per_kernel"].items(): print(k, round(v["ms_per_step"],3), v["GBs"] and round(v["GBs"]))
PY

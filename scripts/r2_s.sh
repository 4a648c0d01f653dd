# call S: shared-reciprocal divisions on by default, MOLOCH without the pai/deltaw exchange; GPU suite + benches
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2q_pytest.log
for w in C4 C4hq C5 C1; do
  st=20; [ $w = C1 ] && st=60
  timeout 300 python bench.py --workload $w --steps $st --warmup 3 --no-cpu > gpurun_out/s_$w.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/s_$w.json").read().strip().splitlines()[-1])
print("%-5s ms/step %.3f e2e %.3f" % ("$w", d['ms_per_step'], d['e2e']['ms_per_step']), {k:round(x["ms_per_step"],3) for k,x in d["roofline"]["per_kernel"].items() if x["ms_per_step"] > 0.02})
PY
done

# call AA: after ipgf = 1 and the semi-Lagrangian advection: full GPU suite (no -x), C4 / C4hq bench lines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2aa_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2aa_pytest.log
for w in C4 C4hq; do
  timeout 400 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/aa_$w.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/aa_$w.json").read().strip().splitlines()[-1])
print("%-5s ms/step %.3f value %.4g" % ("$w", d['ms_per_step'], d['value']))
PY
done

# call Y: after idiffu = 2 and the UW-PBL coupling (TKE, vadv4d ind = 3): full GPU suite, smoke, C4 / C4hq / C1 bench lines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2y_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2y_smoke.log 2>&1; tail -1 gpurun_out/r2y_smoke.log
for w in C4 C4hq C1; do
  timeout 400 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/y_$w.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/y_$w.json").read().strip().splitlines()[-1])
print("%-5s ms/step %.3f value %.4g" % ("$w", d['ms_per_step'], d['value']))
PY
done
timeout 600 python bench.py > gpurun_out/y_default.json 2> gpurun_out/y_default.err; tail -c 600 gpurun_out/y_default.json

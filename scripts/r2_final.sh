# final evidence of the round: GPU suite, smoke, default bench line, reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2final_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2final_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2final_smoke.log 2>&1; tail -1 gpurun_out/r2final_smoke.log
timeout 600 python bench.py > gpurun_out/final_default.json 2> gpurun_out/final_default.err
python - <<PY
import json
d=json.loads(open("gpurun_out/final_default.json").read().strip().splitlines()[-1])
print("default: ms/step %.3f value %.4g e2e %.3f ms roof %s %.3f clocks %s" % (d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac'], d['clocks']['sm_mhz']))
PY

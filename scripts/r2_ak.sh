# call AK: file formats on the GPU library, the semi-Lagrangian tests on the wider-halo build, full GPU suite, default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2ak_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2ak_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2ak_smoke.log 2>&1; tail -1 gpurun_out/r2ak_smoke.log
timeout 400 python bench.py > gpurun_out/r2ak_default.json 2> gpurun_out/r2ak_default.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2ak_default.json").read().strip().splitlines()[-1])
print("default: ms/step %.3f value %.4g e2e %.3f ms roof %s %.3f clocks %s" % (d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac'], d['clocks']['sm_mhz']))
PY

# round 2, call C: full-size oracle comparisons (incl. the C4 headline config), negative-q tests, C4 bench after templating the moisture kernels
mkdir -p gpurun_out
( timeout 1800 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_negq.py -q --durations=5 ) > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r2c_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2c_bench.json").read().strip().splitlines()[-1])
print("ms/step",round(d["ms_per_step"],3),"e2e ms",round(d["e2e"]["ms_per_step"],3),{k:round(v["ms_per_step"],3) for k,v in d["roofline"]["per_kernel"].items()})
PY

# round 2, call B: negative-moisture repair and full-size oracle comparisons on the GPU, C4 bench after the change
mkdir -p gpurun_out
free -g | head -2 > gpurun_out/r2b_host.txt; nproc >> gpurun_out/r2b_host.txt; cat gpurun_out/r2b_host.txt
( timeout 1500 python -m pytest tests/test_gpu_negq.py tests/test_gpu_fullsize.py -x -q --durations=8 ) > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2b_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2b_bench.json").read().strip().splitlines()[-1])
print("ms/step",round(d["ms_per_step"],3),"e2e ms",round(d["e2e"]["ms_per_step"],3),{k:round(v["ms_per_step"],3) for k,v in d["roofline"]["per_kernel"].items()})
PY

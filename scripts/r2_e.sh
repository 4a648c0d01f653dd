# round 2, call E: async API on the GPU, C4 bench with the overlapped e2e leg, C5 and C1 bench lines
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_async.py tests/test_gpu_parity.py -x -q ) > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2e_pytest.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r2e_bench_c4.json 2> gpurun_out/r2e_bench_c4.err; echo "c4 rc=$?"
timeout 300 python bench.py --workload C5 --steps 20 --warmup 3 --no-cpu > gpurun_out/r2e_bench_c5.json 2> gpurun_out/r2e_bench_c5.err; echo "c5 rc=$?"
timeout 300 python bench.py --workload C1 --steps 60 --warmup 10 --no-cpu > gpurun_out/r2e_bench_c1.json 2> gpurun_out/r2e_bench_c1.err; echo "c1 rc=$?"
python - <<PY
import json
for n in ("c4","c5","c1"):
    try:
        d=json.loads(open(f"gpurun_out/r2e_bench_{n}.json").read().strip().splitlines()[-1])
        print(n,"ms/step",round(d["ms_per_step"],3),"e2e ms",round(d["e2e"]["ms_per_step"],3),"value %.4g e2e %.4g"%(d["value"],d["e2e"]["value"]), "cpu", d.get("cpu_baseline") and "%.3g"%d["cpu_baseline"]["value"])
    except Exception as e: print(n,"failed",e); print(open(f"gpurun_out/r2e_bench_{n}.err").read()[-600:])
PY

mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_async.py -x -q ) > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2g_pytest.log
python scripts/e2e_probe.py C4 > gpurun_out/r2g_probe.txt 2>&1; cat gpurun_out/r2g_probe.txt
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2g_bench_c4.json 2> gpurun_out/r2g_bench_c4.err; echo "c4 rc=$?"
timeout 300 python bench.py --workload C1 --steps 60 --warmup 10 --no-cpu > gpurun_out/r2g_bench_c1.json 2> gpurun_out/r2g_bench_c1.err; echo "c1 rc=$?"
python - <<PY
import json
for n in ("c4","c1"):
    try:
        d=json.loads(open(f"gpurun_out/r2g_bench_{n}.json").read().strip().splitlines()[-1])
        print(n,"ms/step",round(d["ms_per_step"],4),"e2e ms",round(d["e2e"]["ms_per_step"],4),"value %.4g e2e %.4g"%(d["value"],d["e2e"]["value"]))
    except Exception as e: print(n,"failed",e); print(open(f"gpurun_out/r2g_bench_{n}.err").read()[-600:])
PY

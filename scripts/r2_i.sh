# round 2, call I: 32-bit index arithmetic: parity subset, C4 / C5 / C4hq bench lines
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_async.py tests/test_gpu_moloch.py -x -q ) > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2i_pytest.log
for w in C4 C5 C4hq; do
timeout 400 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu > gpurun_out/r2i_bench_$w.json 2> gpurun_out/r2i_bench_$w.err; echo "$w rc=$?"
done
python - <<PY
import json
for n in ("C4","C5","C4hq"):
    try:
        d=json.loads(open(f"gpurun_out/r2i_bench_{n}.json").read().strip().splitlines()[-1])
        print(n,"ms/step",round(d["ms_per_step"],3),"e2e ms",round(d["e2e"]["ms_per_step"],3),{k:round(v["ms_per_step"],3) for k,v in d["roofline"]["per_kernel"].items()})
    except Exception as e: print(n,"failed",e); print(open(f"gpurun_out/r2i_bench_{n}.err").read()[-600:])
PY

# round 2, call A: baseline GPU tests, FP64 issue-rate peaks, C4 bench with the exact (-fmad=false) and the FMA build
mkdir -p gpurun_out
./scripts/fp64_peak > gpurun_out/r2a_fp64.json 2>&1; cat gpurun_out/r2a_fp64.json
( timeout 900 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2a_pytest.log
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2a_bench_exact.json 2> gpurun_out/r2a_bench_exact.err; echo "exact rc=$?"
RCM_DYN_LIB=$PWD/regcm_b200/csrc/variants/libregcm_dyn_fma.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu > gpurun_out/r2a_bench_fma.json 2> gpurun_out/r2a_bench_fma.err; echo "fma rc=$?"
python - <<PY
import json
for n in ("exact","fma"):
    try:
        d=json.loads(open(f"gpurun_out/r2a_bench_{n}.json").read().strip().splitlines()[-1])
        print(n,"ms/step",round(d["ms_per_step"],3),"e2e ms",round(d["e2e"]["ms_per_step"],3),{k:round(v["ms_per_step"],3) for k,v in d["roofline"]["per_kernel"].items()})
    except Exception as e: print(n,"failed",e)
PY

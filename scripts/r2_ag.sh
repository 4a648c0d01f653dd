# call AG: k_tr with 4 / 2 / 1 species per thread; tracer parity on the GPU again
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tracers.py tests/test_gpu_tke.py tests/test_sladvec.py tests/test_gpu_idiffu2.py -m gpu -q > gpurun_out/r2ag_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2ag_pytest.log
run() { tag=$1; w=$2; shift 2; opts=""; for o in "$@"; do opts="$opts --opt $o"; done
  timeout 500 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e $opts > gpurun_out/ag_$tag.json 2> gpurun_out/ag_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ag_$tag.json").read().strip().splitlines()[-1])
    pk=d["roofline"]["per_kernel"]
    print("%-12s ms/step %.3f" % ("$tag", d['ms_per_step']), {k:round(x["ms_per_step"],3) for k,x in pk.items() if k in ("k_tr","k_tq","k_bdy","k_negq")})
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/ag_$tag.err").read()[-800:])
PY
}
run ntr4 C4 ntr=4
run ntr2 C4 ntr=2
run ntr1 C4 ntr=1
run ntr8 C4 ntr=8

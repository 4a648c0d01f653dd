# Round-end check on one GPU: the whole gpu-marked suite, smoke(), and the A/B of the vertical WAF grouping.
mkdir -p gpurun_out
( time timeout 600 python -m pytest tests -x -q -m gpu ) > gpurun_out/final_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/final_pytest.log; tail -6 gpurun_out/final_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; tail -3 gpurun_out/final_smoke.log
for z in 7 4 3 2; do
  RCM_MO_WAFZ=$z timeout 200 python bench.py --workload C5 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/mo_abz_$z.json 2> gpurun_out/mo_abz_$z.err; echo "wafz group $z rc=$?"
done
python - <<PY
import json
for n in "7432":
    try:
        d=json.loads(open(f"gpurun_out/mo_abz_{n}.json").read().strip().splitlines()[-1])
        pk=d["roofline"]["per_kernel"]
        print("wafz group",n,"ms/step",round(d["ms_per_step"],3)," ".join(f"{k}={v['ms_per_step']:.3f}" for k,v in pk.items()))
    except Exception as e: print(n,"ERR",e)
PY

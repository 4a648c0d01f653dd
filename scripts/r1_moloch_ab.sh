# A/B of the MOLOCH advection variants on one GPU: parity tests with the default build, then the C5 bench
# per RCM_MO_WAF mode (bit 0: one-cell meridional kernel, bit 1: one-cell zonal kernel).
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_moloch.py -x -q ) > gpurun_out/mo_ab_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/mo_ab_pytest.log; tail -3 gpurun_out/mo_ab_pytest.log
for m in 0 1 2 3; do
  RCM_MO_WAF=$m timeout 200 python bench.py --workload C5 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/mo_ab_$m.json 2> gpurun_out/mo_ab_$m.err; echo "mode $m rc=$?"
done
RCM_MO_WAF=0 timeout 200 python -m pytest tests/test_gpu_moloch.py -x -q -k "parity_vs_oracle or golden or long_run" > gpurun_out/mo_ab_pytest0.log 2>&1; tail -2 gpurun_out/mo_ab_pytest0.log
python - <<PY
import json
for n in "0123":
    try:
        d=json.loads(open(f"gpurun_out/mo_ab_{n}.json").read().strip().splitlines()[-1])
        pk=d["roofline"]["per_kernel"]
        print("mode",n,"ms/step",round(d["ms_per_step"],3)," ".join(f"{k}={v['ms_per_step']:.3f}" for k,v in pk.items()))
    except Exception as e: print(n,"ERR",e)
PY

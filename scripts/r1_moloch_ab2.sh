# A/B of the k_mo_w scratch placement (RCM_MO_WSM=1 shared memory, 0 global memory) on one GPU,
# after the MOLOCH parity tests with the default build.
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_moloch.py -x -q ) > gpurun_out/mo_ab2_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/mo_ab2_pytest.log; tail -3 gpurun_out/mo_ab2_pytest.log
for m in 1 0; do
  RCM_MO_WSM=$m timeout 200 python bench.py --workload C5 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/mo_ab2_$m.json 2> gpurun_out/mo_ab2_$m.err; echo "wsm $m rc=$?"
done
RCM_BLOCK=32,2 RCM_MO_WSM=1 timeout 200 python bench.py --workload C5 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/mo_ab2_b2.json 2> gpurun_out/mo_ab2_b2.err; echo "wsm 1 block 32x2 rc=$?"
python - <<PY
import json
for n in ("1","0","b2"):
    try:
        d=json.loads(open(f"gpurun_out/mo_ab2_{n}.json").read().strip().splitlines()[-1])
        pk=d["roofline"]["per_kernel"]
        print("wsm",n,"ms/step",round(d["ms_per_step"],3)," ".join(f"{k}={v['ms_per_step']:.3f}" for k,v in pk.items()))
    except Exception as e: print(n,"ERR",e)
PY

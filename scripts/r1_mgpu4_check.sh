# 4-GPU check of the MOLOCH decomposition (2x2 tiles): parity test, then one strong-scaling bench line of C5.
mkdir -p gpurun_out
( timeout 120 python -m pytest tests/test_gpu_multirank.py -x -q -k "moloch and 4-layout2" ) > gpurun_out/mg4_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/mg4_pytest.log; tail -3 gpurun_out/mg4_pytest.log
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 5 --warmup 3 --workload C5 --no-cpu --no-e2e > gpurun_out/mg4_bench_c5.json 2> gpurun_out/mg4_bench_c5.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/mg4_bench_c5.json").read().strip().splitlines()[-1])
    pk=d["roofline"]["per_kernel"]
    print("C5 4 GPUs ms/step",round(d["ms_per_step"],3),"value %.4g"%d["value"]," ".join(f"{k}={v['ms_per_step']:.3f}" for k,v in pk.items()))
except Exception as e: print("ERR",e)
PY

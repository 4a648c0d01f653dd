# 2-GPU check: multi-rank parity of the MOLOCH core (product library, NVLink/NCCL transport) and of the
# non-hydrostatic core after this round's ABI changes, then a strong-scaling bench line of the MOLOCH workload C5.
mkdir -p gpurun_out
( timeout 400 python -m pytest tests/test_gpu_multirank.py -x -q ) > gpurun_out/mg2_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/mg2_pytest.log; tail -5 gpurun_out/mg2_pytest.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --workload C5 --no-cpu --no-e2e > gpurun_out/mg2_bench_c5.json 2> gpurun_out/mg2_bench_c5.err; echo "bench rc=$?"; tail -3 gpurun_out/mg2_bench_c5.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/mg2_bench_c5.json").read().strip().splitlines()[-1])
    pk=d["roofline"]["per_kernel"]
    print("C5 2 GPUs ms/step",round(d["ms_per_step"],3),"value %.4g"%d["value"]," ".join(f"{k}={v['ms_per_step']:.3f}" for k,v in pk.items()))
except Exception as e: print("ERR",e)
PY

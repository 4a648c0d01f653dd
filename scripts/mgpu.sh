# multi-GPU check: NCCL/P2P halo parity tests, then the bench under torchrun.  NS="2 4" TRANSPORTS="p2p nccl"
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/mgpu_smi.txt
if [ -z "$NOTEST" ]; then
timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q > gpurun_out/mgpu_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/mgpu_pytest.log
tail -5 gpurun_out/mgpu_pytest.log
fi
for tr in ${TRANSPORTS:-p2p}; do
for n in $NS; do
RCM_HALO=$tr timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps ${STEPS:-10} --warmup 3 --no-cpu ${BENCHARGS} > gpurun_out/bench_${tr}_n$n.json 2> gpurun_out/bench_${tr}_n$n.err; echo "bench $tr n=$n rc=$?"
grep -v "OMP_NUM_THREADS\|^\*\*\*\|^$" gpurun_out/bench_${tr}_n$n.err | tail -5
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${tr}_n$n.json").read().strip().splitlines()[-1])
    pk=d["roofline"]["per_kernel"]
    print("$tr n=$n ms/step", round(d["ms_per_step"],3), "G/s", round(d["value"]/1e9,3), "e2e", d["e2e"] and round(d["e2e"]["value"]/1e9,3), "halo ms", pk.get("halo",{}).get("ms_per_step"), "kernels", round(sum(v["ms_per_step"] for k,v in pk.items() if k!="halo"),3))
except Exception as e: print("no json", e)
PY
done
done

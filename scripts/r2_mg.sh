# multi-GPU call: N = number of GPUs of the box; hardware parity of the N-rank cases, then the strong-scaling bench lines
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
sel="-k ${N}-"
[ $N = 2 ] && sel=""
timeout 1200 python -m pytest tests/test_gpu_multirank.py -x -q $sel > gpurun_out/mg${N}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/mg${N}_pytest.log
for w in C4 C5 C4hq; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --workload $w --steps 20 --warmup 3 --no-cpu > gpurun_out/mg${N}_$w.json 2> gpurun_out/mg${N}_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/mg${N}_$w.json").read().strip().splitlines()[-1])
    print("N=$N %-5s ms/step %.3f e2e %.3f value %.4g" % ("$w", d['ms_per_step'], d['e2e']['ms_per_step'], d['value']), {k:round(x["ms_per_step"],3) for k,x in d["roofline"]["per_kernel"].items() if x["ms_per_step"] > 0.05})
except Exception as e:
    print("N=$N $w failed", e); print(open("gpurun_out/mg${N}_$w.err").read()[-1500:])
PY
done

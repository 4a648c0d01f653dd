# where does the e2e overhead at N ranks come from: both transfers / no staging / no export / neither
N=${1:-8}
mkdir -p gpurun_out
for p in both nostage noexport nostage,noexport; do
  RCM_E2E_PROBE=$p timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --no-cpu > gpurun_out/e2ep_n${N}_$p.json 2> gpurun_out/e2ep.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/e2ep_n${N}_$p.json").read().strip().splitlines()[-1])
print("N=$N %-18s ms/step %.3f e2e %.3f" % ("$p", d['ms_per_step'], d['e2e']['ms_per_step']))
PY
done
nvidia-smi topo -m 2>/dev/null | head -14

# e2e leg with the staging path warmed, default step count, N ranks (C4 only)
N=${1:-4}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --no-cpu > gpurun_out/e2e_n${N}.json 2> gpurun_out/e2e_n${N}.err
python - <<PY
import json
d=json.loads(open("gpurun_out/e2e_n${N}.json").read().strip().splitlines()[-1])
print("N=$N C4 steps %d ms/step %.3f e2e %.3f" % (d['steps'], d['ms_per_step'], d['e2e']['ms_per_step']), d['clocks'])
PY

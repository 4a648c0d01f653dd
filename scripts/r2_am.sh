# call AM (2 GPUs): the driver's launch of the bench at N = 2 on the final build, both arms
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2am_n2.json 2> gpurun_out/r2am_n2.err; echo "own arm rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/r2am_ref_n2.json 2> gpurun_out/r2am_ref_n2.err; echo "reference arm rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r2am_n2.json").read().strip().splitlines()[-1])
print("N=2: ms/step %.3f value %.4g e2e %.3f halo %.3f clk %s" % (d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['per_kernel'].get('halo',{}).get('ms_per_step',-1), d['clocks']['sm_mhz']))
r=json.loads(open("gpurun_out/r2am_ref_n2.json").read().strip().splitlines()[-1])
print("ref:", r.get('impl'), r.get('value'), r.get('cpu_baseline',{}).get('cores'))
PY

# call AI: idiffu = 3 on the GPU, the diffusion / option tests around it, and a C4 line (k_ppw_tma spill check)
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_idiffu2.py tests/test_gpu_parity.py -m gpu -q > gpurun_out/r2ai_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2ai_pytest.log
timeout 200 python bench.py --workload C4 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2ai_c4.json 2> gpurun_out/r2ai_c4.err; tail -c 1500 gpurun_out/r2ai_c4.json
timeout 120 python bench.py --workload C4 --steps 10 --warmup 3 --no-cpu --no-e2e --opt idiffu=3 > gpurun_out/r2ai_c4_d3.json 2> gpurun_out/r2ai_c4_d3.err; python - <<PY
import json
for n in ("r2ai_c4","r2ai_c4_d3"):
    try:
        d=json.loads(open("gpurun_out/%s.json"%n).read().strip().splitlines()[-1]); print(n, d['ms_per_step'], d['clocks']['sm_mhz'], {k: round(v['ms_per_step'],3) for k,v in d['roofline']['per_kernel'].items() if v['ms_per_step']>0.5})
    except Exception as e: print(n, 'ERR', e)
PY

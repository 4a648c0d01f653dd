# call X: k_xkc beside k_prep, k_snd_init beside the slow kernels (RCM_OVERLAP2): A/B on C4 and C4hq, parity tests
mkdir -p gpurun_out
for w in C4 C4hq; do for o in 0 1; do
  RCM_OVERLAP2=$o timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/x_${w}_$o.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/x_${w}_$o.json").read().strip().splitlines()[-1])
print("%-5s overlap2=$o ms/step %.3f" % ("$w", d['ms_per_step']))
PY
done; done
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_hydro.py tests/test_gpu_physics.py tests/test_gpu_restart.py -x -q > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r2x_pytest.log

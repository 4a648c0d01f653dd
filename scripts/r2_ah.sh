# call AH: MOLOCH physics coupling on the GPU; the MOLOCH suite and C5 after the step driver was split
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_physics.py tests/test_gpu_moloch.py -q > gpurun_out/r2ah_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2ah_pytest.log
timeout 400 python bench.py --workload C5 --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/ah_C5.json 2>/dev/null
python - <<PY
import json
d=json.loads(open("gpurun_out/ah_C5.json").read().strip().splitlines()[-1])
print("C5 ms/step %.3f value %.4g" % (d['ms_per_step'], d['value']))
PY

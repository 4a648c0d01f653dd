# Last 1-GPU check of the round: MOLOCH parity tests after the decomposition changes, C5 bench line (value, e2e, cpu_baseline).
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_moloch.py -x -q ) > gpurun_out/last_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/last_pytest.log; tail -3 gpurun_out/last_pytest.log
timeout 300 python bench.py --workload C5 --steps 20 --warmup 3 > gpurun_out/last_bench_c5.json 2> gpurun_out/last_bench_c5.err; echo "c5 rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/last_bench_c5.json").read().strip().splitlines()[-1])
print("C5 ms/step",round(d["ms_per_step"],3),"value %.4g"%d["value"],"e2e %.4g"%d["e2e"]["value"],"cpu %.4g"%d["cpu_baseline"]["value"],"launches",d["gpu_launches"])
PY

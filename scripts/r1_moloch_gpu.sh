# One GPU call for the MOLOCH core (budget-bound): parity tests first, then bench lines
# (C5 = BASELINE configs[4]), smoke, launch list, one ncu --set full capture of one advection step.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version --format=csv,noheader > gpurun_out/mo_gpu.txt 2>&1
( time timeout 420 python -m pytest tests/test_gpu_moloch.py -x -q ) > gpurun_out/mo_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/mo_pytest.log; tail -6 gpurun_out/mo_pytest.log
timeout 300 python bench.py --workload C5 --steps 20 --warmup 3 > gpurun_out/mo_bench_c5.json 2> gpurun_out/mo_bench_c5.err; echo "c5 rc=$?"; tail -3 gpurun_out/mo_bench_c5.err
timeout 120 python bench.py --workload smallm --steps 50 --warmup 3 --no-cpu > gpurun_out/mo_bench_small.json 2> gpurun_out/mo_bench_small.err; echo "small rc=$?"
timeout 150 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/mo_smoke.log 2>&1; tail -3 gpurun_out/mo_smoke.log
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/mo_launches_c5q.csv python bench.py --workload C5q --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/mo_ncu_c5q.log 2>&1; echo "ncu list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_mo_" -s 89 -c 26 -o gpurun_out/mo_c5_full -f python bench.py --workload C5 --steps 1 --warmup 3 --no-e2e --no-cpu > gpurun_out/mo_ncu_c5.log 2>&1; echo "ncu full rc=$?"
python - <<PY
import json
for n in ("c5","small"):
    try:
        d=json.loads(open(f"gpurun_out/mo_bench_{n}.json").read().strip().splitlines()[-1])
        print(n, "value %.4g"%d["value"], "ms/step", round(d["ms_per_step"],4), "e2e", d.get("e2e") and "%.4g"%d["e2e"]["value"], "cpu", d.get("cpu_baseline") and "%.4g"%d["cpu_baseline"]["value"], "launches", d.get("gpu_launches"))
        for k,v in d["roofline"]["per_kernel"].items(): print("   ", k, round(v["ms_per_step"],4), v["launches_per_step"], v["GBs"] and round(v["GBs"]))
    except Exception as e: print(n, "ERR", e)
PY

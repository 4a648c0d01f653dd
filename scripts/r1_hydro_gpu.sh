# One GPU call for the hydrostatic core (budget-bound): parity tests first, then
# bench lines, launch list, NH regression subset, one ncu --set full capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version --format=csv,noheader > gpurun_out/hy_gpu.txt 2>&1
( time timeout 420 python -m pytest tests/test_gpu_hydro.py -x -q ) > gpurun_out/hy_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/hy_pytest.log; tail -4 gpurun_out/hy_pytest.log
timeout 240 python bench.py --workload C1 --steps 100 --warmup 5 > gpurun_out/hy_bench_c1.json 2> gpurun_out/hy_bench_c1.err; echo "c1 rc=$?"
timeout 300 python bench.py --workload C4hq --steps 10 --warmup 3 --no-cpu --no-e2e > gpurun_out/hy_bench_c4h.json 2> gpurun_out/hy_bench_c4h.err; echo "c4h rc=$?"
( time timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "one_step or parity_vs_oracle or golden or staged" ) > gpurun_out/nh_pytest_subset.log 2>&1; echo "rc=$?" >> gpurun_out/nh_pytest_subset.log; tail -3 gpurun_out/nh_pytest_subset.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/hy_smoke.log 2>&1; tail -2 gpurun_out/hy_smoke.log
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/hy_launches_c1.csv python bench.py --workload C1 --steps 3 --warmup 3 --no-e2e --no-cpu > gpurun_out/hy_ncu_c1.log 2>&1; echo "ncu c1 rc=$?"
timeout 240 ncu --set full --clock-control none --import-source on -k regex:"k_h_tq|k_h_uv|k_h_prep|k_sp_proj" -s 12 -c 4 -o gpurun_out/hy_c4h_full -f python bench.py --workload C4hq --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/hy_ncu_c4h.log 2>&1; echo "ncu c4h rc=$?"
python - <<PY
import json
for n in ("c1","c4h"):
    try:
        d=json.loads(open(f"gpurun_out/hy_bench_{n}.json").read().strip().splitlines()[-1])
        print(n, "value %.4g"%d["value"], "ms/step", round(d["ms_per_step"],4), "e2e", d.get("e2e") and "%.4g"%d["e2e"]["value"], "cpu", d.get("cpu_baseline") and "%.4g"%d["cpu_baseline"]["value"], "launches", d.get("gpu_launches"))
        for k,v in d["roofline"]["per_kernel"].items(): print("   ", k, round(v["ms_per_step"],4), v["GBs"] and round(v["GBs"]))
    except Exception as e: print(n, "ERR", e)
PY

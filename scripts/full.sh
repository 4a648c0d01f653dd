mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
( time python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; tail -4 gpurun_out/bench_default.err
( time python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -4 gpurun_out/bench_reference.err
python bench.py --workload C3 --steps 40 --warmup 5 --no-cpu > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python bench.py --workload C2 --steps 40 --warmup 5 --no-cpu > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python - <<PY
import json
for n in ("default","reference","c3","c2"):
    try:
        d=json.loads(open(f"gpurun_out/bench_{n}.json").read().strip().splitlines()[-1])
        print(n, "value %.4g"%d["value"], "ms/step", round(d["ms_per_step"],3), "e2e", d.get("e2e") and "%.4g"%d["e2e"]["value"], "cpu", d.get("cpu_baseline") and "%.4g"%d["cpu_baseline"]["value"], "launches", d.get("gpu_launches"))
    except Exception as e: print(n, "ERR", e)
PY

# call AL: bench lines of the other workloads on the final build; launch list of the default command under ncu (after it ran clean)
mkdir -p gpurun_out
for W in C4hq C5 C1; do
  timeout 200 python bench.py --workload $W --steps 20 --warmup 3 --no-cpu > gpurun_out/r2al_$W.json 2> gpurun_out/r2al_$W.err
done
python - <<PY
import json
for w in ("C4hq","C5","C1"):
    try:
        d=json.loads(open("gpurun_out/r2al_%s.json"%w).read().strip().splitlines()[-1]); print(w, "ms/step %.3f e2e %.3f top %s %.3f clk %s" % (d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac'], d['clocks']['sm_mhz']))
    except Exception as e: print(w, 'ERR', e)
PY
timeout 120 python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2al_plain.json 2> gpurun_out/r2al_plain.err && \
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2al_launches_c4.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/r2al_ncu.log 2>&1
wc -l gpurun_out/r2al_launches_c4.csv

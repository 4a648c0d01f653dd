mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_d.json 2> gpurun_out/plain_d.err || exit 1
RCM_OVERLAP=0 ncu --set full --clock-control none --import-source on \
  --kernel-name 'regex:k_' --launch-skip 60 --launch-count 20 \
  -f -o gpurun_out/prof_r1d python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_d.log 2>&1
tail -2 gpurun_out/ncu_d.log | cut -c1-200
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 60 --launch-count 200 --csv --log-file gpurun_out/launches_r1d.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_d2.log 2>&1
wc -l gpurun_out/launches_r1d.csv

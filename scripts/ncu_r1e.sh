# ONE ncu pass per gpurun call: MODE=full (ncu --set full of one step) or MODE=list (launch list)
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_e.json 2> gpurun_out/plain_e.err || exit 1
if [ "$MODE" = full ]; then
RCM_OVERLAP=0 ncu --set full --clock-control none --import-source on \
  --kernel-name 'regex:k_' --launch-skip 60 --launch-count 21 \
  -f -o gpurun_out/prof_r1e python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_e.log 2>&1
tail -2 gpurun_out/ncu_e.log | cut -c1-200
else
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 60 --launch-count 200 --csv --log-file gpurun_out/launches_r1e.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_e2.log 2>&1
wc -l gpurun_out/launches_r1e.csv
fi

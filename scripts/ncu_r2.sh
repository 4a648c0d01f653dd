# round 2 profiling evidence, ONE ncu pass per gpurun call (each only after the same command exited 0 without ncu):
#   MODE=list : launch list (gpu__time_duration.sum) of 2 timed steps            -> gpurun_out/launches_r2.csv
#   MODE=full : ncu --set full of the kernels of ONE step, slow kernels serialised -> gpurun_out/prof_r2_raw.csv (+ details);
#               the report itself is too large to travel (gpurun_out is capped at 64 MiB), so the csv pages are
#               exported on the box; a second, small report holds one launch of the dominant kernel with its source
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$CMD > gpurun_out/plain_r2.json 2> gpurun_out/plain_r2.err || { echo "plain run failed"; tail -5 gpurun_out/plain_r2.err; exit 1; }
if [ "$MODE" = full ]; then
RCM_OVERLAP=0 ncu --set full --metrics smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum --clock-control none --kernel-name 'regex:k_' --launch-skip ${SKIP:-92} --launch-count ${COUNT:-23} \
  -f -o /tmp/prof_r2 $CMD > gpurun_out/ncu_r2_full.log 2>&1
tail -2 gpurun_out/ncu_r2_full.log | cut -c1-200; ls -la /tmp/prof_r2.ncu-rep
ncu -i /tmp/prof_r2.ncu-rep --page raw --csv > gpurun_out/prof_r2_raw.csv 2> gpurun_out/prof_r2_raw.err
ncu -i /tmp/prof_r2.ncu-rep --page details --csv > gpurun_out/prof_r2_details.csv 2>> gpurun_out/prof_r2_raw.err
RCM_OVERLAP=0 ncu --set full --import-source on --clock-control none --kernel-name 'regex:k_snd_wfwd' --launch-skip 17 --launch-count 1 \
  -f -o gpurun_out/prof_r2_wfwd $CMD > gpurun_out/ncu_r2_wfwd.log 2>&1
ls -la gpurun_out/
else
ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip ${SKIP:-60} --launch-count 200 --csv --log-file gpurun_out/launches_r2.csv $CMD > gpurun_out/ncu_r2_list.log 2>&1
wc -l gpurun_out/launches_r2.csv
fi

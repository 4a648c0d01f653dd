# round 2 profiling evidence for the other two cores: ncu --set full of the kernels of ONE step of C4hq (hydrostatic) and
# C5 (MOLOCH), each only after the same command exited 0 without ncu; the csv page is exported on the box
mkdir -p gpurun_out
for spec in "C4hq 24" "C5 81"; do
  set -- $spec; W=$1; PER=$2
  CMD="python bench.py --workload $W --steps 2 --warmup 3 --no-e2e --no-cpu"
  $CMD > gpurun_out/plain_$W.json 2> gpurun_out/plain_$W.err || { echo "plain run $W failed"; tail -5 gpurun_out/plain_$W.err; continue; }
  RCM_OVERLAP=0 timeout 900 ncu --set full --metrics smsp__inst_executed_pipe_fp64.sum,smsp__inst_executed.sum --clock-control none --kernel-name 'regex:k_' \
    --launch-skip $((PER * 4)) --launch-count $PER -f -o /tmp/prof_$W $CMD > gpurun_out/ncu_$W.log 2>&1
  tail -1 gpurun_out/ncu_$W.log | cut -c1-160
  ncu -i /tmp/prof_$W.ncu-rep --page raw --csv > gpurun_out/prof_${W}_raw.csv 2> gpurun_out/prof_${W}_raw.err
  wc -l gpurun_out/prof_${W}_raw.csv
done

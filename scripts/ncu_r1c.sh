mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_c.json 2> gpurun_out/plain_c.err || exit 1
ncu --set full --clock-control none --import-source on \
  --kernel-name 'regex:k_snd_wback|k_snd_wfwd|k_uv|k_tq|k_ppw|k_prep' --launch-skip 24 --launch-count 12 \
  -f -o gpurun_out/prof_r1c python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_c.log | cut -c1-200

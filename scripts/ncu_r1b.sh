# plain run first, then one ncu --set full capture of one step's heavy kernels
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/plain_b.json 2> gpurun_out/plain_b.err || exit 1
ncu --set full --clock-control none --import-source on \
  --kernel-name 'regex:k_snd_wback|k_snd_wfwd|k_uv|k_tq|k_ppw' --launch-skip 22 --launch-count 11 \
  -f -o gpurun_out/prof_r1b python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_b.log 2>&1
tail -3 gpurun_out/ncu_b.log

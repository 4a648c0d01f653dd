# call AB (2 GPUs): the round-2 options (TKE, kpbl, tracer, idiffu = 2) on two GPUs against one GPU, both sigma cores
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -q -k "uw_idiffu2" > gpurun_out/r2ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2ab_pytest.log

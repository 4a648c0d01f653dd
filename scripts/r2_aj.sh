# call AJ (2 GPUs): semi-Lagrangian advection on a decomposed domain; the round-2 options test again (wider allocation path untouched for them)
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_multirank.py -m gpu -q -k "sladvec or uw_idiffu2" > gpurun_out/r2aj_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2aj_pytest.log

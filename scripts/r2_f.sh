mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_async.py -x -q ) > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2f_pytest.log
python scripts/e2e_probe.py C4 > gpurun_out/r2f_probe.txt 2>&1; cat gpurun_out/r2f_probe.txt
CUDA_DEVICE_MAX_CONNECTIONS=32 python scripts/e2e_probe.py C4 > gpurun_out/r2f_probe32.txt 2>&1; echo "--- 32 connections"; cat gpurun_out/r2f_probe32.txt

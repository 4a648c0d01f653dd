# call Z: the UW-PBL GPU tests after the ind = 3 comparison was made step-synchronous; the suites that follow them in -x order
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tke.py tests/test_gpu_tracers.py tests/test_gpu_restart.py -q > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2z_pytest.log

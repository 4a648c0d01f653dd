# call N: tracer GPU tests, then the wfwd software-pipelining / occupancy variants on C4
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tracers.py tests/test_gpu_hydro.py -x -q > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2n_pytest.log
VARIANTS="pipe2 pipe3 mb4 mb2" bash scripts/r2_variants.sh

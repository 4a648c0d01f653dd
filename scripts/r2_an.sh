# call AN: the file-driven time loop on the GPU library
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_ncio.py -m gpu -q > gpurun_out/r2an_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2an_pytest.log

# call L: whole GPU suite (incl. iboudy 2/3/4 variants), default bench line, then the round-2 launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2l_pytest.log
timeout 600 python bench.py > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; echo "bench rc=$?"; cat gpurun_out/r2l_bench.json | cut -c1-1500
MODE=list bash scripts/ncu_r2.sh

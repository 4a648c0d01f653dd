# call U: final GPU suite (incl. restart tests), smoke(), the default bench line and the reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2u_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2u_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2u_smoke.log
timeout 900 python bench.py > gpurun_out/r2u_bench.json 2> gpurun_out/r2u_bench.err; echo "bench rc=$?"; cut -c1-700 gpurun_out/r2u_bench.json
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2u_ref.json 2> gpurun_out/r2u_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/r2u_ref.json

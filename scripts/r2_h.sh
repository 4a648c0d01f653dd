# round 2, call H (2 GPUs): multi-rank parity (NH + MOLOCH + negative-q seam), physics coupling tests, C4 bench at N=2
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_physics.py -q ) > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2h_pytest.log
NOTEST=1 NS="2" STEPS=20 bash scripts/mgpu.sh
cp gpurun_out/bench_p2p_n2.json gpurun_out/r2h_bench_c4_n2.json 2>/dev/null

# round 2, call D: the allocation-free entry point: alone in a fresh process, then after the whole GPU suite
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q ) > gpurun_out/r2d_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/r2d_alone.log
( timeout 1800 python -m pytest tests -m gpu -x -q ) > gpurun_out/r2d_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/r2d_all.log

mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -3
for T in 0 1; do echo "== WFWD_TMA $T"; RCM_WFWD_TMA=$T RCM_OVERLAP=0 timeout 300 bash scripts/gq.sh | grep -E "^[0-9]|wfwd|wback"; done
echo "== overlap on"; timeout 300 bash scripts/gq.sh | grep -E "^[0-9]"

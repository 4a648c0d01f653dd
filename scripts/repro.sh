echo "== current"; RCM_PFD=2 bash scripts/gq.sh | grep -E "^[0-9]|k_uv|k_tq|k_ppw|wfwd|wback"
echo "== current again"; RCM_PFD=2 bash scripts/gq.sh | grep -E "^[0-9]|k_uv|k_tq|k_ppw|wfwd|wback"
echo "== minb2 (older code)"; RCM_PFD=2 RCM_DYN_LIB=$PWD/regcm_b200/csrc/variants/libregcm_dyn_minb2.so bash scripts/gq.sh | grep -E "^[0-9]|k_uv|k_tq|k_ppw|wfwd|wback"

for V in ${VARS}; do echo "== VAR $V"; RCM_DYN_LIB=$PWD/regcm_b200/csrc/variants/libregcm_dyn_$V.so bash scripts/gq.sh | grep -E "^[0-9]|k_prep|k_xkc|k_uv|k_tq|k_ppw|wfwd|wback"; done

for P in ${PFDS:-0 2}; do echo "== PFD $P"; RCM_PFD=$P bash scripts/gq.sh | grep -E "^[0-9]|k_uv|k_tq|k_ppw|wfwd|wback"; done

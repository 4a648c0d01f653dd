for P in 0 2 3; do echo "== PFD $P"; RCM_PFD=$P bash scripts/gq.sh | grep -E "^[0-9]|wfwd|wback"; done

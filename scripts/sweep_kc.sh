for K in 100 21 11 6 3 1; do echo "== KC_UV $K"; RCM_KC_UV=$K bash scripts/gq.sh | grep -E "^[0-9]|k_uv"; done

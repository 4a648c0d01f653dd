for B in 32,4 64,2 128,1 32,8 64,4 128,2 256,1 32,2 64,1 32,1; do echo "== BLOCK $B"; RCM_BLOCK=$B bash scripts/gq.sh; done

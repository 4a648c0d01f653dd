mkdir -p gpurun_out
for V in b4s3m3 b4s2m4 b4s4m3 b8s3m1 b4s3m4; do echo "== $V"; RCM_DYN_LIB=$PWD/regcm_b200/csrc/variants/libregcm_dyn_$V.so RCM_OVERLAP=0 timeout 300 bash scripts/gq.sh | grep -E "^[0-9]|wfwd"; done

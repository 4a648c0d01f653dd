# call T: tensor-map TMA retry (scripts/tma_tensor_retry.cu); hydrostatic register-cap variants on C4hq
mkdir -p gpurun_out
./scripts/tma_tensor_retry > gpurun_out/tma_tensor_retry.log 2>&1; echo "tma retry rc=$?"; cat gpurun_out/tma_tensor_retry.log
WL=C4hq VARIANTS="htq2 huv3 huv2" bash scripts/r2_variants.sh

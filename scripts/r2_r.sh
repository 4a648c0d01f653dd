# call R: shared-reciprocal (Markstein) divisions in k_snd_wfwd: A/B on C4 and the NH parity tests with the variant library
mkdir -p gpurun_out
VARIANTS="divs" bash scripts/r2_variants.sh
RCM_DYN_LIB=$PWD/regcm_b200/csrc/variants/libregcm_dyn_divs.so timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -k "not c5 and not headline" > gpurun_out/r2r_pytest.log 2>&1; echo "pytest(divs) rc=$?"; tail -3 gpurun_out/r2r_pytest.log

# call O: register-cap variants (wfwd MB 2 = "mb2" is the new baseline candidate) and the L2 prefetch distance
mkdir -p gpurun_out
VARIANTS="mb2 wb3 pr3 st1" bash scripts/r2_variants.sh
for pfd in 1 3 4; do
  RCM_PFD=$pfd RCM_DYN_LIB=$PWD/regcm_b200/csrc/variants/libregcm_dyn_mb2.so timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_pfd$pfd.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/var_pfd$pfd.json").read().strip().splitlines()[-1])
print("mb2 pfd=$pfd ms/step %.3f" % d['ms_per_step'], {k:round(x["ms_per_step"],3) for k,x in d["roofline"]["per_kernel"].items() if x["ms_per_step"] > 0.25})
PY
done

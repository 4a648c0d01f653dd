# A/B of register caps (variant libraries built with scripts/build_variant.sh) on the C5 MOLOCH workload
mkdir -p gpurun_out
for v in base mow5 mow6 wafz4; do
  lib=regcm_b200/csrc/variants/libregcm_dyn_$v.so
  [ $v = base ] && lib=regcm_b200/csrc/libregcm_dyn.so
  RCM_DYN_LIB=$PWD/$lib timeout 200 python bench.py --workload C5 --steps 5 --warmup 3 --no-cpu --no-e2e > gpurun_out/mo_ab3_$v.json 2> gpurun_out/mo_ab3_$v.err; echo "$v rc=$?"
done
python - <<PY
import json
for n in ("base","mow5","mow6","wafz4"):
    try:
        d=json.loads(open(f"gpurun_out/mo_ab3_{n}.json").read().strip().splitlines()[-1])
        pk=d["roofline"]["per_kernel"]
        print(n,"ms/step",round(d["ms_per_step"],3)," ".join(f"{k}={v['ms_per_step']:.3f}" for k,v in pk.items()))
    except Exception as e: print(n,"ERR",e)
PY

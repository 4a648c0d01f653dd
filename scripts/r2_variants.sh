# A/B of library variants built by scripts/build_variant.sh (regcm_b200/csrc/variants/): C4 bench per variant
mkdir -p gpurun_out
for v in base $VARIANTS; do
  if [ $v = base ]; then lib=""; else lib=$PWD/regcm_b200/csrc/variants/libregcm_dyn_$v.so; fi
  RCM_DYN_LIB=$lib timeout 300 python bench.py --workload ${WL:-C4} --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/var_$v.json 2> gpurun_out/var_$v.err; echo "$v rc=$?"
done
python - <<PY
import json, os
for v in ["base"] + os.environ.get("VARIANTS","").split():
    try:
        d=json.loads(open(f"gpurun_out/var_{v}.json").read().strip().splitlines()[-1])
        print(f"{v:10s} ms/step {d['ms_per_step']:.3f}", {k:round(x["ms_per_step"],3) for k,x in d["roofline"]["per_kernel"].items() if x["ms_per_step"] > 0.25})
    except Exception as e: print(v,"failed",e)
PY

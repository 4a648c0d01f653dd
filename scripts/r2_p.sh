# call P: prefetch distances per kernel (C4), hydro kernels with L2 prefetch (C4hq), MOLOCH (C5)
mkdir -p gpurun_out
run() { # name, workload, env...
  n=$1; w=$2; shift 2
  env "$@" timeout 300 python bench.py --workload $w --steps 20 --warmup 3 --no-cpu --no-e2e > gpurun_out/p_$n.json 2>/dev/null
  python - <<PY
import json
d=json.loads(open("gpurun_out/p_$n.json").read().strip().splitlines()[-1])
print("%-14s %-5s ms/step %.3f" % ("$n", "$w", d['ms_per_step']), {k:round(x["ms_per_step"],3) for k,x in d["roofline"]["per_kernel"].items() if x["ms_per_step"] > 0.2})
PY
}
run c4_w4b1 C4 A=1
run c4_w5b1 C4 RCM_PFD_WFWD=5
run c4_w6b1 C4 RCM_PFD_WFWD=6
run c4_w8b1 C4 RCM_PFD_WFWD=8
run c4_w6b0 C4 RCM_PFD_WFWD=6 RCM_PFD_WBACK=0
run c4_w4b1_p4 C4 RCM_PFD=4 RCM_PFD_WFWD=4 RCM_PFD_WBACK=1
run hq_pfd0 C4hq RCM_PFD=0
run hq_pfd2 C4hq A=1
run hq_pfd4 C4hq RCM_PFD=4
run hq_pfd6 C4hq RCM_PFD=6
run c5_pfd0 C5 RCM_PFD=0
run c5_pfd2 C5 A=1
run c5_pfd4 C5 RCM_PFD=4

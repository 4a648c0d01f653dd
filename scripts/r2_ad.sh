# call AD: what the round-2 options cost on C4 (10 timed steps each, device-resident) and ipgf = 1 on C4hq
mkdir -p gpurun_out
run() { tag=$1; w=$2; shift 2; opts=""; for o in "$@"; do opts="$opts --opt $o"; done
  timeout 500 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu --no-e2e $opts > gpurun_out/ad_$tag.json 2> gpurun_out/ad_$tag.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/ad_$tag.json").read().strip().splitlines()[-1])
    pk=d["roofline"]["per_kernel"]
    print("%-12s ms/step %.3f" % ("$tag", d['ms_per_step']), {k:round(x["ms_per_step"],3) for k,x in pk.items() if k in ("k_tr","k_tq","k_uv","k_ppw","k_h_tq","k_h_uv","k_h_prep","k_bdy","k_negq")})
except Exception as e:
    print("$tag failed", e); print(open("gpurun_out/ad_$tag.err").read()[-800:])
PY
}
run base C4
run ntr4 C4 ntr=4
run tke C4 ibltyp=2
run sl C4 isladvec=1
run idiffu2 C4 idiffu=2
run all C4 ntr=4 ibltyp=2 iuwvadv=1 idiffu=2 isladvec=1
run hq_base C4hq
run hq_ipgf1 C4hq ipgf=1

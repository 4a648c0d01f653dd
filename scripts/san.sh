mkdir -p gpurun_out
cat > /tmp/t.py <<'PY'
import sys; sys.path.insert(0,'.')
from regcm_b200 import capi, synth
case=synth.make_case("small"); core=capi.DynCore(capi.cuda_lib(), case.cfg); synth.load_case(core, case)
try:
    core.step(1); print("step ok")
except Exception as e: print("ERR", e)
PY
timeout 600 compute-sanitizer --tool memcheck --print-limit 5 python /tmp/t.py > gpurun_out/san.log 2>&1
grep -v "^=========     at\|Host Frame\|in /usr\|python" gpurun_out/san.log | head -40

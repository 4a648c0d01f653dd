#!/bin/bash
# Tuning aid: build a variant of the library with extra nvcc flags into
# regcm_b200/csrc/variants/libregcm_dyn_<name>.so (select it with RCM_DYN_LIB=<path>).
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d /tmp/rcmvar.XXXX)
mkdir -p $tmp/regcm_b200/csrc $tmp/include $root/regcm_b200/csrc/variants
cp $root/regcm_b200/csrc/*.cu $root/regcm_b200/csrc/*.hpp $root/regcm_b200/csrc/*.h $root/regcm_b200/csrc/*.cuh $root/regcm_b200/csrc/*.inc $root/regcm_b200/csrc/Makefile $tmp/regcm_b200/csrc/
cp $root/include/*.h $tmp/include/
make -s -C $tmp/regcm_b200/csrc -j8 NVFLAGS="-O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math -Xptxas -v $*" >/dev/null
python3 - $tmp/regcm_b200/csrc <<'PY'
import glob, re, sys
for p in sorted(glob.glob(sys.argv[1] + "/*.ptxas.log")):
    t = open(p).read()
    for m in re.finditer(r"Compiling entry function '_Z\d+(\w+?)\d+Geom.*?\n.*?\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores.*?\n.*?Used (\d+) registers", t):
        print(f"{m.group(1):14s} regs={m.group(4):>3s} spill={m.group(3)}", end=" | ")
print()
PY
cp $tmp/regcm_b200/csrc/libregcm_dyn.so $root/regcm_b200/csrc/variants/libregcm_dyn_$name.so
rm -rf $tmp

#!/bin/bash
# builds libfeelgpu.so with different entity-plan geometries on the GPU box and prints the kernel time of each
mkdir -p gpurun_out
run() {
  name=$1; shift
  make clean > /dev/null 2>&1
  make -j16 EPLAN_DEFS="$*" > gpurun_out/build_$name.log 2>&1 || { echo "$name: build failed"; tail -5 gpurun_out/build_$name.log; return; }
  grep -A2 "k_eplanILi2ELi1ELi1" feelpp_b200/lib/kernels_eplan.ptxas.log | grep -i "registers" | head -1
  python bench.py --steps 10 --warmup 3 --algo eplan --no-e2e --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_$name.json")); r=d["roofline"]; e=d["config"]["setup"]["plan"]["eplan"]
    print("$name [$*]: kernel_ms %.4f frac %.4f ctas %d visits %d padded %d"%(r["kernel_ms"],r["frac"],e["ctas"],e["visits"],e["iterations_padded"]))
except Exception as ex:
    print("$name failed", ex); print(open("gpurun_out/bench_$name.err").read()[-600:])
PY
}
"$@"
# reference: the committed sources kept under gpurun_ref/ (same box, same call)
runref() {
  mkdir -p /tmp/cur && cp feelpp_b200/csrc/kernels_eplan.cu feelpp_b200/csrc/eplan_build.cpp feelpp_b200/csrc/eplan.h /tmp/cur/
  cp gpurun_ref/* feelpp_b200/csrc/
  run "$@"
  cp /tmp/cur/* feelpp_b200/csrc/
}

#!/bin/bash
# occupancy sweep of the fused LV kernel: rebuild with __launch_bounds__(128, mb) and time bench.py
run() { python fet-ode_b200/build.py > /dev/null 2>&1; echo "$1 $(python bench.py --steps 2 --warmup 3 | python -c 'import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d["ms_per_step"], d["roofline"]["frac"])')"; }
setmb() { sed -i "s/__launch_bounds__(SN_THREADS, [0-9]*) smallnet_rk_kernel/__launch_bounds__(SN_THREADS, $1) smallnet_rk_kernel/" fet-ode_b200/csrc/smallnet.cu; }
setmb 8; sed -i 's/^#ifndef FETODE_SN_SCALAR/#if 0/' fet-ode_b200/csrc/smallnet.cuh; run "mb=8 scalar"
sed -i 's/^#if 0/#ifndef FETODE_SN_SCALAR/' fet-ode_b200/csrc/smallnet.cuh
setmb 10; run "mb=10 packed"
setmb 12; run "mb=12 packed"

#!/bin/bash
# one-off instrumentation run: where does the MMA thread of the tcgen05 forward kernel wait?
sed -i 's|^#include <cstdio>|#define FETODE_TC_TIMING 1\n#include <cstdio>|' fet-ode_b200/csrc/kanlinear_tc.cu
python fet-ode_b200/build.py > /dev/null 2>&1
python scratch/tc_one.py 2>&1 | tail -6

#!/bin/bash
# occupancy sweep of the LV reverse-sweep kernel
for mb in 1 3 4 5; do
  sed -i "s/__launch_bounds__(SN_THREADS[, 0-9]*) smallnet_rk_bwd_kernel/__launch_bounds__(SN_THREADS, $mb) smallnet_rk_bwd_kernel/" fet-ode_b200/csrc/smallnet_bwd.cu
  python fet-ode_b200/build.py > /dev/null 2>&1
  echo "mb=$mb $(python -c "
import sys; sys.argv=['x','--skip-cpu']
import bench_configs as b
print(b.lv_train(65536, True)['gpu_fwd_bwd_opt_ms'])")"
done

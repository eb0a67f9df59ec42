for d in 0 2 4 6 8 14; do echo "DAU_DEBUG=$d"; DAU_DEBUG=$d timeout 100 python tools/quick_time.py --iters 2 2>&1 | grep -E "^   bwd" | tail -1; done
ncu --set full --clock-control none --import-source on -k regex:dau_grad_tma -c 1 -o gpurun_out/grad_v3 python tools/quick_time.py --iters 1 > gpurun_out/ncu_grad.log 2>&1

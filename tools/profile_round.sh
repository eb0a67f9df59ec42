#!/bin/bash
# One GPU call: bench line, launch list, full ncu capture of the three hot kernels (C2).
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'dau_(grad|gather)_tma' -s 3 -c 3 -o gpurun_out/hot \
    python tools/quick_time.py --iters 2 > gpurun_out/ncu_hot.log 2>&1
ncu --set full --clock-control none -k regex:'blur_pair|error_relayout' -s 4 -c 4 -o gpurun_out/blur \
    python tools/quick_time.py --iters 2 > gpurun_out/ncu_blur.log 2>&1
ls -la gpurun_out

#!/bin/bash
# One GPU call: bench line, reference arm, launch list, full ncu capture of the hot and helper kernels (C2),
# per-config kernel times, smoke().
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'dau_(grad|gather)_tma' -s 3 -c 3 -o gpurun_out/hot \
    python tools/quick_time.py --iters 2 > gpurun_out/ncu_hot.log 2>&1
ncu --set full --clock-control none -k regex:'blur_pair|error_relayout' -s 4 -c 4 -o gpurun_out/blur \
    python tools/quick_time.py --iters 2 > gpurun_out/ncu_blur.log 2>&1
for cfg in c1 c4 l2 l3 l4 l5; do
  python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$cfg.json 2> gpurun_out/bench_$cfg.err
done
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
ls -la gpurun_out | head -40

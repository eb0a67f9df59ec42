python -m pytest tests -m gpu -q -rx > gpurun_out/r2_pytest_gpu_final.log 2>&1; tail -3 gpurun_out/r2_pytest_gpu_final.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_c2_final.json 2> gpurun_out/r2_bench_c2_final.err
for c in c1 st; do python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline --no-alexnet > gpurun_out/r2_bench_${c}_final.json 2>/dev/null; done
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1

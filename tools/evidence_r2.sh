#!/bin/bash
# Round-2 evidence in one GPU call: bench lines (both arms, every config), ncu launch list of the bench command,
# north-star counters of C2 / L4 / L2 / C4, the reference-CUDA comparison table, the GPU test-suite, smoke().
T=${1:-r2}
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_c2_final.json 2> gpurun_out/${T}_bench_c2_final.err || exit 1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${T}_bench_c2_reference_arm.json 2> gpurun_out/${T}_bench_ref.err
for c in c1 st l2 l3 l4 l5; do
  python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-alexnet > gpurun_out/${T}_bench_${c}_final.json 2> gpurun_out/${T}_bench_${c}_final.err
done
python bench.py --config c4 --steps 5 --warmup 3 --no-cpu-baseline --no-alexnet --no-reference-cuda > gpurun_out/${T}_bench_c4_final.json 2> gpurun_out/${T}_bench_c4_final.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-alexnet --no-reference-cuda > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_c2.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-alexnet --no-reference-cuda > gpurun_out/${T}_ncu_launches.log 2>&1
bash tools/ncu_northstar.sh c2
bash tools/ncu_northstar.sh l4 --N 256 --S 384 --F 384 --G 2 --H 13 --W 13
bash tools/ncu_northstar.sh l2 --N 256 --S 96 --F 256 --G 2 --H 27 --W 27
bash tools/ncu_northstar.sh c4 --N 256 --S 512 --F 512 --G 2 --H 28 --W 28 --ks 17
if [ -f oracle/_ref/libdau_ref_cuda.so ]; then
python tests/ref_cuda_time.py --config c1 c2 l2 l3 l4 l5 > gpurun_out/${T}_reference_cuda_vs_this_repo.jsonl 2> gpurun_out/${T}_refcuda.err
python tests/ref_cuda_time.py --config c4 --images 64 >> gpurun_out/${T}_reference_cuda_vs_this_repo.jsonl 2>> gpurun_out/${T}_refcuda.err
fi
python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu_final.log 2>&1; tail -3 gpurun_out/${T}_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
python tools/alexnet_breakdown.py --batch 32 --rows 45 > gpurun_out/${T}_alexnet_breakdown_b32.txt 2>&1
python tools/alexnet_breakdown.py --batch 256 --rows 45 > gpurun_out/${T}_alexnet_breakdown_b256.txt 2>&1

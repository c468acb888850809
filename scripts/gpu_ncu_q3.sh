set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
python -m pytest tests/test_gpu_vector.py -q -m gpu 2>&1 | tail -3
DEG=${DEG:-3}; ND=${ND:-300000000}; TAG=${TAG:-v2}
python bench.py --degree $DEG --ndofs $ND --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stiffness -s 3 -c 1 -o gpurun_out/prof_q${DEG}_$TAG -f python bench.py --degree $DEG --ndofs $ND --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/plain_$TAG.log | cut -c1-300

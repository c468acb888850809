# ncu --set full of the fp32 action kernels at Q3 and Q6 (qmode 1), one launch each.
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
for DEG in 3 6; do
  ARGS="--degree $DEG --ndofs 200000000 --float 32 --qmode 1 --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg"
  timeout 200 python bench.py $ARGS > gpurun_out/plain_f32_q${DEG}.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:stiffness -s 3 -c 1 \
    -o gpurun_out/prof_f32_q${DEG} -f python bench.py $ARGS > gpurun_out/ncu_full_f32_q${DEG}.log 2>&1
  echo "q$DEG rc=$?"
done
tail -1 gpurun_out/plain_f32_q3.log; tail -1 gpurun_out/plain_f32_q6.log

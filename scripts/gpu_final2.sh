# Whole GPU suite + smoke(), then ncu --set full of the shipped pencil kernels (Q7 fp64, Q6 fp32, Q7 fp32).
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
timeout 400 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | cut -c1-400 | tee gpurun_out/final_pytest.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -3 | tee gpurun_out/final_smoke.txt
run() { # tag degree float
  ARGS="--degree $2 --ndofs 100000000 --float $3 --qmode 1 --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg"
  timeout 100 python bench.py $ARGS > gpurun_out/plain_pencil_final_$1.log 2>&1 && \
  timeout 150 ncu --set full --clock-control none --import-source on -k regex:stiffness -s 3 -c 1 \
    -o gpurun_out/prof_pencil_final_$1 -f python bench.py $ARGS > gpurun_out/ncu_pencil_final_$1.log 2>&1
  echo "$1 rc=$?"
}
run q7f64 7 64
run q6f32 6 32
run q7f32 7 32

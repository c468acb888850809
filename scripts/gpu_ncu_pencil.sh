# ncu --set full of the pencil kernel: Q6 fp32 (one cell per CTA) and Q7 fp64.
set -x
mkdir -p gpurun_out
export BD_STIFFNESS_VARIANT=pencil
run() { # tag cpb degree float
  ARGS="--degree $3 --ndofs 100000000 --float $4 --qmode 1 --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg"
  BD_PENCIL_CPB=$2 timeout 120 python bench.py $ARGS > gpurun_out/plain_pencil_$1.log 2>&1 && \
  BD_PENCIL_CPB=$2 timeout 200 ncu --set full --clock-control none --import-source on -k regex:stiffness -s 3 -c 1 \
    -o gpurun_out/prof_pencil_$1 -f python bench.py $ARGS > gpurun_out/ncu_pencil_$1.log 2>&1
  echo "$1 rc=$?"
}
run q6f32 2 6 32
run q7f64 0 7 64

set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
DEG=${DEG:-3}; ND=${ND:-300000000}; TAG=${TAG:-v5}; FL=${FL:-64}
QM=${QM:-1}
ARGS="--degree $DEG --ndofs $ND --float $FL --qmode $QM --steps 2 --warmup 3 --no-cpu --no-e2e"
python bench.py $ARGS > gpurun_out/plain_q${DEG}_$TAG.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_q${DEG}_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_list_q${DEG}_$TAG.log 2>&1
echo "list rc=$?"
python bench.py $ARGS --no-cg > gpurun_out/plain2_q${DEG}_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stiffness -s 3 -c 1 -o gpurun_out/prof_q${DEG}_$TAG -f python bench.py $ARGS --no-cg > gpurun_out/ncu_full_q${DEG}_$TAG.log 2>&1
echo "full rc=$?"

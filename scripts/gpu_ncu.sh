# Launch list + `ncu --set full` of the top kernel for one configuration, each only after the same command has run
# plain (a number printed under ncu is never a bench value).
#   DEG=3 ND=300000000 QM=1 FL=64 GEO=auto TAG=r2 KREGEX=stiffness bash scripts/gpu_ncu.sh
set -x
mkdir -p gpurun_out
DEG=${DEG:-3}; ND=${ND:-300000000}; TAG=${TAG:-r2}; FL=${FL:-64}; QM=${QM:-1}; GEO=${GEO:-auto}; KREGEX=${KREGEX:-stiffness}
ARGS="--degree $DEG --ndofs $ND --float $FL --qmode $QM --geometry $GEO --steps 2 --warmup 3 --no-cpu --no-e2e --no-variants --no-q6 --no-parity --no-reference-gpu-kernel"
python bench.py $ARGS > gpurun_out/${TAG}_plain_q${DEG}.json 2> gpurun_out/${TAG}_plain_q${DEG}.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_q${DEG}.csv python bench.py $ARGS > gpurun_out/${TAG}_ncu_list_q${DEG}.log 2>&1
echo "list rc=$?"
python bench.py $ARGS --no-cg > gpurun_out/${TAG}_plain2_q${DEG}.json 2> gpurun_out/${TAG}_plain2_q${DEG}.err && \
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 3 -c 1 -o gpurun_out/${TAG}_prof_q${DEG} -f python bench.py $ARGS --no-cg > gpurun_out/${TAG}_ncu_full_q${DEG}.log 2>&1
echo "full rc=$?"

set -x
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_default.json; tail -5 gpurun_out/bench_default.err
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:stiffness_kernel -s 3 -c 1 -o gpurun_out/prof_q3_r1 python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-cg > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out

ARGS="--no-cpu --no-e2e --no-cg --no-variants --no-q6 --no-parity --no-reference-gpu-kernel --steps 2 --warmup 3 --geometry onthefly"
ncu --set full --clock-control none --import-source on -k regex:affine -s 3 -c 1 -o gpurun_out/r2_prof_q3_affine -f python bench.py $ARGS > gpurun_out/r2_ncu_q3_affine.log 2>&1; echo "ncu rc=$?"

set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
python -m pytest tests -x -q -m gpu 2>&1 | tail -8

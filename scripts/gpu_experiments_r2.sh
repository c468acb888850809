# The experiments prepared at the end of round 1, in one GPU call (about 5 minutes):
# parity of every opt-in variant first (each under a timeout: two of them wait on mbarriers and have never
# run on a GPU), then the A/B runs.  Results: gpurun_out/experiments_r2.txt
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
BD_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_experimental.py tests/test_gpu_pencil_bulk.py tests/test_gpu_slab_f2.py -q 2>&1 | tail -12 | cut -c1-600 | tee gpurun_out/experiments_r2_parity.txt
one() { # label "ENV=.. ENV=.." degree qmode float ndofs
  env $2 timeout 150 python bench.py --degree $3 --qmode $4 --float $5 --ndofs $6 --no-cpu --no-e2e --no-cg --steps 10 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT %-28s P=$3 q=$4 f$5 n=$6: %.2f GDoF/s frac %.3f kernel_ms %.3f' % ('$1', d['value'], r['frac'], r['kernel_ms_per_apply']))
"
}
{
  one default            "BD_X=0"               3 1 32 500000000
  one slab_f2            "BD_SLAB_F2=1"         3 1 32 500000000
  one default            "BD_X=0"               3 1 32 1000000000
  one slab_f2            "BD_SLAB_F2=1"         3 1 32 1000000000
  one default            "BD_X=0"               6 1 32 500000000
  one pencil_bulk        "BD_PENCIL_BULK=1"     6 1 32 500000000
  one pencil_bulk+ffma2  "BD_PENCIL_BULK=1 BD_PENCIL_PAIRED=1" 6 1 32 500000000
  one default            "BD_X=0"               6 1 64 500000000
  one mma8_value_prefetch "BD_MMA8_PREFETCH=2"  6 1 64 500000000
  one default            "BD_X=0"               7 0 64 300000000
  one mma8_value_prefetch "BD_MMA8_PREFETCH=2"  7 0 64 300000000
  one default            "BD_X=0"               7 1 64 300000000
  one pencil_ring5       "BD_PENCIL_RING=3"     7 1 64 300000000
} 2>&1 | grep RESULT | tee gpurun_out/experiments_r2.txt

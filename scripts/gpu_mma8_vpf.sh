# First run of the DMMA kernel with value prefetch (BD_MMA8_PREFETCH=2): parity, then A/B at Q6 / Q7q0 / Q5 fp64.
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
BD_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_experimental.py -x -q -k dmma 2>&1 | tail -8 | cut -c1-600 || exit 1
one() { # prefetch degree qmode ndofs
  BD_MMA8_PREFETCH=$1 timeout 150 python bench.py --degree $2 --qmode $3 --ndofs $4 --no-cpu --no-e2e --steps 10 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT prefetch=$1 P=$2 q=$3 n=$4 f64: %.2f GDoF/s frac %.3f kernel_ms %.3f cg %.2f' % (d['value'], r['frac'], r['kernel_ms_per_apply'], (d.get('cg') or {}).get('value', 0)))
"
}
for cfg in "6 1 500000000" "7 0 300000000" "5 1 300000000" "6 0 300000000"; do
  set -- $cfg
  one 1 $1 $2 $3
  one 2 $1 $2 $3
done 2>&1 | grep RESULT | tee gpurun_out/mma8_vpf_ab.txt

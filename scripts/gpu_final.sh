# A/B of the pencil defaults against the previous defaults, then the whole GPU suite and smoke().
set -x
mkdir -p gpurun_out
make -C benchmark_dolfinx_b200 -j 16 2>&1 | tail -1
one() { # variant ring degree ndofs float
  BD_STIFFNESS_VARIANT=$1 BD_PENCIL_RING=$2 timeout 150 python bench.py --degree $3 --ndofs $4 --float $5 --qmode 1 --no-cpu --no-e2e --no-cg --steps 8 --warmup 3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT variant=$1 ring=$2 P=$3 n=$4 f=$5: %.2f GDoF/s frac %.3f kernel_ms %.3f | %s' % (d['value'], r['frac'], r['kernel_ms_per_apply'], r['kernel'][:40]))
"
}
for cfg in "7 300000000 64" "6 500000000 32" "7 300000000 32" "5 300000000 32"; do
  set -- $cfg
  one tma 0 $1 $2 $3
  one default 0 $1 $2 $3
  one default 1 $1 $2 $3
  one default 2 $1 $2 $3
done 2>&1 | grep RESULT | tee gpurun_out/pencil_ab3.txt
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -6 | cut -c1-400 | tee gpurun_out/final_pytest.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -3 | tee gpurun_out/final_smoke.txt

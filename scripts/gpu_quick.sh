# quick timing of a few configurations: CFGS="degree:qmode:float:ndofs:geometry ..." bash scripts/gpu_quick.sh
mkdir -p gpurun_out
for cfg in $CFGS; do
IFS=: read d q f n g <<< "$cfg"
timeout 300 python bench.py --degree $d --qmode $q --float $f --ndofs $n --geometry $g $EXTRA --no-cpu --no-e2e --no-cg --no-variants --no-q6 --no-parity --no-reference-gpu-kernel --steps 10 --warmup 3 2>gpurun_out/quick_err.txt | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); r = d['roofline']
        print('RESULT $TAG P=$d q=$q f$f n=$n %-16s: %.2f GDoF/s %s frac %.3f kernel_ms %.3f ynorm %.15e' % (r['geometry'], d['value'], r['bound'], r['frac'], r['kernel_ms_per_apply'], d['y_norm']))
" | tee -a gpurun_out/quick_$TAG.txt
tail -n 2 gpurun_out/quick_err.txt | cut -c1-300
done

B=" --steps 3 --warmup 3 --no-cpu --plonk-log-n 0 --bls12-377-log-n 0"
run() { name=$1; shift; env "$@" python bench.py $B > gpurun_out/ab_$name.json 2> gpurun_out/ab_$name.err; python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/ab_$name.json') if l.startswith('{')][-1]); s=d['sub_metrics']['stage_ms']
print('$name', round(d['value']*1e3,2), round(d['e2e']['value']*1e3,2), d['sub_metrics']['verified'], ' '.join('%s=%.1f'%(k[:-3],v) for k,v in s.items()))"; }

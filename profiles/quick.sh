# Usage: profiles/quick.sh  -> GPU parity tests + two bench regimes
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -4 gpurun_out/pytest_gpu.log
for w in 60 5; do
python bench.py --steps 20 --warmup $w --no-cpu $* 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('warmup $w', 'ms/step %.3f'%d['ms_per_step'], 'Mws/s %.3f'%(d['value']/1e6), 'e2e %.3f'%(d['e2e']['value']/1e6), d['config']['launch'])
    elif 'rror' in l: print(l.strip())
"
done

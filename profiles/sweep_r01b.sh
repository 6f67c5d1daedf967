python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -5 gpurun_out/pytest_gpu.log
for cfg in "64 1" "64 2" "64 4" "64 6" "64 8" "32 8" "32 16" "96 4" "128 2"; do set -- $cfg
  python bench.py --steps 20 --warmup 60 --no-cpu --threads $1 --worlds-per-block $2 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('tpw $1 wpb $2', 'ms/step %.3f'%d['ms_per_step'], 'Mws/s %.3f'%(d['value']/1e6), d['config']['launch'])
    elif 'rror' in l: print(l.strip())
"
done

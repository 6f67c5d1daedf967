for cfg in "1 64 4 0" "0 64 4 0" "0 64 8 0" "0 64 10 0" "0 64 2 0" "0 32 10 0" "0 32 15 0" "1 64 8 1" "0 64 4 3"; do set -- $cfg
  MDR_LOCKSTEP=$1 MDR_MIN_BLOCKS=$4 python bench.py --steps 20 --warmup 60 --no-cpu --threads $2 --worlds-per-block $3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('lockstep $1 tpw $2 wpb $3 minb $4', 'ms/step %.3f'%d['ms_per_step'], 'Mws/s %.3f'%(d['value']/1e6), d['config']['launch'])
    elif 'rror' in l: print(l.strip())
"
done

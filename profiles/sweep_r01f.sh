# wider world groups (no tail passes): lockstep tpw wpb minblocks
for cfg in "1 64 4 0" "1 96 4 0" "1 96 5 0" "1 96 2 0" "1 128 2 0" "1 128 4 0" "0 64 4 0" "0 96 4 0" "1 96 3 2"; do set -- $cfg
  MDR_LOCKSTEP=$1 MDR_MIN_BLOCKS=$4 python bench.py --steps 20 --warmup 60 --no-cpu --threads $2 --worlds-per-block $3 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('lockstep $1 tpw $2 wpb $3 minb $4', 'ms/step %.3f'%d['ms_per_step'], 'Mws/s %.3f'%(d['value']/1e6), d['config']['launch'])
    elif 'rror' in l or 'overflow' in l: print(l.strip()[:200])
"
done

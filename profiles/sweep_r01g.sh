# independent blocks per SM vs worlds per block: lockstep tpw wpb minblocks smem_contacts max_contacts
for cfg in "1 64 2 0 0 640" "1 64 1 0 0 640" "1 64 3 0 0 640" "1 64 2 0 48 160" "1 64 1 0 48 160"; do set -- $cfg
  MDR_LOCKSTEP=$1 MDR_MIN_BLOCKS=$4 python bench.py --steps 20 --warmup 60 --no-cpu --threads $2 --worlds-per-block $3 --smem-contacts $5 --max-contacts $6 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('lockstep $1 tpw $2 wpb $3 minb $4 smemc $5 maxc $6', 'ms/step %.3f'%d['ms_per_step'], 'Mws/s %.3f'%(d['value']/1e6), d['config']['launch'])
    elif 'rror' in l or 'overflow' in l: print(l.strip()[:200])
"
done

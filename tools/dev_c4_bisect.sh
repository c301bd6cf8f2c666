B='--steps 5 --warmup 3 --no-strong --no-cpu-baseline'
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d.get('extra',{}); print('$1', 'c4 us/pos', e.get('config4_us_per_position'), 'c3', e.get('config3_ms_per_iteration'), 'nm', e.get('config3_nm_ms_per_iteration'))"; }
python bench.py $B 2>/dev/null | show default
UAVMVS_CULL_EARLY_AT=0 python bench.py $B 2>/dev/null | show early_off
python bench.py --config 4 --positions 65536 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('standalone --config 4: ms/step', d['ms_per_step'], 'value', d['value'], d.get('config',{}).get('workload','')[:80])"

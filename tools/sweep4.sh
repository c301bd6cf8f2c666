B='--steps 20 --warmup 5 --no-cpu-baseline --no-strong --no-extras'
show() { python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['roofline']['k1_kernels_ms']; print('$1  step %.3f e2e %.3f K1 %.3f cull %.3f trace %.3f retire %.3f parity %s' % (d['ms_per_step'], d['e2e']['ms_per_step'], d['stage_ms']['K1'], k['k_cull_emit_q'], k['k_trace'], k['k_retire'], d['parity_check']['ok']))"; }
for r in 20 12 8 4; do UAVMVS_CULL_SETUP=1 UAVMVS_K1_REFILL=$r python bench.py $B 2>/dev/null | show "cull-setup r$r "; done
for r in 24 28; do UAVMVS_K1_REFILL=$r python bench.py $B 2>/dev/null | show "default    r$r "; done

# A/B runs of environment-switched variants on one box: tools/probe/ab.sh "VAR=val VAR2=val" "..." ...
for v in "$@"; do
env $v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-dense 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
l=d['roofline']['labelled_kernels_ms_per_step']
print('$v', d['ms_per_step'], d['state_sha256'][:12], {k: round(x, 2) for k, x in l.items() if k in ('nh_p_grad_winds', 'dzd_spline', 'update_dz_c')}, d.get('acoustic_substep', {}).get('ms_per_substep'))"
done

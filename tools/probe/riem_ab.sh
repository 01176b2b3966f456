# A/B of the Riemann-solver variants on one box (PACE_B200_RIEM = tmem | smem, PACE_B200_RIEM_U = levels per chunk)
timeout 300 python -m pytest tests/test_modules_gpu.py -m gpu -x -q -k "riem" 2>&1 | tail -2
timeout 600 python -m pytest tests/test_engine_c48.py tests/test_engine_acoustic.py -m gpu -x -q 2>&1 | tail -2
for v in "tmem 2" "tmem 3" "smem 2"; do
set -- $v
PACE_B200_RIEM=$1 PACE_B200_RIEM_U=$2 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-dense 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
l=d['roofline']['labelled_kernels_ms_per_step']
print('$1 U=$2', d['ms_per_step'], d['state_sha256'][:12], 'riem_c', l['riem_solver_c'], 'riem3', l['riem_solver3'])"
done

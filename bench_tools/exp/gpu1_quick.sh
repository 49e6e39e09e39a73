set -x
for m in 1 0 1 0; do
DPS_MASK_X=$m timeout 120 python bench.py --steps 300 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('mask_x=$m', d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
done

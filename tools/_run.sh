timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('%.0f' % d['value'], '%.4f' % d['ms_per_step'], d['stage_ms'])"

python bench.py --steps 40 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['stage_ms'], d['sparse_regime']['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1l_launches.csv python tools/prof_chain.py --steps 3 > /dev/null 2>&1
python tools/launches.py gpurun_out/r1l_launches.csv 60 | tail -12

for g in 2 3 4 5 6 8; do
  python bench.py --steps 100 --warmup 4 --min-seconds 1 --groups $g --skip-cpu-baseline --skip-closed-loop --skip-extras > gpurun_out/sw.json 2> gpurun_out/sw.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/sw.json').read().strip().splitlines()[-1])
print('groups $g', round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']))
"
done

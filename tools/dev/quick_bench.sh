python -m pytest tests/test_gpu_track.py -x -q 2>&1 | tail -2
python bench.py --steps 100 --warmup 4 --min-seconds 1 --skip-cpu-baseline --skip-extras --skip-e2e --skip-closed-loop > gpurun_out/q.json 2> gpurun_out/q.err
python -c "
import json,sys
d=json.loads(open('gpurun_out/q.json').read().strip().splitlines()[-1])
k=d['kernels']
print(round(d['value']), round(d['ms_per_step'],4), 'lk', round(k['lk_track_fwd']['ms_per_step'],4), round(k['lk_track_bwd']['ms_per_step'],4), 'cs', round(k['corner_score_nms']['ms_per_step'],4))
"

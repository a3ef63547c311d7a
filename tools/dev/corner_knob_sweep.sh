run() {
  python bench.py --steps 100 --warmup 4 --min-seconds 1 --skip-cpu-baseline --skip-e2e --skip-closed-loop --skip-extras > gpurun_out/sw.json 2> gpurun_out/sw.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/sw.json').read().strip().splitlines()[-1])
k=d['kernels']
print('$1', round(d['value']), round(d['ms_per_step'],4), 'lk', round(k['lk_track_fwd']['ms_per_step'],4), round(k['lk_track_bwd']['ms_per_step'],4), 'cs', round(k['corner_score_nms']['ms_per_step'],4))
"
}
run base
SVO_CORNER_MIN_BLOCKS=20 run mb20
SVO_CORNER_CPT=1 run cpt1
SVO_CORNER_CPT=1 SVO_CORNER_MIN_BLOCKS=20 run cpt1_mb20
SVO_CORNER_SEGMENTS=4 run seg4
SVO_CORNER_SEGMENTS=5 run seg5
SVO_CORNER_SEGMENTS=8 run seg8
SVO_CORNER_SEGMENTS=8 SVO_CORNER_MIN_BLOCKS=20 run seg8_mb20

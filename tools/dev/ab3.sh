for v in head fix tmplfix fix tmplfix; do
  cp tools/dev/variants/libsvo_$v.so stereo_vo_b200/libsvo.so; cp tools/dev/variants/libsvo_host_$v.so stereo_vo_b200/libsvo_host.so; touch stereo_vo_b200/libsvo_host.so
  python bench.py --steps 100 --warmup 4 --min-seconds 1 --skip-cpu-baseline --skip-extras --skip-e2e --skip-closed-loop > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/ab_$v.json').read().strip().splitlines()[-1])
k=d['kernels']
print('$v', round(d['value']), round(d['ms_per_step'],4), 'lk', round(k['lk_track_fwd']['ms_per_step'],4), round(k['lk_track_bwd']['ms_per_step'],4), 'cs', round(k['corner_score_nms']['ms_per_step'],4))
"
done

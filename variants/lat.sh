for h in 0 1 0 1; do
  SVO_LK_HANDOVER=$h python bench.py --steps 50 --warmup 4 --min-seconds 0.3 --skip-cpu-baseline --skip-closed-loop --skip-e2e > gpurun_out/lat_$h.json 2> gpurun_out/lat_$h.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/lat_$h.json').read().strip().splitlines()[-1])
x=d['extras']
print('handover $h', round(d['value']), 'kitti', x['kitti_latency']['ms_per_frame_median'], x['kitti_latency']['ms_skip_frames'], x['kitti_latency']['ms_keyframes'], 'd435i', x['d435i_latency']['ms_per_frame_median'], 'uhd', x['uhd_stress']['frames_per_s'])
"
done

python -m pytest tests/test_gpu_track.py tests/test_gpu_pipeline.py tests/test_gpu_bench_path.py tests/test_gpu_host_cpp.py tests/test_gpu_golden.py tests/test_gpu_uhd_stress.py -x -q 2>&1 | tail -5
for h in 0 1; do
  SVO_LK_HANDOVER=$h python bench.py --steps 100 --warmup 4 --min-seconds 1 --skip-cpu-baseline --skip-closed-loop --skip-extras > gpurun_out/ho_$h.json 2> gpurun_out/ho_$h.err
  python -c "
import json,sys
d=json.loads(open('gpurun_out/ho_$h.json').read().strip().splitlines()[-1])
k=d['kernels']
print('handover $h', round(d['value']), round(d['ms_per_step'],4), 'e2e', round(d['e2e']['value']), 'lk', round(k['lk_track_fwd']['ms_per_step'],4), round(k['lk_track_bwd']['ms_per_step'],4), 'cs', round(k['corner_score_nms']['ms_per_step'],4))
"
done

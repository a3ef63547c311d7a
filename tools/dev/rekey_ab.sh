# svo_tracker_set re-keying the lk_track hand-over (SVO_LK_REKEY=1, side stream): the new test,
# the closed-loop leg with and without it on the same box, then the whole GPU suite with it on
mkdir -p gpurun_out
( timeout 60 python -m pytest tests/test_gpu_track.py -x -q -m gpu; echo "rc=$?" ) > gpurun_out/rk_track.log 2>&1
tail -2 gpurun_out/rk_track.log
for v in 1 0; do
  SVO_LK_REKEY=$v timeout 60 python bench.py --steps 100 --warmup 4 --min-seconds 2 --skip-cpu-baseline --skip-extras --skip-e2e > gpurun_out/rk2_$v.json 2>> gpurun_out/rk2.err
  python -c "
import json
d=json.loads(open('gpurun_out/rk2_$v.json').read().strip().splitlines()[-1])
print('side-stream rekey=$v', round(d['value']), round(d['closed_loop']['value']), d['closed_loop']['repeats'])
" | tee -a gpurun_out/rk2_ab.txt
done
( SVO_LK_REKEY=1 timeout 120 python -m pytest tests -x -q -m gpu; echo "rc=$?" ) > gpurun_out/rk_all_on.log 2>&1
tail -2 gpurun_out/rk_all_on.log

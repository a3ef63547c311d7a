# last GPU call of the round: the changed paths first, then the whole GPU suite, then a short
# bench with only the resident + closed-loop legs (the closed loop is what svo_tracker_set's
# re-keyed hand-over changes)
mkdir -p gpurun_out
( timeout 170 python -m pytest tests/test_gpu_track.py tests/test_gpu_pnp.py tests/test_gpu_host_cpp.py -x -q -m gpu; echo "rc=$?" ) > gpurun_out/f_changed.log 2>&1
tail -3 gpurun_out/f_changed.log
( timeout 330 python -m pytest tests -x -q -m gpu; echo "rc=$?" ) > gpurun_out/f_all.log 2>&1
tail -3 gpurun_out/f_all.log
timeout 150 python bench.py --steps 100 --warmup 4 --min-seconds 1 --skip-cpu-baseline --skip-extras --skip-e2e > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err
tail -c 600 gpurun_out/f_bench.json

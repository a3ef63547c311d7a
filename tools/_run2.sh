mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_stereo_dense.py -q > gpurun_out/dense_test.log 2>&1; echo "dense_test rc=$?" >> gpurun_out/dense_test.log
tail -6 gpurun_out/dense_test.log
timeout 200 python tools/dense_probe.py 64 20 > gpurun_out/dense_probe.json 2> gpurun_out/dense_probe.err; cat gpurun_out/dense_probe.json; tail -3 gpurun_out/dense_probe.err
timeout 200 python tools/dense_probe.py 1 20 > gpurun_out/dense_probe_1.json 2>> gpurun_out/dense_probe.err; cat gpurun_out/dense_probe_1.json
timeout 200 ncu --set full -k regex:stereo_bm_dense_kernel -s 65 -c 1 --clock-control none --import-source on -f -o gpurun_out/dense_full python tools/dense_probe.py 64 1 > gpurun_out/ncu_dense.log 2>&1; echo "ncu rc=$?"

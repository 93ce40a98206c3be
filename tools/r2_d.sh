cd /root/repo
timeout 600 python -m pytest tests/test_gpu_attack.py -q -x 2>&1 | tail -25 > gpurun_out/r2_gputest_attack.log
timeout 300 python tools/probe_refine2.py 300000 > gpurun_out/r2_probe_300k.jsonl 2> gpurun_out/r2_probe_300k.err
echo done

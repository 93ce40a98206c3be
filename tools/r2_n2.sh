cd /root/repo
timeout 600 python -m pytest tests/test_gpu_comm.py -q -x 2>&1 | tail -8 > gpurun_out/r2_gputest_comm.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err
echo "rc=$?" >> gpurun_out/r2_bench_n2.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --config 4 --steps 2 --warmup 3 > gpurun_out/r2_bench_cfg4_n2.json 2> gpurun_out/r2_bench_cfg4_n2.err
echo "rc=$?" >> gpurun_out/r2_bench_cfg4_n2.err
echo done

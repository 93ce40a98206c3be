cd /root/repo
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_n2_copy.json 2> gpurun_out/r2_bench_n2_copy.err
echo "rc=$?" >> gpurun_out/r2_bench_n2_copy.err
ALINK_FEED_GATHER=nccl python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29516 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_n2_nccl.json 2> gpurun_out/r2_bench_n2_nccl.err
echo "rc=$?" >> gpurun_out/r2_bench_n2_nccl.err
echo done

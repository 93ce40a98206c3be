cd /root/repo
o=gpurun_out/n8
mkdir -p $o
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 5 --warmup 3 --no-target > $o/bench_n8_copy.json 2> $o/bench_n8_copy.err
echo "rc=$?" >> $o/bench_n8_copy.err
ALINK_FEED_GATHER=nccl python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 5 --warmup 3 --no-target > $o/bench_n8_nccl.json 2> $o/bench_n8_nccl.err
echo "rc=$?" >> $o/bench_n8_nccl.err
echo done

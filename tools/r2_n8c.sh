cd /root/repo
o=gpurun_out/n8
mkdir -p $o
for cfg in "copy 2" "nccl 2" "nccl 8" "copy 0"; do
  set -- $cfg
  ALINK_FEED_GATHER=$1 ALINK_RESERVE_SMS=$2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 5 --warmup 3 --no-target > $o/bench_n8_$1_r$2.json 2> $o/bench_n8_$1_r$2.err
  echo "rc=$?" >> $o/bench_n8_$1_r$2.err
done
echo done

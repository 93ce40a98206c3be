cd /root/repo
o=gpurun_out/n8
mkdir -p $o
python bench.py --impl reference --gpus 8 --steps 3 --warmup 1 > $o/bench_ref_n8.json 2> $o/bench_ref_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 5 --warmup 3 > $o/bench_n8.json 2> $o/bench_n8.err
echo "rc=$?" >> $o/bench_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --config 4 --steps 3 --warmup 3 > $o/bench_cfg4_n8.json 2> $o/bench_cfg4_n8.err
echo "rc=$?" >> $o/bench_cfg4_n8.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 8 --config 5 --mode refine --steps 3 --warmup 3 > $o/bench_cfg5_n8.json 2> $o/bench_cfg5_n8.err
echo "rc=$?" >> $o/bench_cfg5_n8.err
echo done

N=$1
for C in 3 5; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$C tools/bench_configs.py --config $C 2>/dev/null | grep '^{' >> gpurun_out/bench_configs_n$N.log
done
cat gpurun_out/bench_configs_n$N.log | cut -c1-700

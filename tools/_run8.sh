for n in 8 4; do
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29500"
timeout 300 $TR bench.py --gpus $n --workload c5 --steps 20 --warmup 3 --no-e2e 2>&1 | grep "\"value\"" | tee gpurun_out/bench_c5_n$n.json
done

timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -k "bf16" 2>&1 | tail -3
for f in "" "--no-overlap"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus 2 --workload c5 --steps 20 --warmup 5 --no-e2e $f 2>&1 | grep "\"value\""
done

timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29500"
timeout 300 $TR bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e 2>&1 | grep "\"value\""
timeout 300 $TR bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e --batch 16384 2>&1 | grep "\"value\""

import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
import tensorflow_examples_b200 as pkg
from tensorflow_examples_b200 import synthetic
from oracle import tf1_graph as oracle
dims=[784,100,10]
for B, steps in ((1024, 6), (4096+40, 4), (30000, 2)):
    x, y = synthetic.make_pool(B*steps)
    Ws, bs = synthetic.make_weights(dims, False)
    g = oracle.TF1Graph(dims, 1, 1, Ws, bs)
    want = [g.train_step(x[s*B:(s+1)*B], y[s*B:(s+1)*B], 0.001) for s in range(steps)]
    for split in (True, False):
        tr = pkg.DenseTrainer(dims, 1, 1, max_local_batch=B)
        tr.tensor_core_fp32(split)
        tr.set_params(Ws, bs)
        lt, ct = tr.train_steps(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), B, steps, lr=0.001)
        torch.cuda.synchronize()
        W, b = tr.get_params()
        print(B, 'split3' if split else 'cuda-core', 'loss rel', max(abs(float(lt[s])-float(want[s][0]))/abs(float(want[s][0])) for s in range(steps)),
              'correct diff', max(abs(int(ct[s])-want[s][1]) for s in range(steps)),
              'W0 err', float(np.abs(W[0]-g.W[0]).max()), 'b0 err', float(np.abs(b[0]-g.b[0]).max()), 'W1 err', float(np.abs(W[1]-g.W[1]).max()))
        tr.close()

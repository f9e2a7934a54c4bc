'''
Drop-in for distributed/distributed.py of manigoswami/tensorflow-examples: the same 784-100-10
sigmoid / softmax network trained with plain gradient descent, the same four flags and the same
printed lines.  What changes underneath (BASELINE.json north_star):
  * the TF1 session graph           -> hand-written sm_100a CUDA kernels behind include/b2mlp.h
  * the parameter server            -> every worker keeps a replica on its own GPU; gradients are
                                       all-reduced (NCCL over NVLink) before each SGD apply, so the
                                       update is synchronous and `--job_name=ps` has nothing to serve
  * feed_dict / MNIST download      -> a device-resident pool: MNIST IDX files if they are already in
                                       --data_dir, else synthetic MNIST-shaped rows (no network here)

To run (as in the reference, distributed.py:11-13; one worker per GPU):

python distributed.py  --ps_hosts=127.0.0.1:2222  --worker_hosts=127.0.0.1:2223,127.0.0.1:2224  --job_name=ps --task_index=0
python distributed.py  --ps_hosts=127.0.0.1:2222  --worker_hosts=127.0.0.1:2223,127.0.0.1:2224  --job_name=worker --task_index=0
python distributed.py  --ps_hosts=127.0.0.1:2222  --worker_hosts=127.0.0.1:2223,127.0.0.1:2224  --job_name=worker --task_index=1

Reference lines are cited as distributed.py:N.
'''
import os
import sys
import time

import numpy as np
from absl import flags

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# flags for input defined here -- names, types, defaults and help text of distributed.py:24-32
flags.DEFINE_string("job_name", "", "Either 'ps' or 'worker'")
flags.DEFINE_integer("task_index", 0, "Index of task within the job")
flags.DEFINE_string("worker_hosts", None,
                    "The worker url list, separated by comma (e.g. tf-worker1:2222,1.2.3.4:2222)")
flags.DEFINE_string("ps_hosts", None,
                    "The ps url list, separated by comma (e.g. tf-ps2:2222,1.2.3.5:2222)")
# additions: the defaults reproduce the constants the reference hard-codes (distributed.py:46-49,134)
flags.DEFINE_integer("batch_size", 100, "rows per worker per step (distributed.py:46)")
flags.DEFINE_float("learning_rate", 0.001, "gradient-descent step size (distributed.py:47)")
flags.DEFINE_integer("training_epochs", 50, "passes over the training pool (distributed.py:48)")
flags.DEFINE_integer("frequency", 100, "print every this many steps (distributed.py:134)")
flags.DEFINE_string("data_dir", "MNIST_data", "directory that may hold the four MNIST IDX files (distributed.py:55)")
flags.DEFINE_integer("train_rows", 55000, "synthetic pool: training rows (mnist.train.num_examples)")
flags.DEFINE_integer("test_rows", 10000, "synthetic pool: test rows (mnist.test)")
flags.DEFINE_boolean("shuffle", True, "reshuffle the training pool every epoch, as next_batch does (distributed.py:146)")
flags.DEFINE_integer("shuffle_seed", 0, "seed of the per-epoch permutation (the reference's is unseeded)")
flags.DEFINE_string("dump_perm", "", "write the per-epoch row permutations the shuffle drew to this .npy ([epochs, rows] "
                    "int64), so that a run can be replayed elsewhere (next_batch's own shuffle is unseeded, distributed.py:146)")
flags.DEFINE_string("logs_path", "", "TensorBoard event directory; the reference hard-codes /tmp/mnist/ (distributed.py:49). "
                    "Empty (default) writes no event files")
flags.DEFINE_integer("async_period", 0, "k > 0: emulate the reference's asynchronous parameter-server updates -- every worker "
                     "applies its own gradients, replicas are averaged every k steps; 0 (default): synchronous all-reduce")
flags.DEFINE_boolean("sharded_apply", False, "several workers, NCCL exchange: every worker owns one slice of the variables (reduce-"
                     "scatter, apply on the slice, all-gather) -- what several ps tasks are in the reference (distributed.py:63-65)")
flags.DEFINE_boolean("peer_exchange", True, "workers on one NVLink box sum gradients through peer memory instead of NCCL")
flags.DEFINE_string("tf_op_seeds", "", "two integers 'a,b': draw W1 and W2 with TensorFlow 1.x's own generator for "
                    "tf.set_random_seed(2) (distributed.py:83) and these op seeds (the graph's op counter at "
                    "distributed.py:85 and :86, which depends on the TensorFlow version). Empty: fixed-seed numpy weights.")
flags.DEFINE_string("restore_path", "", "start from variables saved by --save_path instead of fresh ones")
flags.DEFINE_string("save_path", "", "chief worker: write the variables to this .npz when training ends "
                    "(the reference keeps no checkpoint: Supervisor(logdir=None), distributed.py:129-131)")
flags.DEFINE_boolean("strict_reference_bugs", False,
                     "reproduce distributed.py:165 literally: the 'Time Taken' line raises TypeError")
FLAGS = flags.FLAGS


def load_pools(torch, dev):
    """mnist = input_data.read_data_sets('MNIST_data', one_hot=True)   distributed.py:52-55"""
    from tensorflow_examples_b200 import mnist_idx, synthetic
    if mnist_idx.available(FLAGS.data_dir):
        d = mnist_idx.load(FLAGS.data_dir)
        (tx, ty), (ex, ey) = d["train"], d["test"]
    else:
        x, y = synthetic.make_pool(FLAGS.train_rows + FLAGS.test_rows)
        tx, ty, ex, ey = x[:FLAGS.train_rows], y[:FLAGS.train_rows], x[FLAGS.train_rows:], y[FLAGS.train_rows:]
    to = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return to(tx), to(ty), to(ex), to(ey)


def main(argv):
    FLAGS(argv)
    from tensorflow_examples_b200 import cluster as cl
    from tensorflow_examples_b200 import report, synthetic

    # cluster = tf.train.ClusterSpec({...}); server = tf.train.Server(...)      distributed.py:37-43
    cluster = cl.parse_cluster(FLAGS.ps_hosts, FLAGS.worker_hosts)
    role = cl.resolve_role(cluster, FLAGS.job_name, FLAGS.task_index)

    batch_size = FLAGS.batch_size            # distributed.py:46
    learning_rate = FLAGS.learning_rate      # distributed.py:47
    training_epochs = FLAGS.training_epochs  # distributed.py:48

    if role.job_name == "ps":
        # server.join()   distributed.py:57-58 -- the variables live on the workers' GPUs now
        print("ps task %d: parameters are replicated on the workers and all-reduced every step; "
              "nothing to serve" % role.task_index)
        return 0

    import torch
    import tensorflow_examples_b200 as b2
    ngpu = torch.cuda.device_count()
    if ngpu == 0:
        raise RuntimeError("distributed.py needs a CUDA device (B200); there is no CPU fallback")
    device_index = role.rank % ngpu
    torch.cuda.set_device(device_index)
    dev = torch.device("cuda", device_index)

    train_x, train_y, test_x, test_y = load_pools(torch, dev)
    num_examples = train_x.shape[0]

    # x [None,784], y_ [None,10]; W1 [784,100], W2 [100,10] ~ N(0,1) seed 2; b1, b2 zeros;
    # z2 = x@W1+b1; a2 = sigmoid(z2); z3 = a2@W2+b2; y = softmax(z3);
    # cross_entropy = mean(-sum(y_*log(y), 1)); GradientDescentOptimizer(learning_rate)   distributed.py:68-108
    dims = [784, 100, 10]
    model = b2.DenseTrainer(dims, b2.ACT_SIGMOID, b2.LOSS_SOFTMAX_CE, max_local_batch=max(batch_size, 4096),
                            device=device_index)
    if FLAGS.restore_path:
        model.restore(FLAGS.restore_path)
    elif FLAGS.tf_op_seeds:
        model.init_random_normal(2, [int(t) for t in FLAGS.tf_op_seeds.split(",")])   # distributed.py:83-91
    else:
        Ws, bs = synthetic.make_weights(dims, scaled=False)
        model.set_params(Ws, bs)
    print("Variables initialized ...")                                           # distributed.py:127

    # sv = tf.train.Supervisor(is_chief=(task_index == 0), ...); prepare_or_wait_for_session   distributed.py:129-135
    uid, store = cl.exchange_unique_id(role, b2.DenseTrainer.new_unique_id)
    model.init_comm(role.rank, role.world, uid)
    if role.world > 1 and FLAGS.peer_exchange:
        # replica_device_setter's ps round trip (distributed.py:63-65) -> gradients summed straight out of
        # the other workers' GPU memory over NVLink, fused with the SGD apply.  Only when every worker is a
        # process on THIS node with a GPU of its own; every rank publishes whether each stage worked for it
        # and all of them take the same branch: peer exchange everywhere, or the NCCL all-reduce everywhere.
        ok = cl.single_node(cluster) and role.world <= ngpu
        handle = b"-"                            # "no handle from me"
        if ok:
            try:
                handle = model.peer_export()
            except b2.B2Error:
                ok = False
        store.set("b2_peer_%d" % role.rank, handle)
        handles = [bytes(store.get("b2_peer_%d" % r)) for r in range(role.world)]
        ok = ok and all(len(h) == 64 for h in handles)
        if ok:
            try:
                model.peer_import(handles)
            except b2.B2Error:
                ok = False
        store.set("b2_peer_ok_%d" % role.rank, b"1" if ok else b"0")
        if not all(bytes(store.get("b2_peer_ok_%d" % r)) == b"1" for r in range(role.world)):
            model.use_peer_exchange(False)      # somebody could not map a peer: everyone stays on NCCL
    if role.world > 1 and FLAGS.sharded_apply:
        model.use_peer_exchange(False)
        model.sharded_apply(True)
    if role.world > 1 and FLAGS.async_period > 0:
        model.set_async_period(FLAGS.async_period)
    cl.store_barrier(store, role, "b2_comm_ready")

    # writer = tf.summary.FileWriter(logs_path, graph=...)   distributed.py:138 -- every worker writes its own file
    from tensorflow_examples_b200.summaries import SummaryWriter
    writer = SummaryWriter(FLAGS.logs_path)

    begin_time = time.time()
    frequency = FLAGS.frequency
    batch_count = cl.steps_per_epoch(num_examples, batch_size, role.world)       # distributed.py:142
    global_batch = batch_size * role.world
    gen = torch.Generator(device=dev)
    gen.manual_seed(FLAGS.shuffle_seed)
    cost = float("nan")
    perms = []
    start_time = time.time()
    for epoch in range(training_epochs):                                          # distributed.py:141
        if FLAGS.shuffle:
            # mnist.train.next_batch reshuffles every epoch; every worker draws the same permutation
            perm = torch.randperm(num_examples, device=dev, generator=gen)
            if FLAGS.dump_perm:
                perms.append(perm.cpu().numpy())
            pool_x, pool_y = train_x[perm].contiguous(), train_y[perm].contiguous()
        else:
            pool_x, pool_y = train_x, train_y
        i = 0
        while i < batch_count:                                                    # distributed.py:145
            # up to `frequency` sess.run([train_op, cross_entropy, summary_op, global_step]) calls per launch;
            # cost / accuracy of every step are written to a device trace              distributed.py:148-150
            n = min(frequency, batch_count - i)
            loss_trace, correct_trace = model.train_steps(pool_x, pool_y, batch_size, n,
                                                          first_row=i * global_batch, lr=learning_rate)
            i += n
            torch.cuda.synchronize(dev)
            cost = float(loss_trace[n - 1].item())
            step = model.global_step - np.float32(1)      # the value fetched with the last step (pre-increment)
            if writer.enabled:                            # writer.add_summary(summary, step) for every step   distributed.py:151
                lt, ct = loss_trace[:n].cpu().numpy(), correct_trace[:n].cpu().numpy()
                for k in range(n):
                    writer.add_scalars(int(step) - (n - 1) + k, lt[k], ct[k] / float(global_batch))
            elapsed_time = time.time() - start_time                               # distributed.py:155-156
            start_time = time.time()
            print(report.progress_line(step, epoch, i - 1, batch_count, cost, elapsed_time, frequency))

    # print("Acc: %2.2f" % sess.run(accuracy, feed_dict={x: mnist.test.images, y_: mnist.test.labels}))  distributed.py:164
    _, correct = model.evaluate(test_x, test_y)
    print(report.accuracy_line(correct / float(test_x.shape[0])))
    print(report.time_taken_line(time.time() - begin_time, FLAGS.strict_reference_bugs))   # distributed.py:165
    print(report.final_cost_line(cost))                                                    # distributed.py:166
    writer.close()
    if FLAGS.dump_perm and role.rank == 0:
        np.save(FLAGS.dump_perm, np.stack(perms) if perms else np.zeros((0, num_examples), np.int64))
    if FLAGS.save_path and role.rank == 0:
        model.save(FLAGS.save_path)
    cl.store_barrier(store, role, "b2_done")
    model.close()                                                                          # sv.stop()  distributed.py:168
    print("done with training")                                                            # distributed.py:169
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv))

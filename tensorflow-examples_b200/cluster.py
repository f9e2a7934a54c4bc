"""ps/worker flags -> data-parallel ranks.

Replaces tf.train.ClusterSpec / tf.train.Server / replica_device_setter / Supervisor
(distributed/distributed.py:37-43,63-65,129-135).  The reference places the variables on the `ps`
job and lets every worker pull and push them over gRPC, asynchronously.  Here every worker owns a
replica on its own GPU and gradients are all-reduced every step, so:
  worker task_index        -> rank
  len(worker_hosts)        -> world size
  ps job                   -> nothing to serve (the process reports that and exits)
  chief (task_index == 0)  -> creates the NCCL unique id; the others fetch it from a TCP store that
                              the chief opens on worker_hosts[0] (the reference's own host list: no new flags)
"""
import collections

ClusterSpec = collections.namedtuple("ClusterSpec", ["ps", "worker"])
Role = collections.namedtuple("Role", ["job_name", "task_index", "rank", "world", "is_chief", "store_host", "store_port"])


def parse_cluster(ps_hosts, worker_hosts):
    """distributed.py:37 -- `FLAGS.ps_hosts.split(",")`: a missing flag is an AttributeError there too."""
    return ClusterSpec(ps=ps_hosts.split(","), worker=worker_hosts.split(","))


def resolve_role(cluster, job_name, task_index):
    """distributed.py:41-43,57-65: which process am I, and what is my rank among the workers."""
    if job_name not in ("ps", "worker"):
        raise ValueError("job_name must be 'ps' or 'worker', got %r" % (job_name,))
    members = cluster.ps if job_name == "ps" else cluster.worker
    if not 0 <= task_index < len(members):
        raise ValueError("task_index %d is not in the %s job (%d tasks)" % (task_index, job_name, len(members)))
    host, _, port = cluster.worker[0].rpartition(":")
    world = len(cluster.worker)
    rank = task_index if job_name == "worker" else -1
    return Role(job_name, task_index, rank, world, job_name == "worker" and task_index == 0, host or "127.0.0.1",
                int(port) if port else 2223)


def single_node(cluster):
    """True when every worker address names the same host (peer-memory exchange needs all ranks on one box)."""
    hosts = {w.rpartition(":")[0] or w for w in cluster.worker}
    if len(hosts) == 1:
        return True
    local = {"127.0.0.1", "localhost", "::1", "0.0.0.0"}
    return hosts <= local


def exchange_unique_id(role, make_id, timeout_s=300):
    """Chief publishes the 128-byte communicator id on a TCP store at worker_hosts[0]; everyone reads it.

    Returns (id_bytes, store).  Keep `store` alive until every worker has read the id (a barrier key is
    provided by `store_barrier`).  With one worker nothing is exchanged.
    """
    if role.world == 1:
        return None, None
    import datetime
    from torch.distributed import TCPStore
    store = TCPStore(role.store_host, role.store_port, role.world, is_master=role.is_chief,
                     timeout=datetime.timedelta(seconds=timeout_s), wait_for_workers=False)
    if role.is_chief:
        store.set("b2_unique_id", bytes(make_id()))
    uid = bytes(store.get("b2_unique_id"))
    return uid, store


def store_barrier(store, role, name):
    """All workers reach this point before anyone proceeds (used before the chief closes the store)."""
    if store is None:
        return
    store.add(name, 1)
    import time
    t0 = time.time()
    while int(store.add(name, 0)) < role.world:
        time.sleep(0.01)
        if time.time() - t0 > 300:
            raise RuntimeError("timed out waiting for the other workers at %r" % name)


def steps_per_epoch(num_examples, batch_size, world):
    """distributed.py:142 batch_count = int(num_examples / batch_size); with synchronous data
    parallelism one step consumes batch_size rows on every worker."""
    return int(num_examples / (batch_size * world))

"""TensorBoard event files for the two scalars the reference logs every step
(distributed/distributed.py:121-125 tf.summary.scalar("cost"/"accuracy"), :138 FileWriter(logs_path), :151 add_summary).

Host-side only: the values come out of the device trace that b2_train_steps fills, once per print
window.  Uses tensorboard's own EventFileWriter when it is installed; without it the writer is a no-op
that says so once (the training step never depends on it)."""
import time


class SummaryWriter:
    def __init__(self, logs_path):
        self._w = None
        self._warned = False
        if not logs_path:
            return
        try:
            from tensorboard.compat.proto.event_pb2 import Event  # noqa: F401
            from tensorboard.summary.writer.event_file_writer import EventFileWriter
            self._w = EventFileWriter(logs_path)
        except Exception as e:  # tensorboard missing or unusable: keep training
            print("summaries disabled (%s)" % (repr(e)[:120],))
            self._warned = True

    @property
    def enabled(self):
        return self._w is not None

    def add_scalars(self, step, cost, accuracy):
        """One Event per step carrying both scalars, like the merged summary_op (distributed.py:125,149-151)."""
        if self._w is None:
            return
        from tensorboard.compat.proto.event_pb2 import Event
        from tensorboard.compat.proto.summary_pb2 import Summary
        summ = Summary(value=[Summary.Value(tag="cost", simple_value=float(cost)),
                              Summary.Value(tag="accuracy", simple_value=float(accuracy))])
        self._w.add_event(Event(wall_time=time.time(), step=int(step), summary=summ))

    def close(self):
        if self._w is not None:
            self._w.flush()
            self._w.close()
            self._w = None

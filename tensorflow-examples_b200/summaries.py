"""TensorBoard event files for the two scalars the reference logs every step
(distributed/distributed.py:121-125 tf.summary.scalar("cost"/"accuracy"), :138 FileWriter(logs_path), :151 add_summary).

Host-side only: the values come out of the device trace that b2_train_steps fills, once per print
window.  Uses tensorboard's own EventFileWriter when it is installed; without it a small native writer produces
the same file format (TFRecord framing with masked CRC-32C, hand-encoded Event / Summary protobufs), so the
feature does not depend on the package (the training step never depends on either)."""
import os
import socket
import struct
import time


# ---- native event-file writer ------------------------------------------------------------------------------------
def _crc32c_table():
    t = []
    for i in range(256):
        c = i
        for _ in range(8):
            c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1     # Castagnoli polynomial, reflected
        t.append(c)
    return t


_CRC_TABLE = _crc32c_table()


def crc32c(data):
    c = 0xFFFFFFFF
    for b in data:
        c = _CRC_TABLE[(c ^ b) & 0xFF] ^ (c >> 8)
    return c ^ 0xFFFFFFFF


def masked_crc32c(data):
    """The checksum TFRecord files store: crc32c rotated right by 15 bits plus a constant."""
    c = crc32c(data)
    return (((c >> 15) | (c << 17)) + 0xA282EAD8) & 0xFFFFFFFF


def _varint(n):
    out = bytearray()
    n &= (1 << 64) - 1
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _len_field(field, payload):
    return _varint((field << 3) | 2) + _varint(len(payload)) + payload


def encode_event(wall_time, step, scalars=None, file_version=None):
    """tensorflow.Event: wall_time = 1 (double), step = 2 (int64), file_version = 3 (string), summary = 5 (Summary);
    Summary: repeated Value value = 1; Value: tag = 1 (string), simple_value = 2 (float)."""
    ev = b"\x09" + struct.pack("<d", float(wall_time)) + b"\x10" + _varint(int(step))
    if file_version is not None:
        ev += _len_field(3, file_version.encode("utf-8"))
    if scalars:
        summ = b""
        for tag, value in scalars:
            val = _len_field(1, tag.encode("utf-8")) + b"\x15" + struct.pack("<f", float(value))
            summ += _len_field(1, val)
        ev += _len_field(5, summ)
    return ev


def encode_record(data):
    """One TFRecord: uint64 length, masked crc of the length, the data, masked crc of the data (little endian)."""
    head = struct.pack("<Q", len(data))
    return head + struct.pack("<I", masked_crc32c(head)) + data + struct.pack("<I", masked_crc32c(data))


class NativeEventWriter:
    """events.out.tfevents.<time>.<host> in `logs_path`, first record = the file-version event, like tf.summary.FileWriter."""

    def __init__(self, logs_path):
        os.makedirs(logs_path, exist_ok=True)
        now = time.time()
        self.path = os.path.join(logs_path, "events.out.tfevents.%010d.%s.%d.0" % (int(now), socket.gethostname(), os.getpid()))
        self._f = open(self.path, "wb")
        self._f.write(encode_record(encode_event(now, 0, file_version="brain.Event:2")))

    def add_scalars(self, step, scalars):
        self._f.write(encode_record(encode_event(time.time(), step, scalars=scalars)))

    def flush(self):
        self._f.flush()

    def close(self):
        self._f.close()


class SummaryWriter:
    def __init__(self, logs_path):
        self._w = None
        self._native = None
        if not logs_path:
            return
        if os.environ.get("B2_NATIVE_SUMMARIES") != "1":
            try:
                from tensorboard.compat.proto.event_pb2 import Event  # noqa: F401
                from tensorboard.summary.writer.event_file_writer import EventFileWriter
                self._w = EventFileWriter(logs_path)
                return
            except Exception:   # tensorboard missing or unusable: write the same format ourselves
                pass
        self._native = NativeEventWriter(logs_path)
        self._w = self._native

    @property
    def enabled(self):
        return self._w is not None

    def add_scalars(self, step, cost, accuracy):
        """One Event per step carrying both scalars, like the merged summary_op (distributed.py:125,149-151)."""
        if self._w is None:
            return
        if self._native is not None:
            self._native.add_scalars(step, [("cost", cost), ("accuracy", accuracy)])
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

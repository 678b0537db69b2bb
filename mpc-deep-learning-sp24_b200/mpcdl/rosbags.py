"""rosbag2 -> training rows on the GPU: the reference's data preparation path

    data/rosbags.py:100-124      RosBagConnector.getAllMessagesInTopic*  (sqlite3: SELECT TIMESTAMP, DATA FROM MESSAGES ...)
    (rclpy + mpclab_msgs)        deserialize_message(data, VehicleStateMsg)          -- not importable: private package
    scripts/setup_data.py:52-84  setup_data: sort by t, drop the last second, keep 8 fields per message

with the CDR decode and setup_data as device code (csrc/cdr_kernels.cu: one thread per message walks the CDR stream of
mpclab_msgs/msg/VehicleStateMsg).  The sqlite3 read is the storage engine and stays on the host: a bag's blobs are joined
into one byte buffer + offsets and copied once.  `bags_to_rows` returns the nested-list shape of `*_clean.pkl`
(data[bag][msg][8] python floats), so `RobotData(bags_to_rows(paths))` replaces the pickle round trip."""
from __future__ import annotations

import sqlite3

import numpy as np
import torch

from . import _lib  # noqa: F401  (loads libmpcdl.so; raises if it is missing)

TOPIC_TYPE = "mpclab_msgs/msg/VehicleStateMsg"


class RosBagConnector:
    """the query surface of data/rosbags.py:13-124 this path uses (read-only connection)"""

    def __init__(self, sqlite_file):
        self.conn = sqlite3.connect(f"file:{sqlite_file}?mode=ro", uri=True)
        self.cursor = self.conn.cursor()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.conn.close()

    def countRows(self, table_name):
        return self.cursor.execute(f"SELECT COUNT(*) FROM {table_name}").fetchall()[0][0]

    def getAllTopicsInfo(self):
        return {row[1]: row[2] for row in self.cursor.execute("SELECT * from(topics)").fetchall()}

    def is_topic(self, topic_name):
        return len(self.cursor.execute("SELECT NAME from TOPICS WHERE NAME=?", (topic_name,)).fetchall()) > 0

    def getTopicInfo(self, topic_name):
        if not self.is_topic(topic_name):
            raise KeyError(f"Topic {topic_name} not found! ")
        return self.cursor.execute("SELECT ID, TYPE FROM TOPICS WHERE NAME=?", (topic_name,)).fetchall()[0]

    def getAllMessagesInTopic(self, topic_name, limit=None, offset=0):
        """(TIMESTAMP, DATA) rows of a topic in storage order and the message type (rosbags.py:113-124; there LIMIT / OFFSET
        default to 100 / 100, here to the whole topic)"""
        topic_id, msg_type = self.getTopicInfo(topic_name)
        q = "SELECT TIMESTAMP, DATA FROM MESSAGES WHERE TOPIC_ID=? ORDER BY ID"
        if limit is not None:
            q += f" LIMIT {int(limit)} OFFSET {int(offset)}"
        return self.cursor.execute(q, (topic_id,)).fetchall(), msg_type


def pack_blobs(blobs):
    """list of bytes -> (uint8 tensor of the messages back to back, int64 offsets [n + 1]), pinned host memory"""
    lens = np.fromiter((len(b) for b in blobs), dtype=np.int64, count=len(blobs))
    offs = np.zeros(len(blobs) + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    buf = torch.from_numpy(np.frombuffer(b"".join(blobs), dtype=np.uint8).copy())
    return buf, torch.from_numpy(offs)


def decode_messages(blobs, device="cuda"):
    """CDR blobs of VehicleStateMsg -> float64 [n, 9] = [t, x, y, psi, w_psi, v_long, v_tran, u_a, u_steer] on `device`.
    Raises if a message is truncated or not little-endian CDR."""
    if len(blobs) == 0:
        return torch.empty(0, 9, dtype=torch.float64, device=device)
    buf, offs = pack_blobs(blobs)
    rows, status = torch.ops.mpcdl.cdr_decode(buf.to(device), offs.to(device))
    bad = int((status != 0).sum().item())
    if bad:
        first = int(torch.nonzero(status)[0].item())
        raise RuntimeError(f"mpcdl.rosbags: {bad} of {len(blobs)} messages are truncated or not little-endian CDR (first: #{first})")
    return rows


def setup_rows(rows9, cut_time=1.00):
    """scripts/setup_data.py:64-82 on the device: stable sort by t, keep t <= t_last - cut_time, drop the t column"""
    if rows9.shape[0] == 0:
        return rows9[:, 1:]
    order = torch.argsort(rows9[:, 0], stable=True)
    s = rows9[order]
    keep = s[:, 0] <= (s[-1, 0] - cut_time)
    return s[keep][:, 1:].contiguous()


def bag_rows(sqlite_file, topic_name=None, cut_time=1.00, device="cuda"):
    """one .db3 file -> float64 [m, 8] rows [x, y, psi, w_psi, v_long, v_tran, u_a, u_steer] on `device`"""
    with RosBagConnector(sqlite_file) as bag:
        topics = bag.getAllTopicsInfo()
        if topic_name is None:
            names = [n for n, t in topics.items() if t == TOPIC_TYPE]
            if len(names) != 1:
                raise KeyError(f"{sqlite_file}: expected one {TOPIC_TYPE} topic, found {names}")
            topic_name = names[0]
        msgs, msg_type = bag.getAllMessagesInTopic(topic_name)
    if msg_type != TOPIC_TYPE:
        raise RuntimeError(f"mpcdl.rosbags: topic {topic_name} carries {msg_type}, the decoder is written for {TOPIC_TYPE}")
    return setup_rows(decode_messages([m[1] for m in msgs], device), cut_time)


def bags_to_rows(sqlite_files, cut_time=1.00, device="cuda"):
    """the `*_clean.pkl` structure straight from the bags: data[bag][msg][8] (python floats, as setup_data returns)"""
    return [bag_rows(f, None, cut_time, device).cpu().tolist() for f in sqlite_files]

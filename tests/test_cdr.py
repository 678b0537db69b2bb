"""rosbag2 CDR decode + setup_data.  The reference's deserialiser (rclpy + mpclab_msgs) is absent: parity of
the decode is UNPINNED (oracle/cdr_ref.py header); what is tested is the oracle against an independent CDR writer and
against the shipped bag's own structure, and the CUDA kernel bit for bit against the oracle."""
import os
import random
import struct

import numpy as np
import pytest
import torch

from conftest import GOLD


def _bag():
    g = np.load(os.path.join(GOLD, "cdr_bag.npz"))
    return [bytes(b) for b in g["blobs"]], g


def _synthetic(n, seed):
    """messages with frame_id strings of every length 0..11 (all alignment cases), random payload, random tail"""
    from oracle import cdr_ref as CDR
    rng = random.Random(seed)
    blobs, rows = [], []
    for _ in range(n):
        vals = [rng.uniform(-10, 10) for _ in range(34)]
        fids = tuple("f" * rng.randrange(12) for _ in range(3))
        blobs.append(CDR.encode(vals, fids, stamp=(rng.randrange(2 ** 31), rng.randrange(10 ** 9)), tail=rng.randrange(0, 200)))
        rows.append([vals[i] for i in CDR.ROW_OF_VALS])
    return blobs, np.asarray(rows)


def test_oracle_decodes_what_the_independent_writer_wrote():
    from oracle import cdr_ref as CDR
    blobs, rows = _synthetic(300, 1)
    got = np.asarray([CDR.row9(CDR.decode(b)) for b in blobs])
    assert np.array_equal(got, rows)
    m = CDR.decode(CDR.encode(list(range(34)), ("base_link", "", "imu"), stamp=(12, 34)))
    assert m["header"] == {"stamp": {"sec": 12, "nanosec": 34}, "frame_id": "base_link"} and m["u"]["header"]["frame_id"] == "imu"
    with pytest.raises(ValueError):
        CDR.decode(bytes([0, 0, 0, 0]) + blobs[0][4:])             # big-endian encapsulation header
    with pytest.raises((ValueError, struct.error)):
        CDR.decode(blobs[0][:100])


def test_oracle_on_the_shipped_bag_fixture():
    """DAgger_827 (425 messages of 468 bytes, exported by oracle/pin_reference.py section 8): the fixture's rows, the message
    clock, and setup_data (sorted by t, last second dropped: scripts/setup_data.py:64-69)"""
    from oracle import cdr_ref as CDR
    blobs, g = _bag()
    msgs = [CDR.decode(b) for b in blobs]
    rows = np.asarray([CDR.row9(m) for m in msgs])
    assert np.array_equal(rows, g["rows9"]) and rows.shape == (425, 9)
    t = rows[:, 0]
    stamp = np.array([m["header"]["stamp"]["sec"] + 1e-9 * m["header"]["stamp"]["nanosec"] for m in msgs])
    assert np.abs(t - stamp).max() < 1e-6 and abs(np.median(np.diff(t)) - 0.02) < 1e-3
    clean = CDR.setup_data({"bag": dict(enumerate(msgs))})
    assert np.array_equal(np.asarray(clean[0]), g["clean"]) and len(clean[0]) == int((t <= t.max() - 1.0).sum()) < 425
    shuffled = dict(enumerate(random.Random(0).sample(msgs, len(msgs))))        # setup_data sorts: storage order does not matter
    assert CDR.setup_data({"bag": shuffled}) == clean


def test_cdr_entry_point_without_device():
    from mpcdl import _lib
    assert _lib.lib.mpcdl_cdr_decode(None, None, 10, None, None, None) == -1


@pytest.mark.gpu
def test_kernel_decodes_the_shipped_bag_bitwise():
    from mpcdl import rosbags
    blobs, g = _bag()
    rows = rosbags.decode_messages(blobs, "cuda:0")
    assert rows.dtype == torch.float64 and np.array_equal(rows.cpu().numpy(), g["rows9"])
    clean = rosbags.setup_rows(rows)
    assert np.array_equal(clean.cpu().numpy(), g["clean"])
    perm = torch.randperm(len(blobs), generator=torch.Generator().manual_seed(3)).tolist()      # out-of-order storage
    assert np.array_equal(rosbags.setup_rows(rosbags.decode_messages([blobs[i] for i in perm], "cuda:0")).cpu().numpy(), g["clean"])


@pytest.mark.gpu
def test_kernel_handles_every_string_alignment_and_flags_bad_messages():
    from mpcdl import rosbags
    blobs, rows = _synthetic(5000, 2)
    got = rosbags.decode_messages(blobs, "cuda:0")
    assert np.array_equal(got.cpu().numpy(), rows)
    bad = list(blobs[:50])
    bad[7] = bad[7][:120]                                    # truncated
    bad[30] = bytes([0, 0, 0, 0]) + bad[30][4:]              # big endian
    bad[41] = bad[41][:12] + struct.pack("<I", 1 << 30) + bad[41][16:]      # absurd string length
    buf, offs = rosbags.pack_blobs(bad)
    r, status = torch.ops.mpcdl.cdr_decode(buf.cuda(), offs.cuda())
    s = status.cpu().numpy()
    assert s[7] == 1 and s[30] == 1 and s[41] == 1 and s.sum() == 3
    ok = np.ones(50, bool); ok[[7, 30, 41]] = False
    assert np.array_equal(r.cpu().numpy()[ok], rows[:50][ok])
    with pytest.raises(RuntimeError):
        rosbags.decode_messages(bad, "cuda:0")
    with pytest.raises(RuntimeError):
        torch.ops.mpcdl.cdr_decode(buf, offs)                # CPU tensors: no fallback


@pytest.mark.gpu
def test_bag_file_to_robot_data(tmp_path):
    """a .db3 written here with the fixture's blobs (sqlite3 schema of rosbag2: topics + messages) -> bags_to_rows ->
    RobotData: the `*_clean.pkl` structure without the pickle round trip"""
    import sqlite3
    import mpcdl
    from mpcdl import rosbags
    blobs, g = _bag()
    path = str(tmp_path / "bag_0.db3")
    con = sqlite3.connect(path)
    con.execute("CREATE TABLE topics(id INTEGER PRIMARY KEY, name TEXT NOT NULL, type TEXT NOT NULL, serialization_format TEXT NOT NULL, offered_qos_profiles TEXT NOT NULL)")
    con.execute("CREATE TABLE messages(id INTEGER PRIMARY KEY, topic_id INTEGER NOT NULL, timestamp INTEGER NOT NULL, data BLOB NOT NULL)")
    con.execute("INSERT INTO topics VALUES (1, '/experiment/barc_3/est_state', 'mpclab_msgs/msg/VehicleStateMsg', 'cdr', '')")
    con.executemany("INSERT INTO messages(topic_id, timestamp, data) VALUES (1, ?, ?)", [(int(t), b) for t, b in zip(g["timestamps"], blobs)])
    con.commit(); con.close()
    with rosbags.RosBagConnector(path) as bag:
        assert bag.countRows("messages") == 425 and bag.getAllTopicsInfo() == {"/experiment/barc_3/est_state": "mpclab_msgs/msg/VehicleStateMsg"}
        with pytest.raises(KeyError):
            bag.getTopicInfo("/nope")
    data = rosbags.bags_to_rows([path, path], device="cuda:0")
    assert len(data) == 2 and data[0] == g["clean"].tolist()
    ds = mpcdl.RobotData(data)
    assert len(ds) == 2 * (len(data[0]) - 1)

"""CPU restatement of the rosbag2 -> rows path -- TEST INFRASTRUCTURE ONLY (imported by tests/ and tools, never by mpcdl).

    reference: data/rosbags.py:100-124   getAllMessagesInTopic*: SELECT TIMESTAMP, DATA FROM MESSAGES WHERE TOPIC_ID = ...
               (caller, not in the repo) rclpy.serialization.deserialize_message(data, mpclab_msgs.msg.VehicleStateMsg)
               scripts/setup_data.py:4-32  the resulting dict (field list and nesting), :52-84 setup_data: sort by msg["t"], drop
               the last `cut_time` seconds, keep [x.x, x.y, e.psi, w.w_psi, v.v_long, v.v_tran, u.u_a, u.u_steer]

PARITY UNPINNED for the decode: the deserialiser is rclpy + the private `mpclab_msgs` package, neither present in
/root/reference nor installable here, and the repository holds no output of it for the bags it ships (the validation pickles
come from other bags: no decoded value of the 72 048 shipped messages occurs in them).  What anchors this file instead:
  * the published CDR rules (OMG XCDR1 as rmw_fastrtps / rmw_cyclonedds write it: 4-byte encapsulation header {0x00, 0x01, 0, 0}
    = little endian; every primitive aligned to its own size counted from the start of the body; string = uint32 length
    including the terminating NUL, then the bytes);
  * the message layout: the field order of scripts/setup_data.py's docstring, all members float64, three std_msgs/Header
    members (top level, x, u); confirmed on the shipped bags: every message is 468 bytes, header.stamp and t agree to the
    microsecond, t advances by 20 ms, positions / speeds / controls fall in the BARC ranges of the validation pickles;
  * `encode` below, an independent CDR WRITER for the same layout, so that decode(encode(x)) == x is checked for frame_id
    strings of every length (the alignment cases the shipped bags, whose frame_ids are empty, never exercise).
`setup_data` restates scripts/setup_data.py:52-84 line for line (pinned: pure Python, no dependency)."""
import struct

FIELDS = ("t", "x", "y", "psi", "w_psi", "v_long", "v_tran", "u_a", "u_steer")


class _Reader:
    def __init__(self, body):
        self.b, self.o = body, 0

    def align(self, a):
        self.o = (self.o + a - 1) // a * a

    def u32(self):
        self.align(4)
        v = struct.unpack_from("<I", self.b, self.o)[0]
        self.o += 4
        return v

    def f64(self):
        self.align(8)
        v = struct.unpack_from("<d", self.b, self.o)[0]
        self.o += 8
        return v

    def header(self):
        sec, nsec, n = self.u32(), self.u32(), self.u32()
        s = self.b[self.o:self.o + n]
        self.o += n
        if len(s) != n:
            raise ValueError("truncated string")
        return sec, nsec, s[:-1].decode("latin1")


def decode(blob):
    """one serialised VehicleStateMsg -> dict in the nesting scripts/setup_data.py:4-32 documents (members up to u.u_steer)"""
    if len(blob) < 4 or blob[0] != 0 or blob[1] != 1:
        raise ValueError("not little-endian CDR")
    r = _Reader(bytes(blob[4:]))
    m = {}
    sec, nsec, fid = r.header()
    m["header"] = {"stamp": {"sec": sec, "nanosec": nsec}, "frame_id": fid}
    m["t"] = r.f64()
    sec, nsec, fid = r.header()
    m["x"] = {"header": {"stamp": {"sec": sec, "nanosec": nsec}, "frame_id": fid}, "x": r.f64(), "y": r.f64(), "z": r.f64()}
    m["e"] = {"phi": r.f64(), "theta": r.f64(), "psi": r.f64()}
    m["q"] = [r.f64() for _ in range(4)]
    m["w"] = {"w_phi": r.f64(), "w_theta": r.f64(), "w_psi": r.f64()}
    m["aa"] = {"a_phi": r.f64(), "a_theta": r.f64(), "a_psi": r.f64()}
    m["v"] = {"v_long": r.f64(), "v_tran": r.f64(), "v_n": r.f64()}
    m["a"] = {"a_long": r.f64(), "a_tran": r.f64(), "a_n": r.f64()}
    m["p"] = [r.f64() for _ in range(4)]
    m["pt"] = [r.f64() for _ in range(4)]
    sec, nsec, fid = r.header()
    m["u"] = {"header": {"stamp": {"sec": sec, "nanosec": nsec}, "frame_id": fid}, "t": r.f64(), "u_a": r.f64(), "u_steer": r.f64()}
    return m


def row9(m):
    return [m["t"], m["x"]["x"], m["x"]["y"], m["e"]["psi"], m["w"]["w_psi"], m["v"]["v_long"], m["v"]["v_tran"], m["u"]["u_a"], m["u"]["u_steer"]]


class _Writer:
    def __init__(self):
        self.b = bytearray()

    def align(self, a):
        self.b += b"\0" * (-len(self.b) % a)

    def u32(self, v):
        self.align(4)
        self.b += struct.pack("<I", v)

    def f64(self, v):
        self.align(8)
        self.b += struct.pack("<d", v)

    def header(self, sec, nsec, fid):
        self.u32(sec); self.u32(nsec)
        s = fid.encode("latin1") + b"\0"
        self.u32(len(s))
        self.b += s


def encode(vals, frame_ids=("", "", ""), stamp=(0, 0), tail=144):
    """CDR WRITER for the same layout (an independent statement of the alignment rules): vals = 34 float64 in stream order
    (t; x y z; phi theta psi; q0..3; w3; aa3; v3; a3; p4; pt4; u.t u_a u_steer), `tail` bytes of trailing members."""
    assert len(vals) == 34
    w = _Writer()
    it = iter(vals)
    w.header(stamp[0], stamp[1], frame_ids[0])
    w.f64(next(it))
    w.header(0, 0, frame_ids[1])
    for _ in range(3 + 3 + 4 + 3 + 3 + 3 + 3 + 4 + 4):
        w.f64(next(it))
    w.header(0, 0, frame_ids[2])
    for _ in range(3):
        w.f64(next(it))
    w.align(8)
    return bytes([0, 1, 0, 0]) + bytes(w.b) + b"\0" * tail


ROW_OF_VALS = (0, 1, 2, 6, 13, 17, 18, 32, 33)      # positions of row9's members in encode()'s value list


def setup_data(pkl_data, cut_time=1.00):
    """scripts/setup_data.py:52-84, line for line (print dropped): {bag name: {id: msg dict}} -> data[bag][msg][8]"""
    data_list = []
    for bag in pkl_data.values():
        bag_list = []
        bag_sort = sorted(bag.values(), key=lambda msg: msg["t"])
        max_time = bag_sort[-1]["t"] - cut_time
        for msg in bag_sort:
            if msg["t"] <= max_time:
                bag_list.append([msg["x"]["x"], msg["x"]["y"], msg["e"]["psi"], msg["w"]["w_psi"], msg["v"]["v_long"], msg["v"]["v_tran"],
                                 msg["u"]["u_a"], msg["u"]["u_steer"]])
        data_list.append(bag_list)
    return data_list

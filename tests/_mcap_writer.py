"""Test-side, independent MCAP v0 writer + CDR (XCDR1 little-endian) encoder for the two message types the SLAM node
consumes.  Written from the container / encoding specifications, NOT from the reader under test
(ekf-slam_b200/host/rosbag/mcap_reader.hpp), so that a round trip checks the reader against the format rather than
against itself.  The reference's own bag payload is missing from the checkout (.MISSING_LARGE_BLOBS:2); topic and
type names follow rosbags/rosbag2_2024_03_17-18_35_57/metadata.yaml."""
import ctypes
import struct
import zlib

MAGIC = b"\x89MCAP0\r\n"
OP_HEADER, OP_FOOTER, OP_SCHEMA, OP_CHANNEL, OP_MESSAGE, OP_CHUNK, OP_MESSAGE_INDEX, OP_CHUNK_INDEX = 1, 2, 3, 4, 5, 6, 7, 8
OP_STATISTICS, OP_SUMMARY_OFFSET, OP_DATA_END = 0x0B, 0x0E, 0x0F


# ---- CDR ---------------------------------------------------------------------------------------------------------
class CdrWriter:
    def __init__(self):
        self.b = bytearray(b"\x00\x01\x00\x00")      # CDR_LE encapsulation

    def _align(self, a):
        rel = len(self.b) - 4
        self.b += b"\x00" * ((a - rel % a) % a)

    def put(self, fmt, *v):
        self._align(struct.calcsize("<" + fmt[-1]))
        self.b += struct.pack("<" + fmt, *v)

    def string(self, s):
        raw = s.encode() + b"\x00"
        self.put("I", len(raw))
        self.b += raw

    def header(self, stamp_ns, frame_id):
        self.put("i", stamp_ns // 1_000_000_000)
        self.put("I", stamp_ns % 1_000_000_000)
        self.string(frame_id)

    def seq(self, fmt, values):
        self.put("I", len(values))
        for v in values:
            self.put(fmt, v)


def encode_joint_state(stamp_ns, names, position, velocity=(), effort=(), frame_id=""):
    w = CdrWriter()
    w.header(stamp_ns, frame_id)
    w.put("I", len(names))
    for n in names:
        w.string(n)
    w.seq("d", position)
    w.seq("d", velocity)
    w.seq("d", effort)
    return bytes(w.b)


def encode_marker_array(stamp_ns, markers, frame_id="green/base_scan"):
    """markers: dicts with id, x, y and optionally action, ns, text, points (list of xyz), colors (list of rgba),
    scale, mesh_resource, texture_data - the optional ones exercise every variable-length field of Marker."""
    w = CdrWriter()
    w.put("I", len(markers))
    for m in markers:
        w.header(stamp_ns, frame_id)
        w.string(m.get("ns", "detected"))
        w.put("i", m["id"])
        w.put("i", m.get("type", 3))                 # CYLINDER
        w.put("i", m.get("action", 0))
        for v in (m["x"], m["y"], m.get("z", 0.125), 0.0, 0.0, 0.0, 1.0):
            w.put("d", v)
        for v in m.get("scale", (0.076, 0.076, 0.25)):
            w.put("d", v)
        for v in (0.0, 1.0, 0.0, 1.0):
            w.put("f", v)
        w.put("i", 0); w.put("I", 200_000_000)       # lifetime
        w.put("B", 0)                                # frame_locked
        pts = m.get("points", [])
        w.put("I", len(pts))
        for p in pts:
            for v in p:
                w.put("d", v)
        cols = m.get("colors", [])
        w.put("I", len(cols))
        for c in cols:
            for v in c:
                w.put("f", v)
        w.string(m.get("texture_resource", ""))
        w.header(0, "")                              # texture: CompressedImage
        w.string(m.get("texture_format", ""))
        w.seq("B", m.get("texture_data", b""))
        uv = m.get("uv", [])
        w.put("I", len(uv))
        for c in uv:
            w.put("f", c[0]); w.put("f", c[1])
        w.string(m.get("text", ""))
        w.string(m.get("mesh_resource", ""))
        w.string(m.get("mesh_filename", ""))
        w.seq("B", m.get("mesh_data", b""))
        w.put("B", 0)
    return bytes(w.b)


# ---- compression through the system's runtime libraries --------------------------------------------------------------
def _zstd(data):
    lib = ctypes.CDLL("libzstd.so.1")
    lib.ZSTD_compressBound.restype = ctypes.c_size_t
    lib.ZSTD_compressBound.argtypes = [ctypes.c_size_t]
    lib.ZSTD_compress.restype = ctypes.c_size_t
    lib.ZSTD_compress.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_int]
    cap = lib.ZSTD_compressBound(len(data))
    out = ctypes.create_string_buffer(cap)
    n = lib.ZSTD_compress(out, cap, data, len(data), 3)
    return out.raw[:n]


def _lz4(data):
    lib = ctypes.CDLL("liblz4.so.1")
    lib.LZ4F_compressFrameBound.restype = ctypes.c_size_t
    lib.LZ4F_compressFrameBound.argtypes = [ctypes.c_size_t, ctypes.c_void_p]
    lib.LZ4F_compressFrame.restype = ctypes.c_size_t
    lib.LZ4F_compressFrame.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_size_t, ctypes.c_void_p]
    cap = lib.LZ4F_compressFrameBound(len(data), None)
    out = ctypes.create_string_buffer(cap)
    n = lib.LZ4F_compressFrame(out, cap, data, len(data), None)
    return out.raw[:n]


def compression_available(method):
    try:
        {"": lambda d: d, "zstd": _zstd, "lz4": _lz4}[method](b"probe" * 10)
        return True
    except OSError:
        return False


# ---- MCAP ----------------------------------------------------------------------------------------------------------------
def _s(x):
    raw = x.encode()
    return struct.pack("<I", len(raw)) + raw


def _rec(op, body):
    return struct.pack("<BQ", op, len(body)) + body


def _map(pairs, kfmt, vfmt):
    raw = b"".join(struct.pack("<" + kfmt + vfmt, k, v) for k, v in pairs)
    return struct.pack("<I", len(raw)) + raw


def write_mcap(path, channels, messages, chunk_messages=64, compression="", channels_in_chunks=True):
    """channels: {channel_id: (topic, schema_name)}; messages: [(channel_id, log_time_ns, payload bytes)] in FILE order
    (deliberately not required to be time-sorted).  chunk_messages = 0 writes unchunked Message records."""
    compress = {"": lambda d: d, "zstd": _zstd, "lz4": _lz4}[compression]
    out = bytearray(MAGIC)
    out += _rec(OP_HEADER, _s("ros2") + _s("tests/_mcap_writer.py"))
    schema_recs, channel_recs = {}, {}
    for cid, (topic, schema) in channels.items():
        schema_recs[cid] = _rec(OP_SCHEMA, struct.pack("<H", cid) + _s(schema) + _s("ros2msg") + struct.pack("<I", 0))
        meta = _map([], "", "")
        channel_recs[cid] = _rec(OP_CHANNEL, struct.pack("<HH", cid, cid) + _s(topic) + _s("cdr") + meta)
    data_start = len(out)
    declared = set()
    if not (channels_in_chunks and chunk_messages):
        for cid in channels:
            out += schema_recs[cid] + channel_recs[cid]
            declared.add(cid)
    chunk_indexes, seq = [], 0

    def message_record(cid, t, payload):
        nonlocal seq
        seq += 1
        return _rec(OP_MESSAGE, struct.pack("<HIQQ", cid, seq, t, t) + payload)

    if chunk_messages:
        for i in range(0, len(messages), chunk_messages):
            group = messages[i:i + chunk_messages]
            records, index = bytearray(), {}
            for cid, t, payload in group:
                if cid not in declared:
                    records += schema_recs[cid] + channel_recs[cid]
                    declared.add(cid)
                index.setdefault(cid, []).append((t, len(records)))
                records += message_record(cid, t, payload)
            times = [t for _, t, _ in group]
            comp = compress(bytes(records))
            body = struct.pack("<QQQI", min(times), max(times), len(records), zlib.crc32(bytes(records))) + \
                _s(compression) + struct.pack("<Q", len(comp)) + comp
            chunk_start = len(out)
            out += _rec(OP_CHUNK, body)
            chunk_len = len(out) - chunk_start
            mi_start, offsets = len(out), []
            for cid, entries in index.items():
                offsets.append((cid, len(out)))
                out += _rec(OP_MESSAGE_INDEX, struct.pack("<H", cid) + _map(entries, "Q", "Q"))
            chunk_indexes.append(_rec(OP_CHUNK_INDEX, struct.pack("<QQQQ", min(times), max(times), chunk_start, chunk_len) +
                                      _map(offsets, "H", "Q") + struct.pack("<Q", len(out) - mi_start) + _s(compression) +
                                      struct.pack("<QQ", len(comp), len(records))))
    else:
        for cid, t, payload in messages:
            out += message_record(cid, t, payload)
    out += _rec(OP_DATA_END, struct.pack("<I", zlib.crc32(bytes(out[data_start:]))))
    # summary: schemas, channels, statistics, chunk indexes, then the summary offsets
    summary_start = len(out)
    groups = []
    for op, recs in ((OP_SCHEMA, list(schema_recs.values())), (OP_CHANNEL, list(channel_recs.values()))):
        start = len(out)
        for r in recs:
            out += r
        groups.append((op, start, len(out) - start))
    counts = {}
    for cid, _, _ in messages:
        counts[cid] = counts.get(cid, 0) + 1
    times = [t for _, t, _ in messages] or [0]
    start = len(out)
    out += _rec(OP_STATISTICS, struct.pack("<QHIIIIQQ", len(messages), len(channels), len(channels), 0, 0, len(chunk_indexes),
                                           min(times), max(times)) + _map(sorted(counts.items()), "H", "Q"))
    groups.append((OP_STATISTICS, start, len(out) - start))
    if chunk_indexes:
        start = len(out)
        for r in chunk_indexes:
            out += r
        groups.append((OP_CHUNK_INDEX, start, len(out) - start))
    offset_start = len(out)
    for op, start, length in groups:
        out += _rec(OP_SUMMARY_OFFSET, struct.pack("<BQQ", op, start, length))
    footer_prefix = struct.pack("<BQ", OP_FOOTER, 20) + struct.pack("<QQ", summary_start, offset_start)
    crc = zlib.crc32(bytes(out[summary_start:]) + footer_prefix)
    out += footer_prefix + struct.pack("<I", crc) + MAGIC
    with open(path, "wb") as f:
        f.write(out)
    return len(out)

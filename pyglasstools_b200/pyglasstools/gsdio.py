"""Minimal GSD (General Simulation Data) reader / writer for the HOOMD schema subset this path needs.

Replaces the reference's dependency on ``gsd.hoomd`` (pyglasstools/utils.py:5,41-43), which is not installed
on the GPU box.  File layout (gsd.h:127-190 of the vendored gsd 2.1.2): a 256-byte header, an index of 32-byte
entries ``(frame u64, N u64, location i64, M u32, id u16, type u8, flags u8)``, a name list (64-byte slots in
v1 files, NUL-separated strings in v2 files) and the raw chunks.

Only what ``utils.read_gsd`` touches is modelled: ``traj[k].configuration.box``, ``.configuration.step``,
``traj[k].particles.{N, position, diameter, mass, velocity}`` with the HOOMD schema's fall-backs (a chunk absent
from frame k is taken from frame 0, else the schema default: mass 1, diameter 1, velocity 0).
"""
import builtins
import struct

import numpy as np

_MAGIC = 0x65DF65DF65DF65DF
_HEADER = struct.Struct("<QQQQQII64s64s80s")
_ENTRY = np.dtype([("frame", "<u8"), ("N", "<u8"), ("location", "<i8"), ("M", "<u4"), ("id", "<u2"), ("type", "u1"),
                   ("flags", "u1")])
_TYPES = {1: np.uint8, 2: np.uint16, 3: np.uint32, 4: np.uint64, 5: np.int8, 6: np.int16, 7: np.int32, 8: np.int64,
          9: np.float32, 10: np.float64}
_TYPE_IDS = {np.dtype(v): k for k, v in _TYPES.items()}


class GSDFile(object):
    """Read-only access to the chunks of a GSD file."""

    def __init__(self, name):
        self.name = name
        self._data = np.memmap(name, dtype=np.uint8, mode="r")
        if self._data.size < _HEADER.size:
            raise RuntimeError("{}: not a GSD file (too short)".format(name))
        (magic, iloc, ialloc, nloc, nalloc, self.schema_version, self.gsd_version, app, schema,
         _) = _HEADER.unpack(bytes(self._data[:_HEADER.size]))
        if magic != _MAGIC:
            raise RuntimeError("{}: not a GSD file (bad magic)".format(name))
        self.application = app.split(b"\0")[0].decode()
        self.schema = schema.split(b"\0")[0].decode()
        major = self.gsd_version >> 16
        if major not in (1, 2):
            raise RuntimeError("{}: unsupported GSD file version {}".format(name, major))
        raw_names = bytes(self._data[nloc:nloc + nalloc * 64])
        if major == 1:
            names = [raw_names[k:k + 64].split(b"\0")[0] for k in range(0, len(raw_names), 64)]
            names = [n for n in names if n]
        else:
            names = [n for n in raw_names.split(b"\0") if n]
        self.names = [n.decode() for n in names]
        index = np.frombuffer(self._data, dtype=_ENTRY, count=ialloc, offset=iloc)
        used = np.flatnonzero(index["location"] == 0)
        self._index = index[:used[0]] if used.size else index
        self.nframes = int(self._index["frame"].max()) + 1 if self._index.size else 0
        # (frame, name id) -> row of the index
        self._lookup = {(int(e["frame"]), int(e["id"])): k for k, e in enumerate(self._index)}
        self._name_id = {n: k for k, n in enumerate(self.names)}

    def chunk_exists(self, frame, name):
        return (frame, self._name_id.get(name, -1)) in self._lookup

    def read_chunk(self, frame, name):
        e = self._index[self._lookup[(frame, self._name_id[name])]]
        dt = np.dtype(_TYPES[int(e["type"])])
        n, m = int(e["N"]), int(e["M"])
        a = np.frombuffer(self._data, dtype=dt, count=n * m, offset=int(e["location"]))
        return a.reshape(n, m) if m > 1 else a.reshape(n)

    def close(self):
        self._lookup = {}
        self._index = None
        self._data = None


class _Section(object):
    pass


class Snapshot(object):
    """One frame: ``configuration`` (box, step, dimensions) and ``particles``."""

    def __init__(self):
        self.configuration = _Section()
        self.particles = _Section()


class HOOMDTrajectory(object):
    """Sequence of Snapshots read lazily from a GSD file in the 'hoomd' schema."""

    def __init__(self, gsdfile):
        self.file = gsdfile
        if gsdfile.schema != "hoomd":
            raise RuntimeError("{}: schema '{}' is not 'hoomd'".format(gsdfile.name, gsdfile.schema))

    def __len__(self):
        return self.file.nframes

    def _get(self, frame, name, default):
        if self.file.chunk_exists(frame, name):
            return self.file.read_chunk(frame, name)
        if frame != 0 and self.file.chunk_exists(0, name):
            return self.file.read_chunk(0, name)
        return default() if callable(default) else default

    def _default(self, kind, n):
        """Schema defaults for chunks a file does not hold (velocity 0, mass 1, ...): one read-only array per (kind, N),
        shared by the frames instead of being allocated and touched anew for each."""
        cache = self.__dict__.setdefault("_defaults", {})
        a = cache.get((kind, n))
        if a is None:
            a = np.zeros((n, 3), np.float32) if kind == "zeros3" else np.ones(n, np.float32)
            a.setflags(write=False)
            cache[(kind, n)] = a
        return a

    def read_frame(self, frame):
        frame = int(frame)
        if frame < 0:
            frame += len(self)
        if frame < 0 or frame >= len(self):
            raise IndexError("frame {} outside [0,{})".format(frame, len(self)))
        snap = Snapshot()
        c, p = snap.configuration, snap.particles
        c.step = int(self._get(frame, "configuration/step", np.zeros(1, np.uint64))[0])
        c.dimensions = int(self._get(frame, "configuration/dimensions", np.array([3], np.uint8))[0])
        c.box = np.asarray(self._get(frame, "configuration/box", np.array([1, 1, 1, 0, 0, 0], np.float32)), np.float32)
        p.N = int(self._get(frame, "particles/N", np.zeros(1, np.uint32))[0])
        n = p.N
        p.position = self._get(frame, "particles/position", lambda: self._default("zeros3", n))
        p.velocity = self._get(frame, "particles/velocity", lambda: self._default("zeros3", n))
        p.diameter = self._get(frame, "particles/diameter", lambda: self._default("ones", n))
        p.mass = self._get(frame, "particles/mass", lambda: self._default("ones", n))
        for key, shape in (("position", (n, 3)), ("velocity", (n, 3)), ("diameter", (n,)), ("mass", (n,))):
            if getattr(p, key).shape != shape:
                raise RuntimeError("{}: particles/{} of frame {} has shape {}, expected {}".format(
                    self.file.name, key, frame, getattr(p, key).shape, shape))
        return snap

    def __getitem__(self, key):
        if isinstance(key, slice):
            return [self.read_frame(k) for k in range(*key.indices(len(self)))]
        return self.read_frame(key)

    def __iter__(self):
        return (self.read_frame(k) for k in range(len(self)))

    def close(self):
        self.file.close()


def open(name, mode="rb"):  # noqa: A001  (mirrors gsd.hoomd.open)
    """``gsd.hoomd.open(name, 'rb')`` for reading; writing goes through :func:`write_trajectory`."""
    if mode not in ("rb", "r"):
        raise ValueError("gsdio.open only reads; use gsdio.write_trajectory to create files")
    return HOOMDTrajectory(GSDFile(name))


def write_trajectory(name, frames, application="pyglasstools_b200"):
    """Writes a GSD v2 file in the hoomd schema.

    ``frames`` is an iterable of dicts with keys ``box`` (6 floats or 3: Lx,Ly,Lz), ``position`` (N,3) and
    optionally ``diameter``, ``mass``, ``velocity``, ``step``.  Arrays are stored float32 like HOOMD does.
    """
    chunks = []      # (frame, name, array)

    def put(frame, cname, a, dtype):
        a = np.ascontiguousarray(a, dtype=dtype)
        chunks.append((frame, cname, a))

    for k, fr in enumerate(frames):
        box = np.asarray(fr["box"], np.float32).ravel()
        if box.size == 3:
            box = np.concatenate([box, np.zeros(3, np.float32)])
        pos = np.asarray(fr["position"])
        put(k, "configuration/step", np.array([fr.get("step", k)]), np.uint64)
        put(k, "configuration/dimensions", np.array([fr.get("dimensions", 2)]), np.uint8)
        put(k, "configuration/box", box, np.float32)
        put(k, "particles/N", np.array([pos.shape[0]]), np.uint32)
        put(k, "particles/position", pos, np.float32)
        for key in ("diameter", "mass", "velocity"):
            if fr.get(key) is not None:
                put(k, "particles/" + key, fr[key], np.float32)

    names = []
    for _, cname, _ in chunks:
        if cname not in names:
            names.append(cname)
    name_blob = b"".join(n.encode() + b"\0" for n in names)
    nalloc = max(1, (len(name_blob) + 1 + 63) // 64)
    name_blob = name_blob.ljust(nalloc * 64, b"\0")
    ialloc = max(128, len(chunks) + 1)
    index = np.zeros(ialloc, _ENTRY)
    iloc = _HEADER.size
    nloc = iloc + ialloc * _ENTRY.itemsize
    cursor = nloc + len(name_blob)
    order = sorted(range(len(chunks)), key=lambda k: (chunks[k][0], names.index(chunks[k][1])))
    with builtins.open(name, "wb") as f:
        f.write(_HEADER.pack(_MAGIC, iloc, ialloc, nloc, nalloc, (1 << 16) | 4, (2 << 16) | 0, application.encode()[:63],
                             b"hoomd", b""))
        f.seek(cursor)
        for row, k in enumerate(order):
            frame, cname, a = chunks[k]
            e = index[row]
            e["frame"], e["N"], e["location"] = frame, a.shape[0], cursor
            e["M"] = a.shape[1] if a.ndim > 1 else 1
            e["id"], e["type"], e["flags"] = names.index(cname), _TYPE_IDS[a.dtype], 0
            f.write(a.tobytes())
            cursor += a.nbytes
        f.seek(iloc)
        f.write(index.tobytes())
        f.write(name_blob)

"""Scenario I/O next to the MPPI path.

* `read_obstacle_field`: the reference's circle-obstacle scenarios, `PathPlanning/Scenarios/obstacle_field.mat`
  — a MAT-file level 5 holding one cell array `obstacle_field` of 100 trials, each an n x 3 double matrix
  `[x y r]` — as read by `matread("../Scenarios/obstacle_field.mat")["obstacle_field"][Trial_num]`
  (`PathPlanning/Astar/main.jl:15-22`).  Self-contained reader (struct + zlib + numpy): cell arrays and real
  numeric matrices, compressed or not; that is all the file holds.
* `write_trajectory_csv` (in mppi.py) writes the closed-loop log the way `main.jl:87` does.
"""
import struct
import zlib

import numpy as np

# MAT level-5 data types (miXXX) -> numpy
_MI = {1: "i1", 2: "u1", 3: "i2", 4: "u2", 5: "i4", 6: "u4", 7: "f4", 9: "f8", 12: "i8", 13: "u8"}
_MI_MATRIX, _MI_COMPRESSED = 14, 15
_MX_CELL = 1
_MX_NUMERIC = {6: "f8", 7: "f4", 8: "i1", 9: "u1", 10: "i2", 11: "u2", 12: "i4", 13: "u4", 14: "i8", 15: "u8"}


class MatFormatError(ValueError):
    pass


def _tag(buf, pos, bo):
    """One data-element tag at pos -> (type, nbytes, data offset, offset of the next element)."""
    word, = struct.unpack_from(bo + "I", buf, pos)
    if word >> 16:  # small element: type and size share the first word, data in the second
        n = word >> 16
        return word & 0xFFFF, n, pos + 4, pos + 8
    n, = struct.unpack_from(bo + "I", buf, pos + 4)
    return word, n, pos + 8, pos + 8 + ((n + 7) & ~7)


def _numeric(buf, pos, bo):
    t, n, off, nxt = _tag(buf, pos, bo)
    if t not in _MI:
        raise MatFormatError(f"unsupported element type {t}")
    return np.frombuffer(buf, dtype=np.dtype(_MI[t]).newbyteorder(bo), count=n // np.dtype(_MI[t]).itemsize, offset=off), nxt


def _matrix(buf, pos, end, bo):
    """Body of one miMATRIX element -> (name, value)."""
    flags, pos = _numeric(buf, pos, bo)
    cls = int(flags[0]) & 0xFF
    if int(flags[0]) & 0x0800:
        raise MatFormatError("complex arrays are not supported")
    dims, pos = _numeric(buf, pos, bo)
    dims = [int(d) for d in dims]
    name, pos = _numeric(buf, pos, bo)
    name = name.astype("u1").tobytes().decode("ascii")
    if cls == _MX_CELL:
        cells = []
        for _ in range(int(np.prod(dims))):
            t, n, off, nxt = _tag(buf, pos, bo)
            if t != _MI_MATRIX:
                raise MatFormatError("cell element is not a matrix")
            cells.append(_matrix(buf, off, off + n, bo)[1] if n else np.zeros((0, 0)))
            pos = nxt
        out = np.empty(len(cells), dtype=object)
        out[:] = cells
        return name, out.reshape(dims, order="F")
    if cls in _MX_NUMERIC:
        real, pos = _numeric(buf, pos, bo)   # may be stored in a narrower type than the class
        return name, np.asarray(real, dtype=_MX_NUMERIC[cls]).reshape(dims, order="F")
    raise MatFormatError(f"unsupported array class {cls}")


def read_mat5(path):
    """{name: array} for the cell / real numeric variables of a MAT-file level 5."""
    raw = open(path, "rb").read()
    if len(raw) < 128 or not raw.startswith(b"MATLAB 5.0 MAT-file"):
        raise MatFormatError("not a MAT-file level 5")
    endian = raw[126:128]
    bo = "<" if endian == b"IM" else ">" if endian == b"MI" else None
    if bo is None:
        raise MatFormatError("bad endian indicator")
    out, pos = {}, 128
    while pos + 8 <= len(raw):
        t, n, off, nxt = _tag(raw, pos, bo)
        if t == _MI_COMPRESSED:
            buf = zlib.decompress(raw[off:off + n])
            nxt = off + n   # compressed elements are not padded
            t, n, off, _ = _tag(buf, 0, bo)
        else:
            buf = raw
        if t != _MI_MATRIX:
            raise MatFormatError(f"top-level element of type {t}")
        name, val = _matrix(buf, off, off + n, bo)
        out[name] = val
        pos = nxt
    return out


def read_obstacle_field(path, name="obstacle_field"):
    """The reference's scenario file -> list of trials, each an (n, 3) float64 array of [x, y, r]
    (Astar/main.jl:15-22 indexes it 1-based: Trial_num 53 is element 52 here)."""
    cells = read_mat5(path)[name]
    trials = [np.array(c, dtype=np.float64) for c in cells.ravel(order="F")]
    for t in trials:
        if t.ndim != 2 or t.shape[1] != 3:
            raise MatFormatError(f"trial of shape {t.shape}, expected n x 3")
    return trials

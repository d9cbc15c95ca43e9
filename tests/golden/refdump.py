"""Reader for the flat "named array" files written by oracle/ref/ref_driver.cpp (put()).

Test infrastructure only.  Record layout: u32 name_len, name, u8 dtype ('d' f64 | 'q' i64),
u32 ndim, u64 shape[ndim], raw little-endian data.
"""
import struct
import numpy as np


def read_dump(path):
    out = {}
    with open(path, "rb") as f:
        buf = f.read()
    o = 0
    while o < len(buf):
        (nl,) = struct.unpack_from("<I", buf, o); o += 4
        name = buf[o:o + nl].decode(); o += nl
        dt = chr(buf[o]); o += 1
        (nd,) = struct.unpack_from("<I", buf, o); o += 4
        shape = struct.unpack_from("<%dQ" % nd, buf, o); o += 8 * nd
        n = int(np.prod(shape)) if nd else 1
        dtype = np.float64 if dt == "d" else np.int64
        out[name] = np.frombuffer(buf, dtype=dtype, count=n, offset=o).reshape(shape).copy()
        o += 8 * n
    return out

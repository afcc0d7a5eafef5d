"""Reader/writer for the CTBM0001 named-array container (include/cantera_b200/MechTables.h).

The same little-endian layout holds flat mechanism tables (``cantera_b200/data/*.ctbm``) and the
golden vectors under ``tests/golden/`` (gzip-compressed there).
"""
import gzip
import struct

import numpy as np

_MAGIC = b"CTBM0001"


def read_arrays(path):
    """Return ``{name: ndarray | str}`` for a .ctbm or .ctbm.gz file."""
    opener = gzip.open if str(path).endswith(".gz") else open
    with opener(path, "rb") as f:
        blob = f.read()
    if blob[:8] != _MAGIC:
        raise ValueError(f"{path} is not a CTBM0001 file")
    (n,) = struct.unpack_from("<I", blob, 8)
    off = 12
    out = {}
    for _ in range(n):
        (nl,) = struct.unpack_from("<I", blob, off)
        off += 4
        name = blob[off:off + nl].decode()
        off += nl
        dtype, cnt = struct.unpack_from("<IQ", blob, off)
        off += 12
        if dtype == 0:
            out[name] = np.frombuffer(blob, dtype="<f8", count=cnt, offset=off).copy()
            off += 8 * cnt
        elif dtype == 1:
            out[name] = np.frombuffer(blob, dtype="<i4", count=cnt, offset=off).copy()
            off += 4 * cnt
        else:
            out[name] = blob[off:off + cnt].decode()
            off += cnt
    return out


def write_arrays(path, arrays):
    """Write ``{name: ndarray(float64|int32) | str}`` as a CTBM0001 file."""
    with open(path, "wb") as f:
        f.write(_MAGIC)
        f.write(struct.pack("<I", len(arrays)))
        for name, v in arrays.items():
            nb = name.encode()
            f.write(struct.pack("<I", len(nb)))
            f.write(nb)
            if isinstance(v, str):
                b = v.encode()
                f.write(struct.pack("<IQ", 2, len(b)))
                f.write(b)
            else:
                a = np.ascontiguousarray(v)
                if a.dtype.kind == "f":
                    a = a.astype("<f8")
                    f.write(struct.pack("<IQ", 0, a.size))
                else:
                    a = a.astype("<i4")
                    f.write(struct.pack("<IQ", 1, a.size))
                f.write(a.tobytes())

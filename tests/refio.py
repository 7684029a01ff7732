"""Reader for the record files written by oracle/ref_driver.cpp (Tier-1 reference dumps).

Test infrastructure only.  Record = name[16] | count u64 | elem_size u32 | ncomp u32 | data.
"""
import struct
import numpy as np


def read_records(path):
    out = {}
    with open(path, "rb") as f:
        while True:
            hdr = f.read(32)
            if len(hdr) < 32:
                break
            name = hdr[:16].split(b"\0", 1)[0].decode()
            count, elem, ncomp = struct.unpack("<QII", hdr[16:32])
            real = {4: np.float32, 8: np.float64}[elem]
            data = np.frombuffer(f.read(count * elem * ncomp), dtype=real)
            if ncomp == 2:
                data = data.view({4: np.complex64, 8: np.complex128}[elem])
            out[name] = data.copy()
    return out

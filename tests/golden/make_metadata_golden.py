"""Writes tests/golden/metadata_ref.bin through the reference's OWN writer (src/metadata.cpp compiled unchanged into
oracle/_ref/libref_metadata.so by oracle/build_ref.sh) and reads it back through the reference's reader.
Run in the container that has /root/reference:  python tests/golden/make_metadata_golden.py"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

# (label, filename, x, y, w, h): ordinary records, a long and an empty filename, negative and large values
RECORDS = [(0, "slice_0.png", 0, 0, 224, 17), (1, "slice_1.png", 3, 17, 61, 40), (7, "slice_7.png", 100, 5, 1, 1),
           (12345, "slice_12345.png", 2047, 2047, 1, 1), (-3, "", -1, -2, 0, 0), (2, "a" * 300 + ".png", 5, 6, 7, 8)]
IMAGE_W, IMAGE_H = 2048, 2048


def main():
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_metadata.so"))
    n = len(RECORDS)
    cols = [np.array([r[i] for r in RECORDS], np.int32) for i in (0, 2, 3, 4, 5)]
    names = b"".join(r[1].encode() + b"\0" for r in RECORDS)
    path = os.path.join(HERE, "metadata_ref.bin")
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert L.refmeta_write(path.encode(), n, *[p(c) for c in cols], names, C.c_uint32(IMAGE_W), C.c_uint32(IMAGE_H)) == 0
    back = [np.zeros(n, np.int32) for _ in range(5)]
    nb = C.create_string_buffer(4096)
    iw, ih = C.c_uint32(), C.c_uint32()
    assert L.refmeta_read(path.encode(), n, *[p(b) for b in back], nb, 4096, C.byref(iw), C.byref(ih)) == n
    assert (iw.value, ih.value) == (IMAGE_W, IMAGE_H) and all(np.array_equal(a, b) for a, b in zip(cols, back))
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()

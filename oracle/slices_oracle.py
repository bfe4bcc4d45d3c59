"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference's slice writer, its metadata.bin sidecar and its
decoder.  Only tests/ may import this; the product path never does.

Pinning: the sidecar functions are checked byte for byte against tests/golden/metadata_ref.bin, which the reference's
own src/metadata.cpp wrote (oracle/_ref/libref_metadata.so, tests/golden/make_metadata_golden.py).  The slicer and the
decoder are OpenCV C++ code, which this image cannot compile; they are restated from the source and held against the
reference's own sequence of cv:: calls made through Python cv2, the same library behind a binding
(tests/test_slicer_vs_opencv.py: ROI + clone + copyTo(mask), split + threshold + copyTo(mask), imencode/imdecode).  The
masking they do is the same Rule P the estimator applies, and that one is pinned through the reference's estimator outputs.
"""
import struct

import numpy as np

MAGIC = 0x534C4943          # "SLIC", include/metadata.h:14


def compute_bounding_box(mask_u8):
    """src/image_slicer.cpp:15-42: (x, y, w, h) of the non-zero pixels, or (0, 0, 0, 0) when there are none."""
    ys, xs = np.nonzero(mask_u8)
    if ys.size == 0:
        return 0, 0, 0, 0
    return int(xs.min()), int(ys.min()), int(xs.max() - xs.min() + 1), int(ys.max() - ys.min() + 1)


def slice_image(image_hw4, mask_hw, label):
    """src/image_slicer.cpp:48-74: crop to the bbox of (mask == label), copy the masked pixels onto zeros.
    Returns (slice [h,w,4] or None, (x, y, w, h))."""
    binary = (mask_hw == label).astype(np.uint8)                     # :51
    x, y, w, h = compute_bounding_box(binary)                        # :54
    if w == 0 or h == 0:
        return None, (0, 0, 0, 0)                                    # :55-57
    crop = image_hw4[y:y + h, x:x + w]                               # :59
    m = binary[y:y + h, x:x + w]                                     # :62
    out = np.zeros_like(crop)                                        # :64
    out[m != 0] = crop[m != 0]                                       # :66
    return out, (x, y, w, h)


def write_slices(image_hw4, mask_hw):
    """src/image_slicer.cpp:81-125 without the files: one record per label 0..max that occurs, ascending label
    (the reference's order is whichever thread finishes first; the decoder does not depend on it for a partition).
    Returns [(label, filename, x, y, w, h, slice)]."""
    out = []
    for label in range(int(mask_hw.max()) + 1):                      # :91,96
        sl, (x, y, w, h) = slice_image(image_hw4, mask_hw, label)
        if sl is None:
            continue                                                 # :105
        out.append((label, "slice_%d.png" % label, x, y, w, h, sl))  # :107,111-117
    return out


def metadata_bytes(records, image_w, image_h):
    """src/metadata.cpp:4-34 with the packed structs of include/metadata.h:12-28.  records: (label, filename, x, y, w, h, ...)."""
    b = struct.pack("<IIII", MAGIC, len(records), image_w, image_h)
    for r in records:
        name = r[1].encode()
        b += struct.pack("<iiiiiH", r[0], r[2], r[3], r[4], r[5], len(name)) + name
    return b


def read_metadata(b):
    """src/metadata.cpp:36-80 -> ([(label, filename, x, y, w, h)], image_w, image_h)."""
    magic, count, iw, ih = struct.unpack_from("<IIII", b, 0)
    if magic != MAGIC:
        raise RuntimeError("Invalid metadata file (magic mismatch)")
    off, out = 16, []
    for _ in range(count):
        label, x, y, w, h, n = struct.unpack_from("<iiiiiH", b, off)
        off += 22
        out.append((label, b[off:off + n].decode(), x, y, w, h))
        off += n
    return out, iw, ih


def reassemble(records, slices, image_w, image_h):
    """src/reassemble.cpp:50-107: paste each slice at (x, y), clipped to the canvas, where its alpha > 0, in record order."""
    canvas = np.zeros((image_h, image_w, 4), np.uint8)               # :50
    for (label, name, x, y, w, h), sl in zip(records, slices):
        cw, ch = min(sl.shape[1], image_w - x), min(sl.shape[0], image_h - y)   # :58-59 (the actual slice size is used)
        if cw <= 0 or ch <= 0:
            continue                                                 # :61-64
        src = sl[:ch, :cw]
        m = src[..., 3] > 0                                          # :80-85
        canvas[y:y + ch, x:x + cw][m] = src[m]                       # :97
    return canvas

"""The slicer and the decoder of the reference are OpenCV C++ code, which this image cannot compile.  Python cv2 is the
same library behind a binding, so the numpy restatement (oracle/slices_oracle.py) is held here against the reference's
own sequence of cv:: calls made through cv2: ROI + clone + copyTo(mask) for the slice (src/image_slicer.cpp:59-66),
split + threshold + copyTo(mask) for the paste (src/reassemble.cpp:80-111), imencode/imdecode at the reference's PNG
level for the file (src/image_writer.cpp:7, include/configuration.h:11).  CPU only."""
import numpy as np
import pytest

from image_compression_cpp_b200 import synth
from oracle import slices_oracle as so

cv2 = pytest.importorskip("cv2")


def _cv_slice_image(bgra, mask, label):
    """src/image_slicer.cpp:48-74 with cv2 calls where the reference makes cv:: calls."""
    binary = (mask == label).astype(np.uint8)                                   # :51 (torch in the reference)
    ys, xs = np.nonzero(binary)                                                 # :15-42 is a hand-written scan, not OpenCV
    if ys.size == 0:
        return None, (0, 0, 0, 0)
    x, y, w, h = int(xs.min()), int(ys.min()), int(xs.max() - xs.min() + 1), int(ys.max() - ys.min() + 1)
    cropped = bgra[y:y + h, x:x + w].copy()                                     # input(box).clone(), :59
    cropped_mask = binary[y:y + h, x:x + w].copy()                              # mask_mat(box).clone(), :62
    output = np.zeros_like(cropped)                                             # cv::Scalar(0,0,0,0), :64
    cv2.copyTo(cropped, cropped_mask, output)                                   # cropped.copyTo(output, cropped_mask), :66
    return output, (x, y, w, h)


def _cv_reassemble(records, slices, W, H):
    """src/reassemble.cpp:50-111."""
    canvas = np.zeros((H, W, 4), np.uint8)
    for (label, name, x, y, w, h), sl in zip(records, slices):
        cw, ch = min(sl.shape[1], W - x), min(sl.shape[0], H - y)
        if cw <= 0 or ch <= 0:
            continue
        roi = canvas[y:y + ch, x:x + cw]
        src = np.ascontiguousarray(sl[:ch, :cw])
        alpha = cv2.split(src)[3]                                               # :80-83
        _, alpha_mask = cv2.threshold(alpha, 0, 255, cv2.THRESH_BINARY)         # :87
        out = np.ascontiguousarray(roi)
        cv2.copyTo(src, alpha_mask, out)                                        # :111
        roi[...] = out
    return canvas


@pytest.mark.parametrize("seed,H,W", [(1, 40, 56), (2, 33, 47), (3, 64, 64)])
def test_slice_restatement_equals_the_opencv_call_sequence(seed, H, W):
    rng = np.random.default_rng(seed)
    bgra = rng.integers(0, 256, (H, W, 4), dtype=np.uint8)
    bgra[..., 3] = 255                                                          # cv::COLOR_BGR2BGRA, src/compress.cpp:145
    for mask in (synth.bfs_partition(H, W, 7, seed=seed), synth.voronoi_partition(H, W, cell=9, jitter=3, seed=seed)):
        recs, slices = [], []
        for label in range(int(mask.max()) + 2):                                # one label past the end: the empty case
            want, box = _cv_slice_image(bgra, mask, label)
            got, gbox = so.slice_image(bgra, mask, label)
            assert box == gbox
            assert (want is None) == (got is None)
            if want is None:
                continue
            assert np.array_equal(want, got)
            # the file the reference writes: PNG level 4 through cv::imwrite; what comes back is what went in
            ok, png = cv2.imencode(".png", want, [cv2.IMWRITE_PNG_COMPRESSION, 4])
            assert ok and np.array_equal(cv2.imdecode(png, cv2.IMREAD_UNCHANGED), want)
            recs.append((label, "slice_%d.png" % label) + box)
            slices.append(want)
        assert np.array_equal(_cv_reassemble(recs, slices, W, H), so.reassemble(recs, slices, W, H))
        assert np.array_equal(so.reassemble(recs, slices, W, H), bgra)          # a partition's slices tile the image


def test_paste_restatement_equals_opencv_on_overlapping_and_clipped_slices():
    rng = np.random.default_rng(9)
    W, H = 30, 20
    slices = [rng.integers(0, 256, (6, 8, 4), dtype=np.uint8) for _ in range(4)]
    for s in slices:
        s[..., 3] = np.where(rng.random(s.shape[:2]) < 0.5, 0, s[..., 3])        # holes: alpha 0 leaves the canvas alone
    recs = [(0, "a", 2, 3, 8, 6), (1, "b", 5, 5, 8, 6), (2, "c", W - 3, H - 2, 8, 6), (3, "d", W, 0, 8, 6)]
    assert np.array_equal(_cv_reassemble(recs, slices, W, H), so.reassemble(recs, slices, W, H))

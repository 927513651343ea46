"""Parity of the CUDA extractor (liborbx.so through the C ABI) with the CPU oracle, stage by stage and end to end.
Bit-exact everywhere: pyramid bytes, FAST candidates (order included), per-level keypoint counts, blurred levels,
keypoint fields (angles included: the device fastAtan2 is op-for-op) and descriptor bytes."""
import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth

pytestmark = pytest.mark.gpu

CONFIGS = [  # (w, h, nfeatures)  BASELINE.json configs 1, 2, 3, 5 + the mono-init extractor (Tracking.cc:125)
    (640, 480, 1000), (640, 480, 2000), (752, 480, 1200), (1241, 376, 2000), (1920, 1080, 4000)]


def _compare_frame(ex, oe, img, stages=True, frame=0, result=None):
    okps, odesc = oe(img)
    if result is None:
        kps, desc = ex(img)
    else:
        kps, desc = result
    msgs = []
    if stages:
        for l in range(ex.nlevels):
            assert ex.level_size(l) == oe.level_size(l)
            if not np.array_equal(ex.pyramid_level(l, frame, border=19), oe.level_padded(l)):
                msgs.append("pyramid level %d differs" % l)
            gc, oc = ex.level_candidates(l, frame), oe.level_candidates(l)
            if not np.array_equal(gc, oc):
                msgs.append("FAST candidates level %d differ (%d vs %d)" % (l, len(gc), len(oc)))
            if ex.level_count(l, frame) != oe.level_nkeys(l):
                msgs.append("quadtree count level %d: %d vs %d" % (l, ex.level_count(l, frame), oe.level_nkeys(l)))
            if not np.array_equal(ex.blurred_level(l, frame), oe.level_blurred(l)):
                msgs.append("blurred level %d differs" % l)
    if len(kps) != len(okps):
        msgs.append("keypoint count %d vs %d" % (len(kps), len(okps)))
    else:
        for fld in KPF:
            if not np.array_equal(kps[fld], okps[fld]):
                bad = np.flatnonzero(kps[fld] != okps[fld])
                msgs.append("field %s differs at %d keypoints (first %d: %r vs %r)" % (fld, len(bad), bad[0], kps[fld][bad[0]], okps[fld][bad[0]]))
        if not np.array_equal(desc, odesc):
            bits = int(np.unpackbits(desc ^ odesc).sum())
            msgs.append("descriptors differ: %d bits in %d rows" % (bits, int((desc != odesc).any(axis=1).sum())))
    assert not msgs, "; ".join(msgs)
    return len(kps)


KPF = ["x", "y", "size", "angle", "response", "octave", "class_id"]


def test_tables_match_oracle():
    for nf, sf, nl in [(1000, 1.2, 8), (2000, 1.2, 8), (1200, 1.2, 8), (4000, 1.2, 8), (500, 1.5, 4), (777, 1.1, 10)]:
        ex = orbx.ORBextractor(nf, sf, nl, 20, 7)
        oe = O.OracleExtractor(nf, sf, nl, 20, 7)
        for a, b in zip(ex._tables(), oe.scale_tables()):
            assert a.tobytes() == b.tobytes()
        assert list(ex.features_per_level()) == list(oe.features_per_level())
        assert list(ex.umax()) == list(oe.umax())
        assert ex.GetLevels() == nl


@pytest.mark.parametrize("w,h,nf", CONFIGS)
def test_rects_frame_stage_by_stage(w, h, nf):
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    seed = {640: 1000, 752: 2000, 1241: 3000, 1920: 5000}[w]
    n = _compare_frame(ex, oe, synth.rects(w, h, seed))
    assert n > nf // 2


@pytest.mark.parametrize("kind", ["texture", "lowcontrast", "flat"])
def test_other_kinds_640x480(kind):
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    n = _compare_frame(ex, oe, synth.frame(kind, 640, 480, 1100))
    if kind == "flat":
        assert n == 0


def test_many_frames_end_to_end():
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    for i in range(12):
        _compare_frame(ex, oe, synth.rects(640, 480, 1000 + i), stages=False)


def test_batch_equals_single_and_oracle():
    B = 6
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    kinds = ["rects", "texture", "rects", "lowcontrast", "flat", "rects"]
    imgs = [synth.frame(k, 640, 480, 1200 + i) for i, k in enumerate(kinds)]
    res = ex.extract_batch(imgs)
    for b in range(B):
        _compare_frame(ex, oe, imgs[b], stages=(b in (1, 3)), frame=b, result=res[b])


def test_strided_input_and_shape_change():
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    big = synth.rects(800, 600, 77)
    view = big[40:40 + 480, 60:60 + 640]          # a sub-view, like a cv::Mat ROI
    kps, desc = ex(view)
    _compare_frame(ex, oe, np.ascontiguousarray(view), stages=False, result=(kps, desc))
    _compare_frame(ex, oe, synth.rects(752, 480, 5), stages=False)     # re-plans for the new shape
    _compare_frame(ex, oe, synth.rects(320, 240, 6), stages=True)


def test_small_and_odd_parameters():
    for (w, h, nf, sf, nl, ini, mn) in [(320, 240, 500, 1.2, 8, 20, 7), (401, 301, 300, 1.5, 4, 30, 10), (960, 540, 1500, 1.1, 10, 15, 5),
                                        (640, 480, 50, 1.2, 8, 20, 7), (300, 400, 400, 1.2, 4, 20, 7)]:
        ex = orbx.ORBextractor(nf, sf, nl, ini, mn)
        oe = O.OracleExtractor(nf, sf, nl, ini, mn)
        _compare_frame(ex, oe, synth.rects(w, h, w + h))
        _compare_frame(ex, oe, synth.texture(w, h, w), stages=False)


def test_empty_image_and_errors():
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    kps, desc = ex(np.zeros((0, 0), np.uint8))
    assert len(kps) == 0 and desc.shape == (0, 32)
    with pytest.raises(orbx.OrbxError):      # smaller than one FAST cell at the top level: the reference divides by zero
        ex(np.zeros((64, 64), np.uint8))
    with pytest.raises(orbx.OrbxError):
        orbx.ORBextractor(1000, 1.2, 99, 20, 7)


def test_fma_mode_matches_contracted_oracle():
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    ex.set_fma_mode(1)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7, fma_mode=1)
    _compare_frame(ex, oe, synth.rects(640, 480, 1003), stages=False)


def test_device_resident_input():
    import torch
    B = 4
    imgs = np.stack([synth.rects(640, 480, 1300 + i) for i in range(B)])
    d = torch.from_numpy(imgs).cuda()
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    torch.cuda.synchronize()
    ex.extract_device(d.data_ptr(), B, 640, 480, 640, 640 * 480)
    res = ex.fetch()
    for b in range(B):
        _compare_frame(ex, oe, imgs[b], stages=False, result=res[b])


def test_device_input_with_4_byte_alignment_only():
    """Level 0 given as a device image whose rows are only 4-byte aligned: the FAST cell tiles are then staged by 4-byte
    copies (extract_kernels.cu, fast_cell_body) and must give the same bytes as the 16-byte path."""
    import torch
    B, W, H, pitch = 3, 640, 480, 644
    imgs = np.stack([synth.rects(W, H, 1400 + i) for i in range(B)])
    buf = torch.zeros(4 + B * H * pitch, dtype=torch.uint8, device="cuda")
    view = buf[4:].view(B, H, pitch)                       # base pointer 4 bytes past a 256-byte boundary
    view[:, :, :W] = torch.from_numpy(imgs).cuda()
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    torch.cuda.synchronize()
    assert view.data_ptr() % 16 == 4
    ex.extract_device(view.data_ptr(), B, W, H, pitch, pitch * H)
    res = ex.fetch()
    for b in range(B):
        _compare_frame(ex, oe, imgs[b], stages=False, result=res[b])


def test_levels_with_one_or_two_cell_columns():
    """Levels whose active width is just under 60 px have ONE cell column of up to 59 px; widths just under 90 px give two
    columns of 44-45 px: the widest FAST cell tiles there are (with the reference's 30-px cells a tile row never exceeds
    64 bytes from its 16-byte-aligned origin, so the 128-byte tile form of the kernel is a guard, not a reachable path)."""
    for (w, h) in [(91, 140), (105, 121), (121, 92), (150, 91), (178, 96)]:
        ex = orbx.ORBextractor(200, 1.2, 2, 20, 7)
        oe = O.OracleExtractor(200, 1.2, 2, 20, 7)
        _compare_frame(ex, oe, synth.rects(w, h, 3 * w + h))
        _compare_frame(ex, oe, synth.texture(w, h, w + 7), stages=False)


@pytest.mark.parametrize("form", [1, 2, 3])
def test_every_quadtree_kernel_form(form):
    """The quadtree kernel has three CTA sizes picked per launch (small batches in the test suite always get the 256-thread
    one): force each and compare everything.  The KITTI shape has levels with more candidates than the 128-thread form keeps
    in shared memory (global-workspace path)."""
    for (w, h, nf, seed) in [(640, 480, 1000, 1500), (1241, 376, 2000, 1501)]:
        ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
        ex.set_octree_form(form)
        oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
        _compare_frame(ex, oe, synth.rects(w, h, seed))
        _compare_frame(ex, oe, synth.texture(w, h, seed + 10), stages=False)


def test_batch_call_on_page_locked_arrays_from_the_library():
    """orbx_alloc_pinned / orbx_free_pinned (include/orbx.h): frames back to back in write-combined page-locked memory travel as
    one copy, results land directly in page-locked result arrays laid out with the handle's per-frame capacity."""
    import ctypes as C
    B, W, H = 5, 640, 480
    imgs = np.stack([synth.rects(W, H, 1600 + i) for i in range(B)])
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    ex(imgs[0])                                            # plans the shape: max_keypoints is exact afterwards
    cap = ex.max_keypoints()
    fin = orbx.PinnedArray(imgs.shape, np.uint8, write_combined=True)
    fin.array[...] = imgs
    kp = orbx.PinnedArray((B, cap, 28), np.uint8)
    de = orbx.PinnedArray((B, cap, 32), np.uint8)
    n = np.zeros(B, np.int32)
    L = orbx.lib()
    ptrs = (C.c_void_p * B)(*[fin.ptr + i * W * H for i in range(B)])
    rc = L.orbx_extract_batch(ex._h, ptrs, B, W, H, W, C.c_void_p(kp.ptr), C.c_void_p(de.ptr), cap, n.ctypes.data_as(C.c_void_p))
    assert rc == 0, L.orbx_last_error().decode()
    kps = kp.array.view(orbx.KP_DTYPE).reshape(B, cap)
    for b in range(B):
        ok, od = oe(imgs[b])
        assert n[b] == len(ok) and kps[b, :n[b]].tobytes() == ok.tobytes() and np.array_equal(de.array[b, :n[b]], od)
    for a in (fin, kp, de):
        a.free()


def test_large_batch_runs_as_two_halves_on_two_streams():
    """From 128 VGA frames' worth of pixels on, a call is split over the handle's two streams (orbx_extractor.cu, launch_all): every
    frame of both halves, odd batch size included, must still equal the oracle."""
    distinct = [synth.rects(640, 480, 60 + i) for i in range(5)]
    oe = O.OracleExtractor(500, 1.2, 8, 20, 7)
    want = [oe(f) for f in distinct]
    ex = orbx.ORBextractor(500, 1.2, 8, 20, 7, max_batch=131)
    for B in (131, 128):
        got = ex.extract_batch([distinct[i % 5] for i in range(B)])
        assert ex.last_launches() == 22
        for i, (k, d) in enumerate(got):
            assert k.tobytes() == want[i % 5][0].tobytes() and np.array_equal(d, want[i % 5][1]), (B, i)
    got = ex.extract_batch(distinct)            # a small batch afterwards on the same handle: one chain again
    assert ex.last_launches() == 11
    for i, (k, d) in enumerate(got):
        assert k.tobytes() == want[i][0].tobytes() and np.array_equal(d, want[i][1])

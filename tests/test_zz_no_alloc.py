"""-m gpu test of svo_fast_detect_noalloc (include/svo.h): the allocation-free variant of svo_fast_detect whose
response map and row counts live in a static device array.  It has no counterpart in the reference; the checker is
the existing entry point, and the bar is the one FAST has everywhere else in this suite: identical keypoints and
responses, in the same order.  Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

import synthdata as synth
from stereo_visual_odometry_b200._lib import lib

pytestmark = pytest.mark.gpu

SVO_OK, SVO_ERR_CAPACITY = 0, -4
SCRATCH_BYTES = 4 << 20   # SVO_NOALLOC_SCRATCH_BYTES


def detect_alloc(L, img, thr, cap):
    """The existing entry point (host pointers, cudaMalloc per call)."""
    xy = np.zeros((cap, 2), np.float32)
    resp = np.zeros(cap, np.float32)
    n = L.svo_fast_detect(img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], thr,
                          xy.ctypes.data_as(C.POINTER(C.c_float)), resp.ctypes.data_as(C.POINTER(C.c_float)), cap)
    assert n >= 0, L.svo_last_error()
    return n, xy[:min(n, cap)], resp[:min(n, cap)]


class DeviceBuffers:
    """The caller's side of the new entry point: inputs and outputs are the caller's device memory."""

    def __init__(self, torch, h, pitch, cap):
        self.torch = torch
        self.img = torch.zeros((h, pitch), dtype=torch.uint8, device="cuda")
        self.xy = torch.zeros((cap, 2), dtype=torch.float32, device="cuda")
        self.resp = torch.zeros(cap, dtype=torch.float32, device="cuda")
        self.n = torch.zeros(1, dtype=torch.int32, device="cuda")
        self.cap = cap

    def detect(self, L, img, thr):
        t = self.torch
        h, w = img.shape
        self.img[:h, :w].copy_(t.from_numpy(img))
        self.xy.fill_(-1.0)
        self.resp.fill_(-1.0)
        self.n.fill_(-1)
        t.cuda.synchronize()
        stream = t.cuda.current_stream().cuda_stream
        rc = L.svo_fast_detect_noalloc(self.img.data_ptr(), w, h, self.img.stride(0), thr, self.xy.data_ptr(),
                                       self.resp.data_ptr(), self.cap, self.n.data_ptr(), stream)
        assert rc == SVO_OK, L.svo_last_error()
        t.cuda.synchronize()
        n = int(self.n.item())
        m = min(n, self.cap)
        return n, self.xy[:m].cpu().numpy(), self.resp[:m].cpu().numpy()


def test_noalloc_matches_fast_detect_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    L = lib()
    seq = synth.make_sequence(width=1241, height=376, calib="kitti00", n_frames=2, seed=3)
    big0, big1 = seq.render(0)[0], seq.render(1)[1]
    rng = np.random.default_rng(7)
    small = rng.integers(0, 256, (131, 257), dtype=np.uint8)
    cap = 1 << 16
    dev = DeviceBuffers(torch, 376, 1280, cap)
    # two calls in a row on the shared static scratch, then a smaller image (stale map rows beyond its height and
    # stale row counts stay in the array), then the first image again, then a cap the count exceeds
    for img, thr in ((big0, 20), (big1, 20), (small, 5), (big0, 20), (big1, 60)):
        img = np.ascontiguousarray(img)
        n_ref, xy_ref, resp_ref = detect_alloc(L, img, thr, cap)
        n, xy, resp = dev.detect(L, img, thr)
        assert n_ref > 0 and n == n_ref, (img.shape, thr, n, n_ref)
        assert np.array_equal(xy, xy_ref) and np.array_equal(resp, resp_ref), (img.shape, thr)
    small_cap = DeviceBuffers(torch, 376, 1280, 100)
    n_ref, xy_ref, resp_ref = detect_alloc(L, np.ascontiguousarray(big0), 20, 100)
    n, xy, resp = small_cap.detect(L, np.ascontiguousarray(big0), 20)
    assert n == n_ref > 100 and np.array_equal(xy, xy_ref) and np.array_equal(resp, resp_ref)


def test_noalloc_unaligned_pitch_and_side_stream():
    """A pitch that rules the TMA tile loads out (the kernel's plain-load path) on a non-default stream."""
    torch = pytest.importorskip("torch")
    L = lib()
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (200, 333), dtype=np.uint8)
    n_ref, xy_ref, resp_ref = detect_alloc(L, img, 10, 1 << 15)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        dev = DeviceBuffers(torch, 200, 333, 1 << 15)
        for _ in range(2):
            n, xy, resp = dev.detect(L, img, 10)
            assert n == n_ref and np.array_equal(xy, xy_ref) and np.array_equal(resp, resp_ref)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    torch = pytest.importorskip("torch")
    L = lib()
    d = torch.zeros(16, dtype=torch.uint8, device="cuda")   # never dereferenced: the size check comes first
    w, h = 4096, 2048
    assert ((w + 127) & ~127) * h > SCRATCH_BYTES
    rc = L.svo_fast_detect_noalloc(d.data_ptr(), w, h, w, 20, d.data_ptr(), d.data_ptr(), 4, d.data_ptr(), None)
    assert rc == SVO_ERR_CAPACITY
    msg = L.svo_last_error().decode()
    assert "scratch" in msg and str(SCRATCH_BYTES) in msg, msg
    # the largest documented size passes the check (2048 x 2040: 4177920 + 8164 bytes)
    assert 2048 * 2040 + 4 * 2041 <= SCRATCH_BYTES

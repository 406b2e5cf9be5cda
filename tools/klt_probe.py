"""Debug tool (probe build: make -C stereo_visual_odometry_b200/csrc probe; SVO_LIB_PATH=tools/libsvo_probe.so):
per-feature duration / LK iterations of klt_kernel in the synchronous path."""
import ctypes as C
import sys

import numpy as np

sys.path.insert(0, '.')
import torch  # noqa: E402

import synthdata as synth  # noqa: E402
from stereo_visual_odometry_b200._lib import check, default_params, lib, ptr  # noqa: E402

W, H, N = 1241, 376, 12
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=N, seed=0)
Ls, Rs = zip(*[seq.render(i) for i in range(N)])
Lp = torch.from_numpy(np.stack(Ls)).pin_memory()
Rp = torch.from_numpy(np.stack(Rs)).pin_memory()
L_ = lib()
p = default_params(W, H)
ctx = L_.svo_create(C.byref(p))
Pl, Pr = np.ascontiguousarray(seq.Pl), np.ascontiguousarray(seq.Pr)
check(L_.svo_set_projection(ctx, ptr(Pl, C.c_float), ptr(Pr, C.c_float)))
T = np.zeros(16)
for j in range(N):
    check(L_.svo_process_frame(ctx, Lp[j].data_ptr(), Rp[j].data_ptr(), W, ptr(T, C.c_double)))
buf = np.zeros(4 * 16384, np.int32)
L_.svo_debug_klt_probe.argtypes = [C.c_void_p, C.c_int]
L_.svo_debug_klt_probe(buf.ctypes.data, buf.size)
a = buf.reshape(-1, 4)
a = a[a[:, 1] > 0]
st = a[:, 0].astype(np.uint32).astype(np.int64)
st = (st - st.min()) % (1 << 32)
dur, it = a[:, 1].astype(np.float64), a[:, 2].astype(np.float64)
print("features %d  span %.1f us  dur mean %.1f max %.1f us  iters mean %.1f max %d" %
      (len(a), (st + dur).max() / 1e3, dur.mean() / 1e3, dur.max() / 1e3, it.mean(), it.max()))
A = np.stack([it, np.ones_like(it)], 1)
coef = np.linalg.lstsq(A, dur, rcond=None)[0]
print("fit: dur = %.1f ns * iters + %.1f ns" % (coef[0], coef[1]))
order = np.argsort(-(st + dur))[:12]
for i in order:
    print("  end %.1f us  start %.1f  dur %.1f us  iters %d  sm %d" % ((st[i] + dur[i]) / 1e3, st[i] / 1e3, dur[i] / 1e3, it[i], a[i, 3]))
for lo, hi in ((0, 50), (50, 100), (100, 150), (150, 200), (200, 300), (300, 481)):
    m = (it >= lo) & (it < hi)
    if m.any():
        print("  iters [%3d,%3d): n %5d  dur mean %.1f us  (%.0f ns/iter)" % (lo, hi, m.sum(), dur[m].mean() / 1e3, (dur[m] / np.maximum(it[m], 1)).mean()))
# start-time profile: how many features start late (second features of a warp)
print("  started after 20 us: %d, after 100 us: %d" % ((st > 20e3).sum(), (st > 100e3).sum()))
L_.svo_destroy(ctx)

import sys, time, ctypes as C, numpy as np
sys.path.insert(0, '.')
from oracle import synth
from stereo_visual_odometry_b200 import vo as V
from stereo_visual_odometry_b200._lib import lib, check
W, H = 1241, 376
seq = synth.make_sequence(width=W, height=H, calib="kitti00", n_frames=12, seed=0)
fr = [seq.render(i) for i in range(12)]
vo = V.VisualOdometry(W, H)
vo.initalize_projection_matricies(seq.Pl, seq.Pr)
info = (C.c_int * 8)()
for i, (L, R) in enumerate(fr):
    t0 = time.perf_counter()
    ok, T = vo.stereo_callback(L, R)
    dt = time.perf_counter() - t0
    check(lib().svo_pipeline_info(vo._ctx, info))
    r = vo.last_result()
    print(i, ok, 'bucketed', r.n_bucketed, 'matched', r.n_matched, 'inl', r.n_inliers, 'info', list(info), 'ms %.3f' % (dt * 1e3))

"""The `vo N folder` command-line driver (stereo_visual_odometry_b200/cli; reference: src/main.cpp:315-407).

CPU part: the program builds without OpenCV, its image reader returns cv::imread's pixels, and a run without a
GPU fails loudly.  GPU part: a run over the reference's bundled run1 frames writes the reference's own
run1/result.csv rows (CLI-faithful 8UC3 mode, identity initial pose), with and without frames in flight."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, has_gpu

CLI_DIR = os.path.join(ROOT, "stereo_visual_odometry_b200", "cli")
EXE = os.path.join(CLI_DIR, "vo")


def build():
    subprocess.run(["make", "-C", CLI_DIR], check=True, capture_output=True)


def write_run1(tmp_path, n=8, gt=True):
    z = np.load(os.path.join(ROOT, "tests", "golden", "run1_first8.npz"))
    d = tmp_path / "run1"
    for cam in ("left", "right"):
        (d / cam).mkdir(parents=True, exist_ok=True)
        for i in range(n):
            (d / cam / f"frame{i:06d}.png").write_bytes(z[f"{cam}_{i}"].tobytes())
    if gt:   # layout of the reference's run1/gt.csv
        rows = ["time,x,y,dx,dy"] + [f"{0.5 * i:.4f},{3.8354 + 0.001 * i:.6g},{0.0145 + 0.002 * i:.6g},0,0" for i in range(n)]
        (d / "gt.csv").write_text("\n".join(rows) + "\n")
    return d, z["result_xyz"]


def read_ppm(path):
    import cv2
    return cv2.imread(str(path), cv2.IMREAD_COLOR)


def test_cli_builds_and_decodes_like_imread(tmp_path):
    import cv2
    build()
    d, _ = write_run1(tmp_path, n=2, gt=False)
    src = d / "left" / "frame000001.png"
    subprocess.run([EXE, "--dump-image", str(src), str(tmp_path / "a.ppm")], check=True)
    assert np.array_equal(read_ppm(tmp_path / "a.ppm"), cv2.imread(str(src)))
    # other PNG colour types / filters and PNM, against cv2.imread(IMREAD_COLOR)
    rng = np.random.default_rng(0)
    gray = cv2.GaussianBlur(rng.integers(0, 256, (37, 53), dtype=np.uint8), (0, 0), 2.0)
    rgba = cv2.GaussianBlur(rng.integers(0, 256, (41, 29, 4), dtype=np.uint8), (0, 0), 1.5)
    cv2.imwrite(str(tmp_path / "g.png"), gray)
    cv2.imwrite(str(tmp_path / "rgba.png"), rgba)
    cv2.imwrite(str(tmp_path / "g.pgm"), gray)
    cv2.imwrite(str(tmp_path / "c.ppm"), rgba[:, :, :3])
    for name in ("g.png", "rgba.png", "g.pgm", "c.ppm"):
        subprocess.run([EXE, "--dump-image", str(tmp_path / name), str(tmp_path / "o.ppm")], check=True)
        assert np.array_equal(read_ppm(tmp_path / "o.ppm"), cv2.imread(str(tmp_path / name), cv2.IMREAD_COLOR)), name
    r = subprocess.run([EXE, "--dump-image", str(tmp_path / "missing.png"), str(tmp_path / "o.ppm")], capture_output=True)
    assert r.returncode != 0
    # --gray: cv::cvtColor(BGR2GRAY) bit for bit (15-bit fixed point), on a real frame and on random colours
    cv2.imwrite(str(tmp_path / "rand.png"), rng.integers(0, 256, (64, 96, 3), dtype=np.uint8))
    for path in (src, tmp_path / "rand.png"):
        subprocess.run([EXE, "--dump-gray", str(path), str(tmp_path / "g_out.pgm")], check=True)
        got = cv2.imread(str(tmp_path / "g_out.pgm"), cv2.IMREAD_UNCHANGED)
        assert np.array_equal(got, cv2.cvtColor(cv2.imread(str(path)), cv2.COLOR_BGR2GRAY)), path


@pytest.mark.skipif(has_gpu(), reason="only meaningful without a GPU")
def test_cli_fails_loudly_without_gpu(tmp_path):
    build()
    d, _ = write_run1(tmp_path, n=2)
    r = subprocess.run([EXE, "2", str(d)], capture_output=True, text=True)
    assert r.returncode != 0 and "error" in r.stderr


def load_csv(path):
    lines = open(path).read().strip().splitlines()
    assert lines[0] == "x,y,z,gtx,gty"   # main.cpp:347
    return np.array([[float(v) for v in ln.split(",")] for ln in lines[1:]])


@pytest.mark.gpu
def test_cli_reproduces_reference_result_csv(tmp_path):
    build()
    d, gold = write_run1(tmp_path)
    r = subprocess.run([EXE, "8", str(d), "--pose0", "identity"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    rows = load_csv(d / "result.csv")
    assert rows.shape == (8, 5)
    step = np.linalg.norm(np.diff(gold, axis=0), axis=1).max()
    assert np.abs(rows[:, :3] - gold).max() <= 1e-3 * max(step, 1e-3) * len(gold) + 2e-5
    # gt.csv pass-through: x, y of one row per processed frame (main.cpp:381-392)
    assert np.allclose(rows[:, 3], [3.8354 + 0.001 * i for i in range(8)], rtol=1e-5)
    assert np.allclose(rows[:, 4], [0.0145 + 0.002 * i for i in range(8)], rtol=1e-5)
    # frames in flight: the same file
    r = subprocess.run([EXE, "8", str(d), "--pose0", "identity", "--stream", "--out", str(tmp_path / "s.csv")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert open(tmp_path / "s.csv").read() == open(d / "result.csv").read()
    # today's CLI starts pitched down by 26 degrees (main.cpp:368-373): the same trajectory rotated
    r = subprocess.run([EXE, "8", str(d), "--out", str(tmp_path / "p.csv")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    th = np.deg2rad(26.0)
    P0 = np.array([[1, 0, 0], [0, np.cos(th), np.sin(th)], [0, -np.sin(th), np.cos(th)]])
    assert np.abs(load_csv(tmp_path / "p.csv")[:, :3] - rows[:, :3] @ P0.T).max() < 1e-5


@pytest.mark.gpu
def test_cli_gray_mode_calibration_yaml_and_missing_frames(tmp_path):
    build()
    d, _ = write_run1(tmp_path)
    os.remove(d / "left" / "frame000003.png")   # a missing left image skips the frame (main.cpp:376-378)
    for style in ("fx", "Camera.fx"):
        pre = "Camera." if style.startswith("Camera") else ""
        (tmp_path / "c.yaml").write_text(
            "%YAML:1.0\n# comment\n" + "".join(f"{pre}{k}: {v}\n" for k, v in
                                               [("fx", 322.11376), ("fy", 322.11376), ("cx", 327.47336), ("cy", 176.33722), ("bf", -22.5428)])
            + "ThDepth: 35\n")
        out = tmp_path / f"{style}.csv"
        r = subprocess.run([EXE, "8", str(d), "--gray", "--calib", str(tmp_path / "c.yaml"), "--out", str(out)],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout + r.stderr
    r = subprocess.run([EXE, "8", str(d), "--gray", "--out", str(tmp_path / "default.csv")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0
    a, b, c = (load_csv(tmp_path / n) for n in ("fx.csv", "Camera.fx.csv", "default.csv"))
    assert a.shape == (7, 5) and np.array_equal(a, b) and np.array_equal(a, c)
    assert np.isfinite(a).all() and np.abs(a[-1, :3]).max() > 0

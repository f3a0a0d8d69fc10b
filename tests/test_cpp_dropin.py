"""The C++ drop-in headers (include/Features/*.hpp, include/SlamTypes/Frame.hpp) keep the reference's
class API (Features/ORBextractor.hpp:55-158, Features/ORBmatcher.hpp:7-63, SlamTypes/Frame.hpp:20-136).
CPU: they compile against the C ABI.  GPU: the demo sequence of demo/demo_initialization.cpp:65-108
gives the oracle's keypoint counts, match count and vnMatches12."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _build(tmp_path):
    from orb_slam_tracking_b200 import build
    build.build()
    exe = tmp_path / "dropin_demo"
    lib = ROOT / "orb_slam_tracking_b200"
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-Werror", "-I" + str(ROOT / "include"),
                           str(ROOT / "tests/csrc/dropin_demo.cpp"), "-o", str(exe), "-L" + str(lib), "-lorbx",
                           "-Wl,-rpath," + str(lib)])
    return exe


def test_dropin_headers_compile(tmp_path):
    _build(tmp_path)


@pytest.mark.gpu
def test_dropin_demo_sequence(tmp_path, oracle):
    from orb_slam_tracking_b200 import synth
    exe = _build(tmp_path)
    W, H, nf = 752, 480, 1200
    imgs = synth.frame_pair(5, W, H)
    raw = tmp_path / "pair.raw"
    imgs.tofile(raw)
    out = subprocess.run([str(exe), str(raw), str(W), str(H), str(nf)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    res = [l for l in out.stdout.splitlines() if l.startswith("RESULT")][0].split()
    K4, D4 = synth.DEMO_K4, synth.DEMO_D4
    b = oracle.image_bounds(W, H, K4, D4)
    R = oracle.Extractor(nf, 1.2, 8, 20, 7)
    F = []
    for im in imgs:
        k, d, _ = R(im)
        F.append(oracle.Frame(k, d, K4, D4, b))
    n, m12, cnt = oracle.search_for_initialization(F[0], F[1], 0.9, True, 100)
    chk = 0
    for v in m12:
        chk = (chk * 31 + int(v) + 1) & 0xFFFFFFFFFFFFFFFF
    got_chk = int(res[4]) & 0xFFFFFFFFFFFFFFFF
    assert (int(res[1]), int(res[2]), int(res[3])) == (F[0].N, F[1].N, n)
    assert got_chk == chk
    assert int(res[5]) == oracle.descriptor_distance(F[0].desc[0], F[1].desc[0])
    area = [l for l in out.stdout.splitlines() if l.startswith("AREA")][0].split()
    assert int(area[1]) == len(F[1].features_in_area(300.0, 200.0, 50.0, 0, 0))
    assert int(area[7].split("x")[0]) == 627 and int(area[5]) == 8      # mvImagePyramid is filled like the reference's
    # the reference's own stdout lines (Features/ORBextractor.cpp:549, Features/ORBmatcher.cpp:144-147)
    assert "Sum of features = %d" % nf in out.stdout
    want = ("SearchForInitialization done ----------------------\ninvalidMatchByDistance: %d\ninvalidMatchByRatio: %d\n"
            "invalidMatchByOrientation: %d\n" % tuple(cnt))
    assert want in out.stdout
    st = [l for l in out.stdout.splitlines() if l.startswith("STATICS")][0].split()
    assert (int(st[3]), int(st[4]), int(st[5]), int(st[6])) == tuple(int(v) for v in b)       # Frame::mnMinX ... as data members
    assert abs(float(st[1]) - float(K4[0])) < 1e-3 and st[8] == "2" and st[10:] == ["0", "1"]   # nNextId, mnId
    assert [l for l in out.stdout.splitlines() if l.startswith("COPY")][0].split()[1:] == [str(F[1].N), "1", "1", "1"]
    assert [l for l in out.stdout.splitlines() if l.startswith("NODE")][0].split()[1:] == ["51", "26", "3", "0", "0", "2", "0"]
    assert [l for l in out.stdout.splitlines() if l.startswith("LATENCY")]

"""stdsort_emul() (orb_slam_tracking_b200/csrc/stdsort_emul.cuh) must produce the same PERMUTATION as
libstdc++'s std::sort with the reference's tying comparator (Features/ORBextractor.cpp:684-696,
912-913): the quadtree's split order depends on it.  Host compile of the same header the kernel uses."""
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def test_stdsort_emul_matches_libstdcxx(tmp_path):
    exe = tmp_path / "stdsort_check"
    subprocess.check_call(["g++", "-O2", "-std=c++17", str(ROOT / "tests/csrc/stdsort_check.cpp"), "-o", str(exe)])
    out = subprocess.run([str(exe), "6000"], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "0", out.stdout + out.stderr


def test_stdnth_emul_matches_libstdcxx(tmp_path):
    """stdnth_emul_t() == std::nth_element with retainBest's comparator (ComputeKeyPointsOld)."""
    exe = tmp_path / "stdnth_check"
    subprocess.check_call(["g++", "-O2", "-std=c++17", str(ROOT / "tests/csrc/stdnth_check.cpp"), "-o", str(exe)])
    out = subprocess.run([str(exe), "6000"], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "0", out.stdout + out.stderr

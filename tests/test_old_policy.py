"""ComputeKeyPointsOld (Features/ORBextractor.cpp:1207-1496), the selection policy the reference keeps behind a
commented-out call: the oracle's restatement (CPU) and the CUDA path (orbx_set_keypoint_policy(ctx, 1)) against outputs of
the reference's own function body (tests/golden/old_policy.npz, tools/gen_golden.py) and against each other."""
import zlib

import numpy as np
import pytest

from golden_util import old_policy_cases


def _crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes())


def test_oracle_old_policy_matches_reference_outputs(oracle):
    for name, im, nf, ini, mn, n, kc, dc in old_policy_cases():
        k, d, mono = oracle.Extractor(nf, 1.2, 8, ini, mn, policy=1)(im)
        assert (len(k), _crc(k), _crc(d)) == (n, kc, dc), name
        assert mono == n


@pytest.mark.gpu
def test_cuda_old_policy_matches_reference_outputs():
    from orb_slam_tracking_b200 import frontend as fe
    ctx = {}
    for name, im, nf, ini, mn, n, kc, dc in old_policy_cases():
        key = (im.shape, nf, ini, mn)
        if key not in ctx:
            ctx[key] = fe.ORBextractor(nf, 1.2, 8, ini, mn, im.shape[1], im.shape[0], max_batch=2)
            ctx[key].set_keypoint_policy(1)
        k, d, mono = ctx[key](im)
        assert (len(k), _crc(k), _crc(d)) == (n, kc, dc), name
        assert mono == n
    for E in ctx.values():
        E.close()


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(640, 480, 1000, 20, 7), (752, 480, 1200, 20, 7), (752, 480, 2000, 0, 0), (1241, 376, 2000, 25, 12),
                                 (640, 480, 500, 7, 20)])
def test_cuda_old_policy_against_the_oracle(oracle, cfg):
    """Stage results (level keypoints before orientation order, packed output with a lapping area, batches) under both
    policies on one context: switching back and forth leaves neither policy's results changed."""
    from orb_slam_tracking_b200 import frontend as fe, synth
    W, H, nf, ini, mn = cfg
    E = fe.ORBextractor(nf, 1.2, 8, ini, mn, W, H, max_batch=4)
    R = {p: oracle.Extractor(nf, 1.2, 8, ini, mn, policy=p) for p in (0, 1)}
    imgs = list(synth.frames(4, W, H, first_pair=11))
    imgs.append(np.full((H, W), 77, np.uint8))                                   # flat: no keypoints anywhere
    noise = np.random.default_rng(3).integers(0, 256, (H, W)).astype(np.uint8)   # corners everywhere, every cell over its quota
    imgs.append(noise)
    for policy in (1, 0, 1):
        E.set_keypoint_policy(policy)
        for i, im in enumerate(imgs):
            for lap in ((0, 0), (150, 400)):
                k, d, mono = E(im, vLappingArea=lap)
                rk, rd, rmono = R[policy](im, lap)
                assert len(k) == len(rk) and mono == rmono, (policy, i, lap, len(k), len(rk))
                assert np.array_equal(k.view(np.uint8), rk.view(np.uint8)), (policy, i, lap)
                assert np.array_equal(d, rd), (policy, i, lap)
        res = E.extract_batch(imgs[:4])
        for i in range(4):
            rk, rd, _ = R[policy](imgs[i])
            assert np.array_equal(res[i][0].view(np.uint8), rk.view(np.uint8)) and np.array_equal(res[i][1], rd), (policy, i)
    E.close()


# Shapes found by the wider fuzz run (tools/gpu/fuzz_stream.py).  (W, H, nfeatures, scale, levels, iniTh, minTh, synthetic pair)
WIDE_CELLS = [(875, 227, 900, 1.1, 6, 12, 7, 500), (818, 300, 849, 1.1, 4, 31, 3, 900)]   # cells of ~280 to ~350 tested columns
REFUSED = [(577, 125, 900, 1.1, 2, 12, 0), (816, 142, 900, 1.1, 5, 31, 7), (991, 170, 900, 1.2, 3, 12, 12)]


def test_oracle_old_policy_wide_cells_against_the_reference(oracle):
    """Cells several hundred columns wide (few cell columns on a wide level): the oracle's restatement against the reference's own
    function body (oracle/_ref) -- skipped where /root/reference is not available (the GPU box)."""
    from oracle import ref_orb
    from orb_slam_tracking_b200 import synth
    if not ref_orb.available():
        pytest.skip("oracle/_ref not built")
    for W, H, nf, sf, nl, ini, mn, pair in WIDE_CELLS:
        O = oracle.Extractor(nf, sf, nl, ini, mn, policy=1)
        Rf = ref_orb.OldPolicyExtractor(nf, sf, nl, ini, mn)
        for im in synth.frames(2, W, H, first_pair=pair):
            k, d, m = O(im)
            rk, rd, rm = Rf(im)
            assert len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd) and m == rm


def test_oracle_old_policy_refuses_shapes_outside_the_working_range(oracle):
    """A cell grid that runs past the tested region of a level (very wide levels with many short cell rows): the reference keeps
    keypoints closer than 19 px to the border there, reads their descriptors from outside the blurred level (two runs of oracle/_ref
    give different bits) and would assert in cv::Mat::rowRange with a real OpenCV.  The oracle reports -8."""
    from orb_slam_tracking_b200 import synth
    for W, H, nf, sf, nl, ini, mn in REFUSED:
        with pytest.raises(RuntimeError, match="status -8"):
            oracle.Extractor(nf, sf, nl, ini, mn, policy=1)(synth.frames(1, W, H)[0])


@pytest.mark.gpu
def test_cuda_old_policy_wide_cells_and_refused_shapes(oracle):
    from orb_slam_tracking_b200 import frontend as fe, synth
    from orb_slam_tracking_b200.capi import OrbxError, ERR_ARG
    for W, H, nf, sf, nl, ini, mn, pair in WIDE_CELLS:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H)
        E.set_keypoint_policy(1)
        O = oracle.Extractor(nf, sf, nl, ini, mn, policy=1)
        for im in synth.frames(2, W, H, first_pair=pair):
            k, d, m = E(im)
            rk, rd, rm = O(im)
            assert len(k) == len(rk) and np.array_equal(k.view(np.uint8), rk.view(np.uint8)) and np.array_equal(d, rd) and m == rm
        E.close()
    for W, H, nf, sf, nl, ini, mn in REFUSED:
        E = fe.ORBextractor(nf, sf, nl, ini, mn, W, H)
        with pytest.raises(OrbxError) as ei:
            E.set_keypoint_policy(1)
        assert ei.value.code == ERR_ARG
        k, d, m = E(synth.frames(1, W, H)[0])     # the context stays usable with the default policy
        assert len(k) > 0
        E.close()

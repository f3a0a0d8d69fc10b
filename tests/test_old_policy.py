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

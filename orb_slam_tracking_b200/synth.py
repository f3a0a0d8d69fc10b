"""Deterministic synthetic frames for tests and bench: a smooth low-frequency
background, random rectangles and discs with contrast U[25,120], Gaussian noise sigma 3.  A frame
PAIR is one scene seen through two crops shifted by U[-20,20] px, each with its own noise, so that
SearchForInitialization(window 100) has real matches to find.  numpy only."""
import numpy as np

SEED0 = 0xB200


def _scene(rng, W, H, margin=32):
    cw, ch = W + 2 * margin, H + 2 * margin
    yy, xx = np.mgrid[0:ch, 0:cw].astype(np.float32)
    img = np.full((ch, cw), 110.0, np.float32)
    for _ in range(6):
        fx, fy = rng.uniform(0.5, 3.0, 2) * 2 * np.pi / np.array([cw, ch])
        img += (40.0 / 6) * np.sin(fx * xx + fy * yy + rng.uniform(0, 2 * np.pi))
    n = int(400 * (W * H / 307200.0))
    for _ in range(n):
        c = rng.uniform(25, 120) * (1 if rng.random() < 0.5 else -1)
        x, y = int(rng.integers(0, cw)), int(rng.integers(0, ch))
        if rng.random() < 0.6:
            w, h = int(rng.integers(4, 40)), int(rng.integers(4, 40))
            img[y:y + h, x:x + w] += c
        else:
            r = int(rng.integers(3, 16))
            y0, y1, x0, x1 = max(0, y - r), min(ch, y + r + 1), max(0, x - r), min(cw, x + r + 1)
            m = (yy[y0:y1, x0:x1] - y) ** 2 + (xx[y0:y1, x0:x1] - x) ** 2 <= r * r
            img[y0:y1, x0:x1][m] += c
    return img


def frame_pair(index, W, H):
    """Frames 2*index and 2*index+1 of the synthetic stream, uint8 [2, H, W]."""
    rng = np.random.default_rng(SEED0 + index)
    margin = 32
    s = _scene(rng, W, H, margin)
    out = np.empty((2, H, W), np.uint8)
    for k in range(2):
        dx, dy = (0, 0) if k == 0 else (int(rng.integers(-20, 21)), int(rng.integers(-20, 21)))
        crop = s[margin + dy:margin + dy + H, margin + dx:margin + dx + W]
        out[k] = np.clip(np.rint(crop + rng.normal(0, 3, crop.shape).astype(np.float32)), 0, 255).astype(np.uint8)
    return out


def frames(n, W, H, first_pair=0):
    """n frames (n even: pairs first_pair .. first_pair + n/2 - 1), uint8 [n, H, W]."""
    out = np.empty((n, H, W), np.uint8)
    for p in range((n + 1) // 2):
        pr = frame_pair(first_pair + p, W, H)
        out[2 * p] = pr[0]
        if 2 * p + 1 < n:
            out[2 * p + 1] = pr[1]
    return out


def tile_frames(n, W, H, unique_pairs, first_pair=0):
    """n frames built by repeating `unique_pairs` distinct pairs (cheap to generate for large
    batches; every frame is still processed independently by the pipeline)."""
    base = frames(2 * unique_pairs, W, H, first_pair)
    reps = (n + len(base) - 1) // len(base)
    return np.ascontiguousarray(np.concatenate([base] * reps)[:n])


# camera of the reference's Settings.yaml (752x480); other shapes use zero distortion
DEMO_K4 = np.array([609.2855, 609.3422, 351.4274, 237.7324], np.float32)
DEMO_D4 = np.array([-0.3492, 0.1363, 0.0, 0.0], np.float32)


def camera_for(W, H):
    if (W, H) == (752, 480):
        return DEMO_K4.copy(), DEMO_D4.copy()
    f = 0.8 * W
    return np.array([f, f, W / 2.0, H / 2.0], np.float32), np.zeros(4, np.float32)

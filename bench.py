#!/usr/bin/env python
"""ORB front-end throughput: frames/s of extract + Frame glue + SearchForInitialization.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--workload tum]

A "step" is one pass of the whole hot path over one batch of B synthetic frames (B/2 frame
pairs).  `value` is measured with the batch already resident in HBM; `e2e` goes through the
same C ABI calls with host buffers (pinned H2D of the images, D2H of keypoints, descriptors and
matches inside the timed region).  N > 1: one process per GPU (torchrun), frame pairs sharded per
rank, no collective on the data path (weak scaling); the barrier and max-over-ranks timing use
torch.distributed.  --impl reference times the CPU restatement of the reference (oracle/) on the
host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # BASELINE.json configs[1..4]; configs[0] (demo pair) is a parity-test case
    "tum": dict(W=640, H=480, nfeatures=1000, ini=20, mn=7, batch=512, desc="synthetic TUM-shape 640x480, 1000 kp"),
    "euroc": dict(W=752, H=480, nfeatures=1200, ini=20, mn=7, batch=512, desc="synthetic EuRoC-shape 752x480, 1200 kp"),
    "kitti": dict(W=1241, H=376, nfeatures=2000, ini=20, mn=7, batch=256, desc="synthetic KITTI-shape 1241x376, 2000 kp"),
    # configs[4]: brute-force matching = SearchForInitialization with a window that covers the image
    "uhd": dict(W=3840, H=2160, nfeatures=8000, ini=20, mn=7, batch=64, window=8192,
                desc="synthetic 3840x2160, 8000 kp, all-pairs window"),
}
METRIC = "ORB frames/s (extract+match, 640x480, 1k kp)"
UNIT = "frames/s"
NNRATIO, WINDOW = 0.9, 100  # WINDOW: default; a workload may carry its own
STAGES = ["pyramid", "fast", "quadtree", "blur", "describe", "pack", "glue", "match"]


def level_sizes(W, H, nlevels=8, sf=1.2):
    s, out = np.float32(1.0), []
    for l in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(sf)))
    return out


def algorithmic_bytes(W, H, K):
    """per-frame algorithmic bytes of each stage (u8 images, no pads)."""
    P = [w * h for w, h in level_sizes(W, H)]
    S = sum(P)
    return {"pyramid": 2 * S - P[0] - P[-1], "fast": S, "blur": 2 * S, "describe": K * (749 + 512), "pack": K * 60,
            "frame": 5 * S - P[0] - P[-1] + 1321 * K}


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.QUERY,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def measured_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the newest committed
    ncu --set full capture (profiles/*_traffic.json, written by tools/ncu_summary.py); None if absent.
    Third value: the instruction-issue side of the same capture (warp instructions, issue-slot and pipe utilisation)."""
    files = sorted((ROOT / "profiles").glob("*_traffic.json"))
    if not files:
        return None, None, None
    z = json.loads(files[-1].read_text())
    t = z["bytes_per_launch"].get(kernel)
    return (int(t) if t else None), files[-1].name, z.get("issue", {}).get(kernel)


def host_threads():
    return len(os.sched_getaffinity(0))


def cpu_impl():
    """The timed CPU implementation: oracle/_ref (the reference's own sources compiled in place, kind
    "reference") when it was built, else the oracle restatement (kind "port")."""
    from oracle import ref_orb
    if ref_orb.available():
        return ref_orb, "reference"
    from oracle import orb_oracle
    orb_oracle.build()
    return orb_oracle, "port"


def cpu_sample_frames(wl, threads):
    """Frames per pass of the CPU legs (the cpu_baseline object and --impl reference use the SAME sample)."""
    return max(16 * threads, 64) if wl["W"] * wl["H"] < 2_000_000 else 2 * threads


def opencv_primitives(wl, n_frames=8):
    """The "real OpenCV" data point: the three streaming stages of the reference -- the resize chain with
    its reflect-101 frames (ORBextractor.cpp:1660-1713), per-cell cv::FAST at iniTh with the minTh retry (:1051-1141) and
    the 7x7 GaussianBlur of every level (:1598-1606) -- with the cv2 wheel (IPP / AVX-512 OpenCV), per frame, at 1 thread
    and at OpenCV's default thread count.  The CPU arms above run the reference's sources over the scalar
    oracle/cv_prims.hpp; this says how much faster a production OpenCV makes those stages."""
    import math
    import cv2
    from orb_slam_tracking_b200 import synth
    W, H = wl["W"], wl["H"]
    imgs = synth.tile_frames(n_frames, W, H, unique_pairs=min(4, n_frames // 2))
    sizes = level_sizes(W, H)
    fi = cv2.FastFeatureDetector_create(wl["ini"], True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    fm = cv2.FastFeatureDetector_create(wl["mn"], True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    out = {}
    default_threads = cv2.getNumThreads()
    for label, nt in (("threads_1", 1), ("threads_default", default_threads)):
        cv2.setNumThreads(nt)
        t = {"pyramid": 0.0, "fast": 0.0, "blur": 0.0}
        for im in imgs:
            t0 = time.perf_counter()
            pyr = [im]
            for (w, h) in sizes[1:]:
                pyr.append(cv2.resize(pyr[-1], (w, h), interpolation=cv2.INTER_LINEAR))
            padded = [cv2.copyMakeBorder(p, 19, 19, 19, 19, cv2.BORDER_REFLECT_101) for p in pyr]
            t1 = time.perf_counter()
            for p in pyr:
                h, w = p.shape
                mbx, mby = w - 16, h - 16
                nc, nr = int((mbx - 16) / 35.0), int((mby - 16) / 35.0)
                wc, hc = math.ceil((mbx - 16) / nc), math.ceil((mby - 16) / nr)
                for i in range(nr):
                    y0 = 16 + i * hc
                    if y0 >= mby - 3:
                        continue
                    y1 = min(y0 + hc + 6, mby)
                    for j in range(nc):
                        x0 = 16 + j * wc
                        if x0 >= mbx - 6:
                            continue
                        roi = p[y0:y1, x0:min(x0 + wc + 6, mbx)]
                        if not fi.detect(roi):
                            fm.detect(roi)
            t2 = time.perf_counter()
            for p in padded:
                cv2.GaussianBlur(p[19:-19, 19:-19].copy(), (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
            t3 = time.perf_counter()
            t["pyramid"] += t1 - t0; t["fast"] += t2 - t1; t["blur"] += t3 - t2
        out[label] = {"threads": nt, "ms_per_frame": {k: round(v / n_frames * 1e3, 3) for k, v in t.items()},
                      "frames_per_s_these_stages_only": n_frames / sum(t.values())}
    cv2.setNumThreads(default_threads)
    out["kind"] = "opencv"
    out["version"] = cv2.__version__
    out["sample"] = "%d frames of %s; Python call overhead of the per-cell FAST calls included" % (n_frames, wl["desc"])
    return out


def cpu_baseline(wl, n_frames, threads, target_s=12.0, repeats=None, frames=None):
    """The reference's CPU path on a bounded sample of the same workload (about `target_s` seconds of
    work, or exactly `repeats` passes).  Only the processing is timed (images are generated before)."""
    from orb_slam_tracking_b200 import synth
    impl, kind = cpu_impl()
    # the first n_frames of the same stream, all distinct (run_ours hands over its own batch: no second process pool)
    imgs = frames[:n_frames] if frames is not None and len(frames) >= n_frames else synth.frames_parallel(n_frames, wl["W"], wl["H"])
    K4, D4 = synth.camera_for(wl["W"], wl["H"])
    a = (wl["nfeatures"], 1.2, 8, wl["ini"], wl["mn"], K4, D4, wl.get("window", WINDOW), NNRATIO, threads)
    t = time.perf_counter()
    impl.run_stream(imgs, *a)  # warm-up pass, also calibrates the repeat count
    once = time.perf_counter() - t
    if repeats is None:
        repeats = max(1, int(round(target_s / max(once, 1e-3))))
    t = time.perf_counter()
    for _ in range(repeats):
        impl.run_stream(imgs, *a)
    dt = time.perf_counter() - t
    return repeats * n_frames / dt, dt, repeats, kind


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def run_reference(args, wl):
    """The reference's CPU path (oracle restatement: the reference itself needs OpenCV C++ and
    does not build here) on all host threads; rank 0 only."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    threads = host_threads()
    per_step = cpu_sample_frames(wl, threads)
    fps, dt, _, kind = cpu_baseline(wl, per_step, threads, repeats=args.steps)
    sample = "%d frames (%d pairs) of %s per step, %d host threads, %s" % (
        per_step, per_step // 2, wl["desc"], threads,
        "reference sources compiled in place (oracle/_ref)" if kind == "reference" else "oracle restatement")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": wl["desc"], "frames_per_step": per_step, "nnratio": NNRATIO, "window": wl.get("window", WINDOW)},
        "cpu_baseline": {"value": fps, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": fps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def bind_near_gpu(local):
    """Pin this rank to the CPUs next to its GPU (sysfs local_cpulist of the PCI device) before any page-locked buffer is
    allocated, so that the frames it uploads live on the GPU's own NUMA node.  Returns a short description."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        dev = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        txt = open("/sys/bus/pci/devices/%s/local_cpulist" % dev).read().strip()
        cpus = set()
        for part in txt.split(","):
            if part:
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "no local cpus visible (%s)" % txt
        os.sched_setaffinity(0, cpus)
        node = open("/sys/bus/pci/devices/%s/numa_node" % dev).read().strip()
        return "numa node %s, %d cpus" % (node, len(cpus))
    except Exception as e:  # sysfs not exposed: stay unbound
        return "unbound (%s)" % type(e).__name__


def run_ours(args, wl):
    import ctypes as C
    import torch
    from orb_slam_tracking_b200 import capi, frontend as fe, shard, synth
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    use_dist = world > 1
    B = args.batch or wl["batch"]
    B -= B % 2
    W, H, nf = wl["W"], wl["H"], wl["nfeatures"]
    # synthetic frames: every frame of the batch is DISTINCT (pair p of the stream is generated from seed p; rounds 1 and 2 up to
    # r02o timed 16 distinct pairs tiled to the batch, which now is the variant "tiled_16_pairs").  The job is a stream of
    # world * B/2 pairs; this rank owns one contiguous block of it (shard.py), no collective.  Generated first: the process
    # pool forks, and that is done before this process owns a CUDA context or NCCL threads.
    pair_lo, pair_hi = shard.shard_pairs(world * (B // 2), world, rank)
    assert pair_hi - pair_lo == B // 2
    imgs = synth.frames_parallel(B, W, H, first_pair=pair_lo)
    torch.cuda.set_device(local)
    binding = bind_near_gpu(local) if world > 1 and not os.environ.get("ORBX_BENCH_NO_BIND") else "not bound"
    if use_dist:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    WINDOW = wl.get("window", 100)
    L = capi.lib()
    E = fe.ORBextractor(nf, 1.2, 8, wl["ini"], wl["mn"], W, H, max_batch=B, device=local)
    K4, D4 = synth.camera_for(W, H)
    E.set_camera(K4, D4)
    S = fe.Stream(E, NNRATIO, True, WINDOW)
    nbytes_in = imgs.nbytes
    wc = bool(os.environ.get("ORBX_BENCH_WC"))   # experiment: write-combined upload buffer (bench-side switch, not a library option)
    hp = L.orbx_host_alloc_ex(nbytes_in, 1) if wc else L.orbx_host_alloc(nbytes_in)
    assert hp, "pinned allocation failed"
    pinned = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_ubyte)), shape=(nbytes_in,)).reshape(imgs.shape)
    pinned[...] = imgs
    cap = E.cap
    # pinned result buffers for the e2e path
    sizes = {"kps": B * cap * 28, "kps_un": B * cap * 28, "desc": B * cap * 32, "n": B * 4, "m12": (B // 2) * cap * 4,
             "nm": (B // 2) * 4, "cnt": (B // 2) * 12}
    res = {k: L.orbx_host_alloc(v) for k, v in sizes.items()}
    res_b = {k: L.orbx_host_alloc(v) for k, v in sizes.items()}   # second set: two steps are in flight on the streaming path
    nbytes_out = sum(sizes.values())

    def barrier():
        if use_dist:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        capi.check(L.orbx_stream_run(E._h, B, NNRATIO, 1, WINDOW), E._h)

    def step_e2e():
        # the call a user of the host API makes: pinned host frames in, keypoints / descriptors / matches out;
        # chunked inside so that the H2D and D2H copies overlap the kernels
        capi.check(L.orbx_stream_process(E._h, hp, B, NNRATIO, 1, WINDOW, args.chunk, res["kps"], res["kps_un"], res["desc"], cap,
                                         res["n"], res["m12"], res["nm"], res["cnt"]), E._h)

    flight = {"prev": None, "i": 0}

    def step_e2e_stream():
        # the streaming form of the same call: every step uploads its frames from pinned host memory and lands its
        # results in one of two pinned result sets; step s+1 is enqueued before step s is waited for, so its upload
        # overlaps the kernels and the download of step s
        r = res if flight["i"] % 2 == 0 else res_b
        flight["i"] += 1
        t = capi.check(L.orbx_stream_process_async(E._h, hp, B, NNRATIO, 1, WINDOW, args.chunk, r["kps"], r["kps_un"], r["desc"],
                                                   cap, r["n"], r["m12"], r["nm"], r["cnt"]), E._h)
        if flight["prev"] is not None:
            capi.check(L.orbx_stream_wait(E._h, flight["prev"]), E._h)
        flight["prev"] = t

    def flush_e2e_stream():
        if flight["prev"] is not None:
            capi.check(L.orbx_stream_wait(E._h, flight["prev"]), E._h)
            flight["prev"] = None

    def timed(fn, steps, flush=None):
        """K steps bracketed by barrier + synchronize; device time from CUDA events on the
        context's own stream (max over ranks)."""
        barrier()
        ms = C.c_float(0)
        capi.check(L.orbx_timer_start(E._h), E._h)
        for _ in range(steps):
            fn()
        if flush:
            flush()
        capi.check(L.orbx_timer_stop(E._h, C.byref(ms)), E._h)
        barrier()
        return shard.max_over_ranks(ms.value, device="cuda")

    capi.check(L.orbx_stream_upload(E._h, hp, B), E._h)
    for _ in range(max(args.warmup, 3)):
        step_resident()
    S.sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(step_resident, args.steps)
    launches = L.orbx_last_launch_count(E._h) * args.steps
    assert launches > 0, "no kernel launches counted"
    clocks = sampler.stop() if rank == 0 else None
    # the same resident step repeated for at least a second: the K-step headline region is tens of milliseconds, too
    # short for the clock sampler (100 ms period) and for the sustained clocks to settle
    k_sus = max(args.steps, int(np.ceil(1200.0 / max(ms / args.steps, 1e-3))))
    sampler2 = ClockSampler(local)
    if rank == 0:
        sampler2.start()
    ms_sus = timed(step_resident, k_sus)
    clocks_sus = sampler2.stop() if rank == 0 else None
    for _ in range(2):
        step_e2e()
    ms_e2e_call = timed(step_e2e, args.steps)
    for _ in range(2):
        step_e2e_stream()
    flush_e2e_stream()
    ms_e2e = timed(step_e2e_stream, args.steps, flush_e2e_stream)
    # what the host link alone needs for one e2e step: the same uploads and downloads with no kernels, every rank at once
    barrier()
    link_ms = C.c_float(0)
    hp_out = L.orbx_host_alloc(nbytes_out)
    capi.check(L.orbx_hostlink_probe(E._h, hp, nbytes_in, hp_out, nbytes_out, 5, C.byref(link_ms)), E._h)
    L.orbx_host_free(hp_out)
    link_ms_step = shard.max_over_ranks(link_ms.value, device="cuda")
    barrier()
    # per-stage device times (separate pass: the stage events are not in the timed region above)
    L.orbx_set_stage_timing(E._h, 1)
    acc = np.zeros(8)
    reps = max(3, min(args.steps, 10))
    for _ in range(reps):
        step_resident()
        S.sync()
        st = np.zeros(8, np.float32)
        L.orbx_get_stage_times(E._h, capi.ptr(st))
        acc += st
    stage_ms = acc / reps
    # the step after the path, outside every timed region above: Initializer::CheckHomography /
    # CheckFundamental for a RANSAC set of 200 hypotheses per model on every resident pair
    scoring = None
    if rank == 0:
        import time
        nP, nh = B // 2, 200
        rng = np.random.default_rng(5)
        Hs = np.tile(np.eye(3, dtype=np.float32).reshape(1, 1, 9), (nP, nh, 1))
        Hs[:, :, [2, 5]] += rng.uniform(-20, 20, (nP, nh, 2)).astype(np.float32)
        Hinv = Hs.copy()
        Hinv[:, :, [2, 5]] *= -1
        Fs = rng.normal(0, 1e-3, (nP, nh, 9)).astype(np.float32)
        S.score_models(nP, Hs, Hinv, Fs)   # warm-up (allocates the scratch)
        t0 = time.perf_counter()
        sc = S.score_models(nP, Hs, Hinv, Fs)
        call_ms = (time.perf_counter() - t0) * 1e3
        st = np.zeros(8, np.float32)
        L.orbx_get_stage_times(E._h, capi.ptr(st))
        evals = 2.0 * nh * float(sc["n_list"].sum())
        scoring = {"pairs": nP, "hypotheses_per_model": nh, "mean_matches": float(sc["n_list"].mean()),
                   "kernel_ms": round(float(st[5]), 4), "call_ms_with_copies": round(call_ms, 3),
                   "hypothesis_match_evals_per_s": evals / (float(st[5]) * 1e-3) if st[5] > 0 else None}
        if world == 1 and not args.no_cpu:
            from oracle import orb_oracle as orc, ref_orb
            n_all = np.ctypeslib.as_array(C.cast(res["n"], C.POINTER(C.c_int32)), shape=(B,))
            ku = np.ctypeslib.as_array(C.cast(res["kps_un"], C.POINTER(C.c_ubyte)), shape=(B * cap * 28,)).view(capi.KP_DTYPE).reshape(B, cap)
            m12 = np.ctypeslib.as_array(C.cast(res["m12"], C.POINTER(C.c_int32)), shape=(B // 2, cap))
            use_ref = ref_orb.available() and hasattr(ref_orb.lib(), "ref_check_homography")
            impl = ref_orb if use_ref else orc
            t0 = time.perf_counter()
            ev = 0
            for p in range(min(nP, 8)):
                k1, k2, m = ku[2 * p][: n_all[2 * p]], ku[2 * p + 1][: n_all[2 * p + 1]], m12[p][: n_all[2 * p]]
                impl.check_homography(k1, k2, m, Hs[p], Hinv[p])
                impl.check_fundamental(k1, k2, m, Fs[p])
                ev += 2 * nh * int((m >= 0).sum())
            dt = time.perf_counter() - t0
            scoring["cpu"] = {"hypothesis_match_evals_per_s": ev / dt, "cores": 1, "kind": "reference" if use_ref else "port",
                              "sample": "%d pairs x %d hypotheses x 2 models" % (min(nP, 8), nh)}
    # the same resident step on other inputs (the headline batch is 16 distinct pairs tiled; FAST, quadtree
    # and matcher costs are data dependent): (i) that batch of 16 pairs tiled, now a variant, (ii) the reference's own four
    # 640x480 images (Thirdparty/DBoW2/demo/images, carried in tests/golden/dbow_640x480_nf1000.npz) tiled to the batch
    variants = {}
    L.orbx_set_stage_timing(E._h, 0)   # the stage events serialise the step on one stream: off again for everything timed below
    if args.variants and args.workload == "tum":
        def resident_fps(batch):
            pinned[...] = batch
            capi.check(L.orbx_stream_upload(E._h, hp, B), E._h)
            for _ in range(3):
                step_resident()
            S.sync()
            return world * B * args.steps / (timed(step_resident, args.steps) * 1e-3)
        variants["tiled_16_pairs"] = {"value": resident_fps(synth.tile_frames(B, W, H, unique_pairs=min(16, B // 2), first_pair=pair_lo)),
                                      "unit": UNIT, "frames": B,
                                      "what": "16 distinct synthetic pairs tiled to the batch (the headline batch of round 1 and of round 2 up to r02o)"}
        try:
            import cv2
            z = np.load(ROOT / "tests" / "golden" / "dbow_640x480_nf1000.npz")
            real = np.stack([cv2.imdecode(z["png%d" % i], cv2.IMREAD_GRAYSCALE) for i in range(4)])
            variants["real_images_tiled"] = {"value": resident_fps(np.concatenate([real] * (B // 4 + 1))[:B]), "unit": UNIT,
                                             "what": "the reference's four DBoW2 demo images (640x480 gray) tiled to the batch"}
        except Exception as e:   # fixture or cv2 missing: report, do not fail the run
            variants["real_images_tiled"] = {"unavailable": type(e).__name__}
        pinned[...] = imgs
        capi.check(L.orbx_stream_upload(E._h, hp, B), E._h)
    # single-frame latency through the reference's own call shape ("report both single-frame latency and
    # batched throughput"): one Frame::Frame per image (orbx_frame_construct: host image in, keypoints / descriptors /
    # undistorted keys / grid out, ONE CUDA graph launch and ONE synchronisation) and one SearchForInitialization per pair
    # (host frames in, matches out), wall clock per call, outside every timed region above
    latency = None
    if rank == 0:
        import time
        fr = synth.frames(8, W, H, first_pair=pair_lo)
        Fs = [fe.Frame(E, im, fused=True) for im in fr]        # warm-up: builds the graph
        tf, tm = [], []
        M = fe.ORBmatcher(NNRATIO, True)
        for it in range(60):
            im = fr[it % 8]
            t0 = time.perf_counter()
            F = fe.Frame(E, im, fused=True)
            t1 = time.perf_counter()
            M.SearchForInitialization(Fs[(it % 8) & ~1], F, WINDOW)
            t2 = time.perf_counter()
            if it >= 10:
                tf.append((t1 - t0) * 1e6)
                tm.append((t2 - t1) * 1e6)
        latency = {"frame_us": {"p50": round(float(np.percentile(tf, 50)), 1), "p99": round(float(np.percentile(tf, 99)), 1)},
                   "pair_match_us": {"p50": round(float(np.percentile(tm, 50)), 1), "p99": round(float(np.percentile(tm, 99)), 1)},
                   "api": "orbx_frame_construct (CUDA graph per context) / orbx_search_for_initialization, host buffers, "
                          "ctypes call overhead included", "samples": len(tf)}
        if world == 1 and not args.no_cpu:
            impl, kind = cpu_impl()
            a1 = (wl["nfeatures"], 1.2, 8, wl["ini"], wl["mn"], K4, D4, WINDOW, NNRATIO, 1)
            impl.run_stream(fr[:2], *a1)
            t0 = time.perf_counter()
            for _ in range(5):
                impl.run_stream(fr[:2], *a1)
            latency["cpu_reference_pair_us"] = {"value": round((time.perf_counter() - t0) / 5 * 1e6, 1), "kind": kind, "cores": 1,
                                                "what": "2 x Frame::Frame + SearchForInitialization on one host core"}
    # raw all-pairs Hamming best / second best, 8000 x 8000 descriptors (the matcher's distance matrix at
    # BASELINE.json configs[4] without geometry), outside every timed region above
    brute = None
    if rank == 0:
        rng = np.random.default_rng(7)
        da = rng.integers(0, 256, (8000, 32), dtype=np.uint8)
        db = rng.integers(0, 256, (8000, 32), dtype=np.uint8)
        _, _, _, bf_ms = E.hamming_best2(da, db, repeats=20)
        brute = {"queries": 8000, "candidates": 8000, "kernel": "k_hamming_topk", "kernel_ms": round(bf_ms, 4),
                 "pairs_per_s": 64e6 / (bf_ms * 1e-3), "popc32_per_s": 8 * 64e6 / (bf_ms * 1e-3),
                 "note": "integer / popc pipe utilisation of this kernel: profiles/ (ncu)"}
    L.orbx_set_stage_timing(E._h, 0)
    # sanity of the work done: every frame produced keypoints and pairs produced matches
    n_out = np.ctypeslib.as_array(C.cast(res["n"], C.POINTER(C.c_int32)), shape=(B,)).copy()
    nm = np.ctypeslib.as_array(C.cast(res["nm"], C.POINTER(C.c_int32)), shape=(B // 2,)).copy()
    assert n_out.min() > 0 and nm.sum() > 0, "pipeline produced no features / matches"

    if rank == 0:
        fps = world * B * args.steps / (ms * 1e-3)
        fps_e2e = world * B * args.steps / (ms_e2e * 1e-3)
        ab = algorithmic_bytes(W, H, float(n_out.mean()))
        peak, how = measured_peak()
        hbm_stages = {"pyramid": 0, "fast": 1, "blur": 3}
        dom = max(hbm_stages, key=lambda s: stage_ms[hbm_stages[s]])
        dom_ms = float(stage_ms[hbm_stages[dom]])
        n_launch_dom = 7 if dom == "pyramid" else 1
        achieved = ab[dom] * B / (dom_ms * 1e-3) / 1e9
        dom_kernel = {"pyramid": "k_resize", "fast": "k_fast", "blur": "k_blur"}[dom]
        traffic, traffic_src, issue = measured_traffic(dom_kernel) if (args.workload == "tum" and B == 512) else (None, None, None)
        if issue:
            issue = dict(issue, warp_instr_per_pixel=issue["warp_instr"] / (ab[dom] * B), source=traffic_src,
                         note="the kernel is bound by instruction issue (integer SWAR on the alu pipe), not by HBM: its roofline fraction "
                              "against HBM says how far the exact 8-bit arithmetic is from a pure copy")
        line = {
            "metric": METRIC, "value": fps, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": wl["desc"], "frames_per_step_per_gpu": B, "pairs_per_step_per_gpu": B // 2,
                       "batch": "all frames distinct (pairs %d..%d of the synthetic stream on rank 0)" % (pair_lo, pair_hi - 1),
                       "nlevels": 8, "scale": 1.2, "fast_th": [wl["ini"], wl["mn"]], "nnratio": NNRATIO, "window": WINDOW,
                       "l2": "batch working set (%.0f MB of pyramids) exceeds the 126 MB L2; no flush" % (2 * B * 1.25 * W * H / 307200),
                       "sharding": "independent frame pairs per rank, no collective", "host_binding_rank0": binding,
                       "upload_buffer": "pinned, write-combined" if wc else "pinned"},
            "e2e": {"value": fps_e2e, "unit": UNIT, "h2d_bytes_per_step": int(nbytes_in), "d2h_bytes_per_step": int(nbytes_out),
                    "ms_per_step": ms_e2e / args.steps, "api": "orbx_stream_process_async + orbx_stream_wait, 2 steps in flight",
                    "host_link": {"copy_only_ms_per_step": link_ms_step, "frac_of_e2e_step": link_ms_step / (ms_e2e / args.steps),
                                  "h2d_GBps_per_gpu": nbytes_in / (link_ms_step * 1e-3) / 1e9,
                                  "what": "orbx_hostlink_probe: this step's H2D and D2H bytes, both directions at once, no kernels, all ranks at once"},
                    "chunk_frames": args.chunk if args.chunk > 0 else "auto: 3 equal chunks per step on 3 compute lanes",
                    "blocking_call": {"value": world * B * args.steps / (ms_e2e_call * 1e-3), "ms_per_step": ms_e2e_call / args.steps,
                                      "api": "orbx_stream_process (one step at a time)"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "sustained": {"value": world * B * k_sus / (ms_sus * 1e-3), "unit": UNIT, "steps": k_sus, "seconds": ms_sus * 1e-3,
                          "ms_per_step": ms_sus / k_sus, "clocks": clocks_sus},
            "stage_ms_per_step": {STAGES[i]: round(float(stage_ms[i]), 4) for i in range(8) if i != 5},
            "pipeline_hbm": {"algorithmic_GBps": ab["frame"] * fps / world / 1e9, "frac_of_peak": ab["frame"] * fps / world / 1e9 / peak,
                             "bytes_per_frame": ab["frame"]},
            "roofline": {"kernel": {"pyramid": "k_resize (7 launches)", "fast": "k_fast", "blur": "k_blur"}[dom],
                         "bound": "hbm", "bound_measured": "issue" if issue and issue["issue_active_pct"] > 50 else None, "issue": issue,
                         "achieved": achieved, "peak": peak, "peak_source": how, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "launches_per_step": n_launch_dom,
                         "ms_per_step": dom_ms, "algorithmic_bytes_per_step": int(ab[dom] * B)},
        }
        line["variants"] = variants
        line["latency"] = latency
        line["next_rows"] = {"init_scoring": scoring, "brute_force_hamming": brute}
        if world == 1 and not args.no_cpu:
            threads = host_threads()
            n_cpu = cpu_sample_frames(wl, threads)
            v, dt, reps, kind = cpu_baseline(wl, n_cpu, threads, frames=imgs if pair_lo == 0 else None)
            try:
                line["cpu_opencv_primitives"] = opencv_primitives(wl)
            except Exception as e:   # no cv2 on the box: the reference-source arm below is the baseline
                line["cpu_opencv_primitives"] = {"unavailable": type(e).__name__}
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": kind,
                                    "sample": "%d passes over %d frames (%d pairs) of the same workload, %.1f s, %s" % (
                                        reps, n_cpu, n_cpu // 2, dt, "reference sources compiled in place (oracle/_ref)"
                                        if kind == "reference" else "oracle restatement")}
        print(json.dumps(line), flush=True)
    for p in list(res.values()) + list(res_b.values()) + [hp]:
        L.orbx_host_free(p)
    E.close()
    if use_dist:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="tum", choices=sorted(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--chunk", type=int, default=0,
                    help="frames per pipelined chunk of the e2e path (0 = the library's default: 3 equal chunks per step when streaming, 128 ramped when blocking)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-variants", dest="variants", action="store_false", help="skip the tiled-pairs and real-image batches")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()

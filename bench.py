#!/usr/bin/env python
"""bench.py — headline benchmark of the ORB front end on B200.

Workload (BASELINE.json configs[1]): a monocular stream of 1024 synthetic 640x480 frames per
GPU, ORBextractor(1000, 1.2, 8, 20, 7) on every frame + SearchForInitialization(window 100,
ratio 0.9) between consecutive frames.  One "step" = one pass over the
1024-frame batch.  `value` = frames/s with the frames already resident in HBM; `e2e` = the same
through the host-buffer C-ABI call (pinned host images in, keypoints + descriptors out).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl orbfe|reference]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from orb_slam2_annotate_b200 import synth  # noqa: E402

METRIC = "ORB frames/s @640x480/1000kp/8lvl"
W, H, NFEAT = 640, 480, 1000
KNN_N = 1_000_000            # configs[4]: 1M x 1M 256-bit rows


def workload_config(batch):
    """`config` of both arms (identical on purpose: the driver compares them)."""
    return {"workload": "configs[1]: monocular stream, 1024 synthetic 640x480 frames per GPU, ORBextractor(1000,1.2,8,20,7) on every "
                        "frame + SearchForInitialization(window 100, ratio 0.9) between consecutive frames",
            "frames_per_step_per_gpu": batch, "l2": "inputs_larger_than_l2 (315 MB of frames per step)",
            "e2e": "host-buffer C-ABI calls, two batches in flight per GPU (two handles, two host threads), every copy in the timed region",
            "knn": "configs[4]: 1Mx1M 256-bit kNN-2, queries sharded over the GPUs, 8 B/query NCCL all-gather (strong scaling)"}


def cpu_knn_gcmps(nq=4096, nt=KNN_N):
    """CPU leg of the Hamming metric: the reference's bit-hack DescriptorDistance (src/ORBmatcher.cc:1828-1844) in the
    best / second-best loop (:233-269), one thread per host core, on a nq x nt slice of configs[4]; the 1M x 1M figure
    would be this rate (the loop is linear in both sizes) — labelled extrapolated where it is quoted."""
    from oracle import oracle_py as O
    cores = os.cpu_count() or 1
    t = synth.descriptors(nt, 1)
    q = synth.descriptors(nq, 2)
    parts = np.array_split(np.arange(nq), cores)
    O.knn2(q[:4], t[:1024], 50, 0.7)
    th = [threading.Thread(target=lambda p=p: O.knn2(q[p], t, 50, 0.7)) for p in parts if len(p)]
    t0 = time.perf_counter()
    [x.start() for x in th]
    [x.join() for x in th]
    dt = time.perf_counter() - t0
    return {"gcmps": nq * nt / dt / 1e9, "cores": cores, "sample": f"{nq} x {nt} slice of configs[4], one thread per core, bit-hack popcount", "seconds": dt}


def level_sizes(w, h, nlevels=8, sf=1.2):
    scale = [np.float32(1.0)]
    for _ in range(1, nlevels):
        scale.append(np.float32(float(scale[-1]) * float(np.float32(sf))))
    inv = [np.float32(1.0) / s for s in scale]
    return [(int(np.rint(np.float32(w) * i)), int(np.rint(np.float32(h) * i))) for i in inv]


def algorithmic_bytes(w, h):
    """Compulsory HBM bytes per frame and kernel."""
    px = [a * b for a, b in level_sizes(w, h)]
    return {
        "pyramid": sum(px[:-1]) + sum(px[1:]),       # read levels 0..6, write levels 1..7
        "fast": sum(px),                             # read every level once (+ tiny corner lists)
        "blur": 2 * sum(px),                         # read + write every level
    }


# dram__bytes_read.sum + dram__bytes_write.sum per 1024-frame launch of each stage's kernels, from the
# `ncu --set full` capture summarised in profiles/r01_ncu_full_summary.txt (640x480, 8 levels)
NCU_DRAM_BYTES_PER_1024 = {
    "pyramid": 1.0066e9 + 447.2e6,
    "fast": 938.8e6 + 22.8e6 + 42.2e6,
    "blur": 1.077e9 + 985.6e6,
}


def guided_search_timings(frame_img):
    """The per-call guided searches beside their CPU restatement (part of the cpu_baseline leg):
    SearchByProjection(Frame, MapPoints) at config 4's shape (20 000 map points against one
    1000-keypoint frame) and SearchByProjection(CurrentFrame, LastFrame) (one point per last-frame
    keypoint).  GPU = the host-buffer C-ABI call (copies and synchronisation included), CPU = the
    oracle port on one core; ms per call, median of 5."""
    from oracle import oracle_py as O
    import orb_slam2_annotate_b200 as orbfe
    rng = np.random.default_rng(4)
    ex = orbfe.ORBextractor(NFEAT)
    k, d = ex(frame_img)
    scale = ex.GetScaleFactors()
    n, nmp = len(k), 20000
    src = rng.integers(0, n, nmp)
    mps = dict(track_in_view=np.ones(nmp, np.uint8), proj_x=(k["x"][src] + rng.uniform(-3, 3, nmp)).astype(np.float32),
               proj_y=(k["y"][src] + rng.uniform(-3, 3, nmp)).astype(np.float32), proj_xr=np.zeros(nmp, np.float32),
               level=k["octave"][src].astype(np.int32), view_cos=np.full(nmp, 0.9, np.float32),
               desc=d[src] ^ rng.integers(0, 2, (nmp, 32), dtype=np.uint8), observations=rng.integers(0, 3, nmp).astype(np.int32))
    pts = dict(valid=np.ones(n, np.uint8), u=(k["x"] + rng.uniform(-4, 4, n)).astype(np.float32),
               v=(k["y"] + rng.uniform(-4, 4, n)).astype(np.float32), ur=np.zeros(n, np.float32),
               radius=(np.float32(15.0) * scale[k["octave"]]).astype(np.float32), min_level=(k["octave"] - 1).astype(np.int32),
               max_level=(k["octave"] + 1).astype(np.int32), angle=k["angle"].astype(np.float32),
               desc=d ^ rng.integers(0, 2, (n, 32), dtype=np.uint8), observations=rng.integers(0, 3, n).astype(np.int32))
    F = orbfe.Frame(k, d, W, H, scale)
    fo = O.frame_view(k, d, W, H, scale)
    mt = orbfe.ORBmatcher(0.8, True)

    def med(fn):
        fn()
        ts = []
        for _ in range(5):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        return float(np.median(ts)) * 1e3

    # the same searches against a frame that never left the device (Frame epilogue K11: mvKeysUn + mGrid
    # written by the extractor; camera without k1, so mvKeysUn == mvKeys and the answers are the same)
    ex.set_camera([520.0, 520.0, 320.0, 240.0, 0.0, 0.0, 0.0, 0.0, 0.0], W, H)
    ex.extract_batch(frame_img[None])
    dv = ex.device_frame_view(0, n)
    # Frame epilogue itself: 64 copies of the frame through the extractor with the TUM1 camera, stage timer
    tum1 = [517.306408, 516.469215, 318.643040, 255.313989, 0.262383, -0.953104, -0.005358, 0.002628, 1.163314]
    ex2 = orbfe.ORBextractor(NFEAT)
    ex2.set_camera(tum1, W, H)
    ex2.set_profiling(True)
    rep = np.ascontiguousarray(np.repeat(frame_img[None], 64, 0))
    ex2.extract_batch(rep); ex2.extract_batch(rep)
    epi_us = ex2.stage_times_ms()["frame"] * 1e3 / 64
    xy = np.stack([k["x"], k["y"]], 1)
    cam_f, dist_f = np.array(tum1[:4], np.float32), np.array(tum1[4:], np.float32)
    bnd = O.image_bounds(cam_f, dist_f, W, H)

    def cpu_epilogue():
        un = k.copy()
        u = O.undistort_points(xy, cam_f, dist_f)
        un["x"], un["y"] = u[:, 0], u[:, 1]
        O.assign_grid(O.frame_view(un, d, W, H, scale, bounds=bnd))

    # SearchByBoW(KeyFrame*, Frame&): the frame against a noisy copy of itself, stand-in feature vectors (node =
    # octave | low 4 bits of the first descriptor byte -> ~128 nodes of ~8 features, the shape DBoW2 gives at levelsup 4)
    def fvec(dd):
        fv = {}
        for i in range(n):
            fv.setdefault(int((dd[i, 0] & 15) | (k["octave"][i] << 4)), []).append(i)
        return fv
    d_noisy = d.copy()
    flip = rng.integers(0, 256, (n, 12))
    for j in range(12):
        d_noisy[np.arange(n), flip[:, j] >> 3] ^= (1 << (flip[:, j] & 7)).astype(np.uint8)
    d_noisy[:, 0] = (d_noisy[:, 0] & 0xF0) | (d[:, 0] & 0x0F)
    fv1, fv2 = fvec(d), fvec(d_noisy)
    valid1 = (rng.random(n) < 0.8).astype(np.uint8)
    b1, b2 = orbfe.BowSide(d, k["angle"], fv1, valid1), orbfe.BowSide(d_noisy, k["angle"], fv2)
    o1, o2 = O.bow_view(d, k["angle"], fv1, valid1), O.bow_view(d_noisy, k["angle"], fv2)
    mtb = orbfe.ORBmatcher(0.75, True)
    bow_matches = mtb.SearchByBoW(b1, b2)[0]

    # key-frame guided searches: 20 000 map points fused into one key frame (Fuse, pose
    # variant with its reprojection gate), SearchBySim3 between the frame and its noisy twin, SearchForTriangulation
    # over the stand-in feature vectors with a horizontal-line fundamental matrix
    tb = ex.GetScaleFactors(); sig2 = (tb * tb).astype(np.float32); inv_sig2 = (np.float32(1.0) / sig2).astype(np.float32)
    lvl = k["octave"][src].astype(np.int32)
    wpts = dict(valid=np.ones(nmp, np.uint8), u=mps["proj_x"], v=mps["proj_y"], ur=np.zeros(nmp, np.float32),
                radius=(np.float32(3.0) * scale[lvl]).astype(np.float32), level=lvl, desc=mps["desc"])
    kpts = dict(valid=np.ones(n, np.uint8), u=pts["u"], v=pts["v"], ur=np.zeros(n, np.float32),
                radius=(np.float32(7.5) * scale[k["octave"]]).astype(np.float32), level=k["octave"].astype(np.int32), desc=pts["desc"],
                angle=k["angle"].astype(np.float32))
    F2 = orbfe.Frame(k, d_noisy, W, H, scale); fo2 = O.frame_view(k, d_noisy, W, H, scale)
    F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
    none = np.zeros(n, np.uint8)
    t1, t2 = orbfe.TriSide(k, d, none, fv1), orbfe.TriSide(k, d_noisy, none, fv2)
    ot1, ot2 = O.tri_side(k, d, none, fv1), O.tri_side(k, d_noisy, none, fv2)

    # MapPoint::ComputeDistinctiveDescriptors for 20 000 map points with 3..20 observations each
    sizes = rng.integers(3, 21, 20000).astype(np.int32)
    dstart = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    dobs = d[rng.integers(0, n, int(dstart[-1]))] ^ (rng.random((int(dstart[-1]), 32)) < 0.02).astype(np.uint8)
    mtd = orbfe.ORBmatcher()

    return {
        "distinctive_descriptors": {"map_points": len(sizes), "observations": int(dstart[-1]),
                                    "gpu_ms_per_call": med(lambda: mtd.ComputeDistinctiveDescriptors(dobs, dstart)),
                                    "cpu_ms_per_call": med(lambda: O.distinctive_descriptors(dobs, dstart))},
        "search_by_bow": {"features": n, "nodes": len(fv1), "matches": bow_matches,
                          "gpu_ms_per_call": med(lambda: mtb.SearchByBoW(b1, b2)),
                          "cpu_ms_per_call": med(lambda: O.search_by_bow(o1, o2, 50, 0.75, True)),
                          # relocalisation shape: 32 key frames against the one frame, one call
                          "gpu_ms_per_32_pairs_batched": med(lambda: mtb.SearchByBoWBatch([b1] * 32, [b2] * 32)),
                          "cpu_ms_per_32_pairs": med(lambda: [O.search_by_bow(o1, o2, 50, 0.75, True) for _ in range(32)])},
        "search_local_points": {"map_points": nmp, "keypoints": n,
                                "gpu_ms_per_call": med(lambda: mt.SearchByProjection(F, mps, 3.0, np.zeros(n, np.uint8))),
                                "gpu_resident_frame_ms_per_call": med(lambda: mt.SearchByProjection(dv, mps, 3.0, np.zeros(n, np.uint8))),
                                "cpu_ms_per_call": med(lambda: O.search_by_projection(fo, mps, 3.0, 0.8, np.zeros(n, np.uint8)))},
        "search_frame_points": {"points": n, "keypoints": n,
                                "gpu_ms_per_call": med(lambda: mt.SearchByProjectionLast(F, pts, np.zeros(n, np.uint8))),
                                "gpu_resident_frame_ms_per_call": med(lambda: mt.SearchByProjectionLast(dv, pts, np.zeros(n, np.uint8))),
                                "cpu_ms_per_call": med(lambda: O.search_by_projection_frame(fo, pts, True, np.zeros(n, np.uint8)))},
        "fuse": {"map_points": nmp, "keypoints": n, "fused": int((mt.FuseSearch(F, wpts, inv_sig2)[0] >= 0).sum()),
                 "gpu_ms_per_call": med(lambda: mt.FuseSearch(F, wpts, inv_sig2)),
                 "gpu_resident_frame_ms_per_call": med(lambda: mt.FuseSearch(dv, wpts, inv_sig2)),
                 "cpu_ms_per_call": med(lambda: O.fuse_search(fo, wpts, inv_sig2))},
        "search_by_sim3": {"keypoints": n, "found": mt.SearchBySim3(F, F2, kpts, kpts)[0],
                           "gpu_ms_per_call": med(lambda: mt.SearchBySim3(F, F2, kpts, kpts)),
                           "cpu_ms_per_call": med(lambda: O.search_by_sim3(fo, fo2, kpts, kpts))},
        "search_projection_keyframe": {"points": n, "keypoints": n,
                                       "gpu_ms_per_call": med(lambda: mt.SearchByProjectionKeyFrame(F2, kpts, 100, none)),
                                       "cpu_ms_per_call": med(lambda: O.search_by_projection_keyframe(fo2, kpts, 100, True, none))},
        "search_for_triangulation": {"features": n, "nodes": len(fv1),
                                     "matches": mt.SearchForTriangulation(t1, t2, F12, -5000.0, 0.0, tb, sig2)[0],
                                     "gpu_ms_per_call": med(lambda: mt.SearchForTriangulation(t1, t2, F12, -5000.0, 0.0, tb, sig2)),
                                     "cpu_ms_per_call": med(lambda: O.search_for_triangulation(ot1, ot2, F12, -5000.0, 0.0, tb, sig2))},
        "frame_epilogue": {"what": "UndistortKeyPoints + AssignFeaturesToGrid, TUM1 camera", "keypoints": n,
                           "gpu_us_per_frame_in_batch": epi_us, "cpu_us_per_frame": med(cpu_epilogue) * 1e3},
    }


def single_frame_latency(frames):
    """The shape Tracking uses: ONE 640x480 frame through the host-buffer C ABI (frame in, keypoints + descriptors
    out on the host) and SearchForInitialization on two such frames; beside it the unmodified reference on one core."""
    from oracle import oracle_py as O
    import orb_slam2_annotate_b200 as orbfe
    ex = orbfe.ORBextractor(NFEAT, 1.2, 8, 20, 7)
    mt = orbfe.ORBmatcher(0.9, True)
    res = [ex(np.ascontiguousarray(frames[i])) for i in range(2)]
    for _ in range(20):
        ex(frames[0])
    te = []
    for i in range(200):
        t0 = time.perf_counter(); ex(frames[i & 1]); te.append(time.perf_counter() - t0)
    scale = ex.GetScaleFactors()
    F = [orbfe.Frame(k, d, W, H, scale) for k, d in res]
    prev = np.ascontiguousarray(np.stack([res[0][0]["x"], res[0][0]["y"]], 1), np.float32)
    tm = []
    for i in range(110):
        p = prev.copy()
        t0 = time.perf_counter(); mt.SearchForInitialization(F[0], F[1], p, 100); tm.append(time.perf_counter() - t0)
    kind = "reference" if O.ref_lib("_native") is not None else "port"
    rex = O.RefExtractor(NFEAT, variant="_native") if kind == "reference" else O.OracleExtractor(NFEAT)
    search = O.ref_search_for_initialization if (kind == "reference" and O.mref_lib() is not None) else O.search_for_initialization
    rex(frames[0])
    t0 = time.perf_counter()
    cpu = [rex(frames[i & 1]) for i in range(6)]
    cpu_extract = (time.perf_counter() - t0) / 6
    v = [O.frame_view(k, d, W, H, scale) for k, d in cpu[:2]]
    t0 = time.perf_counter()
    for _ in range(6):
        search(v[0], v[1], np.ascontiguousarray(np.stack([cpu[0][0]["x"], cpu[0][0]["y"]], 1), np.float32), 100, 0.9, True)
    cpu_search = (time.perf_counter() - t0) / 6
    return {"what": "one 640x480 frame through the host-buffer C ABI, and SearchForInitialization(window 100) on two such frames",
            "extract_ms": float(np.median(te)) * 1e3, "extract_ms_p90": float(np.percentile(te, 90)) * 1e3,
            "search_init_ms": float(np.median(tm[10:])) * 1e3, "cpu_extract_ms_one_core": cpu_extract * 1e3,
            "cpu_search_init_ms_one_core": cpu_search * 1e3, "cpu_kind": kind}


def stereo_timings(npairs=128):
    """configs[2] beside the CPU: `npairs` KITTI-shape 1241x376 stereo pairs, nFeatures 2000 — both images extracted
    (one interleaved host batch through the C ABI, copies included) and Frame::ComputeStereoMatches on the
    device-resident results; CPU = the unmodified reference extractor on both images + the oracle's restatement of
    ComputeStereoMatches, one core, 2 pairs."""
    from oracle import oracle_py as O
    import orb_slam2_annotate_b200 as orbfe
    from orb_slam2_annotate_b200 import synth
    distinct = [synth.stereo_pair(1241, 376, seed=s, disparity=20) for s in range(4)]
    imgs = orbfe.pinned_empty((2 * npairs, 376, 1241), np.uint8)
    for p in range(npairs):
        imgs[2 * p], imgs[2 * p + 1] = distinct[p % 4]
    mbf = 386.1448; mb = mbf / 718.856
    ex = orbfe.ORBextractor(2000)
    mt = orbfe.ORBmatcher()
    cap = ex.max_keypoints()
    out = (orbfe.pinned_empty((2 * npairs, cap), orbfe.KP_DTYPE), orbfe.pinned_empty((2 * npairs, cap, 32), np.uint8),
           orbfe.pinned_empty((2 * npairs,), np.int32))
    ex.extract_batch(imgs, out=out)

    def gpu():
        ex.extract_batch(imgs, out=out)
        return mt.ComputeStereoMatchesBatch(ex, npairs, mbf, mb)

    def gpu_match_only():
        return mt.ComputeStereoMatchesBatch(ex, npairs, mbf, mb)
    ur, _ = gpu()
    ts, tm = [], []
    for _ in range(3):
        t0 = time.perf_counter(); gpu(); ts.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); gpu_match_only(); tm.append(time.perf_counter() - t0)
    kind = "reference" if O.ref_lib("_native") is not None else "port"
    t0 = time.perf_counter()
    for p in range(2):
        fr = []
        for im in distinct[p]:
            oc = O.RefExtractor(2000, variant="_native") if kind == "reference" else O.OracleExtractor(2000)
            k, d = oc(im)
            fr.append((k, d, [np.ascontiguousarray(oc.level(l)) for l in range(8)]))
        tb = O.OracleExtractor(2000).tables()
        O.compute_stereo_matches(O.frame_view(fr[0][0], fr[0][1], 1241, 376, tb["scale"]), O.frame_view(fr[1][0], fr[1][1], 1241, 376, tb["scale"]),
                                 O.pyramid_view(fr[0][2]), O.pyramid_view(fr[1][2]), tb["inv_scale"], mbf, mb)
    cpu_ms = (time.perf_counter() - t0) * 1e3 / 2
    return {"workload": "configs[2]: 1241x376 stereo pairs, nFeatures 2000, extract both + ComputeStereoMatches", "pairs": npairs,
            "matched_per_pair": float((ur >= 0).sum() / npairs), "gpu_pairs_per_s": npairs / float(np.median(ts)),
            "gpu_stereo_match_ms_per_pair": float(np.median(tm)) * 1e3 / npairs,
            "cpu_ms_per_pair_one_core": cpu_ms, "cpu_kind": kind}


def make_frames(batch, seed):
    """`batch` distinct frames: crops of one synthetic canvas with per-frame noise."""
    distinct = min(batch, 256)
    frames, _ = synth.sequence(distinct, W, H, seed=seed)
    if distinct < batch:
        frames = np.concatenate([frames] * ((batch + distinct - 1) // distinct))[:batch]
    return np.ascontiguousarray(frames)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.p = gpu_index, [], None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def cpu_reference_fps(frames, seconds_budget=20.0, steps=1, warmup=0):
    """The reference's own CPU path for the workload, one thread per host core: extraction by
    the unmodified ORBextractor.cc (oracle/_ref, reference -O3 flags, OpenCV primitives from the
    cv-shim; the oracle port when _ref was not prebuilt) and SearchForInitialization between
    consecutive frames by the unmodified ORBmatcher.cc + Frame.cc (oracle/_ref/liborbref_matcher.so; the
    oracle's restatement when it was not prebuilt).  Every thread owns one extractor and a contiguous run of frames."""
    from oracle import oracle_py as O
    cores = os.cpu_count() or 1
    kind = "reference" if O.ref_lib("_native") is not None else "port"
    search = O.ref_search_for_initialization if (kind == "reference" and O.mref_lib() is not None) else O.search_for_initialization
    matcher_kind = "unmodified ORBmatcher.cc / Frame.cc on the cv-shim" if search is O.ref_search_for_initialization else "oracle port"

    def make():
        return O.RefExtractor(NFEAT, variant="_native") if kind == "reference" else O.OracleExtractor(NFEAT)

    exs = [make() for _ in range(cores)]
    scale = O.OracleExtractor(NFEAT).tables()["scale"]
    t0 = time.perf_counter()
    exs[0](frames[0])
    per_frame = time.perf_counter() - t0
    # every thread a contiguous run of frames; all of `frames` unless the time budget bounds the sample
    n = int(max(2 * cores, min(len(frames), seconds_budget / max(per_frame, 1e-3))))
    runs = [len(p) for p in np.array_split(np.arange(n), cores)]
    first = np.concatenate([[0], np.cumsum(runs)])
    run = max(runs)

    def run_once():
        def work(t):
            prev = None
            for i in range(runs[t]):
                img = frames[(first[t] + i) % len(frames)]
                k, d = exs[t](img)
                cur = O.frame_view(k, d, W, H, scale)
                if prev is not None:
                    pm = np.ascontiguousarray(np.stack([prev[0][0]["x"], prev[0][0]["y"]], 1), np.float32)
                    search(prev[1], cur, pm, 100, 0.9, True)
                prev = ((k, d), cur)
        th = [threading.Thread(target=work, args=(t,)) for t in range(cores)]
        t0 = time.perf_counter()
        [t.start() for t in th]
        [t.join() for t in th]
        return time.perf_counter() - t0

    for _ in range(warmup):
        run_once()
    times = [run_once() for _ in range(max(steps, 1))]
    dt = float(np.mean(times))
    return {"value": n / dt, "unit": "frames/s", "cores": cores, "kind": kind,
            "sample": f"{n} of the workload's 640x480 frames ({cores} threads x <= {run} consecutive frames: extract + "
                      f"SearchForInitialization between neighbours), mean of {max(steps, 1)} pass(es); extractor = "
                      f"unmodified ORBextractor.cc on the cv-shim, matcher = {matcher_kind}",
            "ms_per_step": dt * 1e3, "frames": n}


def run_reference(args, rank, world):
    if rank != 0:
        return
    frames = make_frames(args.batch, seed=0)
    # the same step as the orbfe arm: one pass over the 1024 frames (about 2 s on 16 cores)
    r = cpu_reference_fps(frames, seconds_budget=1e9, steps=args.steps, warmup=min(args.warmup, 1))
    k = cpu_knn_gcmps()
    base = {kk: r[kk] for kk in ("value", "unit", "cores", "kind", "sample")}
    base.update({"knn_gcmps": k["gcmps"], "knn_sample": k["sample"]})
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "frames/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(r["frames"]),
            "cpu_baseline": base,
            "e2e": {"value": r["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "roofline": {"knn_cpu_gcmps": k["gcmps"]},
            "gpu_launches": 0,
            "note": "unmodified ORBextractor.cc / ORBmatcher.cc / Frame.cc on the scalar cv-shim OpenCV stand-ins (a SIMD OpenCV build would be about 2x faster)"}
    emit(line)


_JSON_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries underneath write there too (NCCL prints its version line when the
    first communicator is created): from here on file descriptor 1 goes to stderr and the JSON line to the real stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="orbfe", choices=["orbfe", "reference"])
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-knn", action="store_true", help="profiling runs only: skip the configs[4] Hamming leg (the line then lacks it)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import orb_slam2_annotate_b200 as orbfe

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: liborbfe has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B = args.batch
    frames = make_frames(B, seed=1000 * rank)           # every rank its own stream (weak scaling)
    ex = orbfe.ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank, max_width=W, max_height=H, max_batch=B)
    mt = orbfe.ORBmatcher(0.9, True, device=local_rank)            # Tracking.cc:641
    d_frames = torch.from_numpy(frames).cuda()
    xs = torch.cuda.ExternalStream(ex.stream, device=local_rank)   # the library's own streams
    ms = torch.cuda.ExternalStream(mt.stream, device=local_rank)
    cap = ex.max_keypoints()
    d_m12 = torch.empty((max(B - 1, 1), cap), dtype=torch.int32, device="cuda")
    d_nm = torch.zeros(max(B - 1, 1), dtype=torch.int32, device="cuda")
    m_ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    x_done, m_done = torch.cuda.Event(), torch.cuda.Event()

    def step(timed=None):
        """extract the batch, then SearchForInitialization between consecutive frames"""
        xs.wait_event(m_done)                            # the matcher of the previous step read these buffers
        ex.extract_batch_device(d_frames.data_ptr(), B, W, H)
        x_done.record(xs)
        ms.wait_event(x_done)
        kp_ptr, desc_ptr, cnt_ptr, dcap = ex.device_results()
        m_ev[0].record(ms)
        mt.search_init_consecutive_device(kp_ptr, desc_ptr, cnt_ptr, dcap, B, W, H, d_m12.data_ptr(), d_nm.data_ptr(), 100)
        m_ev[1].record(ms)
        m_done.record(ms)
        if timed is not None:
            for k, v in ex.stage_times_ms().items():     # syncs on the step's last extractor event
                timed[k] = timed.get(k, 0.0) + v
            m_ev[1].synchronize()
            timed["search_init"] = timed.get("search_init", 0.0) + m_ev[0].elapsed_time(m_ev[1])

    # ---- device-resident timing: K steps, CUDA events on the library's streams
    m_done.record(ms)
    for _ in range(args.warmup):
        step()
    ex.sync(); mt.sync()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stage_sum = {}
    e0.record(xs)
    for _ in range(args.steps):
        step()
    e1.record(ms)
    ex.sync(); mt.sync()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    # the same K steps again with the library's per-stage events on (stages serialised on one
    # stream, so each kernel's duration is its own): feeds roofline.per_kernel
    ex.set_profiling(True)
    for _ in range(args.steps):
        step(stage_sum)
    ex.sync(); mt.sync()
    ex.set_profiling(False)
    clocks = sampler.stop()
    launches = (ex.last_launches + 3) * args.steps        # search_init: grid + top-k + greedy kernels
    kps_h, desc_h, counts = ex.fetch()
    mean_kp = float(counts.mean())
    mean_matches = float(d_nm[:B - 1].float().mean().item()) if B > 1 else 0.0
    # what the timed loop computed, against the oracle, once and outside the timed region: the first four pairs
    bench_path_parity = "not checked"
    if rank == 0 and B >= 5:
        from oracle import oracle_py as O
        scale_f = ex.GetScaleFactors()
        m12_h, nm_h = d_m12[:4].cpu().numpy(), d_nm[:4].cpu().numpy()
        bad = 0
        for p_ in range(4):
            n1, n2 = int(counts[p_]), int(counts[p_ + 1])
            prev_ = np.ascontiguousarray(np.stack([kps_h[p_]["x"][:n1], kps_h[p_]["y"][:n1]], 1), np.float32)
            no_, mo_ = O.search_for_initialization(O.frame_view(kps_h[p_][:n1], desc_h[p_][:n1], W, H, scale_f),
                                                   O.frame_view(kps_h[p_ + 1][:n2], desc_h[p_ + 1][:n2], W, H, scale_f), prev_, 100, 0.9, True)
            bad += int(no_ != int(nm_h[p_])) + int(not np.array_equal(mo_, m12_h[p_][:n1]))
        bench_path_parity = "ok" if bad == 0 else f"MISMATCH in {bad} of 8 checks"
        if bad:
            raise SystemExit("bench.py: SearchForInitialization results of the timed path differ from the oracle: " + bench_path_parity)

    # ---- end to end through the host-buffer C-ABI call: pinned frames in, results out
    h_frames = orbfe.pinned_empty(frames.shape, np.uint8)
    h_frames[...] = frames

    # Two batches in flight, the way a streaming caller uses the synchronous host-buffer call: two extractor / matcher
    # handles driven from two host threads (the reference itself extracts a stereo pair on two threads, src/Frame.cc:78-81),
    # so one batch's upload overlaps the other's pipeline drain and result download.  Every step is still one complete
    # orbfe_extract_batch + SearchForInitialization + result download for 1024 frames of host memory.
    class E2EWorker:
        def __init__(self, ex_, mt_):
            self.ex, self.mt = ex_, mt_
            self.out = (orbfe.pinned_empty((B, cap), orbfe.KP_DTYPE), orbfe.pinned_empty((B, cap, 32), np.uint8),
                        orbfe.pinned_empty((B,), np.int32))
            self.d_m12 = torch.empty((max(B - 1, 1), cap), dtype=torch.int32, device="cuda")
            self.d_nm = torch.zeros(max(B - 1, 1), dtype=torch.int32, device="cuda")
            self.h_m12 = torch.empty((max(B - 1, 1), cap), dtype=torch.int32).pin_memory()
            self.h_nm = torch.empty(max(B - 1, 1), dtype=torch.int32).pin_memory()

        def step(self):
            self.ex.extract_batch(h_frames, out=self.out)   # H2D frames, kernels, D2H keypoints + descriptors
            kp_ptr, desc_ptr, cnt_ptr, dcap = self.ex.device_results()
            self.mt.search_init_consecutive_device(kp_ptr, desc_ptr, cnt_ptr, dcap, B, W, H, self.d_m12.data_ptr(),
                                                   self.d_nm.data_ptr(), 100)
            self.mt.fetch(self.h_m12.data_ptr(), self.d_m12.data_ptr(), self.h_m12.numel() * 4)
            self.mt.fetch(self.h_nm.data_ptr(), self.d_nm.data_ptr(), self.h_nm.numel() * 4)
            self.mt.sync()

    in_flight = 1 if os.environ.get("ORBFE_BENCH_E2E_SINGLE") else int(os.environ.get("ORBFE_BENCH_E2E_INFLIGHT", "2"))
    workers = [E2EWorker(ex, mt)]
    for _ in range(in_flight - 1):
        workers.append(E2EWorker(orbfe.ORBextractor(NFEAT, 1.2, 8, 20, 7, device=local_rank, max_width=W, max_height=H, max_batch=B),
                                 orbfe.ORBmatcher(0.9, True, device=local_rank)))
    out = workers[0].out
    h_m12, h_nm = workers[0].h_m12, workers[0].h_nm

    def run_steps(n):
        """n steps, dealt round-robin to the workers; returns when all have finished"""
        share = [len(range(w, n, len(workers))) for w in range(len(workers))]

        def loop(w):
            for _ in range(share[w]):                        # ctypes calls only: the worker threads never touch torch
                workers[w].step()
        th = [threading.Thread(target=loop, args=(w,)) for w in range(1, len(workers))]
        [t_.start() for t_ in th]
        loop(0)
        [t_.join() for t_ in th]

    run_steps(max(2, min(args.warmup, 2)) * len(workers))
    barrier()
    t0 = time.perf_counter()
    run_steps(args.steps)
    barrier()
    e2e_s = time.perf_counter() - t0
    # the overlapped schedule computes what the resident one does: same match counts from both workers
    for w_ in workers:
        assert torch.equal(w_.h_nm, d_nm.cpu()), "e2e worker results differ from the resident path"

    # ---- the PCIe leg alone: one pinned H2D copy of the step's frames (what bounds e2e from below)
    th = torch.from_numpy(h_frames)
    td = torch.empty_like(th, device="cuda")
    td.copy_(th, non_blocking=True)
    torch.cuda.synchronize()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0.record()
    for _ in range(3):
        td.copy_(th, non_blocking=True)
    c1.record()
    torch.cuda.synchronize()
    h2d_gbs = 3 * th.numel() / (c0.elapsed_time(c1) * 1e-3) / 1e9
    del td

    # ---- Hamming half of the metric: configs[4] as specified — 1M x 1M, identical inputs on every rank, queries sharded,
    # packed 8-byte results all-gathered over NCCL inside the timed region (orbfe_knn2_sharded, strong scaling); parity
    # (sharded == one GPU == oracle subsample, planted property on all queries) checked outside it
    del d_frames
    torch.cuda.empty_cache()
    from tools.knn_sharded import run as knn_run
    knn_n = 8192 if args.no_knn else KNN_N
    knn_line, knn_ok = knn_run(knn_n, knn_n, engine="auto", steps=3, warmup=1, check=True, oracle_sub=256,
                               rank=rank, world=world, local_rank=local_rank)
    knn_line = knn_line or {"value": 0.0, "ms_per_step": 0.0, "gather_us": 0.0, "gather_bytes_total": 0, "parity": "other rank"}
    if not knn_ok:
        raise SystemExit("bench.py: sharded kNN results differ from the single-GPU / oracle results")
    # the integer-pipe engine north_star names, on one GPU's sample (131072 x 131072): its fraction of the popc pipe
    MQ = NT = 131072
    tq = torch.randint(0, 256, (MQ, 32), dtype=torch.uint8, device="cuda")
    tt = torch.randint(0, 256, (NT, 32), dtype=torch.uint8, device="cuda")
    ko = torch.empty((3, MQ), dtype=torch.int32, device="cuda")
    mt.set_knn_engine("popc")
    mt.knn2_device(tq.data_ptr(), MQ, tt.data_ptr(), NT, ko[0].data_ptr(), ko[1].data_ptr(), ko[2].data_ptr(), 50, 0.7)
    mt.sync()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record(ms)
    for _ in range(3):
        mt.knn2_device(tq.data_ptr(), MQ, tt.data_ptr(), NT, ko[0].data_ptr(), ko[1].data_ptr(), ko[2].data_ptr(), 50, 0.7)
    k1.record(ms)
    mt.sync()
    mt.set_knn_engine("auto")
    knn_ms = k0.elapsed_time(k1) / 3

    t = torch.tensor([dev_ms, e2e_s * 1e3, knn_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms, knn_ms = t.tolist()

    if rank == 0:
        total_frames = B * world * args.steps
        ab = algorithmic_bytes(W, H)
        stage_ms = {k: v / args.steps for k, v in stage_sum.items()}
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        dom = max(("pyramid", "fast", "blur"), key=lambda k: stage_ms.get(k, 0.0))
        achieved = ab[dom] * B / (stage_ms[dom] * 1e-3) / 1e9
        per_kernel = {k: {"ms": stage_ms[k], "gbs": ab[k] * B / (stage_ms[k] * 1e-3) / 1e9 if k in ab else None}
                      for k in stage_ms}
        popc_peak, popc_src = 557.1, "tools/microbench on this pool (profiles/r01_microbench.json)"
        try:
            mb = subprocess.run([os.path.join(ROOT, "tools", "microbench")], capture_output=True, text=True, timeout=60)
            popc_peak = float(json.loads(mb.stdout.strip().splitlines()[-1])["popc_gcmp_s"])
            popc_src = "tools/microbench run live"
        except Exception:
            pass
        popc_gcmp = MQ * NT / (knn_ms * 1e-3) / 1e9                       # one GPU, integer-pipe engine
        knn_gcmp = float(knn_line["value"])                               # whole job, configs[4], gather included
        # a 256-bit comparison on the tensor cores is 256 multiply-adds.  The AUTO engine at this size is the block-scaled
        # FP4 one (kind::mxf4), nominally four times the bf16 rate; MEASURED_PEAKS has the sustained bf16 cuBLAS figure.
        # Beside that paper peak: the raw mxf4 issue rate of this very MMA shape family measured on this box, no epilogue,
        # no operand traffic (tools/mma_rate_probe, about 2 s), which is the kernel's own ceiling
        fp4_peak_tops = 4.0 * float(peaks.get("bf16_tflops_sustained", 1400.0))
        i8_peak_tops = fp4_peak_tops                                      # key name kept from the s8 engine's rounds
        probe_tcmps, probe_src = 16.4, "tools/mma_rate_probe on this pool (profiles/r02_mma_rate_probe.json)"
        try:
            pr = subprocess.run([os.path.join(ROOT, "tools", "mma_rate_probe")], capture_output=True, text=True, timeout=60)
            probe_tcmps = float(json.loads(pr.stdout)["mxf4_m128_n128"]["tcmp_s"])
            probe_src = "tools/mma_rate_probe run live"
        except Exception:
            pass
        knn_tops = knn_gcmp / world * 512.0 / 1e3
        desc_bytes = mean_kp * 544.0                                      # N_kp x (512 B gathered + 32 B out)
        frac = {k: ab[k] * B / (stage_ms[k] * 1e-3) / 1e9 / peak for k in ("pyramid", "fast", "blur")}
        frac["desc"] = desc_bytes * B / (stage_ms["orient_desc"] * 1e-3) / 1e9 / peak
        step_bytes = (sum(ab.values()) + mean_kp * (749.0 + 544.0 + 60.0)) * B   # stage-separated model of SURVEY 8d
        hamming = {"metric": "Hamming Gcmp/s (brute-force kNN-2, 256-bit rows)", "value": knn_gcmp, "unit": "Gcmp/s",
                   "sharded": knn_line, "popc_engine_one_gpu": {"gcmps": popc_gcmp, "queries": MQ, "train": NT, "ms": knn_ms,
                                                                "peak": popc_peak, "frac": popc_gcmp / popc_peak, "peak_source": popc_src}}
        cfg = workload_config(B)
        line = {
            "metric": METRIC, "value": total_frames / (dev_ms * 1e-3), "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": cfg,
            "stats": {"mean_keypoints_per_frame": mean_kp, "mean_init_matches_per_pair": mean_matches,
                      "bench_path_parity": bench_path_parity, "knn_parity": knn_line["parity"]},
            "e2e": {"value": total_frames / (e2e_ms * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": int(frames.nbytes), "h2d_gbs_measured": h2d_gbs,
                    "pcie_bound_frames_per_s": h2d_gbs * 1e9 / (W * H) * world, "batches_in_flight": in_flight,
                    "d2h_bytes_per_step": int(out[0].nbytes + out[1].nbytes + out[2].nbytes
                                              + h_m12.numel() * 4 + h_nm.numel() * 4)},
            "hamming": hamming,
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "algorithmic_bytes": ab[dom] * B,
                         "traffic": NCU_DRAM_BYTES_PER_1024[dom] * B / 1024.0,
                         "traffic_source": "ncu --set full at 1024 frames (profiles/), scaled to this batch",
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s",
                         # scalars the driver's record keeps: per-kernel HBM fractions (algorithmic bytes of SURVEY 8d / stage time / peak)
                         "frac_pyramid": frac["pyramid"], "frac_fast": frac["fast"], "frac_blur": frac["blur"], "frac_desc": frac["desc"],
                         "frac_step": step_bytes / (dev_ms / args.steps * 1e-3) / 1e9 / peak,
                         "ms_pyramid": stage_ms["pyramid"], "ms_fast": stage_ms["fast"], "ms_blur": stage_ms["blur"],
                         "ms_octree": stage_ms["octree"], "ms_desc": stage_ms["orient_desc"], "ms_search_init": stage_ms["search_init"],
                         # Hamming half of the metric (configs[4], whole job) and its rooflines, per GPU
                         "knn_gcmps": knn_gcmp, "knn_ms": float(knn_line["ms_per_step"]), "knn_gather_us": float(knn_line["gather_us"]),
                         "knn_gather_bytes": int(knn_line["gather_bytes_total"]),
                         "knn_x_popc_peak": knn_gcmp / world / popc_peak, "knn_tensor_tops": knn_tops, "knn_tensor_peak_tops": i8_peak_tops,
                         "knn_frac_tensor": knn_tops / i8_peak_tops,
                         # the same against four times the BURST bf16 figure (a 0.1 s kernel timed alone sits between the two)
                         "knn_frac_tensor_burst": knn_tops / (4.0 * float(peaks.get("bf16_tflops", 1650.0))),
                         "knn_mma_probe_gcmps": probe_tcmps * 1e3, "knn_frac_mma_probe": knn_gcmp / world / (probe_tcmps * 1e3),
                         "knn_popc_gcmps": popc_gcmp, "knn_popc_peak": popc_peak, "knn_frac_popc": popc_gcmp / popc_peak,
                         "note": "kernels are integer-issue bound, not HBM bound (ncu, profiles/); kNN: tcgen05 mxf4 engine vs 4x sustained bf16 cuBLAS",
                         "per_kernel": per_kernel},
        }
        if not args.no_cpu_baseline:
            r = cpu_reference_fps(frames[:256], seconds_budget=15.0)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
            ck = cpu_knn_gcmps()
            line["cpu_baseline"].update({"knn_gcmps": ck["gcmps"], "knn_sample": ck["sample"],
                                         "knn_1Mx1M_seconds_extrapolated": KNN_N * KNN_N / (ck["gcmps"] * 1e9)})
            line["roofline"]["knn_cpu_gcmps"] = ck["gcmps"]
            line["guided_searches"] = guided_search_timings(frames[0])
            line["stereo"] = stereo_timings()
            line["latency"] = single_frame_latency(frames)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

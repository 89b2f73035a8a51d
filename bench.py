#!/usr/bin/env python
"""Headline benchmark: frames/s of ORB extract + match at 1280x720, 2000 keypoints per eye (BASELINE.json configs[1]).

A step = `--inner` consecutive batches of synthetic ZED-shaped stereo frames through the hot path: both eyes extracted (pyramid,
FAST cells, quadtree, blur, orientation, rBRIEF) and the stereo row matcher run on every pair. (One batch is 1.2 ms: the inner
loop makes 20 steps last more than a second, so that the clock record covers sustained load.)
  value         device-resident batches (inputs already in HBM), CUDA-event timed on the library's streams
  e2e           the same batches through the host-buffer C ABI: ONE host thread per GPU, two batches in flight
                (lpslam_stereo_frames_submit / _collect), pinned host -> device copies and result read-back inside the timed region
  roofline      the dominant kernel against the bound that holds for it (k_fast_cells: instruction issue), plus the two HBM-bound
                stages (pyramid, blur) against MEASURED_PEAKS.json and against the nominal 8 TB/s
  c1 / c3       BASELINE configs[0] (640x480, 1000 kpts, extract + brute-force match of consecutive frames) and configs[2]
                (3840x2160, 4000 kpts, extract + frame-to-frame projection matcher)
  latency_1pair one stereo pair per call (the reference's live call pattern, data/frame.cc:55-93)
  replay_4096   BASELINE configs[3]: 4096 distinct frames split over the ranks (strong scaling), resident and from the host
  knn           BASELINE configs[4]: 2000 queries against 10 M rows sharded over the ranks, planted neighbours asserted
  cpu_baseline / --impl reference   the reference's own sources (oracle/_ref) on the host cores
N > 1 (torchrun): frame-parallel, every rank extracts its own frames, no data-path collective (weak scaling of `value`).
"""
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

W, H, KPTS = 1280, 720, 2000
FX_B, BASELINE = 700.0 * 0.12, 0.12  # ZED-like: fx 700 px, 12 cm baseline -> max disparity 700 px
WORKLOAD = "ZED stereo 1280x720 pair, 2000 kpts/eye, ORB extract (8 levels x1.2) + stereo row-search matching (BASELINE configs[1])"
NOMINAL_HBM_GBS = 8000.0  # north_star: "roughly 8 TB/s"
POPC_WARP_INSTR_PER_S = 144.6e9  # whole-GPU POPC rate measured by tools/pipe_rates.cu (profiles/r1_pipe_rates.txt)


def level_pixels(w, h, levels=8, sf=1.2):
    s, out = np.float32(1.0), [(w, h)]
    for _ in range(1, levels):
        s = np.float32(sf) * s
        out.append((int(round(w * 1.0 / float(s))), int(round(h * 1.0 / float(s)))))
    return out


def algorithmic_bytes(w, h):
    """un-fused compulsory traffic per frame and stage."""
    px = [a * b for a, b in level_pixels(w, h)]
    return {"pyramid": sum(px[l - 1] + px[l] for l in range(1, len(px))), "blur": 2 * sum(px), "fast": sum(px)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [int(r[0]) for r in self.rows if r and r[0].isdigit()]
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].startswith("Active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


# ---- CPU arm: the reference's own sources (oracle/_ref) on the host cores -------------------------------------------------

def cpu_reference_arm(steps, warmup, sample_pairs=1):
    """One process per core, each extracting both eyes of a stereo pair and matching them, as data/frame.cc:55-93 does per frame."""
    import multiprocessing as mp
    cores = max(1, min(os.cpu_count() or 1, 64))
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        pool.map(_cpu_worker, [(0, 1)] * cores)  # builds / loads the oracle in every worker
        times = []
        for s in range(warmup + steps):
            t = time.perf_counter()
            pool.map(_cpu_worker, [(s * cores + k, sample_pairs) for k in range(cores)])
            if s >= warmup:
                times.append(time.perf_counter() - t)
    dt = sum(times)
    return steps * cores * sample_pairs / dt, cores, dt / steps * 1e3, kind_of_oracle(), time.perf_counter() - t0


def kind_of_oracle():
    from oracle import pyoracle
    return "reference" if pyoracle.ref() is not None else "port"


_cpu_state = {}


def _cpu_worker(arg):
    seed, n_pairs = arg
    from oracle import pyoracle
    from lpslam_b200.synth import synth_stereo_pair
    if "o" not in _cpu_state:
        o = pyoracle.ref() or pyoracle.port()
        _cpu_state.update(o=o, el=o.extractor(max_kpts=KPTS), er=o.extractor(max_kpts=KPTS), pair=synth_stereo_pair(H, W, 2000))
    o, el, er = _cpu_state["o"], _cpu_state["el"], _cpu_state["er"]
    left, right = _cpu_state["pair"]
    n = 0
    for _ in range(n_pairs):
        kl, dl = el.extract(left)
        kr, dr = er.extract(right)
        xr, _ = o.match_stereo([el.level(l) for l in range(8)], [er.level(l) for l in range(8)], kl, kr, dl, dr, el.table(0), el.table(1),
                               FX_B, BASELINE)
        n += int((xr >= 0).sum())
    return n


def _cpu_extract_worker(arg):
    """extractor part of configs[0] / configs[2] on one core: (h, w, kpts, frames)"""
    h, w, kpts, frames = arg
    from oracle import pyoracle
    from lpslam_b200.synth import synth_image
    key = ("x", h, w)
    if key not in _cpu_state:
        o = pyoracle.ref() or pyoracle.port()
        _cpu_state[key] = (o.extractor(max_kpts=kpts), synth_image(h, w, 1000))
    e, img = _cpu_state[key]
    t = time.perf_counter()
    for _ in range(frames):
        e.extract(img)
    return time.perf_counter() - t


def cpu_extract_rate(h, w, kpts, frames_per_core):
    import multiprocessing as mp
    cores = max(1, min(os.cpu_count() or 1, 64))
    with mp.get_context("fork").Pool(cores) as pool:
        pool.map(_cpu_extract_worker, [(h, w, kpts, 1)] * cores)
        t = time.perf_counter()
        pool.map(_cpu_extract_worker, [(h, w, kpts, frames_per_core)] * cores)
        dt = time.perf_counter() - t
    return cores * frames_per_core / dt, cores


KERNEL_OF_STAGE = {"pyramid": "k_resize_walk", "blur": "k_gauss7_tma", "fast": "k_fast_cells", "quadtree": "k_quadtree",
                   "describe": "k_describe+k_stereo_match"}
LAUNCHES_PER_BATCH = 17  # 7 x k_resize_walk, k_pad_borders, k_fast_cells, k_quadtree, k_gauss7_tma, k_offsets, k_describe, 4 x k_stereo_*
KNN_ROWS_TOTAL, KNN_QUERIES = 10_000_000, 2000  # BASELINE configs[4]
REPLAY_FRAMES = 4096  # BASELINE configs[3]


def gpu_arm(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from lpslam_b200 import _lib, knn, match, orb, replay  # noqa: F401  (match / knn / replay register their entry points)
    from lpslam_b200._lib import check, lib
    from lpslam_b200.synth import synth_image, synth_stereo_pair

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: liblpslam_cuda has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, inner = args.batch, args.inner
    L = lib()
    props = torch.cuda.get_device_properties(local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def timed_dev(fn, steps, warmup, on_streams):
        """CUDA events on the library's own streams, barrier + synchronize on both sides, max over ranks"""
        for _ in range(warmup):
            fn()
        barrier()
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in on_streams]
        ev0.record(on_streams[0])  # everything is idle after the barrier: the first stream's mark is the start of all
        for _ in range(steps):
            fn()
        for ev, st in zip(ev1, on_streams):
            ev.record(st)
        torch.cuda.synchronize()
        ms = max_over_ranks(max(ev0.elapsed_time(ev) for ev in ev1))
        barrier()
        return ms

    def timed_wall(fn, warm_fn=None):
        if warm_fn:
            warm_fn()
        barrier()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        barrier()
        return ms

    # ---- the batch: B DISTINCT synthetic stereo pairs per rank, in page-locked host memory ----
    host = replay.pinned_array((2 * B, H, W))
    for p in range(B):
        host[2 * p], host[2 * p + 1] = synth_stereo_pair(H, W, 2000 + 1000 * rank + p)

    # ---- device resident: two extractor contexts (own stream, own resident batch), consecutive batches alternate between them, so
    # that a batch's latency-bound tail (quadtree, descriptors, stereo) runs beside the next batch's FAST ----
    n_dev = max(1, int(os.environ.get("LPSLAM_DEV_CONTEXTS", "2")))
    dev_ctx = []
    n_kp = n_match = 0
    for _ in range(n_dev):
        e = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=2 * B)
        res = match.stereo_frames(e, [host[2 * p] for p in range(B)], [host[2 * p + 1] for p in range(B)], FX_B, BASELINE, cap_per_frame=4096)
        n_kp = sum(len(r["keypts_left"]) + len(r["keypts_right"]) for r in res)
        n_match = sum(int((r["stereo_x_right"] >= 0).sum()) for r in res)
        p_in, pitch_in, _ = e.input_dev(W, H)
        dev_ctx.append((e, p_in, pitch_in))
    ext = dev_ctx[0][0]
    dev_streams = [torch.cuda.ExternalStream(L.lpslam_orb_stream(e._h), device=local) for e, _, _ in dev_ctx]
    turn = [0]

    def dev_step():
        for _ in range(inner):
            e, p_in, pitch_in = dev_ctx[turn[0] % len(dev_ctx)]
            turn[0] += 1
            check(L.lpslam_stereo_frames_dev(e._h, p_in, B, W, H, pitch_in, FX_B, BASELINE))

    clocks = ClockSampler(local)
    clocks.start()
    dev_ms = timed_dev(dev_step, args.steps, args.warmup, dev_streams)
    clk = clocks.stop()

    # ---- end to end: ONE host thread, two batches in flight on one context (submit / collect) ----
    e2e_ext = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=2 * B)
    pipe = replay.PipelinedReplay(e2e_ext, H, W, 2 * B, stereo=(FX_B, BASELINE))
    kp_seen = [0]
    e2e_ms = timed_wall(lambda: kp_seen.__setitem__(0, pipe.run(host, cycles=args.steps * inner)[1]),
                        lambda: pipe.run(host, cycles=max(args.warmup, 1) * 2))
    d2h_per_batch = n_kp * (28 + 32 + 8) + 4 * (2 * B + 2)

    # ---- per-stage device time of a few more resident batches (CUDA events inside the library, every stage alone on one stream) ----
    ext.enable_timing(True)
    stage = {k: [] for k in ("pyramid", "fast", "quadtree", "blur", "describe")}
    for _ in range(12):
        check(L.lpslam_stereo_frames_dev(ext._h, dev_ctx[0][1], B, W, H, dev_ctx[0][2], FX_B, BASELINE))
        st = ext.stage_times()
        for k in stage:
            stage[k].append(st[k])
    ext.enable_timing(False)
    stage = {k: float(np.median(v[2:])) for k, v in stage.items()}
    frames_img = 2 * B
    alg = algorithmic_bytes(W, H)
    peaks, traffic = {}, {}
    for name, dst in (("MEASURED_PEAKS.json", peaks), (os.path.join("profiles", "traffic.json"), traffic)):
        try:
            dst.update(json.load(open(os.path.join(ROOT, name))))
        except OSError:
            pass
    peak, peak_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs") if "hbm_gbs" in peaks else (6650.0, "B200_PROFILING.md fallback")
    per_stage = {}
    for k in ("pyramid", "blur", "fast"):
        gbs = alg[k] * frames_img / (stage[k] * 1e-3) / 1e9 if stage[k] > 0 else 0.0
        per_stage[k] = {"kernel": KERNEL_OF_STAGE[k], "bound": "hbm", "ms": round(stage[k], 4), "alg_bytes": alg[k] * frames_img,
                        "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4), "frac_of_nominal_8TBs": round(gbs / NOMINAL_HBM_GBS, 4),
                        "traffic": traffic.get(KERNEL_OF_STAGE[k])}
    per_stage["pyramid"]["launches"] = 7
    # k_fast_cells is bound by instruction issue, not by HBM: its warp instructions per launch (ncu, profiles/warp_instr.json, a
    # property of the workload) over the live stage time, against 4 issue slots per SM per clock at the maximum SM clock
    sm_clock_ghz = props.clock_rate / 1e6 if hasattr(props, "clock_rate") else 1.965
    peak_issue = props.multi_processor_count * 4 * sm_clock_ghz
    issue = None
    try:
        wi = json.load(open(os.path.join(ROOT, "profiles", "warp_instr.json")))
        if stage["fast"] > 0 and frames_img == wi["frames"]:
            ginstr = wi["k_fast_cells"] / (stage["fast"] * 1e-3) / 1e9
            issue = {"warp_instr_per_launch": wi["k_fast_cells"], "achieved": round(ginstr, 1), "peak": round(peak_issue, 1),
                     "unit": "1e9 warp instructions/s", "frac": round(ginstr / peak_issue, 4), "peak_source": "SMs x 4 schedulers x max SM clock"}
            per_stage["fast"].update(bound="issue", issue=issue)
    except (OSError, KeyError, ValueError):
        pass
    dom = max(stage, key=stage.get)
    if dom == "fast" and issue:
        roofline = {"bound": "issue", "kernel": "k_fast_cells", "achieved": issue["achieved"], "peak": issue["peak"], "unit": issue["unit"],
                    "frac": issue["frac"], "peak_source": issue["peak_source"], "traffic": traffic.get("k_fast_cells"),
                    "hbm_view": {"achieved_gbs": per_stage["fast"]["achieved_gbs"], "frac": per_stage["fast"]["frac"]}}
    else:
        rk = dom if dom in per_stage else "blur"
        roofline = {"bound": "hbm", "kernel": per_stage[rk]["kernel"], "achieved": per_stage[rk]["achieved_gbs"], "peak": peak,
                    "peak_source": peak_src, "unit": "GB/s", "frac": per_stage[rk]["frac"], "traffic": per_stage[rk]["traffic"]}
    roofline.update({"dominant_stage": dom, "stage_ms": {k: round(v, 4) for k, v in stage.items()}, "stages": per_stage,
                     "hbm_peak": {"measured_gbs": peak, "source": peak_src, "nominal_gbs": NOMINAL_HBM_GBS},
                     "stage_timing": "stage_ms: a separate pass with every stage alone on one stream (CUDA events inside the library, median "
                                     "of 10); in the timed region of 'value' the level-0 apron and the blur run on a side stream beside FAST"})

    extra = {}
    if world == 1 and not args.headline_only:
        extra["latency_1pair"] = latency_1pair(L, check, orb, match, host, local)
        extra["c1"] = config_line(L, check, orb, match, replay, torch, local, 640, 480, 1000, 64, "bruteforce", props, peak,
                                  cpu_frames_per_core=None if args.no_cpu_baseline else 4)
        extra["c3"] = config_line(L, check, orb, match, replay, torch, local, 3840, 2160, 4000, 8, "projection", props, peak,
                                  cpu_frames_per_core=None if args.no_cpu_baseline else 1)
    if not args.headline_only:
        extra["replay_4096"] = replay_4096(L, check, orb, replay, torch, host, B, rank, world, local, barrier, max_over_ranks)
        extra["knn"] = knn_line(L, check, knn, torch, dist, rank, world, local, timed_dev, args)

    pairs_total = B * inner * args.steps * world
    out = {"metric": "frames/s ORB extract+match @1280x720 2000kpt", "value": pairs_total / (dev_ms * 1e-3),
           "unit": "stereo frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "u8", "data": "synthetic",
           "config": {"workload": WORKLOAD, "pairs_per_batch_per_gpu": B, "batches_per_step": inner, "pairs_per_step_per_gpu": B * inner,
                      "distinct_pairs_per_gpu": B, "keypoints_per_batch": n_kp, "stereo_matches_per_batch": n_match,
                      "l2": "inputs + pyramids of one batch (%.0f MB) exceed the 126 MB L2" % (2 * B * 2 * 2.85),
                      "parallelism": "frame-parallel x%d" % world,
                      "contexts": "value: consecutive batches alternate between %d extractor contexts per GPU (own stream, own resident "
                                  "batch); e2e: one context, one host thread, two batches in flight" % len(dev_ctx)},
           "e2e": {"value": pairs_total / (e2e_ms * 1e-3), "unit": "stereo frames/s", "ms_per_step": e2e_ms / args.steps,
                   "h2d_bytes_per_step": inner * 2 * B * W * H, "d2h_bytes_per_step": inner * d2h_per_batch,
                   "h2d_gbs_achieved": inner * 2 * B * W * H / (e2e_ms / args.steps * 1e-3) / 1e9, "keypoints_read_back": kp_seen[0],
                   "timing": "wall clock around lpslam_stereo_frames_submit / _collect from ONE host thread per GPU (two batches in "
                             "flight: the copy-in of batch n + 1 and the read-back of batch n - 1 run beside the kernels of batch n), "
                             "barrier + synchronize on both sides, max over ranks"},
           "gpu_launches": args.steps * inner * LAUNCHES_PER_BATCH, "clocks": clk, "roofline": roofline}
    out.update(extra)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out, rank


def latency_1pair(L, check, orb, match, host, local):
    """one 1280x720 stereo pair per blocking call (data/frame.cc:55-93: the live tracker's pattern), 200 calls"""
    import ctypes as C
    e = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=2)
    cap = 2 * 4096
    from lpslam_b200._lib import KP_DTYPE
    kp, desc = np.empty(cap, KP_DTYPE), np.empty((cap, 32), np.uint8)
    xr, dep, offs = np.empty(cap, np.float32), np.empty(cap, np.float32), np.zeros(3, np.uint32)
    ts = []
    for i in range(220):
        p = i % (len(host) // 2)
        pl, pr = (C.c_void_p * 1)(host[2 * p].ctypes.data), (C.c_void_p * 1)(host[2 * p + 1].ctypes.data)
        t = time.perf_counter()
        check(L.lpslam_stereo_frames(e._h, pl, pr, 1, W, H, W, FX_B, BASELINE, kp.ctypes.data, desc.ctypes.data, cap, offs.ctypes.data,
                                     xr.ctypes.data, dep.ctypes.data))
        ts.append((time.perf_counter() - t) * 1e3)
    ts = np.sort(ts[20:])
    out = {"call": "lpslam_stereo_frames, n_pairs = 1, pinned host images, results on the host", "calls": len(ts),
           "median_ms": round(float(np.median(ts)), 4), "p99_ms": round(float(ts[int(0.99 * (len(ts) - 1))]), 4),
           "frames_per_s_at_median": round(1e3 / float(np.median(ts)), 1)}
    # the same pair through the C++ drop-in classes (cpp/build/dropin_smoke, built where the reference headers are), the way
    # data::frame's stereo constructor calls them
    binary = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cpp", "build", "dropin_smoke")
    if os.path.exists(binary):
        import subprocess
        import tempfile
        with tempfile.TemporaryDirectory() as d:
            host[0].tofile(os.path.join(d, "l.raw"))
            host[1].tofile(os.path.join(d, "r.raw"))
            env = dict(os.environ, LPSLAM_CUDA_DEVICE=str(local))
            r = subprocess.run([binary, str(W), str(H), str(KPTS), os.path.join(d, "l.raw"), os.path.join(d, "r.raw"), "latency:200",
                                repr(FX_B), repr(BASELINE)], capture_output=True, text=True, env=env, timeout=300)
        out["dropin_cpp"] = json.loads(r.stdout) if r.returncode == 0 else {"unavailable": (r.stderr or "failed")[-200:]}
    else:
        out["dropin_cpp"] = {"unavailable": "cpp/build/dropin_smoke not built"}
    return out


def config_line(L, check, orb, match, replay, torch, local, w, h, kpts, batch, matcher, props, peak, cpu_frames_per_core):
    """BASELINE configs[0] / configs[2]: mono frames, extract + the config's matcher between consecutive frames."""
    from lpslam_b200.synth import synth_image
    e = orb.orb_extractor(max_num_keypts=kpts, device=local, max_batch=batch)
    host = replay.pinned_array((batch, h, w))
    base = synth_image(h, w, 1000)
    for f in range(batch):  # a camera panning over the scene: consecutive frames overlap, so that the matchers find matches
        host[f] = np.roll(base, (2 * f, 3 * f), axis=(0, 1))
    res = e.extract_batch([host[f] for f in range(batch)], cap_per_frame=2 * kpts + 1024)
    n_rows = sum(len(k) for k, _ in res)
    p_in, pitch_in, _ = e.input_dev(w, h)
    stream = torch.cuda.ExternalStream(L.lpslam_orb_stream(e._h), device=local)
    reps = 20

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    ext_ms = timed(lambda: check(L.lpslam_orb_extract_batch_dev(e._h, p_in, batch, w, h, pitch_in)))
    out = {"workload": "%dx%d mono, %d kpts, 8 levels x1.2, batch of %d frames" % (w, h, kpts, batch), "keypoints_per_frame": n_rows // batch,
           "extract": {"ms_per_batch": round(ext_ms, 4), "frames_per_s": round(batch / (ext_ms * 1e-3), 1)}}
    # per-stage roofline of the extraction at this size
    e.enable_timing(True)
    st_acc = {}
    for _ in range(7):
        check(L.lpslam_orb_extract_batch_dev(e._h, p_in, batch, w, h, pitch_in))
        for k, v in e.stage_times().items():
            st_acc.setdefault(k, []).append(v)
    e.enable_timing(False)
    alg = algorithmic_bytes(w, h)
    out["roofline"] = {k: {"ms": round(float(np.median(st_acc[k][2:])), 4),
                           "achieved_gbs": round(alg[k] * batch / (float(np.median(st_acc[k][2:])) * 1e-3) / 1e9, 1),
                           "frac": round(alg[k] * batch / (float(np.median(st_acc[k][2:])) * 1e-3) / 1e9 / peak, 4),
                           "frac_of_nominal_8TBs": round(alg[k] * batch / (float(np.median(st_acc[k][2:])) * 1e-3) / 1e9 / NOMINAL_HBM_GBS, 4)}
                       for k in ("pyramid", "blur", "fast")}
    if matcher == "bruteforce":
        # robust::brute_force_match of every frame against the one before (frame_tracker.cc:162), descriptors resident
        def step():
            check(L.lpslam_orb_extract_batch_dev(e._h, p_in, batch, w, h, pitch_in))
            match.match_consecutive_frames(e, n_rows, batch, 0.8)
        t0 = time.perf_counter()
        for _ in range(3):
            step()
        t0 = time.perf_counter()
        for _ in range(reps):
            step()
        both_ms = (time.perf_counter() - t0) * 1e3 / reps
        m, cnt = match.match_consecutive_frames(e, n_rows, batch, 0.8)
        cmps = sum(len(res[f][0]) * len(res[f - 1][0]) for f in range(1, batch))
        out["value"] = round(batch / (both_ms * 1e-3), 1)
        out["unit"] = "frames/s (extract + brute-force match of consecutive frames, images resident, match indices read back)"
        out["match"] = {"ms_per_batch": round(both_ms - ext_ms, 4), "matches_per_frame": int(cnt[1:].mean()),
                        "gcmp_per_s": round(cmps / ((both_ms - ext_ms) * 1e-3) / 1e9, 1),
                        "popc_frac": round(cmps / ((both_ms - ext_ms) * 1e-3) * 4 / 32 / POPC_WARP_INSTR_PER_S, 4)}
    else:
        # projection::match_current_and_last_frames per frame pair through its host-buffer entry point (poses, landmarks and
        # keypoints are the tracker's host objects): extraction stays batched and resident, the matcher is a call per frame
        kp0, d0 = res[0]
        kp1, d1 = res[1]
        n0 = len(kp0)
        cam = match.make_camera(fx=2100.0, fy=2100.0, cx=w / 2.0, cy=h / 2.0, cols=w, rows=h, is_monocular=True)
        depth = np.random.default_rng(1).uniform(3.0, 9.0, n0)
        pos = np.stack([(kp0["x"] - w / 2.0) / 2100.0 * depth, (kp0["y"] - h / 2.0) / 2100.0 * depth, depth], axis=1)
        Tc = np.eye(4)
        Tc[:3, 3] = [0.01, 0.0, 0.02]
        sf = np.asarray(e.get_scale_factors(), np.float32)
        pm = match.projection(0.9, True)
        args_m = (cam, Tc, np.eye(4), np.ones(n0, np.uint8), np.zeros(n0, np.uint8), pos, d0, kp0["octave"], kp0["angle"], kp1, d1,
                  np.full(len(kp1), -1.0, np.float32), np.zeros(len(kp1), np.uint8), sf, 20.0)
        for _ in range(3):
            got, nm = pm.match_current_and_last_frames(*args_m)
        t0 = time.perf_counter()
        for _ in range(reps):
            got, nm = pm.match_current_and_last_frames(*args_m)
        m_ms = (time.perf_counter() - t0) * 1e3 / reps
        per_frame_ms = ext_ms / batch + m_ms
        out["value"] = round(1e3 / per_frame_ms, 1)
        out["unit"] = "frames/s (batched resident extraction + one frame-to-frame projection matcher call per frame)"
        out["match"] = {"ms_per_call": round(m_ms, 4), "landmarks": int(n0), "matches": int(nm)}
    # end to end: frames from page-locked host memory, two batches in flight, keypoints and descriptors read back
    pipe = replay.PipelinedReplay(e, h, w, batch, cap_per_frame=2 * kpts + 1024)
    pipe.run(host, cycles=2)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    cycles = 10
    pipe.run(host, cycles=cycles)
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / cycles
    out["e2e"] = {"extract_frames_per_s": round(batch / (e2e_ms * 1e-3), 1), "h2d_bytes_per_batch": batch * w * h,
                  "d2h_bytes_per_batch": n_rows * 60}
    if matcher == "bruteforce":
        out["e2e"]["value"] = round(batch / ((e2e_ms + out["match"]["ms_per_batch"]) * 1e-3), 1)
    else:
        out["e2e"]["value"] = round(1e3 / (e2e_ms / batch + out["match"]["ms_per_call"]), 1)
    if cpu_frames_per_core:
        v, cores = cpu_extract_rate(h, w, kpts, cpu_frames_per_core)
        out["cpu_baseline"] = {"value": round(v, 2), "unit": "frames/s (extractor only)", "cores": cores, "kind": kind_of_oracle(),
                               "sample": "%d processes x %d frames" % (cores, cpu_frames_per_core)}
    return out


def replay_4096(L, check, orb, replay, torch, pool_host, B, rank, world, local, barrier, max_over_ranks):
    """BASELINE configs[3]: 4096 DISTINCT 1280x720 frames, contiguous per-rank ranges (strong scaling), extraction only."""
    lo, hi = replay.frame_range(REPLAY_FRAMES, rank, world)
    n = hi - lo
    frames = replay.pinned_array((n, H, W))
    # distinct frames without minutes of numpy: every frame is one of the rank's 2 B synthetic images shifted by its own offset
    pool_dev = torch.from_numpy(np.ascontiguousarray(pool_host)).cuda()
    chunk = 64
    for c0 in range(0, n, chunk):
        c1 = min(c0 + chunk, n)
        stack = torch.stack([torch.roll(pool_dev[(lo + k) % len(pool_dev)], shifts=(((lo + k) * 7) % H, ((lo + k) * 13) % W), dims=(0, 1))
                             for k in range(c0, c1)])
        torch.from_numpy(frames[c0:c1]).copy_(stack)  # device -> page-locked host
    torch.cuda.synchronize()
    bf = 64
    e = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=bf)
    pipe = replay.PipelinedReplay(e, H, W, bf)
    pipe.run(frames, 0, min(n, 4 * bf))  # warm-up
    barrier()
    t0 = time.perf_counter()
    _, n_kp = pipe.run(frames)
    torch.cuda.synchronize()
    e2e_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    # resident: the rank's frames in HBM (n x 0.92 MB), batches of 64 through lpslam_orb_extract_batch_dev
    dev = torch.from_numpy(frames).cuda()
    stream = torch.cuda.ExternalStream(L.lpslam_orb_stream(e._h), device=local)

    def run_dev():
        for c0 in range(0, n, bf):
            check(L.lpslam_orb_extract_batch_dev(e._h, dev.data_ptr() + c0 * H * W, min(bf, n - c0), W, H, W))

    run_dev()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    run_dev()
    e1.record(stream)
    torch.cuda.synchronize()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    barrier()
    out = {"frames": REPLAY_FRAMES, "frames_per_rank": n, "scaling": "strong", "keypoints_this_rank": int(n_kp),
           "frames_per_s": round(REPLAY_FRAMES / (dev_ms * 1e-3), 1), "e2e_frames_per_s": round(REPLAY_FRAMES / (e2e_ms * 1e-3), 1),
           "e2e_h2d_gbs_per_gpu": round(n * H * W / (e2e_ms * 1e-3) / 1e9, 2), "efficiency_vs_n1": None,
           "how": "lpslam_b200.replay.PipelinedReplay (lpslam_orb_extract_batch_submit / _collect, two batches of 64 frames in flight, one "
                  "host thread per GPU); resident = lpslam_orb_extract_batch_dev over the rank's frames in HBM"}
    try:
        n1 = json.load(open(os.path.join(ROOT, "profiles", "r2_bench_n1.json")))["replay_4096"]
        out["efficiency_vs_n1"] = {"frames_per_s": round(out["frames_per_s"] / n1["frames_per_s"] / world, 3),
                                   "e2e_frames_per_s": round(out["e2e_frames_per_s"] / n1["e2e_frames_per_s"] / world, 3),
                                   "n1_source": "profiles/r2_bench_n1.json"}
    except (OSError, KeyError, ValueError):
        pass
    return out


def knn_line(L, check, knn, torch, dist, rank, world, local, timed_dev, args):
    """BASELINE configs[4]: 2000 queries against 10 M rows sharded over the ranks; queries are database rows with Binomial(256, 0.1)
    flipped bits, and every rank asserts that the merged best index is the planted row."""
    lo, hi = knn.shard_range(KNN_ROWS_TOTAL, rank, world)
    rows = hi - lo
    g = torch.Generator(device="cuda").manual_seed(5 + rank)
    shard = torch.randint(0, 256, (rows, 32), dtype=torch.uint8, device="cuda", generator=g)
    gq = torch.Generator(device="cuda").manual_seed(99)  # the same on every rank
    target = torch.randint(0, KNN_ROWS_TOTAL, (KNN_QUERIES,), device="cuda", generator=gq)
    flips = (torch.rand((KNN_QUERIES, 256), device="cuda", generator=gq) < 0.1).to(torch.uint8)
    weights = (2 ** torch.arange(8, device="cuda", dtype=torch.int32)).repeat(32)
    flip_bytes = (flips.to(torch.int32) * weights).view(KNN_QUERIES, 32, 8).sum(dim=2).to(torch.uint8)
    mine = (target >= lo) & (target < hi)
    q32 = torch.zeros((KNN_QUERIES, 32), dtype=torch.int32, device="cuda")
    q32[mine] = shard[(target[mine] - lo)].to(torch.int32)
    if world > 1:
        dist.all_reduce(q32, op=dist.ReduceOp.SUM)  # exactly one rank owns each planted row
    queries = (q32.to(torch.uint8) ^ flip_bytes).contiguous()
    db = knn.KnnDatabase(rows, device=local, index_base=lo)
    db.add_dev(shard.data_ptr(), rows)
    cur = torch.cuda.current_stream()
    merged = knn.sharded_query(db, queries, KNN_QUERIES, cur.cuda_stream)
    torch.cuda.synchronize()
    keys = merged.cpu().numpy().astype(np.uint64)
    idx, bd, sd = knn.unpack_keys_host(keys)
    planted_ok = bool(np.array_equal(idx, target.cpu().numpy()))
    assert planted_ok, "merged kNN result does not return the planted neighbours on rank %d" % rank
    want_d = flips.sum(dim=1).cpu().numpy()
    assert np.array_equal(bd, want_d.astype(np.uint32)), "best distance differs from the number of flipped bits"

    steps = max(args.steps, 5)
    knn_ms = timed_dev(lambda: knn.sharded_query(db, queries, KNN_QUERIES, cur.cuda_stream), steps, 3, [cur])
    cmp_per_step = KNN_QUERIES * KNN_ROWS_TOTAL
    gcmp = cmp_per_step * steps / (knn_ms * 1e-3) / 1e9
    per_gpu = gcmp / world
    return {"metric": "Gdesc-cmp/s kNN", "value": gcmp, "unit": "1e9 256-bit comparisons/s", "ms_per_query_batch": knn_ms / steps,
            "queries": KNN_QUERIES, "db_rows_total": KNN_ROWS_TOTAL, "db_rows_per_gpu": rows, "scaling": "strong",
            "exchange": "nccl all_gather of 2 keys/query + device merge" if world > 1 else "none",
            "planted_neighbours": {"checked_on_every_rank": True, "all_found": planted_ok, "mean_distance": float(want_d.mean()),
                                   "second_best_mean": float(sd.mean())},
            "roofline": {"bound": "popc pipe", "popc_per_comparison": 4, "achieved_gcmp_per_gpu": round(per_gpu, 1),
                         "peak_gcmp_per_gpu": round(POPC_WARP_INSTR_PER_S * 32 / 4 / 1e9, 1),
                         "frac": round(per_gpu * 1e9 * 4 / 32 / POPC_WARP_INSTR_PER_S, 4),
                         "peak_source": "POPC alone: 144.6 G warp instructions/s on one B200 (tools/pipe_rates.cu, profiles/r1_pipe_rates.txt); "
                                        "the carry-save-adder scan of csrc/knn.cu issues 4 POPC per 256-bit comparison"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32, help="stereo pairs per batch per GPU")
    ap.add_argument("--inner", type=int, default=48, help="batches per step (20 steps x 48 batches x 1.2 ms > 1 s)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--headline-only", action="store_true", help="skip c1 / c3 / latency / replay_4096 / knn")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", 0))

    if args.impl == "reference":
        if rank != 0:
            return
        # the driver's --steps / --warmup are honoured; a step is one stereo pair on every host core (about half a second)
        steps, warmup = max(1, args.steps), max(0, args.warmup)
        v, cores, ms, kind, took = cpu_reference_arm(steps, warmup)
        line = {"impl": "reference", "metric": "frames/s ORB extract+match @1280x720 2000kpt", "value": v,
                "unit": "stereo frames/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic", "config": {"workload": WORKLOAD, "parallelism": "%d host processes" % cores},
                "cpu_baseline": {"value": v, "unit": "stereo frames/s", "cores": cores, "kind": kind,
                                 "sample": "%d steps x %d processes x 1 stereo pair each (%.1f s)" % (steps, cores, took)},
                "e2e": {"value": v, "unit": "stereo frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return

    # stdout carries ONE line, the JSON: whatever libraries write to file descriptor 1 while the bench runs (NCCL prints its
    # version banner there at NCCL_DEBUG=VERSION / WARN, its log at INFO) is sent to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        out, rank = gpu_arm(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if rank == 0:
        if args.gpus == 1 and not args.no_cpu_baseline:
            v, cores, ms, kind, took = cpu_reference_arm(2, 0)
            out["cpu_baseline"] = {"value": v, "unit": "stereo frames/s", "cores": cores, "kind": kind,
                                   "sample": "2 steps x %d processes x 1 stereo pair each (%.1f s)" % (cores, took)}
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()

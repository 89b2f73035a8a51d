#!/usr/bin/env python
"""Headline benchmark: frames/s of ORB extract + match at 1280x720, 2000 keypoints per eye (BASELINE.json configs[1]).

A step = one batch of synthetic ZED-shaped stereo frames through the hot path: both eyes extracted (pyramid, FAST cells,
quadtree, blur, orientation, rBRIEF) and the stereo row matcher run on every pair.
  value     device-resident batches (inputs already in HBM), CUDA-event timed on the library's stream
  e2e       the same batches through the host-buffer C ABI (pinned host -> device copies and result read-back inside)
  roofline  the dominant kernel's algorithmic bytes / its measured duration against MEASURED_PEAKS.json
  cpu_baseline / --impl reference   the CPU oracle (the reference's own sources when oracle/_ref is built) on host cores
N > 1 (torchrun): frame-parallel replay, every rank extracts its own frames, no data-path collective (weak scaling).
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


def make_frames(n_pairs, seed0):
    """A small pool of distinct synthetic stereo pairs, cycled to fill the batch (generation is slow in numpy)."""
    from lpslam_b200.synth import synth_stereo_pair
    pool = [synth_stereo_pair(H, W, seed0 + k) for k in range(min(n_pairs, 4))]
    left = [pool[k % len(pool)][0] for k in range(n_pairs)]
    right = [pool[k % len(pool)][1] for k in range(n_pairs)]
    return left, right


def cpu_reference_arm(steps, warmup, sample_pairs=1):
    """The CPU oracle (reference sources in oracle/_ref when built, else the port) on the host cores: one process per
    core, each extracting both eyes of a stereo pair and matching them, as data/frame.cc:55-93 does per frame."""
    import multiprocessing as mp
    cores = max(1, min(os.cpu_count() or 1, 64))
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        pool.map(_cpu_worker, [(0, 1)] * cores)  # warm-up: builds / loads the oracle in every worker
        times = []
        for s in range(warmup + steps):
            t = time.perf_counter()
            pool.map(_cpu_worker, [(s * cores + k, sample_pairs) for k in range(cores)])
            if s >= warmup:
                times.append(time.perf_counter() - t)
    dt = sum(times)
    frames = steps * cores * sample_pairs
    return frames / dt, cores, dt / steps * 1e3, kind_of_oracle(), time.perf_counter() - t0


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
        _cpu_state["o"] = o
        _cpu_state["el"] = o.extractor(max_kpts=KPTS)
        _cpu_state["er"] = o.extractor(max_kpts=KPTS)
        _cpu_state["pair"] = synth_stereo_pair(H, W, 2000)
    o, el, er = _cpu_state["o"], _cpu_state["el"], _cpu_state["er"]
    left, right = _cpu_state["pair"]
    n = 0
    for _ in range(n_pairs):
        kl, dl = el.extract(left)
        kr, dr = er.extract(right)
        pl = [el.level(l) for l in range(8)]
        pr = [er.level(l) for l in range(8)]
        xr, _ = o.match_stereo(pl, pr, kl, kr, dl, dr, el.table(0), el.table(1), FX_B, BASELINE)
        n += int((xr >= 0).sum())
    return n


KERNEL_OF_STAGE = {"pyramid": "k_resize_walk", "blur": "k_gauss7", "fast": "k_fast_cells", "quadtree": "k_quadtree",
                   "describe": "k_describe+k_stereo_match"}
LAUNCHES_PER_STEP = 17  # 7 x k_resize_walk, k_pad_borders, k_fast_cells, k_quadtree, k_gauss7, k_offsets, k_describe, 4 x k_stereo_*
KNN_ROWS_PER_GPU, KNN_QUERIES = 1_250_000, 2000  # BASELINE configs[4]: 10 M rows over 8 GPUs, 2000 queries


def gpu_arm(args):
    import concurrent.futures as cf
    import ctypes as C
    import torch
    import torch.distributed as dist
    from lpslam_b200 import _lib, knn, match, orb  # noqa: F401  (match / knn register their entry points)
    from lpslam_b200._lib import check, lib

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: liblpslam_cuda has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG", "WARN")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch  # stereo pairs per step and per GPU
    L = lib()
    ext = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=2 * B)
    stream = torch.cuda.ExternalStream(L.lpslam_orb_stream(ext._h), device=local)

    lefts, rights = make_frames(B, 2000 + 16 * rank)
    # pinned host copies of the batch (what a capture / replay thread would hand over)
    hp = C.c_void_p()
    check(L.lpslam_cuda_host_alloc(C.byref(hp), 2 * B * W * H))
    host = np.ctypeslib.as_array(C.cast(hp, C.POINTER(C.c_uint8)), shape=(2 * B, H, W))
    for p in range(B):
        host[2 * p], host[2 * p + 1] = lefts[p], rights[p]

    def pinned(nbytes):
        q = C.c_void_p()
        check(L.lpslam_cuda_host_alloc(C.byref(q), nbytes))
        return q

    # ---- end to end: the replay a user runs. Two extractor contexts per GPU, each fed by its own host thread with half
    # of the step's pairs through the blocking host-buffer entry point, so one context's PCIe copies overlap the other's
    # kernels (the reference also extracts on two threads, data/frame.cc:65-66).
    halves = []
    n_ctx = max(1, min(int(os.environ.get("LPSLAM_E2E_CONTEXTS", "4")), B))
    for hx in range(n_ctx):
        lo = B * hx // n_ctx
        n = B * (hx + 1) // n_ctx - lo
        if n <= 0:
            continue
        e = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=2 * n)
        cap = 2 * n * 4096
        bufs = dict(kp=pinned(cap * 28), desc=pinned(cap * 32), xr=pinned(cap * 4), dep=pinned(cap * 4))
        offs = np.zeros(2 * n + 1, np.uint32)
        pl = (C.c_void_p * n)(*[host[2 * (lo + p)].ctypes.data for p in range(n)])
        pr = (C.c_void_p * n)(*[host[2 * (lo + p) + 1].ctypes.data for p in range(n)])
        halves.append((e, n, cap, bufs, offs, pl, pr))

    def run_half(hf):
        e, n, cap, bufs, offs, pl, pr = hf
        check(L.lpslam_stereo_frames(e._h, pl, pr, n, W, H, W, FX_B, BASELINE, bufs["kp"], bufs["desc"], cap, offs.ctypes.data,
                                     bufs["xr"], bufs["dep"]))
        return int(offs[2 * n])

    pool = cf.ThreadPoolExecutor(max_workers=len(halves))

    def e2e_step():
        return sum(pool.map(run_half, halves))

    n_kp = e2e_step()
    n_match = 0
    for e, n, cap, bufs, offs, pl, pr in halves:
        xr = np.ctypeslib.as_array(C.cast(bufs["xr"], C.POINTER(C.c_float)), shape=(cap,))[:int(offs[2 * n])]
        n_match += int((xr >= 0).sum())

    # ---- device resident: the context's own level-0 buffer, filled once through the host entry point ----
    okp, odesc, oxr, odep = pinned(2 * B * 4096 * 28), pinned(2 * B * 4096 * 32), pinned(2 * B * 4096 * 4), pinned(2 * B * 4096 * 4)
    offs_all = np.zeros(2 * B + 1, np.uint32)
    pl_all = (C.c_void_p * B)(*[host[2 * p].ctypes.data for p in range(B)])
    pr_all = (C.c_void_p * B)(*[host[2 * p + 1].ctypes.data for p in range(B)])
    check(L.lpslam_stereo_frames(ext._h, pl_all, pr_all, B, W, H, W, FX_B, BASELINE, okp, odesc, 2 * B * 4096, offs_all.ctypes.data,
                                 oxr, odep))
    dptr, dpitch, dstride = C.c_void_p(), C.c_size_t(), C.c_size_t()
    check(L.lpslam_orb_input_dev(ext._h, W, H, C.byref(dptr), C.byref(dpitch), C.byref(dstride)))

    # consecutive steps alternate between LPSLAM_DEV_CONTEXTS extractor contexts (own stream, own resident batch): a step's
    # latency-bound tail (quadtree, descriptors, stereo) then runs beside the next step's FAST instead of before it
    n_dev = max(1, int(os.environ.get("LPSLAM_DEV_CONTEXTS", "2")))
    dev_ctx = [(ext, dptr, dpitch)]
    for _ in range(n_dev - 1):
        e2 = orb.orb_extractor(max_num_keypts=KPTS, device=local, max_batch=2 * B)
        check(L.lpslam_stereo_frames(e2._h, pl_all, pr_all, B, W, H, W, FX_B, BASELINE, okp, odesc, 2 * B * 4096, offs_all.ctypes.data,
                                     oxr, odep))
        p2, pitch2, stride2 = C.c_void_p(), C.c_size_t(), C.c_size_t()
        check(L.lpslam_orb_input_dev(e2._h, W, H, C.byref(p2), C.byref(pitch2), C.byref(stride2)))
        dev_ctx.append((e2, p2, pitch2))
    dev_streams = [torch.cuda.ExternalStream(L.lpslam_orb_stream(e._h), device=local) for e, _, _ in dev_ctx]
    turn = [0]

    def dev_step():
        e, p_in, pitch_in = dev_ctx[turn[0] % len(dev_ctx)]
        turn[0] += 1
        check(L.lpslam_stereo_frames_dev(e._h, p_in, B, W, H, pitch_in.value, FX_B, BASELINE))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, on_streams):
        for _ in range(warmup):
            fn()
        barrier()
        ev0 = torch.cuda.Event(enable_timing=True)
        ev1 = [torch.cuda.Event(enable_timing=True) for _ in on_streams]
        t0 = time.perf_counter()
        ev0.record(on_streams[0])  # everything is idle after the barrier: the first stream's mark is the start of all
        for _ in range(steps):
            fn()
        for ev, st in zip(ev1, on_streams):
            ev.record(st)
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) * 1e3
        ms = max(ev0.elapsed_time(ev) for ev in ev1)
        t = torch.tensor([ms, wall], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        barrier()
        return float(t[0]), float(t[1])

    clocks = ClockSampler(local)
    clocks.start()
    dev_ms, _ = timed(dev_step, args.steps, args.warmup, dev_streams)
    turn[0] = 0  # the stage pass below uses the first context
    clk = clocks.stop()
    # end to end: every context's host thread replays its share of each step back to back (a replay thread per context,
    # no join between steps), so one context's PCIe copies run beside the others' kernels; wall clock, the calls block
    # until the results are on the host
    def e2e_replay(steps):
        def worker(hf):
            for _ in range(steps):
                run_half(hf)
        list(pool.map(worker, halves))

    e2e_replay(args.warmup)
    barrier()
    t0 = time.perf_counter()
    e2e_replay(args.steps)
    torch.cuda.synchronize()
    t = torch.tensor([(time.perf_counter() - t0) * 1e3], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    barrier()
    e2e_ms = float(t[0])

    # ---- per-stage device time of a few more resident steps (CUDA events on the library's stream) ----
    ext.enable_timing(True)
    stage = {k: 0.0 for k in ("pyramid", "fast", "quadtree", "blur", "describe")}
    reps = 5
    for _ in range(reps):
        check(L.lpslam_stereo_frames_dev(ext._h, dptr, B, W, H, dpitch.value, FX_B, BASELINE))
        st = ext.stage_times()
        for k in stage:
            stage[k] += st[k] / reps
    ext.enable_timing(False)
    frames_img = 2 * B
    alg = algorithmic_bytes(W, H)
    peaks, traffic = {}, {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))  # dram bytes per launch, ncu --set full
    except OSError:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs") if "hbm_gbs" in peaks else (6650.0, "B200_PROFILING.md fallback")
    per_stage = {}
    for k in ("pyramid", "blur", "fast"):
        gbs = alg[k] * frames_img / (stage[k] * 1e-3) / 1e9 if stage[k] > 0 else 0.0
        per_stage[k] = {"kernel": KERNEL_OF_STAGE[k], "ms": round(stage[k], 4), "alg_bytes": alg[k] * frames_img,
                        "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4), "traffic": traffic.get(KERNEL_OF_STAGE[k])}
    # k_fast_cells is bound by instruction issue, not by HBM: its warp instructions per launch (ncu, profiles/warp_instr.json,
    # a property of the workload) over the live stage time, against 4 issue slots per SM per clock at the maximum SM clock
    try:
        wi = json.load(open(os.path.join(ROOT, "profiles", "warp_instr.json")))
        if stage["fast"] > 0 and frames_img == wi["frames"]:
            sm_clock_ghz = torch.cuda.get_device_properties(local).clock_rate / 1e6  # kHz -> GHz
            ginstr = wi["k_fast_cells"] / (stage["fast"] * 1e-3) / 1e9
            peak_issue = torch.cuda.get_device_properties(local).multi_processor_count * 4 * sm_clock_ghz
            per_stage["fast"]["issue"] = {"warp_instr_per_launch": wi["k_fast_cells"], "achieved": round(ginstr, 1),
                                          "peak": round(peak_issue, 1), "unit": "1e9 warp instructions/s", "frac": round(ginstr / peak_issue, 4),
                                          "peak_source": "SMs x 4 schedulers x max SM clock"}
    except (OSError, KeyError, ValueError):
        pass
    dom = max(stage, key=stage.get)
    rk = dom if dom in per_stage else "blur"
    roofline = {"bound": "hbm", "kernel": per_stage[rk]["kernel"], "achieved": per_stage[rk]["achieved_gbs"], "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": per_stage[rk]["frac"], "traffic": per_stage[rk]["traffic"],
                "dominant_stage": dom, "stage_ms": {k: round(v, 4) for k, v in stage.items()}, "stages": per_stage,
                "stage_timing": "stage_ms: a separate pass with every stage alone on one stream (CUDA events inside the library); in "
                                "the timed region of 'value' the level-0 apron and the blur run on a side stream beside FAST / quadtree",
                "note": "k_fast_cells reads each level once (algorithmic bytes = sum of level pixels) but is bound by the "
                        "integer compare pipe, not by HBM; the HBM-bound kernels are k_resize_walk and k_gauss7 (stages)"}

    # ---- relocalisation kNN (BASELINE configs[4]): one database shard per GPU, NCCL all_gather of the top-2 keys, merge ----
    g = torch.Generator(device="cuda").manual_seed(5 + rank)
    shard = torch.randint(0, 256, (KNN_ROWS_PER_GPU, 32), dtype=torch.uint8, device="cuda", generator=g)
    gq = torch.Generator(device="cuda").manual_seed(99)
    queries = torch.randint(0, 256, (KNN_QUERIES, 32), dtype=torch.uint8, device="cuda", generator=gq)
    db = knn.KnnDatabase(KNN_ROWS_PER_GPU, device=local, index_base=rank * KNN_ROWS_PER_GPU)
    db.add_dev(shard.data_ptr(), KNN_ROWS_PER_GPU)
    cur = torch.cuda.current_stream()

    def knn_step():
        knn.sharded_query(db, queries, KNN_QUERIES, cur.cuda_stream)

    knn_ms, _ = timed(knn_step, max(args.steps, 5), 3, [cur])
    knn_steps = max(args.steps, 5)
    cmp_per_step = KNN_QUERIES * KNN_ROWS_PER_GPU * world
    knn_out = {"metric": "Gdesc-cmp/s kNN", "value": cmp_per_step * knn_steps / (knn_ms * 1e-3) / 1e9, "unit": "1e9 256-bit comparisons/s",
               "ms_per_query_batch": knn_ms / knn_steps, "queries": KNN_QUERIES, "db_rows_total": KNN_ROWS_PER_GPU * world,
               "db_rows_per_gpu": KNN_ROWS_PER_GPU, "exchange": "nccl all_gather of 2 keys/query + device merge" if world > 1 else "none",
               "popc_roofline_gcmp_per_gpu": 578.0,
               "note": "popc pipe measured at 16 lanes/clk/SM (profiles/r1_pipe_rates.txt): 8 POPC per comparison caps a GPU at 578 "
                       "Gcmp/s; the carry-save-adder scan needs 4"}

    pairs_total = B * args.steps * world
    out = {"metric": "frames/s ORB extract+match @1280x720 2000kpt", "value": pairs_total / (dev_ms * 1e-3),
           "unit": "stereo frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "u8", "data": "synthetic",
           "config": {"workload": "ZED stereo 1280x720 pair, 2000 kpts/eye, ORB extract (8 levels x1.2) + stereo row-search "
                                  "matching (BASELINE configs[1])", "pairs_per_step_per_gpu": B, "images_per_step_per_gpu": 2 * B,
                      "keypoints_per_step": n_kp, "stereo_matches_per_step": n_match,
                      "l2": "inputs + pyramids of one step (%.0f MB) exceed the 126 MB L2" % (2 * B * 2 * 2.85),
                      "parallelism": "frame-parallel x%d" % world,
                      "contexts": "value: consecutive steps alternate between %d extractor contexts per GPU (own stream, own "
                                  "resident batch); e2e: %d contexts per GPU" % (len(dev_ctx), len(halves))},
           "e2e": {"value": pairs_total / (e2e_ms * 1e-3), "unit": "stereo frames/s", "ms_per_step": e2e_ms / args.steps,
                   "h2d_bytes_per_step": 2 * B * W * H, "d2h_bytes_per_step": n_kp * (28 + 32 + 8) + 4 * (2 * B + 2),
                   # the end-to-end leg is bound by the host link: pinned H2D alone runs at 48-54 GB/s on this box
                   # (tools/pcie_rate.py), i.e. 1.1-1.2 ms for one step's 59 MB of images
                   "h2d_gbs_achieved": 2 * B * W * H / (e2e_ms / args.steps * 1e-3) / 1e9,
                   "timing": "wall clock around the blocking host-buffer C-ABI calls: %d contexts per GPU, each replayed by its own host thread (its share of every step, back to back), max over ranks" % len(halves)},
           "gpu_launches": args.steps * LAUNCHES_PER_STEP, "clocks": clk, "roofline": roofline, "knn": knn_out}
    pool.shutdown()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out, rank


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32, help="stereo pairs per step per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", 0))

    if args.impl == "reference":
        if rank != 0:
            return
        steps = min(args.steps, 5)
        v, cores, ms, kind, _ = cpu_reference_arm(steps, min(args.warmup, 1))
        line = {"impl": "reference", "metric": "frames/s ORB extract+match @1280x720 2000kpt", "value": v,
                "unit": "stereo frames/s", "n_gpus": args.gpus, "steps": steps, "warmup": min(args.warmup, 1),
                "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
                "data": "synthetic",
                "config": {"workload": "ZED stereo 1280x720 pair, 2000 kpts/eye, ORB extract (8 levels x1.2) + stereo "
                                       "row-search matching (BASELINE configs[1])", "parallelism": "%d host processes" % cores},
                "cpu_baseline": {"value": v, "unit": "stereo frames/s", "cores": cores, "kind": kind,
                                 "sample": "%d steps x %d processes x 1 stereo pair each" % (steps, cores)},
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

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

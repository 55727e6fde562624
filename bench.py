#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: fwd + bwd + Gram-loss steps/s @ 16384 samples.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload cfg1|cfg0|cfg2] [--precision bf16|fp32] [--pairs B]

A "step" is one (loss, d loss / d waveform) evaluation of methods.py's L-BFGS-B loop
(methods.py:133-137,167) on synthetic audio with a random-init NSynth encoder (no network: no
checkpoint, no wavs).  N=1 workload = BASELINE.json configs[1] (16384-sample pair, --gatys over all
30 layers, CUDA-graph-captured step).  N>1: one process per GPU, independent pairs per rank, no
data-path collective (weak scaling); --workload cfg2 instead time-shards one 65536-sample clip with
one NCCL all-reduce of the Gram partials per step (strong scaling).

Prints ONE JSON line (rank 0).  `value` = device-resident steps/s (CUDA-graph replay, CUDA events);
`e2e` = the same evaluation through ast_step_host with HOST float64 buffers, as SciPy calls it.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # name: (T, cont_ids, stack, gatys, description)
    "cfg1": (16384, [29], None, True, "BASELINE configs[1]: 16384-sample pair, --gatys, all 30 layers, lambd 100, gamma 0"),
    "cfg0": (16384, [25], 0, False, "BASELINE configs[0]: 16384-sample pair, channel-wise Gram, --cont_lyrs 25 --stack 0"),
    "cfg2": (65536, [25], 0, False, "BASELINE configs[2]: 65536-sample clip time-sharded over the ranks (cfg0 layer set)"),
}
SR = 16000


def algorithmic_flops(T, NL, L, gatys, B=1):
    """2 FLOP/MAC, fwd + dX-only bwd."""
    C, K = 128, 3
    gram = 4 * L * C * C * T if gatys else 4 * C * L * L * T
    return B * (4 * NL * T * C * C * (K + 1) + 4 * T * K * C + gram)


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return dict(tflops=float(d.get("bf16_tflops_sustained", d.get("bf16_tflops"))), hbm=float(d["hbm_gbs"]), src="measured")
    return dict(tflops=1400.0, hbm=6650.0, src="fallback")     # B200_PROFILING.md fallback


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (NVML; nvidia-smi is not on this image)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:      # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:       # noqa: BLE001
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "note": "NVML unavailable: " + self.err}
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def build_workload(name, B, seed_shift=0):
    from audio_style_transfer_b200 import synth
    from audio_style_transfer_b200.mu_law import mu_law_numpy
    T, cont, stack, gatys, desc = WORKLOADS[name]
    style = synth.resolve_style_ids(stack)
    NL = synth.layers_needed(cont, style)
    w = synth.make_weights(NL, seed=0, bias_scale=0.0)          # reference init: biases 0 (masked.py:117)
    cs = [mu_law_numpy(synth.content_audio(T, seed=1 + 10 * (seed_shift + i))).astype(np.float32) for i in range(B)]
    ss = [mu_law_numpy(synth.style_audio(T, seed=2 + 10 * (seed_shift + i), chunks=1)).astype(np.float32) for i in range(B)]
    xs = np.stack([synth.step_input(c, seed=3 + 10 * (seed_shift + i)) for i, c in enumerate(cs)])
    return dict(T=T, cont=cont, style=style, gatys=gatys, NL=NL, L=len(style), w=w, content=np.stack(cs),
                style_a=np.stack(ss), x=xs, desc=desc, lambd=100.0)

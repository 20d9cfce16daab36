#!/usr/bin/env python
"""bench.py — log_prob samples/s of the flow hot path on N B200s (contract in the task brief).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one log_prob pass of the named flow over one batch of synthetic samples.
Default workload `maf_mnist` = BASELINE.json configs[3], the configuration `metric` is quoted on
("MAF-MNIST"): the 8-layer RQS MaskedCoupling flow of src/maf_mnist.py (conditioner 784->512x3->19600,
Uniform base) on 4096 synthetic dequantised 28x28 images per GPU — log_prob (`value`), `train`, `sampling`,
`e2e`, `e2e_uint8`, `sustained`, `parity`, `roofline`, `cpu_baseline`.  The metric's second configuration
("RealNVP-2D", configs[0]) and every other workload are timed into `others`, each with its own `roofline`
and `kernel_ms` (weak scaling; no data-path collective — the only exchanges are the scalar mean-log-likelihood
all-reduce and, in `train`, the parameter-gradient all-reduce north_star names).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PEAKS_FALLBACK = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            with open(p) as f:
                d = json.load(f)
            d["_source"] = "measured"
            return d
        except Exception:
            pass
    d = dict(PEAKS_FALLBACK)
    d["_source"] = "fallback"
    return d


# ----------------------------------------------------------------------------- synthetic data
def gaussian_blobs(n, gen):
    """src/densities/gaussian_blobs.py:34-59: 8 blobs on a circle of radius 2, sigma 0.3."""
    k = torch.randint(0, 8, (n,), generator=gen)
    ang = k.double() * (2 * math.pi / 8)
    mu = torch.stack([2 * torch.cos(ang), 2 * torch.sin(ang)], -1)
    return (mu + 0.3 * torch.randn(n, 2, generator=gen, dtype=torch.float64)).float()


def banana(n, gen):
    """src/densities/banana.py:31-41."""
    x2 = torch.randn(n, generator=gen, dtype=torch.float64) * math.sqrt(3.0)
    x1 = (x2 * x2 / 3 - 2) + torch.randn(n, generator=gen, dtype=torch.float64) * (math.sqrt(3.0) / 2)
    return torch.stack([x1, x2], -1).float()


def mnist_like(n, gen):
    """src/maf_mnist.py:101-106 prepare_data on synthetic uint8 images: (u8 + U[0,1)) / 256."""
    return ((torch.randint(0, 256, (n, 784), generator=gen).float() + torch.rand(n, 784, generator=gen)) / 256.0)


# ----------------------------------------------------------------------------- workloads
class Workload:
    name = ""
    desc = ""
    n = 0               # samples per step per GPU
    bytes_per_sample = 0  # algorithmic HBM bytes
    flops_per_sample = 0  # algorithmic conditioner FLOPs (needed-only), SURVEY §8d
    bound = "hbm"
    dtype = "f32"
    launches_per_step = 1
    tc_launches = 2      # weight pack + fused stack kernel
    tc_dtype = "fp16x3 (fp16 hi/lo split, 3 passes, fp32 accumulate)"
    cpu_sample = 1 << 16
    # the highest roofline fraction the ALGORITHM admits on this path (stated next to `frac`)
    ceiling = 1.0
    ceiling_why = ""

    def make_inputs(self, gen, n):
        raise NotImplementedError

    def setup(self, dev):
        raise NotImplementedError

    def run(self, x):
        """public-API call: device tensor in -> log_prob on device"""
        raise NotImplementedError

    def cpu_run(self, x):
        raise NotImplementedError


class CouplingWorkload(Workload):
    ospec_name = "RQS_TOY"
    data = staticmethod(gaussian_blobs)
    transform = "rqs"
    precision = "tc"   # "tc": tensor-core path (3-pass fp16 hi/lo split at event_dim 2); "fp32": CUDA-core FP32 path
    tc_precision = 1   # FLOWCUDA_PRECISION_3XTF32
    ceiling = 1.0 / 3.0
    ceiling_why = ("error-compensated 3-pass products (hi*hi + hi*lo + lo*hi): every algorithmic FLOP costs three "
                   "tensor-pipe FLOPs; 1- and 2-pass forms miss the parity tolerance (profiles/r02/precision_study.txt)")

    def _ospec(self):
        from oracle import flows_oracle as O

        return getattr(O, self.ospec_name)

    def _params(self):
        from oracle import flows_oracle as O

        # params ~ N(0, 0.1^2)-scaled: non-trivial layers, not the zero init
        return O.init_coupling_params(self._ospec(), seed=1234, dtype=torch.float32, scale=0.1)

    def make_inputs(self, gen, n):
        return self.data(n, gen)

    def setup(self, dev):
        from normalizing_flows_b200 import flows
        from oracle import flows_oracle as O

        sp = self._ospec()
        self._p = self._params()
        flat = O.flatten_layers(self._p).float()
        self.model = flows.make_flow_model((sp.event_dim,), sp.num_layers, sp.hidden, transform=self.transform,
                                           num_bins=sp.num_bins, range_min=sp.range_min, range_max=sp.range_max,
                                           base=sp.base, params=flat.to(dev),
                                           precision=self.tc_precision if self.precision == "tc" else 0)
        if self.precision == "tc":
            self.launches_per_step = self.tc_launches
            self.dtype = self.tc_dtype

    def run(self, x):
        return self.model.log_prob(x)

    def sampling(self, dev, K=5):
        """Transformed.sample with the base noise resident in HBM (configs[1] names log_prob AND sampling): the stack
        run in the generative direction — couplings invert in closed form, so this is one pass of the same kernels."""
        sp = self._ospec()
        n = self.n
        gen = torch.Generator(device=dev).manual_seed(5)
        if str(sp.base).lower().startswith("n"):
            z = torch.randn(n, sp.event_dim, device=dev, generator=gen)
        else:
            z = torch.rand(n, sp.event_dim, device=dev, generator=gen)
        for _ in range(3):
            self.model.sample_from_noise(z)
        torch.cuda.synchronize()
        # per-call events, median: these are eager public-API calls (each allocates its output), and a single
        # allocator / host hiccup otherwise dominates a five-call mean (seen: 1.8 ms typical, 20 ms mean once)
        K = max(K, 9)
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for e0, e1 in evs:
            e0.record()
            x = self.model.sample_from_noise(z)
            e1.record()
        torch.cuda.synchronize()
        times = sorted(e0.elapsed_time(e1) for e0, e1 in evs)
        ms = times[len(times) // 2]
        return {"value": n / (ms * 1e-3), "unit": "samples/s", "samples_per_step_per_gpu": n, "ms_per_step": ms,
                "ms_min_max": [times[0], times[-1]], "calls": K, "finite": bool(torch.isfinite(x).all().item())}

    def cpu_run(self, x):
        from oracle import flows_oracle as O

        return O.coupling_flow_log_prob(self._ospec(), self._p, x)


class MafToy(CouplingWorkload):
    name = "maf_toy"
    desc = "configs[1]: 8x MaskedCoupling+RQS(12 bins, +-8), MLP 2->64->64->74, Normal base (src/maf_toy_densities.py); gaussian_blobs, 2^22 samples/GPU, log_prob"
    n = 1 << 22
    bytes_per_sample = 12
    flops_per_sample = 105_472  # needed-only columns
    bound = "tensor"


class RealNvpToy(CouplingWorkload):
    name = "realnvp_toy"
    desc = "configs[0]: 8x MaskedCoupling+ScalarAffine, MLP 2->64->64->2 (src/real_nvp_toy_densities.py); banana, 2^22 samples/GPU, log_prob"
    ospec_name = "REAL_NVP_TOY"
    data = staticmethod(banana)
    transform = "scalar_affine"
    n = 1 << 22
    bytes_per_sample = 12
    flops_per_sample = 69_632
    bound = "tensor"


class MafMnist(CouplingWorkload):
    name = "maf_mnist"
    desc = "configs[3]: 8x MaskedCoupling+RQS(8 bins,[0,1]), MLP 784->512x3->19600, Uniform base (src/maf_mnist.py); synthetic dequantised 28x28, 4096 samples/GPU, log_prob"
    ospec_name = "RQS_MNIST"
    data = staticmethod(mnist_like)
    n = 4096
    bytes_per_sample = 3140
    flops_per_sample = 95_100_000
    bound = "tensor"
    launches_per_step = 8 * 5 + 1
    tc_precision = 3          # FLOWCUDA_PRECISION_BF16X3
    # weight packs + first activation pack + per layer the hidden chain kernel and the transform GEMM (one wave of
    # clusters at this batch) + finish
    tc_launches = 4 + 1 + 8 * 2 + 1
    tc_dtype = "bf16x3 (fp32 accumulate)"
    cpu_sample = 256


class MafMnistLarge(MafMnist):
    """The same flow at the large synthetic batch SURVEY §8d lists (B = 32768): the per-layer chain of short kernels
    (three 512-wide hidden Linears) stops being latency-bound, so this is the kernels' throughput rather than the
    4096-sample step's."""
    name = "maf_mnist_b32768"
    desc = MafMnist.desc.replace("4096 samples/GPU", "32768 samples/GPU")
    n = 32768
    tc_launches = 4 + 1 + 8 * 4 + 1  # above one wave of clusters the hidden Linears run as three GEMM launches per layer


class Deep16Nvp(CouplingWorkload):
    """configs[4]: deep per-dimension affine coupling stack at D=784."""
    name = "deep16_nvp"
    desc = ("configs[4]: 16x MaskedCoupling + per-dimension affine (shift, log_scale), MLP 784->512->512->1568, Normal base "
            "(SURVEY C5); synthetic dequantised 28x28, 16384 samples/GPU, log_prob (+ sampling)")
    data = staticmethod(mnist_like)
    transform = "affine"
    n = 16384
    bytes_per_sample = 3140
    # needed-only: visible half of the input rows, transformed half of the output columns
    flops_per_sample = 16 * 2 * (392 * 512 + 512 * 512 + 512 * 784)
    bound = "tensor"
    launches_per_step = 16 * 4 + 1
    tc_precision = 3
    tc_launches = 3 + 1 + 16 * 3 + 1  # weight packs + first activation pack + per layer 3 GEMMs + finish
    tc_dtype = "bf16x3 (fp32 accumulate)"
    cpu_sample = 512

    def _ospec(self):
        from oracle import flows_oracle as O

        return O.CouplingSpec(784, 16, [512, 512], "affine")

    def sampling(self, dev, K=5):
        """Transformed.sample with the base noise resident: one forward pass of the stack (couplings invert in closed form)."""
        n = self.n
        z = torch.randn(n, 784, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
        for _ in range(2):
            self.model.sample_from_noise(z)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            x = self.model.sample_from_noise(z)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        return {"value": n / (ms * 1e-3), "unit": "samples/s", "samples_per_step_per_gpu": n, "ms_per_step": ms,
                "finite": bool(torch.isfinite(x).all().item())}


class MafMade(Workload):
    """Row #8: the MADE-conditioned MAF north_star names for the MNIST-shaped configs."""
    name = "maf_made"
    desc = ("configs[3]/north_star(b): 8x MADE-affine MAF, masked MLP 784->512->512->1568, reversal between layers, Normal base "
            "(SURVEY A.7; the reference has no MADE code); synthetic dequantised 28x28, 16384 samples/GPU, log_prob")
    n = 16384
    bytes_per_sample = 3140
    bound = "tensor"
    dtype = "bf16x3 (fp32 accumulate)"
    precision = "tc"
    layers = 8
    ceiling = 1.0 / 3.0
    ceiling_why = "error-compensated 3-pass bf16 products: three tensor-pipe FLOPs per algorithmic FLOP"
    launches_per_step = 3 + 1 + 8 * 3 + 1  # weight packs + first activation pack + per layer 3 GEMMs + finish
    cpu_sample = 1024

    def _ospec(self):
        from oracle import flows_oracle as O

        return O.MadeSpec(784, self.layers, [512, 512])

    @property
    def flops_per_sample(self):
        # needed-only: 2 FLOP per UNMASKED weight (SURVEY §8d counts 23.5 M for the dense shapes)
        sp = self._ospec()
        return int(2 * sp.num_layers * sum(float(m.sum()) for m in sp.masks()))

    def make_inputs(self, gen, n):
        return mnist_like(n, gen)

    def setup(self, dev):
        from normalizing_flows_b200 import flows
        from oracle import flows_oracle as O

        sp = self._ospec()
        self._p = O.init_made_params(sp, seed=1234, dtype=torch.float32, scale=0.1)
        self.model = flows.make_maf_model((784,), sp.num_layers, sp.hidden, params=O.flatten_layers(self._p).float().to(dev))

    def run(self, x):
        return self.model.log_prob(x)

    def cpu_run(self, x):
        from oracle import flows_oracle as O

        return O.maf_log_prob(self._ospec(), self._p, x)

    def train(self, dev, rank, world, dist, ops, K=5, batch=8192):
        """train samples/s: value_and_grad of -mean(log_prob) (hand-written FP32 reverse mode), gradient
        all-reduce across ranks (NCCL sum, the step's only exchange), fused Adam (maf_mnist.py:155-166)."""
        from normalizing_flows_b200 import _lib

        sp = self._ospec()
        spec = ops.make_made_spec(sp.event_dim, sp.num_layers, sp.hidden, precision=_lib.PRECISION_BF16X3)
        p = self.model.params.clone()
        m, v = torch.zeros_like(p), torch.zeros_like(p)
        gen = torch.Generator().manual_seed(77 + rank)
        xs = [self.make_inputs(gen, batch).to(dev) for _ in range(2)]
        need = _lib_ws(ops, spec, batch)
        ws = torch.empty((need + 3) // 4, device=dev, dtype=torch.float32)

        def step(i):
            loss, grad = ops.made_nll_grad(spec, p, xs[i % len(xs)], workspace=ws)
            if dist is not None:
                dist.all_reduce(grad)
            ops.adam_step(p, grad, m, v, i + 1, 1e-4, grad_scale=1.0 / world)
            return loss

        for i in range(2):
            step(i)
        ms_step, per, loss, _ = _train_windows(step, K, dist, dev)
        return {"value": batch * world / (ms_step * 1e-3), "unit": "samples/s", "batch_per_gpu": batch, "steps": K,
                "ms_per_step": ms_step, "windows_ms_per_step": per, "final_loss": float(loss.item()), "optimizer": "adam(1e-4)",
                "grad_allreduce_bytes": int(p.numel() * 4) if dist is not None else 0,
            "grad_allreduce": ("one NCCL sum all-reduce per flow layer (%d buckets), each followed by that layer's Adam update, overlapped with the reverse mode" % sp.num_layers)
            if dist is not None else None,
                "precision": "bf16x3 GEMMs (fp32 accumulate, fp32 stored activations)"}

    def sampling(self, dev, K=3):
        """Sequential sampling (persistent kernel over the D positions), base noise resident in HBM."""
        n = 148 * 8 * 4
        z = torch.randn(n, 784, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
        for _ in range(2):
            self.model.sample_from_noise(z)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            x = self.model.sample_from_noise(z)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        return {"value": n / (ms * 1e-3), "unit": "samples/s", "samples_per_step_per_gpu": n, "ms_per_step": ms,
                "finite": bool(torch.isfinite(x).all().item())}


class Deep16Made(MafMade):
    name = "deep16_made"
    desc = ("configs[4]: 16x MADE-affine MAF at D=784 (SURVEY C5), masked MLP 784->512->512->1568; synthetic dequantised 28x28, "
            "16384 samples/GPU, log_prob (+ sequential sampling)")
    layers = 16
    launches_per_step = 3 + 1 + 16 * 3 + 1
    cpu_sample = 512


class RealNvpHand(Workload):
    name = "realnvp_hand"
    desc = "hand-rolled Real NVP chain, 8 layers [1,4,2] leaky_relu (src/real_nvp.py); banana, 2^22 samples/GPU, log_prob"
    n = 1 << 22
    bytes_per_sample = 12
    flops_per_sample = 192
    bound = "hbm"
    cpu_sample = 1 << 20
    # 161 thread-instructions per 12-byte sample (8 layers x (4 hidden units x 5 + exp + affine)): at 100 % issue
    # (148 SMs x 4 schedulers x 32 lanes x 1.965 GHz) that is 2.31e11 samples/s = 2.77 TB/s = 0.42 of the HBM roof
    ceiling = 0.42
    ceiling_why = "FP32 issue floor: 161 thread-instructions per sample at L=8 (18 us per 2^22) vs the 7.7 us HBM floor"

    def make_inputs(self, gen, n):
        return banana(n, gen)

    def setup(self, dev):
        from normalizing_flows_b200 import real_nvp
        from oracle import flows_oracle as O

        self._p, self._flips = O.init_nvp_chain_like(8, hidden=4, seed=534, dtype=torch.float32, scale=0.1)
        self.flat = real_nvp.pack_params(O.flatten_layers(self._p))
        self.cfg = [(real_nvp.shift_and_log_scale_fn, f) for f in self._flips]
        self.mod = real_nvp

    def run(self, x):
        return self.mod.log_prob_nvp_chains(self.flat, self.cfg, self.mod.log_prob_N01, x, hidden=4)

    def cpu_run(self, x):
        from oracle import flows_oracle as O

        return O.log_prob_nvp_chains(self._p, self._flips, x)


class Bernstein(Workload):
    name = "bernstein"
    desc = "configs[2]: BernsteinBijector forward_and_log_det, order 32 shared theta (src/BernsteinBijector.py); U(0,1), 2^24 elements/GPU"
    n = 1 << 24
    order = 32
    bytes_per_sample = 12
    flops_per_sample = 2 * 32 + 10
    bound = "hbm"
    cpu_sample = 1 << 20
    # order 32: two polynomials x 32 coefficients at 1.25 packed operations per coefficient + setup = ~83 FP32 operations per
    # 12-byte element; ncu (profiles/r01/final/ncu_bern_summary.txt) puts the sustained packed-FMA rate of this mix at
    # ~76 operations/cycle/SM => 65 us per 2^24 elements against the 31 us HBM floor
    ceiling = 0.47
    ceiling_why = ("FMA-bound, not HBM-bound, at order 32: ~83 FP32 operations per 12-byte element at the measured ~76 "
                   "operations/cycle/SM is 65 us per 2^24 elements against the 31 us HBM floor; the reference's own order "
                   "(M = 9, bernstein_m9) IS HBM-bound and runs at 0.68 of the roof")

    def make_inputs(self, gen, n):
        return torch.rand(n, generator=gen)

    def setup(self, dev):
        from normalizing_flows_b200 import BernsteinBijector as B

        g = torch.Generator().manual_seed(7)
        self.raw = torch.randn(self.order, generator=g)
        self.bij = B.BernsteinBijector(self.raw.to(dev))

    def run(self, x):
        y, ld = self.bij.forward_and_log_det(x)
        return ld

    def cpu_run(self, x):
        from oracle import flows_oracle as O

        return O.bernstein_forward_and_log_det(O.constrain_thetas(self.raw), x)[1]


class Bernstein9(Bernstein):
    name = "bernstein_m9"
    desc = "configs[2] at the reference's own order: BernsteinBijector forward_and_log_det, M=9 shared theta (src/bflow.py:24); U(0,1), 2^24 elements/GPU"
    order = 9
    flops_per_sample = 2 * 9 + 10
    ceiling = 1.0
    ceiling_why = ""


WORKLOADS = {w.name: w for w in (MafMnist, MafMnistLarge, RealNvpToy, MafToy, MafMade, Deep16Nvp, Deep16Made, RealNvpHand, Bernstein, Bernstein9)}
# timed into `others` by default: at N=1 everything, at N>1 the configurations BASELINE.json sweeps over GPUs
OTHERS_N1 = ("realnvp_toy", "maf_toy", "maf_mnist_b32768", "maf_made", "deep16_nvp", "deep16_made", "realnvp_hand", "bernstein_m9", "bernstein")
OTHERS_NN = ("realnvp_toy", "maf_toy", "maf_made", "deep16_nvp", "deep16_made")


def _train_windows(step, K, dist, dev, first=2, max_windows=8):
    """K training steps timed on the device, window after window (max over ranks per window), until the last three
    windows agree within 5 % or `max_windows` have run; the median of the last three is the figure and every window is
    reported.  A 5-step window is 30 ms, and at N = 2 the first two windows were seen to run 30-40 % slower than the
    steady state on some boxes and not on others (8.9, 9.5, 6.5 ms on one; 6.54, 6.50, 6.46 on the next, same tree; torch's
    allocator still calls cudaMalloc in the first windows — `windows_device_allocs`).  Returns (ms per step, all windows, last loss)."""
    per, loss, i = [], None, first
    _train_windows.device_allocs = []  # cudaMalloc calls of torch's allocator per window (diagnostic of a slow start)
    while len(per) < max_windows:
        a0 = torch.cuda.memory_stats(dev).get("num_device_alloc", 0)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _k in range(K):
            loss = step(i)
            i += 1
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()  # the same number on every rank, so every rank stops at the same window
        per.append(ms / K)
        _train_windows.device_allocs.append(torch.cuda.memory_stats(dev).get("num_device_alloc", 0) - a0)
        if len(per) >= 3 and max(per[-3:]) <= 1.05 * min(per[-3:]):
            break
    return statistics.median(per[-3:]), per, loss, i


def _lib_ws(ops, spec, n):
    import ctypes as C

    from normalizing_flows_b200 import _lib

    return _lib.load().flowcuda_made_nll_grad_workspace_bytes(C.byref(spec), n)


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for t, line in self.rows:
            f = [s.strip() for s in line.split(",")]
            if len(f) < 7:
                continue
            try:
                if t0 - 0.05 <= t <= t1 + 0.15:
                    sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            if t0 - 0.05 <= t <= t1 + 0.15:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        if not sm:  # timed region shorter than one sample: take everything we saw
            for t, line in self.rows:
                try:
                    sm.append(float(line.split(",")[0]))
                except ValueError:
                    pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- reference arm
def workload_config(w, world):
    """The `config` object: identical for both arms (the workload definition, nothing implementation-specific)."""
    return {"workload": w.name, "desc": w.desc, "samples_per_step_per_gpu": w.n, "parallelism": f"dp{world}"}


def run_reference(args, wl):
    """The reference's CPU implementation of the path.  JAX/distrax are not installable on
    either box, so this is the float32 CPU restatement (oracle/), all host
    threads, each step a bounded sample of the same workload; per-step times, median reported."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    gen = torch.Generator().manual_seed(1)
    w = wl()
    n = min(w.n, w.cpu_sample)
    x = w.make_inputs(gen, n)
    _cpu_setup(w)
    steps = max(5, args.steps)
    with torch.no_grad():
        for _ in range(max(1, min(args.warmup, 3))):
            w.cpu_run(x)
        times = []
        budget = time.perf_counter() + 150.0  # the whole arm ends within a few minutes whatever K is passed
        for _ in range(steps):
            t0 = time.perf_counter()
            w.cpu_run(x)
            times.append(time.perf_counter() - t0)
            if time.perf_counter() > budget and len(times) >= 5:
                break
    med = statistics.median(times)
    val = n / med
    line = {
        "impl": "reference", "metric": "log_prob samples/s", "value": val, "unit": "samples/s", "n_gpus": args.gpus,
        "steps": len(times), "warmup": args.warmup, "ms_per_step": 1e3 * med, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(w, args.gpus),
        "step_s": {"median": med, "min": min(times), "max": max(times)},
        "cpu_baseline": {"value": val, "unit": "samples/s", "cores": cores, "kind": "port",
                         "sample": f"{n} samples/step of the same workload, median of {len(times)} steps; float32 torch-CPU "
                                   f"restatement of the reference with torch.set_num_threads({cores}) (JAX unavailable); rank 0 only"},
        "e2e": {"value": val, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def _cpu_setup(w):
    """CPU-side parameters only (no CUDA import path)."""
    from oracle import flows_oracle as O

    if isinstance(w, CouplingWorkload):
        w._p = w._params()
    elif isinstance(w, MafMade):
        w._p = O.init_made_params(w._ospec(), seed=1234, dtype=torch.float32, scale=0.1)
    elif isinstance(w, RealNvpHand):
        w._p, w._flips = O.init_nvp_chain_like(8, hidden=4, seed=534, dtype=torch.float32, scale=0.1)
    elif isinstance(w, Bernstein):
        w.raw = torch.randn(w.order, generator=torch.Generator().manual_seed(7))


def cpu_baseline(w, seconds_target=12.0):
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    gen = torch.Generator().manual_seed(1)
    n = min(w.n, w.cpu_sample)
    x = w.make_inputs(gen, n)
    times = []
    with torch.no_grad():
        w.cpu_run(x)
        t_end = time.perf_counter() + seconds_target
        while len(times) < 5 or (time.perf_counter() < t_end and len(times) < 50):
            t0 = time.perf_counter()
            w.cpu_run(x)
            times.append(time.perf_counter() - t0)
    med = statistics.median(times)
    return {"value": n / med, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": f"median of {len(times)} x {n} samples of the same workload (float32 torch-CPU restatement of the "
                      f"reference, {cores} threads; JAX unavailable)",
            "step_s": {"median": med, "min": min(times), "max": max(times)}}


def parity_check(w, x_dev, lp_dev, n_check=256):
    """Before anything is timed: the first `n_check` samples of the first batch through the float64 oracle on the
    host, compared per sample with what the timed call returns (tolerances: north_star)."""
    n = min(n_check, x_dev.shape[0])
    x = x_dev[:n].detach().cpu()
    got = lp_dev.reshape(-1)[:n].detach().double().cpu()
    save = getattr(w, "_p", None)
    try:
        conv = lambda t: t.double() if torch.is_tensor(t) else t
        if isinstance(save, (list, tuple)):
            w._p = [[tuple(conv(t) for t in lin) if isinstance(lin, tuple) else [conv(t) for t in lin] for lin in layer] for layer in save]
        ref = w.cpu_run(x.double()).double()
    finally:
        if save is not None:
            w._p = save
    fin = torch.isfinite(ref)
    err = (got[fin] - ref[fin]).abs()
    absolute = w.bytes_per_sample == 3140
    if not absolute:
        err = err / ref[fin].abs().clamp(min=1.0)
    tol = 1e-3 if absolute else 1e-5
    out = {"checked_samples": int(n), "max_err": float(err.max()) if err.numel() else 0.0,
           "p99.9_err": float(torch.quantile(err, 0.999)) if err.numel() else 0.0,
           "tolerance": tol, "form": "absolute |d log_prob|" if absolute else "|d log_prob| / max(1, |ref|)",
           "against": "float64 oracle (oracle/flows_oracle.py) on the host", "ok": bool(err.numel() == 0 or err.max() <= tol)}
    if not out["ok"]:
        raise SystemExit(f"bench: parity check failed before timing: {out}")
    return out


# ----------------------------------------------------------------------------- main arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="flowcuda", choices=["flowcuda", "reference"])
    ap.add_argument("--workload", default="maf_mnist", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--precision", default=None, choices=["tc", "fp32"],
                    help="coupling workloads: tensor-core path (default where built) or the FP32 CUDA-core path")
    ap.add_argument("--no-others", action="store_true", help="skip the other workloads (default: timed into 'others')")
    ap.add_argument("--others", action="store_true", help="(default behaviour; kept for compatibility)")
    ap.add_argument("--only-kernel", action="store_true", help="profiling runs: hot-path steps only (no e2e/train/others/CPU legs)")
    args = ap.parse_args()
    wl = WORKLOADS[args.workload]
    if args.precision is not None and issubclass(wl, CouplingWorkload):
        wl.precision = args.precision
    if args.impl == "reference":
        run_reference(args, wl)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the flow hot path has no CPU fallback")
    affinity = bind_rank_to_cores(local, world)  # before any pinned allocation (first-touch on this rank's cores)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL writes its debug lines ("NCCL version ...") to stdout by default; stdout carries the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist_mod.init_process_group("nccl", device_id=dev)
        dist = dist_mod
    from normalizing_flows_b200 import ops

    peaks = load_peaks()
    W = max(3, args.warmup)
    K = args.steps
    full = not args.only_kernel
    t_start = time.time()
    res = time_workload(wl(), dev, rank, world, dist, W, K, ops, e2e=full, clocks=True, sustained=full, parity=full)
    others = {}
    if full and not args.no_others:
        budget_s = 150.0  # the whole default run stays within a few minutes
        for name in (OTHERS_N1 if world == 1 else OTHERS_NN):
            if name == args.workload:
                continue
            if time.time() - t_start > budget_s:
                others[name] = {"skipped": "time budget of the default run"}
                continue
            cls = WORKLOADS[name]
            r = time_workload(cls(), dev, rank, world, dist, W, max(5, K // 2), ops, e2e=False, clocks=False,
                              extras=("sampling",) if name.startswith("deep16") or world > 1 else
                              (("train", "sampling") if name in ("realnvp_toy", "maf_toy") else ()),
                              parity=(world == 1))
            ow = r["w"]
            others[name] = {"value": r["value"], "unit": "samples/s", "ms_per_step": r["ms_per_step"], "kernel_ms": r["kernel_ms"],
                            "samples_per_step_per_gpu": ow.n, "dtype": ow.dtype, "roofline": roofline(ow, r, peaks)}
            for k in ("train", "sampling", "parity"):
                if r.get(k):
                    others[name][k] = r[k]
            del r, ow
            torch.cuda.empty_cache()
    if rank == 0:
        w = res["w"]
        line = {
            "metric": "log_prob samples/s", "value": res["value"], "unit": "samples/s", "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": w.dtype, "data": "synthetic",
            "config": workload_config(w, world),
            "impl_detail": {"precision": getattr(w, "precision", "fp32"), "launch": getattr(w, "launch_mode", "eager"),
                            "l2": f"inputs rotate over {res['nbuf']} distinct batches ({res['nbuf'] * w.n * w.bytes_per_sample / 1e6:.0f} MB > 126 MB L2)",
                            "cpu_affinity": affinity},
            "roofline": roofline(w, res, peaks),
            "gpu_launches": res["launches"],
            "clocks": res["clocks"],
        }
        for k in ("e2e", "e2e_uint8", "prepared", "train", "sampling", "sustained", "parity"):
            if res.get(k):
                line[k] = res[k]
        if "e2e" not in line:
            line["e2e"] = None
        if world == 1 and full and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(w)
        if others:
            line["others"] = others
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def bind_rank_to_cores(local, world):
    """N ranks on one host: give each rank its own slice of the cores this process may run on, BEFORE it allocates
    pinned staging memory, so that eight ranks do not share one set of cores (e2e scaling)."""
    if world <= 1 or not hasattr(os, "sched_getaffinity"):
        return None
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = len(cores) // world
        if per < 1:
            return None
        mine = cores[local * per:(local + 1) * per]
        os.sched_setaffinity(0, mine)
        torch.set_num_threads(max(1, min(per, 4)))
        return f"cores {mine[0]}-{mine[-1]} ({per} of {len(cores)})"
    except OSError:
        return None


def roofline(w, res, peaks):
    """achieved = ALGORITHMIC bytes (or needed-only FLOPs) of one step / the mean step duration measured with CUDA
    events in this run; peak = MEASURED_PEAKS.json.  `ceiling` = the fraction the algorithm itself admits on this
    path (3-pass products, instruction floors).  `traffic` is only filled from an ncu capture of THIS round's
    kernel (profiles/traffic.json entry tagged with the round) — it is not measured inside this run."""
    dur = res["kernel_ms"] * 1e-3
    src = peaks["_source"]
    if w.bound == "hbm":
        ach = w.n * w.bytes_per_sample / dur / 1e9
        peak = peaks["hbm_gbs"]
        unit = "GB/s"
    else:
        ach = w.n * w.flops_per_sample / dur / 1e12
        peak = peaks["bf16_tflops"]
        unit = "TFLOP/s"
    traffic, tsrc = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            ent = json.load(f).get(w.name, {})
        if ent.get("round") == "r02":
            traffic, tsrc = ent.get("dram_bytes_per_step"), ent.get("source")
    except Exception:
        pass
    out = {"bound": w.bound, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak, "traffic": traffic,
           "ceiling": w.ceiling, "frac_of_ceiling": ach / peak / w.ceiling,
           "peak_source": f"MEASURED_PEAKS.json ({src})" if src == "measured" else "B200_PROFILING.md fallback",
           "kernel_ms": res["kernel_ms"],
           "algorithmic": (f"{w.n} x {w.bytes_per_sample} B" if w.bound == "hbm" else f"{w.n} x {w.flops_per_sample} FLOP (needed-only)")}
    if w.ceiling_why:
        out["ceiling_why"] = w.ceiling_why
    if tsrc:
        out["traffic_source"] = tsrc
    if w.bound != "hbm":
        out["hbm_frac"] = w.n * w.bytes_per_sample / dur / 1e9 / peaks["hbm_gbs"]
    return out


def time_workload(w, dev, rank, world, dist, W, K, ops, e2e, clocks, sustained=False, parity=False, extras=()):
    gen = torch.Generator().manual_seed(1000 + rank)
    w.setup(dev)
    # rotate over enough distinct batches that consecutive steps never hit L2 (126 MB)
    per = w.n * w.bytes_per_sample
    nbuf = max(2, min(16, int(math.ceil(300e6 / per))))
    host = [w.make_inputs(gen, w.n).pin_memory() for _ in range(min(nbuf, 4))]
    bufs = [host[i % len(host)].to(dev) if i < len(host) else None for i in range(nbuf)]
    for i in range(len(host), nbuf):  # distinct device data without holding more pinned memory
        bufs[i] = bufs[i % len(host)].roll(i, 0).contiguous()
    stream = torch.cuda.current_stream()
    # The hot-path call of each distinct batch is captured ONCE in a CUDA graph and replayed in the timed
    # loop: several of these kernels run for 20-100 us, less than the Python/ctypes cost of issuing them,
    # so eager launches would time the host.  (The e2e leg below stays eager through the public API.)
    graphs = None
    if os.environ.get("FLOWCUDA_BENCH_GRAPH", "1") != "0":
        try:
            w.run(bufs[0])
            torch.cuda.synchronize()
            graphs, pool = [], None
            for i in range(nbuf):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, pool=pool):
                    out = w.run(bufs[i])
                pool = g.pool()
                graphs.append((g, out))
            torch.cuda.synchronize()
        except Exception as e:  # capture not possible: fall back to eager launches and say so
            print(f"bench: CUDA-graph capture failed ({e!r}); timing eager launches", file=sys.stderr)
            graphs = None
    w.launch_mode = "cuda-graph replay" if graphs else "eager"
    par = None
    if parity and rank == 0:
        if graphs:
            graphs[0][0].replay()  # capture does not execute: produce batch 0's output first
            torch.cuda.synchronize()
        par = parity_check(w, bufs[0], graphs[0][1] if graphs else w.run(bufs[0]))
    # The scalar mean-log-likelihood exchange (north_star (c)) runs on a side stream: the reduction kernels and the
    # 4-byte NCCL all-reduce of step i overlap step i+1's kernels instead of sitting on the compute stream
    # (+41 us per step at N=8 when serialised).
    side = torch.cuda.Stream(dev) if dist is not None else None
    side_done = []

    def step(i):
        if graphs:
            g, lp = graphs[i % nbuf]
            g.replay()
        else:
            lp = w.run(bufs[i % nbuf])
        if dist is not None:
            ev = torch.cuda.Event()
            ev.record(stream)
            with torch.cuda.stream(side):
                side.wait_event(ev)
                m = ops.nll(lp)
                dist.all_reduce(m)
            side_done.append(m)
            if len(side_done) > 4:
                side_done.pop(0)
        return lp

    def join_side():
        if side is not None:
            stream.wait_stream(side)

    for i in range(W):
        step(i)
    torch.cuda.synchronize()
    # rank 0 starts its clock sampler BEFORE the barrier, so no rank enters the timed region early
    # and then waits in the first all-reduce for rank 0 to wake up
    sampler = ClockSampler(dev.index) if clocks and rank == 0 else None
    if sampler:
        sampler.start()
        t_wait = time.time()
        while not sampler.rows and time.time() - t_wait < 5.0:  # nvidia-smi takes a while to print its first row
            time.sleep(0.05)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    t0 = time.time()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record(stream)
    for i in range(K):
        evs[i][0].record(stream)
        step(W + i)
        evs[i][1].record(stream)
    join_side()  # the last steps' scalar all-reduces are inside the timed region
    end.record(stream)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t1 = time.time()
    clk = sampler.stop(t0, t1) if sampler else None
    total_ms = start.elapsed_time(end)
    if dist is not None:
        t = torch.tensor([total_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = t.item()
    kernel_ms = statistics.mean(a.elapsed_time(b) for a, b in evs)
    out = {"w": w, "value": w.n * world * K / (total_ms * 1e-3), "ms_per_step": total_ms / K, "kernel_ms": kernel_ms,
           "nbuf": nbuf, "clocks": clk, "launches": K * (w.launches_per_step + (2 if dist is not None else 0)), "parity": par}
    if sustained:
        out["sustained"] = time_sustained(w, step, join_side, stream, dev, rank, world, dist)
    for k in extras:
        if k == "train":
            out["train"] = w.train(dev, rank, world, dist, ops) if hasattr(w, "train") else time_train(w, dev, rank, world, dist, ops)
        elif k == "sampling" and hasattr(w, "sampling"):
            out["sampling"] = w.sampling(dev)
    if e2e:
        out["e2e"] = time_e2e(w, dev, host, dist, world, max(3, K // 2))
        if getattr(w, "bytes_per_sample", 0) == 3140 and hasattr(w.model, "prepare"):
            out["prepared"] = time_prepared(w, bufs, K)
        if getattr(w, "bytes_per_sample", 0) == 3140:  # the D=784 image workloads
            out["e2e_uint8"] = time_e2e_u8(w, dev, rank, dist, world, max(3, K // 2), ops)
        out["train"] = w.train(dev, rank, world, dist, ops) if hasattr(w, "train") else time_train(w, dev, rank, world, dist, ops)
        if hasattr(w, "sampling"):
            out["sampling"] = w.sampling(dev)
    return out


def time_sustained(w, step, join_side, stream, dev, rank, world, dist, seconds=2.5):
    """The same step replayed back to back for >= 2 s with the clocks sampled throughout: the 20-step headline is a
    burst figure (tens of ms); this one shows what the kernels hold once power and clocks have settled."""
    steps = max(50, int(seconds * 1e3 / max(w_ms(w, step, stream), 1e-3)))
    if dist is not None:
        t = torch.tensor([steps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        steps = int(t.item())
    sampler = ClockSampler(dev.index) if rank == 0 else None
    if sampler:
        sampler.start()
        t_wait = time.time()
        while not sampler.rows and time.time() - t_wait < 5.0:
            time.sleep(0.05)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record(stream)
    for i in range(steps):
        step(i)
    join_side()
    e1.record(stream)
    torch.cuda.synchronize()
    t1 = time.time()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    clk = sampler.stop(t0, t1) if sampler else None
    return {"value": w.n * world * steps / (ms * 1e-3), "unit": "samples/s", "steps": steps, "seconds": ms * 1e-3,
            "ms_per_step": ms / steps, "clocks": clk}


def w_ms(w, step, stream):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(5):
        step(i)
    e1.record(stream)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5


def time_train(w, dev, rank, world, dist, ops, K=5, batch=1 << 18):
    """train samples/s: one step = value_and_grad of -mean(log_prob) over a batch (hand-written
    reverse mode), the gradient all-reduce across ranks (NCCL sum, only exchange of the step) and the
    fused Adam update (optax.adam, real_nvp_toy_densities.py:162-173).  Batch per GPU is fixed (weak)."""
    if not isinstance(w, CouplingWorkload):
        return None
    from normalizing_flows_b200 import _lib

    sp = w._ospec()
    if sp.event_dim != 2:
        batch = 4096  # stored activations: 88 KB per sample and layer at the maf_mnist shape (2.9 GB)
    tr = {"affine_scalar": 0, "affine": 1, "rqs": 2}[sp.transform]
    tc = sp.event_dim != 2  # wide conditioners: reverse-mode GEMMs on tcgen05 (mgtc.cuh); D=2: fused FP32 kernels
    spec = ops.make_spec(sp.event_dim, sp.num_layers, sp.hidden, tr, sp.num_bins, sp.range_min, sp.range_max,
                         _lib.BASE_UNIFORM01 if sp.base == "uniform" else _lib.BASE_NORMAL,
                         _lib.PRECISION_BF16X3 if tc else 0)
    p = w.model.params.clone()
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    gen = torch.Generator().manual_seed(77 + rank)
    xs = [w.make_inputs(gen, batch).to(dev) for _ in range(4)]

    from normalizing_flows_b200.data_parallel import LayerwiseAllReduce

    def step(i):
        # the reverse mode reports each flow layer as it completes (flowcuda_coupling_nll_grad_layers): that layer's sum
        # all-reduce (N > 1) and its Adam update are enqueued on a second stream at once and run under the
        # differentiation of the remaining layers; the step ends when the last layer's update has run
        t = i + 1

        def adam_layer(layer, lo, hi, g):
            ops.adam_step(p[lo:hi], g, m[lo:hi], v[lo:hi], t, 1e-4, grad_scale=1.0 / world)

        if dist is None:  # one GPU: nothing to exchange, and a concurrent Adam only slows the chain it runs beside (measured)
            loss, grad = ops.coupling_nll_grad(spec, p, xs[i % len(xs)])
            ops.adam_step(p, grad, m, v, t, 1e-4, grad_scale=1.0)
            return loss
        upd = LayerwiseAllReduce(after=adam_layer)
        loss, grad = ops.coupling_nll_grad(spec, p, xs[i % len(xs)], on_layer=upd)
        upd.wait()
        return loss

    for i in range(2):
        step(i)
    ms_step, per, loss, done = _train_windows(step, K, dist, dev)  # done = Adam steps taken so far
    final_loss = float(loss.item())
    allocs = list(_train_windows.device_allocs)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # one GPU, wide flows: the same step with value_and_grad replayed from a CUDA graph (one per distinct batch; the C
    # call is capturable, tests/test_gpu_grad.py) and Adam launched after it — reported beside the eager figure, which
    # stays `value` because the multi-GPU step (host callbacks per layer) cannot be replayed
    replay_ms = None
    if dist is None and tc:
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            caps = []
            for x in xs:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=side):
                    lg = ops.coupling_nll_grad(spec, p, x)
                caps.append((g, lg))
            torch.cuda.current_stream().wait_stream(side)

            def rstep(i):
                g, (_, grad) = caps[i % len(caps)]
                g.replay()
                ops.adam_step(p, grad, m, v, i + 1, 1e-4, grad_scale=1.0)

            for i in range(2):
                rstep(done + i)
            torch.cuda.synchronize()
            e0.record()
            for i in range(K):
                rstep(done + 2 + i)
            e1.record()
            torch.cuda.synchronize()
            replay_ms = e0.elapsed_time(e1) / K
        except Exception as e:  # noqa: BLE001 - the eager figure stands on its own
            replay_ms = None
            print("train graph replay not measured: %r" % (e,), file=sys.stderr)
    return {"value": batch * world / (ms_step * 1e-3), "unit": "samples/s", "batch_per_gpu": batch, "steps": K,
            "ms_per_step": ms_step, "windows_ms_per_step": per, "windows_device_allocs": allocs, "graph_replay_ms_per_step": replay_ms, "final_loss": final_loss, "optimizer": "adam(1e-4)",
            "grad_allreduce_bytes": int(p.numel() * 4) if dist is not None else 0,
            "grad_allreduce": ("one NCCL sum all-reduce per flow layer (%d buckets), each followed by that layer's Adam update, overlapped with the reverse mode" % sp.num_layers)
            if dist is not None else None,
            "precision": "bf16x3 GEMMs (fp32 accumulate, fp32 stored activations)" if tc else "fp32"}


def time_prepared(w, bufs, K):
    """Serving variant of the same step: the weights are packed into tile images ONCE (Transformed.prepare —
    parameters do not change between inference calls) instead of on every call; inputs resident, graph replay."""
    w.model.prepare(w.n)
    side = torch.cuda.Stream()
    try:
        outs = []
        for b in bufs[:2]:
            cap = _capture(lambda b=b: w.run(b), side)
            if cap is None:
                return None
            outs.append(cap)
        for g, _ in outs:
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(K):
            outs[i % 2][0].replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / K
        return {"value": w.n / (ms * 1e-3), "unit": "samples/s per GPU", "ms_per_step": ms,
                "note": "weights pre-packed once (Transformed.prepare); the headline value re-packs them every call"}
    finally:
        w.model.invalidate()


def _capture(fn, stream):
    """Capture `fn()` (the public-API call on fixed device buffers) into a CUDA graph replayed on `stream`;
    None when FLOWCUDA_BENCH_GRAPH=0 or the capture fails (then the call is issued eagerly)."""
    if os.environ.get("FLOWCUDA_BENCH_GRAPH", "1") == "0":
        return None
    try:
        with torch.cuda.stream(stream):
            fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            out = fn()
        torch.cuda.synchronize()
        return g, out
    except Exception as e:
        print(f"bench: e2e CUDA-graph capture failed ({e!r}); eager launches", file=sys.stderr)
        return None


def _e2e_pipeline(w, dev, host, prep, dist, world, K, h2d_bytes, note=None):
    """One API call per step (the D=784 workloads: ~45 short kernels), depth-2 pipeline: two device staging
    buffers, each with its API call captured in a CUDA graph; step i+1's host->device copy runs under step i's
    kernels and step i-1's read-back.  A slot is reissued only after its previous results reached the host."""
    s_in, s_cmp, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    slots = []
    for _ in range(2):
        d = torch.empty_like(host[0], device=dev)
        call = prep(d)
        slots.append({"d": d, "call": call, "cap": _capture(call, s_cmp), "out": torch.empty(w.n, dtype=torch.float32).pin_memory(),
                      "done": None})

    def issue(i):
        sl = slots[i % 2]
        if sl["done"] is not None:
            sl["done"].synchronize()
        with torch.cuda.stream(s_in):
            sl["d"].copy_(host[i % len(host)], non_blocking=True)
            e_in = torch.cuda.Event()
            e_in.record(s_in)
        with torch.cuda.stream(s_cmp):
            s_cmp.wait_event(e_in)
            if sl["cap"]:
                sl["cap"][0].replay()
                lp = sl["cap"][1]
            else:
                lp = sl["call"]()
            e_c = torch.cuda.Event()
            e_c.record(s_cmp)
        with torch.cuda.stream(s_out):
            s_out.wait_event(e_c)
            sl["out"].copy_(lp.reshape(-1), non_blocking=True)
            sl["done"] = torch.cuda.Event()
            sl["done"].record(s_out)
        sl["keep"] = lp

    for i in range(4):
        issue(i)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(K):
        issue(i)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = t.item()
    out = {"value": w.n * world * K / dt, "unit": "samples/s", "h2d_bytes_per_step": int(h2d_bytes),
           "d2h_bytes_per_step": int(w.n * 4), "steps": K, "chunks": 1, "pipeline_depth": 2,
           "finite": bool(torch.isfinite(slots[0]["out"]).all().item()),
           "launch": "cuda-graph replay of the API call" if slots[0]["cap"] else "eager"}
    if note:
        out["input"] = note
    return out


def time_e2e(w, dev, host, dist, world, K):
    """Same metric through the public API with HOST buffers: every step copies its inputs from
    pinned host memory, runs log_prob, and reads the per-sample log_prob back to pinned host
    memory.  Three streams (H2D, compute, D2H); the long one-kernel workloads are split into 8 chunks so
    that the copies hide under the kernels, and two sets of staging buffers alternate between steps so that
    a step's first copy does not wait for the previous step's last read-back."""
    chunks = 8 if w.n >= (1 << 16) else 1
    if chunks == 1:
        return _e2e_pipeline(w, dev, host, lambda d: (lambda: w.run(d)), dist, world, max(K, 6), host[0].numel() * 4)
    cs = w.n // chunks
    s_in, s_cmp, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    slots = [{"d": [torch.empty_like(host[0][:cs], device=dev) for _ in range(chunks)],
              "out": torch.empty(w.n, dtype=torch.float32).pin_memory(), "done": None, "keep": None} for _ in range(2)]

    def step(i):
        sl = slots[i % 2]
        if sl["done"] is not None:
            sl["done"].synchronize()  # this slot's previous results are on the host: its buffers are free
        src = host[i % len(host)]
        lps = []
        for c in range(chunks):
            with torch.cuda.stream(s_in):
                sl["d"][c].copy_(src[c * cs:(c + 1) * cs], non_blocking=True)
                e_in = torch.cuda.Event()
                e_in.record(s_in)
            with torch.cuda.stream(s_cmp):
                s_cmp.wait_event(e_in)
                lp = w.run(sl["d"][c])
                e_c = torch.cuda.Event()
                e_c.record(s_cmp)
            with torch.cuda.stream(s_out):
                s_out.wait_event(e_c)
                sl["out"][c * cs:(c + 1) * cs].copy_(lp.reshape(-1), non_blocking=True)
            lps.append(lp)
        sl["done"] = torch.cuda.Event()
        sl["done"].record(s_out)
        sl["keep"] = lps

    for i in range(4):
        step(i)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    for i in range(K):
        step(i)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([dt], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = t.item()
    return {"value": w.n * world * K / dt, "unit": "samples/s", "h2d_bytes_per_step": int(host[0].numel() * 4),
            "d2h_bytes_per_step": int(w.n * 4), "steps": K, "chunks": chunks, "pipeline_depth": 2, "launch": "eager"}


def time_e2e_u8(w, dev, rank, dist, world, K, ops):
    """The D=784 pipelines as the reference feeds them (src/maf_mnist.py:101-118): the HOST holds uint8 images;
    every step copies the bytes (a quarter of the float32 image) from pinned memory, dequantises on the device
    (`(u8 + U[0,1)) / 256`, Philox, csrc/densities.cu), runs log_prob and reads the result back."""
    gen = torch.Generator().manual_seed(4242 + rank)
    host = [torch.randint(0, 256, (w.n, 784), generator=gen, dtype=torch.uint8).pin_memory() for _ in range(3)]

    def prep(d):
        f = torch.empty(w.n, 784, device=dev, dtype=torch.float32)

        def call():
            ops.dequantize_u8(d, 7, offset=0, out=f)
            return w.run(f)

        return call

    return _e2e_pipeline(w, dev, host, prep, dist, world, max(K, 6), w.n * 784,
                         note="uint8 images in pinned host memory, dequantised on the device")


if __name__ == "__main__":
    # stdout carries exactly one JSON line.  Libraries write to file descriptor 1 behind Python's back (NCCL prints
    # "NCCL version ..." there when NCCL_DEBUG=VERSION and ignores NCCL_DEBUG_FILE at that level), so descriptor 1
    # is pointed at stderr for the whole run and Python's sys.stdout keeps the real one.
    sys.stdout.flush()
    _real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _real_stdout
    main()
    _real_stdout.flush()

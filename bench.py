#!/usr/bin/env python
"""Benchmark of the rollout hot path (BASELINE.json metric: rollout samples*steps/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One bench "step" = one full autoregressive rollout (all lead times) of this rank's shard of
ensemble members, inputs resident in HBM, followed by the on-device RMSE/ACC partial sums and
(N > 1) the single all-reduce of those partials.  ``value`` = members * lead_steps * K / time
summed over all ranks.  ``--scaling strong`` (the default of workload c4 = BASELINE configs[3]: "256-member
ensemble sharded across 1/2/4/8 B200"): the ensemble has ``--members`` members in TOTAL and rank r rolls out
``ensemble.shard_range(members, r, N)``; ``--scaling weak``: every GPU rolls out ``--members`` members.
``e2e`` is the same rollout through ``dlwp_rollout_host`` with pinned HOST buffers: H2D of the inputs
and D2H of the whole forecast inside the timed region.

``--impl reference`` times the reference's CPU path for the same model (the oracle: the
reference's own AFNO source restated for FourCastNet, the neuralop restatement for FNO/TFNO)
on the host cores, on a bounded sample of the members.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

TFNO_C2 = dict(n_modes=[12, 12], constant_channels=4, prescribed_channels=1, prognostic_channels=8,
               hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4, rank=0.8,
               bias=True, context_size=1)
FNO_C1 = dict(n_modes=[12, 12], constant_channels=0, prescribed_channels=0, prognostic_channels=1,
              hidden_channels=32, lifting_channels=256, projection_channels=256, n_layers=4, context_size=1)
AFNO_C3 = dict(img_height=32, img_width=64, patch_size=[1, 1], constant_channels=4, prescribed_channels=1,
               prognostic_channels=3, embed_dim=64, filter="AFNO2D", depth=4, mlp_ratio=4.0, num_blocks=4,
               sparsity_threshold=0.01, hard_thresholding_fraction=1.0, context_size=1, use_pos_embed=True)
FNO_C5 = dict(n_modes=[64, 64], constant_channels=4, prescribed_channels=1, prognostic_channels=8,
              hidden_channels=128, lifting_channels=256, projection_channels=256, n_layers=4, context_size=1)

# name -> (class, kwargs, H, W, lead steps, default members per GPU, description)
WORKLOADS = {
    "c4": ("TFNO2DModule", TFNO_C2, 32, 64, 20, 256,
           "BASELINE configs[3]: TFNO2D (configs/model/fno.yaml, rank 0.8) 32x64, 20 lead steps (5 days, 6-hourly), "
           "256 ensemble members sharded over the GPUs"),
    "c2": ("TFNO2DModule", TFNO_C2, 32, 64, 20, 8, "BASELINE configs[1]: TFNO2D 32x64, batch 8, 20 lead steps"),
    "c1": ("FNO2DModule", FNO_C1, 64, 64, 50, 8, "BASELINE configs[0]: FNO2D 64x64, batch 8, 50 lead steps"),
    "c3": ("FourCastNet", AFNO_C3, 32, 64, 5, 8, "BASELINE configs[2]: FourCastNet AFNO 32x64, batch 8, 5 lead steps"),
    "c5": ("FNO2DModule", FNO_C5, 721, 1440, 8, 2,
           "BASELINE configs[4]: FNO2D 721x1440, 128 channels, 64x64 modes, 8 lead steps, 2 samples per GPU"),
}


def synthetic_inputs(kw, members, T, H, W, seed, pin):
    import torch
    g = torch.Generator().manual_seed(seed)

    def field(*shape):
        t = torch.randn(*shape, generator=g)
        return t.pin_memory() if pin else t

    const = field(members, 1, kw["constant_channels"], H, W) if kw["constant_channels"] else None
    presc = field(members, T, kw["prescribed_channels"], H, W) if kw["prescribed_channels"] else None
    prog = field(members, T, kw["prognostic_channels"], H, W)
    return const, presc, prog


class ClockSampler:
    """SM clock / throttle reasons sampled while the timed region runs: NVML in-process every 10 ms (the timed
    region of the default run is a few tens of ms), `nvidia-smi -lms 100` when NVML is not importable."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    # NVML clocks-event-reason bits
    BITS = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))

    def __init__(self, index):
        self.index, self.proc, self.lines, self.samples = index, None, [], []
        self.stop, self.thread, self.nvml = threading.Event(), None, None

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = self.index
        if vis:
            try:
                idx = int(vis.split(",")[self.index])
            except (ValueError, IndexError):
                pass
        return pynvml, pynvml.nvmlDeviceGetHandleByIndex(idx)

    def _poll(self, pynvml, h):
        mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
        while True:
            try:
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                try:
                    bits = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                except AttributeError:
                    bits = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.samples.append((float(sm), float(mx), int(bits)))
            except Exception:
                pass
            if self.stop.is_set():
                break
            self.stop.wait(0.01)

    def __enter__(self):
        try:
            pynvml, h = self._nvml_handle()
            self.nvml = pynvml
            self.thread = threading.Thread(target=self._poll, args=(pynvml, h), daemon=True)
            self.thread.start()
            return self
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def __exit__(self, *exc):
        if self.nvml is not None:
            self.stop.set()
            self.thread.join(timeout=2)      # (the poll loop takes its last sample after the stop request)
        elif self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        return False

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for s_, m_, bits in self.samples:
            sm.append(s_)
            mx.append(m_)
            for n, bit in self.BITS:
                if bits & bit:
                    reasons.add(n)
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvml" if self.samples else "nvidia-smi"}


# ----------------------------------------------------------------------------------------
# algorithmic work (DESIGN.md "Measurement")
# ----------------------------------------------------------------------------------------
def fno_dims(kw, H, W):
    m1 = min(kw["n_modes"][0], H)
    m2 = min(kw["n_modes"][1] // 2 + 1, W // 2 + 1)
    cin = kw["constant_channels"] + (kw["prescribed_channels"] + kw["prognostic_channels"]) * kw["context_size"]
    return cin, kw["hidden_channels"], kw["lifting_channels"], kw["projection_channels"], kw["prognostic_channels"], \
        kw["n_layers"], m1, m2


def step_flops(cls, kw, B, H, W):
    """Real flops of one model evaluation on B samples (truncated-DFT-GEMM count)."""
    S = H * W
    if cls != "FourCastNet":
        cin, C, L, P, Cg, N, m1, m2 = fno_dims(kw, H, W)
        lift = 2 * B * S * (cin * L + L * C)
        proj = 2 * B * S * (C * P + P * Cg)
        dftw = 2 * (B * C * H) * W * (2 * m2)
        dfth = 8 * B * C * m2 * H * m1
        mix = 8 * B * C * C * m1 * m2
        spectral = 2 * dftw + 2 * dfth + mix
        return {"step": lift + proj + N * (spectral + 2 * B * S * C * C), "spectral_block": spectral,
                "spectral_mid": 2 * dfth + mix}
    p1, p2 = kw["patch_size"]
    h, w = H // p1, W // p2
    E, nb = kw["embed_dim"], kw["num_blocks"]
    cin = kw["constant_channels"] + (kw["prescribed_channels"] + kw["prognostic_channels"]) * kw["context_size"]
    total = h // 2 + 1
    kept = int(total * kw["hard_thresholding_fraction"])
    rows, cols = min(2 * kept, h), min(kept, w // 2 + 1)
    tok = B * h * w
    dftw = 2 * (B * E * h) * w * (2 * cols)
    dfth = 8 * B * E * cols * h * rows
    mix = 2 * 8 * B * rows * cols * nb * (E // nb) ** 2
    spectral = 2 * dftw + 2 * dfth + mix
    mlp = 4 * tok * E * int(E * kw["mlp_ratio"])
    return {"step": 2 * tok * cin * E * p1 * p2 + 2 * tok * E * kw["prognostic_channels"] * p1 * p2
            + kw["depth"] * (spectral + mlp), "spectral_block": spectral}


def step_bytes(cls, kw, B, H, W, b=4):
    """Algorithmic HBM bytes of one model evaluation with per-layer fusion: every layer reads its
    input once and writes its output once, MLP hiddens stay on chip, weights are read once.
    `b` = bytes per ACTIVATION element between kernels (2 in the bf16 mode); the model's inputs and outputs
    (NCHW frames of the reference) stay fp32."""
    S = H * W
    if cls != "FourCastNet":
        cin, C, L, P, Cg, N, m1, m2 = fno_dims(kw, H, W)
        act = B * S * (4 * cin + b * (C + 2 * N * C + C) + 4 * Cg)
        wts = N * (2 * C * C * m1 * m2 + C * C + C) * b + (cin * L + L + L * C + C + C * P + P + P * Cg + Cg) * b
        return act + wts
    p1, p2 = kw["patch_size"]
    E, D = kw["embed_dim"], kw["depth"]
    tok = B * (H // p1) * (W // p2)
    cin = kw["constant_channels"] + (kw["prescribed_channels"] + kw["prognostic_channels"]) * kw["context_size"]
    hid = int(E * kw["mlp_ratio"])
    act = B * S * b * (cin + 2 * kw["prognostic_channels"]) + tok * E * b * (1 + 4 * D + 1)
    wts = b * (cin * p1 * p2 * E + E + tok // B * E + D * (4 * E + 4 * (E // kw["num_blocks"]) * E + 4 * E
                                                         + 2 * E * hid + hid + E) + E * kw["prognostic_channels"] * p1 * p2)
    return act + wts


def kernel_bytes(cls, kw, B, H, W, name, tcw=False, fused_tail=False):
    """Algorithmic bytes ONE model evaluation moves through the kernel called `name` (sum over
    its launches in the step), or None when bench.py has no model for that kernel."""
    S, b = H * W, 4
    if cls == "FourCastNet":
        p1, p2 = kw["patch_size"]
        tok = B * (H // p1) * (W // p2)
        E, D, hid, Cg = kw["embed_dim"], kw["depth"], int(kw["embed_dim"] * kw["mlp_ratio"]), kw["prognostic_channels"]
        cin = kw["constant_channels"] + (kw["prescribed_channels"] + Cg) * kw["context_size"]
        if name == "pw_linear_kernel":
            return b * (B * S * cin + tok * E + D * (tok * E + tok * hid + tok * hid + 2 * tok * E)
                        + tok * E + 2 * B * S * Cg)
        if name == "layernorm_kernel":
            return b * D * 2 * 2 * tok * E
        # fused AFNO kernels (afno_fused.cu, tc_afno.cu; 1x1 patches): retained half spectrum [B, h, Kc, E] complex
        h, w = H // p1, W // p2
        frac = kw.get("hard_thresholding_fraction", 1.0)
        kc = int((h // 2 + 1) * frac)
        rows = min(2 * kc, h)
        spec = B * h * kc * E * 2
        table = {
            # block 0 reads the NCHW step input and writes x instead of reading it
            "afno_fwd_kernel": b * ((B * S * cin + tok * E) + (D - 1) * tok * E + D * spec),
            "afno_mid_kernel": b * D * (2 * spec + 4 * 2 * E * (E // kw["num_blocks"])),
            "afno_inv_kernel": b * D * (spec + 2 * tok * E),
            # the last block writes the NCHW frame (+ reads the residual frame) instead of the token stream
            "afno_mlp_kernel": b * ((D - 1) * 2 * tok * E + tok * E + 2 * B * S * Cg) + D * 2 * 2 * 2 * E * hid,
        }
        # the tcgen05 forms of the W-DFT kernels move the same tensors
        table["afno_fwd_tc_kernel"] = table["afno_fwd_kernel"]
        table["afno_inv_tc_kernel"] = table["afno_inv_kernel"]
        # split-bf16 operand images both ways: the bytes of the fp32 spectrum; (re|im)-stacked filter weights of both layers
        table["afno_mid_tc_kernel"] = b * D * (2 * spec + 2 * 2 * 4 * E * (E // kw["num_blocks"]))
        return table.get(name)
    cin, C, L, P, Cg, N, m1, m2 = fno_dims(kw, H, W)
    px = B * S
    if name.startswith("tcw_") or name.startswith("wide_") or tcw:
        # bf16 tile-image path (tc_wide.cuh): 128-slot tiles, rows padded to a whole number of tiles
        pxp = B * H * ((W + 127) // 128) * 128
        half = B * H * m2 * C * 2               # complex fp32 [B, H, Kc, C], in floats
        modes = B * m1 * m2 * C * 2
        kz = (2 * m2 + 15) // 16 * 16
        table = {
            "wide_gather_kernel": 4 * px * cin + 2 * pxp * 32,
            "tcw_lift": 2 * pxp * (32 + C),
            "tcw_fwd_wdft": N * (2 * pxp * C + 2 * B * m2 * ((H + 63) // 64) * 128 * C),
            "tcw_fwd_hdft": N * (2 * B * m2 * ((H + 63) // 64) * 128 * C + 4 * modes),
            "mix_dense_kernel": N * (4 * modes + 2 * B * m2 * 128 * C + 4 * m1 * m2 * C * C * 2),
            "tcw_inv_hdft": N * (2 * B * m2 * 128 * C + 2 * B * H * kz * C),
            "tcw_layer": N * (2 * 2 * pxp * C + 2 * B * H * kz * C),
            "tcw_proj": 2 * pxp * C + 4 * 2 * px * Cg,
        }
        return table.get(name)
    if name == "pw_linear_kernel":      # lifting 1+2, N skips, projection 1+2 (unfused fp32 path)
        return b * px * ((cin + L) + (L + C) + N * 2 * C + (C + P) + (P + 2 * Cg))
    if name == "tdft_kernel":           # 4 truncated-DFT passes per layer
        half = B * H * m2 * C * 2       # complex [B, H, Kc, C]
        modes = B * m1 * m2 * C * 2
        return b * N * ((px * C + half) + (half + modes) + (modes + half) + (half + 2 * px * C + px * C))
    if name in ("mix_dense_kernel", "mix_few_tc_kernel"):
        return b * N * (2 * B * m1 * m2 * C * 2 + m1 * m2 * C * C * 2)
    if name in ("fno_lift_kernel", "fno_lift4_kernel"):       # fused lifting MLP (+ forward W-DFT of its output)
        return b * px * (cin + C) + b * B * H * m2 * C * 2
    if name == "fno_layer_kernel":      # fused inverse W-DFT + skip + GELU (+ forward W-DFT, not after the last layer)
        if fused_tail:                  # the last layer runs inside fno_layerproj_kernel: N - 1 launches, each with a forward W-DFT
            return b * (N - 1) * (2 * px * C + 2 * B * H * m2 * C * 2)
        return b * (N * 2 * px * C + (2 * N - 1) * B * H * m2 * C * 2)
    if name == "fno_layerproj_kernel":  # last layer (activation image + Z' in) + projection MLP + residual frame (NCHW in / out)
        return b * (px * C + B * H * m2 * C * 2 + 2 * px * Cg)
    if name in ("fno_mid_kernel", "fno_mid_tc_kernel"):   # H-DFT, mode mixing, inverse H-DFT on the retained columns
        return b * N * (2 * B * H * m2 * C * 2 + m1 * m2 * C * C * 2)
    if name == "fno_tail_kernel":       # last layer + projection MLP + residual
        return b * (px * C + B * H * m2 * C * 2 + 2 * px * Cg)
    return None


# ----------------------------------------------------------------------------------------
# CPU arm: the oracle on the host cores (cpu_baseline leg and --impl reference)
# ----------------------------------------------------------------------------------------
def oracle_module(cls, kw):
    import torch
    from oracle import afno as oafno, wrappers as owrap
    torch.manual_seed(1234)
    if cls == "FourCastNet":
        return oafno.AFNONet(**kw).eval()
    return getattr(owrap, cls)(**kw).eval()


def cpu_rollout_rate(cls, kw, H, W, lead, sample_members, repeats, warmup, state_dict=None, inputs=None):
    """samples*steps/s of the CPU oracle on `sample_members` members x `lead` lead steps; also returns the last
    rollout (the GPU arm checks its own output against it: `parity` on the bench line).  `state_dict` / `inputs`:
    the timed module's weights and the first members of its inputs instead of fresh random ones."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    model = oracle_module(cls, kw)
    if state_dict is not None:
        model.load_state_dict(state_dict, strict=True)
    ctx = kw["context_size"]
    if inputs is None:
        inputs = synthetic_inputs(kw, sample_members, ctx + lead, H, W, 1234, pin=False)
    const, presc, prog = (None if t is None else t[:sample_members, :(1 if i == 0 else ctx + lead)]
                          for i, t in enumerate(inputs))
    times, out = [], None
    with torch.no_grad():
        for i in range(warmup + repeats):
            t0 = time.perf_counter()
            out = model(constants=const, prescribed=presc, prognostic=prog)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    total = sum(times)
    return sample_members * lead * len(times) / total, total / len(times), cores, out


def gpu_eager_rate(cls, kw, lead, state_dict, inputs, dev, repeats=3):
    """The oracle modules moved to the GPU (`.cuda()`): stock PyTorch eager (cuFFT, cuBLAS / cuDNN, torch defaults for
    TF32) on the same inputs and weights -- the number BASELINE.md calls the meaningful one to beat, since the reference
    ships no GPU kernels of its own.  Timed with CUDA events, outside the bench's timed region."""
    import torch
    model = oracle_module(cls, kw)
    model.load_state_dict(state_dict, strict=True)
    model = model.to(dev)
    const, presc, prog = inputs
    members = prog.shape[0]
    with torch.no_grad():
        model(constants=const, prescribed=presc, prognostic=prog)
        torch.cuda.synchronize(dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(repeats):
            model(constants=const, prescribed=presc, prognostic=prog)
        e1.record()
        torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1) / repeats
    del model
    torch.cuda.empty_cache()
    return members * lead / (ms * 1e-3), ms


def run_reference(args, wl):
    cls, kw, H, W, lead, members_default, desc = wl
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    total, per_gpu, _, _ = member_split(args, members_default, 0)
    members = min(args.cpu_members, per_gpu)
    if args.workload == "c5":
        lead = 1
    rate, sec, cores, _ = cpu_rollout_rate(cls, kw, H, W, lead, members, max(args.steps, 1), min(args.warmup, 1))
    sample = f"{members} members x {lead} lead steps per bench step (bounded sample of the workload), fp32, torch CPU"
    line = {
        "impl": "reference", "metric": "rollout samples*steps/s", "value": rate, "unit": "samples*steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_block(args, wl, per_gpu, total),
        "cpu_baseline": {"value": rate, "unit": "samples*steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "samples*steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference CPU path = oracle/ (AFNO: restatement pinned to the reference source; FNO/TFNO: "
                "restatement of the un-vendored neuralop, parity unpinned); the reference is pure Python, "
                "so there is no oracle/_ref binary",
    }
    print(json.dumps(line), flush=True)


def member_split(args, members_default, rank):
    """(members in total, members on the largest rank, this rank's first member, this rank's member count)."""
    m = args.members or members_default
    if args.scaling == "strong":
        if args.impl == "ours":
            from dlwp_benchmark_b200.ensemble import shard_range
        else:      # the reference arm imports nothing of the product: same contiguous balanced split
            def shard_range(total, r, world):
                base, extra = divmod(total, world)
                start = r * base + min(r, extra)
                return start, start + base + (1 if r < extra else 0)
        a, b = shard_range(m, rank, args.gpus)
        a0, b0 = shard_range(m, 0, args.gpus)
        return m, b0 - a0, a, b - a
    return m * args.gpus, m, rank * m, m


def config_block(args, wl, members, total):
    cls, kw, H, W, lead, _, desc = wl
    return {"workload": f"{args.workload}: {desc}", "model_class": cls, "grid": [H, W], "lead_steps": lead,
            "members_per_gpu": members, "members_total": total,
            "parallelism": f"ensemble-sharded x{args.gpus} ({args.scaling} scaling), no data-path collective; one in-place "
                           "all-reduce of [T,C,4] fp64 metric partials per rollout",
            "l2": "inputs larger than L2" if members * (lead + 1) * kw["prognostic_channels"] * H * W * 4 > 126e6
                  else "L2 flushed between timed rollouts (256 MiB write)",
            "precision_mode": args.precision}


# ----------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------
NCU_FULL_CSV = "r02_ncu_full_c4.csv"      # `ncu --set full` summary of this command on the current kernels (tools/ncu_summary_csv.py)


def ncu_traffic(kernel):
    """DRAM bytes per launch of `kernel` from the committed `ncu --set full` summary (profiles/), or None."""
    import csv
    path = os.path.join(ROOT, "profiles", NCU_FULL_CSV)
    try:
        with open(path) as f:
            rows = list(csv.reader(f))
        hdr = rows[0]
        kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot, n = 0.0, 0
        for r in rows[2:]:
            if kernel in r[kn]:
                tot += float(r[rd]) * unit[rows[1][rd]] + float(r[wr]) * unit[rows[1][wr]]
                n += 1
        return tot / n if n else None
    except (OSError, ValueError, KeyError, IndexError):
        return None


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p.get("bf16_tflops_sustained", p.get("bf16_tflops"))), "measured"
    except (OSError, KeyError, ValueError):
        return 6650.0, 1400.0, "fallback"


def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    cls, kw, H, W, lead, members_default, desc = wl
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import __graft_entry__ as entry
    if rank == 0:
        entry.build()
    if world > 1:
        dist.barrier()
    import dlwp_benchmark_b200 as pkg
    from dlwp_benchmark_b200 import ensemble, metrics as M, native
    numa_bound = ensemble.bind_to_device_cpus(local) if world > 1 else False   # before any pinned allocation

    members_total, members_max, _, members = member_split(args, members_default, rank)
    ctx = kw["context_size"]
    T = ctx + lead
    torch.manual_seed(1234)                                   # identical random-init weights on every rank
    mod = getattr(pkg, cls)(**kw).to(dev).eval()
    mod.precision = {"fp32": native.PREC_FP32, "fp32_tc": native.PREC_FP32_TC, "bf16": native.PREC_BF16}[args.precision]
    const_h, presc_h, prog_h = synthetic_inputs(kw, members, T, H, W, 1234 + rank, pin=True)
    Cg = kw["prognostic_channels"]
    g = torch.Generator().manual_seed(99 + rank)
    target = torch.randn(members, lead, Cg, H, W, generator=g).to(dev)
    clim = (0.3 * torch.randn(Cg, H, W, generator=torch.Generator().manual_seed(7))).to(dev)
    const, presc, prog = (None if t is None else t.to(dev) for t in (const_h, presc_h, prog_h))
    in_bytes = sum(t.numel() * 4 for t in (const_h, presc_h, prog_h) if t is not None)
    # what dlwp_rollout_host copies: constants, prescribed frames [0, T-1), and the first ctx frames of prognostic
    # whole job (all ranks): constants, the first ctx prognostic frames, prescribed frames [0, T-1) in; the forecast out
    h2d_bytes = members_total * H * W * 4 * (kw["constant_channels"] + ctx * Cg + (T - 1) * kw["prescribed_channels"])
    out_bytes = members_total * lead * Cg * H * W * 4
    flush = None if in_bytes > 126e6 else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    count = members_total * H * W      # (sample, lat, lon) points over all ranks: known without communication
    sums_buf = torch.zeros(lead, Cg, 4, dtype=torch.float64, device=dev)

    def step():
        out = mod(constants=const, prescribed=presc, prognostic=prog)
        sums_buf.zero_()
        M.partial_sums(out, target, clim, out=sums_buf)
        M.reduce_sums(sums_buf)      # in place; NCCL when world > 1
        return out, sums_buf

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    with torch.no_grad():
        for _ in range(max(args.warmup, 3)):
            out, sums = step()
        barrier()
        n0 = native.launch_count()
        pairs = []
        with ClockSampler(local) as clk:
            for _ in range(args.steps):
                if flush is not None:
                    flush.fill_(1)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                out, sums = step()
                e1.record(stream)
                pairs.append((e0, e1))
            barrier()
        launches = native.launch_count() - n0
        ms_total = sum(a.elapsed_time(b) for a, b in pairs)
        rmse, acc = M.finalize(sums, count)

        # end to end through the C ABI with HOST buffers (H2D + rollout + D2H inside the call)
        out_h = torch.empty(members, lead, Cg, H, W, dtype=torch.float32, pin_memory=True)
        for _ in range(2):
            mod.forward_host(constants=const_h, prescribed=presc_h, prognostic=prog_h, out=out_h)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            mod.forward_host(constants=const_h, prescribed=presc_h, prognostic=prog_h, out=out_h)
        barrier()
        e2e_s = time.perf_counter() - t0

        # per-kernel device times of one rollout (events after every launch; separate pass)
        with native.kernel_profile(stream.cuda_stream) as prof:
            step()
    t = torch.tensor([ms_total, e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, e2e_s = float(t[0]), float(t[1])
    if rank == 0:
        hbm_peak, tc_peak, which = measured_peaks()
        units = members_total * lead * args.steps
        ms_step = ms_total / args.steps
        entries = prof.entries
        name, (kms, kn) = max(entries.items(), key=lambda kv: kv[1][0])
        tcw = "tcw_layer" in entries
        kb = kernel_bytes(cls, kw, members, H, W, name, tcw, fused_tail="fno_layerproj_kernel" in entries)
        total_prof = sum(v[0] for v in entries.values())
        roof = {"bound": "hbm", "kernel": name, "achieved": None, "peak": hbm_peak, "unit": "GB/s", "frac": None,
                "traffic": None, "peak_source": f"{which} (MEASURED_PEAKS.json hbm_gbs)",
                "kernel_ms_per_launch": kms / kn, "kernel_launches_per_rollout": kn,
                "kernel_share_of_rollout": kms / total_prof}
        if kb is not None:
            per_launch = kb * lead / kn
            roof["achieved"] = per_launch / (kms / kn * 1e-3) / 1e9
            roof["frac"] = roof["achieved"] / hbm_peak
            roof["algorithmic_bytes_per_launch"] = per_launch
        if args.workload == "c4" and members == members_default and world == 1:
            roof["traffic"] = ncu_traffic(name)
            roof["traffic_source"] = (f"profiles/{NCU_FULL_CSV} (ncu --set full of this command on this kernel set, mean over "
                                      "the kernel's launches; null when the summary does not hold the kernel)")
        roof["note"] = ("achieved = ALGORITHMIC bytes per launch / measured time per launch (a fraction of speed of light, "
                        "not measured DRAM traffic: that is `traffic`)")
        fl = step_flops(cls, kw, members, H, W)
        sb = step_bytes(cls, kw, members, H, W, b=2 if args.precision == "bf16" else 4)
        spec_names = [n for n in entries if n in ("tdft_kernel", "mix_dense_kernel", "mix_afno_kernel", "fno_mid_kernel", "fno_mid_tc_kernel",
                                                  "tcw_fwd_wdft", "tcw_fwd_hdft", "tcw_inv_hdft",
                                                  "afno_fwd_kernel", "afno_mid_kernel", "afno_inv_kernel",
                                                  "afno_fwd_tc_kernel", "afno_inv_tc_kernel")]
        spec_ms = sum(entries[n][0] for n in spec_names)
        if tcw:   # the inverse W-DFT runs inside tcw_layer next to the skip conv: its share of that kernel's K extent
            m2c = fno_dims(kw, H, W)[7]
            kzc = (2 * m2c + 15) // 16 * 16
            spec_ms += entries["tcw_layer"][0] * kzc / (kzc + kw["hidden_channels"])
            spec_names.append("tcw_layer (inverse W-DFT share of its K extent)")
        # on the tensor-core path the W-DFTs of the spectral block run inside the lift / layer kernels; the
        # separately timed part is the mid kernel (H-DFT, mode mixing, inverse H-DFT): count only its flops
        fused_w = "fno_mid_kernel" in entries or "fno_mid_tc_kernel" in entries
        spec_flops = fl["spectral_mid"] if fused_w and "spectral_mid" in fl else fl["spectral_block"]
        layers = kw.get("n_layers", kw.get("depth"))
        line = {
            "metric": "rollout samples*steps/s", "value": units / (ms_total * 1e-3), "unit": "samples*steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": {"fp32": "f32", "fp32_tc": "f32 (split-bf16 tensor-core products, fp32 accumulate)",
                      "bf16": "bf16"}[args.precision],
            "data": "synthetic N(0,1) fields, random-init weights (seed 1234)",
            "config": config_block(args, wl, members_max, members_total),
            "e2e": {"value": units / e2e_s, "unit": "samples*steps/s", "h2d_bytes_per_step": h2d_bytes,
                    "d2h_bytes_per_step": out_bytes, "ms_per_step": e2e_s / args.steps * 1e3,
                    "bytes_are": "whole job (sum over ranks)",
                    "api": "dlwp_rollout_host (pinned host buffers in the reference's [B,T,C,H,W] layout; copies in "
                           "constants and the ctx prescribed / prognostic frames ahead of the first step, streams the "
                           "later prescribed frames in and the whole forecast back frame by frame under the steps' compute)"},
            "gpu_launches": launches,
            "host_affinity": "ranks pinned to their GPU's NUMA-local cores (NVML)" if numa_bound else "default",
            "roofline": roof,
            "step_roofline": {"algorithmic_bytes_per_rollout": sb * lead, "achieved_gbs": sb * lead / (ms_step * 1e-3) / 1e9,
                              "frac_of_hbm_peak": sb * lead / (ms_step * 1e-3) / 1e9 / hbm_peak,
                              "tflops": fl["step"] * lead / (ms_step * 1e-3) / 1e12},
            "spectral_block": {"kernels": spec_names, "tflops": (spec_flops * layers * lead / (spec_ms * 1e-3) / 1e12)
                               if spec_ms > 0 else None, "peak_tflops": tc_peak,
                               "flops_counted": ("H-DFT + mode mixing + inverse H-DFT (fno_mid_tc_kernel: three chained split-bf16 GEMMs on "
                                                 "tcgen05; fno_mid_kernel: fp32 FMA); algorithmic complex flops, not the 3 product terms; the "
                                                 "W-DFT GEMMs of the block run on tcgen05 inside fno_lift / fno_layer_kernel"
                                                 if fused_w else "whole spectral block (truncated-DFT-GEMM count)"
                                                 + (": W- and H-DFT GEMMs on tcgen05 in bf16, mode mixing in fp32 FMA" if tcw else "")),
                               "note": "profiled pass; includes the events' launch gaps"},
            "kernel_ms_per_rollout": {n: round(v[0], 4) for n, v in sorted(entries.items(), key=lambda kv: -kv[1][0])},
            "clocks": clk.summary(),
            "rmse_lead1_lastlead": [float(rmse[0].mean()), float(rmse[-1].mean())],
        }
        if world == 1 and not args.no_cpu_baseline:
            # the CPU oracle on the first `cm` members of the TIMED inputs with the timed module's weights: its time is
            # the cpu_baseline, its output checks the timed rollout (`parity`; outside the timed region)
            cm = min(args.cpu_members, members)
            cl = 1 if args.workload == "c5" else lead
            sd = {k: v.detach().cpu() for k, v in mod.state_dict().items()}
            rate, sec, cores, want = cpu_rollout_rate(cls, kw, H, W, cl, cm, 1, 0, state_dict=sd,
                                                      inputs=(const_h, presc_h, prog_h))
            line["cpu_baseline"] = {"value": rate, "unit": "samples*steps/s", "cores": cores, "kind": "port",
                                    "sample": f"{cm} members x {cl} lead steps, one rollout ({sec:.1f} s), "
                                              "oracle/ in torch fp32 on all host threads"}
            got = out[:cm, :cl].float().cpu().double()
            want = want.double()
            x0 = prog_h[:cm, ctx - 1].double()
            incr = (want[:, 0] - x0).abs().max()
            disp = (want - x0.unsqueeze(1)).pow(2).mean(dim=(0, 2, 3, 4)).sqrt()
            traj = (got - want).pow(2).mean(dim=(0, 2, 3, 4)).sqrt() / disp
            line["parity"] = {
                "max_rel_increment": float((got[:, 0] - want[:, 0]).abs().max() / incr),
                "max_rel_output_step1": float((got[:, 0] - want[:, 0]).abs().max() / want[:, 0].abs().max()),
                "traj_err_over_displacement": {"lead_1": float(traj[0]), f"lead_{cl}": float(traj[-1]), "max": float(traj.max())},
                "members_checked": cm, "lead_steps_checked": cl,
                "what": "the timed rollout's first members against the CPU oracle (oracle/, torch fp32) on the same inputs "
                        "with the timed module's state_dict; increment = out_1 - x_0; FNO/TFNO oracle = restatement of the "
                        "un-vendored neuralop (parity unpinned), FourCastNet oracle pinned to the reference source"}
            if not args.no_eager:
                try:
                    erate, ems = gpu_eager_rate(cls, kw, lead, sd, (const, presc, prog), dev)
                except Exception as exc:      # a baseline leg must not take the bench line down
                    erate, ems = None, None
                    line["gpu_eager_baseline_error"] = repr(exc)[:200]
                line["gpu_eager_baseline"] = {
                    "value": erate, "unit": "samples*steps/s", "ms_per_step": ems, "members": members,
                    "what": "oracle/ modules .cuda(): stock PyTorch eager on this B200 (cuFFT rfft2 / irfft2, einsum, cuDNN 1x1 "
                            "convolutions; torch defaults, i.e. TF32 allowed in cuDNN), same inputs and weights, CUDA events, "
                            "3 rollouts after one warm-up"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default="c4")
    ap.add_argument("--members", type=int, default=0,
                    help="ensemble members: in total with --scaling strong, per GPU with --scaling weak (0 = the workload's default)")
    ap.add_argument("--scaling", choices=["strong", "weak"], default=None,
                    help="strong: --members in total, sharded over the GPUs (default for c4); weak: --members per GPU")
    ap.add_argument("--no-eager", action="store_true", help="skip the PyTorch-eager-on-GPU baseline leg")
    ap.add_argument("--precision", choices=["fp32", "fp32_tc", "bf16"], default="fp32_tc")
    ap.add_argument("--cpu-members", type=int, default=64, help="members in the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.scaling is None:
        args.scaling = "strong" if args.workload == "c4" else "weak"
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()

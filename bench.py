#!/usr/bin/env python
"""Headline benchmark: image pairs / s of the RoMa matching hot path (roma_outdoor, 560 -> 864) on N B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

One "step" = one pass of the hot path over one batch of synthetic input: `model.match(im_A, im_B, batched=True)`
on `--batch` (default 8) pairs per GPU (BASELINE.json configs[1]; configs[2] is the same per-GPU micro-batch on
2/4/8 GPUs) followed by `model.sample_batch(warp, certainty, num=10000)` (= `sample` for every pair) — the call sequence of the
reference's benchmarks.  Pairs are independent, so ranks shard them with no collective in the
timed region ("weak" scaling: per-GPU work fixed); NCCL is used only for the barrier and the max-over-ranks time.

  value  : pairs/s with the uint8 input images already resident in HBM.
  e2e    : the same call with pinned HOST images — H2D copy in, D2H copy of the sampled matches out, per step.
  roofline / kernels : per-kernel CUDA-event timings taken live inside the timed region (ops.KernelTimer).
  cpu_baseline / --impl reference : the UNMODIFIED reference from baseline/_ref (installed by __graft_entry__.build(),
           travels with the snapshot) on the host cores — roma_outdoor(device="cpu") through its own forward_symmetric +
           epilogue + sample, fp32, all host threads, 1 pair per step (bounded sample), one warm-up;
           kind "reference".  Only when baseline/_ref is absent: the CPU oracle port of the same path (kind "port").
  aten_cuda_baseline : the same unmodified reference on THIS GPU through ATen (fp16 autocast as shipped, and all-bf16).
  hot_path : Σ SURVEY.md §8(d) roofline floors ÷ Σ measured time of the kernels north_star names (GP, gather / local
           correlation, update / resample / epilogue).
  configs  : short driver-visible runs of BASELINE.json configs[3] (672 -> 1344) and configs[4] (Tiny-RoMa, 480 x 640).
  parity   : deviation of this run's (16-bit) output from the same model with fp32 bodies, first pair of the batch.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "image_pairs_per_s_roma_outdoor_560_864"
UNIT = "pairs/s"


def peaks():
    p = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "_source": "fallback"}
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            p.update(json.loads(f.read_text()))
            p["_source"] = "measured"
        except Exception:
            pass
    return p


def synthetic_pairs(batch: int, res, seed: int):
    """uint8 RGB pairs at the upsample resolution (rand images, the coarse pass sees the same
    content resized on the device)."""
    g = torch.Generator().manual_seed(seed)
    shape = (batch, 3, res[0], res[1])
    low = torch.rand(batch, 3, res[0] // 16, res[1] // 16, generator=g)
    base = F.interpolate(low, size=res, mode="bilinear", align_corners=False)
    A = (0.7 * base + 0.3 * torch.rand(shape, generator=g))
    B = (0.7 * base.roll(shifts=(res[0] // 20, res[1] // 30), dims=(2, 3)) + 0.3 * torch.rand(shape, generator=g))
    to8 = lambda t: (t.clamp(0, 1) * 255).round().to(torch.uint8)
    return to8(A), to8(B)


# ------------------------------------------------------------------------------------------------------
# algorithmic work per launch (restated in DESIGN.md)
# ------------------------------------------------------------------------------------------------------
def alg_work(name: str, d: dict):
    """-> (bound, amount) with amount in bytes (hbm) or FLOPs (tensor) for ONE launch."""
    if name == "roma_refiner_gather":
        e, hw, K = d["esize"], d["B"] * d["h"] * d["w"], (2 * d["r"] + 1) ** 2 if d["r"] else 0
        if d.get("only") == "corr":
            return "hbm", hw * (2 * e * d["C"] + 8 + e * K)
        if d.get("only") == "xhat":
            return "hbm", hw * (2 * e * d["C"] + 8)
        dch = 2 * d["C"] + d["E"] + K
        return "hbm", hw * (e * d["C"] + e * d["C"] + 8 + e * dch)         # read f0, f1, flow; write d
    if name == "roma_gp_posterior":
        N, M, C, D = d["Nx"], d["Ny"], d["C"], d["D"]
        return "tensor", d["B"] * (2.0 * M * M * C + 2.0 * N * M * C + M ** 3 / 3.0 + 2.0 * M * M * D + 2.0 * N * M * D)
    if name == "roma_flow_update_resample":
        src, dst = d["B"] * d["h"] * d["w"], d["B"] * d["h2"] * d["w2"]
        return "hbm", 4 * (src * (3 + (3 if d["delta"] else 0) + (3 if d["cur"] else 0)) + dst * 3)
    if name == "roma_match_epilogue":
        n = d["B"] * d["H"] * d["Wout"]
        return "hbm", 4 * n * (3 + 4 + 1 + (1 if d["logits"] else 0))
    if name == "roma_cls_to_flow_refine":
        return "hbm", d["B"] * d["H"] * d["W"] * (d["C"] * d["esize"] + 8)
    if name in ("roma_dwconv5x5_bias_act", "roma_dwconv5x5_pointwise"):    # one read + one write of the activation
        return "hbm", 2 * d["B"] * d["h"] * d["w"] * d["C"] * d["esize"]
    if name == "roma_layernorm":
        return "hbm", 2 * d["rows"] * d["D"] * d["esize"]
    if name == "roma_gelu":
        return "hbm", 2 * d["n"] * d["esize"]
    if name in ("roma_prepare_images", "roma_prepare_images_pil"):
        return "hbm", d["B"] * 3 * (d["H"] * d["W"] + 4 * d["h"] * d["w"])
    if name == "roma_pointwise_narrow":
        return "hbm", d["B"] * d["h"] * d["w"] * (d["C"] + d["OUT"]) * d["esize"]
    if name == "roma_maxpool2x2_nhwc":
        return "hbm", d["B"] * d["h"] * d["w"] * d["C"] * d["esize"] * 5 // 4
    if name == "roma_refiner_out_conv":
        return "hbm", d["B"] * d["h"] * d["w"] * (d["C"] * d["esize"] + 16)
    if name == "roma_kde":
        return "sfu", float(d["N"]) * (d["N"] // d["down"])
    return "hbm", 0


def summarise_kernels(summary: dict, steps: int, pk: dict):
    rows = []
    for name, recs in summary.items():
        groups = {}
        for dims, ms in recs:
            key = json.dumps(dims, sort_keys=True)
            groups.setdefault(key, []).append(ms)
        for key, mss in groups.items():
            dims = json.loads(key)
            bound, amount = alg_work(name, dims)
            avg = sum(mss) / len(mss)
            row = {"kernel": name, "dims": dims, "launches_per_step": len(mss) / steps, "avg_ms": round(avg, 4),
                   "ms_per_step": round(sum(mss) / steps, 4), "bound": bound}
            if bound == "hbm" and amount:
                row.update(achieved=round(amount / avg / 1e6, 1), unit="GB/s", peak=pk["hbm_gbs"],
                           frac=round(amount / avg / 1e6 / pk["hbm_gbs"], 4), alg=amount)
            elif bound == "tensor":
                row.update(achieved=round(amount / avg / 1e9, 2), unit="TFLOP/s", peak=pk["bf16_tflops_sustained"],
                           frac=round(amount / avg / 1e9 / pk["bf16_tflops_sustained"], 5), alg=amount)
            elif bound == "sfu":
                row.update(achieved=round(amount / avg / 1e6, 1), unit="Gpair/s", alg=amount)
            rows.append(row)
    rows.sort(key=lambda r: -r["ms_per_step"])
    return rows


# ------------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.proc = None
        try:
            uuid = str(torch.cuda.get_device_properties(device_index).uuid)
            ident = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
            self.proc = subprocess.Popen(["nvidia-smi", "-i", ident, f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            text, _ = self.proc.communicate(timeout=5)
        except Exception:
            return out
        sm, reasons, mx = [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in text.strip().splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx = float(parts[1])
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))
        return out


# ------------------------------------------------------------------------------------------------------
# CPU baseline (oracle port) — also the --impl reference arm
# ------------------------------------------------------------------------------------------------------
def cpu_pair_fn(coarse, upsample, weights, dinov2):
    """-> (run(A8, B8, seed), kind).  kind "reference": the unmodified reference (baseline/_ref) on the CPU;
    kind "port": the oracle port over the repo's host modules (only when the reference is not on this machine)."""
    import roma_b200
    from oracle import reference as R
    mean = torch.tensor((0.485, 0.456, 0.406)).view(1, 3, 1, 1)
    std = torch.tensor((0.229, 0.224, 0.225)).view(1, 3, 1, 1)
    rs = lambda t: F.interpolate(t, size=tuple(coarse), mode="bicubic", antialias=True, align_corners=False)
    if R.available():
        ref = R.reference_roma_outdoor("cpu", weights, dinov2, tuple(coarse), tuple(upsample), amp_dtype=torch.float32)

        def run(A8: torch.Tensor, B8: torch.Tensor, seed: int = 0):
            with torch.inference_mode():
                fA, fB = (A8.float() / 255 - mean) / std, (B8.float() / 255 - mean) / std
                warp, cert, _ = R.reference_forward_batched(ref, rs(fA), rs(fB), fA, fB)
                torch.manual_seed(seed)
                return ref.sample(warp[0], cert[0], num=10000)
        return run, "reference"
    from oracle import roma_model_oracle as MO
    from oracle import roma_oracle as O
    cpu = roma_b200.roma_outdoor("cpu", weights, dinov2, coarse_res=coarse, upsample_res=upsample, amp_dtype=torch.float32)
    O.set_aten(True)      # issue the reference's own ATen calls (F.grid_sample, torch.cdist) at its call sites: faithful CPU speed

    def run(A8: torch.Tensor, B8: torch.Tensor, seed: int = 0):
        with torch.inference_mode():
            fA, fB = cpu._normalise(A8), cpu._normalise(B8)
            warp, cert = MO.match_tensors(cpu, rs(fA), rs(fB), fA, fB)
            g = torch.Generator().manual_seed(seed)
            return O.sample(warp[0], cert[0], num=10000, generator=g)
    return run, "port"


WHAT = {"reference": "the UNMODIFIED reference (baseline/_ref): roma_outdoor(device='cpu') forward_symmetric x2 + epilogue + sample",
        "port": "fp32 oracle port issuing the reference's ATen calls (baseline/_ref not on this machine)"}


def spawn_reference_child(args, steps, warmup, budget, timeout_s):
    """The CPU arm in a process of its own with a watchdog: the reference's batched CPU linear algebra has dead-locked inside
    oneMKL on some hosts; a hung comparand must not take the bench line down.  -> parsed JSON line or None."""
    cmd = [sys.executable, str(Path(__file__).resolve()), "--impl", "reference", "--child", "--gpus", str(args.gpus),
           "--steps", str(steps), "--warmup", str(warmup), "--batch", str(args.batch), "--ref-budget", str(budget),
           "--coarse", str(args.coarse[0]), str(args.coarse[1]), "--upsample", str(args.upsample[0]), str(args.upsample[1])]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID"):
        env.pop(k, None)
    try:
        res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout_s, env=env)
    except subprocess.TimeoutExpired:
        return None
    for line in reversed(res.stdout.strip().splitlines()):
        if line.startswith("{"):
            try:
                return json.loads(line)
            except Exception:
                return None
    return None


def run_reference_arm(args, rank, world, force_port=False):
    if rank != 0:
        return
    if not args.child and not force_port:
        got = spawn_reference_child(args, args.steps, args.warmup, args.ref_budget, args.ref_budget * 2 + 240)
        if got is not None:
            print(json.dumps(got))
            return
        return run_reference_arm(args, rank, world, force_port=True)     # watchdog fired: the oracle port, in this process
    import roma_b200
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    weights, dinov2 = roma_b200.random_init_weights(seed=0)
    if force_port:
        from oracle import reference as R
        R.locate = lambda: None
    run, kind = cpu_pair_fn(args.coarse, args.upsample, weights, dinov2)
    A8, B8 = synthetic_pairs(1, args.upsample, 1234)
    budget_s = args.ref_budget
    t_all = time.perf_counter()
    warm = min(args.warmup, 1)
    for _ in range(warm):
        run(A8, B8)
    times = []
    for i in range(args.steps):
        t0 = time.perf_counter()
        run(A8, B8, seed=i)
        times.append(time.perf_counter() - t0)
        if time.perf_counter() - t_all + times[-1] > budget_s:
            break
    per = sum(times) / len(times)
    val = 1.0 / per
    sample = (f"1 pair per step (of the {args.batch}-pair batch) at {args.coarse[0]}->{args.upsample[0]}, match+sample(10000), "
              f"{WHAT[kind]}, fp32 on {cores} host threads; {warm} warm-up + {len(times)} timed steps "
              f"(of {args.steps} requested, capped at {int(budget_s)} s), best {min(times):.1f} s")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(val, 5), "unit": UNIT, "n_gpus": args.gpus,
        "steps": len(times), "warmup": warm, "ms_per_step": round(per * 1e3, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args, cpu=True),
        "cpu_baseline": {"value": round(val, 5), "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": round(val, 5), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------------
# the reference on THIS GPU (ATen), other configs, hot-path aggregate, parity of the 16-bit run
# ------------------------------------------------------------------------------------------------------
def aten_cuda_baseline(args, dev, weights, dinov2, A8, B8):
    """The unmodified reference (baseline/_ref) on the B200 through its stock ATen path: the same batch, forward_symmetric x2
    + epilogue + sample(10000) per pair; fp16 autocast (as shipped) and all-bf16.  Wall clock around synchronised steps."""
    from oracle import reference as R
    if not R.available():
        return {"unavailable": "baseline/_ref is not on this machine"}
    out = {"what": "unmodified reference (baseline/_ref) on this GPU, ATen kernels, same batch and call sequence",
           "pairs_per_step": args.batch}
    mean = torch.tensor((0.485, 0.456, 0.406), device=dev).view(1, 3, 1, 1)
    std = torch.tensor((0.229, 0.224, 0.225), device=dev).view(1, 3, 1, 1)
    rs = lambda t: F.interpolate(t, size=tuple(args.coarse), mode="bicubic", antialias=True, align_corners=False)
    for name in ("float16", "bfloat16"):
        try:
            model = R.reference_roma_outdoor(dev, weights, dinov2, tuple(args.coarse), tuple(args.upsample), amp_dtype=getattr(torch, name))
            t_mod = {}

            def step(record=False):
                t = [time.perf_counter()]
                fa, fb = (A8.to(dev).float() / 255 - mean) / std, (B8.to(dev).float() / 255 - mean) / std
                warp, cert, _ = R.reference_forward_batched(model, rs(fa), rs(fb), fa, fb)
                if record:
                    torch.cuda.synchronize(); t.append(time.perf_counter())
                for i in range(args.batch):
                    model.sample(warp[i], cert[i], num=args.num)
                if record:
                    torch.cuda.synchronize(); t.append(time.perf_counter())
                    t_mod["match_ms"], t_mod["sample_ms"] = (t[1] - t[0]) * 1e3, (t[2] - t[1]) * 1e3
            step()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            n = 2
            for _ in range(n):
                step()
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / n
            step(record=True)
            out[name] = {"pairs_per_s": round(args.batch / dt, 3), "ms_per_step": round(dt * 1e3, 2),
                         "match_ms_per_step": round(t_mod["match_ms"], 2), "sample_ms_per_step": round(t_mod["sample_ms"], 2)}
            del model
        except Exception as ex:  # noqa: BLE001 — a failing comparand must not take the bench line down
            out[name] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
        torch.cuda.empty_cache()
    return out


def survey_floor_ms(rows, pk):
    """Σ SURVEY.md §8(d) floors vs Σ measured time for the kernels north_star names, per step.
    K2 bytes per launch (§8d verbatim): e·C·hw·[r>0] + e·C·hw + 8·hw + e·K·hw + e·C·hw (no x copy at the gather-only scales, no
    embedding); GP FLOPs: 2·(2N²C) + N³/3 + 2N²D + 2N²D; K3a: bytes read + written once (alg_work)."""
    hbm, tc = pk["hbm_gbs"] * 1e9, pk["bf16_tflops_sustained"] * 1e12
    parts = {"gp": [0.0, 0.0], "gather_local_corr": [0.0, 0.0], "update_resample_epilogue": [0.0, 0.0]}
    for r in rows:
        d, n = r["dims"], r["launches_per_step"]
        if r["kernel"] == "roma_gp_posterior":
            parts["gp"][0] += n * r["alg"] / tc * 1e3
            parts["gp"][1] += r["ms_per_step"]
        elif r["kernel"] == "roma_refiner_gather" and not d.get("only"):
            e, hw, K = d["esize"], d["B"] * d["h"] * d["w"], (2 * d["r"] + 1) ** 2 if d["r"] else 0
            b = hw * (e * d["C"] * (1 if d["r"] else 0) + e * d["C"] + 8 + e * K + e * d["C"])
            parts["gather_local_corr"][0] += n * b / hbm * 1e3
            parts["gather_local_corr"][1] += r["ms_per_step"]
        elif r["kernel"] in ("roma_flow_update_resample", "roma_match_epilogue"):
            parts["update_resample_epilogue"][0] += n * r["alg"] / hbm * 1e3
            parts["update_resample_epilogue"][1] += r["ms_per_step"]
    floor, meas = sum(v[0] for v in parts.values()), sum(v[1] for v in parts.values())
    return {"floor_ms_per_step": round(floor, 4), "measured_ms_per_step": round(meas, 4), "frac": round(floor / meas, 4) if meas else None,
            "parts": {k: {"floor_ms": round(v[0], 4), "measured_ms": round(v[1], 4), "frac": round(v[0] / v[1], 4) if v[1] else None}
                      for k, v in parts.items()},
            "note": "GP measured in-step (it overlaps the upsample-pass encoder); peaks: measured HBM copy and sustained bf16"}


def other_configs(args, dev, model):
    """BASELINE.json configs[3] and configs[4], short runs (2 warm-up + 3 timed steps, CUDA events)."""
    import roma_b200
    from roma_b200.host.tiny import TinyRoMa, XFeatBackbone, random_init_tiny
    out = {}

    def time_steps(fn, warm=2, steps=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            fn()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / steps
    try:
        keep = (model.h_resized, model.w_resized, model.upsample_res)
        model.h_resized, model.w_resized, model.upsample_res = 672, 672, (1344, 1344)      # plain attributes, as in the reference
        A8, B8 = synthetic_pairs(4, (1344, 1344), 99)
        A8, B8 = A8.to(dev), B8.to(dev)

        def step4():
            warp, cert = model.match(A8, B8, batched=True, device=dev)
            model.sample_batch(warp, cert, num=args.num)
        ms = time_steps(step4)
        out["config4_672_1344"] = {"pairs_per_s": round(4 / ms * 1e3, 2), "ms_per_step": round(ms, 2), "pairs_per_step": 4,
                                   "workload": "roma_outdoor 672 -> 1344 (GP N = 2304), match + sample(10000)", "dtype": args.dtype}
        model.h_resized, model.w_resized, model.upsample_res = keep
        del A8, B8
    except Exception as ex:  # noqa: BLE001
        out["config4_672_1344"] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
    torch.cuda.empty_cache()
    try:
        weights, trunk_sd = random_init_tiny(seed=0)
        trunk = XFeatBackbone()
        trunk.load_state_dict(trunk_sd)
        tiny = TinyRoMa(trunk)
        tiny.load_state_dict(weights)
        tiny = tiny.to(dev).train(False)
        g = torch.Generator().manual_seed(4321)
        Bt = 32
        A = torch.rand(Bt, 3, 480, 640, generator=g).to(dev)
        Bm = torch.rand(Bt, 3, 480, 640, generator=g).to(dev)

        def step5():
            warp, cert = tiny.match(A, Bm, batched=True)
            for i in range(Bt):
                tiny.sample(warp[i], cert[i], num=5000)
        ms = time_steps(step5)
        out["config5_tiny_480x640"] = {"pairs_per_s": round(Bt / ms * 1e3, 1), "ms_per_step": round(ms, 2), "pairs_per_step": Bt,
                                       "workload": "tiny_roma_v1_outdoor (XFeat stand-in trunk), 32 pairs/GPU of 480 x 640, match + "
                                                   "sample(5000); configs[4] proper = 8 such replicas", "dtype": "f32 trunk, f16 kde"}
    except Exception as ex:  # noqa: BLE001
        out["config5_tiny_480x640"] = {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}
    torch.cuda.empty_cache()
    return out


def parity_block(args, dev, model, weights, dinov2, A8, B8):
    """How far this run's 16-bit bodies move the output: first pair of the batch against the SAME model with fp32 bodies
    (the configuration tests/test_refparity_e2e_gpu.py holds to 1e-3 / 1e-2 against the reference)."""
    import roma_b200
    try:
        a, b = A8[:1].to(dev), B8[:1].to(dev)
        w16, c16, l16 = model.match(a, b, batched=True, device=dev, return_logits=True)
        tf = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
        torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = False
        m32 = roma_b200.roma_outdoor(dev, weights, dinov2, coarse_res=args.coarse, upsample_res=args.upsample, amp_dtype=torch.float32)
        w32, c32, l32 = m32.match(a, b, batched=True, device=dev, return_logits=True)
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = tf
        del m32
        torch.cuda.empty_cache()
        dw, dl = (w16 - w32).abs().flatten(), (l16 - l32).abs().flatten()
        q = lambda t, f: t.float().kthvalue(max(int(t.numel() * f), 1)).values.item()
        return {"what": f"{args.dtype} bodies vs fp32 bodies of the same model, pair 0, {args.coarse[0]}->{args.upsample[0]}; north_star "
                        "tolerances (warp 1e-3, logits 1e-2) are held by the fp32-body configuration against the reference",
                "warp_abs": {"max": round(dw.max().item(), 5), "p99": round(q(dw, 0.99), 6), "mean": round(dw.mean().item(), 6)},
                "logit_abs": {"max": round(dl.max().item(), 4), "p99": round(q(dl, 0.99), 5), "mean": round(dl.mean().item(), 5)}}
    except Exception as ex:  # noqa: BLE001
        return {"error": f"{type(ex).__name__}: {str(ex)[:200]}"}


def config_dict(args, cpu=False):
    return {"workload": f"roma_outdoor random-init, {args.batch} synthetic pairs/GPU, coarse {args.coarse[0]}x{args.coarse[1]} -> "
                        f"upsample {args.upsample[0]}x{args.upsample[1]}, symmetric, match + sample(10000) per pair",
            "pairs_per_gpu_per_step": 1 if cpu else args.batch, "coarse_res": list(args.coarse), "upsample_res": list(args.upsample),
            "amp_dtype": "float32" if cpu else args.dtype, "weights": "seeded random init (seed 0)",
            "l2": "no explicit flush: each step streams > 10 GB of activations through a 126 MB L2",
            "parallelism": f"pair-sharded x{args.gpus}, no collective in the timed region"}


# ------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="roma_b200", choices=["roma_b200", "reference"])
    ap.add_argument("--batch", type=int, default=8, help="pairs per GPU per step")
    ap.add_argument("--coarse", type=int, nargs=2, default=[560, 560])
    ap.add_argument("--upsample", type=int, nargs=2, default=[864, 864])
    ap.add_argument("--dtype", default="bfloat16", choices=["bfloat16", "float16", "float32"])
    ap.add_argument("--num", type=int, default=10000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--child", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--no-extras", action="store_true", help="skip aten_cuda_baseline / configs / parity (profiling runs)")
    ap.add_argument("--sample-lanes", type=int, default=2, help="streams sample_batch spreads the pairs over")
    ap.add_argument("--decoder-priority", type=int, default=1, help="0: upsample-pass encoder on a side stream, coarse decoder on "
                    "the caller's; 1 (default): coarse decoder on a high-priority stream")
    ap.add_argument("--ref-budget", type=float, default=200.0, help="wall-clock cap (s) of the CPU reference arm")
    args = ap.parse_args()
    args.coarse, args.upsample = tuple(args.coarse), tuple(args.upsample)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import roma_b200
    from roma_b200 import _lib, ops
    from roma_b200.sharding import aggregate

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    if lib.roma_check_device(local) != 0:
        raise RuntimeError(lib.roma_last_error_string().decode())
    amp = getattr(torch, args.dtype)
    weights, dinov2 = roma_b200.random_init_weights(seed=0)
    model = roma_b200.roma_outdoor(dev, weights, dinov2, coarse_res=args.coarse, upsample_res=args.upsample, amp_dtype=amp)

    model.decoder_priority = bool(args.decoder_priority)
    model.sample_lanes = args.sample_lanes
    A8, B8 = synthetic_pairs(args.batch, args.upsample, 1234 + rank)
    A_host, B_host = A8.pin_memory(), B8.pin_memory()
    A_dev, B_dev = A8.to(dev), B8.to(dev)
    out_host = torch.empty((args.batch, args.num, 5), dtype=torch.float32).pin_memory()
    gens = [torch.Generator(device=dev) for _ in range(args.batch)]

    def step(a, b, to_host: bool):
        warp, cert = model.match(a, b, batched=True, device=dev)
        ms, cs = model.sample_batch(warp, cert, num=args.num, generators=gens)      # `sample` for every pair
        if to_host:
            for i in range(args.batch):
                out_host[i, :, :4].copy_(ms[i], non_blocking=True)
                out_host[i, :, 4].copy_(cs[i], non_blocking=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            fn()
        e.record()
        barrier()
        _, ms = aggregate(0, s.elapsed_time(e), dev)         # max over ranks (NCCL all-reduce of one scalar)
        return ms

    for i, g in enumerate(gens):
        g.manual_seed(100 + i)
    for _ in range(args.warmup):
        step(A_dev, B_dev, False)
    # ---- timed region 1: device-resident inputs (value) with live per-kernel events ----------------------
    clocks = ClockSampler(local)
    l0 = lib.roma_kernel_launches()
    ms_dev = timed(lambda: step(A_dev, B_dev, False), args.steps)
    launches = int(lib.roma_kernel_launches() - l0)
    # ---- same steps again with a CUDA-event pair around every own kernel (per-kernel table; not the headline) ---
    with ops.KernelTimer() as kt:
        ms_inst = timed(lambda: step(A_dev, B_dev, False), args.steps)
    # ---- timed region 2: end to end from pinned host memory ----------------------------------------------
    step(A_host, B_host, True)
    ms_e2e = timed(lambda: step(A_host.to(dev, non_blocking=True), B_host.to(dev, non_blocking=True), True), args.steps)
    clk = clocks.stop()

    pairs = args.batch * world * args.steps
    value = pairs / (ms_dev / 1e3)
    e2e = pairs / (ms_e2e / 1e3)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk = peaks()
    rows = summarise_kernels(kt.summary(), args.steps, pk)
    own_ms = sum(r["ms_per_step"] for r in rows)
    # dominant kernel = the kernel (entry point, all its shapes together) with the largest share of the step; the
    # roofline block reports it at the shape it spends most of its time on
    fam = {}
    for r in rows:
        if r["bound"] in ("hbm", "tensor") and r.get("alg"):
            fam[r["kernel"]] = fam.get(r["kernel"], 0.0) + r["ms_per_step"]
    top_name = max(fam, key=fam.get) if fam else None
    top = next((r for r in rows if r["kernel"] == top_name and r.get("alg")), None)
    roofline = None
    if top:
        # DRAM bytes per launch of the same kernel and shape from the committed `ncu --set full` capture
        traffic = None
        try:
            tj = json.loads((ROOT / "profiles" / "traffic.json").read_text())
            traffic = tj.get(top["kernel"] + " " + json.dumps(top["dims"], sort_keys=True))
        except Exception:
            traffic = None
        roofline = {"kernel": top["kernel"], "dims": top["dims"], "bound": top["bound"], "achieved": top["achieved"],
                    "peak": top["peak"], "unit": top["unit"], "frac": top["frac"], "traffic": traffic,
                    "peak_source": pk["_source"] + (" sustained" if top["bound"] == "tensor" else ""),
                    "avg_launch_ms": top["avg_ms"], "share_of_step": round(fam[top_name] / (ms_dev / args.steps), 4),
                    "share_of_step_this_shape": round(top["ms_per_step"] / (ms_dev / args.steps), 4)}
    # the two kernels BASELINE.json's metric names: GP tensor utilisation and local-correlation HBM GB/s
    headline = {}
    gp_row = next((r for r in rows if r["kernel"] == "roma_gp_posterior"), None)
    if gp_row:
        d = gp_row["dims"]
        side_x, side_y = int(round(math.sqrt(d["Nx"]))), int(round(math.sqrt(d["Ny"])))
        if side_x * side_x == d["Nx"] and side_y * side_y == d["Ny"]:
            # in the step the call shares the SMs with the upsample-pass encoder (second stream); time it alone as well
            g = torch.Generator().manual_seed(7)
            mk = lambda n: torch.randn(d["B"], d["C"], n, n, generator=g).to(dev, amp).contiguous(memory_format=torch.channels_last)
            xg, yg = mk(side_x), mk(side_y)
            pw, pb = torch.randn(d["D"], 2, generator=g).to(dev), torch.randn(d["D"], generator=g).to(dev)
            for _ in range(2):
                ops.gp_posterior(xg, yg, pw, pb)
            torch.cuda.synchronize()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            for _ in range(5):
                ops.gp_posterior(xg, yg, pw, pb)
            e.record()
            torch.cuda.synchronize()
            alone = s.elapsed_time(e) / 5
            gp_row = dict(gp_row, alone_ms=round(alone, 4), alone_achieved=round(gp_row["alg"] / alone / 1e9, 2),
                          alone_frac=round(gp_row["alg"] / alone / 1e9 / pk["bf16_tflops_sustained"], 5),
                          note="avg_ms is measured inside the step, where the call overlaps the upsample-pass encoder on a second "
                               "stream; alone_ms is the same shape timed by itself right after the timed region")
        headline["gp"] = gp_row
    headline["local_corr"] = [r for r in rows if r["kernel"] == "roma_refiner_gather" and r["dims"].get("r")]
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_dev / args.steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": {"bfloat16": "bf16", "float16": "f16", "float32": "f32"}[args.dtype], "data": "synthetic",
        "config": config_dict(args),
        "e2e": {"value": round(e2e, 3), "unit": UNIT, "h2d_bytes_per_step": int(A_host.numel() + B_host.numel()) * world,
                "d2h_bytes_per_step": int(out_host.numel() * 4) * world, "ms_per_step": round(ms_e2e / args.steps, 3)},
        "gpu_launches": launches, "clocks": clk, "roofline": roofline,
        "own_kernels_ms_per_step": round(own_ms, 3), "instrumented_ms_per_step": round(ms_inst / args.steps, 3),
        "hot_path": survey_floor_ms(rows, pk), "headline_kernels": headline, "kernels": rows[:12],
    }
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        if not args.no_extras:
            line["parity"] = parity_block(args, dev, model, weights, dinov2, A8, B8)
            line["aten_cuda_baseline"] = aten_cuda_baseline(args, dev, weights, dinov2, A8, B8)
            line["configs"] = other_configs(args, dev, model)
        got = spawn_reference_child(args, 2, 1, 150.0, 600.0)      # the --impl reference arm, bounded: 1 warm-up + 2 steps
        if got is not None and "cpu_baseline" in got:
            line["cpu_baseline"] = got["cpu_baseline"]
        else:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": cores, "kind": "reference",
                                    "sample": "the CPU arm did not finish within its 600 s watchdog on this host"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Hot-path benchmark.  Headline: fused exhaustive descriptor matching, N=M=8192 keypoints, d=768;
the "extract" object carries the second half of BASELINE.json's metric (images/s, ViT-B/14, 518 px).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = matching every pair of one resident block of synthetic images (32 images x 8192
descriptors x 768 bf16 = 403 MB > L2, 496 pairs).  Prints ONE JSON line (see README / DESIGN.md).
`--impl reference` times the CPU restatement of the reference matcher (oracle/) on the host cores.
"""
from __future__ import annotations

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

METRIC = "matched_pairs_per_sec@8192kpts_d768"
UNIT = "pairs/s"
N_IMAGES, KPTS, DIM = 32, 8192, 768
FLOP_PER_PAIR = 2.0 * KPTS * KPTS * DIM  # 2*N*M*d
# dram__bytes_read.sum + dram__bytes_write.sum of one match_fused2_kernel launch (496 pairs), ncu --set full
# (profiles/r02c_match_fused2_ncu_summary.json: the shipped kernel with sixteen epilogue warps)
NCU_TRAFFIC_BYTES = 7.056956e9 + 6.324735e9


def _peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d.get("bf16_tflops_sustained", 1400.0)), float(d.get("hbm_gbs", 6650.0)), "measured"
    return 1400.0, 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_descriptors_device(torch, dev, n_images=N_IMAGES, kpts=KPTS, dim=DIM, seed=0):
    """SURVEY 8(d) C4/C5 recipe, generated on the device: unit-norm rows; image j shares a random
    half of its rows with image j-1, perturbed and re-normalised.  The perturbation has norm 0.35
    (sigma = 0.35/sqrt(d) per component; SURVEY's 0.05 per component would be norm 1.39 at d=768 and
    leave no pair inside max_distance), so shared rows sit at ~0.34 rad and pass the ratio test."""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    out = torch.empty((n_images, kpts, dim), dtype=torch.bfloat16, device=dev)
    prev = None
    for j in range(n_images):
        x = torch.randn((kpts, dim), generator=g, device=dev, dtype=torch.float32)
        x = x / x.norm(dim=1, keepdim=True)
        if prev is not None:
            m = kpts // 2
            src = torch.randperm(kpts, generator=g, device=dev)[:m]
            dst = torch.randperm(kpts, generator=g, device=dev)[:m]
            y = prev[src] + (0.35 / dim ** 0.5) * torch.randn((m, dim), generator=g, device=dev)
            x[dst] = y / y.norm(dim=1, keepdim=True)
        out[j] = x.to(torch.bfloat16)
        prev = x
    return out


def bench_config():
    """The workload both arms (--impl ours / reference) report: identical dict, static keys only."""
    n_pairs = N_IMAGES * (N_IMAGES - 1) // 2
    return {"workload": f"exhaustive matching of one resident block: {N_IMAGES} images x {KPTS} kpts x d={DIM} "
                        f"({n_pairs} pairs/step/GPU), ratio 0.8, distance 0.7, cross-check",
            "kpts": KPTS, "d": DIM, "pairs_per_step_per_gpu": n_pairs,
            "l2": "inputs 403 MB > 126 MB L2"}


def make_descriptors_nonneg(torch, dev, n_images=N_IMAGES, kpts=KPTS, dim=DIM, seed=0):
    """Descriptors shaped like the product path's uint8 rows (vit_extractor.py:243-250: L2-normalise,
    x512, clip negatives to 0): |unit rows| re-normalised, so unrelated rows have cosine ~ 2/pi = 0.64
    instead of ~0 -- at the default options that is AT the pruning bound (cos(0.7/0.8) = 0.64), i.e.
    the bound removes about half of the unrelated scores instead of all of them."""
    x = make_descriptors_device(torch, dev, n_images, kpts, dim, seed).float().abs()
    return (x / x.norm(dim=2, keepdim=True)).to(torch.bfloat16)


def all_pairs(n):
    return np.array([(i, j) for i in range(n) for j in range(i + 1, n)], dtype=np.int32)


# ------------------------------------------------------------------------------------------- CPU
def _host_cores() -> int:
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def cpu_sample(seconds_target: float = 12.0, seed: int = 0):
    """Reference CPU path (oracle port of COLMAP/feature/sift.cc:733-825,922-971) on a bounded
    sample of the same workload: R rows of image 1 against all 8192 rows of image 2, d=768,
    float32 SIMD dot products, OpenMP over all host threads."""
    from oracle import matching as om
    om.set_num_threads(_host_cores())  # launchers such as torchrun export OMP_NUM_THREADS=1
    rng = np.random.default_rng(seed)
    a = rng.standard_normal((KPTS, DIM)).astype(np.float32)
    a /= np.linalg.norm(a, axis=1, keepdims=True)
    b = a[rng.permutation(KPTS)] + (0.35 / DIM ** 0.5) * rng.standard_normal((KPTS, DIM)).astype(np.float32)
    b /= np.linalg.norm(b, axis=1, keepdims=True)
    om.match_c(a[:256], b, fast_float=True)  # warm the thread pool
    n_done, t0 = 0, time.perf_counter()
    while True:
        om.match_c(a, b, fast_float=True)
        n_done += 1
        dt = time.perf_counter() - t0
        if dt >= seconds_target or n_done >= 64:
            break
    return n_done / dt, n_done, dt, om.num_threads()


def run_reference(args):
    """--impl reference: the reference's own algorithm on the host cores (oracle port: COLMAP cannot
    be built in this image).  Each step is a bounded sample; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host core the box has
    import torch
    torch.set_num_threads(_host_cores())
    vals, rows_used = [], 0
    per_step = max(2.0, min(12.0, 90.0 / max(args.steps + args.warmup, 1)))
    for i in range(args.warmup + args.steps):
        v, rows, dt, threads = cpu_sample(per_step, seed=i)
        if i >= args.warmup:
            vals.append(v); rows_used = rows
    value = float(np.mean(vals))
    sample = f"{rows_used} full 8192x8192 d=768 pairs per step, float32 SIMD dot, OpenMP over all host threads"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 / value if value else None,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": bench_config(),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if not args.skip_extract:
        ev, en, edt, eth = cpu_extract_sample(2)
        line["extract"] = {"metric": "extracted_images_per_sec@vitb14_518px", "value": ev, "unit": "images/s",
                           "cpu_baseline": {"value": ev, "unit": "images/s", "cores": eth, "kind": "port",
                                            "sample": f"{en} images ({edt:.1f} s), fp32 PyTorch DINOv2 restatement + "
                                                      "reference post-processing restatement"}}
    return line


# ------------------------------------------------------------------------------------------- GPU
def run_ours(args):
    import torch
    import torch.distributed as dist
    from vit_colmap_b200 import _lib
    from vit_colmap_b200.matching import Matcher

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the product has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    matcher = Matcher(local)
    allgather_ms = None
    if world > 1:
        # north-star exchange: every rank owns N_IMAGES/world images; one NCCL all-gather makes the
        # block resident everywhere.  Timed once (it happens once per job, not per pair batch).
        per = N_IMAGES // world
        full = make_descriptors_device(torch, dev, N_IMAGES, KPTS, DIM, seed=0)
        shard = full[rank * per:(rank + 1) * per].contiguous()
        desc = torch.empty_like(full)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); dist.all_gather_into_tensor(desc, shard); e1.record(); torch.cuda.synchronize()
        allgather_ms = e0.elapsed_time(e1)
        assert torch.equal(desc, full)
        del full
    else:
        desc = make_descriptors_device(torch, dev, N_IMAGES, KPTS, DIM, seed=0)
    counts = torch.full((N_IMAGES,), KPTS, dtype=torch.int32, device=dev)
    pairs_np = all_pairs(N_IMAGES)
    # weak scaling: every rank matches one full 496-pair block per step; in a real job the ranks
    # hold different pair blocks (vit_colmap_b200.pairing.shard_blocks), here the same synthetic one
    pairs = torch.from_numpy(pairs_np).to(dev)
    n_pairs = len(pairs_np)
    matches = torch.empty((n_pairs, KPTS, 2), dtype=torch.int32, device=dev)
    n_matches = torch.empty((n_pairs,), dtype=torch.int32, device=dev)
    opts = dict(max_ratio=0.8, max_distance=0.7, cross_check=True, inv_norm=1.0)

    def step():
        matcher.match_pairs(desc, counts, pairs, out=(matches, n_matches), **opts)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    sampler = ClockSampler(local); sampler.start()
    _lib.launch_count(reset=True)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    launches = _lib.launch_count()
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * n_pairs * args.steps / (ms_total / 1e3)

    # dominant kernel alone (top-2 stage only), CUDA events on the launching stream
    ws, need = matcher._workspace(n_pairs, KPTS)
    import ctypes
    opt = matcher._opts(0.8, 0.7, True, 1.0)

    def main_kernel_only():
        _lib.check(matcher.lib.vc_match_pairs(matcher.handle, desc.data_ptr(), counts.data_ptr(), N_IMAGES,
                                              KPTS, DIM, pairs.data_ptr(), n_pairs, ctypes.byref(opt), None,
                                              None, ws.data_ptr(), need, matcher._stream()))
    main_kernel_only(); torch.cuda.synchronize()
    k0, k1 = torch.cuda.Event(True), torch.cuda.Event(True)
    reps = max(2, min(args.steps, 5))
    k0.record()
    for _ in range(reps):
        main_kernel_only()
    k1.record(); torch.cuda.synchronize()
    clocks = sampler.stop()
    kern_ms = k0.elapsed_time(k1) / reps
    peak_tf, peak_hbm, peak_src = _peaks()
    achieved_tf = FLOP_PER_PAIR * n_pairs / (kern_ms / 1e3) / 1e12

    mean_matches = float(n_matches.float().mean().item())

    # The pruning bound is data dependent: the same stage (a) with the bound switched off,
    # (b) on non-negative rows shaped like the product path's uint8 descriptors, with and without it.
    def time_variant(d_tensor, no_prune):
        if no_prune:
            os.environ["VC_MATCH_NO_PRUNE"] = "1"
        try:
            def run():
                _lib.check(matcher.lib.vc_match_pairs(matcher.handle, d_tensor.data_ptr(), counts.data_ptr(), N_IMAGES,
                                                      KPTS, DIM, pairs.data_ptr(), n_pairs, ctypes.byref(opt), None,
                                                      None, ws.data_ptr(), need, matcher._stream()))
            run(); run(); torch.cuda.synchronize()
            v0, v1 = torch.cuda.Event(True), torch.cuda.Event(True)
            v0.record()
            for _ in range(reps):
                run()
            v1.record(); torch.cuda.synchronize()
        finally:
            os.environ.pop("VC_MATCH_NO_PRUNE", None)
        v_ms = v0.elapsed_time(v1) / reps
        v_tf = FLOP_PER_PAIR * n_pairs / (v_ms / 1e3) / 1e12
        return {"kernel_ms": v_ms, "pairs_per_s": n_pairs / (v_ms / 1e3), "achieved_tflops": v_tf, "frac": v_tf / peak_tf}

    variants = None
    if rank == 0 and not args.skip_variants:
        nonneg = make_descriptors_nonneg(torch, dev, N_IMAGES, KPTS, DIM, seed=0)
        variants = {"signed_unit_rows_no_prune": time_variant(desc, True),
                    "nonneg_rows": time_variant(nonneg, False),
                    "nonneg_rows_no_prune": time_variant(nonneg, True),
                    "note": "same top-2 stage as roofline; nonneg = |unit rows| renormalised (cosine of unrelated rows ~0.64, "
                            "like clipped uint8 ViT descriptors); no_prune = VC_MATCH_NO_PRUNE=1"}
        del nonneg
    if world > 1:
        dist.barrier()

    # end to end through the host-buffer C entry: pinned bf16 descriptors in, matches out
    e2e = None
    if rank == 0 or world > 1:
        host = torch.empty(desc.shape, dtype=torch.bfloat16).pin_memory()
        host.copy_(desc)
        torch.cuda.synchronize()
        host_np = host.view(torch.int16).numpy().view(np.uint16)
        cnt_np = np.full((N_IMAGES,), KPTS, np.int32)
        out_m = torch.empty((n_pairs, KPTS, 2), dtype=torch.int32).pin_memory()
        out_n = torch.empty((n_pairs,), dtype=torch.int32).pin_memory()
        opt_h = matcher._opts(0.8, 0.7, True, 1.0)

        def e2e_step():
            _lib.check(matcher.lib.vc_match_pairs_host(
                matcher.handle, host_np.ctypes.data, _lib.DTYPE_BF16, cnt_np.ctypes.data, N_IMAGES, KPTS, DIM,
                pairs_np.ctypes.data, n_pairs, ctypes.byref(opt_h), out_m.data_ptr(), out_n.data_ptr()))
        e2e_step()
        barrier()
        e2e_steps = max(1, min(args.steps, 3))
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        n_out = int(out_n.sum().item())
        e2e = {"value": world * n_pairs * e2e_steps / float(tt.item()), "unit": UNIT,
               "h2d_bytes_per_step": int(host.numel() * 2 + cnt_np.nbytes + pairs_np.nbytes),
               "d2h_bytes_per_step": int(n_out * 8 + n_pairs * 4),
               "api": "vc_match_pairs_host (pinned bf16 host descriptors -> host match lists)"}

    _lib.launch_count(reset=True)
    extract = run_extract(args, torch, dist, dev, local, world) if not args.skip_extract else None
    extract_launches = _lib.launch_count()
    extract_tr = run_extract_trainable(args, torch, dist, dev, local, world) if not args.skip_extract else None
    if rank == 0:
        cpu_v, cpu_rows, cpu_dt, cpu_threads = (cpu_sample(12.0) if world == 1 else (None, 0, 0, 0))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": bench_config(),
            "observed": {"mean_matches_per_pair": mean_matches},
            "e2e": e2e, "gpu_launches": int(launches),
            "roofline": {"bound": "tensor", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": achieved_tf / peak_tf, "traffic": NCU_TRAFFIC_BYTES,
                         "peak_source": f"{peak_src} sustained bf16",
                         "kernel": "match_fused2_kernel (two-SM cta_group::2 form of match_fused_kernel)", "kernel_ms": kern_ms,
                         "note": "achieved = 2*N*M*d per pair (one MMA sweep gives row and column top-2) "
                                 "over the top-2 stage (memset + fused kernel + row / column partial merges); traffic "
                                 "from profiles/r02c_match_fused2_ncu_summary.json (dram read+write of one launch of the "
                                 "shipped kernel; tensor pipe 82.7 % active in that capture)"},
            "clocks": clocks,
        }
        if extract is not None:
            extract["gpu_launches"] = int(extract_launches)
            if world == 1:
                ev, en, edt, eth = cpu_extract_sample(2)
                extract["cpu_baseline"] = {"value": ev, "unit": "images/s", "cores": eth, "kind": "port",
                                           "sample": f"{en} images ({edt:.1f} s), fp32 PyTorch DINOv2 restatement + "
                                                     "reference post-processing restatement"}
            line["extract"] = extract
            line["extract_trainable"] = extract_tr
        if variants is not None:
            line["roofline_variants"] = variants
        if allgather_ms is not None:
            line["allgather_ms"] = allgather_ms
        if cpu_v is not None:
            line["cpu_baseline"] = {"value": cpu_v, "unit": UNIT, "cores": cpu_threads, "kind": "port",
                                    "sample": f"{cpu_rows} full 8192x8192 d=768 pairs ({cpu_dt:.1f} s), float32 SIMD dot, OpenMP"}
        if world == 1 and not args.skip_variants:
            line["match_exhaustive_db"] = run_db_e2e(torch, local)
    if world > 1:
        dist.destroy_process_group()
    return line if rank == 0 else None


# ------------------------------------------------------------------------------------ extraction
EXTRACT_MODEL, EXTRACT_SIZE = "dinov2_vitb14", 518
EXTRACT_BATCH = int(os.environ.get("VC_BENCH_EXTRACT_BATCH", "32"))
EXTRACT_KPTS, EXTRACT_DIM = 2048, 128  # reference defaults, vit_extractor.py:25-33


def extract_flops_per_image(name=EXTRACT_MODEL, size=EXTRACT_SIZE):
    """L*(24 T C^2 + 4 T^2 C) + 2 HW 588 C  (303.1 GFLOP for ViT-B/14 at 518 px)."""
    C, L = {"dinov2_vits14": (384, 12), "dinov2_vitb14": (768, 12), "dinov2_vitl14": (1024, 24)}[name]
    hw = (size // 14) ** 2
    T = hw + 1
    return L * (24.0 * T * C * C + 4.0 * T * T * C) + 2.0 * hw * 588 * C


def run_extract(args, torch, dist, dev, local, world):
    """Second half of BASELINE.json's metric: extracted images/s, ViT-B/14 at 518x518, through the
    `vit` extractor (DINOv2 forward + score map + selection + sampling / PCA / packing).  Synthetic
    images and seeded random-init weights (no checkpoint can be fetched here)."""
    from vit_colmap_b200.features.vit_extractor import ViTExtractor
    ex = ViTExtractor(model_name=EXTRACT_MODEL, num_keypoints=EXTRACT_KPTS, descriptor_dim=EXTRACT_DIM,
                      device=f"cuda:{local}", synthetic_seed=0, batch_size=EXTRACT_BATCH)
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    imgs = torch.randint(0, 256, (EXTRACT_BATCH, EXTRACT_SIZE, EXTRACT_SIZE, 3), generator=g, device=dev,
                         dtype=torch.uint8)
    gh = EXTRACT_SIZE // 14
    wh = (EXTRACT_SIZE, EXTRACT_SIZE)
    host = [im.cpu().numpy() for im in imgs]
    ex._extract_batch(host[:1], wh, wh)  # fits the PCA projection on the first image (reference semantics)

    def step():
        tokens = ex.model.forward_patch_tokens(imgs, bgr=True)
        return ex.post.dense_to_sparse(tokens, (gh, gh), wh, wh, EXTRACT_KPTS, "harris",
                                       projection=ex.descriptor_projection)

    def forward_only():
        ex.model.forward_patch_tokens(imgs, bgr=True)

    def timed(fn, n):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(max(args.warmup, 3)):
        out = step()
    # three windows of K steps, median: this section starts right after the matcher has held the board at its
    # power cap, and a single 120 ms window then swings by +- 8 % with the clock recovery
    ms = sorted(timed(step, args.steps) for _ in range(3))[1]
    ms_fwd = timed(forward_only, max(2, min(args.steps, 5))) / max(2, min(args.steps, 5))
    # end to end: host uint8 images in, host keypoints + descriptors out (public extractor API)
    list(ex.extract_batches([(host, wh, wh, None)] * 4))  # warm-up: allocator pools for two batches in flight
    torch.cuda.synchronize()
    n_e2e = max(4, min(2 * args.steps, 20))
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    res = None
    ex.pipe_stats.update(stage_s=0.0, launch_s=0.0, finish_s=0.0, batches=0)
    for _, res, _ in ex.extract_batches([(host, wh, wh, None)] * n_e2e):  # every batch: staged, uploaded, read back
        assert res is not None
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    kp_bytes = int(ex.last_d2h_bytes)  # whole [B, cap, .] output buffers + counts, as copied
    peak_tf, _, peak_src = _peaks()
    tf = extract_flops_per_image() * EXTRACT_BATCH / (ms_fwd / 1e3) / 1e12
    return {
        "metric": "extracted_images_per_sec@vitb14_518px", "value": world * EXTRACT_BATCH * args.steps / (ms / 1e3),
        "unit": "images/s", "ms_per_step": ms / args.steps, "dtype": "bf16",
        "config": {"workload": f"{EXTRACT_MODEL} forward + harris score + bin/NMS selection + PCA {EXTRACT_DIM}-d "
                               f"descriptors, batch {EXTRACT_BATCH} x {EXTRACT_SIZE}x{EXTRACT_SIZE} per GPU per step",
                   "mean_keypoints_per_image": float(out.counts.float().mean().item()),
                   "timing": "median of three windows of K steps (CUDA events, max over ranks)",
                   "l2": "activations (134 MB residual stream, 269 MB qkv/MLP per batch) exceed the 126 MB L2"},
        "e2e": {"value": world * EXTRACT_BATCH * n_e2e / float(dt.item()), "unit": "images/s",
                "h2d_bytes_per_step": int(imgs.numel()), "d2h_bytes_per_step": kp_bytes,
                "host_ms_per_step": {k: round(1e3 * v / n_e2e, 3) for k, v in ex.pipe_stats.items() if k != "batches"},
                "api": "ViTExtractor.extract_batches (host BGR uint8 images -> host keypoints f32 + descriptors u8; "
                       "the pipeline extract() runs: two batches queued on the device, staging / upload / read-back on "
                       "their own streams)"},
        "roofline": {"bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
                     "traffic": None, "peak_source": f"{peak_src} sustained bf16", "kernel": "vc_vit_forward (all launches)",
                     "kernel_ms": ms_fwd,
                     "note": "algorithmic 303.1 GFLOP per image (SURVEY 8d) x 32 images over the whole forward; "
                             "per-kernel shares in profiles/"},
    }


def run_extract_trainable(args, torch, dist, dev, local, world):
    """Config C4's extraction side: the `trainable_vit` extractor (DINOv2 ViT-B/14 backbone +
    ViTFeatureModel conv head + max-pool NMS selection + sparse descriptor projection) at 518x518,
    d = 128, num_keypoints 8192.  Synthetic images, seeded random-init weights."""
    from vit_colmap_b200.features.trainable_vit_extractor import TrainableViTExtractor
    B = 16
    ex = TrainableViTExtractor(model_name=EXTRACT_MODEL, num_keypoints=8192, descriptor_dim=128,
                               device=f"cuda:{local}", score_threshold=0.0, nms_radius=1, synthetic_seed=0,
                               batch_size=B)
    g = torch.Generator(device=dev)
    g.manual_seed(2)
    imgs = torch.randint(0, 256, (B, EXTRACT_SIZE, EXTRACT_SIZE, 3), generator=g, device=dev, dtype=torch.uint8)
    gh = EXTRACT_SIZE // 14
    wh = (EXTRACT_SIZE, EXTRACT_SIZE)
    tokens = ex.backbone.forward_patch_tokens(imgs, bgr=True)

    def timed(fn, n):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / n], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = max(2, min(args.steps, 5))
    ms = timed(lambda: ex._forward_device(imgs, wh, wh), n)
    ms_head = timed(lambda: ex.head.forward(tokens, (gh, gh), dense_descriptors=False), n)
    counts = ex._forward_device(imgs, wh, wh)[4]
    host = [im.cpu().numpy() for im in imgs]
    list(ex.extract_batches([(host, wh, wh, None)] * 4))  # warm-up: allocator pools for two batches in flight
    torch.cuda.synchronize()
    n_e2e = max(4, min(2 * args.steps, 20))
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    res = None
    ex.pipe_stats.update(stage_s=0.0, launch_s=0.0, finish_s=0.0, batches=0)
    for _, res, _ in ex.extract_batches([(host, wh, wh, None)] * n_e2e):
        assert res is not None
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e = {"value": world * B * n_e2e / float(dt.item()), "unit": "images/s", "h2d_bytes_per_step": int(imgs.numel()),
           "d2h_bytes_per_step": int(ex.last_d2h_bytes),
           "host_ms_per_step": {k: round(1e3 * v / n_e2e, 3) for k, v in ex.pipe_stats.items() if k != "batches"},
           "api": "TrainableViTExtractor.extract_batches (host BGR uint8 images -> host keypoints [n,6] f32 + "
                  "descriptors u8)"}
    peak_tf, _, peak_src = _peaks()
    tf = ex.head.flops_per_image(gh, gh) * B / (ms_head / 1e3) / 1e12
    return {"metric": "extracted_images_per_sec@trainable_vitb14_518px_d128", "value": world * B / (ms / 1e3),
            "unit": "images/s", "ms_per_step": ms, "dtype": "bf16",
            "config": {"workload": f"{EXTRACT_MODEL} forward + conv head + NMS/top-8192 selection + sparse 128-d "
                                   f"descriptors, batch {B} x {EXTRACT_SIZE}x{EXTRACT_SIZE} per GPU per step",
                       "mean_keypoints_per_image": float(counts.float().mean().item())},
            "e2e": e2e,
            "roofline": {"bound": "tensor", "achieved": tf, "peak": peak_tf, "unit": "TFLOP/s", "frac": tf / peak_tf,
                         "traffic": None, "peak_source": f"{peak_src} sustained bf16",
                         "kernel": "vc_conv_head_forward (conv_gemm_kernel launches)", "kernel_ms": ms_head,
                         "note": f"algorithmic {ex.head.flops_per_image(gh, gh) / 1e9:.1f} GFLOP per image for the "
                                 "head (2*MACs of the seven convolutions, SURVEY 8a T1)"}}


def cpu_extract_sample(n_images=2):
    """Reference CPU path of the extractor: fp32 PyTorch DINOv2 restatement + the reference's
    post-processing restatement (oracle/), ViT-B/14 at 518 px, all host threads."""
    import torch
    from oracle import extract_oracle as eo
    from oracle import vit_oracle as vo
    torch.set_num_threads(_host_cores())
    sd = vo.synthetic_state_dict(EXTRACT_MODEL, 0)
    rng = np.random.default_rng(0)
    t0 = time.perf_counter()
    for i in range(n_images):
        img = torch.from_numpy(rng.integers(0, 256, (1, EXTRACT_SIZE, EXTRACT_SIZE, 3), dtype=np.uint8))
        tok = vo.forward_features(sd, vo.preprocess(img), vo.CONFIGS[EXTRACT_MODEL][2])["x_norm_patchtokens"][0]
        g = EXTRACT_SIZE // 14
        fm = tok.reshape(g, g, -1).permute(2, 0, 1).contiguous()
        eo.dense_to_sparse(fm, (EXTRACT_SIZE,) * 2, (EXTRACT_SIZE,) * 2, EXTRACT_KPTS, EXTRACT_DIM, "harris")
    dt = time.perf_counter() - t0
    return n_images / dt, n_images, dt, torch.get_num_threads()


def run_db_e2e(torch, local):
    """Rank 0 of SURVEY 8(f): the callers either side of the matcher -- `match_exhaustive` over a COLMAP
    database (sqlite3 reads of the descriptor blobs, ONE upload, pair chunks on the resident set, batched
    write of the `matches` rows).  64 images x 2048 SIFT-like uint8 descriptors (d = 128), 2016 pairs."""
    import tempfile
    from vit_colmap_b200.database import ColmapDatabase
    from vit_colmap_b200.matching import FeatureMatchingOptions, Matcher, match_exhaustive
    n_images, k, d = 64, 2048, 128
    rng = np.random.default_rng(5)
    base = rng.standard_normal((4 * k, d)).astype(np.float32)
    with tempfile.TemporaryDirectory() as tmp:
        def make_db(path):
            db = ColmapDatabase(path)
            cam = db.add_pinhole_camera(640, 480, 500.0, 500.0, 320.0, 240.0)
            for i in range(n_images):
                rows = np.abs(base[rng.permutation(4 * k)[:k]] + 0.15 * rng.standard_normal((k, d)).astype(np.float32))
                image_id = db.add_image(f"im{i:03d}.png", camera_id=cam)
                db.add_keypoints(image_id, rng.uniform(0, 480, (k, 2)).astype(np.float32))
                unit = rows / np.linalg.norm(rows, axis=1, keepdims=True)
                # COLMAP's quantiser, min(255, round(512 x)) (COLMAP/feature/utils.cc:56-71): input construction only
                db.add_descriptors(image_id, np.clip(np.round(512.0 * unit), 0, 255).astype(np.uint8))
            db.close()
            return path
        matcher = Matcher(local)
        opts = FeatureMatchingOptions(max_ratio=0.85, max_distance=0.9)
        # first database: warm-up (device staging, page-locked result buffers are allocated once per Matcher)
        match_exhaustive(make_db(os.path.join(tmp, "warm.db")), opts, matcher=matcher)
        path = make_db(os.path.join(tmp, "bench.db"))
        torch.cuda.synchronize()
        stages = {}
        t0 = time.perf_counter()
        res = match_exhaustive(path, opts, matcher=matcher, stats=stages)
        dt = time.perf_counter() - t0
        n_rows = sum(len(m) for m in res.values())
        matcher.close()
    return {"metric": "matched_pairs_per_sec@match_exhaustive_db_2048kpts_d128", "value": len(res) / dt, "unit": "pairs/s",
            "seconds": dt, "stage_seconds": {k: round(v, 4) for k, v in stages.items()}, "pairs": len(res), "match_rows_written": int(n_rows),
            "h2d_bytes": n_images * k * d, "api": "vit_colmap_b200.matching.match_exhaustive(database_path) -- sqlite3 "
            "reads, one upload, vc_match_pairs_resident per pair chunk, one write transaction"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--skip-extract", action="store_true", help="matcher half of the metric only")
    ap.add_argument("--skip-variants", action="store_true", help="skip the unpruned / non-negative-data matcher timings")
    args = ap.parse_args()
    # stdout carries the ONE JSON line; the extractors' progress prints go to stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(os.dup(2), "w")
    out = os.fdopen(real_stdout, "w")
    line = run_reference(args) if args.impl == "reference" else run_ours(args)
    if line is not None:
        out.write(json.dumps(line) + "\n")
        out.flush()


if __name__ == "__main__":
    main()

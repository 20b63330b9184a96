#!/usr/bin/env python
"""bench.py — KR sample-pairs/s of the B200 path (and of the reference CPU path, --impl reference).

One "step" = one pass of the hot path over the workload's mass trees:
  value : masses resident in HBM -> F build (K1) + pair kernel (K2) + epilogue [+ exchange] + D2H -> matrix complete in HOST memory
  e2e   : HOST (pinned) mass trees -> krd_set_masses_binned (H2D + K0) + compute -> matrix in HOST memory, through the C ABI
Workloads: --gpus 1 runs cfg2 (BASELINE.json configs[1], the "1 B200" config) and adds a short 1-GPU measurement of cfg4
(`strong_scaling_base`); --gpus N > 1 runs cfg4 (the north-star's multi-GPU config: 10,000 samples x 49,999 edges,
100,000 placements per sample, bins 1, p = 1) with the cohort FIXED: strong scaling.
Launch: `python bench.py --gpus 1 ...` or under torchrun (one rank per GPU, NCCL).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "KR sample-pairs/sec (p=1, FP64)"   # BASELINE.json's metric; other exponents (--workload cfg3 / cfg5, --p) say theirs


def metric_of(spec):
    return METRIC if float(spec["p"]) == 1.0 else "KR sample-pairs/sec (p=%g, FP64)" % float(spec["p"])
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("KRD_WORKLOAD"))
    ap.add_argument("--samples", type=int)
    ap.add_argument("--edges", type=int)
    ap.add_argument("--placements", type=int)
    ap.add_argument("--bins", type=int)
    ap.add_argument("--p", type=float)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-base", action="store_true", help="N = 1: skip the 1-GPU cfg4 block (strong_scaling_base)")
    ap.add_argument("--exchange", default="p2p_rows", choices=["host_tiles", "p2p_rows", "allgather_rows", "allgather_rank0", "own_tiles_host"],
                    help="N > 1: how the result tiles reach host memory (SURVEY §8e asks for an A/B)")
    ap.add_argument("--late", type=float, default=0.0625, help="--exchange p2p_rows: fraction of the sample blocks treated as 'late' (krd_set_late_blocks): "
                    "the rectangle outside them is copied to the host under the pair launch of the tiles that touch them (0 = no overlap)")
    ap.add_argument("--tail", type=float, default=0.125, help="--exchange host_tiles: fraction of a rank's tiles in the short LAST launch; the copies "
                    "of the long first launch run under it (0 = one launch)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = dict(sm_mhz=None, sm_max_mhz=None, reasons=[], samples=0)
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f:
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0])); mx.append(float(c[1])); pw.append(float(c[2]))
            except ValueError:
                continue
            for nme, v in zip(names, c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        self.f.close(); os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), power_w_max=float(max(pw)),
                       reasons=sorted(reasons), samples=len(sm))
        return out


def config_of(spec, world):
    """The workload as BOTH arms print it (identical keys and values: derived from the spec alone)."""
    N, E, B = spec["samples"], spec["edges"], spec["bins"]
    S = E * (B + 1)
    ld = (N + 127) // 128 * 128
    return {"workload": spec["name"], "samples": N, "edges": E, "placements_per_sample": spec["placements"], "bins": B,
            "p": spec["p"], "segments": S, "pairs": N * (N - 1) // 2,
            "l2": "inputs larger than L2 (dense matrix F = %.0f MB + split workspace)" % ((S + 15) // 16 * 16 * ld * 8 / 1e6),
            "parallelism": "1 gpu" if world == 1 else
                           f"strong scaling: cohort fixed, upper-triangular 128x128 tiles / {world} ranks, F replicated, NCCL all-gather "
                           "of the result tiles or stores into the owners' row blocks over NVLink peer memory (see `exchange`), every rank copies its row block of the matrix to host memory"}


def traffic_from_profile(spec, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of the pair kernel per launch, from the committed
    `ncu --set full` capture of this workload (profiles/pair_kernel_traffic.json); None when there is none."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "pair_kernel_traffic.json")))
        key = "%s_p%g" % (spec["name"], spec["p"])
        return t[key]["bytes_per_launch"] if world == 1 and spec["samples"] == t[key]["samples"] else None
    except (OSError, KeyError, ValueError):
        return None


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference(tree, masses, p, target_seconds, n_fixed=None):
    """Times the reference's own CPU implementation (oracle/_ref = compiled genesis, OpenMP over ALL host threads —
    the count is passed explicitly, torchrun exports OMP_NUM_THREADS=1) — or the C port when _ref is absent — on a
    bounded sample: the first n samples of the same (binified) mass trees.
    Returns (dict(value pairs/s, cores, kind, sample), matrix of those n samples)."""
    import oracle
    from gappa_b200 import workloads
    use_ref = oracle.have_ref()
    kind = "reference" if use_ref else "port"
    cores = host_threads()

    def run(n):
        sub = workloads.head(masses, n)
        t0 = time.perf_counter()
        if use_ref:
            D, sec = oracle.ref_matrix(tree, sub, p, bins=0, threads=cores)   # masses are already binified
        else:
            D = oracle.port_matrix(tree, sub, p, bins=0, threads=cores)
            sec = time.perf_counter() - t0
        return n * (n - 1) / 2 / sec, sec, D

    N = len(masses["off"]) - 1
    n0 = min(N, n_fixed if n_fixed else max(2, 2 * int(np.sqrt(cores)) + 8))
    rate, sec, D = run(n0)
    if not n_fixed:
        n1 = int(min(N, max(n0, np.sqrt(2 * target_seconds * rate))))
        if n1 > n0:
            rate, sec, D = run(n1)
            n0 = n1
    return dict(value=rate, unit=UNIT, cores=int(cores), kind=kind,
                sample=f"first {n0} of {N} samples = {n0 * (n0 - 1) // 2} pairs, {sec:.2f} s, earth_movers_distance(trees,p) only"), D


def parity_of(blocks, D_ref):
    """max relative error of the timed results' leading block against the reference matrix of the same samples."""
    n = D_ref.shape[0]
    off = ~np.eye(n, dtype=bool)
    out = {"vs": "reference (compiled genesis)" if __import__("oracle").have_ref() else "port", "pairs_checked": n * (n - 1) // 2}
    worst = 0.0
    for name, blk in blocks.items():
        err = float(np.max(np.abs(blk[:n, :n][off] - D_ref[off]) / np.maximum(np.abs(D_ref[off]), 1e-300)))
        out["max_rel_err_" + name] = err
        worst = max(worst, err)
    out["max_rel_err"] = worst
    return out


def default_workload(args, world):
    return args.workload or ("cfg2" if world == 1 else "cfg4")


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores, same config/metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from gappa_b200 import workloads
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    name = default_workload(args, world)
    spec = workloads.spec_of(name, samples=args.samples, edges=args.edges, placements=args.placements, bins=args.bins, p=args.p)
    # a bounded sample: only as many samples as the CPU can finish in a few seconds per step
    n_gen = min(spec["samples"], 160)
    _, tree, masses = workloads.build(name, n_samples=n_gen, **{k: spec[k] for k in ("edges", "placements", "bins", "p")})
    rates, info = [], None
    for it in range(args.warmup + args.steps):
        info, _ = cpu_reference(tree, masses, spec["p"], target_seconds=max(1.0, args.cpu_seconds / max(args.steps, 1)))
        if it >= args.warmup:
            rates.append(info["value"])
    value = float(np.mean(rates))
    info["value"] = value
    pairs_total = spec["samples"] * (spec["samples"] - 1) / 2
    print(json.dumps({
        "impl": "reference", "metric": metric_of(spec), "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": pairs_total / value * 1e3,
        "higher_is_better": True, "scaling": "weak" if world == 1 else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(spec, world),
        "note": "ms_per_step extrapolated linearly in pairs from the bounded sample; pairs/s of the CPU path does not depend on the cohort size",
        "cpu_baseline": info,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except (OSError, KeyError, ValueError):
        return 6550.0, "fallback (B200_PROFILING.md copy bandwidth)"


def rooflines(gappa_b200, local, spec, st, pairs, share, ms_step, world):
    """Roofline of the dominant kernel (k_pairs: FP64 pipe) and of the F builder (K1: HBM)."""
    S = st["n_segments"]
    p = spec["p"]
    # 3 flop per (pair, segment) for the direct operator (subtract, multiply / abs, accumulate).  p = 2 on the
    # Gram path contracts len * (a - c)(b - c) on the FP64 tensor cores: ONE fused multiply-add = 2 flop per (pair, segment)
    # (operands centred and scaled once by k_gram_prescale); counting 3 would put the kernel above the DMMA peak.
    gram = (p == 2.0 and int(st.get("gram_fixups", -1)) >= 0 and spec["bins"] > 0)
    per = 2.0 if gram else 3.0
    flops = per * S * pairs * share
    t_pairs = st["ms_pairs"] * 1e-3
    peak_dfma = gappa_b200.measure_fp64_peak(local, 0, 200.0)
    peak_mix = gappa_b200.measure_fp64_peak(local, 1, 200.0)
    peak_dmma = gappa_b200.measure_fp64_peak(local, 2, 200.0)
    peak = max(peak_dfma, peak_dmma)
    achieved = flops / t_pairs / 1e12
    # executed: full tiles + 40/64 of each diagonal tile + qmax/8 of the off-diagonal tiles of a ragged last column block
    # (the short operators skip at 16-column granularity)
    T = st["ld_samples"] // 128
    rag = spec["samples"] % 128
    qmax = (rag + 15) // 16 if rag else 8
    off_tiles = st["n_tiles"] - T
    rag_tiles = (T - 1) if qmax < 8 else 0
    executed = 3.0 * ((S + 15) // 16 * 16) * 128.0 * 128.0 * ((off_tiles - rag_tiles) + rag_tiles * qmax / 8.0 + 0.625 * T) * share
    if gram:      # every tile in full, both triangles of the diagonal ones
        executed = 2.0 * ((S + 15) // 16 * 16) * 128.0 * 128.0 * st["n_tiles"] * share
        peak = peak_dmma
    roof = {"bound": "fp64", "kernel": ("k_pairs<p=2, Gram form on DMMA>" if gram else "k_pairs<p=%g>" % p), "flop_per_pair_segment": per,
            "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
            "frac": achieved / peak, "traffic": traffic_from_profile(spec, world),
            "peak_source": "measured live: max(DFMA loop %.1f, DMMA m8n8k4 loop %.1f) TFLOP/s; MEASURED_PEAKS.json has no FP64 entry"
                           % (peak_dfma, peak_dmma),
            "mix_ceiling": None if gram else peak_mix, "frac_of_mix_ceiling": None if gram else achieved / peak_mix,   # (DADD+DFMA mix: direct operators only)
            "executed_tflops": executed / t_pairs / 1e12,
            "ms_kernel": st["ms_pairs"], "ms_build_F": st["ms_build"], "ms_epilogue": st["ms_epilogue"],
            "ctas": st["pair_ctas"], "split": st["split"]}
    # the F builder (K1) is HBM-bound; algorithmic bytes per build: F written once + the tile-major entries
    # (10 B each) + the passes over the E x ld subtree matrices
    hp, hsrc = hbm_peak()
    X = 8.0 * spec["edges"] * st["ld_samples"]           # one pass over an E x ld matrix
    if spec["bins"] == 1:
        # bins = 1: M and D ARE the rows of F.  Totals + zeros written (2 X = F), kids' M read (X), D written and M
        # read-modify-written for the inner edges (1.5 X); the tile-major entries are read once
        k1_bytes = 10.0 * st["n_masses"] + 4.5 * X
        k1_name = "K1, bins = 1 (k_tile<TOTALS> into F, k_subtree_level/top on F)"
    else:
        # F written once + the tile-major entries read twice + T written, M read-modify-written, kids' M read, D written
        # (inner edges), D read (inner edges): 4 X
        k1_bytes = float(st["f_bytes"]) + 20.0 * st["n_masses"] + 4.0 * X
        k1_name = "K1 (k_tile<TOTALS>, k_subtree_level/top, k_tile<FILL>)"
    k1_gbs = k1_bytes / (st["ms_build"] * 1e-3) / 1e9 if st["ms_build"] > 0 else 0.0
    k1 = {"bound": "hbm", "kernel": k1_name, "achieved": k1_gbs,
          "peak": hp, "unit": "GB/s", "frac": k1_gbs / hp, "bytes": k1_bytes,
          "ms": st["ms_build"], "peak_source": hsrc, "share_of_step": st["ms_build"] / ms_step}
    return roof, k1


def timed(torch, dist, stream, fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        fn()
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms / steps


def pinned(torch, a):
    t = torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a.view(np.int16) if a.dtype == np.uint16 else a)
    return t.pin_memory()


def transport_cells(cell, spec):
    """The cell ids in the narrowest form the C ABI takes: uint16 (krd_set_masses_binned16, 10 B per entry with the mass)
    when edges * bins <= 65536 — every BASELINE config — else uint32 (12 B per entry)."""
    return cell.astype(np.uint16) if spec["edges"] * spec["bins"] <= 65536 else cell


def cell_form(spec):
    return ("krd_set_masses_binned16", 10) if spec["edges"] * spec["bins"] <= 65536 else ("krd_set_masses_binned", 12)


def measure_single(local, spec, tree, comp, steps, warmup, sample_clocks, e2e_steps):
    """One GPU: `value` (resident masses -> matrix in host memory), `e2e` (host mass trees -> matrix in host memory),
    the stats of the last device step, and the leading blocks of both results for the parity check."""
    import torch
    import gappa_b200
    N, p, bins = spec["samples"], spec["p"], spec["bins"]
    pairs = N * (N - 1) / 2
    stream = torch.cuda.current_stream()
    h_off, h_cell, h_mass = pinned(torch, comp["off"]), pinned(torch, transport_cells(comp["cell"], spec)), pinned(torch, comp["mass"])
    h_out = torch.empty((N, N), dtype=torch.float64).pin_memory()
    h2d_bytes = sum(t.numel() * t.element_size() for t in (h_off, h_cell, h_mass))

    plan = gappa_b200.KrdPlan(device=local, bins=bins)
    plan.set_stream(stream.cuda_stream)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses_binned(h_off.numpy(), h_cell, h_mass)
    d_out = torch.empty((N, N), dtype=torch.float64, device="cuda")
    # the end-to-end arm uses the library's pipelined ingest (krd_config.pipeline = 1): sample blocks cross PCIe while
    # the blocks before them are prepared, built into F and paired.  Same calls, same host buffers.
    plan_e2e = gappa_b200.KrdPlan(device=local, bins=bins, pipeline=1)
    plan_e2e.set_stream(stream.cuda_stream)
    plan_e2e.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])

    def step_device():
        plan.invalidate_F()
        plan.compute_device(d_out, p=p)
        h_out.copy_(d_out, non_blocking=True)

    def step_e2e():
        plan_e2e.set_masses_binned(h_off.numpy(), h_cell, h_mass)
        plan_e2e.compute(p=p, out=h_out)

    sampler = ClockSampler(local) if sample_clocks else None
    ms_step = timed(torch, None, stream, step_device, steps, warmup)
    st = plan.stats()
    nb = min(N, 512)
    blk_dev = h_out[:nb, :nb].numpy().copy()
    clocks = None
    if sampler:
        # clocks: nvidia-smi samples every 20 ms; the timed region is tens of ms, so the same step keeps running (untimed)
        # until the sampler has seen ~0.4 s of this load.  It is stopped before the e2e region: polling the driver that
        # often delays the host-side submissions the e2e path depends on (measured +0.9 ms per step).
        for _ in range(int(min(2000, max(1, 400.0 / max(ms_step, 0.05))))):
            step_device()
        torch.cuda.synchronize()
        clocks = sampler.stop()
    # the D2H term of `value` alone
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(3):
        h_out.copy_(d_out, non_blocking=True)
    e1.record(stream); torch.cuda.synchronize()
    ms_d2h = e0.elapsed_time(e1) / 3
    ms_e2e = timed(torch, None, stream, step_e2e, e2e_steps, 3)
    blk_e2e = h_out[:nb, :nb].numpy().copy()
    res = dict(N=N, pairs=pairs, ms_step=ms_step, ms_e2e=ms_e2e, ms_d2h=ms_d2h, st=st, clocks=clocks, h2d_bytes=int(h2d_bytes),
               d2h_bytes=int(N * N * 8), blocks={"value_path": blk_dev, "e2e_path": blk_e2e}, launches=int(st["launches"]))
    plan.close(); plan_e2e.close()
    del d_out, h_out
    torch.cuda.empty_cache()
    return res


def run_single(args):
    from gappa_b200 import workloads
    import torch
    import gappa_b200
    local = int(os.environ.get("LOCAL_RANK", "0"))
    ncpu = host_threads()
    name = default_workload(args, 1)
    over = dict(samples=args.samples, edges=args.edges, placements=args.placements, bins=args.bins, p=args.p)
    spec, tree, comp = workloads.build_compact(name, workers=min(32, ncpu), n_samples=args.samples, cache=True, **over)
    torch.cuda.set_device(local)
    warmup = max(args.warmup, 3)
    r = measure_single(local, spec, tree, comp, args.steps, warmup, True, max(2, min(args.steps, 5)))
    st, pairs = r["st"], r["pairs"]
    roof, k1 = rooflines(gappa_b200, local, spec, st, pairs, 1.0, r["ms_step"], 1)
    line = {
        "metric": metric_of(spec), "value": pairs / (r["ms_step"] * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": warmup, "ms_per_step": r["ms_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_of(spec, 1),
        "value_includes": "F build (K1) + pair kernel + epilogue + D2H of the n x n matrix into pinned host memory (%.3f ms of the step)" % r["ms_d2h"],
        "masses": int(st["n_masses"]),
        "e2e": {"value": pairs / (r["ms_e2e"] * 1e-3), "unit": UNIT, "ms_per_step": r["ms_e2e"],
                "ingest": "%s (%d B per entry: cell = edge*bins+bin, mass) on a pipeline=1 handle: up to 8 sample blocks, "
                          "upload overlapped with K0/K1/pair phases" % cell_form(spec),
                "h2d_bytes_per_step": r["h2d_bytes"], "d2h_bytes_per_step": r["d2h_bytes"]},
        "gpu_launches": r["launches"] * args.steps,
        "clocks": r["clocks"],
        "roofline": roof, "roofline_build_F": k1,
    }
    blocks = r["blocks"]
    del r
    if not args.no_cpu_baseline:
        full = workloads.expand(tree, workloads.head(comp, min(spec["samples"], 512)), spec["bins"])
        info, D_ref = cpu_reference(tree, full, spec["p"], args.cpu_seconds)
        line["cpu_baseline"] = info
        line["parity"] = parity_of(blocks, D_ref)
    del comp
    # ---- the north-star's multi-GPU config on ONE GPU: the base of the strong-scaling numbers `--gpus N` prints ----
    if not args.no_base and name == "cfg2" and args.samples is None:
        try:
            t0 = time.time()
            spec4, tree4, comp4 = workloads.build_compact("cfg4", workers=min(32, ncpu), cache=True)
            t_gen = time.time() - t0
            r4 = measure_single(local, spec4, tree4, comp4, 3, 2, False, 2)
            roof4, k14 = rooflines(gappa_b200, local, spec4, r4["st"], r4["pairs"], 1.0, r4["ms_step"], 1)
            base = {"config": config_of(spec4, 1), "value": r4["pairs"] / (r4["ms_step"] * 1e-3), "unit": UNIT, "ms_per_step": r4["ms_step"],
                    "steps": 3, "warmup": 2, "ms_d2h": r4["ms_d2h"], "masses": int(r4["st"]["n_masses"]),
                    "e2e": {"value": r4["pairs"] / (r4["ms_e2e"] * 1e-3), "unit": UNIT, "ms_per_step": r4["ms_e2e"],
                            "h2d_bytes_per_step": r4["h2d_bytes"], "d2h_bytes_per_step": r4["d2h_bytes"]},
                    "roofline": {k: roof4[k] for k in ("achieved", "peak", "frac", "frac_of_mix_ceiling", "ms_kernel", "ms_build_F", "ms_epilogue")},
                    "roofline_build_F": {k: k14[k] for k in ("achieved", "frac", "bytes", "ms")},
                    "input_generation_s": t_gen}
            if not args.no_cpu_baseline:
                full4 = workloads.expand(tree4, workloads.head(comp4, 48), 1)
                info4, D4 = cpu_reference(tree4, full4, 1.0, 0.0, n_fixed=48)
                base["cpu_baseline"] = info4
                base["parity"] = parity_of(r4["blocks"], D4)
            line["strong_scaling_base"] = base
        except Exception as ex:   # the headline line must survive a failure of the extra block
            line["strong_scaling_base"] = {"error": repr(ex)[:300]}
    print(json.dumps(line))


def run_exact(args):
    """--workload exact1k (bins = 0, exact positions): the mode `gappa analyze krd` runs.  S ~ number of mass entries, so
    the dense matrix does not fit and the library streams the segment axis in slabs (DESIGN.md "Exact mode")."""
    from gappa_b200 import workloads
    import torch
    import gappa_b200
    local = int(os.environ.get("LOCAL_RANK", "0"))
    name = default_workload(args, 1)
    spec, tree, masses = workloads.build(name, workers=min(32, host_threads()), n_samples=args.samples, samples=args.samples,
                                         edges=args.edges, placements=args.placements, p=args.p)
    N, p = spec["samples"], spec["p"]
    pairs = N * (N - 1) / 2
    torch.cuda.set_device(local)
    stream = torch.cuda.current_stream()
    h = [pinned(torch, masses[k]) for k in ("off", "edge", "pos", "mass")]
    h_out = torch.empty((N, N), dtype=torch.float64).pin_memory()
    plan = gappa_b200.KrdPlan(device=local, bins=0)
    plan.set_stream(stream.cuda_stream)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(h[0].numpy(), h[1], h[2], h[3])
    d_out = torch.empty((N, N), dtype=torch.float64, device="cuda")

    def step_device():
        plan.invalidate_F()
        plan.compute_device(d_out, p=p)
        h_out.copy_(d_out, non_blocking=True)

    def step_e2e():
        plan.set_masses(h[0].numpy(), h[1], h[2], h[3])
        plan.compute(p=p, out=h_out)

    steps, warmup = max(1, min(args.steps, 3)), 1
    sampler = ClockSampler(local)
    ms_step = timed(torch, None, stream, step_device, steps, warmup)
    clocks = sampler.stop()
    st = plan.stats()
    nb = min(N, 256)
    blk = h_out[:nb, :nb].numpy().copy()
    ms_e2e = timed(torch, None, stream, step_e2e, 2, 1)
    blk2 = h_out[:nb, :nb].numpy().copy()
    S = st["n_segments"]
    spec_cfg = config_of(spec, 1)
    spec_cfg["segments"] = int(S)
    spec_cfg["l2"] = "inputs larger than L2 (F is streamed: %.0f MB slabs of a %.0f GB matrix)" % (st["f_bytes"] / 1e6, S * st["ld_samples"] * 8 / 1e9)
    peak_dfma = gappa_b200.measure_fp64_peak(local, 0, 200.0); peak_mix = gappa_b200.measure_fp64_peak(local, 1, 200.0)
    peak_dmma = gappa_b200.measure_fp64_peak(local, 2, 200.0)
    peak = max(peak_dfma, peak_dmma)
    achieved = 3.0 * S * pairs / (st["ms_pairs"] * 1e-3) / 1e12
    hp, hsrc = hbm_peak()
    slab_gbs = S * st["ld_samples"] * 8.0 / (st["ms_slabs"] * 1e-3) / 1e9 if st["ms_slabs"] > 0 else 0.0
    line = {
        "metric": metric_of(spec), "value": pairs / (ms_step * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": spec_cfg, "masses": int(st["n_masses"]),
        "value_includes": "per-entry cumulative values (T, subtree passes, cum) + every slab of F built and paired + epilogue + D2H",
        "e2e": {"value": pairs / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e, "ingest": "krd_set_masses (20 B per entry) + K0 exact (radix sort of the (edge, pos) keys, segment table)",
                "h2d_bytes_per_step": int(sum(t.numel() * t.element_size() for t in h)), "d2h_bytes_per_step": int(N * N * 8)},
        "gpu_launches": int(st["launches"]) * steps, "clocks": clocks,
        "roofline": {"bound": "fp64", "kernel": "k_pairs<p=%g> over all slabs" % p, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak, "traffic": None, "mix_ceiling": peak_mix, "frac_of_mix_ceiling": achieved / peak_mix,
                     "peak_source": "measured live: max(DFMA loop %.1f, DMMA m8n8k4 loop %.1f) TFLOP/s" % (peak_dfma, peak_dmma),
                     "ms_kernel": st["ms_pairs"], "ms_prefix": st["ms_build"], "ms_slabs": st["ms_slabs"]},
        "roofline_build_F": {"bound": "hbm", "kernel": "k_exact_slab (F written once, slab by slab)", "achieved": slab_gbs, "peak": hp, "unit": "GB/s",
                             "frac": slab_gbs / hp, "bytes": S * st["ld_samples"] * 8.0, "ms": st["ms_slabs"], "peak_source": hsrc,
                             "share_of_step": st["ms_slabs"] / ms_step},
    }
    if not args.no_cpu_baseline:
        info, D_ref = cpu_reference(tree, workloads.head(masses, nb), p, args.cpu_seconds)
        line["cpu_baseline"] = info
        line["parity"] = parity_of({"value_path": blk, "e2e_path": blk2}, D_ref)
    print(json.dumps(line))


def run_multi(args):
    from gappa_b200 import workloads
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    ncpu = host_threads()
    name = default_workload(args, world)
    over = dict(samples=args.samples, edges=args.edges, placements=args.placements, bins=args.bins, p=args.p)
    # inputs first (generation forks worker processes: before CUDA is touched); every rank generates 1/world of the samples
    spec, tree, comp = workloads.build_compact(name, workers=max(1, min(32, ncpu // world)), n_samples=args.samples, rank=rank,
                                               world=world, cache=True, **over)
    N, p, bins = spec["samples"], spec["p"], spec["bins"]
    pairs = N * (N - 1) / 2

    import torch
    import torch.distributed as dist
    import gappa_b200
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stream = torch.cuda.current_stream()
    cudart = torch.cuda.cudart()

    # ---- the result in host memory: ONE n x n matrix in a shared segment, page-locked by every rank; rank r fills rows
    # [r0, r1) over its own PCIe link ----
    r0, r1 = rank * N // world, (rank + 1) * N // world
    shm = "/dev/shm/krd_bench_out_%s" % os.environ.get("MASTER_PORT", "0")
    if rank == 0:
        with open(shm, "wb") as f:
            f.truncate(N * N * 8)
    dist.barrier()
    h_full = torch.from_numpy(np.memmap(shm, dtype=np.float64, mode="r+", shape=(N, N)))
    rc = cudart.cudaHostRegister(h_full.data_ptr(), N * N * 8, 0)
    assert int(rc) == 0, f"cudaHostRegister failed: {rc}"
    h_rows = h_full[r0:r1]

    plan = gappa_b200.KrdPlan(device=local, bins=bins)
    plan.set_stream(stream.cuda_stream)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])

    # ---- sharded == single, bit for bit, on a 300-sample cohort (before anything is timed) ----
    small = workloads.head(comp, min(N, 300))
    ns = len(small["off"]) - 1
    plan.set_masses_binned(small["off"], small["cell"], small["mass"])
    sh = plan.shard_doubles(world)
    dp = torch.empty(sh, dtype=torch.float64, device="cuda"); dg = torch.empty(sh * world, dtype=torch.float64, device="cuda")
    plan.compute_shard(dp, rank, world, p=p)
    dist.all_gather_into_tensor(dg, dp)
    got = torch.full((ns, ns), -1.0, dtype=torch.float64, device="cuda")
    plan.unpack_shards(dg, got, world)
    one = torch.full((ns, ns), -2.0, dtype=torch.float64, device="cuda")
    plan.compute_device(one, p=p)
    torch.cuda.synchronize()
    # (the two runs cut the segment axis differently between CTAs, so their partial sums round differently: a few ulp)
    shard_diff = float(((got - one).abs() / one.abs().clamp_min(1e-300)).max().item())
    assert shard_diff <= 1e-12, f"sharded result differs from the single-GPU result: {shard_diff}"
    del dp, dg, got, one

    # ---- the cohort ----
    M = int(comp["off"][-1])
    plan.set_masses_binned(comp["off"], comp["cell"], comp["mass"])     # (pageable source: one-time)
    shard = plan.shard_doubles(world)
    d_packed = torch.empty(shard, dtype=torch.float64, device="cuda")
    d_gath = torch.empty(shard * world, dtype=torch.float64, device="cuda")
    d_rows = torch.empty((r1 - r0, N), dtype=torch.float64, device="cuda")
    d_out_full = torch.empty((N, N), dtype=torch.float64, device="cuda") if (args.exchange == "allgather_rank0" or rank == 0) else None
    # p2p_rows: row blocks every rank can write (cudaMalloc + CUDA IPC inside the library), two sets used alternately so that
    # a rank never writes a block its owner may still be copying to the host
    bounds = np.array([r * N // world for r in range(world + 1)], np.int32)
    p2p_ptrs, p2p_own, p2p_mine, step_no = None, None, [], [0]
    if args.exchange == "p2p_rows":
        class _Raw:
            def __init__(self, ptr, shape):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f8", "data": (ptr, False), "version": 2}
        mine = [gappa_b200.krd.p2p_alloc(local, max(1, (r1 - r0) * N * 8)) for _ in range(2)]
        p2p_mine = [m[0] for m in mine]
        handles = [None] * world
        dist.all_gather_object(handles, [m[1] for m in mine])
        p2p_ptrs = [[p2p_mine[k] if r == rank else gappa_b200.krd.p2p_open(local, handles[r][k]) for r in range(world)] for k in range(2)]
        p2p_own = [torch.as_tensor(_Raw(p2p_mine[k], (r1 - r0, N)), device=torch.device("cuda", local)) for k in range(2)]
        d_flag = torch.zeros(1, dtype=torch.float32, device="cuda")
        # two-part tile order: tiles that touch the last `late` sample blocks come last (krd_set_late_blocks)
        Tb = (N + 127) // 128
        late = min(Tb - 1, max(1, int(round(Tb * args.late)))) if args.late > 0 else 0
        t_split = plan.set_late_blocks(late)
        n_tiles_all = Tb * (Tb + 1) // 2
        a0, a1 = gappa_b200.krd.rank_tiles(t_split, rank, world)
        b0, b1 = (t_split + q for q in gappa_b200.krd.rank_tiles(n_tiles_all - t_split, rank, world))
        cA = min(N, (Tb - late) * 128)           # rows / columns below cA are final after the first part
        d_pkA = torch.empty(max(1, a1 - a0) * 128 * 128, dtype=torch.float64, device="cuda")
        d_pkB = torch.empty(max(1, b1 - b0) * 128 * 128, dtype=torch.float64, device="cuda")
        p2p_stream = torch.cuda.Stream()
        rect = gappa_b200.krd.copy_rect_to_host_async
    PH = 1
    if args.exchange == "host_tiles":
        # the rank's contiguous share of the tile list, cut UNEVENLY: a long first launch and a short last one (--tail); per piece
        # the packed tiles + their transposes
        n_tiles_all = len(gappa_b200.krd.tile_list(N)[0])
        t0r, t1r = gappa_b200.krd.rank_tiles(n_tiles_all, rank, world)
        cut = t1r - int(round((t1r - t0r) * max(0.0, min(0.9, args.tail))))
        pieces = [(a, b) for a, b in ((t0r, cut), (cut, t1r)) if b > a]
        PH = len(pieces)
        d_pk = [torch.empty((b - a) * 128 * 128, dtype=torch.float64, device="cuda") for a, b in pieces]
        d_sc = [torch.empty((b - a) * 128 * 128, dtype=torch.float64, device="cuda") for a, b in pieces]
        out_stream = torch.cuda.Stream()
    h_packed = None
    if args.exchange == "own_tiles_host":
        shm2 = shm + "_packed"
        if rank == 0:
            with open(shm2, "wb") as f:
                f.truncate(shard * world * 8)
        dist.barrier()
        h_packed = torch.from_numpy(np.memmap(shm2, dtype=np.float64, mode="r+", shape=(world, shard)))
        assert int(cudart.cudaHostRegister(h_packed.data_ptr(), shard * world * 8, 0)) == 0
    # e2e ingest at N GPUs: every rank uploads 1/N of the mass entries over ITS PCIe link and the ranks all-gather the
    # rest over NVLink (the one exchange step on the input side; F itself is built replicated)
    per = (M + world - 1) // world
    lo, hi = min(rank * per, M), min((rank + 1) * per, M)
    h_cell = pinned(torch, transport_cells(np.ascontiguousarray(comp["cell"][lo:hi]), spec))
    cell_dt = h_cell.dtype                                      # int16 (10 B per entry with the mass) or int32
    h_mass = pinned(torch, np.ascontiguousarray(comp["mass"][lo:hi]))
    off_np = np.ascontiguousarray(comp["off"], np.int64)
    d_part = [torch.zeros(per, dtype=cell_dt, device="cuda"), torch.zeros(per, dtype=torch.float64, device="cuda")]
    d_all = [torch.empty(per * world, dtype=cell_dt, device="cuda"), torch.empty(per * world, dtype=torch.float64, device="cuda")]
    h2d_bytes = (hi - lo) * (8 + h_cell.element_size()) + off_np.nbytes
    copy_stream = torch.cuda.Stream()
    up_events = [torch.cuda.Event() for _ in range(2)]
    head_blk = workloads.head(comp, min(N, 64))
    del comp

    def exchange(ev=None):
        mark = (lambda k: ev[k].record(stream)) if ev else (lambda k: None)
        if args.exchange == "allgather_rows":
            dist.all_gather_into_tensor(d_gath, d_packed)
            mark(1)
            plan.unpack_shards_rows(d_gath, d_rows, world, r0, r1)
            mark(2)
            h_rows.copy_(d_rows, non_blocking=True)
            mark(3)
        elif args.exchange == "allgather_rank0":        # round-1 variant: full unpack everywhere, one D2H over rank 0's link
            dist.all_gather_into_tensor(d_gath, d_packed)
            plan.unpack_shards(d_gath, d_out_full, world)
            if rank == 0:
                h_full.copy_(d_out_full, non_blocking=True)
        else:                                           # no collective: own packed tiles to the host, unpacked there (untimed here)
            h_packed[rank].copy_(d_packed, non_blocking=True)

    def step_host_tiles(collect=None):
        # No collective: every cell of the host matrix is written by exactly one rank, by the copy engines, and the copies
        # of the long first piece run under the pair launch of the short last one (the pair kernel fills every SM; the
        # copy engines are free)
        stream.wait_stream(out_stream)             # (the previous step's copies have read d_pk / d_sc)
        for q, (a, b) in enumerate(pieces):
            plan.compute_tile_range(d_pk[q], a, b, p=p)
            if collect is not None:
                torch.cuda.synchronize()
                collect.append(plan.stats())
            plan.tile_range_to_host_async(a, b, d_pk[q], d_sc[q], h_full, out_stream.cuda_stream)
        stream.wait_stream(out_stream)             # the step ends when the last copy has landed

    def step_p2p(collect=None, ev=None):
        # Tiles (and their mirrors) are stored straight into the owners' row blocks over NVLink; a one-element all-reduce on
        # the same stream is the barrier after which a block is complete.  Two parts: after part A (tiles that do not touch
        # the late blocks) the rectangle [own rows] x [columns < cA] is final and goes to the host on a second stream, UNDER
        # the pair launch of part B; only the thin late strip is copied after it.
        k = step_no[0] % 2
        step_no[0] += 1
        own, hb = p2p_mine[k], h_full.data_ptr()
        rA = min(r1, cA)                           # own rows [r0, rA) are "early", [rA, r1) late
        stream.wait_stream(p2p_stream)
        plan.compute_tile_range(d_pkA, a0, a1, p=p)
        if collect is not None:
            torch.cuda.synchronize(); collect.append(plan.stats())
        plan.scatter_tile_range_rows(a0, a1, d_pkA, bounds, p2p_ptrs[k])
        dist.all_reduce(d_flag)
        if late > 0:
            p2p_stream.wait_stream(stream)
            if ev: ev[0].record(p2p_stream)
            if rA > r0:
                rect(hb + r0 * N * 8, N, own, N, cA, rA - r0, p2p_stream.cuda_stream)
            if ev: ev[1].record(p2p_stream)
            if b1 > b0:
                plan.compute_tile_range(d_pkB, b0, b1, p=p)
                if collect is not None:
                    torch.cuda.synchronize(); collect.append(plan.stats())
                plan.scatter_tile_range_rows(b0, b1, d_pkB, bounds, p2p_ptrs[k])
            dist.all_reduce(d_flag)
            if ev: ev[2].record(stream)
            if rA > r0 and N > cA:
                rect(hb + (r0 * N + cA) * 8, N, own + cA * 8, N, N - cA, rA - r0, stream.cuda_stream)
            if r1 > max(r0, rA):
                rl = max(r0, rA)
                rect(hb + rl * N * 8, N, own + (rl - r0) * N * 8, N, N, r1 - rl, stream.cuda_stream)
            stream.wait_stream(p2p_stream)
            if ev: ev[3].record(stream)
        else:
            if ev: ev[0].record(stream); ev[1].record(stream); ev[2].record(stream)
            rect(hb + r0 * N * 8, N, own, N, N, r1 - r0, stream.cuda_stream)
            if ev: ev[3].record(stream)

    def step_device():
        plan.invalidate_F()
        if args.exchange == "host_tiles":
            return step_host_tiles()
        if args.exchange == "p2p_rows":
            return step_p2p()
        plan.compute_shard(d_packed, rank, world, p=p)
        exchange()

    def step_e2e():
        srcs = (h_cell, h_mass)
        copy_stream.wait_stream(stream)            # (the previous step's all-gathers have read d_part)
        with torch.cuda.stream(copy_stream):
            for k in (1, 0):                       # masses first: the all-gather left exposed after the last upload is the smaller array's
                d_part[k][: hi - lo].copy_(srcs[k], non_blocking=True)
                up_events[k].record(copy_stream)
        for k in (1, 0):
            stream.wait_event(up_events[k])
            dist.all_gather_into_tensor(d_all[k].view(torch.uint8), d_part[k].view(torch.uint8))   # (bytes: NCCL has no 16-bit integer type)
        plan.set_masses_binned_device(off_np, d_all[0], d_all[1])
        if args.exchange == "host_tiles":
            step_host_tiles()
        elif args.exchange == "p2p_rows":
            step_p2p()
        else:
            plan.compute_shard(d_packed, rank, world, p=p)
            exchange()
        torch.cuda.synchronize()

    warmup = max(args.warmup, 3)
    sampler = ClockSampler(local) if rank == 0 else None
    ms_step = timed(torch, dist, stream, step_device, args.steps, warmup)
    st = plan.stats()
    if args.exchange in ("host_tiles", "p2p_rows"):   # one traced step: per-launch times of the parts (K1 runs in the first)
        plan.invalidate_F()
        per = []
        (step_host_tiles if args.exchange == "host_tiles" else step_p2p)(per)
        PH = len(per)
        torch.cuda.synchronize()
        st = dict(per[-1])
        st["ms_pairs"] = float(sum(q["ms_pairs"] for q in per)); st["ms_epilogue"] = float(sum(q["ms_epilogue"] for q in per))
        st["ms_build"] = per[0]["ms_build"]; st["launches"] = int(sum(q["launches"] for q in per)) + PH
        st["ms_pairs_per_launch"] = [q["ms_pairs"] for q in per]
    for _ in range(int(min(2000, max(1, 400.0 / max(ms_step, 0.05))))):    # (same count on every rank: ms_step is the max over ranks)
        step_device()
    torch.cuda.synchronize()
    clocks = sampler.stop() if sampler else None
    dist.barrier()
    nb = min(N, 64)
    blk_dev = None
    ms_host_unpack = None
    if rank == 0 and args.exchange != "own_tiles_host":
        torch.cuda.synchronize()
        blk_dev = h_full[:nb, :nb].numpy().copy()
    if args.exchange == "own_tiles_host" and rank == 0:
        t0 = time.perf_counter()
        Dh = gappa_b200.krd.unpack_shards_host(N, [h_packed[r].numpy() for r in range(world)])
        ms_host_unpack = (time.perf_counter() - t0) * 1e3
        blk_dev = Dh[:nb, :nb].copy()
        del Dh
    # one traced step: where the time after the pair kernel goes (CUDA events on this rank's stream)
    breakdown = None
    if args.exchange == "p2p_rows":
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        dist.barrier(); torch.cuda.synchronize()
        plan.invalidate_F()
        step_p2p(None, ev)
        torch.cuda.synchronize()
        breakdown = {"d2h_early_rectangle_on_side_stream": ev[0].elapsed_time(ev[1]), "exposed_after_last_barrier": ev[2].elapsed_time(ev[3]),
                     "late_blocks": late, "early_columns": cA}
    if args.exchange == "allgather_rows":
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        dist.barrier(); torch.cuda.synchronize()
        plan.invalidate_F()
        plan.compute_shard(d_packed, rank, world, p=p)
        ev[0].record(stream)
        exchange(ev)
        torch.cuda.synchronize()
        names = ("nccl_allgather", "unpack_row_block", "d2h_row_block")
        breakdown = {nm: ev[k].elapsed_time(ev[k + 1]) for k, nm in enumerate(names)}
    ms_e2e = timed(torch, dist, stream, step_e2e, max(2, min(args.steps, 5)), 3)
    dist.barrier()
    blk_e2e = h_full[:nb, :nb].numpy().copy() if (rank == 0 and args.exchange != "own_tiles_host") else None

    # ---- the same cohort on ONE GPU of this job (rank 0; the others wait): the base of the strong-scaling ratio ----
    one_gpu = None
    if rank == 0:
        h_one = torch.empty((N, N), dtype=torch.float64).pin_memory()
        if args.exchange == "p2p_rows":
            plan.set_late_blocks(0)          # the one-GPU base runs on the default tile order

        def step_one():
            plan.invalidate_F()
            plan.compute_device(d_out_full, p=p)
            h_one.copy_(d_out_full, non_blocking=True)
        ms_one = timed(torch, None, stream, step_one, 2, 1)
        st1 = plan.stats()
        one_blk = h_one.numpy()[:nb, :nb]
        same = float(np.max(np.abs(one_blk - blk_dev) / np.maximum(np.abs(blk_dev), 1e-300))) if blk_dev is not None else None
        one_gpu = {"value": pairs / (ms_one * 1e-3), "unit": UNIT, "ms_per_step": ms_one, "steps": 2, "warmup": 1,
                   "ms_kernel": st1["ms_pairs"], "ms_build_F": st1["ms_build"],
                   "note": "same cohort, same definition of value (incl. D2H of the whole matrix over one PCIe link), rank 0 alone",
                   "leading_block_max_rel_diff_to_sharded": same}
        del h_one
    dist.barrier()

    if rank == 0:
        roof, k1 = rooflines(gappa_b200, local, spec, st, pairs, 1.0 / world, ms_step, world)
        line = {
            "metric": metric_of(spec), "value": pairs / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(spec, world),
            "value_includes": "per rank: F build (K1, replicated) + pair kernel on its tile share + epilogue + exchange (%s) + D2H; max over ranks"
                              % args.exchange,
            "masses": int(st["n_masses"]),
            "exchange": args.exchange, "launches_per_step": PH, "tail_fraction": args.tail if args.exchange == "host_tiles" else None,
            "ms_pairs_per_launch": st.get("ms_pairs_per_launch"), "exchange_breakdown_ms_rank0": breakdown, "ms_host_unpack_untimed": ms_host_unpack,
            "shard_check": {"samples": ns, "max_rel_diff_sharded_vs_single_gpu": shard_diff, "bound": 1e-12},
            "e2e": {"value": pairs / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "ingest": "every rank uploads 1/N of the compact entries (%s, %d B each) on a copy stream, NVLink all-gather of the two arrays, "
                              "K0 + K1 replicated, sharded pair kernel, exchange, per-rank D2H" % cell_form(spec),
                    "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int((r1 - r0) * N * 8)},
            "one_gpu_same_job": one_gpu,
            "strong_scaling": {"speedup_vs_one_gpu_same_job": (pairs / (ms_step * 1e-3)) / one_gpu["value"], "n_gpus": world,
                               "note": "pairs/s is proportional to 1/segments, so this line (cfg4, 99,998 segments) is not comparable with the "
                                       "--gpus 1 line (cfg2, 49,995 segments): the one-GPU base of THIS workload is one_gpu_same_job here and "
                                       "strong_scaling_base in the --gpus 1 line"},
            "gpu_launches": (int(st["launches"]) + 1) * args.steps,
            "clocks": clocks,
            "roofline": roof, "roofline_build_F": k1,
        }
        if not args.no_cpu_baseline and blk_dev is not None:
            full = workloads.expand(tree, head_blk, bins)
            info, D_ref = cpu_reference(tree, full, p, 0.0, n_fixed=nb)
            line["parity"] = parity_of({"value_path": blk_dev, **({"e2e_path": blk_e2e} if blk_e2e is not None else {})}, D_ref)
            line["parity"]["reference_sample"] = info["sample"]
        print(json.dumps(line))
    dist.barrier()
    if p2p_ptrs:
        torch.cuda.synchronize()
        p2p_own = None
        for k in range(2):
            for r in range(world):
                if r != rank:
                    gappa_b200.krd.p2p_close(p2p_ptrs[k][r])
        dist.barrier()
        for q in p2p_mine:
            gappa_b200.krd.p2p_free(q)
    cudart.cudaHostUnregister(h_full.data_ptr())
    dist.destroy_process_group()
    if rank == 0:
        for f in (shm, shm + "_packed"):
            try:
                os.unlink(f)
            except OSError:
                pass


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        from gappa_b200 import workloads
        if workloads.WORKLOADS[default_workload(args, 1)]["bins"] == 0 and not args.bins:
            return run_exact(args)
        return run_single(args)
    return run_multi(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py — KR sample-pairs/s of the B200 path (and of the reference CPU path, --impl reference).

One "step" = one pass of the hot path over the workload's mass trees:
  value : masses resident in HBM -> F build (K1) + pair kernel (K2) + epilogue -> matrix in HBM
  e2e   : HOST (pinned) mass trees -> krd_set_masses (H2D + slots) + krd_compute -> matrix on HOST
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

METRIC = "KR sample-pairs/sec (p=1, FP64)"
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("KRD_WORKLOAD", "cfg2"))
    ap.add_argument("--samples", type=int)
    ap.add_argument("--edges", type=int)
    ap.add_argument("--placements", type=int)
    ap.add_argument("--bins", type=int)
    ap.add_argument("--p", type=float)
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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


def traffic_from_profile(spec, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of the pair kernel per launch, from the committed
    `ncu --set full` capture of this workload (profiles/pair_kernel_traffic.json); None when there is none."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "pair_kernel_traffic.json")))
        key = "%s_p%g" % (spec["name"], spec["p"])
        return t[key]["bytes_per_launch"] if world == 1 and spec["samples"] == t[key]["samples"] else None
    except (OSError, KeyError, ValueError):
        return None


def cpu_reference_rate(tree, masses, p, bins, target_seconds, threads=0):
    """Times the reference's own CPU implementation (oracle/_ref = compiled genesis, OpenMP over
    all host threads) — or the C port when _ref is absent — on a bounded sample: the first n
    samples of the same mass trees.  Returns dict(value pairs/s, cores, kind, sample)."""
    import oracle
    from gappa_b200 import workloads
    use_ref = oracle.have_ref()
    kind = "reference" if use_ref else "port"
    cores = (oracle.ref().ref_max_threads() if use_ref else oracle.port().okrd_max_threads()) if threads == 0 else threads

    def run(n):
        sub = workloads.head(masses, n)
        t0 = time.perf_counter()
        if use_ref:
            _, sec = oracle.ref_matrix(tree, sub, p, bins=0, threads=threads)   # masses are already binified
        else:
            oracle.port_matrix(tree, sub, p, bins=0, threads=threads)
            sec = time.perf_counter() - t0
        return n * (n - 1) / 2 / sec, sec

    N = len(masses["off"]) - 1
    n0 = min(N, max(2, 2 * int(np.sqrt(cores)) + 8))
    rate, sec = run(n0)
    n1 = int(min(N, max(n0, np.sqrt(2 * target_seconds * rate))))
    if n1 > n0:
        rate, sec = run(n1)
        n0 = n1
    return dict(value=rate, unit=UNIT, cores=int(cores), kind=kind,
                sample=f"first {n0} of {N} samples = {n0 * (n0 - 1) // 2} pairs, {sec:.2f} s, earth_movers_distance(trees,p) only")


def run_reference(args):
    """--impl reference: the reference's CPU path on this box's host cores, same config/metric."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from gappa_b200 import workloads
    spec = dict(workloads.WORKLOADS[args.workload])
    for k in ("samples", "edges", "placements", "bins", "p"):
        if getattr(args, k) is not None:
            spec[k] = getattr(args, k)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    if args.samples is None and world > 1:     # same weak-scaling rule as the b200 arm
        spec["samples"] = int(round(spec["samples"] * np.sqrt(world) / 128.0)) * 128
    # a bounded sample: only as many samples as the CPU can finish in a few seconds per step
    n_gen = min(spec["samples"], 160)
    spec_b, tree, masses = workloads.build(args.workload, n_samples=n_gen, **{k: spec[k] for k in ("edges", "placements", "bins", "p")})
    rates, info = [], None
    for it in range(args.warmup + args.steps):
        info = cpu_reference_rate(tree, masses, spec["p"], spec["bins"], target_seconds=max(1.0, args.cpu_seconds / max(args.steps, 1)))
        if it >= args.warmup:
            rates.append(info["value"])
    value = float(np.mean(rates))
    info["value"] = value
    pairs_total = spec["samples"] * (spec["samples"] - 1) / 2
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": pairs_total / value * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, **spec, "note": "ms_per_step extrapolated linearly in pairs from the bounded sample"},
        "cpu_baseline": info,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    # ---- inputs (host, before CUDA is touched: generation forks worker processes) ----
    from gappa_b200 import workloads
    ncpu = os.cpu_count() or 1
    # Weak scaling: the cohort grows with the GPU count so that every GPU keeps the pair count of the 1-GPU
    # workload (pairs ~ samples^2): samples = base * sqrt(world), rounded to whole 128-sample tiles.
    n_samples = args.samples
    scaling = "strong" if (args.samples is not None and world > 1) else "weak"   # --samples pins the cohort: total work fixed
    if n_samples is None and world > 1:
        base = workloads.WORKLOADS[args.workload]["samples"]
        n_samples = int(round(base * np.sqrt(world) / 128.0)) * 128
    spec, tree, masses = workloads.build(args.workload, workers=max(1, min(32, ncpu // world)), samples=n_samples,
                                         edges=args.edges, placements=args.placements, bins=args.bins, p=args.p)
    N, p, bins = spec["samples"], spec["p"], spec["bins"]
    pairs = N * (N - 1) / 2

    import torch
    import gappa_b200
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    # pinned host copies of the mass trees (the e2e arm copies from these every step)
    h_off = torch.from_numpy(masses["off"]).pin_memory()
    h_edge = torch.from_numpy(masses["edge"]).pin_memory()
    h_pos = torch.from_numpy(masses["pos"]).pin_memory()
    h_mass = torch.from_numpy(masses["mass"]).pin_memory()
    h_out = torch.empty((N, N) if rank == 0 else (1, 1), dtype=torch.float64).pin_memory()   # only rank 0 receives the matrix
    h2d_bytes = sum(t.numel() * t.element_size() for t in (h_off, h_edge, h_pos, h_mass))

    plan = gappa_b200.KrdPlan(device=local, bins=bins)
    stream = torch.cuda.current_stream()
    plan.set_stream(stream.cuda_stream)
    plan.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    plan.set_masses(h_off.numpy(), h_edge, h_pos, h_mass)
    d_out = torch.empty((N, N), dtype=torch.float64, device="cuda")
    plan_e2e = plan
    if world == 1:
        # the end-to-end arm uses the library's pipelined ingest (krd_config.pipeline = 1): sample blocks cross PCIe while
        # the blocks before them are prepared, built into F and paired.  Same calls, same host buffers.
        plan_e2e = gappa_b200.KrdPlan(device=local, bins=bins, pipeline=1)
        plan_e2e.set_stream(stream.cuda_stream)
        plan_e2e.set_tree(tree["parent"], tree["child"], tree["bl"], tree["root"])
    if world > 1:
        shard = plan.shard_doubles(world)
        d_packed = torch.empty(shard, dtype=torch.float64, device="cuda")
        d_gath = torch.empty(shard * world, dtype=torch.float64, device="cuda")
        # e2e ingest at N GPUs: every rank uploads 1/N of the mass entries over ITS PCIe link and the ranks
        # all-gather the rest over NVLink (the one exchange step on the input side; F itself is built replicated)
        M = int(h_edge.numel())
        per = (M + world - 1) // world
        lo, hi = min(rank * per, M), min((rank + 1) * per, M)
        d_part = [torch.zeros(per, dtype=t.dtype, device="cuda") for t in (h_edge, h_pos, h_mass)]
        d_full = [torch.empty(per * world, dtype=t.dtype, device="cuda") for t in (h_edge, h_pos, h_mass)]
        h2d_bytes = (hi - lo) * 20 + h_off.numel() * 8
        copy_stream = torch.cuda.Stream()
        up_events = [torch.cuda.Event() for _ in range(3)]
        up_order = (1, 2, 0)      # pos, mass, edge: the all-gather left exposed after the last upload is the smallest array's

    def step_device():
        plan.invalidate_F()
        if world == 1:
            plan.compute_device(d_out, p=p)
        else:
            plan.compute_shard(d_packed, rank, world, p=p)
            dist.all_gather_into_tensor(d_gath, d_packed)
            plan.unpack_shards(d_gath, d_out, world)

    def step_e2e():
        if world == 1:
            plan_e2e.set_masses(h_off.numpy(), h_edge, h_pos, h_mass)
            plan_e2e.compute(p=p, out=h_out)
        else:
            # uploads on a copy stream, array by array; the NVLink all-gather of array k runs under the upload of array k+1
            srcs = (h_edge, h_pos, h_mass)
            copy_stream.wait_stream(stream)            # (the previous step's all-gathers have read d_part)
            with torch.cuda.stream(copy_stream):
                for k in up_order:
                    d_part[k][: hi - lo].copy_(srcs[k][lo:hi], non_blocking=True)
                    up_events[k].record(copy_stream)
            for k in up_order:
                stream.wait_event(up_events[k])
                dist.all_gather_into_tensor(d_full[k], d_part[k])
            plan.set_masses_device(h_off.numpy(), d_full[0], d_full[1], d_full[2])
            plan.compute_shard(d_packed, rank, world, p=p)
            dist.all_gather_into_tensor(d_gath, d_packed)
            plan.unpack_shards(d_gath, d_out, world)
            if rank == 0:
                h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()

    def timed(fn, steps, warmup):
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

    sampler = ClockSampler(local) if rank == 0 else None
    ms_step = timed(step_device, args.steps, max(args.warmup, 3))
    st = plan.stats()
    # clocks: nvidia-smi samples every 20 ms; the timed region is tens of ms, so the same step keeps running (untimed) until
    # the sampler has seen ~0.4 s of this load.  It is stopped before the e2e region: polling the driver that often delays
    # the host-side submissions the e2e path depends on (measured +0.9 ms per step).
    for _ in range(int(min(2000, max(1, 400.0 / max(ms_step, 0.05))))):    # (same count on every rank: ms_step is the max over ranks)
        step_device()
    torch.cuda.synchronize()
    clocks = sampler.stop() if sampler else None
    ms_e2e = timed(step_e2e, max(2, min(args.steps, 5)), 3)

    if rank == 0:
        # roofline of the dominant kernel (k_pairs): FP64 pipe
        S = st["n_segments"]
        share = 1.0 / world
        flops = 3.0 * S * pairs * share                      # 3 flop per (pair, segment)
        t_pairs = st["ms_pairs"] * 1e-3
        peak_dfma = gappa_b200.measure_fp64_peak(local, 0, 200.0)
        peak_mix = gappa_b200.measure_fp64_peak(local, 1, 200.0)
        peak_dmma = gappa_b200.measure_fp64_peak(local, 2, 200.0)
        peak = max(peak_dfma, peak_dmma)
        achieved = flops / t_pairs / 1e12
        T = st["ld_samples"] // 128                          # executed: full tiles + 5/8 of each diagonal tile
        executed = 3.0 * ((S + 15) // 16 * 16) * 128.0 * 128.0 * ((st["n_tiles"] - T) + 0.625 * T) * share
        line = {
            "metric": METRIC, "value": pairs / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": spec["name"], "samples": N, "edges": spec["edges"], "placements_per_sample": spec["placements"],
                       "bins": bins, "p": p, "segments": int(S), "pairs": int(pairs), "masses": int(st["n_masses"]),
                       "l2": "inputs larger than L2 (F %.0f MB + split workspace)" % (st["f_bytes"] / 1e6),
                       "parallelism": (f"upper-triangular tiles/{world}, F replicated, NCCL all-gather of result tiles; "
                                       + (f"weak scaling: samples = {workloads.WORKLOADS[args.workload]['samples']}*sqrt({world}) "
                                          "rounded to 128" if scaling == "weak" else "strong scaling: cohort fixed by --samples"))
                                      if world > 1 else "1 gpu"},
            "e2e": {"value": pairs / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "ingest": "pipelined (krd_config.pipeline=1): up to 8 sample blocks, upload overlapped with K0/K1/pair phases" if world == 1
                              else "sharded upload (copy stream) + NVLink all-gather of the entries, array by array, overlapped",
                    "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(N * N * 8)},
            "gpu_launches": int(st["launches"]) * args.steps,
            "clocks": clocks,
            "roofline": {"bound": "fp64", "kernel": "k_pairs<p=%g>" % p, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak, "traffic": traffic_from_profile(spec, world),
                         "peak_source": "measured live: max(DFMA loop %.1f, DMMA m8n8k4 loop %.1f) TFLOP/s; MEASURED_PEAKS.json has no FP64 entry"
                                        % (peak_dfma, peak_dmma),
                         "mix_ceiling": peak_mix, "frac_of_mix_ceiling": achieved / peak_mix,
                         "executed_tflops": executed / t_pairs / 1e12,
                         "ms_kernel": st["ms_pairs"], "ms_build_F": st["ms_build"], "ms_epilogue": st["ms_epilogue"],
                         "ctas": st["pair_ctas"], "split": st["split"]},
        }
        # second kernel group of the step: the F builder (K1) is HBM-bound; algorithmic bytes per build:
        # F written once + the tile-major entries read twice (10 B each) + ~5 passes over the E x ld subtree matrices
        try:
            hbm_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]); hbm_src = "MEASURED_PEAKS.json hbm_gbs"
        except (OSError, KeyError, ValueError):
            hbm_peak, hbm_src = 6550.0, "fallback (B200_PROFILING.md copy bandwidth)"
        k1_bytes = float(st["f_bytes"]) + 20.0 * st["n_masses"] + 40.0 * spec["edges"] * st["ld_samples"]
        k1_gbs = k1_bytes / (st["ms_build"] * 1e-3) / 1e9 if st["ms_build"] > 0 else 0.0
        line["roofline_build_F"] = {"bound": "hbm", "kernel": "K1 (k_tile<TOTALS>, k_subtree_level/top, k_tile<FILL>)", "achieved": k1_gbs,
                                    "peak": hbm_peak, "unit": "GB/s", "frac": k1_gbs / hbm_peak, "bytes": k1_bytes,
                                    "ms": st["ms_build"], "peak_source": hbm_src, "share_of_step": st["ms_build"] / ms_step}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_reference_rate(tree, masses, p, bins, args.cpu_seconds)
        print(json.dumps(line))
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

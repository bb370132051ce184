#!/usr/bin/env python
# -*- coding: utf-8 -*-
"""
bench.py -- Serra09 (OTI + CRP + Qmax) all-pairs throughput on a synthetic Da-TACOS-benchmark-shape
set (BASELINE.json configs[1]; SURVEY.md section 8(d) "C2"), one process per GPU.

    python bench.py --gpus 1 --steps 3 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the reference's CPU path (oracle port) on host cores

A "step" = one pass of the hot path over the whole pair grid: OTI -> cross-similarity ->
percentile thresholds -> binary CRP -> Qmax -> (all-gather of score vectors, N > 1) ->
N x N scatter/mirror -> length normalisation -> MR/MRR/MDR/MAP/Top-k.  The pair grid of the
FIXED 1000-track job is sharded across ranks by cost (strong scaling).  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "serra09_allpairs_comparisons_per_sec"
UNIT = "pairs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C2", choices=["C1", "C2", "C3", "C4", "C5"],
                    help="BASELINE.json config (SURVEY 8(d)): C1 Serra09 160 tracks, C2 Serra09 1000 tracks (default, the driver line), "
                         "C3 EarlyFusionTralie 1000 tracks, C4 LateFusionChen 1000 tracks, C5 SiMPle + FTM2D 15 000 tracks")
    ap.add_argument("--cpu-full", action="store_true", help="cpu_baseline over the WHOLE pair grid (C1: 12 720 pairs) instead of a bounded sample")
    ap.add_argument("--pairs-per-step", type=int, default=0, help="C3 / C5: pairs of the grid processed per step (0 = config default)")
    ap.add_argument("--shape", default=None, help="synth.SHAPES key (default: by --config)")
    ap.add_argument("--regime", default="L", choices=["S", "R", "L"],
                    help="track-length regime (SURVEY 8(d)); L = n~U[1600,2400] frames at the DP stage")
    ap.add_argument("--scale", type=float, default=1.0, help="DEBUG ONLY: shrink the number of works")
    ap.add_argument("--crp-mode", type=int, default=0)
    ap.add_argument("--easy", action="store_true", help="the round-1 cover model (retrieval saturates: MAP = 1); default is synth.HARD")
    ap.add_argument("--parity-pairs", type=int, default=64, help="random pairs of the grid re-computed by the oracle outside the timed region")
    ap.add_argument("--no-e2e", action="store_true", help="DEBUG ONLY: skip the end-to-end leg")
    ap.add_argument("--no-cpu", action="store_true", help="DEBUG ONLY: skip the cpu_baseline leg")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU work of the cpu_baseline sample")
    args = ap.parse_args()
    if args.shape is None:
        args.shape = {"C1": "covers80", "C5": "datacos15000"}.get(args.config, "datacos1000")
    return args


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=float(d["hbm_gbs"]), bf16=float(d["bf16_tflops"]), bf16_sus=float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                    src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback (B200_PROFILING.md)")


def measured_traffic():
    """DRAM bytes per algorithmic unit from the committed `ncu --set full` capture of this round
    (profiles/traffic.json, written by tools/ncu_traffic.py from dram__bytes_read.sum +
    dram__bytes_write.sum of one launch); scaled below to the units one bench launch processes."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p))
        except Exception:
            return None
    return None


def make_workload(args):
    """Downsampled chroma of every track (what Serra09.load_features returns) + labels."""
    from acoss_b200 import synth
    ds = synth.make_named(args.shape, args.regime, scale=args.scale, hard=not args.easy)
    fac = ds["fac"]
    return ds, fac


class ClockSampler(object):
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.path = os.path.join(tempfile.gettempdir(), "acoss_clocks_%d_%d.csv" % (os.getpid(), index))
        self.proc = None

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        try:
            self.proc.terminate()
            self.proc.wait(timeout=5)
            self.f.close()
            sm, mx, reasons = [], [], set()
            for line in open(self.path):
                p = [x.strip() for x in line.split(",")]
                if len(p) < 9:
                    continue
                try:
                    sm.append(float(p[1])); mx.append(float(p[2]))
                except ValueError:
                    continue
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], p[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            if sm:
                out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
            os.remove(self.path)
        except Exception:
            pass
        return out


def cpu_baseline_sample(tracks, all_pairs, target_seconds, seed=0, full=False, want_dmax=False):
    """The oracle port (single-threaded per pair like essentia) fanned out over all host threads
    on a bounded random sample of the workload's pairs (or, with full=True, on the whole grid)."""
    from oracle import oracle as O
    O.build()
    cores = os.cpu_count() or 1
    rng = np.random.default_rng(seed)
    if full:
        n, sample = len(all_pairs), all_pairs
    else:
        probe = all_pairs[rng.choice(len(all_pairs), size=min(len(all_pairs), cores), replace=False)]
        t0 = time.perf_counter()
        O.serra09_pairs(tracks, probe, n_threads=cores, want_dmax=want_dmax)
        t_probe = time.perf_counter() - t0
        per_round = max(t_probe, 1e-4)                      # seconds per `cores` pairs
        n = int(min(len(all_pairs), max(cores, cores * round(target_seconds / per_round))))
        sample = all_pairs[rng.choice(len(all_pairs), size=n, replace=False)]
    t0 = time.perf_counter()
    q, _, cells = O.serra09_pairs(tracks, sample, n_threads=cores, want_dmax=want_dmax)
    dt = time.perf_counter() - t0
    return dict(value=n / dt, unit=UNIT, cores=cores, kind="port",
                sample="%s of the %d-pair grid (%.1f s, %d threads, oracle/acoss_oracle.c -O2 "
                       "-ffp-contract=off%s); %.4f GCUPS" % ("ALL pairs" if full else "%d random pairs" % n, len(all_pairs), dt, cores,
                                                              ", Qmax + Dmax" if want_dmax else "", cells / dt / 1e9),
                seconds=dt, gcups=cells / dt / 1e9)


def add_cpu_snf(cpu, n_tracks, n_pairs):
    """C4: the N x N similarity network fusion of two matrices (oracle/ports.py, numpy) timed once and added to the
    extrapolated whole-grid CPU time."""
    from oracle import ports as P
    rng = np.random.default_rng(4)
    A = rng.random((n_tracks, n_tracks)).astype(np.float32) + 0.1
    B = rng.random((n_tracks, n_tracks)).astype(np.float32) + 0.1
    t0 = time.perf_counter()
    P.snf([A, B], K=20, niters=20, reg_diag=1)
    t_snf = time.perf_counter() - t0
    total = n_pairs / cpu["value"] + t_snf
    cpu = dict(cpu)
    cpu["sample"] += "; + SNF of two %d x %d matrices (numpy port) %.2f s; value = pairs / (extrapolated pair time + SNF)" % (n_tracks, n_tracks, t_snf)
    cpu["value"] = n_pairs / total
    return cpu


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path (oracle port: essentia is
    not installable here) with all host threads; each step = a bounded sample of the workload."""
    if rank != 0:
        return
    ds, fac = make_workload(args)
    from oracle import oracle as O
    tracks = [O.sync_median(h, fac) if fac > 1 else h for h in ds["hpcp"]]
    n = len(tracks)
    i, j = np.triu_indices(n, 1)
    all_pairs = np.stack([i, j], 1).astype(np.int32)
    per_step = max(2.0, min(30.0, 150.0 / max(1, args.steps + args.warmup)))
    vals, last = [], None
    for s in range(args.warmup + args.steps):
        last = cpu_baseline_sample(tracks, all_pairs, per_step, seed=s, want_dmax=args.config == "C4")
        if s >= args.warmup:
            vals.append(last)
    tot_pairs = sum(v["value"] * v["seconds"] for v in vals)
    tot_s = sum(v["seconds"] for v in vals)
    value = tot_pairs / tot_s
    line = {"impl": "reference", "metric": "latefusionchen_allpairs_comparisons_per_sec" if args.config == "C4" else METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / max(1, len(vals)), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, n, len(all_pairs)),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": last["cores"], "kind": "port", "sample": last["sample"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def workload_config(args, n_tracks, n_pairs):
    return {"workload": "%s-%s: %s all-pairs, %d synthetic %s HPCP tracks "
                        "(%s), %d pairs, m=9 tau=1 kappa=0.095" % (
                            args.config, args.regime,
                            "LateFusionChen (OTI+CRP+Qmax+Dmax, distance normalisation, SNF K=20 x 20 iterations)" if args.config == "C4"
                            else "Serra09 (OTI+CRP+Qmax)", n_tracks,
                            "covers80-shape" if args.shape == "covers80" else "Da-TACOS-benchmark-shape",
                            {"L": "n~U[1600,2400] frames at the DP stage, downsample_fac=1",
                             "R": "raw T~U[15000,25000], downsample_fac=40",
                             "S": "raw T~U[1600,2400], downsample_fac=40"}[args.regime], n_pairs),
            "shape": args.shape, "regime": args.regime, "tracks": n_tracks, "pairs": n_pairs, "cover_model": "easy" if args.easy else "hard",
            "l2_policy": "inputs larger than L2: every step re-streams the CRP working set (GBs) through HBM",
            "debug_scale": args.scale}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import acoss_b200
    from acoss_b200.algorithm_template import shard_pairs

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    ctx = acoss_b200.Context(local_rank)          # raises if the CUDA library / device is missing
    ds, fac = make_workload(args)
    raw = ds["hpcp"]
    tracks = ctx.median_downsample(raw, fac) if fac > 1 else raw
    tracks = [np.ascontiguousarray(t, np.float32) for t in tracks]
    n = len(tracks)
    lengths_np = np.array([t.shape[0] for t in tracks], np.int32)
    i, j = np.triu_indices(n, 1)
    all_pairs = np.stack([i, j], 1).astype(np.int32)
    P = len(all_pairs)
    cost = lengths_np[all_pairs[:, 0]].astype(np.int64) * lengths_np[all_pairs[:, 1]]
    mine = shard_pairs(all_pairs, cost, rank, world)
    my_pairs = np.ascontiguousarray(all_pairs[mine])
    prm = ctx.params(crp_mode=args.crp_mode)

    # clique layout for the device evaluation (same construction as getEvalStatistics)
    cl = {}
    for k, l in enumerate(ds["labels"]):
        cl.setdefault(l, []).append(k)
    cliques = list(cl.values())
    Ks = np.array([len(c) for c in cliques])
    o = np.argsort(-Ks)
    cliques = [cliques[k] for k in o]
    Ks = Ks[o]
    order = np.concatenate(cliques).astype(np.int32)
    csize = np.repeat(Ks, Ks).astype(np.int32)
    cstart = np.repeat(np.concatenate([[0], np.cumsum(Ks)[:-1]]), Ks).astype(np.int32)
    tops = np.array([1, 10, 100, 1000], np.int32)

    # ---- resident state: features in HBM before the timed region
    frames, offs = ctx.pack_tracks(tracks)
    ctx.upload_tracks(frames=frames, offsets=offs)
    stream = torch.cuda.ExternalStream(ctx.lib.acoss_ctx_stream(ctx.h), device=dev)
    cap = (P + world - 1) // world + 1
    chen = args.config == "C4"
    d_scores_local = torch.zeros(cap, dtype=torch.float32, device=dev)
    d_dmax_local = torch.zeros(cap, dtype=torch.float32, device=dev) if chen else None
    d_idx_local = torch.full((cap,), -1, dtype=torch.int64, device=dev)
    d_idx_local[:len(mine)] = torch.from_numpy(mine).to(dev)
    d_pairs_all = torch.from_numpy(all_pairs).to(dev)
    d_scores_all = torch.zeros(P, dtype=torch.float32, device=dev)
    d_D = torch.zeros((n, n), dtype=torch.float32, device=dev)
    d_D2 = torch.zeros((n, n), dtype=torch.float32, device=dev) if chen else None
    d_Df = torch.zeros((n, n), dtype=torch.float32, device=dev) if chen else None
    d_dmax_all = torch.zeros(P, dtype=torch.float32, device=dev) if chen else None
    d_len = torch.from_numpy(lengths_np).to(dev)
    if world > 1:
        g_scores = [torch.empty_like(d_scores_local) for _ in range(world)]
        g_idx = [torch.empty_like(d_idx_local) for _ in range(world)]
        dist.all_gather(g_idx, d_idx_local)          # the pair->rank map is static: gathered once
    torch.cuda.synchronize(dev)
    stats_acc = []
    launches = [0]
    result = {}

    def step_resident():
        ctx.serra09_pairs_dev(my_pairs, prm, d_scores_local.data_ptr(), d_dmax_local.data_ptr() if chen else 0)     # syncs its stream at the end
        st = ctx.pair_stats()
        launches[0] += st["n_launches"]
        if world > 1:
            for loc, dst in ((d_scores_local, d_scores_all),) + (((d_dmax_local, d_dmax_all),) if chen else ()):
                dist.all_gather(g_scores, loc)
                for gi, gs in zip(g_idx, g_scores):
                    ok = gi >= 0
                    dst[gi[ok]] = gs[ok]
                launches[0] += 2 * world
            torch.cuda.synchronize(dev)
            src, src_d = d_scores_all, d_dmax_all
        else:
            src, src_d = d_scores_local, d_dmax_local      # `mine` is the identity permutation for world == 1
        ctx.scores_to_matrix_dev(d_pairs_all.data_ptr(), src.data_ptr(), P, n, True, d_D.data_ptr())
        ctx.normalize_by_length_dev(d_D.data_ptr(), n, d_len.data_ptr(), 1 if chen else 0)
        launches[0] += 2
        if chen:
            # LateFusionChen (latefusion_chen.py:75-91): both matrices become distances sqrt(len_j) / score, SNF (K = 20, 20
            # iterations) of the two, evaluation of -qmax-distance, -dmax-distance and the fused matrix
            ctx.scores_to_matrix_dev(d_pairs_all.data_ptr(), src_d.data_ptr(), P, n, True, d_D2.data_ptr())
            ctx.normalize_by_length_dev(d_D2.data_ptr(), n, d_len.data_ptr(), 1)
            ctx.synchronize()
            _, fused = ctx.snf([d_D.cpu().numpy(), d_D2.cpu().numpy()], K=20, niters=20, reg_diag=1, want_Ws=False)
            d_Df.copy_(torch.from_numpy(fused.astype(np.float32)))
            torch.cuda.synchronize(dev)
            result["eval_dmax"], _ = ctx.eval_statistics((-d_D2).data_ptr(), n, order, csize, cstart, tops)
            result["eval_late"], _ = ctx.eval_statistics(d_Df.data_ptr(), n, order, csize, cstart, tops)
            d_Dn = -d_D
            torch.cuda.synchronize(dev)
            out, ranks = ctx.eval_statistics(d_Dn.data_ptr(), n, order, csize, cstart, tops)
            launches[0] += 2 + 8 + 3
        else:
            out, ranks = ctx.eval_statistics(d_D.data_ptr(), n, order, csize, cstart, tops)
            launches[0] += 1
        result["eval"] = out
        return st

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        ctx.synchronize()

    for _ in range(args.warmup):
        step_resident()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    launches[0] = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        stats_acc.append(step_resident())
    e1.record(stream)
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    dev_ms = e0.elapsed_time(e1)
    n_launch = launches[0]
    # max over ranks of the device-event time of the timed region
    t = torch.tensor([dev_ms, wall * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms = float(t[0]), float(t[1])
    ms_per_step = dev_ms / args.steps
    value = P / (ms_per_step * 1e-3)

    # ---- end-to-end leg: reference-facing class API with HOST buffers, H2D/D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, ctx, ds, tracks, all_pairs, mine, world, rank, local_rank)

    # ---- per-kernel roofline (this rank; DP kernel = HBM-bound at 1 B/cell)
    pk = peaks()
    ms_dp = float(np.mean([s["ms_dp"] for s in stats_acc]))
    ms_crp = float(np.mean([s["ms_crp"] for s in stats_acc]))
    ms_tot = float(np.mean([s["ms_total"] for s in stats_acc]))
    st = stats_acc[-1]
    n_dp_launch = max(1, st["n_dp_launches"])
    n_dp_kinds = 2 if chen else 1
    dp_bytes = n_dp_kinds * (st["crp_cells"] + 4 * len(mine))              # ALGORITHMIC: M*N CRP bytes read once + 4 B score per pair
    dp_gbs = dp_bytes / (ms_dp * 1e-3) / 1e9
    dp_streamed = n_dp_kinds * (st["crp_bytes"] + 4 * len(mine))           # what the kernel streams: M*pitch (row pitch padded to the DP chunk)
    gcups = st["dp_cells"] / (ms_dp * 1e-3) / 1e9
    crp_flop = 2.0 * 108 * st["crp_cells"] + 5.0 * st["crp_cells"]
    crp_tflops = crp_flop / (ms_crp * 1e-3) / 1e12
    tr = measured_traffic() or {}
    split = "split" in st.get("crp_kernels", [])
    # CRP launches of one step: two strip launches per batch (monolithic kernels) or the phase-split pipeline's
    # pivot / count / select per chunk + edges + binarise (everything acoss_serra09_pairs launched except the DP and OTI kernels)
    n_crp_launch = max(1, (st["n_launches"] - n_dp_launch - st["n_batches"]) if split else 2 * st["n_batches"])
    dp_traffic = tr["dp_dram_bytes_per_crp_byte"] * st["crp_bytes"] / n_dp_launch if "dp_dram_bytes_per_crp_byte" in tr else None
    crp_traffic = tr["crp_dram_bytes_per_cell"] * st["crp_cells"] / n_crp_launch if "crp_dram_bytes_per_cell" in tr else None
    kernels = [
        {"kernel": "dp_rowpar_kernel (Qmax)", "bound": "hbm", "achieved": dp_gbs, "peak": pk["hbm"], "unit": "GB/s",
         "frac": dp_gbs / pk["hbm"], "traffic": dp_traffic, "ms_per_step": ms_dp, "launches_per_step": n_dp_launch,
         "streamed_bytes_per_launch": dp_streamed / n_dp_launch, "streamed_over_algorithmic": dp_streamed / dp_bytes,
         "gcups": gcups, "share_of_step": ms_dp / ms_tot},
        {"kernel": ("crp_pivot / crp_stream<0> / crp_select / crp_stream<1> kernels (phase-split: cross-similarity + thresholds + binarise)" if split
                    else "crp_strip_gemm_kernel x2 (cross-similarity + thresholds + binarise)" if "gemm" in st.get("crp_kernels", [])
                    else "crp_strip kernels (cross-similarity + thresholds + binarise)"),
         "bound": "tensor", "achieved": crp_tflops,
         "peak": pk["bf16_sus"], "unit": "TFLOP/s", "frac": crp_tflops / pk["bf16_sus"], "traffic": crp_traffic,
         "traffic_note": "DRAM bytes per launch: ncu dram bytes per CRP cell over all CRP launches (profiles/traffic.json) x cells / launches per step; "
                         "the CRP itself is 1 B per cell, the rest is the list scratch between the count and the selection kernels",
         "launches_per_step": n_crp_launch, "ms_per_step": ms_crp, "share_of_step": ms_crp / ms_tot,
         # what the tensor pipe actually executes per cell and streaming pass (column count, row count, binarise; tile padding and
         # the sample MMAs not counted): monolithic kernel 41 TF32 MMAs of 128x256x8 per 128x256 tile = 656 flop; phase-split
         # pipeline 14 TF32 MMAs + 14 bf16 MMAs of 128x256x16 = 672 flop.  MMA instructions issue at 128 clk each at best
         # (tools/mma_probe.py), so the tensor pipe's own floor for the three passes is mma_floor_ms.
         "executed_tflops": 3.0 * (672.0 if split else 656.0) * st["crp_cells"] / (ms_crp * 1e-3) / 1e12,
         "mma_per_tile_and_pass": 28 if split else 41,
         "mma_floor_ms": 3.0 * (28 if split else 41) * 128.0 * (st["crp_cells"] / (128.0 * 256.0)) / (148.0 * (clocks.get("sm_mhz") or 1965.0) * 1e3),
         "note": "algorithmic flop = (2*108+5)*M*N per pair (SURVEY 8(d)), peak = measured dense bf16. The kernels run the "
                 "contraction three times (column thresholds, row thresholds, binarise) on tcgen05 -- hi.hi in TF32, the cross "
                 "products x_hi.y_lo + x_lo.y_hi in bf16 (phase-split pipeline) or TF32 (monolithic kernel) -- and decide "
                 "everything near a threshold with exact fp32 chains; see DESIGN.md 4.1"},
    ]
    dominant = max(kernels, key=lambda k: k["ms_per_step"])
    roofline = {k: dominant[k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")}
    roofline["kernel"] = dominant["kernel"]
    roofline["peak_source"] = pk["src"]
    roofline["traffic_source"] = tr.get("source")

    # ---- parity sample (outside the timed region): random pairs of THIS rank's shard, CUDA scores vs the oracle
    parity = None
    if args.parity_pairs > 0:
        from oracle import oracle as O
        O.build()
        rng = np.random.default_rng(20261105 + rank)
        pick = np.sort(rng.choice(len(mine), size=min(args.parity_pairs, len(mine)), replace=False))
        got = d_scores_local[torch.from_numpy(pick).to(dev)].cpu().numpy()
        exp, _, _ = O.serra09_pairs(tracks, my_pairs[pick], n_threads=os.cpu_count() or 1)
        parity = {"pairs": int(len(pick)), "mismatches": int(np.sum(got != exp)),
                  "checker": "oracle/acoss_oracle.c (C restatement; essentia itself is not installable: parity unpinned)"}
        if world > 1:
            tt = torch.tensor([parity["pairs"], parity["mismatches"]], dtype=torch.int64, device=dev)
            dist.all_reduce(tt)
            parity["pairs"], parity["mismatches"] = int(tt[0]), int(tt[1])

    if rank != 0:
        return
    cpu = None
    if world == 1 and not args.no_cpu:
        cpu = cpu_baseline_sample(tracks, all_pairs, args.cpu_seconds, full=args.cpu_full, want_dmax=chen)
        if chen:
            cpu = add_cpu_snf(cpu, n, len(all_pairs))
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
    ev = result["eval"]
    line = {"metric": "latefusionchen_allpairs_comparisons_per_sec" if chen else METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args, n, P),
            "dp_gcups": gcups * world if world > 1 else gcups, "wall_ms_per_step": wall_ms / args.steps,
            "roofline": roofline, "roofline_kernels": kernels, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": int(n_launch), "clocks": clocks,
            "eval": {"MR": ev[0], "MRR": ev[1], "MDR": ev[2], "MAP": ev[3], "Top-1": ev[4]},
            "eval_late_fusion": ({"MAP_dmax": result["eval_dmax"][3], "MAP_late": result["eval_late"][3]} if chen else None),
            "parity_sample": parity,
            "batches_per_step": int(st["n_batches"]), "crp_mode": args.crp_mode,
            "debug_counters": ctx.debug_counters(),
            "exact_cell_fraction": st["n_exact_cells"] / max(1, 2 * st["crp_cells"]), "fallback_rows": int(st["n_fallback_rows"])}
    print(json.dumps(line))


def run_e2e(args, ctx, ds, tracks, all_pairs, mine, world, rank, local_rank):
    """Same metric through the reference-facing call: host numpy features in, host score matrix and
    eval tuple out.  Every step: H2D of all features + pair list, D2H of the scores, N x N matrix
    round trip for the normalisation and the evaluation (what Serra09.all_pairwise /
    normalize_by_length / getEvalStatistics do)."""
    import contextlib
    import io
    import torch
    import torch.distributed as dist
    import acoss_b200
    from acoss_b200 import synth
    dev = torch.device("cuda", local_rank)
    tmp = tempfile.mkdtemp(prefix="acoss_bench_r%d_" % rank)
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        # feature files are not re-read inside the timed region (disk is out of scope): the class's
        # host-side cache is primed, the device upload is forced every step
        csv = os.path.join(tmp, "ds.csv")
        with open(csv, "w") as f:
            f.write("work_id,track_id\n")
            for w, t in zip(ds["work_ids"], ds["track_ids"]):
                f.write("%s,%s\n" % (w, t))
        sink = io.StringIO()
        chen = args.config == "C4"
        cls = acoss_b200.ChenFusion if chen else acoss_b200.Serra09
        with contextlib.redirect_stdout(sink):
            s = cls(csv, os.path.join(tmp, "feat") + "/", shortname="bench", downsample_fac=ds["fac"],
                    cachedir=os.path.join(tmp, "cache"), device=local_rank)
        s._ctx = ctx
        for k, t in enumerate(tracks):
            s.all_feats[k] = t
            s.cliques.setdefault(ds["labels"][k], set([])).add(k)
        n = len(tracks)
        nm = 2 if chen else 1               # score matrices
        h2d = sum(t.nbytes for t in tracks) + 8 * (n + 1) + all_pairs[mine].nbytes + nm * 2 * 4 * n * n + 4 * n + (nm * 4 * n * n if chen else 0)
        d2h = nm * 4 * len(mine) + nm * 4 * n * n + 8 * n + (8 * n * n if chen else 0)
        times = []
        for it in range(1 + args.steps):          # 1 warm-up
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(sink):
                s._uploaded = False
                if chen:
                    s.Ds.pop("Late", None)
                    for key in ("qmax", "dmax"):
                        s.Ds[key][:, :] = 0
                else:
                    s.Ds["main"][:, :] = 0
                s.all_pairwise(1, symmetric=True)
                s.normalize_by_length()
                if chen:
                    s.do_late_fusion()
                res = s.getEvalStatistics("Late" if chen else "main")
            torch.cuda.synchronize(dev)
            if world > 1:
                dist.barrier()
            if it > 0:
                times.append(time.perf_counter() - t0)
        tt = torch.tensor([float(np.mean(times))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sec = float(tt[0])
        s.cleanup_memmap()
        return {"value": len(all_pairs) / sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": sec * 1e3,
                "api": ("ChenFusion.all_pairwise + normalize_by_length + do_late_fusion + getEvalStatistics('Late')" if chen
                        else "Serra09.all_pairwise + normalize_by_length + getEvalStatistics"),
                "MAP": res[3]}
    finally:
        os.chdir(cwd)
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)


# =====================================================================================================================
# C3 (EarlyFusionTralie) and C5 (SiMPle + FTM2D): the other BASELINE.json configs, same JSON schema
# =====================================================================================================================
def _clique_layout(labels):
    cl = {}
    for k, l in enumerate(labels):
        cl.setdefault(l, []).append(k)
    cliques = list(cl.values())
    Ks = np.array([len(c) for c in cliques])
    o = np.argsort(-Ks)
    cliques = [cliques[k] for k in o]
    Ks = Ks[o]
    order = np.concatenate(cliques).astype(np.int32)
    csize = np.repeat(Ks, Ks).astype(np.int32)
    cstart = np.repeat(np.concatenate([[0], np.cumsum(Ks)[:-1]]), Ks).astype(np.int32)
    return order, csize, cstart


def _c3_workload(args):
    from acoss_b200 import synth
    nw, pw, ns = synth.SHAPES["datacos1000"]
    if args.scale != 1.0:
        nw, ns = max(1, int(round(nw * args.scale))), int(round(ns * args.scale))
    return synth.make_block_features(nw, pw, ns, 1001, 300, 600, noise=2.0)


def _c5_workload(args):
    """SiMPle features (12, L) float64 with L ~ U[150, 250], columns L2-normalised, covers share a warped latent; FTM2D
    shingles (900,) unit vectors."""
    from acoss_b200 import synth
    nw, pw, ns = synth.SHAPES["datacos15000"]
    if args.scale != 1.0:
        nw, ns = max(1, int(round(nw * args.scale))), int(round(ns * args.scale))
    rng = np.random.default_rng(15000)
    feats, shingles, labels = [], [], []
    for w in range(nw + ns):
        lat = np.abs(np.cumsum(rng.standard_normal((12, 270)), 1)) + 0.05
        sh = rng.random(900)
        for c in range(pw if w < nw else 1):
            L = int(rng.integers(150, 251))
            s0 = int(rng.integers(0, 270 - L + 1))
            f = np.roll(lat[:, s0:s0 + L], int(rng.integers(0, 12)), axis=0) + 0.3 * rng.random((12, L))
            feats.append(f / np.sqrt(np.sum(f * f, 0, keepdims=True)))
            v = sh + 0.2 * rng.random(900)
            shingles.append(v / np.sqrt(np.sum(v * v)))
            labels.append("W_%d" % w)
    return feats, np.stack(shingles), labels


def _pool_rate(fn, items, cores, target_seconds):
    """items / second of fn over a fork pool with `cores` workers on a bounded prefix of `items` (about target_seconds)."""
    import multiprocessing as mp
    t0 = time.perf_counter()
    for it in items[:2]:
        fn(it)
    per = max((time.perf_counter() - t0) / 2, 1e-5)
    n = int(min(len(items), max(cores, cores * target_seconds / per)))
    with mp.get_context("fork").Pool(cores, initializer=_one_blas_thread) as pool:
        pool.map(fn, items[:cores], chunksize=1)          # start-up outside the timed region
        t0 = time.perf_counter()
        pool.map(fn, items[:n], chunksize=max(1, n // (8 * cores)))
        dt = time.perf_counter() - t0
    return n / dt, n, dt


_POOL_STATE = {}


def _one_blas_thread():
    """one BLAS thread per worker process: `cores` processes x `cores` BLAS threads would oversubscribe the host"""
    try:
        from threadpoolctl import threadpool_limits
        _POOL_STATE["_limit"] = threadpool_limits(1)
    except Exception:
        pass


def _c3_cpu_one(pr):
    from oracle import ports as P
    bf = _POOL_STATE["bf"]
    return P.early_pair(bf[pr[0]], bf[pr[1]], 0.1, 10)[3]


def _c5_cpu_one(pr):
    from oracle import ports as P
    f = _POOL_STATE["feats"]
    return P.simple_pair(f[pr[0]], f[pr[1]])


def _finish_line(args, world, metric, unit, value, ms_per_step, workload, roofline, kernels, cpu, e2e, launches, clocks, dtype, extra):
    line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": dtype,
            "data": "synthetic", "config": workload, "roofline": roofline, "roofline_kernels": kernels, "cpu_baseline": cpu,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}
    line.update(extra)
    print(json.dumps(line))


def run_c3(args, rank, world, local_rank):
    """EarlyFusionTralie on 1000 synthetic tracks of 300-600 blocks (mfccs 1000-d, ssms 1225-d, chromas 480-d): per pair 3
    CSMs, 4 kNN binarisations, WCSM early fusion, 4 constrained Smith-Waterman; then the two N x N similarity network fusions
    and the evaluation of all six matrices.  A step = `pairs_per_step` pairs of the grid (default: the whole grid)."""
    import torch
    import torch.distributed as dist
    import acoss_b200
    from acoss_b200.algorithm_template import shard_pairs
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    ctx = acoss_b200.Context(local_rank)
    ds = _c3_workload(args)
    bf = ds["block_feats"]
    n = len(bf)
    nb = np.array([f["mfccs"].shape[0] for f in bf], np.int64)
    i, j = np.triu_indices(n, 1)
    all_pairs = np.stack([i, j], 1).astype(np.int32)
    if args.pairs_per_step and args.pairs_per_step < len(all_pairs):
        all_pairs = all_pairs[np.sort(np.random.default_rng(3).choice(len(all_pairs), args.pairs_per_step, replace=False))]
    P = len(all_pairs)
    mine = shard_pairs(all_pairs, nb[all_pairs[:, 0]] * nb[all_pairs[:, 1]], rank, world)
    my_pairs = np.ascontiguousarray(all_pairs[mine])
    order, csize, cstart = _clique_layout(ds["labels"])
    tops = np.array([1, 10, 100, 1000], np.int32)
    ctx.earlyfusion_upload(bf)
    h2d_feat = sum(f[k].nbytes for f in bf for k in ("mfccs", "ssms", "chromas", "chroma_med"))
    stream = torch.cuda.ExternalStream(ctx.lib.acoss_ctx_stream(ctx.h), device=dev)
    res = {}

    def nxn(scores_local):
        """(4, P) scores of all ranks -> four N x N matrices, the two late fusions, six evaluations."""
        full = np.zeros((4, P), np.float64)
        full[:, mine] = scores_local.T
        if world > 1:
            t = torch.from_numpy(full).to(dev)
            dist.all_reduce(t)
            full = t.cpu().numpy()
        Ds = {}
        for k, s_ in enumerate(("mfccs", "ssms", "chromas", "early")):
            D = np.zeros((n, n), np.float32)
            D[all_pairs[:, 0], all_pairs[:, 1]] = full[k]
            Ds[s_] = D + D.T
        Ds["late"] = ctx.snf([1.0 / (1.0 + Ds[s_]) for s_ in ("chromas", "ssms", "mfccs")], K=20, niters=20, reg_diag=1, want_Ws=False)[1]
        Ds["early+late"] = ctx.snf([1.0 / (1.0 + Ds[s_]) for s_ in ("chromas", "ssms", "mfccs", "early")], K=20, niters=20, reg_diag=1, want_Ws=False)[1]
        ev = {}
        for s_, D in Ds.items():
            Dd = torch.from_numpy(np.ascontiguousarray(D, np.float32)).to(dev)
            torch.cuda.synchronize(dev)
            ev[s_] = ctx.eval_statistics(Dd.data_ptr(), n, order, csize, cstart, tops)[0]
        return ev

    def step():
        sc = ctx.earlyfusion_pairs(my_pairs, kappa=0.1, K=10)
        res["stats"] = ctx.earlyfusion_stats()
        res["eval"] = nxn(sc)
        res["scores"] = sc

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        ctx.synchronize()

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    t0 = time.perf_counter()
    acc = []
    for _ in range(args.steps):
        step()
        acc.append(res["stats"])
    e1.record(stream)
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    tt = torch.tensor([max(e0.elapsed_time(e1), wall * 1e3)], dtype=torch.float64, device=dev)      # the call copies host buffers: wall bounds it
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    ms_per_step = float(tt[0]) / args.steps
    value = P / (ms_per_step * 1e-3)
    # e2e: the class-level call with host features every step (upload inside the timed region)
    e2e = None
    if not args.no_e2e:
        times = []
        for it in range(2):
            barrier()
            t0 = time.perf_counter()
            ctx.earlyfusion_upload(bf)
            sc = ctx.earlyfusion_pairs(my_pairs, kappa=0.1, K=10)
            ev = nxn(sc)
            barrier()
            if it:
                times.append(time.perf_counter() - t0)
        tt = torch.tensor([float(np.mean(times))], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = {"value": P / float(tt[0]), "unit": UNIT, "h2d_bytes_per_step": int(h2d_feat + my_pairs.nbytes + 7 * 8 * n * n),
               "d2h_bytes_per_step": int(32 * len(my_pairs) + 2 * 8 * n * n), "ms_per_step": float(tt[0]) * 1e3,
               "api": "acoss_earlyfusion_upload + acoss_earlyfusion_pairs + acoss_snf x2 + acoss_eval_statistics x6 (EarlyFusion.similarity / do_late_fusion / getEvalStatistics)",
               "MAP_early+late": float(ev["early+late"][3])}
    pk = peaks()
    st = acc[-1]
    ms_csm, ms_sel, ms_sw = (float(np.mean([a[k] for a in acc])) for k in ("ms_csm", "ms_select", "ms_sw"))
    cells = float(st["cells"])
    tfl = st["gemm_flop"] / (ms_csm * 1e-3) / 1e12
    sw_bytes = 4.0 * cells          # four uint8 binary matrices read once
    kernels = [
        {"kernel": "ef_csm kernels (3 cross-similarity GEMMs, d = 1000 / 1225 / 480)", "bound": "tensor", "achieved": tfl, "peak": pk["bf16_sus"],
         "unit": "TFLOP/s", "frac": tfl / pk["bf16_sus"], "traffic": None, "ms_per_step": ms_csm, "share_of_step": ms_csm / ms_per_step},
        {"kernel": "ef_row / ef_col / ef_early kernels (kNN binarise, WCSM, early fusion)", "bound": "hbm", "achieved": 16.0 * cells / (ms_sel * 1e-3) / 1e9,
         "peak": pk["hbm"], "unit": "GB/s", "frac": 16.0 * cells / (ms_sel * 1e-3) / 1e9 / pk["hbm"], "traffic": None, "ms_per_step": ms_sel,
         "share_of_step": ms_sel / ms_per_step, "note": "algorithmic bytes 3*4*M*N + 4*M*N per pair (SURVEY 8(d))"},
        {"kernel": "sw_constrained kernel x4", "bound": "hbm", "achieved": sw_bytes / (ms_sw * 1e-3) / 1e9, "peak": pk["hbm"], "unit": "GB/s",
         "frac": sw_bytes / (ms_sw * 1e-3) / 1e9 / pk["hbm"], "traffic": None, "ms_per_step": ms_sw, "share_of_step": ms_sw / ms_per_step,
         "gcups": 4.0 * cells / (ms_sw * 1e-3) / 1e9},
    ]
    dom = max(kernels, key=lambda k: k["ms_per_step"])
    roofline = {k: dom[k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")}
    roofline["kernel"] = dom["kernel"]
    roofline["peak_source"] = pk["src"]
    if rank != 0:
        return
    cpu = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        _POOL_STATE["bf"] = bf
        rate, ns_, dt = _pool_rate(_c3_cpu_one, [tuple(p) for p in all_pairs[np.random.default_rng(0).permutation(P)[:4096]]], cores, args.cpu_seconds)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d random pairs of the %d-pair grid (%.1f s, %d processes, oracle/ports.py early_pair: numpy CSMs / kNN / WCSM + the C "
                         "Smith-Waterman; the N x N fusion is not included)" % (ns_, P, dt, cores)}
        # parity sample: CUDA scores vs the port on a few pairs (float GEMM order differs: compare to 1e-4 relative)
        from oracle import ports as PT
        pick = np.random.default_rng(1).choice(len(mine), size=min(8, len(mine)), replace=False)
        exp = np.stack([PT.early_pair(bf[a], bf[b], 0.1, 10) for a, b in my_pairs[pick]])
        got = res["scores"][pick]
        parity = {"pairs": int(len(pick)), "mismatches": int(np.sum(np.abs(got - exp) > 1e-4 * np.maximum(1.0, np.abs(exp)))),
                  "checker": "oracle/ports.py (numpy port pinned against the reference run verbatim)"}
    else:
        parity = None
    workload = {"workload": "C3: EarlyFusionTralie all-pairs on %d synthetic tracks of 300-600 blocks (mfccs 50x20, ssms 1225, chromas 480), "
                            "%d pairs per step, kappa=0.1 K=10; SNF K=20 x 20 iterations on 3 and 4 N x N matrices" % (n, P),
                "tracks": n, "pairs": P, "debug_scale": args.scale,
                "l2_policy": "inputs larger than L2: %.1f GB of block features, 4 x M x N float32 scratch per pair" % (h2d_feat / 1e9)}
    ev = res["eval"]
    _finish_line(args, world, "earlyfusion_allpairs_comparisons_per_sec", UNIT, value, ms_per_step, workload, roofline, kernels, cpu, e2e,
                 (4 + 8) * args.steps, clocks, "f32",
                 {"eval": {k: {"MR": v[0], "MAP": v[3]} for k, v in ev.items()}, "parity_sample": parity,
                  "gpu_launches_note": "per step: >= 12 kernels per batch of pairs (3 CSM, 4 row-kNN, 1 column, 1 early, 4 SW) + SNF + eval"})


def run_c5(args, rank, world, local_rank):
    """SiMPle on 15 000 synthetic tracks (12 x L float64, L ~ U[150, 250]): a step = `pairs_per_step` ORDERED pairs of the
    2.25e8-pair grid (default 2^24, a bounded sample: the whole grid is 100 s per step); FTM2D 15 000 x 900 all-pairs once per step."""
    import torch
    import torch.distributed as dist
    import acoss_b200
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    ctx = acoss_b200.Context(local_rank)
    feats, shingles, labels = _c5_workload(args)
    n = len(feats)
    P = args.pairs_per_step or (1 << 24)
    rng = np.random.default_rng(5)
    P = int(min(P, n * (n - 1)))
    a = rng.integers(0, n, size=P).astype(np.int32)
    b = (a + 1 + rng.integers(0, n - 1, size=P).astype(np.int32)) % n
    all_pairs = np.stack([a, b], 1).astype(np.int32)
    mine = np.arange(rank, P, world)
    my_pairs = np.ascontiguousarray(all_pairs[mine])
    ctx.simple_upload(feats)
    stream = torch.cuda.ExternalStream(ctx.lib.acoss_ctx_stream(ctx.h), device=dev)
    res = {}

    def step():
        res["scores"] = ctx.simple_pairs(my_pairs, 10)
        t0 = time.perf_counter()
        res["ftm"] = ctx.ftm2d_allpairs(shingles) if rank == 0 else None
        res["t_ftm"] = time.perf_counter() - t0

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        ctx.synchronize()

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    t_ftm = []
    for _ in range(args.steps):
        step()
        t_ftm.append(res["t_ftm"])
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    tt = torch.tensor([wall], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    sec_step = float(tt[0]) / args.steps
    t_f = float(np.mean(t_ftm))
    sec_simple = max(sec_step - t_f, 1e-9)
    value = P / sec_simple
    La = np.array([f.shape[1] for f in feats], np.int64)
    flop = 26.0 * float(np.sum(La[my_pairs[:, 0]] * La[my_pairs[:, 1]])) * world
    pk = peaks()
    fp64_peak = 40.0          # B200 FP64 (non-tensor) ~ 40 TFLOP/s nominal; no measured figure in MEASURED_PEAKS.json
    kernels = [
        {"kernel": "simple_pair_kernel (matrix-profile join + median)", "bound": "fp64", "achieved": flop / sec_simple / 1e12, "peak": fp64_peak,
         "unit": "TFLOP/s", "frac": flop / sec_simple / 1e12 / fp64_peak, "traffic": None, "ms_per_step": sec_simple * 1e3,
         "note": "algorithmic flop 26 * La * Lb per ordered pair (SURVEY 8(d)); peak = nominal FP64 FMA rate (fallback: not in MEASURED_PEAKS.json)"},
        {"kernel": "ftm2d_allpairs_kernel (15 000 x 900 x 15 000 fp64 Gram + exp epilogue)", "bound": "fp64", "achieved": 2.0 * 900 * n * n / t_f / 1e12,
         "peak": fp64_peak, "unit": "TFLOP/s", "frac": 2.0 * 900 * n * n / t_f / 1e12 / fp64_peak, "traffic": None, "ms_per_step": t_f * 1e3,
         "pairs_per_s": n * (n - 1) / 2 / t_f, "note": "includes the N x N float64 result copy to the host"},
    ]
    roofline = {k: kernels[0][k] for k in ("bound", "achieved", "peak", "unit", "frac", "traffic")}
    roofline["kernel"] = kernels[0]["kernel"]
    roofline["peak_source"] = "nominal FP64 (no measured FP64 peak on this pool)"
    if rank != 0:
        return
    cpu = parity = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        _POOL_STATE["feats"] = feats
        rate, ns_, dt = _pool_rate(_c5_cpu_one, [tuple(p) for p in all_pairs[:65536]], cores, args.cpu_seconds)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "%d ordered pairs of the step's %d (%.1f s, %d processes, oracle/ports.py simple_pair: numpy Gram + diagonal sums; the "
                         "reference's Python row loop is ~10x slower per pair)" % (ns_, P, dt, cores)}
        from oracle import ports as PT
        pick = np.random.default_rng(1).choice(len(mine), size=32, replace=False)
        exp = np.array([PT.simple_pair(feats[x], feats[y]) for x, y in my_pairs[pick]])
        got = res["scores"][pick]
        parity = {"pairs": 32, "mismatches": int(np.sum(np.abs(got - exp) > 1e-9 * np.maximum(1e-3, np.abs(exp)))),
                  "checker": "oracle/ports.py (float64; tolerance 1e-9 relative)"}
    e2e = {"value": value, "unit": UNIT, "h2d_bytes_per_step": int(my_pairs.nbytes), "d2h_bytes_per_step": int(8 * len(my_pairs)),
           "ms_per_step": sec_simple * 1e3, "api": "acoss_simple_pairs (host pair list in, host float64 scores out: the timed call itself)",
           "note": "features stay resident (Simple.preload uploads once per dataset); the pair list and the scores cross PCIe every step"}
    workload = {"workload": "C5: SiMPle matrix-profile similarity, %d ORDERED pairs per step sampled from the %d x %d grid of %d synthetic tracks "
                            "(12 x L float64, L~U[150,250], SSLEN 10); FTM2D %d x 900 all-pairs" % (P, n, n - 1, n, n),
                "tracks": n, "pairs": P, "debug_scale": args.scale, "l2_policy": "290 MB of features, random pairs: inputs larger than L2"}
    _finish_line(args, world, "simple_ordered_pairs_per_sec", UNIT, value, sec_simple * 1e3, workload, roofline, kernels, cpu, e2e,
                 2 * args.steps, clocks, "f64", {"parity_sample": parity, "ftm2d_seconds_per_step": t_f})


def run_reference_other(args, rank, world):
    """--impl reference for C3 / C5: the numpy ports (oracle/ports.py) on all host cores, bounded samples."""
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = max(2.0, min(30.0, 150.0 / max(1, args.steps + args.warmup)))
    if args.config == "C3":
        ds = _c3_workload(args)
        _POOL_STATE["bf"] = ds["block_feats"]
        n = len(ds["block_feats"])
        i, j = np.triu_indices(n, 1)
        items = [tuple(p) for p in np.stack([i, j], 1)[np.random.default_rng(0).permutation(len(i))[:8192]]]
        fn, metric, dtype, wl = _c3_cpu_one, "earlyfusion_allpairs_comparisons_per_sec", "f32", "C3: EarlyFusionTralie, %d tracks" % n
    else:
        feats, _, _ = _c5_workload(args)
        _POOL_STATE["feats"] = feats
        n = len(feats)
        rng = np.random.default_rng(5)
        a = rng.integers(0, n, size=65536)
        items = [(int(x), int((x + 1 + y) % n)) for x, y in zip(a, rng.integers(0, n - 1, size=65536))]
        fn, metric, dtype, wl = _c5_cpu_one, "simple_ordered_pairs_per_sec", "f64", "C5: SiMPle, %d tracks" % n
    rates = []
    for s_ in range(args.warmup + args.steps):
        r, ns_, dt = _pool_rate(fn, items, cores, per_step)
        if s_ >= args.warmup:
            rates.append((r, ns_, dt))
    value = sum(x[1] for x in rates) / sum(x[2] for x in rates)
    line = {"impl": "reference", "metric": metric, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * sum(x[2] for x in rates) / max(1, len(rates)), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": dtype, "data": "synthetic", "config": {"workload": wl},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": "%d pairs per step, oracle/ports.py" % rates[-1][1]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if args.config in ("C3", "C5"):
            run_reference_other(args, rank, world)
        else:
            run_reference(args, rank, world)
        return
    if args.config in ("C3", "C5"):
        if world > 1:
            import torch
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        try:
            (run_c3 if args.config == "C3" else run_c5)(args, rank, world, local_rank)
        finally:
            if world > 1:
                import torch.distributed as dist
                dist.destroy_process_group()
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()

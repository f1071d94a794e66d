#!/usr/bin/env python
"""bench.py - headline benchmark of the KZG hot path (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config.workload): BASELINE config 2 - KZG commit + open of one random
degree-2^20 polynomial (scale 20, machines_scale 0) on synthetic seeded inputs.
A "step" = one worker_commit + one worker_open of that polynomial.

  value       ms per step with the coefficients already resident in HBM, timed with
              CUDA events on the launching stream (max over ranks)
  e2e         the same through the public host API (PianoBackend.worker_commit /
              worker_open on pinned HOST buffers; H2D + D2H inside the timed region)
  roofline    dominant kernel (MSM bucket accumulation), timed live per launch
  cpu_baseline the C oracle (oracle/c/oracle.c) on the host cores, same inputs + SRS
  extra       G1 MSM Mpts/s and Fr NTT Gelem/s (the other two headline metrics)

N > 1: the MSM behind commit and open is sharded by point (column) ranges, one
process per GPU; every rank runs the (cheap) field pipeline, does the MSM over its
SRS slice and the 96-byte partial sums are all-gathered over NCCL and added.
--impl reference: the CPU implementation of the same step (C oracle port - the
reference's own crates cannot be built here: no Rust toolchain), all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCALE = int(os.environ.get("FB_BENCH_SCALE", "20"))
SECRET_X, SECRET_Y = bytes([1] * 32), bytes([2] * 32)
SEED_COEFFS, SEED_ALPHA = 31, 33
METRIC = "kzg_commit_open_ms_deg2^%d" % SCALE
WINDOW = int(os.environ.get("FB_BENCH_WINDOW", "0"))  # 0: the library's own choice
# algorithmic work per unit
MSM_BYTES_PER_POINT = 128          # 32-byte scalar + 96-byte affine point
PHASE_KERNELS = {"msm.round_pass2": "k_round_pass2 (batched-affine rounds: peel inverses, add)",
                 "msm.round_pass1": "k_round_pass1 (batched-affine rounds: gather, running products, warp inversion)",
                 "msm.accumulate": "k_accumulate (XYZZ += affine over equal chunks of the sorted entries)"}
MSM_MAD_PER_POINT = {12: 66000, 20: 48000, 24: 39000}
NTT_BYTES_PER_ELEM = 64


from fourier_b200.benchutil import ClockSampler, load_int_peak, load_peaks, load_traffic  # noqa: E402


def reference_arm(args):
    """The step on the host CPU (C oracle port of the reference's path), all threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from fourier_b200 import synth
    from oracle import cref
    cref.build()
    n = 1 << SCALE
    cores = cref.threads()
    tau = int.from_bytes(SECRET_X, "big") % synth.R
    t0 = time.time()
    lag = cref.lagrange_at(tau.to_bytes(32, "big"), SCALE)  # U[0][j] = g^(L_j(tau_X)) (m = 0: R_0 = 1)
    pts96 = cref.g1_mul_gen(lag)
    setup_s = time.time() - t0
    coeffs = synth.fr_bytes(SEED_COEFFS, n)
    alpha = synth.fr_bytes(SEED_ALPHA, 1).tobytes()
    if args.warmup:
        cref.bench_commit_open(pts96, coeffs, alpha, min(args.warmup, 1))
    tc, to, c, y, p = cref.bench_commit_open(pts96, coeffs, alpha, args.steps)
    ms = (tc + to) * 1e3
    line = {
        "impl": "reference", "metric": METRIC, "value": ms, "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64 limbs (Fr 255-bit, Fp 381-bit Montgomery)", "data": "synthetic",
        "config": {"workload": "commit+open, degree 2^%d, scale=%d machines_scale=0" % (SCALE, SCALE),
                   "commit_ms": tc * 1e3, "open_ms": to * 1e3, "srs_setup_s": setup_s},
        "cpu_baseline": {"value": ms, "unit": "ms", "cores": cores, "kind": "port",
                         "sample": "full workload (2^%d), %d timed step(s); C restatement, not blst "
                                   "(reference crates unbuildable here: no Rust toolchain)" % (SCALE, args.steps)},
        "e2e": {"value": ms, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "result": {"commit": c.hex(), "y": y.hex(), "proof": p.hex()},
    }
    print(json.dumps(line), flush=True)


def own_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from fourier_b200 import PianoBackend, shard, synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = 1 << SCALE
    T = n

    # ---- setup (untimed): SRS resident in HBM with its window tables
    be = PianoBackend(SCALE, 0, device=local, window_bits=WINDOW)
    # a non-default torch stream: the library launches on it and torch.cuda.Event times it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    be.set_stream(stream.cuda_stream)
    col0, ncols = shard.split_range(T, world, rank)  # this rank's point (column) range of the MSM
    if world > 1:
        be.set_col_shard(col0, ncols)
    t0 = time.time()
    be.setup(SECRET_X, SECRET_Y)
    setup_s = time.time() - t0

    # ---- synthetic inputs: identical on every rank (seeded)
    coeffs_np = synth.fr_bytes(SEED_COEFFS, n)
    alpha_b = synth.fr_bytes(SEED_ALPHA, 1).tobytes()
    h_coeffs = torch.from_numpy(coeffs_np.reshape(-1)).pin_memory()
    d_be = h_coeffs.to(dev)
    d_coeffs = torch.empty(n * 32, dtype=torch.uint8, device=dev)
    be.dev_fr_from_be(d_be, d_coeffs, n)
    d_alpha_be = torch.frombuffer(bytearray(alpha_b), dtype=torch.uint8).to(dev)
    d_alpha = torch.empty(32, dtype=torch.uint8, device=dev)
    be.dev_fr_from_be(d_alpha_be, d_alpha, 1)
    d_commit = torch.zeros(96, dtype=torch.uint8, device=dev)
    d_open = torch.zeros(128, dtype=torch.uint8, device=dev)
    gathered = torch.zeros(world * 224, dtype=torch.uint8, device=dev) if world > 1 else None

    def step_device():
        """commit + open with the coefficients resident in HBM"""
        if world == 1:
            be.dev_worker_commit(0, d_coeffs, n, d_commit)
            be.dev_worker_open(0, d_coeffs, n, d_alpha, d_open)
        else:
            be.dev_worker_commit_partial(0, d_coeffs, n, d_commit)
            be.dev_worker_open_partial(0, d_coeffs, n, d_alpha, d_open)
            dist.all_gather_into_tensor(gathered, torch.cat([d_commit, d_open]))

    def combine(g: bytes):
        parts_c = [g[k * 224:k * 224 + 96] for k in range(world)]
        parts_p = [g[k * 224 + 128:k * 224 + 224] for k in range(world)]
        return be.g1_sum(b"".join(parts_c), fmt=96), g[96:128], be.g1_sum(b"".join(parts_p), fmt=96)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    launches0 = be.launch_count()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    barrier()
    ms_dev = e0.elapsed_time(e1) / args.steps
    launches = be.launch_count() - launches0
    clk = clocks.stop()
    if world == 1:
        res_commit, res_y, res_proof = bytes(d_commit[:48].cpu().numpy()), bytes(d_open[:32].cpu().numpy()), \
            bytes(d_open[32:80].cpu().numpy())
    else:
        res_commit, res_y, res_proof = combine(bytes(gathered.cpu().numpy()))

    # ---- e2e: public host API, pinned host inputs, H2D + D2H inside the timed region
    def step_host():
        if world == 1:
            c = be.worker_commit(0, h_coeffs)
            y, p = be.worker_open(0, h_coeffs, alpha_b)
            return c, y, p
        c = be.worker_commit_partial(0, h_coeffs)
        y, p = be.worker_open_partial(0, h_coeffs, alpha_b)
        mine = torch.frombuffer(bytearray(c + y + p), dtype=torch.uint8).to(dev)
        dist.all_gather_into_tensor(gathered, mine)
        return combine(bytes(gathered.cpu().numpy()))

    for _ in range(2):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hc, hy, hp = step_host()
    barrier()
    ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e = float(t[0]), float(t[1])
    assert (hc, hy, hp) == (res_commit, res_y, res_proof), "host-API and device-resident results differ"

    # ---- per-kernel timing (CUDA events inside the library, on the launching stream)
    be.profile(True)
    be.profile_read()
    prof_steps = max(1, min(args.steps, 5))
    for _ in range(prof_steps):
        step_device()
    torch.cuda.synchronize()
    prof = be.profile_read()
    be.profile(False)

    line = None
    if rank == 0:
        peaks = load_peaks()
        int_peak, int_src = load_int_peak()
        nw = lambda c: (255 + c - 1) // c
        wbits = WINDOW if WINDOW else min(range(4, 23), key=lambda c: ncols * nw(c) * 10 + (1 << (c - 1)) * 60)  # msm_pick_window
        table_windows = (255 + wbits - 1) // wbits
        n_msm = 2 * prof_steps  # MSMs profiled (commit + open per step)
        # dominant MSM phase (all its launches inside one MSM count as "the kernel")
        dom = max((k for k in prof if k.startswith("msm.")), key=lambda k: prof[k][0])
        dom_s = prof[dom][0] / n_msm * 1e-3
        pts_per_launch = ncols
        msm_total_ms = sum(v[0] for k, v in prof.items() if k.startswith("msm.")) / n_msm
        ntt_ms = sum(v[0] for k, v in prof.items() if k.startswith("ntt.")) / prof_steps / 2  # 2 NTTs per step
        step_ms_prof = sum(v[0] for v in prof.values()) / prof_steps
        ach_gbs = pts_per_launch * MSM_BYTES_PER_POINT / dom_s / 1e9 if dom_s else 0.0
        traffic = load_traffic().get(dom)
        mad_pp = MSM_MAD_PER_POINT.get(SCALE, 48000)
        ach_mad = pts_per_launch * mad_pp / (msm_total_ms * 1e-3) if msm_total_ms else 0.0
        field_ms = sum(v[0] for k, v in prof.items() if k.startswith(("ntt.", "open.", "poly."))) / prof_steps
        field_gbs = n * NTT_BYTES_PER_ELEM / (field_ms * 1e-3) / 1e9 if field_ms else 0.0
        line = {
            "metric": METRIC, "value": ms_dev, "unit": "ms", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "u32 limbs (Fr 255-bit, Fp 381-bit Montgomery, exact integer)", "data": "synthetic",
            "config": {"workload": "KZG commit+open of one random degree-2^%d polynomial (scale=%d, machines_scale=0), "
                                   "bit-exact vs CPU oracle" % (SCALE, SCALE),
                       "parallelism": "1 GPU" if world == 1 else "MSM point ranges over %d GPUs + all-gather of 224-byte partials" % world,
                       "l2": "inputs larger than L2: window tables %.2f GB per GPU + 32 MiB coefficients" %
                             (table_windows * ncols * 128 / 1e9), "window_bits": wbits, "srs_setup_s": round(setup_s, 2), "seed": [SEED_COEFFS, SEED_ALPHA]},
            "e2e": {"value": ms_e2e, "unit": "ms", "h2d_bytes_per_step": (2 * n if world == 1 else n + ncols) * 32 + 32,
                    "d2h_bytes_per_step": 48 + 32 + 48 if world == 1 else 96 + 32 + 96 + world * 224},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"kernel": PHASE_KERNELS.get(dom, dom), "bound": "hbm",
                         "achieved": ach_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach_gbs / peaks["hbm_gbs"],
                         "traffic": traffic, "peak_source": peaks["source"],
                         "note": "dominant MSM phase, its launches inside one MSM taken together; the binding limit is the "
                                 "integer pipe, see roofline_int; algorithmic bytes = 128 B/point (32-B scalar + 96-B affine "
                                 "point) x points of the MSM; traffic = dram bytes of those launches from the ncu capture in "
                                 "profiles/ (each point is read once per window: 15 table entries per scalar)"},
            "roofline_int": {"kernel": "whole MSM (digits+sort+accumulate+reduce)", "bound": "int (IMAD.WIDE)",
                             "achieved": ach_mad / 1e12, "peak": int_peak / 1e12, "unit": "T wide-MAD/s",
                             "frac": ach_mad / int_peak, "peak_source": int_src,
                             "work": "%d wide-MAD/point at the fixed accounting window (SURVEY 8d)" % mad_pp,
                             "note": "work-equivalent rate, not pipe utilisation: the accounting credits 10 Fp multiplies per point and window (XYZZ mixed addition), the batched-affine rounds spend 6, so frac can reach or pass 1; ncu puts the round kernels at 74-82 % of the multiplier pipe (profiles/r1_v7_ncu_k_round_pass.txt)"},
            "roofline_field": {"kernel": "field half of worker_open: " + ("evalform_open (y = p(alpha) and the quotient's evaluations, "
                               "no transform)" if field_ms and not ntt_ms else "iNTT + Horner/quotient scan + NTT"),
                               "bound": "hbm", "achieved": field_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                               "frac": field_gbs / peaks["hbm_gbs"], "ms": field_ms,
                               "note": "algorithmic bytes = 64 B/coefficient (read the row once, write the quotient once)"},
            "phases_ms_per_step": {k: round(v[0] / prof_steps, 4) for k, v in sorted(prof.items())},
            "extra": {"msm_mpts_per_s": ncols * world / (msm_total_ms * 1e-3) / 1e6 if msm_total_ms else None,
                      "msm_ms": msm_total_ms, "step_ms_profiled": step_ms_prof},
            "result": {"commit": res_commit.hex(), "y": res_y.hex(), "proof": res_proof.hex()},
        }
        # ---- CPU baseline beside it: the C oracle on the same inputs and SRS (rank 0, N = 1 only)
        if world == 1 and not args.no_cpu:
            from oracle import cref
            cref.build()
            blob = be.setup_to_bytes(False)
            off = 96 + 8 + T * 96 + 8 + 96 + 16
            pts96 = blob[off:off + T * 96]
            tc, to, cc, cy, cp = cref.bench_commit_open(pts96, coeffs_np, alpha_b, 1)
            line["cpu_baseline"] = {"value": (tc + to) * 1e3, "unit": "ms", "cores": cref.threads(), "kind": "port",
                                    "sample": "full workload (2^%d), 1 step; C restatement (oracle/c/oracle.c), not blst" % SCALE,
                                    "commit_ms": tc * 1e3, "open_ms": to * 1e3}
            line["parity"] = {"bit_exact_vs_cpu_oracle": (cc, cy, cp) == (res_commit, res_y, res_proof)}
        print(json.dumps(line), flush=True)
    be.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--workload", default="kzg", choices=["kzg", "ntt", "msm", "bipoly"],
                    help="kzg = the headline commit+open at 2^20 (default); the others are BASELINE configs 3-5")
    ap.add_argument("--check", action="store_true", help="(msm, bipoly) also check the result against the oracle's known-dlog identity")
    args = ap.parse_args()
    if args.workload != "kzg":
        if args.impl == "reference":
            raise SystemExit("--impl reference is defined for the headline workload only")
        import bench_workloads
        bench_workloads.RUN[args.workload](args)
        return
    if args.impl == "reference":
        reference_arm(args)
    else:
        own_arm(args)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py - headline benchmark of the KZG hot path (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (config.workload): BASELINE config 2 - KZG commit + open of one random
degree-2^20 polynomial (scale 20, machines_scale 0) on synthetic seeded inputs.
A "step" = the commitment AND the opening (y, proof) of that polynomial.

  value        ms per step with the coefficients already resident in HBM, timed with
               CUDA events on the caller's stream (max over ranks).  The step is measured
               three ways and the best is the headline (config.step says which):
                 separate  fb_dev_worker_commit then fb_dev_worker_open, one stream
                 overlap   the same two calls on two lanes (fb_ctx_set_lanes(2) + join)
                 fused     fb_dev_worker_commit_open: one call; rows of >= 2^18 points run the two MSMs
                           on two streams of the lane, smaller rows as one batch of two scalar
                           vectors over the same resident row
               all three give identical bytes (checked) - extra.step_variants_ms
  e2e          the same through the public host API on HOST buffers (H2D + D2H inside
               the timed region): pinned and pageable sources, separate and fused calls
  roofline     the binding bound is the integer (IMAD.WIDE) pipe: whole-MSM wide-MAD/s at
               SURVEY 8d's fixed accounting (frac), and frac_real = Fp multiplications
               actually issued / time / the measured fe_mul rate, for the MSM and for the
               dominant kernel.  roofline_hbm keeps the HBM view with the ncu traffic.
  cpu_baseline the C oracle (oracle/c/oracle.c) on the host cores, same inputs + SRS
  extra        the other two headline metrics, measured in the same run: one G1 MSM of
               2^24 points (Mpts/s, sharded by point ranges over the ranks) and the Fr NTT
               sweep 2^16..2^24 (Gelem/s); setup-file loader throughput (C2 and C5 files)

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
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCALE = int(os.environ.get("FB_BENCH_SCALE", "20"))
SECRET_X, SECRET_Y = bytes([1] * 32), bytes([2] * 32)
SEED_COEFFS, SEED_ALPHA = 31, 33
METRIC = "kzg_commit_open_ms_deg2^%d" % SCALE
WORKLOAD = ("KZG commit+open of one random degree-2^%d polynomial (scale=%d, machines_scale=0), "
            "synthetic seeded inputs, known-tau SRS" % (SCALE, SCALE))
WINDOW = int(os.environ.get("FB_BENCH_WINDOW", "0"))  # 0: the library's own choice
# algorithmic work per unit
MSM_BYTES_PER_POINT = 128          # 32-byte scalar + 96-byte affine point
PHASE_KERNELS = {"msm.round_pass2": "k_round_pass2 (batched-affine rounds: peel inverses, add)",
                 "msm.round_pass1": "k_round_pass1 (batched-affine rounds: gather, running products, CTA inversion)",
                 "msm.accumulate": "k_accumulate (XYZZ += affine over equal chunks of the sorted entries)"}
MSM_MAD_PER_POINT = {12: 66000, 20: 48000, 24: 39000}
NTT_BYTES_PER_ELEM = 64
# Fp multiplications per unit of work as the kernels issue them (msm.cu): an affine-round slot costs 1 (running product,
# pass 1) + 5 (pass 2: two to peel the inverse, slope, x3 = lambda^2.., y3); an XYZZ mixed addition 10 (8M + 2S); a
# bucket goes through two full XYZZ additions (14 each) in the row/column + bit-sliced reduction
MULS_PER_SLOT, MULS_PASS2_PER_SLOT, MULS_PER_MIXED_ADD, MULS_PER_BUCKET = 6, 5, 10, 28


from fourier_b200.benchutil import ClockSampler, host_cores, load_int_peak, load_peaks, load_traffic  # noqa: E402


def cpu_step(pts96, coeffs, alpha, reps):
    """One commit + open on the host CPU through the C oracle with every core the process may use (never the
    inherited OMP_NUM_THREADS: torchrun exports 1)."""
    from oracle import cref
    cref.build()
    cores = host_cores()
    cref.set_threads(cores)
    tc, to, c, y, p = cref.bench_commit_open(pts96, coeffs, alpha, reps)
    return tc, to, c, y, p, cref.threads()


def reference_arm(args):
    """The step on the host CPU (C oracle port of the reference's path), all threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from fourier_b200 import synth
    from oracle import cref
    cref.build()
    cref.set_threads(host_cores())
    n = 1 << SCALE
    tau = int.from_bytes(SECRET_X, "big") % synth.R
    t0 = time.time()
    lag = cref.lagrange_at(tau.to_bytes(32, "big"), SCALE)  # U[0][j] = g^(L_j(tau_X)) (m = 0: R_0 = 1)
    pts96 = cref.g1_mul_gen(lag)
    setup_s = time.time() - t0
    coeffs = synth.fr_bytes(SEED_COEFFS, n)
    alpha = synth.fr_bytes(SEED_ALPHA, 1).tobytes()
    if args.warmup:
        cpu_step(pts96, coeffs, alpha, min(args.warmup, 1))
    tc, to, c, y, p, cores = cpu_step(pts96, coeffs, alpha, args.steps)
    ms = (tc + to) * 1e3
    line = {
        "impl": "reference", "metric": METRIC, "value": ms, "unit": "ms", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "u64 limbs (Fr 255-bit, Fp 381-bit Montgomery, exact integer)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "commit_ms": tc * 1e3, "open_ms": to * 1e3, "srs_setup_s": setup_s},
        "cpu_baseline": {"value": ms, "unit": "ms", "cores": cores, "kind": "port",
                         "sample": "full workload (2^%d), %d timed step(s); C restatement, not blst: upper bound on the "
                                   "speed-up, the reference engine is unavailable on this box (no Rust toolchain)" % (SCALE, args.steps)},
        "e2e": {"value": ms, "unit": "ms", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "result": {"commit": c.hex(), "y": y.hex(), "proof": p.hex()},
    }
    print(json.dumps(line), flush=True)


def own_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import bench_workloads as bw
    from fourier_b200 import PianoBackend, shard, synth

    env = bw.Env()
    rank, world, local, dev, stream = env.rank, env.world, env.local, env.dev, env.stream
    n = 1 << SCALE
    T = n

    # ---- setup (untimed): SRS resident in HBM with its window tables
    be = PianoBackend(SCALE, 0, device=local, window_bits=WINDOW)
    be.set_stream(stream.cuda_stream)  # the caller's stream: torch.cuda.Event times it
    col0, ncols = shard.split_range(T, world, rank)  # this rank's point (column) range of the MSM
    if world > 1:
        be.set_col_shard(col0, ncols)
    t0 = time.time()
    be.setup(SECRET_X, SECRET_Y)
    setup_s = time.time() - t0
    wbits, table_windows = be.window()
    msm_split = int(os.environ.get("FB_BENCH_SPLIT", "1"))
    be.set_msm_split(msm_split)

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
    # one result buffer, the layout of the fused call: point | y | point (points 48 B, or 96 B partials when sharded)
    P = 48 if world == 1 else 96
    d_res = torch.zeros(2 * P + 32, dtype=torch.uint8, device=dev)
    d_open = torch.zeros(32 + P, dtype=torch.uint8, device=dev)
    gathered = torch.zeros(world * (2 * P + 32), dtype=torch.uint8, device=dev) if world > 1 else None
    # N > 1, field half sharded too: this rank's slice of the coefficients, the 128-byte states of all ranks, the final record
    d_slice = d_coeffs[col0 * 32:(col0 + ncols) * 32]
    d_state = torch.zeros(128, dtype=torch.uint8, device=dev)
    d_states = torch.zeros(world * 128, dtype=torch.uint8, device=dev)
    d_final = torch.zeros(128, dtype=torch.uint8, device=dev)

    def step_separate():
        if world == 1:
            be.dev_worker_commit(0, d_coeffs, n, d_res[:P])
            be.dev_worker_open(0, d_coeffs, n, d_alpha, d_open)
        else:
            be.dev_worker_commit_partial(0, d_coeffs, n, d_res[:P])
            be.dev_worker_open_partial(0, d_coeffs, n, d_alpha, d_open)

    def step_fused():
        if world == 1:
            be.dev_worker_commit_open(0, d_coeffs, n, d_alpha, d_res)
        else:
            be.dev_worker_commit_open_partial(0, d_coeffs, n, d_alpha, d_res)

    def step_sharded():
        """N > 1 only: every rank runs the field half over ITS slice; one all-gather of 128 bytes in the middle"""
        be.dev_commit_open_shard_begin(0, d_slice, n, d_alpha, d_state)
        dist.all_gather_into_tensor(d_states, d_state)
        be.dev_commit_open_shard_finish(0, d_slice, n, d_alpha, d_states, world, d_res)

    def finish(separate):
        """outputs ordered on the caller's stream; N > 1: the all-gather of the partials and their sum, on the device"""
        be.join()
        if separate:
            d_res[P:].copy_(d_open)
        if world > 1:
            dist.all_gather_into_tensor(gathered, d_res)
            be.dev_combine_commit_open(gathered, world, d_final)

    def result():
        r = bytes((d_res if world == 1 else d_final).cpu().numpy())
        return r[:48], r[48:80], r[80:128]

    variants = {"separate": (1, step_separate, True), "overlap": (2, step_separate, True), "fused": (1, step_fused, False)}
    if world > 1:
        variants["sharded"] = (1, step_sharded, False)
    only = os.environ.get("FB_BENCH_STEP")  # measure one variant only (A/B runs, ncu)
    if only:
        variants = {only: variants[only]}
    warm = max(args.warmup, 3)
    timed, results, launches_by = {}, {}, {}
    clocks = ClockSampler(local)
    for name, (lanes, fn, sep) in variants.items():
        be.set_lanes(lanes)
        for _ in range(warm):
            fn()
            finish(sep)
        env.barrier()
        l0 = be.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            fn()
            finish(sep)
        e1.record(stream)
        env.barrier()
        timed[name] = env.max_over_ranks([e0.elapsed_time(e1) / args.steps])[0]
        launches_by[name] = be.launch_count() - l0
        results[name] = result()
    clk = clocks.stop()
    be.set_lanes(1)
    assert len(set(results.values())) == 1, "the step variants disagree: %r" % results
    best = min(timed, key=timed.get)
    ms_dev = timed[best]
    res_commit, res_y, res_proof = results[best]
    launches = launches_by[best]

    # ---- e2e: public host API, H2D + D2H inside the timed region
    h_slice = h_coeffs[col0 * 32:(col0 + ncols) * 32]  # (a view of the pinned buffer)

    def host_step(kind, src):
        if world == 1:
            if kind == "fused":
                return be.worker_commit_open(0, src, alpha_b)
            c = be.worker_commit(0, src)
            y, p = be.worker_open(0, src, alpha_b)
            return c, y, p
        if kind == "sharded":  # host slice up once, 128-byte states exchanged, partials gathered
            st = be.commit_open_shard_begin(0, src, n, alpha_b)
            mine = torch.frombuffer(bytearray(st), dtype=torch.uint8).to(dev)
            dist.all_gather_into_tensor(d_states, mine)
            c, y, p = be.commit_open_shard_finish(0, bytes(d_states.cpu().numpy()), alpha_b)
        elif kind == "fused":
            c, y, p = be.worker_commit_open_partial(0, src, alpha_b)
        else:
            c = be.worker_commit_partial(0, src)
            y, p = be.worker_open_partial(0, src, alpha_b)
        mine = torch.frombuffer(bytearray(c + y + p), dtype=torch.uint8).to(dev)
        dist.all_gather_into_tensor(gathered, mine)
        be.dev_combine_commit_open(gathered, world, d_final)
        torch.cuda.current_stream().synchronize()
        r = bytes(d_final.cpu().numpy())
        return r[:48], r[48:80], r[80:128]

    pageable = coeffs_np.tobytes()  # a plain bytes object: what a Rust Vec<u8> / Python caller hands over
    e2e = {}
    host_cases = [("separate_pinned", "separate", h_coeffs), ("fused_pinned", "fused", h_coeffs),
                  ("separate_pageable", "separate", pageable), ("fused_pageable", "fused", pageable),
                  ("fused_pageable_driver_staged", "fused", pageable)]
    if world > 1:
        host_cases += [("sharded_pinned", "sharded", h_slice), ("sharded_pageable", "sharded", pageable[col0 * 32:(col0 + ncols) * 32])]
    for name, fused, src in host_cases:
        if only and not name.startswith(only if only != "overlap" else "separate"):
            continue
        be.set_staging(name.endswith("driver_staged"))  # A/B: the driver's own pageable path instead of the library's ring
        for _ in range(2):
            host_step(fused, src)
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            hres = host_step(fused, src)
        env.barrier()
        e2e[name] = env.max_over_ranks([(time.perf_counter() - t0) * 1e3 / args.steps])[0]
        assert hres == (res_commit, res_y, res_proof), "host-API and device-resident results differ (%s)" % name
    be.set_staging(False)
    e2e_key = min((k for k in e2e if k.endswith("_pinned")), key=e2e.get)  # the host call a user would make
    ms_e2e = e2e[e2e_key]
    fused_best = best == "fused"
    if world == 1:
        h2d = (n if e2e_key.startswith("fused") else 2 * n) * 32 + 32
        d2h = 48 + 32 + 48
    elif e2e_key.startswith("sharded"):
        h2d = ncols * 32 + 64 + 128 + world * 128 + 224
        d2h = 128 + world * 128 + 224 + 128
    else:
        h2d = (n if e2e_key.startswith("fused") else n + ncols) * 32 + 32 + 224
        d2h = 96 + 32 + 96 + 128

    # ---- per-kernel timing (CUDA events inside the library, on the launching stream) + work counters
    be.profile(True)
    prof_steps = max(1, min(args.steps, 5))
    # Which step to profile.  Under profiling every call runs on ONE stream, so the fused call of a row of >= 2^18 points
    # (timed as two MSMs on two streams) would be profiled as one batch-of-two pass, a pipeline that is not the timed one:
    # profile the two plain calls instead; smaller rows run the batch-of-two pass when timed, too.
    fused_prof = fused_best and ncols < (1 << 18)
    prof_fn, prof_sep = (step_sharded, False) if best == "sharded" else (step_fused, False) if fused_prof else (step_separate, True)
    # one untimed step first: under profiling a call runs on ONE stream (no two-stream pair, no split), whose scratch may
    # have to grow for that shape - allocations in the middle of a step would be charged to whatever phase they interrupt
    prof_fn()
    finish(prof_sep)
    torch.cuda.synchronize()
    be.profile_read()
    be.msm_stats()
    for _ in range(prof_steps):
        prof_fn()
        finish(prof_sep)
    torch.cuda.synchronize()
    prof = be.profile_read()
    stats = be.msm_stats()
    be.profile(False)

    line = None
    if rank == 0:
        peaks = load_peaks()
        int_peak, int_src, fp_mul_rate = load_int_peak()
        msm_ms = sum(v[0] for k, v in prof.items() if k.startswith("msm.")) / prof_steps  # both MSMs of a step
        step_ms_prof = sum(v[0] for v in prof.values()) / prof_steps
        field_ms = sum(v[0] for k, v in prof.items() if k.startswith(("ntt.", "open.", "poly."))) / prof_steps
        dom = max((k for k in prof if k.startswith("msm.")), key=lambda k: prof[k][0])
        dom_ms = prof[dom][0] / prof_steps
        mad_pp = MSM_MAD_PER_POINT.get(SCALE, 48000)
        pts_step = 2 * ncols  # points through this rank's MSMs per step (commit + open)
        ach_mad = pts_step * mad_pp / (msm_ms * 1e-3) if msm_ms else 0.0
        # Fp multiplications actually issued per step (from the library's work counters)
        slots, surv, buckets = (stats[k] / prof_steps for k in ("round_slots", "survivors", "buckets"))
        muls_step = slots * MULS_PER_SLOT + surv * MULS_PER_MIXED_ADD + buckets * MULS_PER_BUCKET
        frac_real = muls_step / (msm_ms * 1e-3) / fp_mul_rate if msm_ms else 0.0
        p2_ms = prof.get("msm.round_pass2", (0.0, 0))[0] / prof_steps
        dom_real = slots * MULS_PASS2_PER_SLOT / (p2_ms * 1e-3) / fp_mul_rate if p2_ms else None
        traffic = load_traffic()
        ach_gbs = pts_step * MSM_BYTES_PER_POINT / (dom_ms * 1e-3) / 1e9 if dom_ms else 0.0
        field_gbs = n * NTT_BYTES_PER_ELEM / (field_ms * 1e-3) / 1e9 if field_ms else 0.0
        line = {
            "metric": METRIC, "value": ms_dev, "unit": "ms", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_dev, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
            "dtype": "u32 limbs (Fr 255-bit, Fp 381-bit Montgomery, exact integer)", "data": "synthetic",
            "config": {"workload": WORKLOAD,
                       "step": {"separate": "fb_dev_worker_commit + fb_dev_worker_open, one stream",
                                "overlap": "fb_dev_worker_commit + fb_dev_worker_open on two lanes (fb_ctx_set_lanes(2)), joined",
                                "fused": "fb_dev_worker_commit_open (one call: the two MSMs on two streams of the lane for rows of >= 2^18 points, "
                                         "one batch of two scalar vectors over the resident row below that)",
                                "sharded": "fb_dev_commit_open_shard_begin + all-gather of 128-byte states + _finish (field half and "
                                           "MSMs over this rank's slice only)"}[best]
                               + ("" if world == 1 else "; then all-gather of 224-byte partials + fb_dev_combine_commit_open on every rank"),
                       "parallelism": "1 GPU" if world == 1 else "MSM point ranges over %d GPUs + all-gather of %d-byte partials + device-side sum" % (world, 2 * P + 32),
                       "l2": "inputs larger than L2: window tables %.2f GB per GPU + 32 MiB coefficients" %
                             (table_windows * ncols * 128 / 1e9), "window_bits": wbits, "msm_split": msm_split,
                       "srs_setup_s": round(setup_s, 2),
                       "seed": [SEED_COEFFS, SEED_ALPHA]},
            "e2e": {"value": ms_e2e, "unit": "ms", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "call": e2e_key, "all_ms": e2e,
                    "note": "host API on host buffers; *_pageable = a plain bytes object staged through the library's "
                            "pinned double buffer, *_pinned = cudaHostAlloc'ed source copied directly"},
            "gpu_launches": int(launches),
            "clocks": clk,
            "roofline": {"kernel": "whole MSM pipeline (digits, sort, batched-affine rounds, XYZZ accumulate, bucket reduction), "
                                   "both MSMs of the step", "bound": "int",
                         "achieved": ach_mad / 1e12, "peak": int_peak / 1e12, "unit": "T wide-MAD/s", "frac": ach_mad / int_peak,
                         "frac_real": frac_real, "frac_real_step": muls_step / (ms_dev * 1e-3) / fp_mul_rate, "peak_source": int_src,
                         "traffic": sum(traffic.values()) * 2 if traffic else None,
                         "work": "%d wide-MAD/point at the fixed accounting window (SURVEY 8d) x %d points" % (mad_pp, pts_step),
                         "real_work": {"fp_muls_per_step": muls_step, "affine_round_slots": slots, "xyzz_survivors": surv,
                                       "buckets": buckets, "fp_mul_per_s_peak": fp_mul_rate,
                                       "formula": "6 x slots + 10 x survivors + 28 x buckets"},
                         "dominant_kernel": {"name": PHASE_KERNELS.get(dom, dom), "ms_per_step": dom_ms,
                                             "share_of_step": dom_ms / step_ms_prof if step_ms_prof else None,
                                             "frac_real": dom_real if dom == "msm.round_pass2" else None},
                         "note": "frac_real is measured on the profiled single-stream run (phases_ms_per_step); frac_real_step divides the same "
                                 "multiply count by the timed step (value), in which the two MSMs overlap.  "
                                 "frac credits 10 Fp multiplies per (point, window) whatever the kernels spend (they spend 6 in the "
                                 "affine rounds), so it is a work-equivalent rate; frac_real is multiplier utilisation: Fp multiplies "
                                 "issued / time / the fe_mul microbenchmark rate.  traffic = dram bytes of the round kernels of both "
                                 "MSMs from the ncu capture in profiles/ (traffic.json: r2_v7_ncu_round_kernels.txt)"},
            "roofline_hbm": {"kernel": PHASE_KERNELS.get(dom, dom), "bound": "hbm", "achieved": ach_gbs, "peak": peaks["hbm_gbs"],
                             "unit": "GB/s", "frac": ach_gbs / peaks["hbm_gbs"], "traffic": traffic.get(dom),
                             "peak_source": peaks["source"],
                             "note": "algorithmic bytes = 128 B/point (32-B scalar + 96-B affine point); not the binding bound"},
            "roofline_field": {"kernel": "field half of worker_open: evalform_open (y = p(alpha) and the quotient's evaluations, no transform)",
                               "bound": "hbm", "achieved": field_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                               "frac": field_gbs / peaks["hbm_gbs"], "ms": field_ms,
                               "note": "algorithmic bytes = 64 B/coefficient (read the row once, write the quotient once)"},
            "phases_ms_per_step": {k: round(v[0] / prof_steps, 4) for k, v in sorted(prof.items())},
            "extra": {"step_variants_ms": timed, "step_ms_profiled": step_ms_prof, "msm_ms_per_step": msm_ms,
                      "msm_2^%d_mpts_per_s_inside_step" % SCALE: pts_step * world / (msm_ms * 1e-3) / 1e6 if msm_ms else None},
            "result": {"commit": res_commit.hex(), "y": res_y.hex(), "proof": res_proof.hex()},
        }
        # ---- CPU baseline beside it: the C oracle on the same inputs and SRS (rank 0)
        if not args.no_cpu:
            from oracle import cref
            cref.build()
            cref.set_threads(host_cores())
            if world == 1:
                blob = be.setup_to_bytes(False)
                off = 96 + 8 + T * 96 + 8 + 96 + 16
                pts96 = blob[off:off + T * 96]
                del blob
            else:  # the rank holds a slice only: rebuild the row on the host (closed-form Lagrange values x G)
                tau = int.from_bytes(SECRET_X, "big") % synth.R
                pts96 = cref.g1_mul_gen(cref.lagrange_at(tau.to_bytes(32, "big"), SCALE))
            tc, to, cc, cy, cp, cores = cpu_step(pts96, coeffs_np, alpha_b, 1)
            line["cpu_baseline"] = {"value": (tc + to) * 1e3, "unit": "ms", "cores": cores, "kind": "port",
                                    "sample": "full workload (2^%d), 1 step; C restatement (oracle/c/oracle.c), not blst: the ratio "
                                              "is an upper bound on the speed-up over the real reference engine" % SCALE,
                                    "commit_ms": tc * 1e3, "open_ms": to * 1e3}
            line["parity"] = {"bit_exact_vs_cpu_oracle": (cc, cy, cp) == (res_commit, res_y, res_proof)}
            del pts96
    # ---- setup-file loader (north_star: "the setup-point loader that keeps the SRS resident in HBM")
    if not args.no_extra and world == 1:
        line["extra"]["srs_load"] = bw.srs_load_lines(be, SCALE, 0)
    be.close()
    del be, d_coeffs, d_be
    torch.cuda.empty_cache()
    # ---- the other two headline metrics in the same run
    if not args.no_extra:
        msm_line = bw.run_msm(args, env, emit=False)
        ntt_line = bw.run_ntt(args, env, emit=False)
        if world == 1:
            c5 = bw.srs_load_c5(env)
        if rank == 0:
            line["extra"]["msm_2^24"] = {"mpts_per_s": msm_line["value"], "ms": msm_line["ms_per_step"],
                                         "e2e_mpts_per_s": msm_line["e2e"]["value"], "int_frac": msm_line["roofline"]["frac"],
                                         "frac_real": msm_line["roofline"]["frac_real"], "parity": msm_line.get("parity"),
                                         "window_bits": msm_line["config"]["window_bits"], "phases_ms": msm_line["phases_ms"]}
            line["extra"]["ntt_sweep"] = {k: {"gelem_per_s": v["gelem_per_s"], "fwd_ms": v["fwd_ms"], "inv_ms": v["inv_ms"],
                                              "hbm_frac": v["hbm_frac"], "int_frac": v["int_frac"], "frac_real": v["frac_real"],
                                              "roundtrip_identity": v["roundtrip_identity"]}
                                          for k, v in ntt_line["sweep"].items()}
            line["extra"]["ntt_2^24_e2e_gelem_per_s"] = ntt_line["e2e"]["value"]
            if world == 1:
                line["extra"]["srs_load"].update(c5)
    if rank == 0:
        print(json.dumps(line), flush=True)
    env.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip extra.msm_2^24 / extra.ntt_sweep / extra.srs_load")
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

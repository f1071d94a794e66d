"""The other BASELINE.json workloads, behind `bench.py --workload {ntt,msm,bipoly}`:

  ntt     config 3: Fr NTT / iNTT sweep 2^16 .. 2^24, single transform resident in HBM
          (N GPUs = N independent replicas: a single NTT has no exchange step on this path)
  msm     config 4: one G1 MSM of 2^24 points / random scalars, sharded by point ranges
  bipoly  config 5: commit + open of 2^10 rows x 2^12 coefficients, rows sharded over the GPUs,
          master_commit + master_open on rank 0

Same JSON-line contract as bench.py (metric / value / e2e / roofline / clocks ...).
"""
from __future__ import annotations

import json
import os
import time

from fourier_b200.benchutil import ClockSampler, load_fr_mul_rate, load_int_peak, load_peaks

R = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
SECRET_X, SECRET_Y = bytes([1] * 32), bytes([2] * 32)
DTYPE = "u32 limbs (Fr 255-bit, Fp 381-bit Montgomery, exact integer)"


class Env:
    """One process per GPU: device, the caller's (non-default) torch stream, the NCCL group when N > 1."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench needs a CUDA device: the product has no CPU path (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.stream = torch.cuda.Stream(device=self.dev)
        torch.cuda.set_stream(self.stream)
        assert self.stream.cuda_stream != 0

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, vals):
        if self.world == 1:
            return list(vals)
        t = self.torch.tensor(list(vals), device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def _env():
    e = Env()
    return e.torch, e.dist, e.rank, e.world, e.local, e.dev, e.stream


def _barrier(torch, dist, world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(torch, dist, world, dev, vals):
    if world == 1:
        return vals
    t = torch.tensor(vals, device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t]


def _dev_fr(torch, be, dev, be_bytes_np):
    """host (n,32) big-endian uint8 -> device Montgomery Fr tensor (n*32 bytes)"""
    n = be_bytes_np.shape[0]
    d_be = torch.from_numpy(be_bytes_np.reshape(-1)).to(dev)
    d = torch.empty(n * 32, dtype=torch.uint8, device=dev)
    be.dev_fr_from_be(d_be, d, n)
    return d


# ------------------------------------------------------------------ config 3: NTT sweep
def ntt_plan(lg: int, max_r: int = 10):
    """log2 of the radix of each pass (ntt_make_plan, csrc/ntt.cu): as even as possible over ceil(lg / 10) passes"""
    npass = max(1, -(-lg // max_r))
    out, left = [], lg
    for j in range(npass):
        r = -(-left // (npass - j))
        out.append(r)
        left -= r
    return out


def ntt_fr_muls_per_element(lg: int) -> float:
    """Fr multiplications the kernel issues per element: half a butterfly multiply per stage, minus the multiplier-free
    butterflies of the last two stages of every pass (0.75), plus one inter-pass twiddle multiply per non-last pass"""
    plan = ntt_plan(lg)
    return sum(max(r * 0.5 - 0.75, 0.0) for r in plan) + (len(plan) - 1)


def run_ntt(args, env=None, emit=True):
    own_env = env is None
    env = env or Env()
    torch, dist, rank, world, local, dev, stream = env.torch, env.dist, env.rank, env.world, env.local, env.dev, env.stream
    from fourier_b200 import PianoBackend, synth
    be = PianoBackend(24, 0, device=local)
    be.set_stream(stream.cuda_stream)
    peaks = load_peaks()
    int_peak, int_src, _ = load_int_peak()
    fr_rate = load_fr_mul_rate()
    sweep = {}
    clocks = ClockSampler(local)
    total_launches0 = be.launch_count()
    ms_headline = None
    for lg in (16, 18, 20, 22, 24):
        n = 1 << lg
        # enough distinct buffers that consecutive transforms never find their input in L2 (126 MB)
        nbuf = max(1, min(64, (256 << 20) // (n * 32)))
        host = synth.fr_bytes(40 + lg, n)
        bufs = [_dev_fr(torch, be, dev, host) for _ in range(nbuf)]
        out = torch.empty(n * 32, dtype=torch.uint8, device=dev)
        res = {}
        for inverse in (False, True):
            for k in range(max(args.warmup, 3)):
                be.dev_ntt(bufs[k % nbuf], out, n, inverse)
            _barrier(torch, dist, world)
            reps = max(args.steps, 10)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for k in range(reps):
                be.dev_ntt(bufs[k % nbuf], out, n, inverse)
            e1.record(stream)
            _barrier(torch, dist, world)
            ms = _max_over_ranks(torch, dist, world, dev, [e0.elapsed_time(e1) / reps])[0]
            res["inverse" if inverse else "forward"] = ms
        # round trip is the identity (size-independent parity property)
        tmp = torch.empty_like(out)
        be.dev_ntt(bufs[0], out, n, False)
        be.dev_ntt(out, tmp, n, True)
        torch.cuda.synchronize()
        ok = bool(torch.equal(tmp, bufs[0]))
        # SURVEY 8(d)'s fixed accounting: log2(N)/2 Fr multiplies of 136 wide MADs per element (butterflies only; the
        # inter-pass twiddles are overhead, not credited)
        mad_per_elem = lg / 2 * 136
        sweep["2^%d" % lg] = {
            "fwd_ms": res["forward"], "inv_ms": res["inverse"],
            "gelem_per_s": world * n / (res["forward"] * 1e-3) / 1e9,
            "hbm_gbs": n * 64 / (res["forward"] * 1e-3) / 1e9, "hbm_frac": n * 64 / (res["forward"] * 1e-3) / 1e9 / peaks["hbm_gbs"],
            "int_frac": n * mad_per_elem / (res["forward"] * 1e-3) / int_peak,
            "frac_real": n * ntt_fr_muls_per_element(lg) / (res["forward"] * 1e-3) / fr_rate,
            "fr_muls_per_element": ntt_fr_muls_per_element(lg),
            "roundtrip_identity": ok, "buffers": nbuf}
        if lg == 24:
            ms_headline = res["forward"]
            # e2e: the public host call (fb_ntt) on pinned host buffers: H2D of the 32-byte big-endian scalars, decode,
            # transform, encode, D2H of the result, all inside the timed region
            h_in = torch.from_numpy(host.reshape(-1).copy()).pin_memory()
            h_out = torch.empty(n * 32, dtype=torch.uint8).pin_memory()
            be.ntt(h_in, False, out=h_out)
            _barrier(torch, dist, world)
            t0 = time.perf_counter()
            for _ in range(args.steps):
                be.ntt(h_in, False, out=h_out)
            _barrier(torch, dist, world)
            ms_e2e = _max_over_ranks(torch, dist, world, dev, [(time.perf_counter() - t0) * 1e3 / args.steps])[0]
            be.dev_fr_to_be(out, tmp, n)
            torch.cuda.synchronize()
            e2e_ok = bool(torch.equal(tmp.cpu(), h_out))
            del h_in, h_out
        del bufs, out, tmp
    clk = clocks.stop()
    line = None
    if rank == 0:
        n = 1 << 24
        line = {
            "metric": "fr_ntt_gelem_per_s_2^24", "value": world * n / (ms_headline * 1e-3) / 1e9, "unit": "Gelem/s",
            "n_gpus": world, "steps": max(args.steps, 10), "warmup": max(args.warmup, 3), "ms_per_step": ms_headline,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": {"workload": "Fr NTT sweep 2^16..2^24, natural order in/out (BASELINE config 3)",
                       "parallelism": "1 GPU" if world == 1 else "%d independent replicas (a single NTT does not shard on this path)" % world,
                       "l2": "inputs rotate over >= 256 MB of distinct buffers (larger than L2)"},
            "e2e": {"value": world * n / (ms_e2e * 1e-3) / 1e9, "unit": "Gelem/s", "h2d_bytes_per_step": n * 32,
                    "d2h_bytes_per_step": n * 32, "ms": ms_e2e, "matches_device_path": e2e_ok,
                    "note": "fb_ntt on pinned host buffers; bounded by the two 512 MiB PCIe copies"},
            "gpu_launches": int(be.launch_count() - total_launches0), "clocks": clk,
            "roofline": {"kernel": "k_ntt_pass (3 passes at 2^24)", "bound": "hbm", "achieved": n * 64 / (ms_headline * 1e-3) / 1e9,
                         "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": n * 64 / (ms_headline * 1e-3) / 1e9 / peaks["hbm_gbs"],
                         "traffic": None, "peak_source": peaks["source"],
                         "note": "algorithmic bytes = 64 B/element (one read + one write); the binding limit is the integer "
                                 "pipe (SURVEY 8d: log2(N)/2 Fr multiplies of 136 wide MADs per element): see sweep[*].int_frac "
                                 "against " + int_src + "; sweep[*].frac_real = Fr multiplies actually issued / time / the "
                                 "fe_mul<Fr> microbenchmark rate"},
            "sweep": sweep,
        }
        if emit:
            print(json.dumps(line), flush=True)
    be.close()
    if own_env:
        env.close()
    return line


# ------------------------------------------------------------------ config 4: one big MSM
def run_msm(args, env=None, emit=True):
    own_env = env is None
    env = env or Env()
    torch, dist, rank, world, local, dev, stream = env.torch, env.dist, env.rank, env.world, env.local, env.dev, env.stream
    from fourier_b200 import PianoBackend, shard, synth
    lg = int(os.environ.get("FB_MSM_LOG", "24"))
    n = 1 << lg
    start, per = shard.split_range(n, world, rank)  # this rank's point range
    be = PianoBackend(10, 0, device=local)
    be.set_stream(stream.cuda_stream)
    t0 = time.time()
    bases = be.bases_from_scalars(synth.fr_bytes(21, per, start))  # P_k = s_k * G
    bases.precompute(int(os.environ.get("FB_BENCH_WINDOW", "0")))
    be.set_msm_split(int(os.environ.get("FB_BENCH_SPLIT", "1")))
    torch.cuda.synchronize()
    setup_s = time.time() - t0
    h_scalars = torch.from_numpy(synth.fr_bytes(22, per, start).reshape(-1)).pin_memory()
    d_scalars = torch.empty(per * 32, dtype=torch.uint8, device=dev)
    be.dev_fr_from_be(h_scalars.to(dev), d_scalars, per)
    d_out = torch.zeros(96, dtype=torch.uint8, device=dev)
    gathered = torch.zeros(world * 96, dtype=torch.uint8, device=dev)

    def step_device():
        be.dev_msm_partial(bases, d_scalars, per, d_out)
        if world > 1:
            dist.all_gather_into_tensor(gathered, d_out)

    def result():
        parts = bytes((gathered if world > 1 else d_out).cpu().numpy())
        return be.g1_sum(parts, fmt=96)

    for _ in range(max(args.warmup, 3)):
        step_device()
    _barrier(torch, dist, world)
    l0 = be.launch_count()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(torch, dist, world)
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    _barrier(torch, dist, world)
    ms_dev = e0.elapsed_time(e1) / args.steps
    launches = be.launch_count() - l0
    clk = clocks.stop()
    res = result()

    # e2e: scalars from pinned host memory through the public call, result read back
    def step_host():
        part = be.msm_partial(bases, h_scalars)
        if world == 1:
            return be.g1_sum(part, fmt=96)
        mine = torch.frombuffer(bytearray(part), dtype=torch.uint8).to(dev)
        dist.all_gather_into_tensor(gathered, mine)
        return be.g1_sum(bytes(gathered.cpu().numpy()), fmt=96)

    step_host()
    _barrier(torch, dist, world)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hres = step_host()
    _barrier(torch, dist, world)
    ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps
    ms_dev, ms_e2e = _max_over_ranks(torch, dist, world, dev, [ms_dev, ms_e2e])
    assert hres == res, "host-API and device-resident results differ"

    be.profile(True)
    step_device()  # (scratch of the single-stream shape grows here, untimed)
    torch.cuda.synchronize()
    be.profile_read()
    be.msm_stats()
    step_device()
    torch.cuda.synchronize()
    prof = be.profile_read()
    stats = be.msm_stats()
    be.profile(False)
    wbits, nwin = bases.window()
    line = None
    if rank == 0:
        peaks = load_peaks()
        int_peak, int_src, fp_mul_rate = load_int_peak()
        mad_pp = {24: 39000, 20: 48000}.get(lg, 48000)
        msm_ms = sum(v[0] for k, v in prof.items() if k.startswith("msm."))
        muls = stats["round_slots"] * 6 + stats["survivors"] * 10 + stats["buckets"] * 28
        dom = max((k for k in prof if k.startswith("msm.")), key=lambda k: prof[k][0])
        dom_s = prof[dom][0] / max(prof[dom][1], 1) * 1e-3
        line = {
            "metric": "g1_msm_mpts_per_s_2^%d" % lg, "value": n / (ms_dev * 1e-3) / 1e6, "unit": "Mpts/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": {"workload": "one G1 MSM of 2^%d points P_k = s_k*G with random 255-bit scalars (BASELINE config 4)" % lg,
                       "parallelism": "1 GPU" if world == 1 else "point ranges over %d GPUs + all-gather of 96-byte partial sums" % world,
                       "l2": "inputs larger than L2 (window tables + scalars)", "setup_s": round(setup_s, 2),
                       "window_bits": wbits, "windows": nwin, "msm_split": int(os.environ.get("FB_BENCH_SPLIT", "1"))},
            "e2e": {"value": n / (ms_e2e * 1e-3) / 1e6, "unit": "Mpts/s", "h2d_bytes_per_step": per * 32,
                    "d2h_bytes_per_step": 96 + (world * 96 if world > 1 else 0)},
            "gpu_launches": int(launches), "clocks": clk,
            "roofline": {"kernel": "whole MSM pipeline", "bound": "int", "achieved": n * mad_pp / (ms_dev * 1e-3) / 1e12,
                         "peak": world * int_peak / 1e12, "unit": "T wide-MAD/s", "frac": n * mad_pp / (ms_dev * 1e-3) / (world * int_peak),
                         "frac_real": muls / (msm_ms * 1e-3) / fp_mul_rate if msm_ms else None,
                         "peak_source": int_src, "traffic": None,
                         "work": "%d wide-MAD/point at the fixed accounting window (SURVEY 8d)" % mad_pp,
                         "real_work": {"fp_muls_rank0": muls, **stats, "formula": "6 x slots + 10 x survivors + 28 x buckets"},
                         "note": "frac credits 10 Fp multiplies per (point, window) whatever the kernels spend (6 in the affine "
                                 "rounds): a work-equivalent rate; frac_real = Fp multiplies issued / time / the fe_mul rate (rank 0)"},
            "roofline_hbm": {"kernel": dom, "bound": "hbm", "achieved": per * 128 / dom_s / 1e9, "peak": peaks["hbm_gbs"],
                             "unit": "GB/s", "frac": per * 128 / dom_s / 1e9 / peaks["hbm_gbs"], "traffic": None,
                             "peak_source": peaks["source"], "note": "algorithmic bytes = 128 B/point; not the binding bound"},
            "phases_ms": {k: round(v[0], 4) for k, v in sorted(prof.items())},
            "result": {"msm": res.hex()},
        }
        if args.check or not emit:  # known discrete logs: sum k_i * s_i in the exponent, one scalar multiplication in the oracle
            from oracle import bls12_381 as bls
            acc = 0
            for lo in range(0, n, 1 << 18):
                cnt = min(1 << 18, n - lo)
                sb = synth.fr_words(21, cnt, lo).astype("<u8").tobytes()
                kb = synth.fr_words(22, cnt, lo).astype("<u8").tobytes()
                acc = (acc + sum(int.from_bytes(sb[32 * i:32 * i + 32], "little") * int.from_bytes(kb[32 * i:32 * i + 32], "little")
                                 for i in range(cnt))) % R
            line["parity"] = {"known_dlog_match": bls.g1_compress(bls.g1_mul(bls.G1_GEN, acc)) == res}
        if emit:
            print(json.dumps(line), flush=True)
    bases.free()
    be.close()
    if own_env:
        env.close()
    return line


# ------------------------------------------------------------------ config 5: bivariate rows
def run_bipoly(args, env=None, emit=True):
    own_env = env is None
    env = env or Env()
    torch, dist, rank, world, local, dev, stream = env.torch, env.dist, env.rank, env.world, env.local, env.dev, env.stream
    from fourier_b200 import PianoBackend, shard, synth
    n_scale = int(os.environ.get("FB_BIPOLY_SCALE", "22"))
    m_scale = int(os.environ.get("FB_BIPOLY_MSCALE", "10"))
    M, T = 1 << m_scale, 1 << (n_scale - m_scale)
    row0, rows = shard.split_range(M, world, rank)  # this rank's rows (M and world are powers of two: equal shares)
    be = PianoBackend(n_scale, m_scale, device=local, window_bits=int(os.environ.get("FB_BENCH_WINDOW", "0")))
    be.set_stream(stream.cuda_stream)
    t0 = time.time()
    be.setup(SECRET_X, SECRET_Y, row0, rows)  # this rank keeps rows [row0, row0 + rows) of U resident
    setup_s = time.time() - t0
    coeffs_np = synth.fr_bytes(51, rows * T, row0 * T)  # the local rows of the 2^m x 2^t matrix (Lagrange basis)
    alpha_b = synth.fr_bytes(53, 1).tobytes()
    beta_b = synth.fr_bytes(54, 1).tobytes()
    h_coeffs = torch.from_numpy(coeffs_np.reshape(-1)).pin_memory()
    d_coeffs = _dev_fr(torch, be, dev, coeffs_np)
    d_alpha = _dev_fr(torch, be, dev, synth.fr_bytes(53, 1))
    d_out = torch.zeros(rows * 128, dtype=torch.uint8, device=dev)  # per row: 48 commit | 32 y | 48 proof
    d_c, d_y, d_p = d_out[:rows * 48], d_out[rows * 48:rows * 80], d_out[rows * 80:]
    gathered = torch.zeros(world * rows * 128, dtype=torch.uint8, device=dev) if world > 1 else None

    be.set_lanes(int(os.environ.get("FB_BENCH_LANES", "2")))
    be.set_msm_split(int(os.environ.get("FB_BENCH_SPLIT", "1")))

    fused_rows = os.environ.get("FB_BENCH_ROWS", "fused") == "fused"  # "two": dev_commit_rows + dev_open_rows on two lanes

    def step_device():
        if fused_rows:
            be.dev_commit_open_rows(row0, rows, d_coeffs, T, d_alpha, d_c, d_y, d_p)
        else:
            be.dev_commit_rows(row0, rows, d_coeffs, T, d_c)
            be.dev_open_rows(row0, rows, d_coeffs, T, d_alpha, d_y, d_p)
        be.join()
        if world > 1:
            dist.all_gather_into_tensor(gathered, d_out)

    def unpack(buf: bytes):
        blobs = [buf[k * rows * 128:(k + 1) * rows * 128] for k in range(world)]
        return shard.unpack_rows(blobs, [rows] * world)

    for _ in range(max(args.warmup, 3)):
        step_device()
    _barrier(torch, dist, world)
    l0 = be.launch_count()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _barrier(torch, dist, world)
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    _barrier(torch, dist, world)
    ms_dev = e0.elapsed_time(e1) / args.steps
    launches = be.launch_count() - l0
    clk = clocks.stop()
    cs, ys, ps = unpack(bytes((gathered if world > 1 else d_out).cpu().numpy()))

    # e2e: host coefficient rows through commit_rows / open_rows, gather, master step on rank 0
    def step_host():
        if fused_rows:
            c, y, p = be.commit_open_rows(row0, rows, h_coeffs, alpha_b)
        else:
            c = be.commit_rows(row0, rows, h_coeffs)
            y, p = be.open_rows(row0, rows, h_coeffs, alpha_b)
        blob = shard.pack_rows(c, y, p)
        if world > 1:
            mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(dev)
            dist.all_gather_into_tensor(gathered, mine)
            blob = bytes(gathered.cpu().numpy())
        if rank != 0:
            return None
        hc, hy, hp = unpack(blob)
        com = be.master_commit(hc)
        z, (pi0, pi1) = be.master_open(hy, hp, beta_b)
        return hc, hy, hp, com, z, pi0, pi1

    step_host()
    _barrier(torch, dist, world)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        hres = step_host()
    _barrier(torch, dist, world)
    ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps
    ms_dev, ms_e2e = _max_over_ranks(torch, dist, world, dev, [ms_dev, ms_e2e])

    be.set_lanes(1)
    be.profile(True)
    step_device()  # (scratch of the single-stream shape grows here, untimed)
    torch.cuda.synchronize()
    be.profile_read()
    be.msm_stats()
    step_device()
    torch.cuda.synchronize()
    prof = be.profile_read()
    stats = be.msm_stats()
    be.profile(False)
    wbits, nwin = be.window()
    line = None
    if rank == 0:
        assert (hres[0], hres[1], hres[2]) == (cs, ys, ps), "host-API and device-resident results differ"
        peaks = load_peaks()
        int_peak, int_src, fp_mul_rate = load_int_peak()
        muls = stats["round_slots"] * 6 + stats["survivors"] * 10 + stats["buckets"] * 28
        pts = 2 * M * T  # points through the MSMs of one step (commit + open)
        msm_ms = sum(v[0] for k, v in prof.items() if k.startswith("msm."))
        dom = max((k for k in prof if k.startswith("msm.")), key=lambda k: prof[k][0])
        dom_s = prof[dom][0] / max(prof[dom][1], 1) * 1e-3
        line = {
            "metric": "bipoly_commit_open_ms_2^%dx2^%d" % (m_scale, n_scale - m_scale), "value": ms_dev, "unit": "ms",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_dev,
            "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
            "config": {"workload": "bivariate (piano bench) worker commit+open of 2^%d rows x 2^%d coefficients "
                                   "(scale=%d, machines_scale=%d, BASELINE config 5)" % (m_scale, n_scale - m_scale, n_scale, m_scale),
                       "parallelism": "1 GPU" if world == 1 else "rows sharded over %d GPUs (%d rows each), all-gather of "
                                      "128 bytes per row, master step on rank 0" % (world, rows),
                       "l2": "inputs larger than L2 (window tables of the local rows + coefficient matrix)",
                       "srs_setup_s": round(setup_s, 2), "window_bits": wbits, "windows": nwin,
                       "lanes": int(os.environ.get("FB_BENCH_LANES", "2")),
                       "step": "fb_dev_commit_open_rows (one call; commit MSMs on a sub-stream)" if fused_rows else "fb_dev_commit_rows + fb_dev_open_rows on two lanes", "msm_split": int(os.environ.get("FB_BENCH_SPLIT", "1"))},
            "e2e": {"value": ms_e2e, "unit": "ms", "h2d_bytes_per_step": (1 if fused_rows else 2) * rows * T * 32 + 32,
                    "d2h_bytes_per_step": rows * 128 + (world * rows * 128 if world > 1 else 0) + 128,
                    "includes": "master_commit + master_open on rank 0"},
            "gpu_launches": int(launches), "clocks": clk,
            "roofline": {"kernel": "row MSMs (commit + open), whole pipeline", "bound": "int",
                         "achieved": (pts / world) * 66000 / (msm_ms * 1e-3) / 1e12 if msm_ms else None, "peak": int_peak / 1e12,
                         "unit": "T wide-MAD/s", "frac": (pts / world) * 66000 / (msm_ms * 1e-3) / int_peak if msm_ms else None,
                         "frac_real": muls / (msm_ms * 1e-3) / fp_mul_rate if msm_ms else None, "traffic": None,
                         "peak_source": int_src, "work": "66000 wide-MAD/point at the fixed accounting window for 2^12 rows (SURVEY 8d)",
                         "real_work": {"fp_muls_rank0": muls, **stats, "formula": "6 x slots + 10 x survivors + 28 x buckets"},
                         "note": "frac credits 10 Fp multiplies per (point, window) whatever the kernels spend (6 in the affine "
                                 "rounds): a work-equivalent rate; frac_real = Fp multiplies issued / time / the fe_mul rate (rank 0)"},
            "roofline_hbm": {"kernel": dom, "bound": "hbm", "achieved": rows * T * 128 / dom_s / 1e9, "peak": peaks["hbm_gbs"],
                             "unit": "GB/s", "frac": rows * T * 128 / dom_s / 1e9 / peaks["hbm_gbs"], "traffic": None,
                             "peak_source": peaks["source"], "note": "algorithmic bytes = 128 B/point; not the binding bound"},
            "phases_ms": {k: round(v[0], 4) for k, v in sorted(prof.items())},
            "extra": {"msm_mpts_per_s": pts / (ms_dev * 1e-3) / 1e6},
            "result": {"master_commit": hres[3].hex(), "z": hres[4].hex(), "pi0": hres[5].hex(), "pi1": hres[6].hex()},
        }
        if args.check:
            # known-tau identities on a sample of rows + the master pairing equation in the oracle
            from oracle import bls12_381 as bls
            from oracle import piano as opiano
            o = opiano.PianoOracle(n_scale, m_scale)
            o.generate_trusted_setup(SECRET_X, SECRET_Y, with_points=False)
            alpha = int.from_bytes(alpha_b, "big")
            beta = int.from_bytes(beta_b, "big")
            okc = oko = True
            for i in (0, 1, M // 2 + 3, M - 1):
                row = synth.fr_ints(51, T, i * T)
                okc &= cs[i] == bls.g1_compress(o.worker_commit_known_tau(i, row))
                oy, opi = o.worker_open_known_tau(i, row, alpha)
                oko &= ys[i] == oy.to_bytes(32, "big") and ps[i] == bls.g1_compress(opi)
            line["parity"] = {"rows_commit_known_tau": bool(okc), "rows_open_known_tau": bool(oko)}
            try:
                line["parity"]["master_verify"] = bool(o.master_verify(
                    bls.g1_decompress(hres[3]), beta, alpha, int.from_bytes(hres[4], "big"),
                    (bls.g1_decompress(hres[5]), bls.g1_decompress(hres[6]))))
            except Exception as exc:  # the oracle's pairing needs the G2 side of the setup
                line["parity"]["master_verify"] = "skipped: %s" % exc
        if emit:
            print(json.dumps(line), flush=True)
    be.close()
    if own_env:
        env.close()
    return line


# ------------------------------------------------------------------ the setup-point loader
def srs_load_lines(be, n_scale, m_scale, tag=None):
    """Throughput of PianoSettings::load_setup_from_file's replacement (piano.rs:711-846): the context `be` (whole SRS
    resident) writes its setup in both encodings to a file, and a fresh context loads each: file read, H2D through the
    pinned ring, on-GPU decompression (one Fp square root per point) / deserialisation, window-table build."""
    import tempfile
    from fourier_b200 import PianoBackend
    M, T = 1 << m_scale, 1 << (n_scale - m_scale)
    npts = 1 + T + M + M * T
    out = {}
    tag = tag or ("n%d_m%d" % (n_scale, m_scale))
    with tempfile.TemporaryDirectory() as d:
        for compressed in (True, False):
            path = os.path.join(d, "setup")
            be.save_setup_to_file(path, compressed)
            size = os.path.getsize(path)
            b2 = PianoBackend(n_scale, m_scale)
            t0 = time.perf_counter()
            b2.load_setup_from_file(path, compressed)
            ms = (time.perf_counter() - t0) * 1e3
            b2.set_window(-1)  # no tables: isolates read + upload + decode
            t0 = time.perf_counter()
            b2.load_setup_from_file(path, compressed)
            ms_nt = (time.perf_counter() - t0) * 1e3
            same = b2.setup_to_bytes(True)[:4096] == be.setup_to_bytes(True)[:4096] if npts < (1 << 21) else None
            b2.close()
            out["%s_%s" % (tag, "compressed" if compressed else "uncompressed")] = {
                "file_mib": size / 2**20, "points": npts, "ms": ms, "points_per_s": npts / (ms * 1e-3),
                "ms_without_tables": ms_nt, "points_per_s_without_tables": npts / (ms_nt * 1e-3),
                "gib_per_s_without_tables": size / 2**30 / (ms_nt * 1e-3), "prefix_matches": same}
            os.unlink(path)
    return out


def srs_load_c5(env):
    """The C5 file (n=22, m=10: 2^22 + 2^12 + 2^10 + 1 points, 192 MiB compressed)."""
    from fourier_b200 import PianoBackend
    be = PianoBackend(22, 10, device=env.local)
    be.setup(SECRET_X, SECRET_Y)
    out = srs_load_lines(be, 22, 10)
    be.close()
    env.torch.cuda.empty_cache()
    return out


RUN = {"ntt": run_ntt, "msm": run_msm, "bipoly": run_bipoly}

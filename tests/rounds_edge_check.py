"""Helper of test_gpu_rounds_edge.py (run in a subprocess: the library reads its tuning knobs once per process).
With FB_MSM_ROUNDS forced, small MSMs go through the batched-affine round kernels, whose special slots - doubling
(P + P), cancellation (P + (-P)), identity inputs, single-entry buckets - random large inputs almost never produce.
Prints "ok <count>" or raises."""
import random
import sys

sys.path.insert(0, __import__("os").path.dirname(__file__))
from util import R, bls  # noqa: E402

from fourier_b200 import PianoBackend  # noqa: E402


def expect(ss, ks):
    return bls.g1_compress(bls.g1_mul(bls.G1_GEN, sum(s * k for s, k in zip(ss, ks)) % R))


def large():
    """A 2^20-point MSM with a table against its known discrete log: the size at which a forced FB_MSM_WAVES takes effect
    (smaller jobs fall back to one chain per thread)."""
    from fourier_b200 import synth
    be = PianoBackend(4, 0)
    n = 1 << 20
    bases = be.bases_from_scalars(synth.fr_bytes(41, n))
    bases.precompute()
    got = be.msm(bases, synth.fr_bytes(42, n))
    assert got == expect(synth.fr_ints(41, n), synth.fr_ints(42, n))
    # skewed scalars: a few very deep buckets, many single-entry ones
    ks = [(k * k) % 1000 for k in range(n)]
    assert be.msm(bases, ks) == expect(synth.fr_ints(41, n), ks)
    bases.free()
    be.close()
    print("ok 2")


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "large":
        return large()
    rng = random.Random(2024)
    be = PianoBackend(4, 0)
    n = 512
    ss = [rng.randrange(1, R) for _ in range(n)]
    ss[1] = ss[0]                # duplicate point
    ss[2] = (R - ss[0]) % R      # its negative
    ss[3] = 0                    # the identity as a base
    for k in range(100, 164):    # a run of 64 equal points
        ss[k] = ss[100]
    for k in range(200, 232):    # P, -P, P, -P, ...
        ss[k] = ss[200] if k % 2 == 0 else (R - ss[200]) % R
    scalar_sets = [
        [rng.randrange(R) for _ in range(n)],
        [12345] * n,                                   # one bucket per window: doubling cascades on the equal run
        [1] * n,
        [R - 1] * n,
        [7 if k % 2 == 0 else R - 7 for k in range(n)],  # equal points with opposite digits cancel inside the rounds
        [0] * n,
        [(k * k) % 5 for k in range(n)],               # many zero digits, tiny buckets
    ]
    checked = 0
    for tab, c in ((False, 0), (True, 0), (True, 5), (True, 9)):
        bases = be.bases_from_scalars(ss)
        if tab:
            bases.precompute(c)
        for ks in scalar_sets:
            got = be.msm(bases, ks)
            assert got == expect(ss, ks), "msm mismatch tab=%s c=%d set=%d" % (tab, c, scalar_sets.index(ks))
            checked += 1
        # fewer scalars than bases, and a single one
        assert be.msm(bases, scalar_sets[0][:37]) == expect(ss, scalar_sets[0][:37])
        assert be.msm(bases, [5]) == expect(ss, [5])
        checked += 2
        bases.free()
    # commit + open through the same kernels (fused pair and the batch-of-two pass)
    from util import SECRET_X, SECRET_Y, fr_list_to_bytes, opiano
    be2 = PianoBackend(8, 0)
    be2.setup(SECRET_X, SECRET_Y)
    o = opiano.PianoOracle(8, 0)
    o.generate_trusted_setup(SECRET_X, SECRET_Y, with_points=False)
    coeffs = [rng.randrange(R) for _ in range(256)]
    coeffs[5] = coeffs[6] = 0
    alpha = rng.randrange(R)
    cb = fr_list_to_bytes(coeffs)
    ab = alpha.to_bytes(32, "big")
    c1 = be2.worker_commit(0, cb)
    y1, p1 = be2.worker_open(0, cb, ab)
    c2, y2, p2 = be2.worker_commit_open(0, cb, ab)
    assert (c1, y1, p1) == (c2, y2, p2)
    oy, opi = o.worker_open_known_tau(0, coeffs, alpha)
    assert c1 == bls.g1_compress(o.worker_commit_known_tau(0, coeffs))
    assert y1 == oy.to_bytes(32, "big") and p1 == bls.g1_compress(opi)
    be2.close()
    be.close()
    print("ok %d" % checked)


if __name__ == "__main__":
    main()

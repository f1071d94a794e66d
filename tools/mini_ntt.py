"""Times dev_ntt at the sizes given (log2) with the library's current FB_NTT_* knobs; prints one line per size.
Also checks the result against a second configuration-independent property: iNTT(NTT(x)) == x."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from fourier_b200 import PianoBackend, synth

sizes = [int(x) for x in sys.argv[1:]] or [16, 18, 20, 22, 24]
dev = torch.device("cuda:0")
torch.cuda.set_device(0)
stream = torch.cuda.Stream(device=dev)  # the library launches on this stream; events are recorded on it too
torch.cuda.set_stream(stream)
be = PianoBackend(max(sizes), 0, device=0)
be.set_stream(stream.cuda_stream)
res = []
for lg in sizes:
    n = 1 << lg
    nbuf = max(1, min(64, (256 << 20) // (n * 32)))
    host = synth.fr_bytes(40 + lg, n)
    d_be = torch.from_numpy(host.reshape(-1)).to(dev)
    bufs = []
    for _ in range(nbuf):
        d = torch.empty(n * 32, dtype=torch.uint8, device=dev)
        be.dev_fr_from_be(d_be, d, n)
        bufs.append(d)
    out = torch.empty(n * 32, dtype=torch.uint8, device=dev)
    tmp = torch.empty_like(out)
    for k in range(3):
        be.dev_ntt(bufs[k % nbuf], out, n, False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record(stream)
    for k in range(reps):
        be.dev_ntt(bufs[k % nbuf], out, n, False)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    be.dev_ntt(bufs[0], out, n, False)
    chk = int(out.view(torch.int64).sum().item()) & 0xFFFFFFFF  # equal across configurations
    be.dev_ntt(out, tmp, n, True)
    torch.cuda.synchronize()
    ok = bool(torch.equal(tmp, bufs[0]))
    res.append("2^%d %.4f ms %.2f Ge/s %s %08x" % (lg, ms, n / ms / 1e6, "ok" if ok else "ROUNDTRIP-FAIL", chk))
    del bufs, out, tmp, d_be
print(" | ".join(res), flush=True)
be.close()

"""Smallest end-to-end call (debug aid): one worker_commit at scale FB_MINI_SCALE."""
import os, random, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fourier_b200 import PianoBackend
n = int(os.environ.get("FB_MINI_SCALE", "10"))
be = PianoBackend(n, 0)
be.setup(bytes([1] * 32), bytes([2] * 32))
rng = random.Random(1)
R = 0x73EDA753299D7D483339D80809A1D80553BDA402FFFE5BFEFFFFFFFF00000001
c = be.worker_commit(0, [rng.randrange(R) for _ in range(1 << n)])
print("commit", c.hex())
be.close()

"""The reference's own Python client, unmodified, against this repo's server.

/root/reference/fourier/fourier.py spawns `bin run --host .. --port .. --scale .. [--machines-scale ..]`, waits a
second, pings, and test_routine (fourier.py:472-533) then drives randomPoly / randomPoint / fft / workerCommit /
workerOpen / workerVerify / masterCommit / masterOpen / masterVerify and asserts every `valid`.  The file is imported
from where it lies (it exists in the build container only: the tests skip elsewhere); nothing is copied.
CPU: `bin` = tests/oracle_fourier (the shim's command line over the oracle).  GPU: `bin` = bin/fourier."""
import importlib.util
import os
import socket

import pytest

from util import ROOT

REF_CLIENT = "/root/reference/fourier/fourier.py"


def _load_client():
    if not os.path.exists(REF_CLIENT):
        pytest.skip("the reference tree is not mounted here")
    pytest.importorskip("requests")
    spec = importlib.util.spec_from_file_location("reference_fourier_client", REF_CLIENT)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _run_routine(mod, bin_path, scale, machines_scale):
    made = []

    class Recording(mod.Client):
        def __init__(self, *a, **kw):
            super().__init__(*a, **kw)
            made.append(self)

    mod.Client = Recording  # test_routine looks the name up in its module
    try:
        mod.test_routine(host="127.0.0.1", port=_free_port(), bin=bin_path, scale=scale, machines_scale=machines_scale)
        assert made and made[0].cli.is_running(), "the server died during the routine"
        # quirks the server must tolerate: params null on ping, any other method name is an error object
        assert made[0].ping().json() is None
    finally:
        for c in made:
            if c.cli.process is not None:
                c.cli.process.terminate()
                c.cli.process.wait(timeout=30)


def test_reference_client_routine_on_the_oracle_backed_shim():
    mod = _load_client()
    _run_routine(mod, os.path.join(ROOT, "tests", "oracle_fourier"), 4, 1)


@pytest.mark.gpu
def test_reference_client_routine_on_the_gpu_backend(gpu):
    mod = _load_client()
    _run_routine(mod, os.path.join(ROOT, "bin", "fourier"), 6, 2)

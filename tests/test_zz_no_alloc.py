"""pgm_format_noalloc (include/pgm_gmove.h): the allocation-free variant of the device text formatter, with its scan
scratch in a static __device__ array. No counterpart in the reference.

Checked against the existing entry point pgm_debug_format, byte for byte (the text is integer work: tolerance 0),
on two calls in a row that share the static scratch (the second smaller than the first, so status words the first
call left behind would show), through the C ABI with torch only as the owner of the device buffers.
Collected last on purpose: a failure here hides nothing after it.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _values(n, seed):
    rng = np.random.default_rng(seed)
    v = rng.normal(90.0, 25.0, n)
    k = max(1, n // 50)
    v[rng.integers(0, n, k)] = rng.normal(0.0, 1e-6, k)                 # values that round to +-0.00000000
    v[rng.integers(0, n, k)] = -rng.uniform(0, 2000, k)
    v[rng.integers(0, n, k)] = rng.uniform(1e11, 1e300, k)              # the multi-limb path
    v[rng.integers(0, n, 3)] = [np.inf, -np.inf, np.nan]
    v[rng.integers(0, n, 2)] = [0.0, -0.0]
    return v


def _noalloc(lib, torch, vals, cap=None):
    """One call of the new entry point on torch's current stream; returns (text bytes, length)."""
    dev = torch.cuda.current_device()
    x = torch.from_numpy(vals).cuda()
    cap = int(cap if cap is not None else vals.size * 330 + 16)
    text = torch.full((cap + 1,), 0x7e, dtype=torch.uint8, device="cuda")           # '~' canary everywhere
    ln = torch.zeros(1, dtype=torch.int64, device="cuda")
    rc = lib.pgm_format_noalloc(dev, torch.cuda.current_stream().cuda_stream, x.data_ptr(), vals.size, text.data_ptr(), cap, ln.data_ptr())
    assert rc == 0, lib.pgm_last_error(None)
    torch.cuda.synchronize()
    n = int(ln.item())
    t = text.cpu().numpy().tobytes()
    assert t[min(n, cap):] == b"~" * (cap + 1 - min(n, cap))                        # nothing past the text / past cap
    return t[:min(n, cap)], n


def test_noalloc_matches_debug_format_twice():
    import torch
    from poregen_b200 import _capi as capi, gmove as G, workloads
    lib = capi.load()
    cfg, opts = workloads.c2(n_reads=4)
    with G.GmoveEngine(opts) as g:
        a, b = _values(200_000, 1), _values(5_000, 2)
        want_a, want_b = g.format_doubles(a), g.format_doubles(b)
        got_a, n_a = _noalloc(lib, torch, a)
        free1 = torch.cuda.mem_get_info()[0]
        got_b, n_b = _noalloc(lib, torch, b)                     # second call in a row: the shared scratch holds call 1's words
        assert (n_a, n_b) == (len(want_a), len(want_b))
        assert got_a == want_a and got_b == want_b
        assert want_b.endswith(b";") and want_b.count(b";") == 1
        got_a2, _ = _noalloc(lib, torch, a)                      # and a larger one after the smaller
        assert got_a2 == want_a
        free2 = torch.cuda.mem_get_info()[0]
        assert free2 >= free1 - (64 << 20)                       # torch's caching allocator owns the test's buffers; the entry point adds nothing


def test_noalloc_limits_and_cap():
    import torch
    from poregen_b200 import _capi as capi
    lib = capi.load()
    dev = torch.cuda.current_device()
    ln = torch.full((1,), -1, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    assert lib.pgm_format_noalloc(dev, s, None, 0, None, 0, ln.data_ptr()) == 0     # n = 0: empty text
    torch.cuda.synchronize()
    assert int(ln.item()) == 0
    x = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = lib.pgm_format_noalloc(dev, s, x.data_ptr(), capi.PGM_NOALLOC_MAX_VALUES + 1, None, 0, ln.data_ptr())
    assert rc == capi.PGM_ERR_ARG and b"PGM_NOALLOC_MAX_VALUES" in lib.pgm_last_error(None)
    assert lib.pgm_format_noalloc(dev, s, x.data_ptr(), 8, None, 0, None) == capi.PGM_ERR_ARG
    # cap below the text: whole values only, the length is still the full one
    v = np.array([1.5, -22.25, 333.125], dtype=np.float64)
    full, n = _noalloc(lib, torch, v)
    assert full == b"1.50000000,-22.25000000,333.12500000;" and n == len(full)
    part, n2 = _noalloc(lib, torch, v, cap=30)
    assert n2 == n and part[:24] == full[:24] and part[24:] == b"~" * 6


def test_noalloc_under_graph_capture():
    import torch
    from poregen_b200 import _capi as capi, gmove as G, workloads
    lib = capi.load()
    dev = torch.cuda.current_device()
    v = _values(30_000, 3)
    x = torch.from_numpy(v).cuda()
    text = torch.zeros(v.size * 330, dtype=torch.uint8, device="cuda")
    ln = torch.zeros(1, dtype=torch.int64, device="cuda")
    assert lib.pgm_format_noalloc(dev, torch.cuda.current_stream().cuda_stream, None, 0, None, 0, ln.data_ptr()) == 0   # per-device set-up outside capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        rc = lib.pgm_format_noalloc(dev, torch.cuda.current_stream().cuda_stream, x.data_ptr(), v.size, text.data_ptr(), text.numel(), ln.data_ptr())
    assert rc == 0, lib.pgm_last_error(None)
    cfg, opts = workloads.c2(n_reads=4)
    with G.GmoveEngine(opts) as g:
        for _ in range(2):                                       # two replays: the memset node clears the scratch each time
            text.zero_(); ln.zero_()
            graph.replay()
            torch.cuda.synchronize()
            n = int(ln.item())
            assert text[:n].cpu().numpy().tobytes() == g.format_doubles(v)

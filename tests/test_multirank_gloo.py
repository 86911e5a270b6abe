"""World-size-2 gloo test of the multi-GPU quota rule on CPU (host logic only): each rank counts
the passing events of its contiguous record range, the count vectors are all-gathered, the
exclusive scan over ranks gives each rank the quota already consumed before it, and the per-rank
selections concatenated in rank order must equal the serial run."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle_lib as ol
    from cases import _o
    from poregen_b200 import synth
    from poregen_b200.gmove import quota_exchange
    from poregen_b200.synth import SynthConfig
    import copy
    opts = _o(kmer_size=4, sample_limit=40, scaling=1)
    batch = synth.make_batch(SynthConfig(n_reads=64, median_samples=3000, seed=78))
    per = batch.n_reads // world
    shard = batch.slice(rank * per, (rank + 1) * per)
    ns = opts.index_end - opts.index_start + 1
    # counts of passing events = selection with an unreachable limit
    big = copy.copy(opts)
    big.sample_limit = 2 ** 31 - 1
    counts = ol.oracle_run(big, shard).slot_sel.astype(np.int64)
    gathered = [torch.zeros(ns, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(counts))
    prior, consumed = quota_exchange(np.stack([g.numpy() for g in gathered]), np.zeros(ns, np.uint32), opts.sample_limit)
    totals = prior[rank].copy()
    res = ol.oracle_run(opts, shard, totals)
    q.put((rank, res.slot_sel, res.ev_len, res.ev_start, res.ev_read + rank * per, res.samples, consumed))
    dist.barrier()
    dist.destroy_process_group()


def test_quota_exchange_world2():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from cases import _o
    from poregen_b200 import synth
    from poregen_b200.synth import SynthConfig
    ol.build_oracle()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    opts = _o(kmer_size=4, sample_limit=40, scaling=1)
    batch = synth.make_batch(SynthConfig(n_reads=64, median_samples=3000, seed=78))
    ref = ol.oracle_run(opts, batch)
    ns = ref.n_slots
    np.testing.assert_array_equal(got[0][6], ref.slot_total)
    np.testing.assert_array_equal(got[0][1] + got[1][1], ref.slot_sel)
    offs = [np.concatenate(([0], np.cumsum(g[1].astype(np.int64)))) for g in got]
    soffs = [np.concatenate(([0], np.cumsum(g[2].astype(np.int64)))) for g in got]
    ev_len, ev_start, ev_read, samples = [], [], [], []
    for s in range(ns):
        for g in got:
            r = g[0]
            a, b = offs[r][s], offs[r][s + 1]
            ev_len.append(g[2][a:b]); ev_start.append(g[3][a:b]); ev_read.append(g[4][a:b]); samples.append(g[5][soffs[r][a]:soffs[r][b]])
    np.testing.assert_array_equal(np.concatenate(ev_len), ref.ev_len)
    np.testing.assert_array_equal(np.concatenate(ev_start), ref.ev_start)
    np.testing.assert_array_equal(np.concatenate(ev_read), ref.ev_read)
    np.testing.assert_array_equal(np.concatenate(samples).view(np.uint64), ref.samples.view(np.uint64))

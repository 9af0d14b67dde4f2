"""World-size-2 test of the multi-GPU host logic on CPU (gloo): shard by window centre, each rank takes the reads of its
own halo, results gathered in window order must equal the single-process result.  The per-rank 'compute' here is the
CPU oracle (the CUDA path needs a GPU); what is under test is the partitioning, the shard construction and the gather."""
import os
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rank_ranges_cover_exactly():
    from deepore_b200.sharding import all_ranges
    for n in (0, 1, 7, 8, 1000, 90000):
        for w in (1, 2, 3, 8):
            rr = all_ranges(n, w)
            assert rr[0][0] == 0 and rr[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(rr, rr[1:]))
            sizes = [b - a for a, b in rr]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import pileup_oracle as po
    from deepore_b200 import synth
    from deepore_b200.sharding import gather_on_host, rank_range
    from deepore_b200.workload import make_shard
    ref = synth.random_reference(30_000, seed=1)
    rb = synth.simulate_reads(ref, coverage=20, seed=2)
    wq = synth.window_starts(500, 29_500, stride=100)
    w0, w1 = rank_range(wq.shape[0], world, rank)
    sh = make_shard(rb, ref, 0, wq[w0:w1])                       # this rank's reads + reference halo only
    img, st = po.encode_windows(sh.reads, sh.win_qs, sh.win_lo, sh.win_hi, sh.ref, sh.ref_base)
    digest = np.stack([img.reshape(img.shape[0], -1).astype(np.uint64).sum(axis=1), st.astype(np.uint64)], axis=1)
    full = gather_on_host(digest, dst=0)
    if rank == 0:
        np.save(out_path, full)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shards_equal_single_process(tmp_path):
    import torch.multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pileup_oracle as po
    from deepore_b200 import synth
    out = str(tmp_path / "gathered.npy")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    ref = synth.random_reference(30_000, seed=1)
    rb = synth.simulate_reads(ref, coverage=20, seed=2)
    wq = synth.window_starts(500, 29_500, stride=100)
    lo, hi = rb.window_ranges(wq, 100)
    img, st = po.encode_windows(rb, wq, lo, hi, ref, 0)
    want = np.stack([img.reshape(img.shape[0], -1).astype(np.uint64).sum(axis=1), st.astype(np.uint64)], axis=1)
    assert got.shape == want.shape and np.array_equal(got, want)

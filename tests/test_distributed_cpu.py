"""CPU coverage of the multi-GPU host logic under gloo, world_size 2.

The GPU render cannot run here, so each rank's share of the samples is rendered by the ORACLE (checker only) with the
same sample_offset the CUDA path would get; what is under test is pathtrace_b200.distributed: the spp split and the
single SUM-reduce.  The property: the union of the ranks' samples is exactly the single-process sample set, so the
reduced sums equal a one-process render of all spp up to floating-point summation order.
"""
import os
import socket

import numpy as np
import pytest

from conftest import oracle_camera


def test_split_spp_partitions_every_sample_exactly_once():
    from pathtrace_b200.distributed import split_spp
    for spp in (0, 1, 2, 7, 8, 1024, 16384, 4097):
        for world in (1, 2, 3, 4, 8):
            parts = [split_spp(spp, world, r) for r in range(world)]
            assert sum(n for n, _ in parts) == spp
            nxt = 0
            for n, off in parts:                       # contiguous, disjoint, in rank order
                assert off == nxt
                nxt += n
            sizes = [n for n, _ in parts]
            assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    with pytest.raises(ValueError):
        split_spp(8, 2, 2)
    with pytest.raises(ValueError):
        split_spp(8, 0, 0)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, spp, out_path):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "tests"))
    import torch
    import torch.distributed as dist
    from oracle import pyoracle as O
    from pathtrace_b200 import scenes
    from pathtrace_b200.distributed import reduce_sum, split_spp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh, mt, cam = scenes.default_scene(24, 16)
        oc = oracle_camera(O, cam)
        n, off = split_spp(spp, world, rank)
        acc = np.zeros((16, 24, 3), np.float64)
        if n:
            acc = O.render(oc, sh, mt, O.RenderParams(spp=n, sample_offset=off), "f64", recursive=False, nthreads=1)["sum"]
        t = torch.from_numpy(np.ascontiguousarray(acc))
        reduce_sum(t, 0)
        if rank == 0:
            np.save(out_path, t.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("spp", [5, 8])
def test_spp_split_reduce_equals_single_process(spp, oracle, tmp_path):
    import torch.multiprocessing as mp
    from pathtrace_b200 import scenes
    out = str(tmp_path / "sum.npy")
    mp.spawn(_worker, args=(2, _free_port(), spp, out), nprocs=2, join=True)
    got = np.load(out)
    sh, mt, cam = scenes.default_scene(24, 16)
    ref = oracle.render(oracle_camera(oracle, cam), sh, mt, oracle.RenderParams(spp=spp, sample_offset=0), "f64",
                        recursive=False, nthreads=1)["sum"]
    assert got.shape == ref.shape
    assert np.allclose(got, ref, rtol=1e-12, atol=1e-12), np.abs(got - ref).max()
    assert ref.sum() > 0


def test_reduce_sum_is_a_noop_without_a_process_group():
    import torch
    from pathtrace_b200.distributed import reduce_sum
    t = torch.arange(6, dtype=torch.float32)
    assert reduce_sum(t) is t and t.tolist() == [0, 1, 2, 3, 4, 5]

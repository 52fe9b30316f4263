"""ShardedPairHmm on real GPUs: two ranks (one process per GPU, NCCL over NVLink) must return what a single engine
returns -- posterior pairs merged in input order (the reference consumes cigars positionally,
src/margin/marginAlignLib.py:334) and the all-reduced expectation vector (replaces the file sums of
submodules/cPecan/cPecanEm.py:166-188).  Skipped on a box with fewer than two GPUs."""
import os
import socket

import numpy as np
import pytest

import marginalign_trna_b200 as mab

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out_path):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda:%d" % rank))
    try:
        b = mab.synth.make_trna_batch(3000, seed=321)
        model = mab.PairHmmModel.from_file(mab.synth.default_hmm_path())
        eng = mab.PairHmmEngine(device="cuda:%d" % rank)
        sh = mab.shard.ShardedPairHmm(eng)
        assert sh.world == world
        res = sh.get_aligned_pairs_using_anchors(model, b)
        T, E, lik = sh.get_expectations_using_anchors(model, b)
        cig = sh.realign(model, b)
        if rank == 0:
            np.savez(out_path, offsets=res.offsets, x=res.x, y=res.y, p=res.p, exp=np.concatenate([T, E, [lik]]),
                     c_off=cig.off_ops, c_ops=cig.ops, c_lens=cig.lens)
        else:
            assert res is None
            np.save(out_path + ".rank%d.npy" % rank, np.concatenate([T, E, [lik]]))
    finally:
        dist.destroy_process_group()


def test_two_ranks_equal_one_engine(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "sharded.npz")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    b = mab.synth.make_trna_batch(3000, seed=321)
    model = mab.PairHmmModel.from_file(mab.synth.default_hmm_path())
    eng = mab.PairHmmEngine(device="cuda:0")
    one = eng.get_aligned_pairs_using_anchors(model, b)
    for k in ("offsets", "x", "y", "p"):
        assert np.array_equal(got[k], getattr(one, k)), k
    cig = eng.realign(model, b)
    assert np.array_equal(got["c_off"], cig.off_ops) and np.array_equal(got["c_ops"], cig.ops)
    assert np.array_equal(got["c_lens"], cig.lens)
    T, E, lik = eng.get_expectations_using_anchors(model, b)
    want = np.concatenate([T, E, [lik]])
    np.testing.assert_allclose(got["exp"], want, rtol=1e-11, atol=0)
    # every rank holds the same reduced vector
    np.testing.assert_array_equal(np.load(out + ".rank1.npy"), got["exp"])

"""Host-side sharding logic on CPU: partition, merge in input order, and the expectation all-reduce
under torch.distributed with the gloo backend, world size 2.

The compute inside each rank is stubbed with the CPU oracle here (tests may do that; the product engine is CUDA-only):
what is under test is shard.py -- that a sharded run returns exactly what an unsharded run returns.
"""
import os
import socket
import sys

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from marginalign_trna_b200 import shard
from marginalign_trna_b200.pairhmm import PosteriorPairs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleEngine:
    """Test double with PairHmmEngine's two methods, computing with the CPU oracle."""
    device = "cpu"

    def __init__(self):
        from oracle import pyoracle as po
        self.po = po
        hmm = mab.Hmm.loadHmm(os.path.join(ROOT, "tests", "golden", "input_model.hmm"))
        self.om = po.model_from_hmm(hmm._modelTypeInt(), hmm.transitions, hmm.emissions)

    def _oparams(self, P):
        return self.po.make_params(expansion=P.diagonal_expansion, split=P.split_matrix_bigger_than ** 2)

    def get_aligned_pairs_using_anchors(self, model, batch, params):
        po, offs, xs, ys, ps = self.po, [0], [], [], []
        for i in range(batch.n):
            anc = po.anchors_from_cigar(*batch.cigar(i), batch.x(i), batch.y(i))
            pr = po.aligned_pairs(self.om, batch.x(i), batch.y(i), anc, self._oparams(params))
            offs.append(offs[-1] + len(pr))
            ps.append(pr[:, 0]); xs.append(pr[:, 1]); ys.append(pr[:, 2])
        cat = lambda l: (np.concatenate(l) if l else np.zeros(0)).astype(np.int32)
        return PosteriorPairs(np.array(offs, dtype=np.int64), cat(xs), cat(ys), cat(ps))

    def realign(self, model, batch, params, gap_gamma=0.5, match_gamma=0.0):
        po, off, ops, lens = self.po, [0], [], []
        for i in range(batch.n):
            anc = po.anchors_from_cigar(*batch.cigar(i), batch.x(i), batch.y(i))
            rows, _chain = po.realign(self.om, batch.x(i), batch.y(i), anc, self._oparams(params), gap_gamma, match_gamma)
            ops += rows[:, 0].tolist()
            lens += rows[:, 1].tolist()
            off.append(len(ops))
        return mab.RealignedCigars(np.array(off, dtype=np.int64), np.array(ops, dtype=np.uint8),
                                   np.array(lens, dtype=np.int32))

    def get_expectations_using_anchors(self, model, batch, params, pseudo_count=1e-12, plan_cache=None):
        po = self.po
        if plan_cache is not None:  # what a PairHmmEngine would key its kept device plans on
            self.cache_calls = getattr(self, "cache_calls", []) + [(batch, plan_cache)]
        T, E, lik = np.full(25, pseudo_count), np.full(80, pseudo_count), 0.0
        for i in range(batch.n):
            anc = po.anchors_from_cigar(*batch.cigar(i), batch.x(i), batch.y(i))
            T, E, lik = po.expectations(self.om, batch.x(i), batch.y(i), anc, self._oparams(params), T=T, E=E, lik=lik)
        return T, E, lik


def test_partition_is_balanced_deterministic_and_complete():
    rng = np.random.default_rng(3)
    w = rng.integers(100, 5000, size=10_001)
    for world in (1, 2, 3, 8):
        parts = shard.partition(w, world)
        assert len(parts) == world
        allidx = np.concatenate(parts)
        assert np.array_equal(np.sort(allidx), np.arange(len(w)))
        for p in parts:
            assert np.all(np.diff(p) > 0)  # ascending: a shard preserves input order
            assert len(p) == 0 or p[-1] - p[0] + 1 == len(p)  # ... and is a slice of it
        sums = np.array([w[p].sum() for p in parts])
        assert np.abs(sums - w.sum() / world).max() <= w.max()  # each shard within one item of the ideal share
        again = shard.partition(w, world)
        assert all(np.array_equal(a, b) for a, b in zip(parts, again))
    assert [len(p) for p in shard.partition(np.array([5, 1]), 4)] == [1, 0, 0, 1]
    assert np.array_equal(shard.partition_bounds(np.array([3, 3, 3, 3]), 2), [0, 2, 4])
    assert [len(p) for p in shard.partition(np.zeros(0), 2)] == [0, 0]


def test_merge_pairs_restores_input_order():
    eng = OracleEngine()
    b = mab.synth.make_trna_batch(12, seed=21)
    P = mab.AlignmentParameters()
    whole = eng.get_aligned_pairs_using_anchors(None, b, P)
    shards = shard.partition(shard.estimate_cells(b, P), 3)
    parts = []
    for idx in shards:
        r = eng.get_aligned_pairs_using_anchors(None, b.subset(idx), P)
        parts.append((r.offsets, r.x, r.y, r.p))
    merged = shard.merge_pairs(parts, shards, b.n)
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(getattr(merged, a), getattr(whole, a)), a


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        eng = OracleEngine()
        b = mab.synth.make_trna_batch(10, seed=33)
        P = mab.AlignmentParameters()
        sh = shard.ShardedPairHmm(eng, comm_device="cpu")
        assert (sh.rank, sh.world) == (rank, world)
        res = sh.get_aligned_pairs_using_anchors(None, b, P, dst=0)
        T, E, lik = sh.get_expectations_using_anchors(None, b, P)
        np.savez(os.path.join(out_dir, "exp_%d.npz" % rank), T=T, E=E, lik=lik)
        cig = sh.realign(None, b, P)
        if rank == 0:
            np.savez(os.path.join(out_dir, "cigars.npz"), off=cig.off_ops, ops=cig.ops, lens=cig.lens)
        else:
            assert cig is None
        if rank == 0:
            np.savez(os.path.join(out_dir, "pairs.npz"), offsets=res.offsets, x=res.x, y=res.y, p=res.p)
        else:
            assert res is None
        # EM over the sharded runner: every rank must end with the same model
        opts = mab.em.EmOptions(inputModel=os.path.join(ROOT, "tests", "golden", "input_model.hmm"), iterations=2)
        hmm = mab.em.expectation_maximisation_trials(_NoModelRunner(sh), b, None, opts)
        np.savez(os.path.join(out_dir, "em_%d.npz" % rank), T=hmm.transitions, E=hmm.emissions,
                 L=np.array(hmm.runningLikelihoods))
        # the EM loop hands one PlanCache to every iteration, and the sharder gives the engine the SAME shard object
        # each time -- that identity is what lets a PairHmmEngine re-use the device plans of its shard
        calls = eng.cache_calls
        assert len(calls) == 2 and calls[0][0] is calls[1][0] and calls[0][1] is calls[1][1]
        assert isinstance(calls[0][1], mab.PlanCache)
    finally:
        dist.destroy_process_group()


class _NoModelRunner:
    """EM hands a PairHmmModel (needs the CUDA library's host half only) to the runner; the oracle stub rebuilds its
    own model from the same numbers, so the handle is used to refresh the oracle model each iteration."""

    def __init__(self, sharded):
        self.sh = sharded

    def get_expectations_using_anchors(self, model, batch, params, plan_cache=None):
        T = np.zeros((13, 4))
        eng = self.sh.engine
        # rebuild the oracle model from the CURRENT hmm held by the PairHmmModel's source arrays
        eng.om = eng.po.model_from_hmm(model.type, model.src_transitions, model.src_emissions)
        return self.sh.get_expectations_using_anchors(model, batch, params, plan_cache=plan_cache)


@pytest.mark.timeout(300)
def test_world_size_2_gloo_matches_single_process(tmp_path):
    import torch.multiprocessing as tmp_mp
    port = _free_port()
    tmp_mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    eng = OracleEngine()
    b = mab.synth.make_trna_batch(10, seed=33)
    P = mab.AlignmentParameters()
    whole = eng.get_aligned_pairs_using_anchors(None, b, P)
    got = np.load(tmp_path / "pairs.npz")
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(got[a], getattr(whole, a)), a
    want_c, got_c = eng.realign(None, b, P), np.load(tmp_path / "cigars.npz")
    assert np.array_equal(got_c["off"], want_c.off_ops) and np.array_equal(got_c["ops"], want_c.ops)
    assert np.array_equal(got_c["lens"], want_c.lens) and len(want_c.ops) > 10
    T, E, lik = eng.get_expectations_using_anchors(None, b, P)
    for r in range(2):
        e = np.load(tmp_path / ("exp_%d.npz" % r))
        np.testing.assert_allclose(e["T"], T, rtol=1e-12)
        np.testing.assert_allclose(e["E"], E, rtol=1e-12)
        np.testing.assert_allclose(float(e["lik"]), lik, rtol=1e-12)
    e0, e1 = np.load(tmp_path / "em_0.npz"), np.load(tmp_path / "em_1.npz")
    for k in ("T", "E", "L"):
        assert np.array_equal(e0[k], e1[k]), k
    assert len(e0["L"]) == 2 and e0["L"][1] > e0["L"][0]  # EM likelihood is monotone (reference test_em_*)

"""PlanCache bookkeeping (pairhmm.py) without a GPU: what counts as "the same batch", what refills the cache.  The
device-side behaviour (cached == uncached expectations over several chunks) is tests/test_gpu_em.py."""
import marginalign_trna_b200 as mab


class _Plan:
    def __init__(self, tag):
        self.tag = tag


def test_lookup_needs_the_same_batch_object_parameters_mode_and_device():
    b1 = mab.synth.make_trna_batch(5, seed=1)
    b2 = mab.synth.make_trna_batch(5, seed=1)  # equal content, another object: not the same batch
    P = mab.AlignmentParameters()
    c = mab.PlanCache()
    assert c.lookup(b1, P, True, "cuda:0") is None and c.hits == 0
    plans = [_Plan(0), _Plan(1)]
    c.store(b1, P, True, "cuda:0", plans)
    assert c.lookup(b1, P, True, "cuda:0") is plans and c.hits == 1
    assert c.lookup(b1, mab.AlignmentParameters(), True, "cuda:0") is plans  # equal parameters, another object: fine
    assert c.lookup(b2, P, True, "cuda:0") is None
    assert c.lookup(b1, mab.AlignmentParameters(diagonal_expansion=12), True, "cuda:0") is None
    assert c.lookup(b1, mab.AlignmentParameters(split_matrix_bigger_than=300), True, "cuda:0") is None
    assert c.lookup(b1, P, False, "cuda:0") is None   # posterior-mode plans carry another workspace size
    assert c.lookup(b1, P, True, "cuda:1") is None    # device pointers of another GPU
    assert c.hits == 2
    c.clear()
    assert c.lookup(b1, P, True, "cuda:0") is None and c.plans is None and c.batch is None


def test_the_cache_keeps_its_batch_alive():
    """Identity is only a safe key while the object exists: the cache holds a reference, so a new batch cannot be
    handed the id of the cached one."""
    import gc
    import weakref
    c = mab.PlanCache()
    b = mab.synth.make_trna_batch(3, seed=2)
    ref = weakref.ref(b)
    c.store(b, mab.AlignmentParameters(), True, "cuda:0", [_Plan(0)])
    del b
    gc.collect()
    assert ref() is not None
    c.clear()
    gc.collect()
    assert ref() is None

"""mab_posteriors_compute_noalloc (include/margin_b200.h): posterior mode for callers that may not allocate device
memory.  No counterpart in the reference; collected last on purpose.

The call is checked through the C ABI against mab_posteriors_compute on the same device plan, bit for bit (offsets, pair
count and the emitted pairs), on calls in a row that share the static scratch: an over-wide batch (the case where the
existing entry point calls cudaMallocAsync and the new one uses the static ring), a second over-wide batch with a
different ring width (stale contents of the first call must not matter), the first one again, and a narrow batch that
needs no scratch at all.  The rejection path (a band wider than 69905 cells for five states) would need a gap box of
several 10^9 cells and is not exercised here.
"""
import numpy as np
import pytest

import marginalign_trna_b200 as mab
from marginalign_trna_b200 import capi
from test_gpu_parity import _with_gap_box, gpu_model

pytestmark = pytest.mark.gpu


def _async_pool_high_watermark(device_index):
    """Reserved-bytes high watermark of the device's default cudaMallocAsync pool, or None when it cannot be read."""
    try:
        from cuda.bindings import runtime as cudart
        err, pool = cudart.cudaDeviceGetDefaultMemPool(device_index)
        if int(err) != 0:
            return None
        err, high = cudart.cudaMemPoolGetAttribute(pool, cudart.cudaMemPoolAttr.cudaMemPoolAttrReservedMemHigh)
        return int(high) if int(err) == 0 else None
    except Exception:  # cuda-python missing or a different API level: the parity checks below do not depend on it
        return None


def _compute_and_emit(engine, entry, model, dplan):
    """Posterior mode through `entry` (a C-ABI compute function) + mab_posteriors_emit; host arrays."""
    torch, L = engine.torch, capi.load()
    n = dplan.plan.info.n_alignments
    d_off = torch.full((n + 1,), -1, dtype=torch.int64, device=engine.device)
    d_np = torch.full((1,), -1, dtype=torch.int64, device=engine.device)
    capi.check(entry(model._h, dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes, d_off.data_ptr(),
                     d_np.data_ptr(), engine._stream_ptr()))
    total = int(d_np.item())
    d_x = torch.empty(total, dtype=torch.int32, device=engine.device)
    d_y = torch.empty(total, dtype=torch.int32, device=engine.device)
    d_p = torch.empty(total, dtype=torch.int32, device=engine.device)
    capi.check(L.mab_posteriors_emit(dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes, d_x.data_ptr(),
                                     d_y.data_ptr(), d_p.data_ptr(), total, engine._stream_ptr()))
    torch.cuda.synchronize()
    return total, d_off.cpu().numpy(), d_x.cpu().numpy(), d_y.cpu().numpy(), d_p.cpu().numpy()


def test_no_alloc_entry_point_matches_posteriors_compute(golden):
    engine = mab.PairHmmEngine()
    torch, L = engine.torch, capi.load()
    assert hasattr(L, "mab_posteriors_compute_noalloc")
    model = gpu_model(golden.model_arrays("fixture5asym"))
    P = mab.AlignmentParameters()
    wide_a = _with_gap_box(mab.synth.make_trna_batch(20, seed=71), 2200)
    wide_b = _with_gap_box(mab.synth.make_trna_batch(12, seed=72), 1950)
    narrow = mab.synth.make_trna_batch(200, seed=73)
    plans = {}
    for name, b in (("wide_a", wide_a), ("wide_b", wide_b), ("narrow", narrow)):
        plan = mab.Plan.from_batch(b, P)
        plans[name] = (plan, mab.DevicePlan(plan, engine.device))
    assert plans["wide_a"][0].info.max_width > 1900 and plans["wide_b"][0].info.max_width > 1900
    assert plans["wide_a"][0].info.max_width != plans["wide_b"][0].info.max_width
    assert plans["narrow"][0].info.max_width < 64

    index = engine.device.index if engine.device.index is not None else torch.cuda.current_device()
    want = {}
    for call, name in enumerate(("wide_a", "wide_b", "wide_a", "narrow", "wide_b")):
        dplan = plans[name][1]
        torch.cuda.synchronize()
        high0 = _async_pool_high_watermark(index)
        got = _compute_and_emit(engine, L.mab_posteriors_compute_noalloc, model, dplan)
        high1 = _async_pool_high_watermark(index)
        if high0 is not None and high1 is not None:
            assert high1 == high0, ("the no-allocation call grew the cudaMallocAsync pool", call, name, high0, high1)
        if name not in want:
            want[name] = _compute_and_emit(engine, L.mab_posteriors_compute, model, dplan)
        ref = want[name]
        assert got[0] == ref[0] and got[0] > 0, (call, name, got[0], ref[0])
        for k, what in ((1, "offsets"), (2, "x"), (3, "y"), (4, "p_int")):
            assert np.array_equal(got[k], ref[k]), (call, name, what)

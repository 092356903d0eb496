"""hb_keyswitch_batch_noalloc (include/helm_b200.h; no counterpart in the reference):
the tensor-core keyswitch with its scratch (job descriptors + int8 digit matrix) in a static device array -- no cudaMalloc in the call.
Checked through the C ABI against the existing entry point hb_keyswitch_batch on the same inputs, bit for bit (integer arithmetic:
the tolerance of every keyswitch test here is equality), on two calls in a row that share the static array (the second is smaller,
so stale rows of the first are still in the array), and beyond the stated maximum the call must fail with a clear error.
Collected last (file name) so that a failure cannot hide the other tests under pytest -x."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["toy32", "shortint_m1c1"])
def test_keyswitch_noalloc_matches_keyswitch_batch_twice(name, keys_factory, engine_factory):
    import torch
    K, eng = keys_factory(name), engine_factory(name)
    tdt = torch.int32 if K.bits == 32 else torch.int64
    info = np.iinfo(K.dtype)
    rng = np.random.default_rng(77)
    dev = torch.device("cuda", 0)
    for call, cnt in enumerate((203, 70)):
        cts = rng.integers(0, info.max, size=(cnt, K.kN + 1), dtype=K.dtype, endpoint=True)
        want = eng.keyswitch_batch(cts)                                   # existing entry point (host buffers, growable scratch)
        d_in = torch.from_numpy(cts.view(np.int32 if K.bits == 32 else np.int64)).to(dev)
        d_out = torch.full((cnt, K.n + 1), -1, dtype=tdt, device=dev)
        torch.cuda.synchronize()                                          # inputs visible before the engine's stream reads them
        eng.keyswitch_batch_noalloc(d_in.data_ptr(), cnt, d_out.data_ptr())
        eng.sync()
        got = d_out.cpu().numpy().view(K.dtype)
        assert np.array_equal(got, want), "call %d: rows %s differ" % (call, np.nonzero((got != want).any(axis=1))[0][:8])
        assert np.array_equal(got[0], K.keyswitch(cts[0]))               # and the oracle, for one row


def test_keyswitch_noalloc_rejects_oversize(keys_factory, engine_factory):
    import torch
    import helm_b200
    K, eng = keys_factory("toy32"), engine_factory("toy32")
    buf = torch.zeros(16, dtype=torch.int32, device="cuda")
    with pytest.raises(helm_b200.EngineError, match="exceeds the static scratch"):
        eng.keyswitch_batch_noalloc(buf.data_ptr(), 1345, buf.data_ptr())     # limit for kN * l = 4096 digits per row: 1344
    with pytest.raises(helm_b200.EngineError, match="null buffer"):
        eng.keyswitch_batch_noalloc(0, 4, buf.data_ptr())

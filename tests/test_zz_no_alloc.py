"""ra_detect_events_noalloc (no counterpart in the reference): the no-allocation
form of the event detection, scratch in a static device array.  Collected last on purpose.

Checked through the C ABI against the existing entry point ra_detect_events_batch_dev, bit for bit (the bar of every
event test here), on two calls in a row with different inputs -- the second call finds the first call's scratch."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(seed, n_chunks, max_len):
    rng = np.random.default_rng(seed)
    lens = rng.integers(max_len // 3, max_len + 1, n_chunks).astype(np.uint32)
    lens[0] = max_len; lens[-1] = 7                                  # a full chunk and a tiny one
    offs = np.zeros(n_chunks, np.uint64)
    offs[1:] = np.cumsum((lens[:-1].astype(np.int64) + 3) // 4 * 4)
    total = int(offs[-1]) + int(lens[-1])
    # piecewise-constant levels + noise: the shape the detector is built for
    sig = (np.repeat(rng.normal(90.0, 12.0, total // 8 + 1), 8)[:total] + rng.normal(0.0, 1.5, total)).astype(np.float32)
    return sig, offs, lens


def _run_both(ra, L, ctx, torch, sig, offs, lens, max_len):
    dev = torch.device("cuda:0")
    n = len(lens)
    d_sig = torch.from_numpy(sig).to(dev)
    d_off = torch.from_numpy(offs.astype(np.int64)).to(dev)
    d_len = torch.from_numpy(lens.astype(np.int32)).to(dev)
    cap = int(lens.sum()) + 16
    out = []
    for which in ("existing", "noalloc"):
        d_ev = torch.full((cap,), float("nan"), dtype=torch.float32, device=dev)
        d_eo = torch.full((n + 1,), -1, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        if which == "existing":
            tot = C.c_uint64()
            rc = L.ra_detect_events_batch_dev(ctx._h, d_sig.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), n, max_len,
                                              d_ev.data_ptr(), cap, d_eo.data_ptr(), C.byref(tot))
        else:
            rc = L.ra_detect_events_noalloc(ctx._h, d_sig.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), n, max_len,
                                            d_ev.data_ptr(), cap, d_eo.data_ptr(), None)
        assert rc == 0, L.ra_last_error().decode()
        torch.cuda.synchronize()
        eo = d_eo.cpu().numpy().astype(np.uint64)
        out.append((eo, d_ev.cpu().numpy()[: int(eo[-1])]))
    return out


def test_noalloc_event_detection_equals_the_existing_entry_point_twice_in_a_row():
    import torch
    import rawalign_b200 as ra
    from rawalign_b200 import _lib
    L = _lib.load()
    with ra.Context(0) as ctx:
        # first call: 96 chunks of up to 4000 samples (near the 8 MiB limit); second: other data, other shape, stale scratch
        for seed, n_chunks, max_len in ((1, 96, 4000), (2, 37, 2500)):
            sig, offs, lens = _batch(seed, n_chunks, max_len)
            (eo_a, ev_a), (eo_b, ev_b) = _run_both(ra, L, ctx, torch, sig, offs, lens, max_len)
            assert int(eo_a[-1]) > 20 * n_chunks                     # events were found
            assert np.array_equal(eo_a, eo_b)
            assert np.array_equal(ev_a.view(np.uint32), ev_b.view(np.uint32))     # bit-exact, as every event test here


def test_noalloc_rejects_what_does_not_fit_and_handles_empty_input():
    import torch
    import rawalign_b200 as ra
    from rawalign_b200 import _lib
    L = _lib.load()
    dev = torch.device("cuda:0")
    with ra.Context(0) as ctx:
        sig, offs, lens = _batch(3, 8, 1000)
        d_sig = torch.from_numpy(sig).to(dev); d_off = torch.from_numpy(offs.astype(np.int64)).to(dev); d_len = torch.from_numpy(lens.astype(np.int32)).to(dev)
        d_ev = torch.zeros(16, dtype=torch.float32, device=dev); d_eo = torch.full((9,), -1, dtype=torch.int64, device=dev)
        # 200 chunks of 4096 samples would need 16 MiB of scratch: refused before anything is launched
        rc = L.ra_detect_events_noalloc(ctx._h, d_sig.data_ptr(), d_off.data_ptr(), d_len.data_ptr(), 200, 4096, d_ev.data_ptr(), 16, d_eo.data_ptr(), None)
        assert rc != 0 and b"scratch" in L.ra_last_error()
        rc = L.ra_detect_events_noalloc(ctx._h, None, None, None, 0, 0, None, 0, d_eo.data_ptr(), None)
        assert rc == 0, L.ra_last_error().decode()
        torch.cuda.synchronize()
        assert int(d_eo[0]) == 0

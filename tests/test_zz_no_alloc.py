"""avr_streams_encode_static (no-allocation variant, include/avrecode_b200.h): output == avr_streams_encode's, bit for
bit (the tolerance of every test of this path: exact), on two calls in a row that share the static scratch array.

Collected last on purpose.  Device buffers of the caller come from torch (plumbing of
the TEST; the entry point itself gets raw device pointers and allocates nothing)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bits(seed, nstreams, nbins):
    rng = np.random.RandomState(seed)
    p = np.array([0.5, 0.1, 0.01, 0.001])[np.arange(nstreams) % 4][:, None]
    b = (rng.random_sample((nstreams, nbins)) < p).astype(np.uint8)
    b[min(5, nstreams - 1)] = 1           # all-one / all-zero rows: carry-heavy and empty-ish streams
    b[min(6, nstreams - 1)] = 0
    return b


def _static_encode(A, torch, ctx, bits, digit_bits, model):
    nstreams, nbins = bits.shape
    cap = nstreams * A.streams_coded_bound(nbins, model)
    d_bits = torch.from_numpy(bits).cuda()
    d_coded = torch.full((cap + 16,), 0xA5, dtype=torch.uint8, device="cuda")      # poisoned: stale bytes must not be read back
    d_off = torch.full((nstreams + 1,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    rc = A.streams_encode_static(ctx, d_bits.data_ptr(), nstreams, nbins, digit_bits, model, d_coded.data_ptr(), cap, d_off.data_ptr())
    assert rc == 0, (rc, A.lib().avr_last_error())
    torch.cuda.synchronize()
    off = d_off.cpu().numpy().astype(np.uint64)
    blob = d_coded.cpu().numpy().tobytes()
    assert blob[cap:] == b"\xa5" * 16, "wrote beyond coded_cap"
    return [blob[int(off[s]):int(off[s + 1])] for s in range(nstreams)], int(off[nstreams])


@pytest.mark.parametrize("digit_bits,model", [(16, 0), (8, 1)])
def test_static_entry_point_matches_the_allocating_one_twice_in_a_row(digit_bits, model):
    import torch
    import avr_ctypes as A
    with A.Context(0) as ctx:
        # two calls in a row, the second smaller and with other data: what the first left in the shared array must not show
        for seed, nstreams, nbins in ((1, 1000, 3001), (2, 257, 777)):
            bits = _bits(seed, nstreams, nbins)
            want, _ = A.streams_encode(ctx, bits.tobytes(), nstreams, nbins, digit_bits, model)
            got, total = _static_encode(A, torch, ctx, bits, digit_bits, model)
            assert total == sum(len(w) for w in want)
            assert got == want
            back, _ = A.streams_decode(ctx, got, nstreams, nbins, digit_bits, model)
            assert back == bits.tobytes()


def test_static_entry_point_rejects_what_does_not_fit():
    import torch
    import avr_ctypes as A
    with A.Context(0) as ctx:
        nstreams, nbins = 65536, 2048          # 65536 x 2080 bytes of scratch with the adaptive model: far beyond 8 MiB
        d = torch.zeros(16, dtype=torch.uint8, device="cuda")
        rc = A.streams_encode_static(ctx, d.data_ptr(), nstreams, nbins, 8, 1, d.data_ptr(), 1 << 40, d.data_ptr())
        assert rc == -1                         # AVR_E_ARG, nothing launched
        msg = A.lib().avr_last_error().decode()
        assert "static array holds" in msg and str(A.STREAMS_STATIC_SCRATCH_BYTES) in msg
        # an output buffer below the documented bound is refused as well
        bits = _bits(3, 8, 64)
        db = torch.from_numpy(bits).cuda()
        out = torch.zeros(64, dtype=torch.uint8, device="cuda")
        off = torch.zeros(9, dtype=torch.int64, device="cuda")
        assert A.streams_encode_static(ctx, db.data_ptr(), 8, 64, 16, 0, out.data_ptr(), 64, off.data_ptr()) == -1
        assert "coded_cap" in A.lib().avr_last_error().decode()

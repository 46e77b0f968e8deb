"""abub_gate_entropy_static (include/abub_cuda.h, csrc/gate_static.cu): the Trainer's entropy gate with its bin-count scratch
in a static __device__ array -- the entry point for callers that cannot call cudaMalloc.  No counterpart in the reference;
its arithmetic is that of abub_train_add_events, the existing entry point it
is compared with here: `kept` must be equal, the float entropies within 1e-6 (the tolerance every entropy comparison of
this suite uses: the 16 terms p*log2(p) are formed on the device here and with the host's log2f there), the 16-bin tables
exactly equal to numpy's.  Two calls in a row share the scratch array: the second call is smaller and quieter than the
first, so counts left behind by the first would show.

This file is named to be collected last."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pairs(rng, n, rows, cols, loud_every):
    """n frame pairs [n, 2, rows, cols]: quiet pairs (identical but for a few pixels), every loud_every-th one changed
    over a whole patch or everywhere."""
    base = rng.integers(10, 200, (n, rows, cols)).astype(np.int64)
    f0 = base.copy()
    f1 = base.copy()
    for e in range(n):
        kind = e % loud_every
        if kind == 0:
            f1[e] = rng.integers(0, 256, (rows, cols))                       # every bin populated
        elif kind == 1:
            f1[e, rows // 3:rows // 3 + 9, cols // 4:cols // 4 + 11] += 16 * (1 + e % 3) + 5    # a bubble-sized patch
        elif kind == 2:
            f1[e, e % rows, (7 * e) % cols] += 40                            # one pixel: stays under the gate
            f0[e, (3 * e) % rows, (5 * e) % cols] += 90                      # darker in frame 1: saturates to bin 0
        # else: identical frames, entropy 0
    return np.clip(np.stack([f0, f1], axis=1), 0, 255).astype(np.uint8)


def _hist16(pairs):
    d = np.clip(pairs[:, 1].astype(np.int16) - pairs[:, 0].astype(np.int16), 0, 255) >> 4
    return np.stack([np.bincount(x.ravel(), minlength=16) for x in d]).astype(np.uint32)


def _run_static(capi, torch, pairs, bufs=None):
    n, _, rows, cols = pairs.shape
    d_pairs = torch.from_numpy(pairs).cuda()
    if bufs is None:
        bufs = (torch.full((capi.ABUB_GATE_STATIC_MAX_PAIRS,), -1.0, dtype=torch.float32, device="cuda"),
                torch.full((capi.ABUB_GATE_STATIC_MAX_PAIRS,), -1, dtype=torch.int32, device="cuda"),
                torch.full((capi.ABUB_GATE_STATIC_MAX_PAIRS, 16), 7, dtype=torch.int32, device="cuda"))
    ent, kept, hist = bufs
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc, err = capi.gate_entropy_static(d_pairs.data_ptr(), n, rows, cols, ent.data_ptr(), kept.data_ptr(), hist.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream)
    assert rc == 0, err
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return (ent[:n].cpu().numpy(), kept[:n].cpu().numpy().astype(bool), hist[:n].cpu().numpy().astype(np.uint32),
            free0, free1, bufs)


@pytest.mark.parametrize("shape", [(264, 432), (37, 45)])
def test_static_scratch_gate_equals_train_add_events_twice_in_a_row(shape):
    import torch
    from autobub3hs_b200 import capi
    capi.load()
    rows, cols = shape
    rng = np.random.default_rng(rows * 1000 + cols)
    first = _pairs(rng, 12, rows, cols, loud_every=4)           # 3 loud, 3 patch, 3 single-pixel, 3 identical pairs
    second = _pairs(rng, 7, rows, cols, loud_every=5)           # fewer pairs, quieter: stale counts would show
    bufs = None
    for call, pairs in enumerate((first, second)):
        n = len(pairs)
        # the existing entry point (context arena, cudaMalloc'd scratch, host-side entropy)
        with capi.Context(rows, cols, 1, 4, max_slots=1) as ctx:
            ctx.train_begin(0)
            e_ref, k_ref = ctx.train_add_events(0, pairs)
        assert (np.abs(e_ref - 0.0005) > 1e-5).all()             # no pair sits on the gate within the float tolerance
        assert k_ref.any() and not k_ref.all()
        e, k, h, free0, free1, bufs = _run_static(capi, torch, pairs, bufs)
        assert np.array_equal(h, _hist16(pairs)), f"call {call}: bin counts"
        assert (h.sum(axis=1) == rows * cols).all()
        assert np.array_equal(k, k_ref), f"call {call}: kept"
        assert np.allclose(e, e_ref, rtol=0, atol=1e-6), (call, np.abs(e - e_ref).max())
        if call == 1:
            assert free1 == free0, "device memory in use changed across an allocation-free call"


def test_static_scratch_gate_rejects_more_pairs_than_the_scratch_holds():
    import torch
    from autobub3hs_b200 import capi
    capi.load()
    ent = torch.zeros(4, dtype=torch.float32, device="cuda")
    src = torch.zeros(2 * 16 * 16, dtype=torch.uint8, device="cuda")
    rc, err = capi.gate_entropy_static(src.data_ptr(), capi.ABUB_GATE_STATIC_MAX_PAIRS + 1, 16, 16, ent.data_ptr())
    assert rc == -104 and "too many frame pairs" in err and str(capi.ABUB_GATE_STATIC_MAX_PAIRS) in err   # ABUB_E_CAPACITY
    rc, err = capi.gate_entropy_static(src.data_ptr(), 1, 0, 16, ent.data_ptr())
    assert rc == -101 and err                                                                             # ABUB_E_INVALID
    rc, err = capi.gate_entropy_static(0, 1, 16, 16, ent.data_ptr())
    assert rc == -101 and "NULL" in err
    # the largest batch the scratch holds goes through
    n = capi.ABUB_GATE_STATIC_MAX_PAIRS
    big = torch.zeros((n, 2, 16, 16), dtype=torch.uint8, device="cuda")
    big[5, 1, 3, 3] = 200
    out = torch.full((n,), -1.0, dtype=torch.float32, device="cuda")
    kept = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    rc, err = capi.gate_entropy_static(big.data_ptr(), n, 16, 16, out.data_ptr(), kept.data_ptr())
    assert rc == 0, err
    torch.cuda.synchronize()
    o = out.cpu().numpy()
    assert (np.delete(o, 5) == 0).all() and o[5] > 0.0005
    assert kept.cpu().numpy().sum() == n - 1

"""vitk_hmm_decode_counts_static: the Viterbi decode that allocates nothing (scratch in a fixed device
array inside libvitk.so) against the existing vitk_hmm_decode_counts, through the C ABI.

Collected last on purpose.  Passes alone in a fresh process and
after the other GPU tests: it creates its own handle and owns every buffer it passes."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_FRAMES = 131072   # VITK_HMM_STATIC_MAX_FRAMES (include/vitk.h)
MAX_K = 128           # VITK_HMM_STATIC_MAX_K


@pytest.fixture(scope="module")
def v():
    import vit_research_b200 as v
    assert torch.cuda.is_available()
    v.load()
    return v


@pytest.fixture(scope="module")
def emb(v):
    # a handle for the EXISTING entry point (the new one needs none)
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
    e = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
    yield e
    e.close()


def _counts(n, k, seed):
    """Vote counts of n frames among k neighbours with runs, so the smoother has something to do."""
    rng = np.random.default_rng(seed)
    cls = np.repeat(rng.integers(0, 3, size=n // 37 + 1), 37)[:n]
    flip = rng.random(n) < 0.2
    cls = np.where(flip, rng.integers(0, 3, size=n), cls)
    strong = rng.integers(k // 2, k + 1, size=n)
    out = np.zeros((n, 3), dtype=np.int32)
    out[np.arange(n), cls] = strong
    rest = k - strong
    other = (cls + 1 + rng.integers(0, 2, size=n)) % 3
    part = (rng.random(n) * (rest + 1)).astype(np.int64)
    out[np.arange(n), other] += part
    out[np.arange(n), 3 - cls - other] += rest - part
    assert (out.sum(1) == k).all()
    return out


def _static(v, counts, k, states):
    from vit_research_b200._lib import ptr, stream_ptr
    lib = v.load()
    return lib.vitk_hmm_decode_counts_static(ptr(counts), int(counts.shape[0]), int(k), ptr(states), stream_ptr(None))


def test_static_scratch_decode_equals_the_allocating_entry_point_on_calls_in_a_row(v, emb):
    lib = v.load()
    # every buffer is allocated up front: nothing is created between the calls under test
    cases = [(5000, 25, 1), (1237, 5, 2), (MAX_FRAMES, 50, 3), (2, 25, 4), (1, 7, 5), (3, MAX_K, 6)]
    bufs = []
    for n, k, seed in cases:
        c = torch.from_numpy(_counts(n, k, seed)).cuda()
        bufs.append((c, k, torch.full((n,), -7, dtype=torch.int32, device="cuda"),
                     v.viterbi_from_counts(emb, c, k)))
    torch.cuda.synchronize()
    # first call on this device (resolves the array's address), then the rest back to back; the
    # shorter second call must not see what the longer first one left in the scratch
    assert _static(v, bufs[0][0], bufs[0][1], bufs[0][2]) == 0, lib.vitk_last_error(None)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    held0 = torch.cuda.memory_allocated()
    for c, k, st, _want in bufs[1:] + bufs[:2]:
        assert _static(v, c, k, st) == 0, lib.vitk_last_error(None)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == held0
    assert free1 >= free0, "device memory was allocated by the no-allocation entry point"
    for c, k, st, want in bufs:
        assert torch.equal(st, want), (c.shape[0], k)     # int32 states: exact, as in test_gpu_hmm.py
    assert (bufs[2][2][2:] >= 0).all() and bufs[2][2][0] == -1


def test_static_scratch_limits_are_errors_not_truncation(v):
    lib = v.load()
    c = torch.zeros((8, 3), dtype=torch.int32, device="cuda")
    st = torch.full((8,), -7, dtype=torch.int32, device="cuda")
    from vit_research_b200._lib import ptr, stream_ptr
    assert lib.vitk_hmm_decode_counts_static(ptr(c), MAX_FRAMES + 1, 25, ptr(st), stream_ptr(None)) != 0
    msg = lib.vitk_last_error(None)
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert str(MAX_FRAMES) in msg and "frames" in msg
    assert lib.vitk_hmm_decode_counts_static(ptr(c), 8, MAX_K + 1, ptr(st), stream_ptr(None)) != 0
    assert lib.vitk_hmm_decode_counts_static(ptr(c), 8, 0, ptr(st), stream_ptr(None)) != 0
    assert lib.vitk_hmm_decode_counts_static(None, 8, 25, ptr(st), stream_ptr(None)) != 0
    torch.cuda.synchronize()
    assert (st == -7).all()                                   # a rejected call touches nothing
    assert lib.vitk_hmm_decode_counts_static(ptr(c), 0, 25, ptr(st), stream_ptr(None)) == 0   # no frames: no work

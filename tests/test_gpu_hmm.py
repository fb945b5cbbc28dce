"""GPU parity of the vote -> HMM smoother (nba_proj/hmm.py:5-135) through the C ABI: decoded
states identical to the vectors produced by the reference's own class, lattice values bit-exact
(float64; logs come from the host's libm like CPython's math.log), batched videos independent,
and the full 108 000-frame stream (BASELINE configs[2]) against the oracle restatement."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def v():
    import vit_research_b200 as v
    assert torch.cuda.is_available()
    v.load()
    return v


@pytest.fixture(scope="module")
def emb(v):
    from oracle import vit_ref
    model = vit_ref.build_hf_model("vit_b16_l2", 0, randomise_ln=True)
    e = v.ViTEmbedder.from_hf(model, device=0, max_batch=8)
    yield e
    e.close()


@pytest.fixture(scope="module")
def golden(golden_dir):
    return np.load(os.path.join(golden_dir, "hmm_golden.npz"))


def test_counts_path_matches_reference_vectors(v, emb, golden):
    for c in [str(x) for x in golden["cases"] if str(x) + "_counts" in golden.files]:
        counts = torch.from_numpy(golden[c + "_counts"]).cuda()
        st, dp = v.viterbi_from_counts(emb, counts, int(golden[c + "_k"]), return_dp=True)
        assert np.array_equal(st.cpu().numpy(), golden[c + "_states"]), c
        assert np.array_equal(dp.cpu().numpy(), golden[c + "_dp"]), c


def test_probs_path_and_class_interface_match_reference_vectors(v, emb, golden):
    for c in ("float_a", "float_b"):
        p = golden[c + "_probs"]
        st, dp = v.viterbi_from_probs(emb, p, return_dp=True)
        assert np.array_equal(st, golden[c + "_states"]), c
        assert np.array_equal(dp, golden[c + "_dp"]), c
        h = v.hmm(emb, len(p) + 2)               # the reference's call sequence
        h.add_first({"left": p[0, 0], "right": p[0, 1], "none": p[0, 2]})
        for row in p[1:]:
            h.add_col_to_lattice({"left": row[0], "right": row[1], "none": row[2]})
        want = [v.hmm_module.DIRS[int(s)] for s in golden[c + "_states"]]
        assert h.decode_sequence() == want


def test_batched_videos_are_independent(v, emb, golden):
    names = ["k25_a", "t1", "k25_b", "t3", "t2"]
    parts = [golden[n + "_counts"] for n in names]
    off = np.concatenate([[0], np.cumsum([len(p) for p in parts])]).astype(np.int32)
    st = v.viterbi_from_counts(emb, torch.from_numpy(np.concatenate(parts)).cuda(), 25, seq_offsets=off)
    st = st.cpu().numpy()
    for i, n in enumerate(names):
        assert np.array_equal(st[off[i]:off[i + 1]], golden[n + "_states"]), n


def test_window_cap_raises_instead_of_blocking(v, emb):
    h = v.hmm(emb, cap_count=4)
    h.add_first({"left": 1, "right": 0, "none": 0})
    h.add_col_to_lattice({"left": 1, "right": 0, "none": 0})
    h.add_col_to_lattice({"left": 1, "right": 0, "none": 0})
    with pytest.raises(v.VitkError):
        h.add_col_to_lattice({"left": 1, "right": 0, "none": 0})


def test_hour_long_stream_matches_oracle(v, emb):
    """108 000 frames (1 h at 30 fps): labels equal the CPU restatement everywhere."""
    from oracle import hmm_ref
    rng = np.random.default_rng(8)
    T, k = 108_000, 25
    state = np.zeros(T, np.int64)
    s = 0
    u = rng.random(T)
    for t in range(T):
        if s == 2:
            s = 0 if u[t] < 0.1 else 1 if u[t] < 0.2 else 2
        elif u[t] < 0.01:
            s = 2
        state[t] = s
    p = np.full((T, 3), 0.12)
    p[np.arange(T), state] = 0.76
    counts = np.stack([rng.multinomial(k, row) for row in p[:2000]] +
                      [rng.multinomial(k, p[t]) for t in range(2000, T)]).astype(np.int32)
    want, want_dp = hmm_ref.viterbi(hmm_ref.probs_from_counts(counts, k))
    got, dp = v.viterbi_from_counts(emb, torch.from_numpy(counts).cuda(), k, return_dp=True)
    assert np.array_equal(got.cpu().numpy(), want)
    assert np.array_equal(dp.cpu().numpy(), want_dp)


def test_streamed_viterbi_equals_one_shot_in_any_segmentation(v, emb):
    """vitk_hmm_stream_*: a sequence decoded while its counts arrive in ragged pieces gives bit for
    bit the states of the one-shot kernel (itself pinned on the reference's own class) -- T = 1, 2,
    3 and a long stream cut at random points, including single-frame and empty pieces."""
    rng = np.random.default_rng(5)
    for T in (1, 2, 3, 17, 5000):
        side = np.repeat(rng.integers(0, 3, T // 40 + 1), 40)[:T]
        counts = np.zeros((T, 3), np.int32)
        for t in range(T):
            c = rng.multinomial(25, [0.8 if j == side[t] else 0.1 for j in range(3)])
            counts[t] = c
        dev = torch.from_numpy(counts).cuda()
        want = v.viterbi_from_counts(emb, dev, 25).cpu().numpy()
        for cuts in ([], [1], list(np.sort(rng.integers(0, T + 1, 9)))):
            vs = v.ViterbiStream(emb, T, 25)
            edges = [0] + [int(c) for c in cuts if 0 <= c <= T] + [T]
            for a, b in zip(edges[:-1], edges[1:]):
                vs.append(dev[a:b])
            got = vs.finish().cpu().numpy()
            assert np.array_equal(got, want), (T, cuts)
    with pytest.raises(v.VitkError):
        vs = v.ViterbiStream(emb, 4, 25)
        vs.append(torch.zeros((5, 3), dtype=torch.int32, device="cuda"))


def test_label_and_smooth_stream_equals_label_then_decode(v, emb, banks):
    """shard.label_and_smooth_stream (BASELINE configs[2], single process): chunks streamed through
    the overlapped host API with the Viterbi lattice advancing during the stream == embed + vote of
    all frames followed by one decode."""
    from oracle import vit_ref
    rows, labels = banks
    index = v.BankIndex(emb, rows, labels, 3, metric="l2")
    frames = vit_ref.synthetic_frames_u8(61, seed=4)              # max_batch 8: 7 full chunks + 5 frames
    states, lab = v.label_and_smooth_stream(emb, index, lambda a, b: frames[a:b], len(frames), k=25,
                                            chunk=8, gather_every=3)
    _e, want_lab, _i, _s, want_cnt = emb.embed_and_vote(frames[:8], index, 25)
    all_cnt, all_lab = [], []
    for a in range(0, 61, 8):
        _e, l, _i, _s, c = emb.embed_and_vote(frames[a:a + 8], index, 25)
        all_cnt.append(c.copy()); all_lab.append(l.copy())
    want_states = v.viterbi_from_counts(emb, torch.from_numpy(np.concatenate(all_cnt)).cuda(), 25)
    assert np.array_equal(lab, np.concatenate(all_lab))
    assert torch.equal(states, want_states)

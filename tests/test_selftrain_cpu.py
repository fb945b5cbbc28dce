"""CPU tests of the mutable-bank host logic: determine_class
(generate_clips_hmm.py:179-310) against vectors produced by the reference's own function, the
Chroma where-clause evaluator, and the oracle's sequential upsert loop."""
import json
import os

import numpy as np
import pytest

from oracle import selftrain_ref as ref


@pytest.fixture(scope="module")
def golden(golden_dir):
    return json.load(open(os.path.join(golden_dir, "determine_class_golden.json")))["cases"]


def _same(res, want):
    assert res[0] == want[0] and float(res[1]) == want[1] and len(res) == len(want)
    if len(want) == 3:
        assert {k: (float(v) if isinstance(v, (float, np.floating)) else v) for k, v in res[2].items()} == want[2]


def test_oracle_determine_class_matches_reference_vectors(golden):
    cases = ref.golden_cases()
    assert len(cases) == len(golden) and sum(len(g["result"]) == 3 for g in golden) >= 16
    for case, want in zip(cases, golden):
        lat = []
        res = ref.determine_class_ref(case["ids"], case["metadatas"], case["distances"], case["add_first"],
                                      case["top_n_closest"], case["cur_vid"], lat)
        _same(res, want["result"])
        assert lat[0][0] == want["lattice"]
        assert {k: float(v) for k, v in lat[0][1].items()} == want["col"]


class _Lattice:
    def __init__(self):
        self.cols = []

    def add_first(self, c):
        self.cols.append(("first", c))

    def add_col_to_lattice(self, c):
        self.cols.append(("col", c))


def test_product_determine_class_matches_reference_vectors(golden):
    from vit_research_b200.knn import determine_class
    for case, want in zip(ref.golden_cases(), golden):
        lat = _Lattice()
        res = determine_class(case["ids"], case["metadatas"], case["distances"], case["add_first"],
                              top_n_closest=case["top_n_closest"], cur_vid=case["cur_vid"], hmm_matrix=lat)
        _same(res, want["result"])
        assert lat.cols[0][0] == want["lattice"]
        assert {k: float(v) for k, v in lat.cols[0][1].items()} == want["col"]


def test_where_clauses():
    from vit_research_b200.knn import VitkError, match_where
    m = {"vid_num": "vid3", "side": "left", "t_norm": 0.42, "t_center": 10.0}
    assert match_where(m, None) and match_where(m, {})
    assert match_where(m, {"side": "left"}) and not match_where(m, {"side": "right"})
    assert match_where(m, {"vid_num": {"$ne": "vid0"}}) and not match_where(m, {"vid_num": {"$ne": "vid3"}})
    # retrieval/frame_retriever.py:44-51
    w = {"$and": [{"vid_num": {"$ne": "vid1"}}, {"side": "left"}, {"t_norm": {"$gte": 0.4}},
                  {"t_norm": {"$lte": 0.5}}]}
    assert match_where(m, w)
    assert not match_where(dict(m, t_norm=0.51), w) and not match_where(dict(m, vid_num="vid1"), w)
    assert match_where(m, {"$or": [{"side": "right"}, {"t_center": {"$gt": 9.0}}]})
    assert match_where(m, {"side": {"$in": ["left", "none"]}}) and not match_where(m, {"side": {"$nin": ["left"]}})
    assert not match_where(m, {"missing": "x"}) and not match_where({}, {"side": {"$ne": "left"}})
    with pytest.raises(VitkError):
        match_where(m, {"side": {"$regex": "l.*"}})


def _clusters(rng, n_per, d=64, spread=0.05):
    centres = rng.standard_normal((3, d)) * 2.0
    rows = np.concatenate([centres[c] + spread * rng.standard_normal((n_per, d)) for c in range(3)])
    labels = np.repeat(np.arange(3), n_per)
    return centres, rows.astype(np.float32), labels


def test_sequential_loop_upserts_every_16_confident_frames():
    rng = np.random.default_rng(3)
    centres, rows, labels = _clusters(rng, 40)
    metas = [{"label": ref.NAMES[c], "left_prob": float(c == 0), "right_prob": float(c == 1),
              "none_prob": float(c == 2)} for c in labels]
    col = ref.NumpyCollection(rows, ["b%d" % i for i in range(len(rows))], metas)
    seq = rng.integers(0, 3, 100)
    frames = (centres[seq] + 0.05 * rng.standard_normal((100, 64))).astype(np.float32)
    out = ref.sequential_loop(col.query, col.upsert, frames, ["f%d" % i for i in range(100)], 25, "vidX")
    assert [s[0] for s in out["sides"]] == [ref.NAMES[c] for c in seq]
    assert all(len(s) == 3 and s[1] == 0.999998 for s in out["sides"])      # unanimous neighbourhoods
    assert out["upserts"] == 6 and out["left_over"] == 4 and len(col.ids) == 120 + 96
    assert out["lattice"][0][0] == "first" and all(k == "col" for k, _ in out["lattice"][1:])
    assert col.metadatas[-1]["video"] == "vidX" and col.ids[120] == "f0"

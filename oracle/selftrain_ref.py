"""TEST INFRASTRUCTURE ONLY — CPU restatement of the reference's mutable-bank classification loop.

Follows nba_proj/generate_clips_hmm.py:
  determine_class            :179-310  (counts, mean stored probabilities, (mean+fraction)/2 argmax,
                                         0.7 confidence rule, the all-agree 0.999998 metadata)
  the per-frame loop         :384-470  (query top_n_closest, determine_class, collect confident
                                         frames, upsert every 16)
and chromadb's 'l2' space (squared L2, ascending) as exact brute force in float64.
Pinned by tests/golden/determine_class_golden.json, which tools/make_golden_determine_class.py
produces by executing the reference's own determine_class.  Only tests/ may import this module.
"""
import numpy as np

NAMES = ("left", "right", "none")


def determine_class_ref(ids, metadatas, distances, add_first, top_n_closest, cur_vid, lattice=None):
    """generate_clips_hmm.py:179-310.  Returns what the reference returns; `lattice` receives the
    column the reference hands to hmm_matrix (:242-245)."""
    k = len(ids)
    counts = {"left": 0, "right": 0, "none": 0}
    for md in metadatas:                                   # :191-197 (anything else counts as none)
        counts[md["label"] if md["label"] in ("left", "right") else "none"] += 1
    frac = {s: counts[s] / k for s in NAMES}               # :202-204
    mean = {s: np.mean(np.array([md[s + "_prob"] for md in metadatas])) for s in NAMES}  # :213-223
    if lattice is not None:                                # :242-245
        lattice.append(("first" if add_first else "col", dict(mean)))
    blend = [(mean[s] + frac[s]) / 2 for s in NAMES]       # :238
    win = NAMES[int(np.array(blend).argmax())]             # :246
    if mean[win] < 0.7:                                    # :262-263 and twins
        return [win, mean[win]]
    if counts[win] == top_n_closest:                       # :268-273
        md = {"label": win, "video": cur_vid}
        for s in NAMES:
            md[s + "_prob"] = 0.999998 if s == win else 0.000001
        return [win, 0.999998, md]
    md = {"label": win, "video": cur_vid}                  # :275-279
    for s in NAMES:
        md[s + "_prob"] = mean[s]
    return [win, mean[win], md]


class NumpyCollection:
    """Exact squared-L2 brute force with Chroma's query/upsert surface (float64, stable order)."""

    def __init__(self, rows, ids, metadatas):
        self.rows = [np.asarray(r, np.float64) for r in rows]
        self.ids = list(ids)
        self.metadatas = [dict(m) for m in metadatas]

    def query(self, q, n_results):
        b = np.stack(self.rows)
        d = ((b - np.asarray(q, np.float64)[None, :]) ** 2).sum(1)
        order = np.argsort(d, kind="stable")[:n_results]
        return ([self.ids[j] for j in order], [self.metadatas[j] for j in order], [float(d[j]) for j in order])

    def upsert(self, embeddings, ids, metadatas):
        for e, i, m in zip(embeddings, ids, metadatas):
            if i in self.ids:
                r = self.ids.index(i)
                self.rows[r], self.metadatas[r] = np.asarray(e, np.float64), dict(m)
            else:
                self.rows.append(np.asarray(e, np.float64))
                self.ids.append(i)
                self.metadatas.append(dict(m))


def sequential_loop(query_fn, upsert_fn, embeddings, names, top_n_closest, cur_vid, upsert_batch=16):
    """generate_clips_hmm.py:384-470, one frame at a time.  query_fn(q, k) -> (ids, metadatas,
    distances); upsert_fn(embeddings, ids, metadatas)."""
    sides, lattice = [], []
    rows, ids, metas, upserts = [], [], [], 0
    for i, (e, name) in enumerate(zip(embeddings, names)):
        nid, nmeta, ndist = query_fn(e, top_n_closest)
        side = determine_class_ref(nid, nmeta, ndist, i == 0, top_n_closest, cur_vid, lattice)
        sides.append(side)
        if len(side) == 3:                                 # :411-449
            rows.append(e); ids.append(name); metas.append(side[2])
        if len(rows) == upsert_batch:                      # :456-470
            upsert_fn(np.stack(rows), ids, metas)
            rows, ids, metas = [], [], []
            upserts += 1
    return {"sides": sides, "lattice": lattice, "upserts": upserts, "left_over": len(rows)}


def golden_cases(seed=20240611, n=64):
    """Seeded inputs for the determine_class golden file (regenerated identically in the tests)."""
    rng = np.random.default_rng(seed)
    cases = []
    for c in range(n):
        k = int(rng.choice([5, 25, 50]))
        mode = c % 4
        metas = []
        for j in range(k):
            if mode == 0:                                  # unanimous, one-hot
                lab = NAMES[c // 4 % 3]
            elif mode == 1:                                # strong majority
                lab = NAMES[c // 4 % 3] if rng.random() < 0.85 else NAMES[int(rng.integers(3))]
            else:
                lab = NAMES[int(rng.integers(3))]
            if mode == 3 or rng.random() < 0.3:            # a previously upserted row: soft probs
                p = rng.dirichlet([6.0 if s == lab else 1.0 for s in NAMES])
                p = [float(x) for x in p]
            else:
                p = [1.0 if s == lab else 0.0 for s in NAMES]
            metas.append({"label": lab, "video": "vid%d" % (c % 5), "left_prob": p[0], "right_prob": p[1],
                          "none_prob": p[2]})
        cases.append({"ids": ["f%d_%d" % (c, j) for j in range(k)], "metadatas": metas,
                      "distances": [float(x) for x in np.sort(rng.random(k) * 100)],
                      "add_first": c == 0, "top_n_closest": k if c % 3 else 50, "cur_vid": "vid%d" % (c % 5)})
    return cases

"""Golden vectors for determine_class (generate_clips_hmm.py:179-310), produced by the reference's
OWN function: the script it lives in cannot be imported (TensorFlow / cv2 / chromadb at module
level), so the FunctionDef is cut out of the file with `ast`, compiled as it stands and run with
the three module globals it reads (top_n_closest, cur_vid, hmm_matrix) supplied here.  Nothing of
the reference is copied into the repo: only the inputs' seed and the outputs are stored.

Run in the build container (needs /root/reference):  python tools/make_golden_determine_class.py
"""
import ast
import contextlib
import io
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.selftrain_ref import golden_cases  # noqa: E402

SRC = "/root/reference/nba_proj/generate_clips_hmm.py"


class _Lattice:
    def __init__(self):
        self.cols = []

    def add_first(self, col):
        self.cols.append(("first", dict(col)))

    def add_col_to_lattice(self, col):
        self.cols.append(("col", dict(col)))


def main():
    tree = ast.parse(open(SRC).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "determine_class")
    mod = ast.Module(body=[fn], type_ignores=[])
    g = {"np": np}
    exec(compile(mod, SRC, "exec"), g)
    out = []
    for case in golden_cases():
        lat = _Lattice()
        g.update(top_n_closest=case["top_n_closest"], cur_vid=case["cur_vid"], hmm_matrix=lat)
        with contextlib.redirect_stdout(io.StringIO()):
            res = g["determine_class"](case["ids"], case["metadatas"], case["distances"], case["add_first"])
        res = [res[0], float(res[1])] + ([{k: (float(v) if isinstance(v, (float, np.floating)) else v)
                                           for k, v in res[2].items()}] if len(res) == 3 else [])
        kind, col = lat.cols[0]
        out.append({"result": res, "lattice": kind, "col": {k: float(v) for k, v in col.items()}})
    path = os.path.join(ROOT, "tests", "golden", "determine_class_golden.json")
    json.dump({"source": "generate_clips_hmm.py:179-310 executed as written", "cases": out}, open(path, "w"))
    print("wrote", path, len(out), "cases;", sum(len(o["result"]) == 3 for o in out), "confident")


def main_pass12():
    """determine_class_pass1 / determine_class_pass2 (chroma.py:36-71,102-134) cut out of the
    reference script the same way and run on seeded neighbour lists (k = 25 like chroma.py:27 and
    a few other k): the dir and the temp_smax probabilities they return."""
    src = "/root/reference/nba_proj/chroma.py"
    tree = ast.parse(open(src).read())
    want = {"softmax", "temp_smax", "determine_class_pass1", "determine_class_pass2"}
    fns = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in want]
    g = {"np": np}
    exec(compile(ast.Module(body=fns, type_ignores=[]), src, "exec"), g)
    rng = np.random.default_rng(2025)
    cases = []
    for c in range(96):
        k = [25, 25, 25, 5, 50, 30][c % 6]
        p = rng.dirichlet([0.4, 0.4, 0.25])
        if c % 5 == 0:
            p = np.array([0.5, 0.5, 0.0]) if c % 10 == 0 else np.array([0.0, 0.5, 0.5])   # ties
        labs = [["left", "right", "none", "weird"][min(int(x), 3)] for x in rng.choice(3, k, p=p)]
        if c % 7 == 0:
            labs[0] = "weird"          # any other label counts as none
        metas = [{"label": lab} for lab in labs]
        ids = [f"id{j}" for j in range(k)]
        dist = [float(x) for x in np.sort(rng.random(k) * 100)]
        with contextlib.redirect_stdout(io.StringIO()):
            r1 = g["determine_class_pass1"](ids, metas, dist)
            r2 = g["determine_class_pass2"](ids, metas, dist)
        enc = lambda r: {"dir": r["dir"], **({"probs": [float(x) for x in r["probs"]]} if "probs" in r else {})}
        cases.append({"labels": labs, "distances": dist, "pass1": enc(r1), "pass2": enc(r2)})
    path = os.path.join(ROOT, "tests", "golden", "determine_class_pass12_golden.json")
    json.dump({"source": "chroma.py:28-71,102-134 executed as written", "cases": cases}, open(path, "w"))
    print("wrote", path, len(cases), "cases;", sum(c["pass1"]["dir"] == "pass2" for c in cases), "deferred by pass 1")


if __name__ == "__main__":
    main()
    main_pass12()

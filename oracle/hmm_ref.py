"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatement of the reference's left/right/none Viterbi smoother, nba_proj/hmm.py:5-135
(class hmm), as driven by nba_proj/generate_clips_hmm.py:238-245,484-490.  Plain Python float64
loops, the same operation order, the same `round(x, 3)` per step, the same tie rules and the same
backtrace as written (it starts from column count-1 and reads backpointers[cur][i-1]).

Pinned: tools/make_golden_hmm.py imports the reference's own hmm.py in the build container, feeds
it seeded probability columns and stores inputs + outputs in tests/golden/hmm_golden.npz;
tests/test_oracle_cpu.py checks this restatement against those vectors.
"""
from __future__ import annotations

import math

import numpy as np

STATES = ("left", "right", "none")
_T = [[0.985, 0, 0.015], [0, 0.985, 0.015], [0.15, 0.15, 0.7]]  # hmm.py:26


def viterbi(probs):
    """probs: (T,3) float64 [left,right,none] per frame (the dicts given to add_first /
    add_col_to_lattice).  Returns (states (T,) int32 in {0,1,2,-1}, dp (T,3) float64) where dp[t]
    is lattice column t+1 (hmm.py keeps index 0 unused)."""
    probs = np.asarray(probs, np.float64)
    T = probs.shape[0]
    dp = np.zeros((T, 3), np.float64)
    bp = np.full((T, 3), -1, np.int64)
    states = np.full((T,), -1, np.int32)
    if T == 0:
        return states, dp
    for s in range(3):  # add_first, hmm.py:36-38
        dp[0, s] = math.log(1 / 3) + math.log(float(probs[0, s]) + 0.000001)
    for t in range(1, T):  # add_col_to_lattice, hmm.py:41-106
        l, r, n = (float(x) if x != 0 else 0.000001 for x in probs[t])
        pl, pr, pn = float(dp[t - 1, 0]), float(dp[t - 1, 1]), float(dp[t - 1, 2])
        a = pl + math.log(_T[0][0]) + math.log(l)
        b = pn + math.log(_T[2][0]) + math.log(l)
        bp[t, 0] = 0 if a > b else 2
        dp[t, 0] = round(max(a, b), 3)
        a = pr + math.log(_T[1][1]) + math.log(r)
        b = pn + math.log(_T[2][1]) + math.log(r)
        bp[t, 1] = 1 if a > b else 2
        dp[t, 1] = round(max(a, b), 3)
        c = [pl + math.log(_T[0][2]) + math.log(n), pr + math.log(_T[1][2]) + math.log(n),
             pn + math.log(_T[2][2]) + math.log(n)]
        bp[t, 2] = int(np.argmax(c))
        dp[t, 2] = round(max(c), 3)
    # decode_sequence, hmm.py:108-135, count == T; lattice column c == dp[c-1], bp[c-1]
    count = T
    col = count - 1                       # dp[.][count-1]
    cur = int(np.argmax([dp[col - 1, 0], dp[col - 1, 1], dp[col - 1, 2]])) if col >= 1 else 0
    for i in range(count - 1, 1, -1):
        states[i] = cur
        cur = int(bp[i - 2, cur]) if i - 1 >= 2 else -1   # backpointers[cur][i-1]
    return states, dp


def probs_from_counts(counts, k):
    """generate_clips_hmm.py:188-224 on static one-hot banks: lmean = count/k."""
    return np.asarray(counts, np.float64) / float(k)

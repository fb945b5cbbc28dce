"""Vote -> HMM smoothing with the reference's call shapes, backed by libvitk (sm_100a).

Mirrors nba_proj/hmm.py:5-135 (class hmm: add_first / add_col_to_lattice / decode_sequence) as
driven by nba_proj/generate_clips_hmm.py:238-245,484-490.  Columns are buffered on the host and the
whole lattice is decoded by one CUDA kernel (hmm.cuh) that reproduces the reference's float64
arithmetic, per-step round(x, 3), tie rules and backtrace, so the decoded labels are identical.
`viterbi_from_counts` is the device-resident form: it consumes the (T,3) vote counts that
BankIndex.search() left on the GPU, for any number of videos at once.  No CPU fallback.
"""
from __future__ import annotations

import numpy as np

from . import _lib
from ._lib import VitkError, check, ptr, stream_ptr

DIRS = {-1: -1, 0: "left", 1: "right", 2: "none"}  # hmm.py:128


def _offsets(seq_offsets, total):
    off = np.ascontiguousarray([0, total] if seq_offsets is None else seq_offsets, dtype=np.int32)
    if off.ndim != 1 or off.size < 2 or off[0] != 0 or off[-1] != total:
        raise VitkError("seq_offsets must run from 0 to the number of frames")
    return off


def viterbi_from_counts(embedder_or_handle, counts, k, seq_offsets=None, return_dp=False, stream=None):
    """counts: CUDA int32 (T,3) [left,right,none] votes among k neighbours (BankIndex.search()'s
    4th output).  Returns CUDA int32 (T,) states in {0,1,2,-1} (and the (T,3) float64 lattice)."""
    import torch

    lib = _lib.load()
    handle = getattr(embedder_or_handle, "handle", embedder_or_handle)
    if not counts.is_cuda or counts.dtype != torch.int32 or counts.dim() != 2 or counts.shape[1] != 3:
        raise VitkError("viterbi_from_counts() wants a CUDA int32 (T,3) tensor")
    counts = counts.contiguous()
    T = counts.shape[0]
    off = _offsets(seq_offsets, T)
    states = torch.full((T,), -1, dtype=torch.int32, device=counts.device)
    dp = torch.zeros((T, 3), dtype=torch.float64, device=counts.device) if return_dp else None
    check(lib.vitk_hmm_decode_counts(handle, ptr(counts), int(k), ptr(off), off.size - 1, ptr(states),
                                     ptr(dp), stream_ptr(stream)), handle, "vitk_hmm_decode_counts")
    return (states, dp) if return_dp else states


class ViterbiStream:
    """One long sequence decoded while its vote counts still arrive (vitk_hmm_stream_*): the
    lattice of hmm.py:5-135 advances segment by segment, so a game sharded over several GPUs is
    smoothed DURING the stream instead of after it.  Any segmentation gives bit for bit the
    states of viterbi_from_counts on the whole sequence."""

    def __init__(self, embedder_or_handle, capacity_frames, k):
        self.lib = _lib.load()
        self.handle = getattr(embedder_or_handle, "handle", embedder_or_handle)
        self.k = int(k)
        self.frames = 0
        check(self.lib.vitk_hmm_stream_begin(self.handle, int(capacity_frames)), self.handle,
              "vitk_hmm_stream_begin")

    def append(self, counts, stream=None):
        """counts: CUDA int32 (n,3) votes of the NEXT n frames in time order."""
        import torch
        if not counts.is_cuda or counts.dtype != torch.int32 or counts.dim() != 2 or counts.shape[1] != 3:
            raise VitkError("ViterbiStream.append() wants a CUDA int32 (n,3) tensor")
        counts = counts.contiguous()
        check(self.lib.vitk_hmm_stream_append_counts(self.handle, ptr(counts), counts.shape[0], self.k,
                                                     stream_ptr(stream)), self.handle,
              "vitk_hmm_stream_append_counts")
        self.frames += counts.shape[0]
        self._keep = counts          # alive until the next call on the same stream

    def finish(self, device=None, stream=None):
        """States of every frame appended, CUDA int32 (T,) in {0,1,2,-1}."""
        import ctypes as C
        import torch
        states = torch.full((self.frames,), -1, dtype=torch.int32, device=device or "cuda")
        n = C.c_longlong(0)
        check(self.lib.vitk_hmm_stream_finish(self.handle, ptr(states) if self.frames else None, C.byref(n),
                                              stream_ptr(stream)), self.handle, "vitk_hmm_stream_finish")
        return states


def viterbi_from_probs(embedder_or_handle, probs, seq_offsets=None, return_dp=False):
    """probs: host (T,3) float64 [left,right,none].  Returns numpy int32 (T,) states (and dp)."""
    lib = _lib.load()
    handle = getattr(embedder_or_handle, "handle", embedder_or_handle)
    probs = np.ascontiguousarray(probs, dtype=np.float64).reshape(-1, 3)
    T = probs.shape[0]
    off = _offsets(seq_offsets, T)
    states = np.full((T,), -1, np.int32)
    dp = np.zeros((T, 3), np.float64) if return_dp else None
    if T:
        check(lib.vitk_hmm_decode_probs_host(handle, ptr(probs), ptr(off), off.size - 1, ptr(states),
                                             ptr(dp)), handle, "vitk_hmm_decode_probs_host")
    return (states, dp) if return_dp else states


class hmm:
    """Drop-in for the reference's `hmm.hmm(cap_count)` (hmm.py:5-135)."""

    def __init__(self, embedder_or_handle, cap_count=501):
        self._owner = embedder_or_handle
        self.cap_count = cap_count
        self.count = 1
        self._cols = []
        self.decoded_sequence = []

    @staticmethod
    def _row(col):
        return [float(col["left"]), float(col["right"]), float(col["none"])]

    def add_first(self, first):
        self._cols = [self._row(first)]
        self.count = 1

    def add_col_to_lattice(self, col):
        if self.count >= self.cap_count - 1:
            # the reference blocks on input('stop') here (hmm.py:52-58); a library raises instead
            raise VitkError("window is maxed out: need to call the shift method first")
        self.count += 1
        self._cols.append(self._row(col))

    def decode_sequence(self):
        states = viterbi_from_probs(self._owner, np.array(self._cols, np.float64)) if self._cols else []
        self.decoded_sequence = [DIRS[int(s)] for s in states]
        return self.decoded_sequence

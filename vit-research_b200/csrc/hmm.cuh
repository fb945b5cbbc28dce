// hmm.cuh — left/right/none Viterbi smoothing of the per-frame vote (the step right after the
// hot path).  Reference: nba_proj/hmm.py:5-135 (class hmm: add_first, add_col_to_lattice,
// decode_sequence), driven by nba_proj/generate_clips_hmm.py:238-245,484-490.
//
// The recursion is sequential in time, so one WARP walks one video: lanes 0..2 own the three
// states, exchange the previous column by shuffles, and every step reproduces the reference's
// float64 arithmetic — including its per-step Python round(x, 3), its tie rules and the
// off-by-one of its backtrace — so decoded labels are identical, not merely close.  Videos
// (sequences) are independent: one warp each.  Logarithms come from a host-built table
// (std::log == CPython's math.log on glibc), never from the device's log().
#pragma once
#include "vitk_common.cuh"

namespace vitk {

struct HmmParams {
  // log of the reference's transition matrix entries that are ever used (hmm.py:26)
  double log_stay;        // log(0.985)  left->left, right->right
  double log_to_none;     // log(0.015)  left->none, right->none
  double log_from_none;   // log(0.15)   none->left, none->right
  double log_none_none;   // log(0.7)
  double log_third;       // log(1/3)
};

// Python's round(x, 3) for |x| < 2^40: nearest multiple of 1/1000 to the EXACT value of x (ties
// to even), returned as the nearest double — what CPython's dtoa-based float.__round__ yields.
__device__ __forceinline__ double py_round3(double x) {
  const double p = x * 1000.0;
  const double e = fma(x, 1000.0, -p);   // exact: x*1000 == p + e
  double n = rint(p);                    // ties to even
  const double diff = p - n;             // exact
  if (diff == 0.5 && e > 0.0) n += 1.0;
  else if (diff == -0.5 && e < 0.0) n -= 1.0;
  return n / 1000.0;
}

// loge: [T,3] log emission per frame, already computed by the rule of the column it feeds
// (first column: log(p + 1e-6), hmm.py:36-38; later columns: log(p == 0 ? 1e-6 : p),
// hmm.py:45-50,63-83).  offsets: [n_seq + 1] frame ranges.  bp: [T,3] scratch.
// states_out[t] in {0 left, 1 right, 2 none, -1 undecoded}: per sequence, entries 0 and 1 are
// -1 exactly as decode_sequence leaves them (generate_clips_hmm.py:485-486 patches them).
// dp_out (optional) [T,3]: the rounded lattice, for the parity tests.
__global__ void __launch_bounds__(32)
hmm_viterbi_kernel(const double* __restrict__ loge, const int* __restrict__ offsets, int n_seq,
                   HmmParams hp, int8_t* __restrict__ bp, int32_t* __restrict__ states_out,
                   double* __restrict__ dp_out) {
  const int seq = blockIdx.x;
  const int lane = threadIdx.x;
  if (seq >= n_seq) return;
  const int f0 = offsets[seq], T = offsets[seq + 1] - f0;
  if (T <= 0) return;
  const int s = lane < 3 ? lane : 2;  // lanes >= 3 shadow state 2 (keeps shuffles full-warp)
  const double* le = loge + (size_t)f0 * 3;
  // lattice column c (1-based, hmm.py keeps index 0 unused) holds frame c-1
  double dp = hp.log_third + le[s];                 // add_first: NOT rounded (hmm.py:36-38)
  if (dp_out && lane < 3) dp_out[(size_t)f0 * 3 + s] = dp;
  // np.argmax over the three states of the current column (first maximum wins)
  auto col_argmax = [&]() {
    const double a = __shfl_sync(0xffffffffu, dp, 0), b = __shfl_sync(0xffffffffu, dp, 1),
                 d = __shfl_sync(0xffffffffu, dp, 2);
    return (a >= b && a >= d) ? 0 : (b >= d ? 1 : 2);
  };
  int start = T - 1 == 1 ? col_argmax() : 0;        // decode starts at column count-1 == T-1
  for (int c = 2; c <= T; ++c) {
    const double pl = __shfl_sync(0xffffffffu, dp, 0);
    const double pr = __shfl_sync(0xffffffffu, dp, 1);
    const double pn = __shfl_sync(0xffffffffu, dp, 2);
    const double em = le[(size_t)(c - 1) * 3 + s];
    double best;
    int from;
    if (s == 0) {            // hmm.py:63-72: ties -> from none
      const double a = (pl + hp.log_stay) + em, b = (pn + hp.log_from_none) + em;
      best = fmax(a, b);
      from = a > b ? 0 : 2;
    } else if (s == 1) {     // hmm.py:74-83
      const double a = (pr + hp.log_stay) + em, b = (pn + hp.log_from_none) + em;
      best = fmax(a, b);
      from = a > b ? 1 : 2;
    } else {                 // hmm.py:85-92: np.argmax, first maximum wins
      const double a = (pl + hp.log_to_none) + em, b = (pr + hp.log_to_none) + em,
                   d = (pn + hp.log_none_none) + em;
      best = fmax(a, fmax(b, d));
      from = (a >= b && a >= d) ? 0 : (b >= d ? 1 : 2);
    }
    dp = py_round3(best);    // hmm.py:104-106
    if (lane < 3) {
      bp[((size_t)f0 + (c - 1)) * 3 + s] = (int8_t)from;  // backpointers[s][c]
      if (dp_out) dp_out[((size_t)f0 + (c - 1)) * 3 + s] = dp;
    }
    if (c == T - 1) start = col_argmax();
  }
  __syncwarp();
  // decode_sequence (hmm.py:108-135) with count == T: it starts from column T-1 (not T) and
  // steps with backpointers[cur][i-1] (not [i]) — both reproduced as written; entries 0 and 1
  // stay -1.
  if (lane == 0) {
    int32_t* out = states_out + f0;
    out[0] = -1;
    if (T > 1) out[1] = -1;
    int cur = start;
    for (int i = T - 1; i >= 2; --i) {
      out[i] = cur;
      // backpointers[cur][i-1]; column 1 was never written by add_col_to_lattice (-1 there,
      // only read after the last assignment)
      cur = (i - 1 >= 2 && cur >= 0) ? bp[((size_t)f0 + (i - 2)) * 3 + cur] : -1;
    }
  }
}

// The same recursion for a sequence that ARRIVES IN PIECES (a stream sharded over several GPUs
// delivers its vote counts segment by segment, in time order): one launch per segment continues
// from the lattice column the previous launch left in `carry` -- carry[0..2] = the last column,
// carry[3..5] = the one before it (decode_sequence starts from column T-1, known only at the
// end).  Frame t is lattice column t+1; `t0` frames were consumed before this segment.  The
// arithmetic per step is the code above, so a stream decoded in any segmentation is bit-identical
// to the one-shot kernel.
__global__ void __launch_bounds__(32)
hmm_viterbi_segment_kernel(const double* __restrict__ loge, long long t0, int n, HmmParams hp,
                           double* __restrict__ carry, int8_t* __restrict__ bp) {
  const int lane = threadIdx.x;
  const int s = lane < 3 ? lane : 2;
  if (n <= 0) return;
  double dp, prev;
  int i = 0;
  if (t0 == 0) {
    dp = hp.log_third + loge[s];        // add_first: not rounded
    prev = dp;
    i = 1;
  } else {
    dp = carry[s];
    prev = carry[3 + s];
  }
  for (; i < n; ++i) {
    const double pl = __shfl_sync(0xffffffffu, dp, 0);
    const double pr = __shfl_sync(0xffffffffu, dp, 1);
    const double pn = __shfl_sync(0xffffffffu, dp, 2);
    const double em = loge[(size_t)i * 3 + s];
    double best;
    int from;
    if (s == 0) {
      const double a = (pl + hp.log_stay) + em, b = (pn + hp.log_from_none) + em;
      best = fmax(a, b);
      from = a > b ? 0 : 2;
    } else if (s == 1) {
      const double a = (pr + hp.log_stay) + em, b = (pn + hp.log_from_none) + em;
      best = fmax(a, b);
      from = a > b ? 1 : 2;
    } else {
      const double a = (pl + hp.log_to_none) + em, b = (pr + hp.log_to_none) + em,
                   d = (pn + hp.log_none_none) + em;
      best = fmax(a, fmax(b, d));
      from = (a >= b && a >= d) ? 0 : (b >= d ? 1 : 2);
    }
    prev = dp;
    dp = py_round3(best);
    if (lane < 3) bp[((size_t)t0 + i) * 3 + s] = (int8_t)from;
  }
  if (lane < 3) {
    carry[s] = dp;
    carry[3 + s] = prev;
  }
}

// decode_sequence of a streamed lattice of T frames (see hmm_viterbi_kernel for the reference's
// conventions: start at column T-1, backpointers read one column off, entries 0 and 1 stay -1).
__global__ void __launch_bounds__(32)
hmm_backtrace_kernel(const int8_t* __restrict__ bp, long long T, const double* __restrict__ carry,
                     int32_t* __restrict__ out) {
  if (threadIdx.x != 0 || T <= 0) return;
  out[0] = -1;
  if (T > 1) out[1] = -1;
  int cur = 0;
  if (T >= 2) {   // np.argmax over column T-1 (first maximum wins); for T == 2 that is the first column
    const double a = carry[3], b = carry[4], d = carry[5];
    cur = (a >= b && a >= d) ? 0 : (b >= d ? 1 : 2);
  }
  for (long long i = T - 1; i >= 2; --i) {
    out[i] = cur;
    cur = (i - 1 >= 2 && cur >= 0) ? bp[(size_t)(i - 2) * 3 + cur] : -1;
  }
}

// Emissions of one streamed segment: frame t0 + i; only global frame 0 takes the first-column rule.
__global__ void __launch_bounds__(256)
hmm_emission_segment_kernel(const int32_t* __restrict__ counts, long long t0, int n, int k,
                            const double* __restrict__ table, double* __restrict__ loge) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)n * 3;
       i += (long long)gridDim.x * blockDim.x) {
    int c = counts[i];
    c = c < 0 ? 0 : (c > k ? k : c);
    const bool first = t0 == 0 && i < 3;
    loge[i] = table[(first ? 0 : k + 1) + c];
  }
}

// counts [T,3] (left,right,none votes among k neighbours) -> log emissions.  With static one-hot
// banks the reference's lmean/rmean/nmean are count/k (generate_clips_hmm.py:188-224), so the
// emission takes one of k+1 values: table[0][c] = log(c/k + 1e-6) feeds a sequence's first
// column, table[1][c] = log(c == 0 ? 1e-6 : c/k) every later one.  One block per sequence.
__global__ void __launch_bounds__(256)
hmm_emission_from_counts_kernel(const int32_t* __restrict__ counts, const int* __restrict__ offsets,
                                int n_seq, int k, const double* __restrict__ table,
                                double* __restrict__ loge) {
  const int seq = blockIdx.x;
  if (seq >= n_seq) return;
  const int f0 = offsets[seq], f1 = offsets[seq + 1];
  for (long long i = (long long)f0 * 3 + threadIdx.x; i < (long long)f1 * 3; i += blockDim.x) {
    int c = counts[i];
    c = c < 0 ? 0 : (c > k ? k : c);
    const bool first = i < (long long)f0 * 3 + 3;
    loge[i] = table[(first ? 0 : k + 1) + c];
  }
}

}  // namespace vitk

// Candidate draws on the device, bit-identical to the reference's generator, for any block of rows.
//
// Reference: x_tries = random_generator.uniform(bounds[:, 0], bounds[:, 1], size=(n_warmup, d))
// (acquisition_function.py:88-89) on numpy.random.RandomState, i.e. MT19937 with
// next_double = ((w0 >> 5) * 2^26 + (w1 >> 6)) / 2^53 and x = low + (high - low) * next_double, values filled
// candidate-major.  At C3 that is 1.07e9 32-bit outputs: ~4.5 s of host time per call.
//
// The generator is sequential over 624-word blocks (inside a block: three data-parallel sweeps, k < 227 depends
// on old words only, 227 <= k < 454 and 454 <= k < 624 on words of the previous sweep).  One CTA therefore produces
// ~0.6 M doubles per ms, and the whole C3 matrix took 0.86 s.  Here the matrix is cut into row chunks, one per CTA,
// and every CTA first JUMPS to the generator state at the start of its chunk:
//
//   With T the one-word step of the recurrence, T^J s = g(T) s for g = x^J mod phi (phi: the characteristic
//   polynomial of T, degree 19937), and g(T) s = XOR of the 624-word windows of the stream from s that start at the
//   set bits of g.  g = x^(2^k) mod phi is tabulated for k = 16 .. 40 (mt19937_jump.inc, generated and checked by
//   tools/make_mt_jump_table.py); J is taken bit by bit from the top (one window convolution per set bit: 20 561 stream
//   words in shared memory, each thread XORs the ~10 k windows of one output word), the low 16 bits of J are walked
//   with the recurrence.  A window obtained this way has 31 meaningless bits (the low bits of its first word are not
//   part of the 19937-bit state), so a CTA jumps to ONE WORD BEFORE the 624-block that holds its first output and walks
//   the last word, which also keeps its blocks aligned with numpy's: the last chunk of the matrix ends in exactly the
//   (key, pos) numpy would hold.
//
// So a rank of a multi-GPU job draws only its own rows, and a single GPU draws the C3 matrix with 2 x 148 CTAs.  The
// state after the WHOLE matrix (which the host generator is moved to, so that the x_seeds drawn next are the
// reference's) is computed by one more CTA with the same jump.  Multiplication and addition of the scaling are kept
// separate (no FMA), as in numpy's C code.  Requires an even stream position (true whenever only doubles were drawn).
#pragma once
#include "sweep.cuh"

namespace plx {

#include "mt19937_jump.inc"

constexpr int MT_N = 624;
constexpr int MT_DEG = 19937;
constexpr int MT_SEQ = 33 * MT_N;                 // stream words held in shared memory: >= MT_DEG + MT_N - 1
constexpr int MT_THREADS = 640;                   // >= MT_N: one thread per output word of a window convolution

__device__ __forceinline__ unsigned mt_twist(unsigned a, unsigned b) {
  const unsigned y = (a & 0x80000000u) | (b & 0x7fffffffu);
  return (y >> 1) ^ ((0u - (y & 1u)) & 0x9908b0dfu);
}
__device__ __forceinline__ unsigned mt_temper(unsigned y) {
  y ^= (y >> 11);
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= (y >> 18);
  return y;
}

// x[have .. want) of the stream whose first `have` >= 624 words are in x: x[i + 624] = x[i + 397] ^ twist(x[i], x[i + 1]),
// 227 independent words per step.  All threads of the CTA call this.
__device__ __forceinline__ void mt_extend(unsigned* x, int have, int want) {
  for (int f = have; f < want; f += 227) {
    const int t = threadIdx.x;
    if (t < 227 && f + t < want) x[f + t] = x[f + t - 227] ^ mt_twist(x[f + t - MT_N], x[f + t - MT_N + 1]);
    __syncthreads();
  }
}

// seq[0 .. 624) <- g(T) applied to the state in seq[0 .. 624) (only the top bit of word 0 is meaningful, before and after)
__device__ __forceinline__ void mt_apply(unsigned* seq, const unsigned* __restrict__ g) {
  mt_extend(seq, MT_N, MT_DEG + MT_N);
  const int j = threadIdx.x;
  unsigned acc = 0;
  if (j < MT_N) {
    for (int w = 0; w < MT_N; ++w) {
      unsigned gw = __ldg(g + w);                  // same word for every thread
      const unsigned* base = seq + 32 * w + j;
      while (gw) {
        const int b = __ffs(gw) - 1;
        acc ^= base[b];
        gw &= gw - 1;
      }
    }
  }
  __syncthreads();
  if (j < MT_N) seq[j] = acc;
  __syncthreads();
}

// seq[0 .. 624) <- the window r >= 0 words further down the stream
__device__ __forceinline__ void mt_advance(unsigned* seq, long long r) {
  while (r > 0) {
    const int step = static_cast<int>(min(r, static_cast<long long>(MT_SEQ - MT_N)));
    mt_extend(seq, MT_N, MT_N + step);
    const int j = threadIdx.x;
    const unsigned v = (j < MT_N) ? seq[step + j] : 0u;
    __syncthreads();
    if (j < MT_N) seq[j] = v;
    __syncthreads();
    r -= step;
  }
}

// seq[0 .. 624) <- the exact 624 words x[o .. o + 624) of the stream with x[0 .. 624) = seq on entry; o a multiple of 624
__device__ __forceinline__ void mt_jump_to_block(unsigned* seq, long long o) {
  if (o == 0) return;
  const long long J = o - 1;
  for (int k = MT_JUMP_KMAX; k >= MT_JUMP_KMIN; --k)
    if ((J >> k) & 1) mt_apply(seq, g_mt_jump[k - MT_JUMP_KMIN]);
  mt_advance(seq, (J & ((1LL << MT_JUMP_KMIN) - 1)) + 1);
}

struct MtDrawParams {
  const unsigned* key; int pos;          // generator state on entry: x[0 .. 624) = key, next output = x[pos]
  const double* low; const double* range; int d;
  long long first_row;                   // global row index of out's first row
  long long rows;                        // rows written by this launch
  long long rows_per_cta;
  double* out;                           // rows x d
  long long total_words;                 // pos + 2 N d: stream position after the whole N x d matrix
  unsigned* key_out; int* pos_out;       // state after the whole matrix, in numpy's representation (last CTA)
};

// grid = chunks + 1: CTA c < chunks writes rows [c * rows_per_cta, ...) of this launch, the last CTA the end state.
__global__ void __launch_bounds__(MT_THREADS) mt19937_draw_kernel(const MtDrawParams p) {
  extern __shared__ unsigned mt_smem[];
  unsigned* seq = mt_smem;                                        // MT_SEQ words
  double* lr = reinterpret_cast<double*>(mt_smem + MT_SEQ);       // low[d], range[d]
  const int tid = threadIdx.x, d = p.d;
  for (int i = tid; i < MT_N; i += MT_THREADS) seq[i] = p.key[i];
  for (int i = tid; i < d; i += MT_THREADS) { lr[i] = p.low[i]; lr[d + i] = p.range[i]; }
  __syncthreads();
  if (blockIdx.x == gridDim.x - 1) {
    // numpy regenerates a block only when an output beyond it is asked for: after W words the array holds block
    // (W - 1) / 624 and pos = W - 624 * that (1 .. 624); W <= 624 leaves the array as it was
    const long long W = p.total_words;
    const long long ob = W <= MT_N ? 0 : (W - 1) / MT_N * MT_N;
    mt_jump_to_block(seq, ob);
    for (int i = tid; i < MT_N; i += MT_THREADS) p.key_out[i] = seq[i];
    if (tid == 0) *p.pos_out = static_cast<int>(W - ob);
    return;
  }
  const long long row0 = static_cast<long long>(blockIdx.x) * p.rows_per_cta;
  const long long count = min(p.rows_per_cta, p.rows - row0) * d;        // doubles of this CTA
  if (count <= 0) return;
  const long long o = p.pos + 2LL * d * (p.first_row + row0);            // stream offset of the first output word (even)
  const long long ob = o / MT_N * MT_N;
  mt_jump_to_block(seq, ob);
  // Threads 0-255 regenerate the NEXT 624-word block (three data-parallel sweeps, barriers among themselves only)
  // while threads 256-639 temper, scale and store the doubles of the CURRENT block; one CTA-wide barrier per block.
  unsigned* buf[2] = {seq, seq + MT_N};
  double* out = p.out + row0 * d;
  int cur = 0, pp = static_cast<int>(o - ob);
  int qbase = 0;                                 // (doubles produced so far) mod d: a chunk starts at a row
  long long produced = 0;
  while (produced < count) {
    const long long left = count - produced;
    const int pairs = (pp >= MT_N) ? 0 : static_cast<int>(min(static_cast<long long>((MT_N - pp) / 2), left));
    const bool more = left > pairs;              // another block will be needed after this one
    if (tid < 256) {
      if (more) {
        const unsigned* o_ = buf[cur];
        unsigned* nw = buf[cur ^ 1];
        if (tid < 227) nw[tid] = o_[tid + 397] ^ mt_twist(o_[tid], o_[tid + 1]);
        named_bar_sync(3, 256);
        if (tid < 227) { const int k = tid + 227; nw[k] = nw[k - 227] ^ mt_twist(o_[k], o_[k + 1]); }
        named_bar_sync(3, 256);
        if (tid < 169) { const int k = tid + 454; nw[k] = nw[k - 227] ^ mt_twist(o_[k], o_[k + 1]); }
        named_bar_sync(3, 256);
        if (tid == 0) nw[623] = nw[396] ^ mt_twist(o_[623], nw[0]);
      }
    } else {
      const unsigned* w = buf[cur];
      for (int t = tid - 256; t < pairs; t += MT_THREADS - 256) {
        const unsigned a = mt_temper(w[pp + 2 * t]) >> 5, b = mt_temper(w[pp + 2 * t + 1]) >> 6;
        const double u = (a * 67108864.0 + b) / 9007199254740992.0;
        const int q = (qbase + t) % d;
        out[produced + t] = __dadd_rn(lr[q], __dmul_rn(lr[d + q], u));
      }
    }
    produced += pairs;
    qbase = (qbase + pairs) % d;
    pp += 2 * pairs;
    __syncthreads();
    if (more) { cur ^= 1; pp = 0; }
  }
}

}  // namespace plx

// Stand-alone emission probabilities of the gene-structure emitters for ANY number of classes, states and gene-model
// copies: the 7-state SimpleGenePredHMMEmitter (class probabilities only, gene_pred_hmm_emitter.py:93-121) and
// GenePredHMMEmitter with num_copies >= 1 (class part x 3-mer codon factors, :231-263; the codon columns are repeated
// per copy, :254-257).  The 15-state single-copy model inside MsaHmmLayer never comes here (gene15.cu fuses its
// emission into the recursion); these emissions feed the generic dense recurrence (dense.cu).
//
// One thread per output element (sequence, position, state): the class dot product reads the position's input row
// (served by L1 for the q threads of a position) and one row of B; the codon factor of a constrained state is one entry
// of each 5-symbol 3-mer table (prepare_codon_kernel) for hard nucleotides, the soft 64-term contraction otherwise
// (kmer.py:3-47).  Writes are coalesced.
#include "common.cuh"

namespace hmm {
namespace emission_general {

struct GArgs {
  const float* inputs;   // [k][B][T][s + (has_codon ? 5 : 0)]
  const float* Bfull;    // [k][q][s]  one row per state (intron rows already repeated, make_B_full)
  const float* codon5;   // [2][125][12] or NULL
  const float* hints;    // [k][B][2][q] or NULL
  float* out;            // [k][B][T][q]
  int num_models, batch, T, s, q, num_copies, has_codon;
  unsigned flags;
};

// 0..3 exactly one-hot base, 4 uniform (N flag exactly 1 with no base set / outside the sequence), -1 anything else
__device__ __forceinline__ int nuc_code_at(const float* row_nuc, int t, int T, int stride) {
  if (t < 0 || t >= T) return 4;
  const float* p = row_nuc + (size_t)t * stride;
  const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2), w = __ldg(p + 3), f = __ldg(p + 4);
  const int zeros = (x == 0.f) + (y == 0.f) + (z == 0.f) + (w == 0.f);
  if (f == 1.f) return zeros == 4 ? 4 : -1;
  const int ones = (x == 1.f) + (y == 1.f) + (z == 1.f) + (w == 1.f);
  if (ones != 1 || zeros != 3) return -1;
  return (y == 1.f) ? 1 : (z == 1.f) ? 2 : (w == 1.f) ? 3 : 0;
}

// distribution over A,C,G,T at position t (kmer.py:21-33): N spreads 1/4, outside the sequence 1/4 each
__device__ __forceinline__ void nuc_dist_at(const float* row_nuc, int t, int T, int stride, float (&d)[4]) {
  if (t < 0 || t >= T) { d[0] = d[1] = d[2] = d[3] = 0.25f; return; }
  const float* p = row_nuc + (size_t)t * stride;
  const float add = (__ldg(p + 4) == 1.0f) ? 0.25f : 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) d[i] = __ldg(p + i) + add;
}

// one side's codon factor for column col: pivot position t, near / far neighbours at t + dir, t + 2 dir
__device__ float codon_side(const float* __restrict__ T5side, const float* row_nuc, int t, int T, int stride, int dir, int col) {
  const int c0 = nuc_code_at(row_nuc, t, T, stride), c1 = nuc_code_at(row_nuc, t + dir, T, stride),
            c2 = nuc_code_at(row_nuc, t + 2 * dir, T, stride);
  if ((c0 | c1 | c2) >= 0) return __ldg(T5side + ((c1 * 5 + c2) * 5 + c0) * 12 + col);
  float pv[4], nr[4], fr[4];
  nuc_dist_at(row_nuc, t, T, stride, pv);
  nuc_dist_at(row_nuc, t + dir, T, stride, nr);
  nuc_dist_at(row_nuc, t + 2 * dir, T, stride, fr);
  float acc = 0.f;
  for (int i = 0; i < 64; ++i) {
    const int a = i >> 4, b = (i >> 2) & 3, c = i & 3;
    const float w = nr[a] * fr[b] * pv[c];
    if (w != 0.f) acc = fmaf(w, __ldg(T5side + ((a * 5 + b) * 5 + c) * 12 + col), acc);
  }
  return acc;
}

__global__ void __launch_bounds__(256) emission_general_kernel(const GArgs a) {
  const int q = a.q, s = a.s, stride = s + (a.has_codon ? 5 : 0);
  const size_t total = (size_t)a.num_models * a.batch * a.T * q;
  const float train = (a.has_codon && (a.flags & HMM_B200_TRAINING)) ? 1e-7f : 0.f;
  for (size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
    const int j = (int)(idx % q);
    const size_t pos = idx / q;
    const int t = (int)(pos % a.T);
    const size_t n = pos / a.T;                 // sequence index over (model, batch)
    const int m = (int)(n / a.batch);
    const float* x = a.inputs + pos * stride;
    const float* b = a.Bfull + ((size_t)m * q + j) * s;
    float acc0 = 0.f, acc1 = 0.f;
    int c = 0;
    for (; c + 1 < s; c += 2) { acc0 = fmaf(__ldg(x + c), __ldg(b + c), acc0); acc1 = fmaf(__ldg(x + c + 1), __ldg(b + c + 1), acc1); }
    if (c < s) acc0 = fmaf(__ldg(x + c), __ldg(b + c), acc0);
    float e = acc0 + acc1;
    if (a.has_codon) {
      const int free_states = 1 + 5 * a.num_copies;          // Ir, I*, E0*, E1*: no 3-mer constraint (:258)
      float cod = 1.0f / 4096.0f;
      if (j >= free_states) {
        const int col = (j - free_states) / a.num_copies;    // E2, START, EI0-2, IE0-2, STOP, each repeated per copy
        const float* row_nuc = a.inputs + n * (size_t)a.T * stride + s;
        cod = codon_side(a.codon5, row_nuc, t, a.T, stride, +1, col) * codon_side(a.codon5 + 125 * 12, row_nuc, t, a.T, stride, -1, col);
      }
      e *= cod + train;
    }
    if (a.hints != nullptr) {
      if (t == 0) e *= __ldg(a.hints + (n * 2 + 0) * q + j);
      if (t == a.T - 1) e *= __ldg(a.hints + (n * 2 + 1) * q + j);
    }
    a.out[idx] = e;
  }
}

int run(const float* inputs, int num_models, int batch, int T, int s, int q, int num_copies, int has_codon, const float* Bfull,
        const float* codon5, const float* hints, unsigned flags, float* out, cudaStream_t st) {
  HMM_REQUIRE(inputs && Bfull && out, "emission_general: null inputs / Bfull / output");
  HMM_REQUIRE(num_models > 0 && batch > 0 && T > 0 && s > 0 && q > 0, "emission_general: sizes must be positive");
  HMM_REQUIRE(!has_codon || (codon5 && num_copies >= 1 && q == 1 + 14 * num_copies),
              "emission_general: the codon emitter needs its tables and q == 1 + 14 * num_copies (got q=%d, copies=%d)", q, num_copies);
  GArgs a{inputs, Bfull, codon5, hints, out, num_models, batch, T, s, q, num_copies, has_codon, flags};
  const size_t total = (size_t)num_models * batch * T * q;
  const int sms = device_sm_count();
  size_t blocks = (total + 255) / 256;
  const size_t cap = (size_t)sms * 16;
  if (blocks > cap) blocks = cap;
  emission_general_kernel<<<(unsigned)blocks, 256, 0, st>>>(a);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace emission_general
}  // namespace hmm

extern "C" int hmm_b200_emission_general(const float* inputs, int num_models, int batch, int T, int num_classes, int num_states,
                                         int num_copies, int has_codon, const float* Bfull, const float* codon5,
                                         const float* end_hints, unsigned flags, float* out_E, void* stream) {
  return hmm::emission_general::run(inputs, num_models, batch, T, num_classes, num_states, num_copies, has_codon, Bfull, codon5,
                                    end_hints, flags, out_E, static_cast<cudaStream_t>(stream));
}

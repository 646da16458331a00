// 15-state gene-structure HMM: fused emission + chunked scaled forward / backward / posterior for sm_100a.
//
// Three kernels per call:
//   products_kernel  K_A  per chunk of L positions: emission (never written to HBM) and the chunk's transfer
//                         matrix prod_t (A diag(e_t)) as 15 conditional forward rows, row-normalised
//                         (the reference's parallel_factor>1 form, MsaHmmCell.py:121-142, keeping only the
//                         chunk-end matrix);
//   scan_kernel      K_B  per sequence: prefix (forward) and suffix (backward) boundary vectors over the chunk
//                         matrices (the role of TotalProbabilityCell, TotalProbabilityCell.py:30-49, done in
//                         scaled probability space), log-likelihood;
//   finalize_kernel  K_C  per chunk: the exact sequential recursions of MsaHmmCell.py:73-106 (same floors, same
//                         order) started from the boundary vectors, both directions concurrently, then
//                         log posterior / log alpha / log beta written once.
//
// State order (gene_pred_hmm_transitioner.py:187-188): 0 Ir | 1-3 I0-2 | 4-6 E0-2 | 7 START | 8-10 EI0-2 |
// 11-13 IE0-2 | 14 STOP.  Only the 23 structural edges of A are read (gene_pred_hmm_transitioner.py:200-221).
#include "common.cuh"

namespace hmm {
namespace gene15 {

constexpr int Q = 15;      // states
constexpr int QP = 16;     // padded row
constexpr int SIN = 20;    // input floats per position: 15 class probabilities + 5 nucleotide one-hot
constexpr int NEDGE = 23;
constexpr int MAT = QP * QP;  // floats per stored chunk matrix

// edges in the reference's kernel order (from, to)
__device__ constexpr int ESRC[NEDGE] = {0, 0, 14, 7, 5, 4, 4, 8, 1, 1, 11, 5, 5, 9, 2, 2, 12, 6, 6, 10, 3, 3, 13};
__device__ constexpr int EDST[NEDGE] = {0, 7, 0, 5, 14, 5, 8, 1, 1, 11, 4, 6, 9, 2, 2, 12, 5, 4, 10, 3, 3, 13, 6};

struct Args {
  const float* inputs;
  const float* A;
  const float* pi;
  const float* Bfull;
  const float* codon;
  const float* hints;
  int num_models, batch, T;
  unsigned flags;
  int L, C;  // chunk length, chunks per sequence
  int GA, GC;
  float* Mn;    // [NB*C][16][16] row-normalised chunk matrices, common-scaled so the largest row scale is 1
  float* gsc;   // [NB*C]         log of the common scale
  float* fvec;  // [NB*C][16]     normalised alpha entering chunk c (pi for c = 0)
  float* rvec;  // [NB*C][16]     normalised R (backward variable) at the last position of chunk c
  float* fcum;  // [NB*C]         log scale belonging to fvec
  float* rcum;  // [NB*C]         log scale belonging to rvec
  float* llws;  // [NB]
  float* out_post;
  float* out_alpha;
  float* out_beta;
  float* loglik;
};

// ------------------------------------------------------------------------------------------------------------
// emission (gene_pred_hmm_emitter.py:93-121 class part, :231-263 3-mer part, kmer.py:3-47 soft 3-mers)
// ------------------------------------------------------------------------------------------------------------
struct EmisParams {
  float B[Q][QP];          // [state][class], column 15 = 0
  float codon[2][9][64];   // [left|right][E2,START,EI0-2,IE0-2,STOP][3-mer]
};

__device__ __forceinline__ void load_emis_params(EmisParams& sp, const float* __restrict__ Bm,
                                                 const float* __restrict__ codon) {
  for (int i = threadIdx.x; i < Q * QP; i += blockDim.x) {
    const int q = i >> 4, s = i & 15;
    sp.B[q][s] = (s < Q) ? __ldg(Bm + q * Q + s) : 0.f;
  }
  float* cd = &sp.codon[0][0][0];
  for (int i = threadIdx.x; i < 2 * 9 * 64; i += blockDim.x) cd[i] = __ldg(codon + i);
}

// distribution over A,C,G,T at position t; N (flag exactly 1) and out-of-range positions count 1/4 each
__device__ __forceinline__ float4 nuc_dist(const float* __restrict__ seq, int t, int T) {
  if (t < 0 || t >= T) return make_float4(.25f, .25f, .25f, .25f);
  const float* p = seq + (size_t)t * SIN + Q;
  const float add = (__ldg(p + 4) == 1.0f) ? 0.25f : 0.f;
  return make_float4(__ldg(p) + add, __ldg(p + 1) + add, __ldg(p + 2) + add, __ldg(p + 3) + add);
}

// 0..3 when p is exactly one-hot, -1 otherwise
__device__ __forceinline__ int hard_code(const float4 p) {
  const int ones = (p.x == 1.f) + (p.y == 1.f) + (p.z == 1.f) + (p.w == 1.f);
  const int zeros = (p.x == 0.f) + (p.y == 0.f) + (p.z == 0.f) + (p.w == 0.f);
  if (ones != 1 || zeros != 3) return -1;
  return (p.y == 1.f) ? 1 : (p.z == 1.f) ? 2 : (p.w == 1.f) ? 3 : 0;
}

__device__ __forceinline__ float comp(const float4 p, int i) { return i == 0 ? p.x : i == 1 ? p.y : i == 2 ? p.z : p.w; }

// <soft 3-mer, table[s]> for the 9 constrained states. flat 3-mer index = (near*4 + far)*4 + pivot
__device__ __forceinline__ void codon_side(const float (*tab)[64], const float4 pv, const float4 nr, const float4 fr,
                                           float (&out)[9]) {
  const int c0 = hard_code(pv), c1 = hard_code(nr), c2 = hard_code(fr);
  if ((c0 | c1 | c2) >= 0) {
    const int idx = (c1 * 4 + c2) * 4 + c0;
#pragma unroll
    for (int s = 0; s < 9; ++s) out[s] = tab[s][idx];
  } else {
#pragma unroll
    for (int s = 0; s < 9; ++s) out[s] = 0.f;
#pragma unroll 1
    for (int i = 0; i < 64; ++i) {
      const float w = comp(nr, i >> 4) * comp(fr, (i >> 2) & 3) * comp(pv, i & 3);
      if (w != 0.f) {
#pragma unroll
        for (int s = 0; s < 9; ++s) out[s] = fmaf(w, tab[s][i], out[s]);
      }
    }
  }
}

// emission probabilities of one position, floored at eps when FLOOR; dst[0..15] (slot 15 = 0)
template <bool FLOOR>
__device__ __forceinline__ void emit_position(const EmisParams& sp, const float* __restrict__ seq, int t, int T,
                                              const float4* __restrict__ P /* P[-2..2] around t */,
                                              const float* __restrict__ hints, unsigned flags, float* dst) {
  const float* row = seq + (size_t)t * SIN;
  const float4 x0 = ldg4(row), x1 = ldg4(row + 4), x2 = ldg4(row + 8), x3 = ldg4(row + 12);
  float fl[9], fr[9];
  codon_side(sp.codon[0], P[0], P[1], P[2], fl);
  codon_side(sp.codon[1], P[0], P[-1], P[-2], fr);
  const float train = (flags & HMM_B200_TRAINING) ? 1e-7f : 0.f;
  float e[QP];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const float4* b = reinterpret_cast<const float4*>(sp.B[q]);
    float acc = dot4(x0, b[0], 0.f);
    acc = dot4(x1, b[1], acc);
    acc = dot4(x2, b[2], acc);
    acc = dot4(x3, b[3], acc);  // b[3].w == 0 removes x3.w (a nucleotide flag)
    const float cod = ((q < 6) ? (1.0f / 4096.0f) : fl[q - 6] * fr[q - 6]) + train;
    e[q] = acc * cod;
  }
  if (hints != nullptr) {
    if (t == 0) {
#pragma unroll
      for (int q = 0; q < Q; ++q) e[q] *= __ldg(hints + q);
    }
    if (t == T - 1) {
#pragma unroll
      for (int q = 0; q < Q; ++q) e[q] *= __ldg(hints + Q + q);
    }
  }
  e[15] = 0.f;
  const float lo = FLOOR ? kEps : -INFINITY;
  float4* d4 = reinterpret_cast<float4*>(dst);
#pragma unroll
  for (int v = 0; v < 4; ++v)
    d4[v] = make_float4(fmaxf(e[4 * v], lo), fmaxf(e[4 * v + 1], lo), fmaxf(e[4 * v + 2], lo),
                        v == 3 ? 0.f : fmaxf(e[4 * v + 3], lo));
}

// emission of every position of the G chunks [c_base, c_base+G) of sequence n into Es (chunk stride CS floats)
template <bool FLOOR>
__device__ __forceinline__ void emission_stage(const Args& a, const EmisParams& sp, float4* Ps, float* Es, int n,
                                               int c_base, int G, int CS) {
  const int L = a.L, T = a.T;
  const float* seq = a.inputs + (size_t)n * T * SIN;
  const float* hints = a.hints ? a.hints + (size_t)n * 2 * Q : nullptr;
  const int LH = L + 4;
  for (int i = threadIdx.x; i < G * LH; i += blockDim.x) {
    const int ci = i / LH, u = i - ci * LH;
    const int c = c_base + ci;
    Ps[i] = (c < a.C) ? nuc_dist(seq, c * L - 2 + u, T) : make_float4(.25f, .25f, .25f, .25f);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < G * L; i += blockDim.x) {
    const int ci = i / L, u = i - ci * L;
    const int c = c_base + ci, t = c * L + u;
    if (c < a.C && t < T) emit_position<FLOOR>(sp, seq, t, T, Ps + ci * LH + u + 2, hints, a.flags, Es + ci * CS + u * QP);
  }
  __syncthreads();
}

// ------------------------------------------------------------------------------------------------------------
// sparse 15-state transition products
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void load_edges(const float* __restrict__ Am, float (&w)[NEDGE]) {
#pragma unroll
  for (int e = 0; e < NEDGE; ++e) w[e] = __ldg(Am + ESRC[e] * Q + EDST[e]);
}
// R = a . A   (forward, MsaHmmCell.py:80 with transitioner.reverse == False)
__device__ __forceinline__ void vec_A(const float (&a)[Q], const float (&w)[NEDGE], float (&R)[Q]) {
#pragma unroll
  for (int j = 0; j < Q; ++j) R[j] = 0.f;
#pragma unroll
  for (int e = 0; e < NEDGE; ++e) R[EDST[e]] = fmaf(a[ESRC[e]], w[e], R[EDST[e]]);
}
// R = A . b   (backward: b . A^T, transitioner.reverse == True)
__device__ __forceinline__ void A_vec(const float (&b)[Q], const float (&w)[NEDGE], float (&R)[Q]) {
#pragma unroll
  for (int i = 0; i < Q; ++i) R[i] = 0.f;
#pragma unroll
  for (int e = 0; e < NEDGE; ++e) R[ESRC[e]] = fmaf(w[e], b[EDST[e]], R[ESRC[e]]);
}

__device__ __forceinline__ void load_row16(const float* p, float (&v)[QP]) {
  const float4* p4 = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float4 t = p4[k];
    v[4 * k] = t.x; v[4 * k + 1] = t.y; v[4 * k + 2] = t.z; v[4 * k + 3] = t.w;
  }
}
__device__ __forceinline__ void store_row16(float* p, const float (&v)[Q], float last) {
  float4* p4 = reinterpret_cast<float4*>(p);
  p4[0] = make_float4(v[0], v[1], v[2], v[3]);
  p4[1] = make_float4(v[4], v[5], v[6], v[7]);
  p4[2] = make_float4(v[8], v[9], v[10], v[11]);
  p4[3] = make_float4(v[12], v[13], v[14], last);
}
__device__ __forceinline__ float sum15(const float (&v)[Q]) {
  return ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7])) + (((v[8] + v[9]) + (v[10] + v[11])) +
         ((v[12] + v[13]) + v[14]));
}

// one scaled step: a <- normalise(E ⊙ max(R, eps)), returns the scale S   (MsaHmmCell.py:87-92)
__device__ __forceinline__ float scaled_step(const float* Erow, const float (&R)[Q], float (&a)[Q]) {
  float e[QP];
  load_row16(Erow, e);
#pragma unroll
  for (int j = 0; j < Q; ++j) a[j] = e[j] * fmaxf(R[j], kEps);
  const float S = sum15(a);
  const float inv = __frcp_rn(S);
#pragma unroll
  for (int j = 0; j < Q; ++j) a[j] *= inv;
  return S;
}

// ------------------------------------------------------------------------------------------------------------
// K_A: chunk transfer matrices
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) products_kernel(const Args a) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int G = blockDim.x >> 4, L = a.L;
  const int CS = L * QP + 4;
  EmisParams& sp = *reinterpret_cast<EmisParams*>(smem);
  float4* Ps = reinterpret_cast<float4*>(smem + sizeof(EmisParams));
  float* Es = reinterpret_cast<float*>(Ps + G * (L + 4));
  const int n = blockIdx.y, m = n / a.batch, c_base = blockIdx.x * G;
  load_emis_params(sp, a.Bfull + (size_t)m * Q * Q, a.codon);
  __syncthreads();
  emission_stage<true>(a, sp, Ps, Es, n, c_base, G, CS);

  const int ci = threadIdx.x >> 4, row = threadIdx.x & 15, c = c_base + ci;
  const bool valid = (row < Q) && (c < a.C);
  float av[Q];
  float ls = 0.f;
  if (valid) {
    float w[NEDGE];
    load_edges(a.A + (size_t)m * Q * Q, w);
    const int len = min(L, a.T - c * L);
    const float* Ep = Es + ci * CS;
#pragma unroll
    for (int j = 0; j < Q; ++j) av[j] = (j == row) ? 1.f : 0.f;
    for (int u = 0; u < len; ++u) {
      float R[Q];
      if (c == 0 && u == 0) {
#pragma unroll
        for (int j = 0; j < Q; ++j) R[j] = av[j];  // init step: no transition (MsaHmmCell.py:77-78)
      } else {
        vec_A(av, w, R);
      }
      ls += logf(scaled_step(Ep + u * QP, R, av));
    }
  }
  const float g = half_max(valid ? ls : -INFINITY);
  if (c < a.C) {
    float* dst = a.Mn + ((size_t)n * a.C + c) * MAT + row * QP;
    if (row < Q) {
      const float sc = expf(ls - g);
#pragma unroll
      for (int j = 0; j < Q; ++j) av[j] *= sc;
      store_row16(dst, av, 0.f);
      if (row == 0) a.gsc[(size_t)n * a.C + c] = g;
    } else {
      float4* d4 = reinterpret_cast<float4*>(dst);
      d4[0] = d4[1] = d4[2] = d4[3] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// K_B: boundary scan over the chunk matrices of one sequence
// ------------------------------------------------------------------------------------------------------------
constexpr int SCAN_WARPS = 16;

// v'_l = sum_k v_k * coef[k]; lanes of a half warp hold v; returns the un-normalised result
__device__ __forceinline__ float half_matvec(float v, const float (&coef)[Q]) {
  float acc = 0.f;
#pragma unroll
  for (int k = 0; k < Q; ++k) acc = fmaf(__shfl_sync(0xffffffffu, v, k, 16), coef[k], acc);
  return acc;
}

__global__ void __launch_bounds__(SCAN_WARPS * 32) scan_kernel(const Args a) {
  __shared__ __align__(16) float Xs[SCAN_WARPS][QP][QP];
  __shared__ float Pg[SCAN_WARPS];
  __shared__ float fseg[SCAN_WARPS + 1][QP], rseg[SCAN_WARPS][QP];
  __shared__ double fcs[SCAN_WARPS + 1], rcs[SCAN_WARPS];
  const int n = blockIdx.x, m = n / a.batch, C = a.C;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, hl = lane & 15, hi = lane >> 4;
  const int SL = (C + SCAN_WARPS - 1) / SCAN_WARPS;
  const int c_lo = min(C, w * SL), c_hi = min(C, c_lo + SL);
  const float* Mb = a.Mn + (size_t)n * C * MAT;
  const float* gb = a.gsc + (size_t)n * C;

  // phase 1: ordered product of the segment's matrices, X <- X . M_c, rescaled by its largest entry
  {
    float(*X)[QP] = Xs[w];
    for (int i = lane; i < MAT; i += 32) X[i >> 4][i & 15] = ((i >> 4) == (i & 15)) ? 1.f : 0.f;
    __syncwarp();
    float gs = 0.f;
    for (int c = c_lo; c < c_hi; ++c) {
      float col[Q];
#pragma unroll
      for (int k = 0; k < Q; ++k) col[k] = __ldg(Mb + (size_t)c * MAT + k * QP + hl);
      float acc[8];
      float mx = 0.f;
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        float x[QP];
        load_row16(&X[hi * 8 + r][0], x);
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < Q; ++k) s = fmaf(x[k], col[k], s);
        acc[r] = s;
        mx = fmaxf(mx, s);
      }
      mx = fmaxf(warp_max(mx), 1e-37f);
      const float inv = __frcp_rn(mx);
      __syncwarp();
#pragma unroll
      for (int r = 0; r < 8; ++r) X[hi * 8 + r][hl] = (hl < Q && hi * 8 + r < Q) ? acc[r] * inv : 0.f;
      __syncwarp();
      gs += __ldg(gb + c) + logf(mx);
    }
    if (lane == 0) Pg[w] = gs;
  }
  __syncthreads();

  // phase 2: warp 0 scans the segment products; lanes 0-15 forward (row vector), lanes 16-31 backward (column)
  if (w == 0) {
    float v = (hl < Q) ? (hi == 0 ? __ldg(a.pi + m * Q + hl) : 1.f) : 0.f;
    double cs = 0.0;
    for (int it = 0; it < SCAN_WARPS; ++it) {
      const int s = hi == 0 ? it : SCAN_WARPS - 1 - it;
      if (hi == 0) { fseg[s][hl] = v; if (hl == 0) fcs[s] = cs; }
      else         { rseg[s][hl] = v; if (hl == 0) rcs[s] = cs; }
      float coef[Q];
#pragma unroll
      for (int k = 0; k < Q; ++k) coef[k] = hi == 0 ? Xs[s][k][hl] : Xs[s][hl][k];
      float nv = half_matvec(v, coef);
      const float S = fmaxf(half_sum(nv), 1e-37f);
      v = nv * __frcp_rn(S);
      cs += (double)Pg[s] + (double)logf(S);
    }
    if (hi == 0) { fseg[SCAN_WARPS][hl] = v; if (hl == 0) fcs[SCAN_WARPS] = cs; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const float ll = (float)fcs[SCAN_WARPS];
    a.llws[n] = ll;
    if (a.loglik) a.loglik[n] = ll;
  }

  // phase 3: every warp replays its segment from the segment boundary, emitting per-chunk boundary vectors
  {
    float v = hi == 0 ? fseg[w][hl] : rseg[w][hl];
    double cs = hi == 0 ? fcs[w] : rcs[w];
    float* vout = (hi == 0 ? a.fvec : a.rvec) + (size_t)n * C * QP;
    float* cout = (hi == 0 ? a.fcum : a.rcum) + (size_t)n * C;
    for (int it = 0; it < c_hi - c_lo; ++it) {
      const int c = hi == 0 ? c_lo + it : c_hi - 1 - it;
      vout[(size_t)c * QP + hl] = v;
      if (hl == 0) cout[c] = (float)cs;
      float coef[Q];
      const float* Mc = Mb + (size_t)c * MAT;
#pragma unroll
      for (int k = 0; k < Q; ++k) coef[k] = hi == 0 ? __ldg(Mc + k * QP + hl) : __ldg(Mc + hl * QP + k);
      float nv = half_matvec(v, coef);
      const float S = fmaxf(half_sum(nv), 1e-37f);
      v = nv * __frcp_rn(S);
      cs += (double)__ldg(gb + c) + (double)logf(S);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// K_C: exact in-chunk recursions from the boundary vectors + outputs
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) finalize_kernel(const Args a) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int G = a.GC, L = a.L, T = a.T, C = a.C;
  const int CS = L * QP + 4;
  EmisParams& sp = *reinterpret_cast<EmisParams*>(smem);
  float4* Ps = reinterpret_cast<float4*>(smem + sizeof(EmisParams));
  float* Es = reinterpret_cast<float*>(Ps + G * (L + 4));
  float* Fa = Es + G * CS;
  float* Rb = Fa + G * CS;
  const int n = blockIdx.y, m = n / a.batch, c_base = blockIdx.x * G;
  const bool want_f = a.out_post || a.out_alpha, want_b = a.out_post || a.out_beta;
  load_emis_params(sp, a.Bfull + (size_t)m * Q * Q, a.codon);
  __syncthreads();
  emission_stage<true>(a, sp, Ps, Es, n, c_base, G, CS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp < 2 && lane < G && c_base + lane < C) {
    const int c = c_base + lane, t0 = c * L, len = min(L, T - t0);
    const float* Ep = Es + lane * CS;
    float w[NEDGE];
    load_edges(a.A + (size_t)m * Q * Q, w);
    float bv[QP];
    if (warp == 0 && want_f) {
      // forward: alpha-hat_t, MsaHmmCell.py:73-106 with reverse == False
      load_row16((a.fvec + ((size_t)n * C + c) * QP), bv);
      float av[Q], R[Q];
#pragma unroll
      for (int j = 0; j < Q; ++j) av[j] = bv[j];
      float* out = Fa + lane * CS;
      for (int u = 0; u < len; ++u) {
        if (c == 0 && u == 0) {
#pragma unroll
          for (int j = 0; j < Q; ++j) R[j] = av[j];
        } else {
          vec_A(av, w, R);
        }
        const float S = scaled_step(Ep + u * QP, R, av);
        store_row16(out + u * QP, av, S);
      }
    } else if (warp == 1 && want_b) {
      // backward: R_t = beta_t before the emission at t, MsaHmmCell.py:96-100 with reverse == True
      load_row16((a.rvec + ((size_t)n * C + c) * QP), bv);
      float R[Q], b[Q];
#pragma unroll
      for (int j = 0; j < Q; ++j) R[j] = bv[j];
      float* out = Rb + lane * CS;
      for (int u = len - 1; u >= 0; --u) {
        float Rf[Q];
#pragma unroll
        for (int j = 0; j < Q; ++j) Rf[j] = fmaxf(R[j], kEps);
        const float S = scaled_step(Ep + u * QP, Rf, b);
        store_row16(out + u * QP, Rf, S);
        A_vec(b, w, R);
      }
    }
  }
  __syncthreads();

  // running log scales for log alpha / log beta (slot 15 of each row: S_t -> cumulative log scale)
  if (a.out_alpha || a.out_beta) {
    const int ci = threadIdx.x >> 1, dir = threadIdx.x & 1;
    if (ci < G && c_base + ci < C) {
      const int c = c_base + ci, len = min(L, T - c * L);
      if (dir == 0 && a.out_alpha) {
        double cs = (double)a.fcum[(size_t)n * C + c];
        float* p = Fa + ci * CS + 15;
        for (int u = 0; u < len; ++u) { cs += (double)logf(p[u * QP]); p[u * QP] = (float)cs; }
      } else if (dir == 1 && a.out_beta) {
        double cs = (double)a.rcum[(size_t)n * C + c];
        float* p = Rb + ci * CS + 15;
        for (int u = len - 1; u >= 0; --u) { const float S = p[u * QP]; p[u * QP] = (float)cs; cs += (double)logf(S); }
      }
    }
    __syncthreads();
  }

  // posterior normaliser per position: log sum_q alpha-hat * R-hat, kept in Es[.][0]
  if (a.out_post) {
    const float shift = (a.flags & HMM_B200_NO_LOGLIK) ? a.llws[n] : 0.f;
    for (int i = threadIdx.x; i < G * L; i += blockDim.x) {
      const int ci = i / L, u = i - ci * L, c = c_base + ci;
      if (c < C && c * L + u < T) {
        float f[QP], r[QP];
        load_row16(Fa + ci * CS + u * QP, f);
        load_row16(Rb + ci * CS + u * QP, r);
        float z = 0.f;
#pragma unroll
        for (int j = 0; j < Q; ++j) z = fmaf(f[j], r[j], z);
        Es[ci * CS + u * QP] = shift - logf(z);
      }
    }
    __syncthreads();
  }

  for (int ci = 0; ci < G; ++ci) {
    const int c = c_base + ci;
    if (c >= C) break;
    const int t0 = c * L, len = min(L, T - t0);
    const size_t obase = ((size_t)n * T + t0) * Q;
    const float* f = Fa + ci * CS;
    const float* r = Rb + ci * CS;
    const float* z = Es + ci * CS;
    for (int i = threadIdx.x; i < len * Q; i += blockDim.x) {
      const int u = i / Q, j = i - u * Q;
      float lf = 0.f, lr = 0.f;
      if (want_f) lf = logf(f[u * QP + j]);
      if (want_b) lr = logf(r[u * QP + j]);
      if (a.out_post) a.out_post[obase + i] = lf + lr + z[u * QP];
      if (a.out_alpha) a.out_alpha[obase + i] = lf + f[u * QP + 15];
      if (a.out_beta) a.out_beta[obase + i] = lr + r[u * QP + 15];
    }
  }
}

// stand-alone emission (cell.emission_probs API parity); unfloored
__global__ void __launch_bounds__(256) emission_kernel(const Args a, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int G = a.GC, L = a.L, T = a.T;
  const int CS = L * QP + 4;
  EmisParams& sp = *reinterpret_cast<EmisParams*>(smem);
  float4* Ps = reinterpret_cast<float4*>(smem + sizeof(EmisParams));
  float* Es = reinterpret_cast<float*>(Ps + G * (L + 4));
  const int n = blockIdx.y, m = n / a.batch, c_base = blockIdx.x * G;
  load_emis_params(sp, a.Bfull + (size_t)m * Q * Q, a.codon);
  __syncthreads();
  emission_stage<false>(a, sp, Ps, Es, n, c_base, G, CS);
  for (int ci = 0; ci < G; ++ci) {
    const int c = c_base + ci;
    if (c >= a.C) break;
    const int t0 = c * L, len = min(L, T - t0);
    const size_t obase = ((size_t)n * T + t0) * Q;
    for (int i = threadIdx.x; i < len * Q; i += blockDim.x) {
      const int u = i / Q, j = i - u * Q;
      out[obase + i] = Es[ci * CS + u * QP + j];
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
static int g_L = 64, g_GA = 8, g_GC = 8;

// optional per-kernel timing (bench.py's roofline leg): events around each launch of the last profiled call
static bool g_prof = false;
static cudaEvent_t g_ev[4] = {nullptr, nullptr, nullptr, nullptr};
static int g_ev_n = 0;

int profile_enable(int on) {
  if (on && !g_ev[0])
    for (int i = 0; i < 4; ++i) HMM_CUDA_OK(cudaEventCreate(&g_ev[i]));
  g_prof = on != 0;
  g_ev_n = 0;
  return 0;
}
int profile_read(float* ms, int n) {
  if (g_ev_n < 2) return 0;
  HMM_CUDA_OK(cudaEventSynchronize(g_ev[g_ev_n - 1]));
  int k = 0;
  for (; k < g_ev_n - 1 && k < n; ++k) HMM_CUDA_OK(cudaEventElapsedTime(&ms[k], g_ev[k], g_ev[k + 1]));
  return k;
}

struct Layout {
  size_t Mn, gsc, fvec, rvec, fcum, rcum, llws, total;
};
static Layout layout(int NB, int C) {
  Layout l;
  size_t off = 0;
  const size_t NC = (size_t)NB * C;
  l.Mn = off;   off = align_up(off + NC * MAT * sizeof(float), 256);
  l.gsc = off;  off = align_up(off + NC * sizeof(float), 256);
  l.fvec = off; off = align_up(off + NC * QP * sizeof(float), 256);
  l.rvec = off; off = align_up(off + NC * QP * sizeof(float), 256);
  l.fcum = off; off = align_up(off + NC * sizeof(float), 256);
  l.rcum = off; off = align_up(off + NC * sizeof(float), 256);
  l.llws = off; off = align_up(off + (size_t)NB * sizeof(float), 256);
  l.total = off;
  return l;
}

static size_t smem_products(int G, int L) { return sizeof(EmisParams) + (size_t)G * (L + 4) * 16 + (size_t)G * (L * QP + 4) * 4; }
static size_t smem_finalize(int G, int L) { return sizeof(EmisParams) + (size_t)G * (L + 4) * 16 + 3 * (size_t)G * (L * QP + 4) * 4; }

size_t workspace_size(int num_models, int batch, int T) {
  if (num_models <= 0 || batch <= 0 || T <= 0) return 0;
  return layout(num_models * batch, cdiv(T, g_L)).total;
}

int set_tuning(int L, int GA, int GC) {
  if (L < 0 || GA < 0 || GC < 0 || GA > 16 || (GA & 1) || GC > 32 || L > 1024) return HMM_B200_EINVAL;
  if (L) g_L = L;
  if (GA) g_GA = GA;
  if (GC) g_GC = GC;
  if (smem_finalize(g_GC, g_L) > 227 * 1024 || smem_products(g_GA, g_L) > 227 * 1024) {
    g_L = 64; g_GA = 8; g_GC = 8;
    set_error("tuning rejected: shared memory footprint exceeds 227 KB");
    return HMM_B200_EINVAL;
  }
  return 0;
}

static int fill_args(Args& a, const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                     const float* Bfull, const float* codon, const float* hints, unsigned flags) {
  HMM_REQUIRE(inputs && Bfull && codon, "genepred: null input pointer");
  HMM_REQUIRE(num_models > 0 && batch > 0 && T > 0, "genepred: num_models, batch and T must be positive (got %d,%d,%d)",
              num_models, batch, T);
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(inputs) & 15) == 0, "genepred: inputs must be 16-byte aligned");
  HMM_REQUIRE((long long)num_models * batch <= 65535, "genepred: num_models*batch must be <= 65535");
  a.inputs = inputs; a.A = A; a.pi = pi; a.Bfull = Bfull; a.codon = codon; a.hints = hints;
  a.num_models = num_models; a.batch = batch; a.T = T; a.flags = flags;
  a.L = g_L; a.C = cdiv(T, g_L); a.GA = g_GA; a.GC = g_GC;
  return 0;
}

int run(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
        const float* codon, const float* hints, unsigned flags, float* out_post, float* out_alpha, float* out_beta,
        float* loglik, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  int rc = fill_args(a, inputs, num_models, batch, T, A, pi, Bfull, codon, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE(A && pi, "genepred: null A / pi");
  HMM_REQUIRE(out_post || out_alpha || out_beta || loglik, "genepred: no output requested");
  const int NB = num_models * batch;
  const Layout l = layout(NB, a.C);
  if (!workspace || workspace_bytes < l.total) {
    set_error("genepred: workspace too small (%zu < %zu bytes)", workspace_bytes, l.total);
    return HMM_B200_EWORKSPACE;
  }
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "genepred: workspace must be 256-byte aligned");
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  a.Mn = reinterpret_cast<float*>(ws + l.Mn);     a.gsc = reinterpret_cast<float*>(ws + l.gsc);
  a.fvec = reinterpret_cast<float*>(ws + l.fvec); a.rvec = reinterpret_cast<float*>(ws + l.rvec);
  a.fcum = reinterpret_cast<float*>(ws + l.fcum); a.rcum = reinterpret_cast<float*>(ws + l.rcum);
  a.llws = reinterpret_cast<float*>(ws + l.llws);
  a.out_post = out_post; a.out_alpha = out_alpha; a.out_beta = out_beta; a.loglik = loglik;

  const size_t smA = smem_products(a.GA, a.L), smC = smem_finalize(a.GC, a.L);
  static bool attr_done = false;
  static size_t attrA = 0, attrC = 0;
  if (!attr_done || smA > attrA || smC > attrC) {
    HMM_CUDA_OK(cudaFuncSetAttribute(products_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smA));
    HMM_CUDA_OK(cudaFuncSetAttribute(finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smC));
    attr_done = true; attrA = smA; attrC = smC;
  }
  if (g_prof) { g_ev_n = 0; cudaEventRecord(g_ev[g_ev_n++], st); }
  products_kernel<<<dim3(cdiv(a.C, a.GA), NB), a.GA * 16, smA, st>>>(a);
  if (g_prof) cudaEventRecord(g_ev[g_ev_n++], st);
  scan_kernel<<<NB, SCAN_WARPS * 32, 0, st>>>(a);
  if (g_prof) cudaEventRecord(g_ev[g_ev_n++], st);
  if (out_post || out_alpha || out_beta) {
    finalize_kernel<<<dim3(cdiv(a.C, a.GC), NB), 256, smC, st>>>(a);
    if (g_prof) cudaEventRecord(g_ev[g_ev_n++], st);
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int emission(const float* inputs, int num_models, int batch, int T, const float* Bfull, const float* codon,
             const float* hints, unsigned flags, float* out_E, cudaStream_t st) {
  Args a{};
  int rc = fill_args(a, inputs, num_models, batch, T, nullptr, nullptr, Bfull, codon, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE(out_E, "genepred_emission: null output");
  a.L = 64; a.GC = 8; a.C = cdiv(T, a.L);
  const size_t sm = smem_products(a.GC, a.L);
  static bool done = false;
  if (!done) {
    HMM_CUDA_OK(cudaFuncSetAttribute(emission_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    done = true;
  }
  emission_kernel<<<dim3(cdiv(a.C, a.GC), num_models * batch), 256, sm, st>>>(a, out_E);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace gene15
}  // namespace hmm

extern "C" {

size_t hmm_b200_genepred_workspace_size(int num_models, int batch, int T) {
  return hmm::gene15::workspace_size(num_models, batch, T);
}

int hmm_b200_profile_enable(int on) { return hmm::gene15::profile_enable(on); }
int hmm_b200_profile_read(float* ms_host, int n) { return hmm::gene15::profile_read(ms_host, n); }

int hmm_b200_set_tuning(int chunk_len, int chunks_per_products_cta, int chunks_per_finalize_cta) {
  return hmm::gene15::set_tuning(chunk_len, chunks_per_products_cta, chunks_per_finalize_cta);
}

int hmm_b200_genepred_run(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                          const float* Bfull, const float* codon, const float* end_hints, unsigned flags,
                          float* out_post, float* out_alpha, float* out_beta, float* loglik, void* workspace,
                          size_t workspace_bytes, void* stream) {
  return hmm::gene15::run(inputs, num_models, batch, T, A, pi, Bfull, codon, end_hints, flags, out_post, out_alpha,
                          out_beta, loglik, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_emission(const float* inputs, int num_models, int batch, int T, const float* Bfull,
                               const float* codon, const float* end_hints, unsigned flags, float* out_E, void* stream) {
  return hmm::gene15::emission(inputs, num_models, batch, T, Bfull, codon, end_hints, flags, out_E,
                               static_cast<cudaStream_t>(stream));
}

}  // extern "C"

// Viterbi decoding of the 15-state gene model as a chunked (max,+) scan: the counterpart of the sum-product path's
// chunk matrices / scan / in-chunk recursion (gene15.cu), so that a sequence is decoded by thousands of threads instead of
// one warp walking T dependent steps (viterbi_kernel, gene15.cu: 0.73 ms at config 2).
//
// The recursion is the one viterbi_kernel documents -- the max-product form of MsaHmmCell.py:73-106 over the same floored
// quantities, ties to the lowest source state:
//   delta_0 = log max(pi, eps) + log max(E_0, eps),   delta_t(j) = max_i (delta_{t-1}(i) + log A[i][j]) + log max(E_t(j), eps)
// A sequence is cut into C <= 64 chunks of L positions.
//   viterbi_rows_kernel    thread = (chunk, start state i): row i of the chunk's (max,+) transfer matrix -- the best log score
//                          of reaching every state at the chunk's last position from state i at the position in front of
//                          it -- as a 15-vector in registers over the model's 23 edges; the 16 lanes of a chunk take a common
//                          maximum out every 16 steps (kept apart in double), so the stored rows stay small numbers.
//   viterbi_chunks_kernel  one CTA per sequence.  (1) one warp folds the chunk matrices left to right: the exact delta
//                          vector entering every chunk (up to a constant).  (2) half-warp = chunk, lane = state: the plain
//                          recursion inside every chunk from its entering vector, 2 back-pointer bits per state and step kept in
//                          shared memory, and beside it the map "state at the chunk's end -> state in front of the chunk".
//                          (3) one thread chains the 64 maps backwards from the best final state.  (4) thread = chunk walks its
//                          back-pointers from its now known end state; the path leaves coalesced.
// The emissions come from emission_kernel's log-floored buffer as before.
#include "gene15_dev.cuh"

namespace hmm {
namespace gene15 {

constexpr int VC = 64;              // chunks per sequence (at most)
constexpr int VROW_THREADS = 256;   // rows kernel: 16 chunks x 16 lanes
constexpr int VCH_THREADS = 1024;   // chunks kernel: 64 chunks x 16 lanes
constexpr int VPF = 8;              // emission rows requested ahead in the in-chunk recursion

// in-edges of every state in ascending source order (ESRC / EDST of gene15_dev.cuh regrouped by destination)
__device__ constexpr int VIN_CNT[Q] = {2, 2, 2, 2, 2, 3, 2, 1, 1, 1, 1, 1, 1, 1, 1};
__device__ constexpr int VIN_SRC[Q][3] = {{0, 14, 0}, {1, 8, 0}, {2, 9, 0}, {3, 10, 0}, {6, 11, 0}, {4, 7, 12}, {5, 13, 0}, {0, 0, 0},
                                          {4, 0, 0},  {5, 0, 0}, {6, 0, 0}, {1, 0, 0},  {2, 0, 0},  {3, 0, 0},  {5, 0, 0}};

__device__ __forceinline__ float log_or_ninf(float a) { return a > 0.f ? logf(a) : -INFINITY; }
__device__ __forceinline__ float half_max(float v, unsigned mask = 0xffffffffu) {   // maximum over the 16 lanes of a half warp
#pragma unroll
  for (int o = 8; o >= 1; o >>= 1) v = fmaxf(v, __shfl_xor_sync(mask, v, o, 16));
  return v;
}

// ---- rows of the chunk matrices ----
// Mv: [NB][VC][16 rows][16] floats (row 15 and column 15 unused), offs: [NB][VC] doubles (the maximum taken out of chunk c's
// rows).  Chunk 0 has no matrix: its row 0 is the delta vector at its last position.
__global__ void __launch_bounds__(VROW_THREADS) viterbi_rows_kernel(const float* __restrict__ E, const float* __restrict__ A,
                                                                    const float* __restrict__ pi, int batch, int T, int L, int C,
                                                                    float* __restrict__ Mv, double* __restrict__ offs) {
  const int n = blockIdx.y, m = n / batch, tid = threadIdx.x;
  const int c_raw = blockIdx.x * (VROW_THREADS / 16) + (tid >> 4), i = tid & 15;
  const bool valid = c_raw < C;
  const unsigned hmask = 0xffffu << (tid & 16);   // the two half warps of a warp walk chunks of different lengths: each syncs itself
  const int c = valid ? c_raw : C - 1;   // a half warp behind the last chunk repeats it (the shuffles are warp-wide), stores nothing
  float la[Q][3];
#pragma unroll
  for (int j = 0; j < Q; ++j)
#pragma unroll
    for (int k = 0; k < 3; ++k) la[j][k] = k < VIN_CNT[j] ? log_or_ninf(__ldg(A + (size_t)m * Q * Q + VIN_SRC[j][k] * Q + j)) : -INFINITY;
  const float* Eseq = E + (size_t)n * T * Q;
  float d[Q];
  int t0;
  if (c == 0) {
#pragma unroll
    for (int j = 0; j < Q; ++j) d[j] = logf(fmaxf(__ldg(pi + m * Q + j), kEps)) + __ldg(Eseq + j);
    t0 = 1;
  } else {
    const int ii = min(i, Q - 1);   // lane 15 mirrors lane 14 (its row is not stored)
#pragma unroll
    for (int j = 0; j < Q; ++j) d[j] = (j == ii) ? 0.f : -INFINITY;
    t0 = c * L;
  }
  const int t1 = min(T, (c + 1) * L);
  double off = 0.0;
  auto renorm = [&]() {
    float mx = d[0];
#pragma unroll
    for (int j = 1; j < Q; ++j) mx = fmaxf(mx, d[j]);
    mx = half_max(mx, hmask);
    if (mx > -INFINITY) {
#pragma unroll
      for (int j = 0; j < Q; ++j) d[j] -= mx;
      off += (double)mx;
    }
  };
  float e[Q], en[Q];
  if (t0 < t1) {
#pragma unroll
    for (int j = 0; j < Q; ++j) e[j] = __ldg(Eseq + (size_t)t0 * Q + j);
  }
  for (int t = t0; t < t1; ++t) {
    const int tn = min(t + 1, t1 - 1);
#pragma unroll
    for (int j = 0; j < Q; ++j) en[j] = __ldg(Eseq + (size_t)tn * Q + j);
    float nd[Q];
#pragma unroll
    for (int j = 0; j < Q; ++j) {
      float b = d[VIN_SRC[j][0]] + la[j][0];
      if (VIN_CNT[j] > 1) b = fmaxf(b, d[VIN_SRC[j][1]] + la[j][1]);
      if (VIN_CNT[j] > 2) b = fmaxf(b, d[VIN_SRC[j][2]] + la[j][2]);
      nd[j] = b + e[j];
    }
#pragma unroll
    for (int j = 0; j < Q; ++j) { d[j] = nd[j]; e[j] = en[j]; }
    if (((t - t0) & 15) == 15) renorm();
  }
  renorm();
  if (valid && i < Q) {
    float* row = Mv + (((size_t)n * VC + c) * 16 + i) * 16;
#pragma unroll
    for (int j = 0; j < Q; ++j) row[j] = d[j];
    row[15] = -INFINITY;
  }
  if (valid && i == 0) offs[(size_t)n * VC + c] = off;
}

// ---- scan, in-chunk recursion, backtracking ----
struct VSmemHead {
  float vin[VC][16];            // delta vector entering every chunk (a constant per chunk taken out)
  unsigned char maps[VC][16];   // state at the chunk's last position -> state at the position in front of the chunk
  int endst[VC];                // state at every chunk's last position on the best path
  unsigned char insrc[Q][4];    // VIN_SRC for the replay's data-dependent look-ups
};
__host__ __device__ inline int viterbi_lp(int L) { return L | 1; }   // odd pitch: the replay's 64-bit reads hit 32 different banks
__host__ __device__ inline size_t viterbi_smem_bytes(int T, int L) {
  return ((sizeof(VSmemHead) + 15) / 16) * 16 + (size_t)(VC / 2) * viterbi_lp(L) * sizeof(uint2) + (((size_t)T + 15) / 16) * 16;
}

__global__ void __launch_bounds__(VCH_THREADS, 1) viterbi_chunks_kernel(const float* __restrict__ E, const float* __restrict__ A,
                                                                        const float* __restrict__ pi, int batch, int T, int L, int C,
                                                                        const float* __restrict__ Mv, const double* __restrict__ offs,
                                                                        unsigned char* __restrict__ path, float* __restrict__ score) {
  extern __shared__ __align__(16) unsigned char vsm[];
  VSmemHead& S = *reinterpret_cast<VSmemHead*>(vsm);
  const int Lp = viterbi_lp(L);
  uint2* psi = reinterpret_cast<uint2*>(vsm + ((sizeof(VSmemHead) + 15) / 16) * 16);   // [pair of chunks][step]: bit masks of the 2 x 15 arg bits
  unsigned char* obuf = reinterpret_cast<unsigned char*>(psi + (size_t)(VC / 2) * Lp);
  const int n = blockIdx.x, m = n / batch, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* Eseq = E + (size_t)n * T * Q;
  if (tid < Q * 4) S.insrc[tid >> 2][tid & 3] = (unsigned char)((tid & 3) < 3 ? VIN_SRC[tid >> 2][tid & 3] : 0);

  // (1) fold the chunk matrices: v enters chunk c; lane = state j
  if (warp == 0) {
    const int j = min(lane, 15);
    const float* M0 = Mv + (size_t)n * VC * 256;
    float v = (lane < Q) ? __ldg(M0 + j) : -INFINITY;   // row 0 of chunk 0
    double offset = __ldg(offs + (size_t)n * VC);
    float mc[Q], mn[Q];
    if (C > 1) {
#pragma unroll
      for (int i = 0; i < Q; ++i) mc[i] = __ldg(M0 + (size_t)1 * 256 + i * 16 + j);
    }
    for (int c = 1; c < C; ++c) {
      if (lane < 16) S.vin[c][lane] = v;
      const int cn = min(c + 1, C - 1);
#pragma unroll
      for (int i = 0; i < Q; ++i) mn[i] = __ldg(M0 + (size_t)cn * 256 + i * 16 + j);
      float b0 = -INFINITY, b1 = -INFINITY, b2 = -INFINITY;
#pragma unroll
      for (int i = 0; i < Q; ++i) {
        const float cand = __shfl_sync(0xffffffffu, v, i) + mc[i];
        if (i % 3 == 0) b0 = fmaxf(b0, cand); else if (i % 3 == 1) b1 = fmaxf(b1, cand); else b2 = fmaxf(b2, cand);
      }
      v = (lane < Q) ? fmaxf(fmaxf(b0, b1), b2) : -INFINITY;
      offset += __ldg(offs + (size_t)n * VC + c);
      float mx = v;
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      if (mx > -INFINITY) { v -= mx; offset += (double)mx; }
#pragma unroll
      for (int i = 0; i < Q; ++i) mc[i] = mn[i];
    }
    float mx = v;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0 && score) score[n] = (float)(offset + (double)mx);
  }
  __syncthreads();

  // (2) the recursion inside every chunk: half warp = chunk, lane = state
  {
    const int c = warp * 2 + (lane >> 4), j = lane & 15;
    const bool chunk_ok = c < C, live = chunk_ok && j < Q;
    const int jj = min(j, Q - 1);
    const int s0 = VIN_SRC[jj][0], s1 = VIN_SRC[jj][1], s2 = VIN_SRC[jj][2], cnt = VIN_CNT[jj];
    const float* Am = A + (size_t)m * Q * Q;
    const float la0 = live ? log_or_ninf(__ldg(Am + s0 * Q + jj)) : -INFINITY;
    const float la1 = (live && cnt > 1) ? log_or_ninf(__ldg(Am + s1 * Q + jj)) : -INFINITY;
    const float la2 = (live && cnt > 2) ? log_or_ninf(__ldg(Am + s2 * Q + jj)) : -INFINITY;
    float d = -INFINITY;
    if (live) d = (c == 0) ? logf(fmaxf(__ldg(pi + m * Q + jj), kEps)) + __ldg(Eseq + jj) : S.vin[c][j];
    int org = j;
    const int tb = min(c, C - 1) * L;                 // position of step 0 of this chunk
    // step s works on position tb + s; chunk 0 has no step 0 (position 0 is the initial vector)
    auto act = [&](int s) { const int t = tb + s; return chunk_ok && t >= 1 && t < T && s < L; };
    auto eload = [&](int s) { const int t = tb + s; return (live && t < T && s < L) ? __ldg(Eseq + (size_t)t * Q + jj) : 0.f; };
    float ea[VPF], eb[VPF];
#pragma unroll
    for (int u = 0; u < VPF; ++u) ea[u] = eload(u);
    auto block = [&](const float (&use)[VPF], float (&fill)[VPF], int sb) {
#pragma unroll
      for (int u = 0; u < VPF; ++u) fill[u] = eload(sb + VPF + u);
#pragma unroll
      for (int u = 0; u < VPF; ++u) {
        const int s = sb + u;
        const float c0 = __shfl_sync(0xffffffffu, d, s0, 16) + la0;
        const float c1 = __shfl_sync(0xffffffffu, d, s1, 16) + la1;
        const float c2 = __shfl_sync(0xffffffffu, d, s2, 16) + la2;
        const float best = fmaxf(fmaxf(c0, c1), c2);
        const int arg = (c0 == best) ? 0 : (c1 == best) ? 1 : 2;
        const int srcl = arg == 0 ? s0 : (arg == 1 ? s1 : s2);
        const int o2 = __shfl_sync(0xffffffffu, org, srcl, 16);
        const bool on = act(s);
        const unsigned m0 = __ballot_sync(0xffffffffu, on && live && (arg & 1));
        const unsigned m1 = __ballot_sync(0xffffffffu, on && live && (arg & 2));
        if (on) { d = best + use[u]; org = o2; }
        if (lane == 0 && s < L) psi[(size_t)warp * Lp + s] = make_uint2(m0, m1);
      }
      const float mx = half_max(d);
      if (mx > -INFINITY) d -= mx;
    };
    for (int sb = 0; sb < L; sb += 2 * VPF) {
      block(ea, eb, sb);
      block(eb, ea, sb + VPF);
    }
    if (live) S.maps[c][j] = (unsigned char)org;
    // best final state (lowest index on ties): the last chunk's half warp
    float bv = live ? d : -INFINITY;
    int bs = j;
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o, 16);
      const int os = __shfl_xor_sync(0xffffffffu, bs, o, 16);
      if (ov > bv || (ov == bv && os < bs)) { bv = ov; bs = os; }
    }
    if (c == C - 1 && j == 0) S.endst[C - 1] = bs;
  }
  __syncthreads();
  // (3) the state at every chunk's last position, from the last chunk backwards
  if (tid == 0) {
    int s = S.endst[C - 1];
    for (int c = C - 1; c >= 1; --c) { s = S.maps[c][s]; S.endst[c - 1] = s; }
  }
  __syncthreads();
  // (4) thread = chunk: back along the chunk's pointers
  if (tid < C) {
    const int c = tid, tb = c * L, sh = (c & 1) * 16;
    const uint2* pp = psi + (size_t)(c >> 1) * Lp;
    int s = S.endst[c];
    const int s_hi = min(L, T - tb) - 1, s_lo = (c == 0) ? 1 : 0;
    for (int k = s_hi; k >= s_lo; --k) {
      obuf[tb + k] = (unsigned char)s;
      const uint2 w = pp[k];
      const int arg = (int)((w.x >> (sh + s)) & 1u) | ((int)((w.y >> (sh + s)) & 1u) << 1);
      s = S.insrc[s][arg];
    }
    if (c == 0) obuf[0] = (unsigned char)s;
  }
  __syncthreads();
  unsigned char* out = path + (size_t)n * T;
  for (int t = tid; t < T; t += VCH_THREADS) out[t] = obuf[t];
}

// ---- host side ----
enum { SLOT_VITERBI = 26 };
bool viterbi_scan_applies(int T) {
  static const bool off = getenv("HMM_B200_VITERBI_SCAN") && atoi(getenv("HMM_B200_VITERBI_SCAN")) == 0;
  return !off && T >= 1024 && T <= 32768;
}
size_t viterbi_scan_scratch_bytes(int NB) { return (size_t)NB * VC * 256 * sizeof(float) + (size_t)NB * VC * sizeof(double); }

int launch_viterbi_scan(const float* E, const float* A, const float* pi, int NB, int batch, int T, void* scratch,
                        unsigned char* path, float* score, cudaStream_t st) {
  const int L = cdiv(T, VC), C = cdiv(T, L);
  float* Mv = static_cast<float*>(scratch);
  double* offs = reinterpret_cast<double*>(static_cast<unsigned char*>(scratch) + (size_t)NB * VC * 256 * sizeof(float));
  const size_t smem = viterbi_smem_bytes(T, L);
  int rc = ensure_dynamic_smem(viterbi_chunks_kernel, SLOT_VITERBI, smem);
  if (rc) return rc;
  viterbi_rows_kernel<<<dim3(cdiv(C, VROW_THREADS / 16), NB), VROW_THREADS, 0, st>>>(E, A, pi, batch, T, L, C, Mv, offs);
  viterbi_chunks_kernel<<<NB, VCH_THREADS, smem, st>>>(E, A, pi, batch, T, L, C, Mv, offs, path, score);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace gene15
}  // namespace hmm

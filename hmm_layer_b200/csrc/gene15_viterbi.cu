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
#include <cooperative_groups.h>
#include "gene15_dev.cuh"

namespace hmm {
namespace gene15 {

constexpr int VC = 64;              // chunks per sequence (at most)
constexpr int VROW_THREADS = 256;   // rows kernel: 16 chunks x 16 lanes
constexpr int VCL = 4;              // CTAs (one cluster) per sequence in the chunks kernel
constexpr int VCPC = VC / VCL;      // chunks per CTA
constexpr int VCH_THREADS = VCPC * 16;   // chunks kernel: half warp = chunk
constexpr int VPF = 8;              // emission rows requested ahead in the in-chunk recursion
constexpr int VRING = 8;            // ... and in the rows kernel

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
  const int t1 = min(T, (c + 1) * L), tb = c * L;
  double off = 0.0;
  // Every warp walks the same L step slots (step slot s = position tb + s), so that the two chunks of a warp stay in one
  // instruction stream and the shuffles are plain full-mask ones; a slot outside the chunk's positions (slot 0 of chunk 0,
  // the slots behind a short last chunk) computes and does not commit.
  auto renorm = [&]() {
    float mx = d[0];
#pragma unroll
    for (int j = 1; j < Q; ++j) mx = fmaxf(mx, d[j]);
    mx = half_max(mx);
    if (mx > -INFINITY) {
#pragma unroll
      for (int j = 0; j < Q; ++j) d[j] -= mx;
      off += (double)mx;
    }
  };
  // emissions: lane j of the half warp fetches E_t(j) (one coalesced 60-byte row per step) VRING steps ahead of its use, the
  // 15 values of a step come by shuffles: one load per step and thread instead of fifteen, and the latency is covered
  const int jl = min(i, Q - 1);
  auto eload = [&](int sl) { const int t = tb + sl; return (sl < L && t < T) ? __ldg(Eseq + (size_t)t * Q + jl) : 0.f; };
  float ring[VRING];
#pragma unroll
  for (int u = 0; u < VRING; ++u) ring[u] = eload(u);
  for (int sb = 0; sb < L; sb += VRING) {
    const bool mine = tb + sb >= t0 && tb + sb + VRING <= t1;
    const bool full = __all_sync(0xffffffffu, mine);   // warp-uniform: every slot of the block is a step of both chunks
#pragma unroll
    for (int u = 0; u < VRING; ++u) {
      const int t = tb + sb + u;
      float nd[Q];
#pragma unroll
      for (int j = 0; j < Q; ++j) {
        float b = d[VIN_SRC[j][0]] + la[j][0];
        if (VIN_CNT[j] > 1) b = fmaxf(b, d[VIN_SRC[j][1]] + la[j][1]);
        if (VIN_CNT[j] > 2) b = fmaxf(b, d[VIN_SRC[j][2]] + la[j][2]);
        nd[j] = b + __shfl_sync(0xffffffffu, ring[u], j, 16);
      }
      if (full) {
#pragma unroll
        for (int j = 0; j < Q; ++j) d[j] = nd[j];
      } else {
        const bool on = t >= t0 && t < t1;
#pragma unroll
        for (int j = 0; j < Q; ++j) d[j] = on ? nd[j] : d[j];
      }
      ring[u] = eload(sb + u + VRING);
    }
    renorm();   // every VRING steps
  }
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
// the back-pointer region first holds the sequence's chunk matrices (64 KB, phase 1), so it is at least that large
__host__ __device__ inline size_t viterbi_psi_bytes(int L) {
  const size_t p = (size_t)(VCPC / 2) * viterbi_lp(L) * sizeof(uint2), m = (size_t)VC * 256 * sizeof(float);
  return ((p > m ? p : m) + 15) / 16 * 16;
}
__host__ __device__ inline size_t viterbi_smem_bytes(int T, int L) {
  (void)T;
  return ((sizeof(VSmemHead) + 15) / 16) * 16 + viterbi_psi_bytes(L) + (((size_t)VCPC * L + 15) / 16) * 16;
}

__global__ void __launch_bounds__(VCH_THREADS, 1) viterbi_chunks_kernel(const float* __restrict__ E, const float* __restrict__ A,
                                                                        const float* __restrict__ pi, int batch, int T, int L, int C,
                                                                        const float* __restrict__ Mv, const double* __restrict__ offs,
                                                                        unsigned char* __restrict__ path, float* __restrict__ score) {
  extern __shared__ __align__(16) unsigned char vsm[];
  VSmemHead& S = *reinterpret_cast<VSmemHead*>(vsm);
  const int Lp = viterbi_lp(L);
  uint2* psi = reinterpret_cast<uint2*>(vsm + ((sizeof(VSmemHead) + 15) / 16) * 16);   // [pair of chunks][step]: bit masks of the 2 x 15 arg bits
  unsigned char* obuf = reinterpret_cast<unsigned char*>(psi) + viterbi_psi_bytes(L);
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = blockIdx.x;            // CTA `rank` of the sequence's cluster owns chunks [rank * VCPC, (rank + 1) * VCPC)
  const int n = blockIdx.y, m = n / batch, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float* Eseq = E + (size_t)n * T * Q;
  if (tid < Q * 4) S.insrc[tid >> 2][tid & 3] = (unsigned char)((tid & 3) < 3 ? VIN_SRC[tid >> 2][tid & 3] : 0);
#ifdef HMM_VIT_MARKS
  long long mk[8];
  int nmk = 0;
#define VMARK() do { if (tid == 0) mk[nmk++] = clock_fenced(); } while (0)
#else
#define VMARK() do { } while (0)
#endif
  VMARK();

  // (1) fold the chunk matrices: v enters chunk c; lane = state j.  The matrices are staged in shared memory first (all
  // loads in flight at once; the region becomes the back-pointer table afterwards)
  {
    float4* dst = reinterpret_cast<float4*>(psi);
    const float4* src = reinterpret_cast<const float4*>(Mv + (size_t)n * VC * 256);
    for (int k = tid; k < C * 64; k += VCH_THREADS) dst[k] = __ldg(src + k);
  }
  __syncthreads();
  VMARK();
  if (warp == 0) {
    const int j = lane & 15;
    const float* M0 = reinterpret_cast<const float*>(psi);
    float v = ((lane & 15) < Q) ? M0[lane & 15] : -INFINITY;   // row 0 of chunk 0
    double offset = 0.0;
    for (int c = lane; c < C; c += 32) offset += __ldg(offs + (size_t)n * VC + c);   // the constants taken out of the chunks' rows
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) offset += __shfl_xor_sync(0xffffffffu, offset, o);
    // both half warps hold v (lane j and lane 16 + j the same value); the lower half takes source states 0-7, the upper 8-14
    const int hi = lane >> 4, i0 = hi * 8;
    for (int c = 1; c < C; ++c) {
      if (lane < 16) S.vin[c][lane] = v;
      const float* Mc = M0 + (size_t)c * 256 + i0 * 16 + j;
      float b0 = -INFINITY, b1 = -INFINITY;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float vi = __shfl_sync(0xffffffffu, v, i0 + k);      // (i0 + k == 15 in the upper half: the -inf of lane 15)
        const float cand = vi + Mc[min(k, 14 - i0) * 16];
        if (k & 1) b1 = fmaxf(b1, cand); else b0 = fmaxf(b0, cand);
      }
      float b = fmaxf(b0, b1);
      b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, 16));
      v = (j < Q) ? b : -INFINITY;
      if ((c & 3) == 3 || c == C - 1) {   // take the maximum out now and then (the magnitudes grow by a chunk's spread at most)
        float mx = v;
#pragma unroll
        for (int o = 8; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        if (mx > -INFINITY) { v -= mx; offset += (double)mx; }
      }
    }
    float mx = v;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if (lane == 0 && score && rank == 0) score[n] = (float)(offset + (double)mx);
  }
  __syncthreads();
  VMARK();

  // (2) the recursion inside every chunk: half warp = chunk, lane = state
  {
    const int c = rank * VCPC + warp * 2 + (lane >> 4), j = lane & 15;
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
    // one block of VPF steps.  `full`: every step of the block is a real step of both chunks of the warp (all but the
    // first block of chunk 0 and the blocks at a chunk's / the sequence's end) -- no activity predicates then.  Dead lanes
    // (state 15, chunks behind the last) carry -inf on every in-edge: their comparisons pick source 0 and set no bits.
    auto block = [&](const float (&use)[VPF], float (&fill)[VPF], int sb, bool full) {
#pragma unroll
      for (int u = 0; u < VPF; ++u) fill[u] = eload(sb + VPF + u);
#pragma unroll
      for (int u = 0; u < VPF; ++u) {
        const int s = sb + u;
        const float c0 = __shfl_sync(0xffffffffu, d, s0, 16) + la0;
        const float c1 = __shfl_sync(0xffffffffu, d, s1, 16) + la1;
        const float c2 = __shfl_sync(0xffffffffu, d, s2, 16) + la2;
        const float best = fmaxf(fmaxf(c0, c1), c2);
        const bool p0 = c0 == best, p1 = c1 == best;
        const int srcl = p0 ? s0 : (p1 ? s1 : s2);
        const int o2 = __shfl_sync(0xffffffffu, org, srcl, 16);
        if (full) {
          const unsigned m0 = __ballot_sync(0xffffffffu, !p0 && p1);
          const unsigned m1 = __ballot_sync(0xffffffffu, !p0 && !p1);
          d = best + use[u];
          org = o2;
          if (lane == 0) psi[(size_t)warp * Lp + s] = make_uint2(m0, m1);
        } else {
          const bool on = act(s);
          const unsigned m0 = __ballot_sync(0xffffffffu, on && !p0 && p1);
          const unsigned m1 = __ballot_sync(0xffffffffu, on && !p0 && !p1);
          if (on) { d = best + use[u]; org = o2; }
          if (lane == 0 && s < L) psi[(size_t)warp * Lp + s] = make_uint2(m0, m1);
        }
      }
      const float mx = half_max(d);
      if (mx > -INFINITY) d -= mx;
    };
    auto is_full = [&](int sb) {   // warp-uniform
      const bool mine = !chunk_ok || (tb + sb >= 1 && tb + sb + VPF <= T && sb + VPF <= L);
      return __all_sync(0xffffffffu, mine);
    };
    for (int sb = 0; sb < L; sb += 2 * VPF) {
      block(ea, eb, sb, is_full(sb));
      block(eb, ea, sb + VPF, is_full(sb + VPF));
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
  // the maps of the other CTAs' chunks and the final state come over distributed shared memory
  cluster.sync();
  for (int k = tid; k < C * 4; k += VCH_THREADS) {
    const int c = k >> 2, r = c / VCPC;
    if (r != rank) {
      const unsigned* remote = reinterpret_cast<const unsigned*>(cluster.map_shared_rank(&S.maps[0][0], r));
      reinterpret_cast<unsigned*>(&S.maps[0][0])[k] = remote[k];
    }
  }
  if (tid == 0 && (C - 1) / VCPC != rank) S.endst[C - 1] = *cluster.map_shared_rank(&S.endst[C - 1], (C - 1) / VCPC);
  cluster.sync();   // (nobody reads this CTA's shared memory any more: it may run ahead and exit)
  VMARK();
  // (3) the state at every chunk's last position, from the last chunk backwards
  if (tid == 0) {
    int s = S.endst[C - 1];
    for (int c = C - 1; c >= 1; --c) { s = S.maps[c][s]; S.endst[c - 1] = s; }
  }
  __syncthreads();
  VMARK();
  // (4) thread = chunk: back along the chunk's pointers
  const int pos0 = rank * VCPC * L;       // first position of this CTA
  if (tid < VCPC && rank * VCPC + tid < C) {
    const int lc = tid, c = rank * VCPC + lc, tb = c * L, sh = (lc & 1) * 16;
    const uint2* pp = psi + (size_t)(lc >> 1) * Lp;
    int s = S.endst[c];
    const int s_hi = min(L, T - tb) - 1, s_lo = (c == 0) ? 1 : 0;
    for (int k = s_hi; k >= s_lo; --k) {
      obuf[tb - pos0 + k] = (unsigned char)s;
      const uint2 w = pp[k];
      const int arg = (int)((w.x >> (sh + s)) & 1u) | ((int)((w.y >> (sh + s)) & 1u) << 1);
      s = S.insrc[s][arg];
    }
    if (c == 0) obuf[0] = (unsigned char)s;
  }
  __syncthreads();
  VMARK();
  unsigned char* out = path + (size_t)n * T + pos0;
  const int cnt = max(0, min(T - pos0, VCPC * L));
  for (int t = tid; t < cnt; t += VCH_THREADS) out[t] = obuf[t];
#ifdef HMM_VIT_MARKS
  VMARK();
  if (tid == 0 && n == 0)
    printf("viterbi_chunks_kernel cycles: stage %lld, fold %lld, in-chunk %lld, chain %lld, replay %lld, out %lld\n", mk[1] - mk[0],
           mk[2] - mk[1], mk[3] - mk[2], mk[4] - mk[3], mk[5] - mk[4], mk[6] - mk[5]);
#endif
#undef VMARK
}

// ---- host side ----
enum { SLOT_VITERBI = 26 };
bool viterbi_scan_applies(int T) {
  static const bool off = getenv("HMM_B200_VITERBI_SCAN") && atoi(getenv("HMM_B200_VITERBI_SCAN")) == 0;
  return !off && T >= 1024 && T <= 131072;
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
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(VCL, NB); cfg.blockDim = dim3(VCH_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = VCL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    HMM_CUDA_OK(cudaLaunchKernelEx(&cfg, viterbi_chunks_kernel, E, A, pi, batch, T, L, C, (const float*)Mv, (const double*)offs, path, score));
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace gene15
}  // namespace hmm

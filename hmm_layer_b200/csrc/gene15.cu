// 15-state gene-structure HMM: fused emission + chunked scaled forward / backward / posterior for sm_100a.
//
// Two kernels per call (plus a boundary-scan kernel for very long sequences / sequence sharding), see DESIGN.md §3:
//
//   products_kernel  K1  one CTA per tile = G(32) chunks x L positions of one sequence.  The tile's input rows arrive
//                        in shared memory by one bulk asynchronous copy (cp.async.bulk + mbarrier).  Emission
//                        (computed once; the floored rows go to an L2-resident workspace for K2), then the chunk
//                        transfer matrices prod_t (A diag(e_t)) as 15 conditional rows (the reference's
//                        parallel_factor>1 form, MsaHmmCell.py:121-142, keeping only the chunk-end matrix).  The 8
//                        single-in/single-out states (START, EI*, IE*, STOP) are eliminated algebraically, so a row
//                        step is a 15-FMA two-step recurrence over the 7 core states instead of a 23-edge product;
//                        scaling is lazy and by exact powers of two.  A 5-level in-CTA product tree over the 32 chunk
//                        matrices is kept whole (heap order) and leaves by one bulk store: its root is the tile
//                        matrix, its inner nodes let K2 reach every chunk boundary in 5 parallel steps.  The last CTA
//                        of a sequence to finish scans the sequence's tile matrices (the role of
//                        TotalProbabilityCell, TotalProbabilityCell.py:30-49, in scaled probability space): boundary
//                        vectors entering / leaving every tile and the log-likelihood.
//   boundary_kernel  K_B only when a sequence has more tiles than the in-K1 scan stages, or when the sequence is
//                        sharded over GPUs: the same prefix / suffix scan as a kernel of its own.
//   finalize_kernel  K2  one CTA per tile: emission rows and the tile's tree arrive by bulk copies, a 5-step
//                        down-sweep of the tree gives every chunk's boundary vectors, then the exact recursions of
//                        MsaHmmCell.py:73-106 (same floors, same order) inside every chunk, both directions
//                        concurrently meeting in the middle, then log posterior / log alpha / log beta written
//                        once, coalesced.
//
// State order (gene_pred_hmm_transitioner.py:187-188): 0 Ir | 1-3 I0-2 | 4-6 E0-2 | 7 START | 8-10 EI0-2 |
// 11-13 IE0-2 | 14 STOP.  Only the 23 structural edges of A are read (gene_pred_hmm_transitioner.py:200-221).
// "core" = states 0..6, "pass" = states 7..14 (one predecessor, one successor, both core).
//
// Stored matrix block (16x16 floats): rows 0-14 = [row-normalised transfer row M-hat | relative row scale s'],
// row 15 = [column sums of diag(s') M-hat | block exponent E as int bits]; true matrix = 2^E diag(s') M-hat.
// All scales are exact powers of two (integer bookkeeping), so no log/exp rounding enters the posteriors.
//
// The numpy model tests/chunk_algo_model.py mirrors the arithmetic and is checked against the oracle.
#include "gene15_dev.cuh"

namespace hmm {
namespace gene15 {

// ------------------------------------------------------------------------------------------------------------
// K1: chunk transfer matrices, group and tile products
// ------------------------------------------------------------------------------------------------------------
struct K1Smem {
  EmisSmem sp;
  float Aw[24];        // [0..6] A[src1(j)][j], [7..14] A[pred p][p], [15..22] A[p][succ p]
  float EF[G][QP];     // emission at every chunk's first transition step
  float EL[G][QP];     // emission at every chunk's last position
  float carry[8];      // pass-state emissions of the last position of the previous emission round
  unsigned long long bar;   // mbarrier: the tile's input rows have landed
  unsigned last;       // this CTA finished the last tile of its sequence
  float ones[QP];
};

// shared memory after K1Smem: the staged input rows ((G*L + 4) * 20 floats) and the coefficient rows (G*CS floats) share
// ONE region: an emission round reads its input rows into registers, the CTA meets at a barrier, and the round's
// coefficient rows (16 floats at a stride of 20, like the input rows) overwrite rows that have all been read by then --
// coefficient row i sits on the input row of position i or i - 2.  Reused as the product tree (HEAP blocks) once the conditional rows are done, then as the staged
// tile matrices of the top scan.  At chunk length 35 this is 90 KB instead of 162 KB: two CTAs per SM.
__host__ __device__ inline size_t k1_coef_bytes(int L) { return (size_t)G * L * 20 * sizeof(float); }
__host__ __device__ inline size_t k1_in_bytes(int L) { return ((size_t)G * L + 4) * SIN * sizeof(float); }
__host__ __device__ inline size_t k1_union_bytes(int L, int P) {
  size_t b = k1_coef_bytes(L) > k1_in_bytes(L) ? k1_coef_bytes(L) : k1_in_bytes(L);
  const size_t tree = (size_t)HEAP * MAT * sizeof(float);
  if (tree > b) b = tree;
  const size_t scan = (size_t)(P <= MAX_SCAN_TILES ? P : 0) * MAT * sizeof(float);
  return scan > b ? scan : b;
}
__host__ __device__ inline size_t k1_smem_bytes(int L, int P) {
  return ((sizeof(K1Smem) + 15) / 16) * 16 + k1_union_bytes(L, P) + (((size_t)G * L + 4 + 15) / 16) * 16;
}

// BLOCKS = CTAs per SM the register budget is sized for: 2 (64 registers) when two tiles fit an SM's shared memory, 1 (128
// registers, no spills) for the long chunks that leave room for one
template <int BLOCKS>
__global__ void __launch_bounds__(K1_THREADS, BLOCKS) products_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem[];
  K1Smem& S = *reinterpret_cast<K1Smem*>(smem);
  float* U = reinterpret_cast<float*>(smem + ((sizeof(K1Smem) + 15) / 16) * 16);
  signed char* codes = reinterpret_cast<signed char*>(U) + k1_union_bytes(a.L, a.P);
  const int L = a.L, T = a.T, CS = L * QP + 4;
  float* IN = U;   // staged input rows: the same region as the coefficient rows (see k1_union_bytes)
  const int tid = threadIdx.x;
  const int p = blockIdx.x, n = blockIdx.y, m = n / a.batch;
  const int c0 = p * G, t_lo = c0 * L, npos = min(T, t_lo + G * L) - t_lo;
  const bool seq_start = (a.seq_off == 0);  // chunk 0 of this shard is the start of the whole sequence
  const float* Am = a.A + (size_t)m * Q * Q;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  const size_t tile = (size_t)n * a.P + p;
  HMM_MARK(cta, 0);
  HMM_WALL(cta, 14);
  if (tid == 0) griddep_launch_dependents();   // every CTA of this grid is resident or done when the last one says so

  // ---- parameters first (their loads would otherwise queue behind this CTA's own bulk copy), then the tile's input
  // rows (with the two neighbours either side the 3-mer windows reach) by one bulk copy ----
  const int t_in0 = max(t_lo - 2, 0), t_in1 = min(t_lo + npos + 2, T), n_in = t_in1 - t_in0;
  {
    float aw = 0.f;
    if (tid < 7) aw = __ldg(Am + SRC1[tid] * Q + tid);
    else if (tid < 15) aw = __ldg(Am + PRED[tid - 7] * Q + tid);
    else if (tid < 23) aw = __ldg(Am + (tid - 8) * Q + SUCC[tid - 15]);
    // B rows and the 5-symbol codon tables: at most two 16-byte pieces and one B entry per thread, all in flight at once
    const float4* tsrc = reinterpret_cast<const float4*>(a.codon5);
    float4 tv0 = make_float4(0.f, 0.f, 0.f, 0.f), tv1 = tv0;
    if (tid < T5_FLOATS / 4) tv0 = __ldg(tsrc + tid);
    if (tid + K1_THREADS < T5_FLOATS / 4) tv1 = __ldg(tsrc + tid + K1_THREADS);
    const int bq = tid >> 4, bs = tid & 15;
    float bv = 0.f;
    if (tid < Q * QP && bs < Q) bv = __ldg(a.Bfull + (size_t)m * Q * Q + bq * Q + bs);
    if (tid == 0) {
      mbar_init(&S.bar, 1);
      mbar_init_fence();
      if (!a.cls16) {
        const unsigned bytes = (unsigned)n_in * SIN * sizeof(float);
        mbar_expect_tx(&S.bar, bytes);
        bulk_g2s(IN, a.inputs + ((size_t)n * T + t_in0) * SIN, bytes, &S.bar);
      }
    }
    HMM_MARK(cta, 9);
    if (tid < 23) S.Aw[tid] = aw;
    if (tid >= 32 && tid < 32 + QP) S.ones[tid - 32] = 1.f;
    HMM_MARK(cta, 10);
    float4* tdst = reinterpret_cast<float4*>(&S.sp.T5[0][0][0]);
    if (tid < T5_FLOATS / 4) tdst[tid] = tv0;
    if (tid + K1_THREADS < T5_FLOATS / 4) tdst[tid + K1_THREADS] = tv1;
    if (tid < Q * QP) S.sp.B[bq][bs] = bv;
    HMM_MARK(cta, 11);
  }
  __syncthreads();            // mbarrier initialised, parameters in place
  HMM_MARK(cta, 7);
  if (a.cls16) {
    // compact wire format: bf16 class probabilities -> the fp32 rows the emission reads, nucleotide codes as they are
    const unsigned short* src = a.cls16 + ((size_t)n * T + t_in0) * Q;
    for (int i = tid; i < n_in * Q; i += K1_THREADS) {
      const int pos = i / Q, c = i - pos * Q;
      IN[pos * SIN + c] = __uint_as_float((unsigned)__ldg(src + i) << 16);
    }
    for (int i = tid; i < n_in; i += K1_THREADS) IN[i * SIN + Q] = 0.f;   // slot 15 meets B[q][15] = 0: must not be NaN
    const unsigned char* nc = a.nuc8 + (size_t)n * T;
    for (int i = tid; i < npos + 4; i += K1_THREADS) {
      const int t = t_lo - 2 + i;
      codes[i] = (signed char)((t >= 0 && t < T) ? min((int)__ldg(nc + t), 4) : 4);   // outside the sequence: 1/4 each
    }
  } else {
    mbar_wait(&S.bar, 0);
    HMM_MARK(cta, 8);
    codes_from_smem(a, codes, IN, n, t_lo, npos, t_in0, n_in);
  }
  __syncthreads();
  HMM_MARK(cta, 1);

  // ---- emission + per-position coefficients of the two-step recurrence ----
  // slots 0-6: k1[j] = E_t[j] A[src1(j)][j];  slots 7-14: k2[p] = E_t[succ p] A[p][succ p] * E_{t-1}[p] A[pred p][p].
  // Every chunk's first transition step uses the full 23-edge form, so k2 is never read at a chunk's first
  // position and no emission from outside the tile is needed.  E_{t-1}[pass] travels through slots 8-15 of row t-1.
  // The floored emission row itself goes to the workspace in K2's shared-memory layout.
  float* Eg = a.Ews ? a.Ews + tile * (size_t)G * CS : nullptr;
  for (int ib = 0; ib < npos; ib += K1_THREADS) {   // uniform trip count: the body holds barriers
    const int i0 = ib + tid;
    const bool one = i0 < npos;
    float e0[Q];
    float* r0 = U;
    int ci0 = 0;
    float4 xin[1][4];
    if (one) {
      const float4* row = reinterpret_cast<const float4*>(IN + (size_t)(t_lo + i0 - t_in0) * SIN);
      xin[0][0] = row[0]; xin[0][1] = row[1]; xin[0][2] = row[2]; xin[0][3] = row[3];
    }
    __syncthreads();   // every input row of this round is in registers: its coefficient rows may land on top of them
    if (one) {
      const int t[1] = {t_lo + i0};
      const signed char* const cd[1] = {codes + 2 + i0};
      float e[1][Q];
      emit_x<true, 1>(a, S.sp, n, t, cd, xin, e);
#pragma unroll
      for (int q = 0; q < Q; ++q) e0[q] = e[0][q];
      ci0 = divL(a, i0);
      const int roff = ci0 * CS + (i0 - ci0 * L) * QP;   // the workspace keeps K2's layout
      r0 = U + i0 * KST;
      reinterpret_cast<float4*>(r0)[2] = make_float4(e0[7], e0[8], e0[9], e0[10]);
      reinterpret_cast<float4*>(r0)[3] = make_float4(e0[11], e0[12], e0[13], e0[14]);
      if (Eg) store_row16(Eg + roff, e0, 0.f);
    }
    __syncthreads();
    float ep0[8];
    if (one) {
      const float* q0 = (i0 == 0) ? r0 + 8 : (tid == 0) ? S.carry : r0 - KST + 8;
      const float4 u0 = reinterpret_cast<const float4*>(q0)[0], u1 = reinterpret_cast<const float4*>(q0)[1];
      ep0[0] = u0.x; ep0[1] = u0.y; ep0[2] = u0.z; ep0[3] = u0.w; ep0[4] = u1.x; ep0[5] = u1.y; ep0[6] = u1.z; ep0[7] = u1.w;
    }
    __syncthreads();
    if (one) {
      const int t = t_lo + i0, c = c0 + ci0;
      const int ts = c * L + ((c == 0 && seq_start) ? 1 : 0);
      const int t1 = min(T, (c + 1) * L) - 1;
      if (t == ts) store_row16(S.EF[ci0], e0, 0.f);
      if (t == t1) store_row16(S.EL[ci0], e0, 0.f);
      if (tid == K1_THREADS - 1) {
#pragma unroll
        for (int q = 0; q < 8; ++q) S.carry[q] = e0[7 + q];
      }
      if (t == 0 && seq_start) {
        store_row16(r0, e0, 0.f);   // position 0 keeps its emission row (no transition into it)
      } else {
        float kc[Q];
#pragma unroll
        for (int j = 0; j < 7; ++j) kc[j] = e0[j] * S.Aw[j];
#pragma unroll
        for (int q = 0; q < 8; ++q) kc[7 + q] = (e0[SUCC[q]] * S.Aw[15 + q]) * (ep0[q] * S.Aw[7 + q]);
        store_row16(r0, kc, 0.f);
      }
    }
  }
  __syncthreads();
  HMM_MARK(cta, 2);

  // ---- conditional rows: thread = (chunk, row) ----
  const int ci = tid >> 4, row = tid & 15, c = c0 + ci;
  float vec[Q];
  float sm = 1.f;
  int esum = 0;
#pragma unroll
  for (int j = 0; j < Q; ++j) vec[j] = (j == row) ? 1.f : 0.f;   // chunks beyond the sequence: identity
  if (c < a.C && row < Q) {
    const bool first = (c == 0 && seq_start);
    const int t0 = c * L, t1 = min(T, t0 + L) - 1, ts = t0 + (first ? 1 : 0);
    const float* Kp = U + ci * L * KST;
    const float init = first ? Kp[row] : 1.f;   // E_0[row]
    float xm[7], ym[8];
#pragma unroll
    for (int j = 0; j < 7; ++j) xm[j] = (row == j) ? init : 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) ym[q] = (row == 7 + q) ? init : 0.f;
    if (ts > t1) {
#pragma unroll
      for (int j = 0; j < 7; ++j) vec[j] = xm[j];
#pragma unroll
      for (int q = 0; q < 8; ++q) vec[7 + q] = ym[q];
    } else {
      float x[7], xp[7];
      {
        float k1[QP];
        load_row16(Kp + (ts - t0) * KST, k1);
#pragma unroll
        for (int j = 0; j < 7; ++j) x[j] = k1[j] * xm[SRC1[j]];
#pragma unroll
        for (int q = 0; q < 8; ++q) x[SUCC[q]] = fmaf(S.EF[ci][SUCC[q]] * S.Aw[15 + q], ym[q], x[SUCC[q]]);
#pragma unroll
        for (int j = 0; j < 7; ++j) xp[j] = xm[j];
      }
      int cnt = 0;
#pragma unroll 4
      for (int t = ts + 1; t <= t1; ++t) {
        float k[QP];
        load_row16(Kp + (t - t0) * KST, k);
        float xn[7];
        xn[0] = fmaf(k[7 + 7], xp[5], k[0] * x[0]);
        xn[1] = fmaf(k[7 + 1], xp[4], k[1] * x[1]);
        xn[2] = fmaf(k[7 + 2], xp[5], k[2] * x[2]);
        xn[3] = fmaf(k[7 + 3], xp[6], k[3] * x[3]);
        xn[4] = fmaf(k[7 + 4], xp[1], k[4] * x[6]);
        xn[5] = fmaf(k[7 + 5], xp[2], fmaf(k[7 + 0], xp[0], k[5] * x[4]));
        xn[6] = fmaf(k[7 + 6], xp[3], k[6] * x[5]);
#pragma unroll
        for (int j = 0; j < 7; ++j) { xp[j] = x[j]; x[j] = xn[j]; }
        if ((++cnt & 3) == 0) {
          const float s = sum7(x) + sum7(xp);
          const int ex = pexp(s);
          const float sc = pow2neg(ex);
#pragma unroll
          for (int j = 0; j < 7; ++j) { x[j] *= sc; xp[j] *= sc; }
          esum += clamp_exp(ex);
        }
      }
#pragma unroll
      for (int j = 0; j < 7; ++j) vec[j] = x[j];
#pragma unroll
      for (int q = 0; q < 8; ++q) vec[7 + q] = (S.EL[ci][7 + q] * S.Aw[7 + q]) * xp[PRED[q]];
    }
    sm = fmaxf(sum15(vec), 1e-37f);
    const float inv = __frcp_rn(sm);
#pragma unroll
    for (int j = 0; j < Q; ++j) vec[j] *= inv;
  }
  // relative row scales: common power of two = the largest row exponent of the chunk
  const int erow = (row < Q) ? esum + fexp(sm) : -100000;
  const int emax = half_max_int(erow);
  const float sprime = (row < Q) ? sm * pow2neg(emax - esum) : 0.f;
  __syncthreads();   // every row is done with the coefficient rows: U becomes the block / tree scratch
  HMM_MARK(cta, 3);

  float* H = U;   // product tree in heap order: chunk blocks at [G, 2G), node k = node 2k . node 2k+1, root at 1
  {
    float* b = H + (G + ci) * MAT;
    if (row < Q) store_row16(b + row * QP, vec, sprime);
    __syncwarp();
    finish_block(b, row, sprime, emax);
  }
  __syncthreads();
  HMM_MARK(cta, 4);
  for (int np = G / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  HMM_MARK(cta, 5);

  // the tree (nodes 1 .. 2G-1) and the tile matrix leave by bulk stores
  fence_async_smem();
  __syncthreads();
  if (tid == 0) {
    bulk_s2g(a.Mh + (tile * HEAP + 1) * MAT, H + MAT, (HEAP - 1) * MAT * sizeof(float));
    bulk_s2g(a.Mt + tile * MAT, H + MAT, MAT * sizeof(float));
    bulk_commit();
    bulk_wait_all();
  }
  HMM_MARK(cta, 6);
  HMM_WALL(cta, 15);
  if (a.top_scan != 1) return;   // 0: boundary_kernel scans the tile matrices, 2: every K2 CTA does (scan_tiles_global)

  // ---- top-level scan by the last CTA of the sequence to get here (TotalProbabilityCell.py:30-49 in scaled
  // probability space): boundary vectors entering / leaving every tile, log-likelihood ----
  if (tid == 0) {
    __threadfence();
    const unsigned done = atomicAdd(a.cnt + n, 1u);
    S.last = (done == (unsigned)a.P - 1u) ? 1u : 0u;
    if (S.last) __threadfence();
  }
  __syncthreads();
  if (!S.last) return;
  const int P = a.P;
  {
    const float4* src = reinterpret_cast<const float4*>(a.Mt + (size_t)n * P * MAT);
    float4* dst = reinterpret_cast<float4*>(U);
    for (int i = tid; i < P * (MAT / 4); i += K1_THREADS) dst[i] = __ldcg(src + i);
  }
  __syncthreads();
  if (tid < 32) {
    const int hl = tid & 15, hi = tid >> 4;
    HV h;
    h.v = (hl < Q) ? (hi == 0 ? __ldg(a.pi + m * Q + hl) : 1.f) : 0.f;
    h.e = 0;
    float* vout = (hi == 0 ? a.bvf : a.bvr) + (size_t)n * P * QP;
    for (int it = 0; it < P; ++it) {
      const int s = hi == 0 ? it : P - 1 - it;
      vout[(size_t)s * QP + hl] = (hl < Q) ? h.v : __int_as_float(h.e);
      matvec_raw(U + (size_t)s * MAT, hl, hi, true, h);
    }
    const float tot = half_sum(h.v);
    if (tid == 0) {
      const float ll = (float)((double)h.e * LN2D + (double)logf(tot));
      a.llws[n] = ll;
      if (a.loglik) a.loglik[n] = ll;
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// K_B: boundary vectors over the tile matrices of one sequence (long sequences / log-likelihood only / sharding)
// ------------------------------------------------------------------------------------------------------------
constexpr int KB_WARPS = 16;

struct KBSmem {
  float seg[2][KB_WARPS][MAT];   // segment products (ping-pong for the in-warp chain), row stride QP
  float st[KB_WARPS][2][BSZ];    // per half warp staging of one block (row stride BST)
  float fseg[KB_WARPS + 1][QP], rseg[KB_WARPS][QP];
  int fse[KB_WARPS + 1], rse[KB_WARPS];
  int which[KB_WARPS];
  float ones[QP];
};

// C = A . B by one warp (blocks with row stride QP in shared memory)
__device__ __forceinline__ void warp_block_mm(const float* __restrict__ Ab, const float* __restrict__ Bb, float* __restrict__ Cb, int lane) {
  for (int item = lane; item < 64; item += 32) {
    const int r = item >> 2, jq = item & 3;
    if (r < Q) mm_strip(Ab, Bb, Cb, r, jq);
  }
  __syncwarp();
  mm_finish(Ab, Bb, Cb, lane & 15, lane < 16);   // both half warps hold identical values; one writes
  __syncwarp();
}

// copy one stored block (global, stride QP) into a staged block (shared, stride BST) by one half warp
__device__ __forceinline__ void hw_stage_block(float* dst, const float* __restrict__ src, int hl) {
#pragma unroll
  for (int r = 0; r < QP; ++r) dst[r * BST + hl] = __ldg(src + r * QP + hl);
}

// mode 0: full scan (writes bvf/bvr for every tile and the log-likelihood)
// mode 1: reduce only: the ordered product of all tile matrices of the sequence -> total_out[n] (multi-GPU exchange)
__global__ void __launch_bounds__(KB_WARPS * 32) boundary_kernel(const Args a, int mode, float* __restrict__ total_out) {
  extern __shared__ __align__(128) unsigned char smem[];
  KBSmem& S = *reinterpret_cast<KBSmem*>(smem);
  const int n = blockIdx.x, m = n / a.batch, P = a.P;
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31, hl = lane & 15, hi = lane >> 4;
  const int SL = (P + KB_WARPS - 1) / KB_WARPS;
  const int s_lo = min(P, w * SL), s_hi = min(P, s_lo + SL);
  const float* Mt = a.Mt + (size_t)n * P * MAT;
  if (threadIdx.x < QP) S.ones[threadIdx.x] = 1.f;

  // phase 1: ordered product of the segment's tile matrices
  {
    int cur = 0;
    float* acc = S.seg[0][w];
    if (s_lo < s_hi) {
      for (int i = lane; i < MAT / 4; i += 32) reinterpret_cast<float4*>(acc)[i] = __ldg(reinterpret_cast<const float4*>(Mt + (size_t)s_lo * MAT) + i);
    } else {
      identity_block(acc, lane);
    }
    __syncwarp();
    float* stg = &S.st[w][0][0];   // 2*BSZ floats >= MAT: scratch for the right operand (16-byte aligned)
    for (int s = s_lo + 1; s < s_hi; ++s) {
      for (int i = lane; i < MAT / 4; i += 32) reinterpret_cast<float4*>(stg)[i] = __ldg(reinterpret_cast<const float4*>(Mt + (size_t)s * MAT) + i);
      __syncwarp();
      warp_block_mm(S.seg[cur][w], stg, S.seg[cur ^ 1][w], lane);
      cur ^= 1;
    }
    if (lane == 0) S.which[w] = cur;
  }
  __syncthreads();

  if (mode == 1) {
    if (w == 0) {
      float* accA = &S.st[0][0][0];
      float* accB = &S.st[1][0][0];
      for (int i = lane; i < MAT; i += 32) accA[i] = S.seg[S.which[0]][0][i];
      __syncwarp();
      for (int s = 1; s < KB_WARPS; ++s) {
        warp_block_mm(accA, S.seg[S.which[s]][s], accB, lane);
        float* t = accA; accA = accB; accB = t;
      }
      for (int i = lane; i < MAT / 4; i += 32) reinterpret_cast<float4*>(total_out + (size_t)n * MAT)[i] = reinterpret_cast<const float4*>(accA)[i];
    }
    return;
  }

  // phase 2: warp 0 scans the segment products; lanes 0-15 forward, lanes 16-31 backward
  if (w == 0) {
    HV h;
    const float* inb = hi == 0 ? a.in_f : a.in_r;
    if (inb) {
      h.v = (hl < Q) ? __ldg(inb + (size_t)n * QP + hl) : 0.f;
      h.e = __float_as_int(__ldg(inb + (size_t)n * QP + 15));
    } else {
      h.v = (hl < Q) ? (hi == 0 ? __ldg(a.pi + m * Q + hl) : 1.f) : 0.f;
      h.e = 0;
    }
    float* stg = S.st[0][hi];
    for (int it = 0; it < KB_WARPS; ++it) {
      const int s = hi == 0 ? it : KB_WARPS - 1 - it;
      if (hi == 0) { S.fseg[s][hl] = h.v; if (hl == 0) S.fse[s] = h.e; }
      else         { S.rseg[s][hl] = h.v; if (hl == 0) S.rse[s] = h.e; }
      const float* sb = S.seg[S.which[s]][s];
#pragma unroll
      for (int r = 0; r < QP; ++r) stg[r * BST + hl] = sb[r * QP + hl];
      __syncwarp();
      matvec(stg, S.ones, hl, hi, true, h);
      __syncwarp();
    }
    const float tot = half_sum(h.v);
    if (lane == 0 && !a.ll_external) {
      const float ll = (float)((double)h.e * LN2D + (double)logf(tot));
      a.llws[n] = ll;
      if (a.loglik) a.loglik[n] = ll;
    }
  }
  __syncthreads();

  // phase 3: every warp replays its segment, emitting the boundary vector of every tile
  {
    HV h;
    h.v = hi == 0 ? S.fseg[w][hl] : S.rseg[w][hl];
    h.e = hi == 0 ? S.fse[w] : S.rse[w];
    float* vout = (hi == 0 ? a.bvf : a.bvr) + (size_t)n * P * QP;
    float* stg = S.st[w][hi];
    for (int it = 0; it < s_hi - s_lo; ++it) {
      const int s = hi == 0 ? s_lo + it : s_hi - 1 - it;
      vout[(size_t)s * QP + hl] = (hl < Q) ? h.v : __int_as_float(h.e);
      hw_stage_block(stg, Mt + (size_t)s * MAT, hl);
      __syncwarp();
      matvec(stg, S.ones, hl, hi, true, h);
      __syncwarp();
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// K_G: two-level boundary scan for very long sequences (config 5: 1 Mbp = 1737 tiles).  boundary_kernel walks a
// sequence's tile matrices with one CTA; beyond a few hundred tiles that walk is the whole run time.  Here the tile
// matrices are grouped by G: group_reduce_kernel multiplies every group down to one block (the same tensor-core
// product tree as K1), boundary_kernel then runs on the group matrices (a copy of Args with Mt / P / bvf / bvr pointing
// at the group level), and group_down_kernel replays each group from its boundary vectors to the tile level.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(K1_THREADS) group_reduce_kernel(const Args a, float* __restrict__ Gt) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* H = reinterpret_cast<float*>(smem);   // heap order: leaves [G, 2G), root at 1
  const int g = blockIdx.x, n = blockIdx.y, P = a.P, PG = gridDim.x, tid = threadIdx.x;
  const int nb = min(G, P - g * G);
  const float4* src = reinterpret_cast<const float4*>(a.Mt + ((size_t)n * P + (size_t)g * G) * MAT);
  float4* dst = reinterpret_cast<float4*>(H + G * MAT);
  for (int i = tid; i < nb * (MAT / 4); i += K1_THREADS) dst[i] = __ldg(src + i);
  for (int b = nb + (tid >> 5); b < G; b += K1_THREADS / 32) identity_block(H + (G + b) * MAT, tid & 31);
  __syncthreads();
  for (int np = G / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  float4* out = reinterpret_cast<float4*>(Gt + ((size_t)n * PG + g) * MAT);
  for (int i = tid; i < MAT / 4; i += K1_THREADS) out[i] = reinterpret_cast<const float4*>(H + MAT)[i];
}

__global__ void __launch_bounds__(256) group_down_kernel(const Args a, const float* __restrict__ gvf, const float* __restrict__ gvr) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* stg = reinterpret_cast<float*>(smem);       // G staged blocks, row stride BST
  float* ones = stg + G * BSZ;
  const int g = blockIdx.x, n = blockIdx.y, P = a.P, PG = gridDim.x, tid = threadIdx.x;
  const int nb = min(G, P - g * G);
  stage_blocks_async(stg, a.Mt + ((size_t)n * P + (size_t)g * G) * MAT, nb);
  cp_async_commit();
  if (tid < QP) ones[tid] = 1.f;
  cp_async_wait_all();
  __syncthreads();
  if (tid >= 32) return;
  const int hl = tid & 15, hi = tid >> 4;
  const float* gv = (hi == 0 ? gvf : gvr) + ((size_t)n * PG + g) * QP;
  HV h;
  h.v = (hl < Q) ? __ldg(gv + hl) : 0.f;
  h.e = __float_as_int(__ldg(gv + 15));
  float* vout = (hi == 0 ? a.bvf : a.bvr) + ((size_t)n * P + (size_t)g * G) * QP;
  for (int it = 0; it < nb; ++it) {
    const int s = hi == 0 ? it : nb - 1 - it;
    vout[(size_t)s * QP + hl] = (hl < Q) ? h.v : __int_as_float(h.e);
    matvec(stg + s * BSZ, ones, hl, hi, true, h);
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------------------
// K_S: sequence sharding over GPUs.  blocks[w][n] = ordered product of all tile matrices of shard w of sequence n
// (boundary_kernel mode 1 on every rank, then one all-gather).  One warp per sequence folds them into the boundary
// vector entering this shard (pi . M_0 ... M_{rank-1}), the one leaving it (M_{rank+1} ... M_{W-1} . 1) and the
// log-likelihood of the whole sequence: the TotalProbabilityCell scan (TotalProbabilityCell.py:30-49) across ranks.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) shard_boundary_kernel(const float* __restrict__ blocks, int rank, int world, int NB, int batch,
                                                            const float* __restrict__ pi, float* __restrict__ in_f,
                                                            float* __restrict__ in_r, float* __restrict__ llws,
                                                            float* __restrict__ loglik) {
  __shared__ __align__(16) float stg[2][BSZ];
  __shared__ float ones[QP];
  const int n = blockIdx.x, m = n / batch, lane = threadIdx.x, hl = lane & 15, hi = lane >> 4;
  if (lane < QP) ones[lane] = 1.f;
  HV h;
  h.v = (hl < Q) ? (hi == 0 ? __ldg(pi + m * Q + hl) : 1.f) : 0.f;
  h.e = 0;
  __syncwarp();
  for (int it = 0; it < world; ++it) {
    // forward half: shards 0, 1, ... (snapshot when reaching this rank); backward half: shards W-1 ... rank+1
    if (hi == 0 && it == rank) in_f[(size_t)n * QP + hl] = (hl < Q) ? h.v : __int_as_float(h.e);
    const int s = hi == 0 ? it : world - 1 - it;
    const bool act = hi == 0 ? true : (s > rank);
    hw_stage_block(stg[hi], blocks + ((size_t)(act ? s : rank) * NB + n) * MAT, hl);
    __syncwarp();
    matvec(stg[hi], ones, hl, hi, act, h);
    __syncwarp();
  }
  if (hi == 1) in_r[(size_t)n * QP + hl] = (hl < Q) ? h.v : __int_as_float(h.e);
  const float tot = half_sum(h.v);
  if (lane == 0) {
    const float ll = (float)((double)h.e * LN2D + (double)logf(tot));
    llws[n] = ll;
    if (loglik) loglik[n] = ll;
  }
}

// ------------------------------------------------------------------------------------------------------------
// K_X: sequence sharding with the exchange inside the kernels (peer memory over NVLink).  Two kernels replace
// group_reduce + boundary(mode 1) + [NCCL all-gather] + shard_boundary + boundary + group_down:
//   shard_product_kernel  one CTA per 32 tile matrices: group product (tensor-core tree, as group_reduce_kernel); the
//                         last CTA of a sequence multiplies the group matrices to the shard's block and stores it --
//                         with a step tag -- into the block table of EVERY rank (its own included).
//   shard_scan_kernel     one CTA per 32 tiles: waits for the tagged blocks of all ranks (ld.acquire.sys on its own
//                         flags), folds the ranks in front / behind (TotalProbabilityCell.py:30-49 across ranks), the
//                         groups in front / behind, and replays its own tiles: boundary vectors of every tile; the
//                         log-likelihood of the whole sequence comes out of the forward fold over all ranks.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(K1_THREADS) shard_product_kernel(const Args a, float* __restrict__ Gt, const Peers pr,
                                                                   float* __restrict__ block_out) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* H = reinterpret_cast<float*>(smem);   // heap order: leaves [G, 2G), root at 1
  __shared__ unsigned s_last;
  const int g = blockIdx.x, n = blockIdx.y, P = a.P, PG = gridDim.x, NB = gridDim.y, tid = threadIdx.x;
  const int nb = min(G, P - g * G);
  if (tid == 0) griddep_launch_dependents();
  for (int b = nb + (tid >> 5); b < G; b += K1_THREADS / 32) identity_block(H + (G + b) * MAT, tid & 31);
  griddep_wait();   // the tile matrices come from products_kernel (programmatic dependent launch)
  {
    const float4* src = reinterpret_cast<const float4*>(a.Mt + ((size_t)n * P + (size_t)g * G) * MAT);
    float4* dst = reinterpret_cast<float4*>(H + G * MAT);
    for (int i = tid; i < nb * (MAT / 4); i += K1_THREADS) dst[i] = __ldcg(src + i);
  }
  __syncthreads();
  for (int np = G / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  if (PG > 1) {
    float4* out = reinterpret_cast<float4*>(Gt + ((size_t)n * PG + g) * MAT);
    for (int i = tid; i < MAT / 4; i += K1_THREADS) out[i] = reinterpret_cast<const float4*>(H + MAT)[i];
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      const unsigned done = atomicAdd(a.cnt + n, 1u);
      s_last = (done == (unsigned)PG - 1u) ? 1u : 0u;
      if (s_last) { a.cnt[n] = 0u; __threadfence(); }   // left at zero for the next call
    }
    __syncthreads();
    if (!s_last) return;
    const float4* src = reinterpret_cast<const float4*>(Gt + (size_t)n * PG * MAT);
    // a tree over the next power of two of PG leaves only: its leaves sit at heap slots [W, 2W), its root at 1
    int W = 1;
    while (W < PG) W <<= 1;
    float4* dst = reinterpret_cast<float4*>(H + (size_t)W * MAT);
    for (int i = tid; i < PG * (MAT / 4); i += K1_THREADS) dst[i] = __ldcg(src + i);
    for (int b = PG + (tid >> 5); b < W; b += K1_THREADS / 32) identity_block(H + (W + b) * MAT, tid & 31);
    __syncthreads();
    for (int np = W / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  }
  // the shard's block: to the caller's buffer (NCCL route) or into every rank's table (peer stores over NVLink)
  const float4* root = reinterpret_cast<const float4*>(H + MAT);
  if (block_out != nullptr) {
    for (int i = tid; i < MAT / 4; i += K1_THREADS) reinterpret_cast<float4*>(block_out + (size_t)n * MAT)[i] = root[i];
  }
  if (pr.world > 0) {
    const size_t half = (size_t)(pr.tag & 1u) * pr.world * NB;
    for (int i = tid; i < pr.world * (MAT / 4); i += K1_THREADS) {
      const int r = i / (MAT / 4), j = i - r * (MAT / 4);
      reinterpret_cast<float4*>(pr.blocks[r] + (half + (size_t)pr.rank * NB + n) * MAT)[j] = root[j];
    }
    __threadfence_system();
    __syncthreads();
    if (tid < pr.world) st_release_sys_u32(pr.flags[tid] + half + (size_t)pr.rank * NB + n, pr.tag);
  }
}

// very long shards (more than G * G tiles per rank): the boundary kernels of the NCCL route do the scans, these two move
// the block into every rank's table and wait for the tables
__global__ void __launch_bounds__(64) shard_publish_kernel(const float* __restrict__ block, const Peers pr, int NB) {
  const int n = blockIdx.x, tid = threadIdx.x;
  const size_t half = (size_t)(pr.tag & 1u) * pr.world * NB;
  const float4 v = reinterpret_cast<const float4*>(block + (size_t)n * MAT)[tid];
  for (int r = 0; r < pr.world; ++r) reinterpret_cast<float4*>(pr.blocks[r] + (half + (size_t)pr.rank * NB + n) * MAT)[tid] = v;
  __threadfence_system();
  __syncthreads();
  if (tid < pr.world) st_release_sys_u32(pr.flags[tid] + half + (size_t)pr.rank * NB + n, pr.tag);
}
__global__ void __launch_bounds__(256) shard_wait_kernel(const unsigned* __restrict__ flags, int count, unsigned tag) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x)
    while (ld_acquire_sys_u32(flags + i) != tag) __nanosleep(100);
}

struct XSmem {
  float rb[MAX_PEERS][MAT];   // the ranks' blocks
  float gm[G][MAT];           // this sequence's group matrices
  float tm[G][MAT];           // this group's tile matrices
};

__global__ void __launch_bounds__(256) shard_scan_kernel(const Args a, const float* __restrict__ Gt, const float* __restrict__ blocks,
                                                         const unsigned* __restrict__ flags, int rank, int world, unsigned tag,
                                                         float* __restrict__ loglik) {
  extern __shared__ __align__(128) unsigned char smem[];
  XSmem& S = *reinterpret_cast<XSmem*>(smem);
  const int g = blockIdx.x, n = blockIdx.y, P = a.P, PG = gridDim.x, NB = gridDim.y, m = n / a.batch;
  const int tid = threadIdx.x, lane = tid & 31, hl = lane & 15, hi = lane >> 4;
  const int nb = min(G, P - g * G);
  if (tid == 0) griddep_launch_dependents();
  griddep_wait();   // tile / group matrices come from the kernels in front (programmatic dependent launch)
  // local data first (they do not depend on the other ranks), then the wait, then the ranks' blocks
  {
    const float4* src = reinterpret_cast<const float4*>(a.Mt + ((size_t)n * P + (size_t)g * G) * MAT);
    for (int i = tid; i < nb * (MAT / 4); i += 256) cp_async16(reinterpret_cast<float4*>(&S.tm[0][0]) + i, src + i);
    if (PG > 1) {
      const float4* gs = reinterpret_cast<const float4*>(Gt + (size_t)n * PG * MAT);
      for (int i = tid; i < PG * (MAT / 4); i += 256) cp_async16(reinterpret_cast<float4*>(&S.gm[0][0]) + i, gs + i);
    }
    cp_async_commit();
  }
  if (flags != nullptr && tid < world) {
    const unsigned* f = flags + (size_t)tid * NB + n;
    while (ld_acquire_sys_u32(f) != tag) __nanosleep(100);
  }
  __syncthreads();
  for (int i = tid; i < world * (MAT / 4); i += 256) {
    const int r = i / (MAT / 4), j = i - r * (MAT / 4);
    reinterpret_cast<float4*>(&S.rb[r][0])[j] = __ldcg(reinterpret_cast<const float4*>(blocks + ((size_t)r * NB + n) * MAT) + j);
  }
  cp_async_wait_all();
  __syncthreads();
  if (tid >= 32) return;
  // lanes 0-15: forward fold (row vector), lanes 16-31: backward fold (column vector); the shuffles are warp-wide, so both
  // halves run every loop to the longer trip count and commit only their own steps
  HV h;
  h.v = (hl < Q) ? (hi == 0 ? __ldg(a.pi + (size_t)m * Q + hl) : 1.f) : 0.f;
  h.e = 0;
  HV mine = h;                       // forward: entering this rank's shard
  for (int it = 0; it < world; ++it) {
    if (hi == 0 && it == rank) mine = h;
    const int s = hi == 0 ? it : world - 1 - it;
    matvec_raw(S.rb[s], hl, hi, hi == 0 ? true : s > rank, h);
  }
  {
    const float tot = half_sum(h.v);   // (warp-wide shuffle: both halves)
    if (g == 0 && lane == 0) {
      const float ll = (float)((double)h.e * LN2D + (double)logf(tot));
      a.llws[n] = ll;
      if (loglik) loglik[n] = ll;
    }
    if (hi == 0) h = mine;
  }
  if (PG > 1) {
    for (int it = 0; it < PG; ++it) {
      const int s = hi == 0 ? it : PG - 1 - it;
      matvec_raw(S.gm[s], hl, hi, hi == 0 ? s < g : s > g, h);
    }
  }
  float* vout = (hi == 0 ? a.bvf : a.bvr) + ((size_t)n * P + (size_t)g * G) * QP;
  for (int it = 0; it < nb; ++it) {
    const int s = hi == 0 ? it : nb - 1 - it;
    vout[(size_t)s * QP + hl] = (hl < Q) ? h.v : __int_as_float(h.e);
    matvec_raw(S.tm[s], hl, hi, true, h);
  }
}

// ------------------------------------------------------------------------------------------------------------
// K2: boundary vectors of every chunk, exact in-chunk recursions, outputs
// ------------------------------------------------------------------------------------------------------------
struct K2Smem {
  float Aw[24];            // the 23 edge weights in the reference's kernel order
  float hvf[HEAP][QP], hvr[HEAP][QP];   // forward vector entering / backward vector leaving every tree node
  int hef[HEAP], her[HEAP];             // their power-of-two exponents
  float ll;
  unsigned long long bar[2];            // mbarriers: [0] the tile's tree, [1] the tile's emission rows
};

// shared memory after K2Smem: X = G*CS emission rows, then U = the tile's tree (raw blocks, nodes 1 .. 2G-1) which
// becomes Y (and Yr) once the down-sweep is done
__host__ __device__ inline size_t k2_tree_bytes() { return (size_t)(HEAP - 1) * MAT * sizeof(float); }
__host__ __device__ inline size_t k2_y_bytes(int L, bool full) { return (size_t)(full ? 2 : 1) * G * (L * QP + 4) * sizeof(float); }
__host__ __device__ inline size_t k2_union_bytes(int L, bool full) {
  const size_t y = k2_y_bytes(L, full), b = k2_tree_bytes();
  return y > b ? y : b;
}
__host__ __device__ inline size_t k2_smem_bytes(int L, bool full) {
  return ((sizeof(K2Smem) + 15) / 16) * 16 + (size_t)G * (L * QP + 4) * sizeof(float) + k2_union_bytes(L, full);
}

template <bool FULL>
__global__ void __launch_bounds__(512, 1) finalize_kernel(const Args a) {
  extern __shared__ __align__(128) unsigned char smem[];
  K2Smem& S = *reinterpret_cast<K2Smem*>(smem);
  const int L = a.L, T = a.T, C = a.C, P = a.P, CS = L * QP + 4;
  float* X = reinterpret_cast<float*>(smem + ((sizeof(K2Smem) + 15) / 16) * 16);   // G*CS emissions
  float* U = X + (size_t)G * CS;                                                    // tree, then Y (and Yr)
  const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5, hl = lane & 15, hi = lane >> 4;
  const int p = blockIdx.x, n = blockIdx.y, m = n / a.batch;
  const int c0 = p * G, t_lo = c0 * L, npos = min(T, t_lo + G * L) - t_lo;
  const int nch = min(G, C - c0);                      // chunks in this tile
  const bool seq_start = (a.seq_off == 0);
  const size_t tile = (size_t)n * P + p;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x + gridDim.x * gridDim.y;
  HMM_MARK(cta, 0);
  HMM_WALL(cta, 14);

  // ---- bulk copies: the tile's tree (heap nodes 1 .. 2G-1 land at Bh[1 ..]) and its emission rows ----
  float* Bh = U - MAT;     // heap indexing: node k at Bh + k * MAT (node 0 does not exist and is never touched)
  const bool need_ll = (a.flags & HMM_B200_NO_LOGLIK) != 0;
  {
    // small loads first: they would otherwise queue behind this CTA's own bulk copies
    float aw = 0.f, bvv = 0.f, llv = 0.f;
    int bve = 0;
    if (tid < NEDGE) aw = __ldg(a.A + (size_t)m * Q * Q + ESRC[tid] * Q + EDST[tid]);
    griddep_wait();     // everything below reads what the kernels before this one wrote
    if (warp == 1) {
      if (a.top_scan == 2) {
        // short sequences: this CTA scans the sequence's tile matrices itself (under the bulk copies issued by warp 0)
        const HV b = scan_tiles_global(a.Mt + (size_t)n * P * MAT, P, p, a.pi + (size_t)m * Q, hl, hi, &llv);
        bvv = b.v; bve = b.e;
        if (lane == 0 && p == 0 && a.loglik) a.loglik[n] = llv;
      } else {
        // boundary vectors of the tile (K1's top scan or boundary_kernel): lanes 0-15 forward, lanes 16-31 backward
        const float* bv = (hi == 0 ? a.bvf : a.bvr) + tile * QP;
        bvv = __ldcg(bv + hl);
        bve = __float_as_int(__ldcg(bv + 15));
        if (lane == 0) llv = __ldcg(a.llws + n);
      }
    }
    if (tid == 0) {
      mbar_init(&S.bar[0], 1);
      mbar_init(&S.bar[1], 1);
      mbar_init_fence();
      mbar_expect_tx(&S.bar[0], (unsigned)k2_tree_bytes());
      bulk_g2s(U, a.Mh + (tile * HEAP + 1) * MAT, (unsigned)k2_tree_bytes(), &S.bar[0]);
      const unsigned ebytes = (unsigned)((size_t)G * CS * sizeof(float));
      mbar_expect_tx(&S.bar[1], ebytes);
      bulk_g2s(X, a.Ews + tile * (size_t)G * CS, ebytes, &S.bar[1]);
    }
    if (tid < NEDGE) S.Aw[tid] = aw;
    if (warp == 1) {
      if (hi == 0) { S.hvf[1][hl] = (hl < Q) ? bvv : 0.f; if (hl == 0) S.hef[1] = bve; }
      else         { S.hvr[1][hl] = (hl < Q) ? bvv : 0.f; if (hl == 0) S.her[1] = bve; }
      if (lane == 0) S.ll = llv;
    }
  }
  __syncthreads();   // mbarriers initialised, root vectors in place
  HMM_MARK(cta, 1);

  // ---- down-sweep of the tile's tree: node k hands its forward vector to child 2k and (through child 2k's matrix) to
  // child 2k+1; its backward vector to child 2k+1 and (through child 2k+1's matrix) to child 2k.  The two directions
  // are independent: the first half of the CTA's warps runs the forward sweep, the second half the backward sweep; a
  // half warp handles one node, so a level of `lo` nodes takes ceil(lo / (2 wpd)) rounds.
  {
    const int wpd = nt >> 6;                 // warps per direction
    const int dir = warp >= wpd ? 1 : 0;     // warp-uniform
    const int wd = warp - dir * wpd;         // warp index within its direction
    const int row = min(hl, Q - 1);
    mbar_wait(&S.bar[0], 0);
    for (int lo = 1; lo < G; lo <<= 1) {
      for (int k0 = lo + 2 * wd; k0 < 2 * lo; k0 += 2 * wpd) {
        const int k = k0 + hi;
        const bool act = k < 2 * lo;
        const int kc = act ? k : k0;
        if (dir == 0) {
          float v = S.hvf[kc][hl];
          int e = S.hef[kc];
          const float* blk = Bh + (size_t)(2 * kc) * MAT;
          const float u = (hl < Q) ? v * blk[row * QP + 15] : 0.f;
          const int Eb = __float_as_int(blk[MAT - 1]);
          const float s = half_sum(u);
          float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
          for (int j = 0; j < Q; ++j) {
            const float uj = __shfl_sync(0xffffffffu, u, j, 16);
            const float c = blk[j * QP + hl];
            if (j & 1) acc1 = fmaf(uj, c, acc1); else acc0 = fmaf(uj, c, acc0);
          }
          const int ex = pexp(s);
          if (act) {
            S.hvf[2 * k][hl] = v;
            S.hvf[2 * k + 1][hl] = (hl < Q) ? (acc0 + acc1) * pow2neg(ex) : 0.f;
            if (hl == 0) { S.hef[2 * k] = e; S.hef[2 * k + 1] = e + clamp_exp(ex) + Eb; }
          }
        } else {
          float v = S.hvr[kc][hl];
          int e = S.her[kc];
          const float* blk = Bh + (size_t)(2 * kc + 1) * MAT;
          const float sp = blk[15 + row * QP];
          const int Eb = __float_as_int(blk[MAT - 1]);
          const float u = (hl < Q) ? v : 0.f;
          const float s = half_sum((hl < Q) ? blk[15 * QP + hl] * v : 0.f);
          const float* rp = blk + row * QP;
          float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
          for (int j = 0; j < QP; ++j) {
            const int kk = (hl + j) & 15;            // diagonal walk: 16 lanes, 16 banks; slot 15 meets u[15] = 0
            const float uj = __shfl_sync(0xffffffffu, u, kk, 16);
            const float c = rp[kk];
            if (j & 1) acc1 = fmaf(uj, c, acc1); else acc0 = fmaf(uj, c, acc0);
          }
          const int ex = pexp(s);
          if (act) {
            S.hvr[2 * k + 1][hl] = v;
            S.hvr[2 * k][hl] = (hl < Q) ? sp * (acc0 + acc1) * pow2neg(ex) : 0.f;
            if (hl == 0) { S.her[2 * k + 1] = e; S.her[2 * k] = e + clamp_exp(ex) + Eb; }
          }
        }
      }
      if (dir == 0) named_barrier_sync(2, wpd * 32); else named_barrier_sync(3, wpd * 32);
    }
  }
  __syncthreads();   // boundary vectors complete, tree dead: U becomes Y (and Yr)
  mbar_wait(&S.bar[1], 0);   // every thread observes the emission rows' arrival itself
  HMM_MARK(cta, 2); HMM_MARK(cta, 3); HMM_MARK(cta, 4);

  // ---- exact recursions: warp 0 forward, warp 1 backward, lane = chunk ----
  // Each direction walks its chunk; the first half of its walk stores its own rows, the second half meets the rows
  // the other direction stored and leaves their product (the un-normalised posterior) in place.  States are kept
  // as kappa * alpha-hat with kappa a power of two chosen per step from the exponent bits of the row sum, so the
  // floor max(R, eps) of MsaHmmCell.py:88 becomes max(R', eps * sum) exactly.
  float* Yf = U;
  float* Yr = FULL ? U + (size_t)G * CS : U;
  if (warp < 2) {
    const int ci = lane, c = c0 + ci;
    const bool act = ci < nch;
    const int t0 = c * L, len = act ? min(L, T - t0) : 0, h = (len + 1) >> 1;
    const float* Xc = X + ci * CS;
    float w[NEDGE];
    load_edges(S.Aw, w);
    if (warp == 0) {
      float av[Q];
      {
        float t[QP];
        load_row16(S.hvf[G + ci], t);
#pragma unroll
        for (int j = 0; j < Q; ++j) av[j] = t[j];
      }
      float kap = sum15(av);
      int nsum = 0;
      float* Yc = Yf + ci * CS;
      const bool first = (c == 0 && seq_start);
      float en[QP];
      if (len > 0) load_row16(Xc, en);
      for (int ph = 0; ph < 2; ++ph) {
        const int u_lo = ph == 0 ? 0 : h, u_hi = ph == 0 ? h : len;
        for (int u = u_lo; u < u_hi; ++u) {
          float e[QP], R[Q];
#pragma unroll
          for (int j = 0; j < QP; ++j) e[j] = en[j];
          if (u + 1 < len) load_row16(Xc + (u + 1) * QP, en);   // next row in flight during this step
          if (first && u == 0) {
#pragma unroll
            for (int j = 0; j < Q; ++j) R[j] = av[j];
          } else {
            vec_A(av, w, R);
          }
          const float fl = kEps * kap;
#pragma unroll
          for (int j = 0; j < Q; ++j) av[j] = e[j] * fmaxf(R[j], fl);
          const float g = sum15(av);
          int ex;
          const float sc = recip_pow2(g, ex);
          nsum += ex;
#pragma unroll
          for (int j = 0; j < Q; ++j) av[j] *= sc;
          kap = g * sc;
          if (FULL || ph == 0) {
            // own row (sum in [1,2)); slot 15: the exponent of the true value relative to the chunk boundary
            store_row16(Yc + u * QP, av, __int_as_float(nsum));
          } else {
            float o[QP];
            load_row16(Yc + u * QP, o);
            float gm[Q];
#pragma unroll
            for (int j = 0; j < Q; ++j) gm[j] = av[j] * o[j];
            store_row16(Yc + u * QP, gm, 0.f);
          }
        }
        if (ph == 0) named_barrier_sync(1, 64);
      }
    } else {
      // backward: state bh = kappa * b-hat, output R = max(A.bh, eps*kappa), b = E * R  (MsaHmmCell.py:96-100)
      float R[Q];
      {
        float t[QP];
        load_row16(S.hvr[G + ci], t);
#pragma unroll
        for (int j = 0; j < Q; ++j) R[j] = t[j];
      }
      float kap = 1.f;           // the boundary vector is all-ones (sequence end, exact) or has its sum in [1,2)
      int nsum = 0;
      float* Yc = Yr + ci * CS;
      float en[QP];
      if (len > 0) load_row16(Xc + (len - 1) * QP, en);
      for (int ph = 0; ph < 2; ++ph) {
        const int u_hi = ph == 0 ? len : h, u_lo = ph == 0 ? h : 0;
        for (int u = u_hi - 1; u >= u_lo; --u) {
          float e[QP], Rf[Q], b[Q];
#pragma unroll
          for (int j = 0; j < QP; ++j) e[j] = en[j];
          if (u > 0) load_row16(Xc + (u - 1) * QP, en);
          const float fl = kEps * kap;
#pragma unroll
          for (int j = 0; j < Q; ++j) Rf[j] = fmaxf(R[j], fl);
#pragma unroll
          for (int j = 0; j < Q; ++j) b[j] = e[j] * Rf[j];
          if (FULL || ph == 0) {
            store_row16(Yc + u * QP, Rf, __int_as_float(nsum));
          } else {
            float o[QP];
            load_row16(Yc + u * QP, o);
            float gm[Q];
#pragma unroll
            for (int j = 0; j < Q; ++j) gm[j] = Rf[j] * o[j];
            store_row16(Yc + u * QP, gm, 0.f);
          }
          const float g = sum15(b);
          int ex;
          const float sc = recip_pow2(g, ex);
          nsum += ex;
#pragma unroll
          for (int j = 0; j < Q; ++j) b[j] *= sc;
          kap = g * sc;
          A_vec(b, w, R);
        }
        if (ph == 0) named_barrier_sync(1, 64);
      }
    }
  }
  __syncthreads();
  HMM_MARK(cta, 5);

  // ---- outputs ----
  const size_t obase = ((size_t)n * T + t_lo) * Q;
  const int nel = npos * Q;
  const float ll_shift = need_ll ? S.ll : 0.f;
  if (!FULL) {
    // Y holds alpha-hat * R (any positive scale).  Thread = position: log posterior = log g - log sum(g), staged in X
    // (dead now) in the output's own layout, with the same 16-byte phase as the destination, so that all but <= 3
    // leading / trailing floats leave by one bulk store.
    float* dstg = a.out_post + obase;
    const int shift = (int)((reinterpret_cast<uintptr_t>(dstg) >> 2) & 3);
    float* OUT = X + shift;
    constexpr float LN2F = 0.69314718055994531f;
    for (int i = tid; i < npos; i += nt) {
      float g[QP];
      load_row16(Yf + tile_row_offset(a, i, CS), g);
      float lg[Q];
      float z = 0.f;
#pragma unroll
      for (int j = 0; j < Q; ++j) { g[j] = fmaxf(g[j], 1e-37f); z += g[j]; }
      const float lz = __log2f(z);
#pragma unroll
      for (int j = 0; j < Q; ++j) lg[j] = fmaf(__log2f(g[j]) - lz, LN2F, ll_shift);
      float* o = OUT + i * Q;
#pragma unroll
      for (int j = 0; j < Q; ++j) o[j] = lg[j];
    }
    fence_async_smem();
    __syncthreads();
    HMM_MARK(cta, 6);
    const int head = min((4 - shift) & 3, nel), mid = (nel - head) & ~3, tail = nel - head - mid;
    if (tid == 0 && mid > 0) {
      bulk_s2g(dstg + head, OUT + head, (unsigned)mid * sizeof(float));
      bulk_commit();
    }
    if (tid >= 32 && tid < 32 + head) dstg[tid - 32] = OUT[tid - 32];
    if (tid >= 64 && tid < 64 + tail) dstg[head + mid + tid - 64] = OUT[head + mid + tid - 64];
    if (tid == 0) bulk_wait_read();
  } else {
    const bool norm_alpha = (a.flags & HMM_B200_ALPHA_NORMALIZED) != 0;
    if (a.out_post || norm_alpha) {
      for (int i = tid; i < npos; i += nt) {
        float f[QP], r[QP];
        const int off = tile_row_offset(a, i, CS);
        load_row16(Yf + off, f);
        load_row16(Yr + off, r);
        const float fsum = sum15p(f);
        const float sf = pow2neg(pexp(fsum)), sr = pow2neg(pexp(sum15p(r)));
        X[off + 3] = __logf(fmaxf(fsum, 1e-37f));   // slot 3: log of the row sum of alpha-hat (normalised alpha output)
        float z = 0.f;
#pragma unroll
        for (int j = 0; j < Q; ++j) z += fmaxf((f[j] * sf) * (r[j] * sr), 1e-37f);
        // X slot 0: -log z + log-likelihood shift; slots 1,2: the two power-of-two scales
        X[off] = (need_ll ? S.ll : 0.f) - __logf(z);
        X[off + 1] = sf;
        X[off + 2] = sr;
      }
      __syncthreads();
    }
    HMM_MARK(cta, 6);
    for (int i = tid; i < nel; i += nt) {
      const int pos = i / Q, j = i - pos * Q;
      const int ci = divL(a, pos);
      const int off = ci * CS + (pos - ci * L) * QP;
      const float f = Yf[off + j], r = Yr[off + j];
      // two logs: the product of a floor-level alpha-hat (>= 1e-32) and a floor-level R (>= 1e-16) leaves the
      // float range, its logarithm does not
      if (a.out_post)
        a.out_post[obase + i] = (__logf(fmaxf(f * X[off + 1], 1e-37f)) + __logf(fmaxf(r * X[off + 2], 1e-37f))) + X[off];
      if (a.out_alpha)
        a.out_alpha[obase + i] = norm_alpha ? __logf(fmaxf(f, 1e-37f)) - X[off + 3]
            : (float)((double)__logf(fmaxf(f, 1e-37f)) + (double)(__float_as_int(Yf[off + 15]) + S.hef[G + ci]) * LN2D);
      if (a.out_beta)
        a.out_beta[obase + i] = (float)((double)__logf(fmaxf(r, 1e-37f)) + (double)(__float_as_int(Yr[off + 15]) + S.her[G + ci]) * LN2D);
    }
  }
  HMM_MARK(cta, 7);
  HMM_WALL(cta, 15);
}

// stand-alone emission (cell.emission_probs API parity); unfloored
__global__ void __launch_bounds__(512) emission_kernel(const Args a, float* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char smem[];
  EmisSmem& sp = *reinterpret_cast<EmisSmem*>(smem);
  const int L = a.L, T = a.T, CS = L * QP + 4;
  float* X = reinterpret_cast<float*>(smem + ((sizeof(EmisSmem) + 15) / 16) * 16);
  signed char* codes = reinterpret_cast<signed char*>(X + (size_t)G * CS);
  const int tid = threadIdx.x, nt = blockDim.x;
  const int p = blockIdx.x, n = blockIdx.y, m = n / a.batch;
  const int t_lo = p * G * L, npos = min(T, t_lo + G * L) - t_lo;
  codes_stage(a, codes, n, t_lo, npos);
  load_emis_params(sp, a.Bfull + (size_t)m * Q * Q, a.codon5);
  __syncthreads();
  for (int i0 = tid; i0 < npos; i0 += 2 * nt) {
    const int i1 = i0 + nt;
    const bool two = i1 < npos;
    float e0[Q], e1[Q];
    emit_pair<false>(a, sp, n, t_lo + i0, t_lo + i1, two, codes + 2 + i0, codes + 2 + i1, e0, e1);
    store_row16(X + tile_row_offset(a, i0, CS), e0, 0.f);
    if (two) store_row16(X + tile_row_offset(a, i1, CS), e1, 0.f);
  }
  __syncthreads();
  const size_t obase = ((size_t)n * T + t_lo) * Q;
  const bool as_log = (a.flags & EMIT_LOG_FLOORED) != 0;   // Viterbi's operand: log max(E, eps)
  for (int i = tid; i < npos * Q; i += nt) {
    const int pos = i / Q, j = i - pos * Q;
    const float v = X[tile_row_offset(a, pos, CS) + j];
    out[obase + i] = as_log ? logf(fmaxf(v, kEps)) : v;
  }
}

// ------------------------------------------------------------------------------------------------------------
// Gradient of the log-likelihood (training through the layer: the reference differentiates its Python loop with
// autograd, MsaHmmCell.py:73-106 / gene_pred_hmm_emitter.py:93-121,231-263).  Inputs: the normalised log posterior
// gamma and the normalised forward vector alpha-hat of the fused evaluation.  With R_t = alpha-hat_{t-1} A,
//   d ll / d A[i][j]   = sum_t gamma_t(j) alpha-hat_{t-1}(i) / R_t(j)           (xi_t(i,j) / A[i][j]; the emission cancels)
//   d ll / d pi[i]     = gamma_0(i) / pi(i)
//   d ll / d dot_t(q)  = gamma_t(q) / dot_t(q),  dot_t(q) = sum_c x_t[c] B[q][c]  (codon factor and end hint cancel)
//   d ll / d x_t[c]    = sum_q r_t(q) B[q][c],   d ll / d B[q][c] = sum_t r_t(q) x_t[c],   r_t(q) = gamma_t(q) / dot_t(q)
// A floored factor (max(E, eps), max(R, eps), MsaHmmCell.py:87-88) passes no gradient, like torch.maximum.
// One thread per position for the per-position part, then the CTA's 256 x (15 x 15) outer-product sum; per-CTA partial
// sums go to the workspace and grad_reduce_kernel adds them in a fixed order (deterministic, no atomics).
// ------------------------------------------------------------------------------------------------------------
constexpr int GR_POS = 256;        // positions (= threads) per CTA
constexpr int GR_PART = 320;       // floats per partial record: [0,256) dB[q][c] (stride 16), [256,279) dA edges, [288,303) dpi

struct GradArgs {
  const float* log_post;   // [NB][T][15] normalised log posterior
  const float* log_ahat;   // [NB][T][15] log of the normalised forward vector (HMM_B200_ALPHA_NORMALIZED)
  const float* gll;        // [NB] upstream gradient of every log-likelihood
  float* grad_inputs;      // [NB][T][20] or NULL
  float* part;             // [NB * PT][GR_PART] partial sums
  int PT;                  // CTAs per sequence
};

struct GradSmem {
  EmisSmem sp;
  float A[QP * QP];               // row stride Q; padded so that what follows stays 16-byte aligned
  float pi[QP];
  float WA[GR_POS / 32][24];
  float GP[QP];
  float X[GR_POS * SIN];          // input rows
  float R[GR_POS][QP];            // r_t(q) * upstream gradient
  float PS[GR_POS * Q];           // log posterior rows
  float AS[(GR_POS + 1) * Q];     // log alpha-hat rows of positions t0 - 1 .. t0 + 255
  signed char codes[GR_POS + 16];
};

__global__ void __launch_bounds__(GR_POS, 2) grad_kernel(const Args a, const GradArgs g) {
  extern __shared__ __align__(128) unsigned char smem[];
  GradSmem& S = *reinterpret_cast<GradSmem*>(smem);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int p = blockIdx.x, n = blockIdx.y, m = n / a.batch, T = a.T;
  const int t0 = p * GR_POS, npos = min(T, t0 + GR_POS) - t0;
  codes_stage(a, S.codes, n, t0, npos);
  load_emis_params(S.sp, a.Bfull + (size_t)m * Q * Q, a.codon5);
  if (tid < Q * Q) S.A[tid] = __ldg(a.A + (size_t)m * Q * Q + tid);
  if (tid < QP) { S.pi[tid] = (tid < Q) ? __ldg(a.pi + m * Q + tid) : 0.f; S.GP[tid] = 0.f; }
  {
    const float4* src = reinterpret_cast<const float4*>(a.inputs + ((size_t)n * T + t0) * SIN);
    float4* dst = reinterpret_cast<float4*>(S.X);
    for (int i = tid; i < npos * (SIN / 4); i += GR_POS) dst[i] = __ldg(src + i);
    const float* ps = g.log_post + ((size_t)n * T + t0) * Q;
    for (int i = tid; i < npos * Q; i += GR_POS) S.PS[i] = __ldg(ps + i);
    const int lead = (t0 > 0) ? Q : 0;    // row t0 - 1 exists
    const float* as = g.log_ahat + ((size_t)n * T + t0) * Q - lead;
    for (int i = tid; i < npos * Q + lead - Q; i += GR_POS) S.AS[Q - lead + i] = __ldg(as + i);   // rows t0-1 .. t0+npos-2
  }
  __syncthreads();

  const float w = __ldg(g.gll + n);
  float gA[NEDGE];
#pragma unroll
  for (int e = 0; e < NEDGE; ++e) gA[e] = 0.f;
  if (tid < npos) {
    const int i = tid, t = t0 + i;
    const float4* row = reinterpret_cast<const float4*>(S.X + i * SIN);
    const float4 x0 = row[0], x1 = row[1], x2 = row[2], x3 = row[3];
    float cod[9];
    codon_factors(a, S.sp, n, t, S.codes + 2 + i, cod);
    const float train = (a.flags & HMM_B200_TRAINING) ? 1e-7f : 0.f;
    const float free_cod = (1.0f / 4096.0f) + train;
    const int gpos = a.seq_off + t;
    const float* hint = nullptr;
    if (a.hints != nullptr) {
      if (gpos == 0) hint = a.hints + (size_t)n * 2 * Q;
      else if (gpos == a.seq_total - 1) hint = a.hints + (size_t)n * 2 * Q + Q;
    }
    const bool both_hints = a.hints != nullptr && a.seq_total == 1;
    float gam[Q], r[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const float4* b = reinterpret_cast<const float4*>(S.sp.B[q]);
      const float dot = dot4(x3, b[3], dot4(x2, b[2], dot4(x1, b[1], dot4(x0, b[0], 0.f))));
      float e = dot * ((q < 6) ? free_cod : cod[q < 6 ? 0 : q - 6] + train);
      if (hint) e *= __ldg(hint + q);
      if (both_hints) e *= __ldg(a.hints + (size_t)n * 2 * Q + Q + q);
      gam[q] = __expf(S.PS[i * Q + q]);
      r[q] = (e > kEps) ? w * gam[q] / dot : 0.f;
    }
    store_row16(S.R[i], r, 0.f);
    if (g.grad_inputs) {
      float gx[QP];
#pragma unroll
      for (int c = 0; c < QP; ++c) gx[c] = 0.f;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        const float4* b = reinterpret_cast<const float4*>(S.sp.B[q]);
        const float4 b0 = b[0], b1 = b[1], b2 = b[2], b3 = b[3];
        gx[0] = fmaf(r[q], b0.x, gx[0]); gx[1] = fmaf(r[q], b0.y, gx[1]); gx[2] = fmaf(r[q], b0.z, gx[2]); gx[3] = fmaf(r[q], b0.w, gx[3]);
        gx[4] = fmaf(r[q], b1.x, gx[4]); gx[5] = fmaf(r[q], b1.y, gx[5]); gx[6] = fmaf(r[q], b1.z, gx[6]); gx[7] = fmaf(r[q], b1.w, gx[7]);
        gx[8] = fmaf(r[q], b2.x, gx[8]); gx[9] = fmaf(r[q], b2.y, gx[9]); gx[10] = fmaf(r[q], b2.z, gx[10]); gx[11] = fmaf(r[q], b2.w, gx[11]);
        gx[12] = fmaf(r[q], b3.x, gx[12]); gx[13] = fmaf(r[q], b3.y, gx[13]); gx[14] = fmaf(r[q], b3.z, gx[14]);
      }
      float4* o = reinterpret_cast<float4*>(g.grad_inputs + ((size_t)n * T + t) * SIN);
      o[0] = make_float4(gx[0], gx[1], gx[2], gx[3]);
      o[1] = make_float4(gx[4], gx[5], gx[6], gx[7]);
      o[2] = make_float4(gx[8], gx[9], gx[10], gx[11]);
      o[3] = make_float4(gx[12], gx[13], gx[14], 0.f);   // the nucleotide flags are data: no gradient
      o[4] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (gpos == 0) {
#pragma unroll
      for (int q = 0; q < Q; ++q) S.GP[q] = (S.pi[q] > kEps) ? w * gam[q] / S.pi[q] : 0.f;
    } else {
      float ah[Q], den[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) { ah[q] = __expf(S.AS[i * Q + q]); den[q] = 0.f; }
#pragma unroll
      for (int e = 0; e < NEDGE; ++e) den[EDST[e]] = fmaf(ah[ESRC[e]], S.A[ESRC[e] * Q + EDST[e]], den[EDST[e]]);
      float gd[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) gd[q] = (den[q] > kEps) ? w * gam[q] / den[q] : 0.f;
#pragma unroll
      for (int e = 0; e < NEDGE; ++e) gA[e] = gd[EDST[e]] * ah[ESRC[e]];
    }
  }
  // edge sums: warp shuffle tree, then the 8 warps in order
#pragma unroll
  for (int e = 0; e < NEDGE; ++e) {
    float v = gA[e];
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) S.WA[warp][e] = v;
  }
  __syncthreads();
  float* part = g.part + ((size_t)n * g.PT + p) * GR_PART;
  {
    // dB[q][c] = sum over the CTA's positions of r(q) x[c]: lane = 2 x 4 block of (q, c), warp w takes positions w, w + 8, ...;
    // the 8 warp sums are added in order through RED (the log posterior rows are dead by now)
    float* RED = S.PS;
    const int qt = lane >> 2, ct = lane & 3;
    float acc[2][4];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k >> 2][k & 3] = 0.f;
#pragma unroll 4
    for (int i = warp; i < npos; i += GR_POS / 32) {
      const float2 rv = *reinterpret_cast<const float2*>(&S.R[i][2 * qt]);
      const float4 xv = *reinterpret_cast<const float4*>(&S.X[i * SIN + 4 * ct]);
      acc[0][0] = fmaf(rv.x, xv.x, acc[0][0]); acc[0][1] = fmaf(rv.x, xv.y, acc[0][1]);
      acc[0][2] = fmaf(rv.x, xv.z, acc[0][2]); acc[0][3] = fmaf(rv.x, xv.w, acc[0][3]);
      acc[1][0] = fmaf(rv.y, xv.x, acc[1][0]); acc[1][1] = fmaf(rv.y, xv.y, acc[1][1]);
      acc[1][2] = fmaf(rv.y, xv.z, acc[1][2]); acc[1][3] = fmaf(rv.y, xv.w, acc[1][3]);
    }
#pragma unroll
    for (int h = 0; h < 2; ++h)
      *reinterpret_cast<float4*>(&RED[warp * 256 + (2 * qt + h) * QP + 4 * ct]) = make_float4(acc[h][0], acc[h][1], acc[h][2], acc[h][3]);
    __syncthreads();
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < GR_POS / 32; ++k) v += RED[k * 256 + tid];
    part[tid] = ((tid & 15) < Q) ? v : 0.f;
  }
  if (tid < 32) {
    float v = 0.f;
    if (tid < NEDGE) {
#pragma unroll
      for (int k = 0; k < GR_POS / 32; ++k) v += S.WA[k][tid];
    }
    part[256 + tid] = v;
    part[288 + tid] = (tid < Q) ? S.GP[tid] : 0.f;
  }
}

// partial records of model blockIdx.y -> grad_A (23 edges; the rest of the matrix is zeroed by the caller), grad_pi,
// grad_Bfull.  Block x owns 32 consecutive record slots; warp w adds records w, w + 32, ... and warp 0 adds the 32 sums.
__global__ void __launch_bounds__(1024) grad_reduce_kernel(const float* __restrict__ part, int nrec, float* __restrict__ gA,
                                                           float* __restrict__ gpi, float* __restrict__ gB) {
  __shared__ float red[32][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, m = blockIdx.y, slot = blockIdx.x * 32 + lane;
  const float* src = part + (size_t)m * nrec * GR_PART + slot;
  float v = 0.f;
  for (int j = warp; j < nrec; j += 32) v += __ldg(src + (size_t)j * GR_PART);
  red[warp][lane] = v;
  __syncthreads();
  if (warp != 0) return;
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 32; ++k) s += red[k][lane];
  if (slot < 256) {
    const int q = slot >> 4, c = slot & 15;
    if (gB && q < Q && c < Q) gB[(size_t)m * Q * Q + q * Q + c] = s;
  } else if (slot < 256 + NEDGE) {
    const int e = slot - 256;
    if (gA) gA[(size_t)m * Q * Q + ESRC[e] * Q + EDST[e]] = s;
  } else if (slot >= 288 && slot < 288 + Q) {
    if (gpi) gpi[(size_t)m * Q + slot - 288] = s;
  }
}

// ------------------------------------------------------------------------------------------------------------
// Viterbi decoding (most probable state path).  The reference names it (MsaHmmCell.py:13, log_A_dense :46-47) but does
// not implement it; the recursion is the max-product form of MsaHmmCell.py:73-106 over the same floored quantities:
//   delta_0 = log max(pi, eps) + log max(E_0, eps),  delta_t(j) = max_i (delta_{t-1}(i) + log A[i][j]) + log max(E_t(j), eps),
// ties to the lowest state index.  One warp per sequence (lane = state; a gene-model state has at most 3 in-edges):
// the recursion is a chain of T dependent steps, so this kernel is latency-bound by construction; emissions come from
// the emission kernel's buffer 16 steps ahead of their use, back-pointers go to a byte table that the same warp walks
// backwards through shared memory, 2048 steps at a time.
// ------------------------------------------------------------------------------------------------------------
constexpr int VT_PF = 16;       // steps per block = emission prefetch depth
constexpr int VT_BLK = 2048;    // back-pointer steps staged per backtracking round (64 per lane)
constexpr int VT_SEG = VT_BLK / 32;

// Back-pointer table: step k (position t = k + 1) of state j is byte (k & 3) of word (k >> 2) * 16 + j, so that a lane
// stores one word per 4 steps.
__global__ void __launch_bounds__(32) viterbi_kernel(const float* __restrict__ E, const float* __restrict__ A,
                                                     const float* __restrict__ pi, int batch, int T,
                                                     unsigned* __restrict__ psi, size_t psi_words, unsigned char* __restrict__ path,
                                                     float* __restrict__ score) {
  __shared__ unsigned pbuf[VT_BLK / 4 * 16];     // 2048 steps of back-pointers
  __shared__ unsigned char comp[32][16];         // per segment: state at its first step given the state after its last step
  __shared__ unsigned char obuf[VT_BLK];
  const int n = blockIdx.x, m = n / batch, lane = threadIdx.x;
  const bool live = lane < Q;
  const int j = live ? lane : 0;
  const float NEG = -INFINITY;
  int src[3] = {0, 0, 0};
  float la[3] = {NEG, NEG, NEG};   // dead lanes keep -inf on every in-edge: their delta stays -inf without a select
  if (live) {
    int ns = 0;
    for (int i = 0; i < Q; ++i) {
      const float a = __ldg(A + (size_t)m * Q * Q + i * Q + j);
      if (a > 0.f && ns < 3) {
        if (ns == 0) { src[0] = i; la[0] = logf(a); }
        else if (ns == 1) { src[1] = i; la[1] = logf(a); }
        else { src[2] = i; la[2] = logf(a); }
        ++ns;
      }
    }
  }
  const float* Erow = E + (size_t)n * T * Q + j;
  unsigned* pw = psi + (size_t)n * psi_words + (lane & 15);
  float delta = live ? logf(fmaxf(__ldg(pi + m * Q + j), kEps)) + __ldg(Erow) : NEG;
  double offset = 0.0;
  const int K = T - 1;   // steps
  // one step: the delta chain is shuffle -> add -> max -> max -> add; the back-pointer (lowest source among the maxima)
  // is worked out beside it
  auto step = [&](float e, unsigned& word, int byte) {
    const float c0 = __shfl_sync(0xffffffffu, delta, src[0]) + la[0];
    const float c1 = __shfl_sync(0xffffffffu, delta, src[1]) + la[1];
    const float c2 = __shfl_sync(0xffffffffu, delta, src[2]) + la[2];
    const float best = fmaxf(fmaxf(c0, c1), c2);
    const int arg = (c0 == best) ? src[0] : (c1 == best) ? src[1] : src[2];
    word |= (unsigned)arg << (8 * byte);
    delta = best + e;
  };
  auto renorm = [&]() {   // keep the magnitudes small: take the running maximum out (exact bookkeeping in double)
    float mx = delta;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    delta -= mx;
    offset += (double)mx;
  };
  // E holds log max(E_t, eps) (emission_kernel, EMIT_LOG_FLOORED).  Two register sets alternate: while the chain consumes
  // one block of VT_PF steps, the loads of the next block are in flight into the other set (no register moves)
  float ea[VT_PF], eb[VT_PF];
  auto run_block = [&](const float (&use)[VT_PF], float (&fill)[VT_PF], int kb) {   // steps kb .. kb + VT_PF - 1, all < K
#pragma unroll
    for (int u = 0; u < VT_PF; ++u) fill[u] = (kb + VT_PF + u < K) ? __ldg(Erow + (size_t)(kb + VT_PF + u + 1) * Q) : 0.f;
#pragma unroll
    for (int g4 = 0; g4 < VT_PF / 4; ++g4) {
      unsigned word = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) step(use[g4 * 4 + b], word, b);
      if (lane < 16) pw[(size_t)((kb >> 2) + g4) * 16] = word;
    }
    renorm();
  };
#pragma unroll
  for (int u = 0; u < VT_PF; ++u) ea[u] = (u < K) ? __ldg(Erow + (size_t)(u + 1) * Q) : 0.f;
  int kb = 0;
  for (; kb + 2 * VT_PF <= K; kb += 2 * VT_PF) {
    run_block(ea, eb, kb);
    run_block(eb, ea, kb + VT_PF);
  }
  for (; kb < K; kb += 4) {   // tail: fewer than 2 * VT_PF steps, plain loads
    unsigned word = 0;
    for (int b = 0; b < 4 && kb + b < K; ++b) step(__ldg(Erow + (size_t)(kb + b + 1) * Q), word, b);
    if (lane < 16) pw[(size_t)(kb >> 2) * 16] = word;
    renorm();
  }
  // best final state, lowest index on ties
  float bv = delta;
  int bs = live ? lane : Q;
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
    const int os = __shfl_xor_sync(0xffffffffu, bs, o);
    if (ov > bv || (ov == bv && os < bs)) { bv = ov; bs = os; }
  }
  if (lane == 0 && score) score[n] = (float)(offset + (double)bv);

  // ---- backtracking: state before step k = psi_k(state after step k).  Rounds of VT_BLK steps from the end; inside a
  // round every lane first composes the back-pointer maps of its own 64-step segment for all 15 end states (parallel),
  // lane 0 chains the 32 segment maps, then every lane replays its segment from its now known end state ----
  __threadfence_block();
  __syncwarp();
  int s_after = bs;                          // state after the last step of the round = state at position hi_k + 1
  unsigned char* out = path + (size_t)n * T;
  if (lane == 0) out[T - 1] = (unsigned char)bs;
  const unsigned* pg = psi + (size_t)n * psi_words;
  const unsigned char* pb = reinterpret_cast<const unsigned char*>(pbuf);
  for (int hi = K; hi > 0;) {                // steps [lo, hi) of this round; lo is a multiple of 4
    const int lo = (hi > VT_BLK) ? ((hi - VT_BLK + 3) & ~3) : 0, cnt = hi - lo;
    const int nw = ((cnt + 3) >> 2) * 16;
    for (int i = lane; i < nw; i += 32) pbuf[i] = pg[(size_t)(lo >> 2) * 16 + i];
    __syncwarp();
    // segment of this lane: local steps [a, b)
    const int a = min(lane * VT_SEG, cnt), b = min(a + VT_SEG, cnt);
    {
      int st[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) st[q] = q;
      for (int k = b - 1; k >= a; --k) {
        const unsigned char* row = pb + (size_t)(k >> 2) * 64 + (k & 3);
#pragma unroll
        for (int q = 0; q < Q; ++q) st[q] = row[st[q] * 4];
      }
#pragma unroll
      for (int q = 0; q < Q; ++q) comp[lane][q] = (unsigned char)st[q];
    }
    __syncwarp();
    // state after each segment's last step, from the last segment backwards
    int mine = 0;
    {
      int cur = s_after;
      for (int g = 31; g >= 0; --g) {
        if (g == lane) mine = cur;
        cur = comp[g][cur];
      }
      s_after = cur;                         // state before step lo = state after the previous round's last step
    }
    // replay: position of local step k is lo + k + 1; the state BEFORE step k is the state at position lo + k
    {
      int cur = mine;
      for (int k = b - 1; k >= a; --k) {
        cur = pb[(size_t)(k >> 2) * 64 + cur * 4 + (k & 3)];
        obuf[k] = (unsigned char)cur;
      }
    }
    __syncwarp();
    for (int i = lane; i < cnt; i += 32) out[lo + i] = obuf[i];
    __syncwarp();
    hi = lo;
  }
}

// ------------------------------------------------------------------------------------------------------------
// Parameter construction of the gene model in one launch (training steps rebuild A, pi, B from the kernels every call,
// gene_pred_hmm_transitioner.py:66-112 via Transitioner.py:337-380, gene_pred_hmm_emitter.py:65,87-91,115-116; as ~20 small
// PyTorch ops plus their autograd that is the bulk of a training step's host time):
//   A[i][j]   = (s + 1e-16) / (sum over i's out-edges of (s + 1e-16) + 1e-16),  s = softmax over the 15 dense entries of
//               row i, absent entries at -1000, kernel values clamped at -999 (the reference's guards)
//   pi        = softmax(starting distribution kernel);   Bfull = softmax rows of the emission kernel, intron row x 3
// and its backward.  One CTA; thread = one output row.  The same kernels serve every model (A, pi) / one per model (B).
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) params_kernel(const float* __restrict__ tk, const float* __restrict__ sk,
                                                    const float* __restrict__ ek, int num_models, float* __restrict__ A,
                                                    float* __restrict__ pi, float* __restrict__ Bfull) {
  const int tid = threadIdx.x;
  if (tid < Q) {   // row tid of A
    float kv[3]; int dst[3]; int deg = 0;
    for (int e = 0; e < NEDGE; ++e)
      if (ESRC[e] == tid && deg < 3) { kv[deg] = fmaxf(__ldg(tk + e), -999.f); dst[deg] = EDST[e]; ++deg; }
    float mx = -1000.f;
    for (int d = 0; d < deg; ++d) mx = fmaxf(mx, kv[d]);
    float den = (float)(Q - deg) * expf(-1000.f - mx), p[3], z = 1e-16f;
    for (int d = 0; d < deg; ++d) { p[d] = expf(kv[d] - mx); den += p[d]; }
    for (int d = 0; d < deg; ++d) { p[d] = p[d] / den + 1e-16f; z += p[d]; }
    for (int m = 0; m < num_models; ++m) {
      float* row = A + ((size_t)m * Q + tid) * Q;
      for (int j = 0; j < Q; ++j) row[j] = 0.f;
      for (int d = 0; d < deg; ++d) row[dst[d]] = p[d] / z;
    }
  } else if (tid == 32) {
    float mx = -INFINITY, den = 0.f, v[Q];
    for (int j = 0; j < Q; ++j) mx = fmaxf(mx, __ldg(sk + j));
    for (int j = 0; j < Q; ++j) { v[j] = expf(__ldg(sk + j) - mx); den += v[j]; }
    for (int m = 0; m < num_models; ++m)
      for (int j = 0; j < Q; ++j) pi[m * Q + j] = v[j] / den;
  }
  // emission rows: 13 kernel rows per model -> 15 state rows (rows 1,2,3 = kernel row 1)
  for (int r = tid; r < num_models * Q; r += blockDim.x) {
    const int m = r / Q, st = r - m * Q, kr = st == 0 ? 0 : (st <= 3 ? 1 : st - 2);
    const float* src = ek + ((size_t)m * 13 + kr) * Q;
    float mx = -INFINITY, den = 0.f, v[Q];
    for (int j = 0; j < Q; ++j) mx = fmaxf(mx, __ldg(src + j));
    for (int j = 0; j < Q; ++j) { v[j] = expf(__ldg(src + j) - mx); den += v[j]; }
    for (int j = 0; j < Q; ++j) Bfull[(size_t)r * Q + j] = v[j] / den;
  }
}

__global__ void __launch_bounds__(64) params_grad_kernel(const float* __restrict__ tk, const float* __restrict__ sk,
                                                         const float* __restrict__ ek, int num_models,
                                                         const float* __restrict__ gA, const float* __restrict__ gpi,
                                                         const float* __restrict__ gB, float* __restrict__ gtk,
                                                         float* __restrict__ gsk, float* __restrict__ gek) {
  const int tid = threadIdx.x;
  if (tid < Q && gtk) {
    float kv[3], raw[3], g[3]; int eid[3]; int deg = 0;
    for (int e = 0; e < NEDGE; ++e)
      if (ESRC[e] == tid && deg < 3) {
        raw[deg] = __ldg(tk + e); kv[deg] = fmaxf(raw[deg], -999.f); eid[deg] = e;
        float acc = 0.f;
        if (gA) for (int m = 0; m < num_models; ++m) acc += __ldg(gA + ((size_t)m * Q + tid) * Q + EDST[e]);
        g[deg] = acc; ++deg;
      }
    float mx = -1000.f;
    for (int d = 0; d < deg; ++d) mx = fmaxf(mx, kv[d]);
    float den = (float)(Q - deg) * expf(-1000.f - mx), sft[3], p[3], z = 1e-16f;
    for (int d = 0; d < deg; ++d) { sft[d] = expf(kv[d] - mx); den += sft[d]; }
    for (int d = 0; d < deg; ++d) { sft[d] /= den; p[d] = sft[d] + 1e-16f; z += p[d]; }
    // A_e = p_e / z:  dL/dp_e = (g_e - sum_e' g_e' A_e') / z;  p = s + const;  s = softmax (absent entries carry no gradient)
    float dot = 0.f;
    for (int d = 0; d < deg; ++d) dot += g[d] * (p[d] / z);
    float ds[3], sdot = 0.f;
    for (int d = 0; d < deg; ++d) { ds[d] = (g[d] - dot) / z; sdot += sft[d] * ds[d]; }
    for (int d = 0; d < deg; ++d) gtk[eid[d]] = (raw[d] >= -999.f) ? sft[d] * (ds[d] - sdot) : 0.f;
  } else if (tid == 32 && gsk) {
    float mx = -INFINITY, den = 0.f, v[Q], g[Q], dot = 0.f;
    for (int j = 0; j < Q; ++j) mx = fmaxf(mx, __ldg(sk + j));
    for (int j = 0; j < Q; ++j) { v[j] = expf(__ldg(sk + j) - mx); den += v[j]; }
    for (int j = 0; j < Q; ++j) {
      float acc = 0.f;
      if (gpi) for (int m = 0; m < num_models; ++m) acc += __ldg(gpi + m * Q + j);
      g[j] = acc; v[j] /= den; dot += v[j] * acc;
    }
    for (int j = 0; j < Q; ++j) gsk[j] = v[j] * (g[j] - dot);
  }
  if (gek) {
    for (int r = tid; r < num_models * 13; r += blockDim.x) {
      const int m = r / 13, kr = r - m * 13;
      const float* src = ek + (size_t)r * Q;
      float mx = -INFINITY, den = 0.f, v[Q], g[Q], dot = 0.f;
      for (int j = 0; j < Q; ++j) mx = fmaxf(mx, __ldg(src + j));
      for (int j = 0; j < Q; ++j) { v[j] = expf(__ldg(src + j) - mx); den += v[j]; }
      const int s0 = kr == 0 ? 0 : (kr == 1 ? 1 : kr + 2), ns = kr == 1 ? 3 : 1;
      for (int j = 0; j < Q; ++j) {
        float acc = 0.f;
        if (gB) for (int u = 0; u < ns; ++u) acc += __ldg(gB + ((size_t)m * Q + s0 + u) * Q + j);
        g[j] = acc; v[j] /= den; dot += v[j] * acc;
      }
      for (int j = 0; j < Q; ++j) gek[(size_t)r * Q + j] = v[j] * (g[j] - dot);
    }
  }
}

// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
// single-launch forward + posterior (gene15_fused.cu)
int fused_capacity(int L);
int fused_run_len(int L);
int launch_fused(Args a, int NB, cudaStream_t st);
constexpr int FUSED_NOT_RESIDENT = 1001;   // launch_fused: the cooperative launch was refused for size (gene15_fused.cu)
int launch_boundary_groups(const Args& a, int NB, float* Gt, unsigned* gflg, cudaStream_t st);
int launch_fused_shard(Args a, int NB, const unsigned long long* peer_blocks, const unsigned long long* peer_flags, int rank, int world,
                       unsigned tag, float* Gt, unsigned* gflg, unsigned* bvflg, cudaStream_t st);
static int g_fused_mode = 0; // 0: whenever the call qualifies, 1: never (two-kernel path)

static int g_L = 0;          // 0: choose per call
static int g_k2_threads = 0; // 0: choose per call
static long long* g_dbg = nullptr;  // device buffer for phase timestamps (debug)

// optional per-kernel timing (bench.py's roofline leg): events around each launch of the last profiled call, kept per
// device (an event belongs to the device it was created on)
static bool g_prof = false;
struct ProfEvents { cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}; int n = 0; };
static ProfEvents g_pe[kMaxDevices];
static int g_prof_dev = 0;   // device of the last profiled call

static int prof_events(ProfEvents** out) {
  ProfEvents& p = g_pe[current_device()];
  std::lock_guard<std::mutex> lk(state_mutex());
  if (!p.ev[0])
    for (int i = 0; i < 5; ++i) HMM_CUDA_OK(cudaEventCreate(&p.ev[i]));
  *out = &p;
  return 0;
}
int profile_enable(int on) {
  g_prof = on != 0;
  for (int d = 0; d < kMaxDevices; ++d) g_pe[d].n = 0;
  return 0;
}
int profile_read(float* ms, int n) {
  ProfEvents& p = g_pe[g_prof_dev];
  if (p.n < 2) return 0;
  HMM_CUDA_OK(cudaEventSynchronize(p.ev[p.n - 1]));
  int k = 0;
  for (; k < p.n - 1 && k < n; ++k) HMM_CUDA_OK(cudaEventElapsedTime(&ms[k], p.ev[k], p.ev[k + 1]));
  return k;
}

static int num_sms() { return device_sm_count(); }

static int max_L(bool full) {
  int L = MAX_L;
  while (L > 1 && (k2_smem_bytes(L, full) > 227 * 1024 || k1_smem_bytes(L, MAX_SCAN_TILES) > 227 * 1024)) --L;
  return L;
}

// resident CTAs per SM by shared memory (228 KB per SM, 1 KB reserved per CTA), capped by the kernel's thread budget
static int ctas_per_sm(size_t smem_bytes, int cap) {
  const int c = (int)((size_t)228 * 1024 / (smem_bytes + 1024));
  return c < 1 ? 1 : (c > cap ? cap : c);
}

// Chunk length from a cost model (cycles per SM): both kernels run ceil(tiles / resident CTAs) rounds of a per-tile
// latency chain whose length grows with L (emission + rows + recursion steps) plus a fixed part (loads, tree, scans);
// the tree work per position falls with L.  Coefficients fitted to profiles/r1c_sweep.txt.
static int choose_L(int NB, int T, bool full) {
  const int Lmax = max_L(full);
  if (g_L > 0) return g_L < Lmax ? g_L : Lmax;
  const int sms = num_sms();
  double best = 1e300;
  int bestL = Lmax;
  for (int L = 4; L <= Lmax; ++L) {
    const int P = cdiv(cdiv(T, L), G);
    const long long tiles = (long long)NB * P;
    const int c1 = ctas_per_sm(k1_smem_bytes(L, P), HMM_K1_BLOCKS), c2 = ctas_per_sm(k2_smem_bytes(L, full), 4);
    const long long r1 = (tiles + (long long)sms * c1 - 1) / ((long long)sms * c1);
    const long long r2 = (tiles + (long long)sms * c2 - 1) / ((long long)sms * c2);
    // CTAs that actually share an SM: fewer tiles than SMs run alone whatever the occupancy limit says.  Co-resident
    // CTAs serialise their per-position work and overlap about half of their fixed part (r1i sweeps: B = 8 / 32 / 64)
    const long long per_sm = (tiles + sms - 1) / sms;
    const int e1 = (int)(per_sm < c1 ? per_sm : c1), e2 = (int)(per_sm < c2 ? per_sm : c2);
    const double t1 = 7000.0 * (1.0 + 0.5 * (e1 - 1)) + 330.0 * L * e1, t2 = 4500.0 + 200.0 * L + 60.0 * L * (e2 - 1);
    const double cost = (double)r1 * t1 + (double)r2 * t2;
    if (cost < best) { best = cost; bestL = L; }
  }
  return bestL;
}

struct Layout {
  size_t Mh, Mt, bvf, bvr, llws, inf, inr, cnt, flg, Gt, gvf, gvr, gflg, bvflg, Ews, total;
};
static Layout layout(int NB, int P, int L) {
  Layout l;
  size_t off = 0;
  const size_t NP = (size_t)NB * P;
  l.Mh = off;   off = align_up(off + NP * HEAP * MAT * sizeof(float), 256);
  l.Mt = off;   off = align_up(off + NP * MAT * sizeof(float), 256);
  l.bvf = off;  off = align_up(off + NP * QP * sizeof(float), 256);
  l.bvr = off;  off = align_up(off + NP * QP * sizeof(float), 256);
  l.llws = off; off = align_up(off + (size_t)NB * sizeof(float), 256);
  l.inf = off;  off = align_up(off + (size_t)NB * QP * sizeof(float), 256);
  l.inr = off;  off = align_up(off + (size_t)NB * QP * sizeof(float), 256);
  l.cnt = off;  off = align_up(off + (size_t)NB * sizeof(unsigned), 256);
  l.flg = off;  off = align_up(off + NP * sizeof(unsigned), 256);
  const size_t NG = (size_t)NB * cdiv(P, G);   // tile groups of the two-level boundary scan
  l.Gt = off;   off = align_up(off + NG * MAT * sizeof(float), 256);
  l.gvf = off;  off = align_up(off + NG * QP * sizeof(float), 256);
  l.gvr = off;  off = align_up(off + NG * QP * sizeof(float), 256);
  l.gflg = off; off = align_up(off + NG * sizeof(unsigned), 256);     // single-launch shard mode: group matrix published
  l.bvflg = off; off = align_up(off + NG * sizeof(unsigned), 256);    //                          boundary vectors published
  l.Ews = off;  off = align_up(off + NP * G * ((size_t)L * QP + 4) * sizeof(float), 256);
  l.total = off;
  return l;
}
static Layout layout_for(int NB, int T, int L) { return layout(NB, cdiv(cdiv(T, L), G), L); }

// Chunk length of the single-launch path: every tile of the call must be resident at once (a tile waits inside the
// kernel for the other tiles of its sequence), so the smallest chunk length whose tile count fits the device; 0 = the
// call does not qualify.  A tuning override (set_tuning) is honoured when it fits.
constexpr int FUSED_MIN_L = 12, FUSED_MAX_TILES_PER_SEQ = 64, FUSED_SCAN_SLOTS = 24;   // (FSCAN of gene15_fused.cu)
static int choose_fused_L(int NB, int T) {
  if (g_fused_mode == 1) return 0;
  static const bool env_off = getenv("HMM_B200_FUSED") && atoi(getenv("HMM_B200_FUSED")) == 0;
  if (env_off) return 0;
  const int sms = num_sms();
  // cost model (kilo-cycles per tile, fused_phases.py on B200): a tile's phases are latency chains of about 44 + 0.85 L
  // when its CTA has the SM to itself, 1.36 x that when two tiles share an SM; more tiles than resident CTAs do not fit
  double best = 1e300;
  int bestL = 0;
  for (int L = (g_L > 0 ? g_L : FUSED_MIN_L); L <= (g_L > 0 ? g_L : MAX_L); ++L) {
    const int cap = fused_capacity(L);
    if (cap <= 0) break;             // shared memory: longer chunks do not fit either
    const int P = cdiv(cdiv(T, L), G);
    const long long tiles = (long long)NB * P;
    if (P > FUSED_MAX_TILES_PER_SEQ || tiles > cap) continue;
    const double cost = (44.0 + 0.85 * L) * (tiles <= sms ? 1.0 : 1.36);
    if (cost < best) { best = cost; bestL = L; }
  }
  return bestL;
}

size_t workspace_size(int num_models, int batch, int T) {
  if (num_models <= 0 || batch <= 0 || T <= 0) return 0;
  // run() picks the chunk length from (NB, T, which outputs): cover every choice
  const int NB = num_models * batch;
  size_t best = layout_for(NB, T, choose_L(NB, T, false)).total;
  const size_t b = layout_for(NB, T, choose_L(NB, T, true)).total;
  if (b > best) best = b;
  for (int L = 4; L <= MAX_L; ++L) {   // whichever chunk length the single-launch path may take (mode / tuning may change)
    const size_t f = layout_for(NB, T, L).total;
    if (f > best) best = f;
  }
  return best;
}

int set_fused_mode(int mode) {
  if (mode < 0 || mode > 1) { set_error("fused mode must be 0 (auto) or 1 (off)"); return HMM_B200_EINVAL; }
  g_fused_mode = mode;
  return 0;
}

int set_debug_buffer(void* p) { g_dbg = static_cast<long long*>(p); hmm::dense_tc::set_debug(p); return 0; }

int set_tuning(int L, int k2_threads, int dense_path) {
  hmm::dense::set_path(dense_path);
  if (L < 0 || L > MAX_L || k2_threads < 0 || k2_threads > 512 || (k2_threads & 63) || (k2_threads > 0 && k2_threads < 128)) {
    set_error("tuning rejected: chunk_len in [0,%d], finalize threads a multiple of 64 in [128,512] or 0", MAX_L);
    return HMM_B200_EINVAL;
  }
  g_L = L;
  g_k2_threads = k2_threads;
  return 0;
}

static int fill_args(Args& a, const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                     const float* Bfull, const float* codon5, const float* hints, unsigned flags) {
  HMM_REQUIRE(Bfull && codon5, "genepred: null input pointer");
  HMM_REQUIRE(num_models > 0 && batch > 0 && T > 0, "genepred: num_models, batch and T must be positive (got %d,%d,%d)",
              num_models, batch, T);
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(inputs) & 15) == 0, "genepred: inputs must be 16-byte aligned");
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(codon5) & 15) == 0, "genepred: codon5 must be 16-byte aligned");
  HMM_REQUIRE((long long)num_models * batch <= 65535, "genepred: num_models*batch must be <= 65535");
  a.inputs = inputs; a.A = A; a.pi = pi; a.Bfull = Bfull; a.codon5 = codon5; a.hints = hints;
  a.num_models = num_models; a.batch = batch; a.T = T; a.flags = flags;
  a.seq_off = 0; a.seq_total = T; a.halo = nullptr; a.dbg = g_dbg;
  a.in_f = nullptr; a.in_r = nullptr;
  a.cls16 = nullptr; a.nuc8 = nullptr;
  return 0;
}

static void set_L(Args& a, int L) {
  a.L = L;
  a.mL = (1u << DIV_SHIFT) / (unsigned)L + 1u;
  a.C = cdiv(a.T, L);
  a.P = cdiv(a.C, G);
}

// raises a kernel's dynamic shared memory limit when this call needs more than any earlier one on this device
template <typename K>
static int set_smem(K kernel, int slot, size_t bytes) { return ensure_dynamic_smem(kernel, slot, bytes); }
enum { SLOT_K1 = 0, SLOT_K2P = 1, SLOT_K2F = 2, SLOT_KB = 3, SLOT_EM = 4, SLOT_GRAD = 8 };

static int k2_threads_for(int L) {
  if (g_k2_threads > 0) return g_k2_threads;
  (void)L;
  return 256;   // 2 recursion warps + 6 more for the output stage; several CTAs share an SM
}

// workspace pointers for a chunk length already set in `a`
static int bind_workspace(Args& a, Layout& l, int NB, void* workspace, size_t workspace_bytes, const char* who) {
  l = layout(NB, a.P, a.L);
  if (!workspace || workspace_bytes < l.total) {
    set_error("%s: workspace too small (%zu < %zu bytes)", who, workspace_bytes, l.total);
    return HMM_B200_EWORKSPACE;
  }
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "%s: workspace must be 256-byte aligned", who);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  a.Mh = reinterpret_cast<float*>(ws + l.Mh);
  a.Mt = reinterpret_cast<float*>(ws + l.Mt);
  a.bvf = reinterpret_cast<float*>(ws + l.bvf); a.bvr = reinterpret_cast<float*>(ws + l.bvr);
  a.llws = reinterpret_cast<float*>(ws + l.llws);
  a.cnt = reinterpret_cast<unsigned*>(ws + l.cnt);
  a.flg = reinterpret_cast<unsigned*>(ws + l.flg);
  a.Ews = reinterpret_cast<float*>(ws + l.Ews);
  return 0;
}

int prepare_codon(const float* codon64, float* codon5, cudaStream_t st) {
  HMM_REQUIRE(codon64 && codon5, "prepare_codon: null pointer");
  prepare_codon_kernel<<<1, 256, 0, st>>>(codon64, codon5);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

enum { SLOT_K1B = 7 };
static int launch_products(const Args& a, int NB, cudaStream_t st) {
  const size_t sm1 = k1_smem_bytes(a.L, a.P);
  if (ctas_per_sm(sm1, HMM_K1_BLOCKS) >= 2) {
    int rc = set_smem(products_kernel<2>, SLOT_K1, sm1); if (rc) return rc;
    products_kernel<2><<<dim3(a.P, NB), K1_THREADS, sm1, st>>>(a);
  } else {
    int rc = set_smem(products_kernel<1>, SLOT_K1B, sm1); if (rc) return rc;
    products_kernel<1><<<dim3(a.P, NB), K1_THREADS, sm1, st>>>(a);
  }
  return 0;
}

// Boundary vectors of every tile (mode 0) or the product of all tile matrices (mode 1) after products_kernel.
// Long tile lists go through the two-level scan (K_G); `groups_ready`: Gt still holds this call pair's group matrices.
constexpr int GROUP_MIN_TILES = 96;
enum { SLOT_KG1 = 5, SLOT_KG2 = 6 };
static int launch_boundary(const Args& a, const Layout& l, int NB, int mode, float* total_out, bool groups_ready, void* workspace,
                           cudaStream_t st) {
  int rc = set_smem(boundary_kernel, SLOT_KB, sizeof(KBSmem)); if (rc) return rc;
  if (a.P < GROUP_MIN_TILES) {
    boundary_kernel<<<NB, KB_WARPS * 32, sizeof(KBSmem), st>>>(a, mode, total_out);
    return 0;
  }
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  float* Gt = reinterpret_cast<float*>(ws + l.Gt);
  if (mode == 0 && !groups_ready && a.in_f == nullptr && a.in_r == nullptr) {
    // one cooperative launch instead of three kernels when the groups fit the device at once (gene15_fused.cu)
    rc = launch_boundary_groups(a, NB, Gt, reinterpret_cast<unsigned*>(ws + l.gflg), st);
    if (rc != FUSED_NOT_RESIDENT) return rc;
  }
  float* gvf = reinterpret_cast<float*>(ws + l.gvf);
  float* gvr = reinterpret_cast<float*>(ws + l.gvr);
  const int PG = cdiv(a.P, G);
  const size_t smg = (size_t)HEAP * MAT * sizeof(float), smd = ((size_t)G * BSZ + QP) * sizeof(float);
  rc = set_smem(group_reduce_kernel, SLOT_KG1, smg); if (rc) return rc;
  rc = set_smem(group_down_kernel, SLOT_KG2, smd); if (rc) return rc;
  if (!groups_ready) group_reduce_kernel<<<dim3(PG, NB), K1_THREADS, smg, st>>>(a, Gt);
  Args top = a;
  top.Mt = Gt; top.P = PG; top.bvf = gvf; top.bvr = gvr;
  boundary_kernel<<<NB, KB_WARPS * 32, sizeof(KBSmem), st>>>(top, mode, total_out);
  if (mode == 0) group_down_kernel<<<dim3(PG, NB), 256, smd, st>>>(a, gvf, gvr);
  return 0;
}

// K2 launch.  `pdl`: programmatic dependent launch behind the kernel just queued on `st` -- K2's CTAs are placed and run
// their prologue while that kernel drains; everything that reads its results sits behind griddepcontrol.wait.
static int launch_finalize(const Args& a, int NB, bool full, size_t sm2, cudaStream_t st, bool pdl) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(a.P, NB);
  cfg.blockDim = dim3(k2_threads_for(a.L));
  cfg.dynamicSmemBytes = sm2;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  if (full) HMM_CUDA_OK(cudaLaunchKernelEx(&cfg, finalize_kernel<true>, a));
  else HMM_CUDA_OK(cudaLaunchKernelEx(&cfg, finalize_kernel<false>, a));
  return 0;
}

int run(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
        const float* codon5, const float* hints, unsigned flags, float* out_post, float* out_alpha, float* out_beta,
        float* loglik, void* workspace, size_t workspace_bytes, cudaStream_t st, const unsigned short* cls16 = nullptr,
        const unsigned char* nuc8 = nullptr) {
  Args a{};
  int rc = fill_args(a, inputs, num_models, batch, T, A, pi, Bfull, codon5, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE((inputs != nullptr) != (cls16 != nullptr), "genepred: give either fp32 inputs or the compact pair");
  HMM_REQUIRE(!cls16 || nuc8, "genepred: compact inputs need the nucleotide codes");
  a.cls16 = cls16; a.nuc8 = nuc8;
  HMM_REQUIRE(A && pi, "genepred: null A / pi");
  HMM_REQUIRE(out_post || out_alpha || out_beta || loglik, "genepred: no output requested");
  const int NB = num_models * batch;
  const bool full = out_alpha || out_beta;
  // posterior (+ log-likelihood) from fp32 inputs: the single-launch kernel when every tile of a launch fits the device
  // at once; larger batches go as a few back-to-back launches over slices of the sequences (they are independent)
  if (out_post && !full && inputs) {
    int nb = NB, Lf = choose_fused_L(nb, T);
    for (int parts = 2; Lf == 0 && parts <= 8 && NB / parts >= 8; ++parts) {
      nb = cdiv(NB, parts);
      Lf = choose_fused_L(nb, T);
    }
    if (Lf > 0 && (nb == NB || num_models == 1 || batch % nb == 0)) {   // a slice must not straddle two models
      ProfEvents* pe = nullptr;
      if (g_prof) { rc = prof_events(&pe); if (rc) return rc; g_prof_dev = current_device(); pe->n = 0; cudaEventRecord(pe->ev[pe->n++], st); }
      bool fused_refused = false;
      for (int n0 = 0; n0 < NB; n0 += nb) {
        const int cnt = NB - n0 < nb ? NB - n0 : nb;
        Args s = a;
        const int m0 = n0 / batch;                 // model of this slice (slices never straddle models, see above)
        s.inputs = inputs + (size_t)n0 * T * SIN;
        s.hints = hints ? hints + (size_t)n0 * 2 * Q : nullptr;
        s.A = A + (size_t)m0 * Q * Q; s.pi = pi + (size_t)m0 * Q; s.Bfull = Bfull + (size_t)m0 * Q * Q;
        s.num_models = 1; s.batch = cnt;
        if (nb == NB) { s.A = A; s.pi = pi; s.Bfull = Bfull; s.num_models = num_models; s.batch = batch; }
        set_L(s, Lf);
        s.top_scan = 2;
        Layout lf;
        rc = bind_workspace(s, lf, cnt, workspace, workspace_bytes, "genepred");
        if (rc) return rc;
        s.out_post = out_post + (size_t)n0 * T * Q; s.out_alpha = nullptr; s.out_beta = nullptr;
        s.loglik = loglik ? loglik + n0 : nullptr;
        rc = launch_fused(s, cnt, st);
        if (rc == FUSED_NOT_RESIDENT && n0 == 0) { fused_refused = true; break; }   // nothing launched yet: the two-kernel path below
        if (rc) return rc == FUSED_NOT_RESIDENT ? HMM_B200_ECUDA : rc;
      }
      if (!fused_refused) {
        if (pe) cudaEventRecord(pe->ev[pe->n++], st);
        return 0;
      }
      if (pe) pe->n = 0;
    }
  }
  set_L(a, choose_L(NB, T, full));
  const bool need_k2 = out_post || out_alpha || out_beta;
  // short tile lists are scanned by K1's last CTA of every sequence, long ones by the boundary kernel
  a.top_scan = (need_k2 && a.P <= MAX_K2_SCAN) ? 2 : (a.P <= MAX_SCAN_TILES) ? 1 : 0;
  Layout l;
  rc = bind_workspace(a, l, NB, workspace, workspace_bytes, "genepred");
  if (rc) return rc;
  if (!need_k2) a.Ews = nullptr;   // log-likelihood only: nobody reads the emission rows
  a.out_post = out_post; a.out_alpha = out_alpha; a.out_beta = out_beta; a.loglik = loglik;

  const size_t sm1 = k1_smem_bytes(a.L, a.P), sm2 = k2_smem_bytes(a.L, full);
  if (need_k2 && !full) { rc = set_smem(finalize_kernel<false>, SLOT_K2P, sm2); if (rc) return rc; }
  if (need_k2 && full) { rc = set_smem(finalize_kernel<true>, SLOT_K2F, sm2); if (rc) return rc; }
  if (a.top_scan == 1) HMM_CUDA_OK(cudaMemsetAsync(a.cnt, 0, (size_t)NB * sizeof(unsigned), st));

  ProfEvents* pe = nullptr;
  if (g_prof) { rc = prof_events(&pe); if (rc) return rc; g_prof_dev = current_device(); pe->n = 0; cudaEventRecord(pe->ev[pe->n++], st); }
  rc = launch_products(a, NB, st); if (rc) return rc;
  if (pe) cudaEventRecord(pe->ev[pe->n++], st);
  if (!a.top_scan) {
    rc = launch_boundary(a, l, NB, 0, nullptr, false, workspace, st); if (rc) return rc;
    if (pe) cudaEventRecord(pe->ev[pe->n++], st);
  }
  if (need_k2) {
    // directly behind products_kernel and not timed kernel by kernel: as a programmatic dependent launch
    rc = launch_finalize(a, NB, full, sm2, st, a.top_scan != 0 && !pe); if (rc) return rc;
    if (pe) cudaEventRecord(pe->ev[pe->n++], st);
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- host buffers: batch-sliced H2D -> kernels -> D2H pipeline --------------------------------------------------------
// Sequences are independent, so the batch is cut into slices; slice i's kernels run on the caller's stream while slice
// i+1's inputs come up and slice i-1's outputs go down on two copy streams (PCIe is full duplex).  Device memory comes
// from the caller (`scratch`, host_scratch_size); the copy streams and events are created once per thread and device.
struct HostPipe {
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_k[2] = {nullptr, nullptr}, ev_out[2] = {nullptr, nullptr};
  cudaEvent_t ev_start = nullptr;
  bool ready = false;
};
static int host_pipe(HostPipe** out) {
  static thread_local HostPipe pipes[64];
  int dev = 0;
  HMM_CUDA_OK(cudaGetDevice(&dev));
  HMM_REQUIRE(dev >= 0 && dev < 64, "genepred run_host: device index %d out of range", dev);
  HostPipe& p = pipes[dev];
  if (!p.ready) {
    HMM_CUDA_OK(cudaStreamCreateWithFlags(&p.s_in, cudaStreamNonBlocking));
    HMM_CUDA_OK(cudaStreamCreateWithFlags(&p.s_out, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      HMM_CUDA_OK(cudaEventCreateWithFlags(&p.ev_in[i], cudaEventDisableTiming));
      HMM_CUDA_OK(cudaEventCreateWithFlags(&p.ev_k[i], cudaEventDisableTiming));
      HMM_CUDA_OK(cudaEventCreateWithFlags(&p.ev_out[i], cudaEventDisableTiming));
    }
    HMM_CUDA_OK(cudaEventCreateWithFlags(&p.ev_start, cudaEventDisableTiming));
    p.ready = true;
  }
  *out = &p;
  return 0;
}

static int g_host_slices = 0;   // 0 = by batch size; HMM_HOST_SLICES overrides (experiments)
static int host_slice_len(int batch) {
  static const int env = getenv("HMM_HOST_SLICES") ? atoi(getenv("HMM_HOST_SLICES")) : 0;
  int ns = env > 0 ? env : g_host_slices;
  // measured on config 2 (profiles/r2n_e2e_breakdown.txt, r2n_direct.txt): 2 to 6 slices, uniform or ramped towards the
  // ends, land within 2 % (0.735-0.755 ms), 8 are slower (every cross-stream hop costs microseconds).  Letting the kernels
  // store straight into the caller's pinned buffers instead of the download copies does not pay either: the stores run at
  // 51 GB/s alone but starve the concurrent upload (copy-engine traffic both ways arbitrates better): 0.725 ms at best.
  if (ns <= 0) ns = batch >= 16 ? 4 : (batch >= 4 ? 2 : 1);
  if (ns > batch) ns = batch;
  return cdiv(batch, ns);
}

struct HostLayout {
  size_t in[2], post[2], alpha[2], beta[2], ws[2], ll, hints, total;
  size_t ws_bytes;
  int slice;
};
static HostLayout host_layout(int num_models, int batch, int T) {
  HostLayout h;
  h.slice = host_slice_len(batch);
  const size_t rows = (size_t)h.slice * T;
  h.ws_bytes = workspace_size(1, h.slice, T);
  size_t off = 0;
  for (int b = 0; b < 2; ++b) { h.in[b] = off; off = align_up(off + rows * SIN * sizeof(float), 256); }
  for (int b = 0; b < 2; ++b) { h.post[b] = off; off = align_up(off + rows * Q * sizeof(float), 256); }
  for (int b = 0; b < 2; ++b) { h.alpha[b] = off; off = align_up(off + rows * Q * sizeof(float), 256); }
  for (int b = 0; b < 2; ++b) { h.beta[b] = off; off = align_up(off + rows * Q * sizeof(float), 256); }
  for (int b = 0; b < 2; ++b) { h.ws[b] = off; off = align_up(off + h.ws_bytes, 256); }
  h.ll = off; off = align_up(off + (size_t)num_models * batch * sizeof(float), 256);
  h.hints = off; off = align_up(off + (size_t)num_models * batch * 2 * Q * sizeof(float), 256);
  h.total = off;
  return h;
}
size_t host_scratch_size(int num_models, int batch, int T) {
  if (num_models <= 0 || batch <= 0 || T <= 0) return 0;
  return host_layout(num_models, batch, T).total;
}

int run_host(const float* h_inputs, int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
             const float* codon5, const float* h_hints, unsigned flags, float* h_post, float* h_alpha, float* h_beta,
             float* h_loglik, void* scratch, size_t scratch_bytes, cudaStream_t st, const unsigned short* h_cls16 = nullptr,
             const unsigned char* h_nuc8 = nullptr) {
  HMM_REQUIRE(((h_inputs != nullptr) != (h_cls16 != nullptr)) && (!h_cls16 || h_nuc8) && A && pi && Bfull && codon5,
              "genepred run_host: null input pointer (fp32 inputs, or the compact pair)");
  HMM_REQUIRE(num_models > 0 && batch > 0 && T > 0, "genepred run_host: num_models, batch and T must be positive");
  HMM_REQUIRE(h_post || h_alpha || h_beta || h_loglik, "genepred run_host: no output requested");
  const HostLayout h = host_layout(num_models, batch, T);
  if (!scratch || scratch_bytes < h.total) {
    set_error("genepred run_host: device scratch too small (%zu < %zu bytes)", scratch_bytes, h.total);
    return HMM_B200_EWORKSPACE;
  }
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(scratch) & 255) == 0, "genepred run_host: scratch must be 256-byte aligned");
  HostPipe* pp = nullptr;
  int rc = host_pipe(&pp); if (rc) return rc;
  HostPipe& p = *pp;
  unsigned char* ws = static_cast<unsigned char*>(scratch);
  auto F = [&](size_t o) { return reinterpret_cast<float*>(ws + o); };
  float* d_ll = F(h.ll);
  float* d_hints = h_hints ? F(h.hints) : nullptr;
  // everything queued on the copy streams starts after the work already on the caller's stream
  HMM_CUDA_OK(cudaEventRecord(p.ev_start, st));
  HMM_CUDA_OK(cudaStreamWaitEvent(p.s_in, p.ev_start, 0));
  HMM_CUDA_OK(cudaStreamWaitEvent(p.s_out, p.ev_start, 0));
  if (h_hints)
    HMM_CUDA_OK(cudaMemcpyAsync(d_hints, h_hints, (size_t)num_models * batch * 2 * Q * sizeof(float), cudaMemcpyHostToDevice, p.s_in));
  int it = 0;
  for (int m = 0; m < num_models; ++m) {
    for (int b0 = 0; b0 < batch; b0 += h.slice, ++it) {
      const int nb = batch - b0 < h.slice ? batch - b0 : h.slice, buf = it & 1;
      const size_t seq0 = (size_t)m * batch + b0, rows = (size_t)nb * T;
      if (it >= 2) HMM_CUDA_OK(cudaStreamWaitEvent(p.s_in, p.ev_k[buf], 0));      // the kernels of slice it-2 read this buffer
      // compact wire format: 30 + 1 B / position up instead of 80; both pieces land in the slice's input buffer
      unsigned short* d_cls = reinterpret_cast<unsigned short*>(F(h.in[buf]));
      unsigned char* d_nuc = reinterpret_cast<unsigned char*>(F(h.in[buf])) + align_up(rows * Q * sizeof(unsigned short), 256);
      if (h_cls16) {
        HMM_CUDA_OK(cudaMemcpyAsync(d_cls, h_cls16 + seq0 * T * Q, rows * Q * sizeof(unsigned short), cudaMemcpyHostToDevice, p.s_in));
        HMM_CUDA_OK(cudaMemcpyAsync(d_nuc, h_nuc8 + seq0 * T, rows, cudaMemcpyHostToDevice, p.s_in));
      } else {
        HMM_CUDA_OK(cudaMemcpyAsync(F(h.in[buf]), h_inputs + seq0 * T * SIN, rows * SIN * sizeof(float), cudaMemcpyHostToDevice, p.s_in));
      }
      HMM_CUDA_OK(cudaEventRecord(p.ev_in[buf], p.s_in));
      HMM_CUDA_OK(cudaStreamWaitEvent(st, p.ev_in[buf], 0));
      if (it >= 2) HMM_CUDA_OK(cudaStreamWaitEvent(st, p.ev_out[buf], 0));        // slice it-2's outputs have left
      rc = run(h_cls16 ? nullptr : F(h.in[buf]), 1, nb, T, A + (size_t)m * Q * Q, pi + (size_t)m * Q, Bfull + (size_t)m * Q * Q, codon5,
               d_hints ? d_hints + seq0 * 2 * Q : nullptr, flags, h_post ? F(h.post[buf]) : nullptr,
               h_alpha ? F(h.alpha[buf]) : nullptr, h_beta ? F(h.beta[buf]) : nullptr, d_ll + seq0, ws + h.ws[buf], h.ws_bytes, st,
               h_cls16 ? d_cls : nullptr, h_cls16 ? d_nuc : nullptr);
      if (rc) return rc;
      HMM_CUDA_OK(cudaEventRecord(p.ev_k[buf], st));
      HMM_CUDA_OK(cudaStreamWaitEvent(p.s_out, p.ev_k[buf], 0));
      if (h_post) HMM_CUDA_OK(cudaMemcpyAsync(h_post + seq0 * T * Q, F(h.post[buf]), rows * Q * sizeof(float), cudaMemcpyDeviceToHost, p.s_out));
      if (h_alpha) HMM_CUDA_OK(cudaMemcpyAsync(h_alpha + seq0 * T * Q, F(h.alpha[buf]), rows * Q * sizeof(float), cudaMemcpyDeviceToHost, p.s_out));
      if (h_beta) HMM_CUDA_OK(cudaMemcpyAsync(h_beta + seq0 * T * Q, F(h.beta[buf]), rows * Q * sizeof(float), cudaMemcpyDeviceToHost, p.s_out));
      HMM_CUDA_OK(cudaEventRecord(p.ev_out[buf], p.s_out));
    }
  }
  if (h_loglik) {
    HMM_CUDA_OK(cudaStreamWaitEvent(p.s_out, p.ev_k[(it - 1) & 1], 0));
    HMM_CUDA_OK(cudaMemcpyAsync(h_loglik, d_ll, (size_t)num_models * batch * sizeof(float), cudaMemcpyDeviceToHost, p.s_out));
    HMM_CUDA_OK(cudaEventRecord(p.ev_out[(it - 1) & 1], p.s_out));
  }
  // completion is ordered on the caller's stream: it waits for the last copies of both buffers
  HMM_CUDA_OK(cudaStreamWaitEvent(st, p.ev_out[(it - 1) & 1], 0));
  if (it >= 2) HMM_CUDA_OK(cudaStreamWaitEvent(st, p.ev_out[it & 1], 0));
  return 0;
}

// ---- sequence sharding (multi-GPU): this rank holds positions [seq_off, seq_off + T) of every sequence -----------
static int shard_args(Args& a, Layout& l, const float* inputs, int num_models, int batch, int T, int seq_off, int seq_total,
                      const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                      const float* hints, unsigned flags, void* workspace, size_t workspace_bytes) {
  int rc = fill_args(a, inputs, num_models, batch, T, A, pi, Bfull, codon5, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE(inputs && A && pi, "genepred shard: null inputs / A / pi");
  HMM_REQUIRE(seq_off >= 0 && seq_total >= seq_off + T, "genepred shard: positions [%d, %d) outside a sequence of length %d",
              seq_off, seq_off + T, seq_total);
  HMM_REQUIRE(halo || (seq_off == 0 && seq_total == T), "genepred shard: a proper shard needs the 2+2 halo nucleotides");
  a.seq_off = seq_off; a.seq_total = seq_total; a.halo = halo;
  const int NB = num_models * batch;
  set_L(a, choose_L(NB, T, true));   // one chunk length for both calls, whatever outputs are requested later
  a.top_scan = 0;   // the scan crosses the shards: boundary_kernel with the gathered blocks
  rc = bind_workspace(a, l, NB, workspace, workspace_bytes, "genepred shard");
  if (rc) return rc;
  return 0;
}

int shard_reduce(const float* inputs, int num_models, int batch, int T, int seq_off, int seq_total, const float* halo,
                 const float* A, const float* pi, const float* Bfull, const float* codon5, const float* hints, unsigned flags,
                 float* block_out, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  Layout l;
  int rc = shard_args(a, l, inputs, num_models, batch, T, seq_off, seq_total, halo, A, pi, Bfull, codon5, hints, flags, workspace,
                      workspace_bytes);
  if (rc) return rc;
  HMM_REQUIRE(block_out, "genepred shard_reduce: null output");
  const int NB = num_models * batch;
  const size_t sm1 = k1_smem_bytes(a.L, a.P);
  rc = launch_products(a, NB, st); if (rc) return rc;
  rc = launch_boundary(a, l, NB, 1, block_out, false, workspace, st); if (rc) return rc;
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int shard_finalize(const float* inputs, int num_models, int batch, int T, int seq_off, int seq_total, const float* halo,
                   const float* A, const float* pi, const float* Bfull, const float* codon5, const float* hints, unsigned flags,
                   const float* all_blocks, int rank, int world, float* out_post, float* out_alpha, float* out_beta,
                   float* loglik, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  Layout l;
  int rc = shard_args(a, l, inputs, num_models, batch, T, seq_off, seq_total, halo, A, pi, Bfull, codon5, hints, flags, workspace,
                      workspace_bytes);
  if (rc) return rc;
  HMM_REQUIRE(all_blocks && world > 0 && rank >= 0 && rank < world, "genepred shard_finalize: bad rank %d of %d", rank, world);
  HMM_REQUIRE(out_post || out_alpha || out_beta || loglik, "genepred shard_finalize: no output requested");
  const int NB = num_models * batch;
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  float* in_f = reinterpret_cast<float*>(ws + l.inf);
  float* in_r = reinterpret_cast<float*>(ws + l.inr);
  shard_boundary_kernel<<<NB, 32, 0, st>>>(all_blocks, rank, world, NB, batch, pi, in_f, in_r, a.llws, loglik);
  a.in_f = in_f; a.in_r = in_r; a.ll_external = 1;
  a.out_post = out_post; a.out_alpha = out_alpha; a.out_beta = out_beta; a.loglik = nullptr;
  if (out_post || out_alpha || out_beta) {
    const bool full = out_alpha || out_beta;
    const size_t sm2 = k2_smem_bytes(a.L, full);
    rc = full ? set_smem(finalize_kernel<true>, SLOT_K2F, sm2) : set_smem(finalize_kernel<false>, SLOT_K2P, sm2); if (rc) return rc;
    rc = launch_boundary(a, l, NB, 0, nullptr, true, workspace, st); if (rc) return rc;   // Gt: from shard_reduce
    rc = launch_finalize(a, NB, full, sm2, st, false); if (rc) return rc;
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- sequence sharding with the exchange inside the kernels (peer memory): 4 launches per step --------------------
// peer_blocks / peer_flags: `world` device pointers each (host arrays), the block / flag tables of every rank in rank
// order, mapped into this process (CUDA IPC or symmetric memory); NULL = no peer stores: the block goes to block_out and
// the caller gathers (NCCL).  `tag` = step counter, the same on every rank, never 0.
int shard_reduce_x(const float* inputs, int num_models, int batch, int T, int seq_off, int seq_total, const float* halo,
                   const float* A, const float* pi, const float* Bfull, const float* codon5, const float* hints, unsigned flags,
                   float* block_out, const unsigned long long* peer_blocks, const unsigned long long* peer_flags, int rank,
                   int world, unsigned tag, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  Layout l;
  int rc = shard_args(a, l, inputs, num_models, batch, T, seq_off, seq_total, halo, A, pi, Bfull, codon5, hints, flags, workspace,
                      workspace_bytes);
  if (rc) return rc;
  HMM_REQUIRE(block_out || peer_blocks, "genepred shard_reduce_x: neither an output block nor peer tables");
  HMM_REQUIRE(!peer_blocks || (peer_flags && world > 0 && world <= MAX_PEERS && rank >= 0 && rank < world && tag != 0),
              "genepred shard_reduce_x: bad peer tables (world %d, rank %d, tag %u)", world, rank, tag);
  const int NB = num_models * batch, PG = cdiv(a.P, G);
  Peers pr{};
  if (peer_blocks) {
    for (int r = 0; r < world; ++r) {
      pr.blocks[r] = reinterpret_cast<float*>(static_cast<uintptr_t>(peer_blocks[r]));
      pr.flags[r] = reinterpret_cast<unsigned*>(static_cast<uintptr_t>(peer_flags[r]));
    }
    pr.rank = rank; pr.world = world; pr.tag = tag;
  }
  rc = launch_products(a, NB, st); if (rc) return rc;
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  if (PG > G) {
    // more than G * G tiles per rank: the general boundary kernels reduce the shard, a small kernel publishes the block
    float* blk = block_out ? block_out : reinterpret_cast<float*>(ws + l.bvf);   // (bvf is written only by finalize_x)
    rc = launch_boundary(a, l, NB, 1, blk, false, workspace, st); if (rc) return rc;
    if (peer_blocks) shard_publish_kernel<<<NB, 64, 0, st>>>(blk, pr, NB);
    HMM_CUDA_OK(cudaGetLastError());
    return 0;
  }
  const size_t smg = (size_t)HEAP * MAT * sizeof(float);
  rc = ensure_dynamic_smem(shard_product_kernel, 24, smg); if (rc) return rc;
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(PG, NB); cfg.blockDim = dim3(K1_THREADS); cfg.dynamicSmemBytes = smg; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    HMM_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_product_kernel, a, reinterpret_cast<float*>(ws + l.Gt), pr, block_out));
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

// all_blocks: (world, NB, 256) -- this rank's block table half for `tag` (peer route; then block_flags = the matching
// flag table half) or the NCCL-gathered blocks (block_flags NULL: nothing to wait for)
int shard_finalize_x(const float* inputs, int num_models, int batch, int T, int seq_off, int seq_total, const float* halo,
                     const float* A, const float* pi, const float* Bfull, const float* codon5, const float* hints, unsigned flags,
                     const float* all_blocks, const unsigned* block_flags, int rank, int world, unsigned tag, float* out_post,
                     float* out_alpha, float* out_beta, float* loglik, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  Layout l;
  int rc = shard_args(a, l, inputs, num_models, batch, T, seq_off, seq_total, halo, A, pi, Bfull, codon5, hints, flags, workspace,
                      workspace_bytes);
  if (rc) return rc;
  HMM_REQUIRE(all_blocks && world > 0 && world <= MAX_PEERS && rank >= 0 && rank < world, "genepred shard_finalize_x: bad rank %d of %d", rank, world);
  HMM_REQUIRE(out_post || out_alpha || out_beta || loglik, "genepred shard_finalize_x: no output requested");
  const int NB = num_models * batch, PG = cdiv(a.P, G);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  a.ll_external = 1;
  if (PG > G) {
    if (block_flags) shard_wait_kernel<<<1, 256, 0, st>>>(block_flags, world * NB, tag);
    float* in_f = reinterpret_cast<float*>(ws + l.inf);
    float* in_r = reinterpret_cast<float*>(ws + l.inr);
    shard_boundary_kernel<<<NB, 32, 0, st>>>(all_blocks, rank, world, NB, batch, pi, in_f, in_r, a.llws, loglik);
    a.in_f = in_f; a.in_r = in_r;
    a.out_post = out_post; a.out_alpha = out_alpha; a.out_beta = out_beta; a.loglik = nullptr;
    if (out_post || out_alpha || out_beta) {
      const bool full = out_alpha || out_beta;
      const size_t sm2 = k2_smem_bytes(a.L, full);
      rc = full ? set_smem(finalize_kernel<true>, SLOT_K2F, sm2) : set_smem(finalize_kernel<false>, SLOT_K2P, sm2); if (rc) return rc;
      rc = launch_boundary(a, l, NB, 0, nullptr, true, workspace, st); if (rc) return rc;
      rc = launch_finalize(a, NB, full, sm2, st, false); if (rc) return rc;
    }
    HMM_CUDA_OK(cudaGetLastError());
    return 0;
  }
  rc = ensure_dynamic_smem(shard_scan_kernel, 25, sizeof(XSmem)); if (rc) return rc;
  {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(PG, NB); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = sizeof(XSmem); cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    HMM_CUDA_OK(cudaLaunchKernelEx(&cfg, shard_scan_kernel, a, reinterpret_cast<const float*>(ws + l.Gt), all_blocks, block_flags, rank,
                                   world, tag, loglik));
  }
  a.out_post = out_post; a.out_alpha = out_alpha; a.out_beta = out_beta; a.loglik = nullptr;
  if (out_post || out_alpha || out_beta) {
    const bool full = out_alpha || out_beta;
    const size_t sm2 = k2_smem_bytes(a.L, full);
    rc = full ? set_smem(finalize_kernel<true>, SLOT_K2F, sm2) : set_smem(finalize_kernel<false>, SLOT_K2P, sm2); if (rc) return rc;
    rc = launch_finalize(a, NB, full, sm2, st, true); if (rc) return rc;
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- sequence sharding in ONE launch per step: the single-launch kernel with reducer CTAs (gene15_fused.cu) ----------
// Chunk length for a shard of T positions: the tile CTAs and the reducer CTAs (one per 32 tiles + one) of every sequence must
// be resident at once; 0 = the shard does not qualify (the 4-launch route above takes it).
int shard_fused_chunk(int NB, int T, int world) {
  if (g_fused_mode == 1 || NB <= 0 || T <= 0 || world <= 0 || world > MAX_PEERS) return 0;
  static const bool env_off = getenv("HMM_B200_FUSED") && atoi(getenv("HMM_B200_FUSED")) == 0;
  if (env_off) return 0;
  const int sms = num_sms();
  double best = 1e300;
  int bestL = 0;
  for (int L = (g_L > 0 ? g_L : FUSED_MIN_L); L <= (g_L > 0 ? g_L : MAX_L); ++L) {
    const int cap = fused_capacity(L);
    if (cap <= 0) break;
    const int P = cdiv(cdiv(T, L), G), PG = cdiv(P, G);
    const long long ctas = (long long)NB * (P + PG + 1);   // tiles, group reducers, the block reducer
    if (PG > G || world + PG > FUSED_SCAN_SLOTS || ctas > cap) continue;   // (ranks' blocks + group matrices share the scan slots)
    const double cost = (44.0 + 0.85 * L) * (ctas <= sms ? 1.0 : 1.36);   // choose_fused_L's model
    if (cost < best) { best = cost; bestL = L; }
  }
  return bestL;
}

int shard_run_x(const float* inputs, int num_models, int batch, int T, int seq_off, int seq_total, const float* halo,
                const float* A, const float* pi, const float* Bfull, const float* codon5, const float* hints, unsigned flags,
                const unsigned long long* peer_blocks, const unsigned long long* peer_flags, int rank, int world, unsigned tag,
                float* out_post, float* loglik, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  int rc = fill_args(a, inputs, num_models, batch, T, A, pi, Bfull, codon5, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE(inputs && A && pi && out_post, "genepred shard_run_x: null inputs / A / pi / posterior output");
  HMM_REQUIRE(seq_off >= 0 && seq_total >= seq_off + T, "genepred shard_run_x: positions [%d, %d) outside a sequence of length %d",
              seq_off, seq_off + T, seq_total);
  HMM_REQUIRE(halo || (seq_off == 0 && seq_total == T), "genepred shard_run_x: a proper shard needs the 2+2 halo nucleotides");
  HMM_REQUIRE(peer_blocks && peer_flags && world > 0 && world <= MAX_PEERS && rank >= 0 && rank < world && tag != 0,
              "genepred shard_run_x: bad peer tables (world %d, rank %d, tag %u)", world, rank, tag);
  const int NB = num_models * batch;
  const int L = shard_fused_chunk(NB, T, world);
  HMM_REQUIRE(L > 0, "genepred shard_run_x: %d sequences x %d positions do not fit one launch on this device (use shard_reduce_x / "
              "shard_finalize_x; hmm_b200_genepred_shard_fused_chunk tells beforehand)", NB, T);
  a.seq_off = seq_off; a.seq_total = seq_total; a.halo = halo;
  set_L(a, L);
  a.top_scan = 0;
  Layout l;
  rc = bind_workspace(a, l, NB, workspace, workspace_bytes, "genepred shard_run_x");
  if (rc) return rc;
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  a.out_post = out_post; a.out_alpha = nullptr; a.out_beta = nullptr; a.loglik = loglik;
  return launch_fused_shard(a, NB, peer_blocks, peer_flags, rank, world, tag, reinterpret_cast<float*>(ws + l.Gt),
                            reinterpret_cast<unsigned*>(ws + l.gflg), reinterpret_cast<unsigned*>(ws + l.bvflg), st);
}

int emission(const float* inputs, int num_models, int batch, int T, const float* Bfull, const float* codon5,
             const float* hints, unsigned flags, float* out_E, cudaStream_t st) {
  Args a{};
  int rc = fill_args(a, inputs, num_models, batch, T, nullptr, nullptr, Bfull, codon5, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE(inputs && out_E, "genepred_emission: null inputs / output");
  // Tile length (positions per CTA = 32 L; the results do not depend on it): the one that needs the fewest waves of resident
  // CTAs -- at config 2, L = 32 gives 320 CTAs for 296 places (a second wave of 24 CTAs doubles the time), L = 36 gives 288.
  auto smem_for = [](int L) {
    return ((sizeof(EmisSmem) + 15) / 16) * 16 + (size_t)G * (L * QP + 4) * sizeof(float) + (((size_t)G * L + 4 + 15) / 16) * 16;
  };
  int bestL = 32;
  {
    static int per_sm[kMaxDevices][MAX_L + 1] = {};   // resident CTAs per SM by tile length (0: not asked yet, -1: does not fit)
    const int dev = current_device(), sms = num_sms();
    double best = 1e300;
    for (int L = 24; L <= MAX_L; L += 2) {
      int c;
      { std::lock_guard<std::mutex> lk(state_mutex()); c = per_sm[dev][L]; }
      if (c == 0) {
        c = -1;
        if (smem_for(L) <= (size_t)227 * 1024 && set_smem(emission_kernel, SLOT_EM, smem_for(L)) == 0) {
          int v = 0;
          if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&v, emission_kernel, 512, smem_for(L)) == cudaSuccess && v > 0) c = v;
          else (void)cudaGetLastError();
        }
        std::lock_guard<std::mutex> lk(state_mutex());
        per_sm[dev][L] = c;
      }
      if (c <= 0) continue;
      const long long tiles = (long long)num_models * batch * cdiv(cdiv(T, L), G);
      const long long waves = (tiles + (long long)c * sms - 1) / ((long long)c * sms);
      const double cost = (double)waves * (L + 6) * (c > 1 ? 1.0 + 0.5 * (c - 1) : 1.0);   // co-resident CTAs share the SM's issue slots
      if (cost < best) { best = cost; bestL = L; }
    }
  }
  set_L(a, bestL);
  const size_t sm = smem_for(a.L);
  rc = set_smem(emission_kernel, SLOT_EM, sm); if (rc) return rc;
  emission_kernel<<<dim3(a.P, num_models * batch), 512, sm, st>>>(a, out_E);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

size_t grad_workspace_size(int num_models, int batch, int T) {
  if (num_models <= 0 || batch <= 0 || T <= 0) return 0;
  return (size_t)num_models * batch * cdiv(T, GR_POS) * GR_PART * sizeof(float);
}

int grad(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
         const float* codon5, const float* hints, unsigned flags, const float* log_post, const float* log_ahat,
         const float* grad_loglik, float* grad_inputs, float* grad_A, float* grad_pi, float* grad_Bfull, void* workspace,
         size_t workspace_bytes, cudaStream_t st) {
  Args a{};
  int rc = fill_args(a, inputs, num_models, batch, T, A, pi, Bfull, codon5, hints, flags);
  if (rc) return rc;
  HMM_REQUIRE(inputs && A && pi, "genepred_grad: null inputs / A / pi");
  HMM_REQUIRE(log_post && log_ahat && grad_loglik, "genepred_grad: null log_post / log_alpha_hat / grad_loglik");
  HMM_REQUIRE(grad_inputs || grad_A || grad_pi || grad_Bfull, "genepred_grad: no gradient requested");
  HMM_REQUIRE(!grad_inputs || (reinterpret_cast<uintptr_t>(grad_inputs) & 15) == 0, "genepred_grad: grad_inputs must be 16-byte aligned");
  const int NB = num_models * batch;
  GradArgs g{};
  g.log_post = log_post; g.log_ahat = log_ahat; g.gll = grad_loglik; g.grad_inputs = grad_inputs;
  g.PT = cdiv(T, GR_POS);
  const size_t need = grad_workspace_size(num_models, batch, T);
  if (!workspace || workspace_bytes < need) {
    set_error("genepred_grad: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
    return HMM_B200_EWORKSPACE;
  }
  g.part = static_cast<float*>(workspace);
  set_L(a, 32);   // codes_stage / emission helpers only need a valid chunk length
  rc = set_smem(grad_kernel, SLOT_GRAD, sizeof(GradSmem)); if (rc) return rc;
  if (grad_A) HMM_CUDA_OK(cudaMemsetAsync(grad_A, 0, (size_t)num_models * Q * Q * sizeof(float), st));
  grad_kernel<<<dim3(g.PT, NB), GR_POS, sizeof(GradSmem), st>>>(a, g);
  grad_reduce_kernel<<<dim3(GR_PART / 32, num_models), 1024, 0, st>>>(g.part, batch * g.PT, grad_A, grad_pi, grad_Bfull);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

static size_t viterbi_psi_words(int T) { return (size_t)((T + 2) / 4 + 1) * 16; }   // words per sequence
// chunked (max,+) scan (gene15_viterbi.cu)
bool viterbi_scan_applies(int T);
size_t viterbi_scan_scratch_bytes(int NB);
int launch_viterbi_scan(const float* E, const float* A, const float* pi, int NB, int batch, int T, void* scratch,
                        unsigned char* path, float* score, cudaStream_t st);

size_t viterbi_workspace_size(int num_models, int batch, int T) {
  if (num_models <= 0 || batch <= 0 || T <= 0) return 0;
  const size_t NBT = (size_t)num_models * batch * T;
  size_t second = (size_t)num_models * batch * viterbi_psi_words(T) * 4;           // back-pointer table of viterbi_kernel
  const size_t scan = viterbi_scan_scratch_bytes(num_models * batch);                // chunk matrices of the scan
  if (scan > second) second = scan;
  return align_up(NBT * Q * sizeof(float), 256) + align_up(second, 256);
}

int viterbi(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
            const float* codon5, const float* hints, unsigned flags, unsigned char* path, float* score, void* workspace,
            size_t workspace_bytes, cudaStream_t st) {
  HMM_REQUIRE(A && pi && path, "genepred_viterbi: null A / pi / path");
  const size_t need = viterbi_workspace_size(num_models, batch, T);
  if (!workspace || workspace_bytes < need) {
    set_error("genepred_viterbi: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
    return HMM_B200_EWORKSPACE;
  }
  HMM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "genepred_viterbi: workspace must be 256-byte aligned");
  float* E = static_cast<float*>(workspace);
  unsigned* psi = reinterpret_cast<unsigned*>(static_cast<unsigned char*>(workspace) + align_up((size_t)num_models * batch * T * Q * sizeof(float), 256));
  int rc = emission(inputs, num_models, batch, T, Bfull, codon5, hints, flags | EMIT_LOG_FLOORED, E, st);
  if (rc) return rc;
  if (viterbi_scan_applies(T)) return launch_viterbi_scan(E, A, pi, num_models * batch, batch, T, psi, path, score, st);
  viterbi_kernel<<<num_models * batch, 32, 0, st>>>(E, A, pi, batch, T, psi, viterbi_psi_words(T), path, score);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int params(const float* tk, const float* sk, const float* ek, int num_models, float* A, float* pi, float* Bfull, cudaStream_t st) {
  HMM_REQUIRE(tk && sk && ek && A && pi && Bfull && num_models > 0, "genepred_params: null pointer");
  params_kernel<<<1, 64, 0, st>>>(tk, sk, ek, num_models, A, pi, Bfull);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int params_grad(const float* tk, const float* sk, const float* ek, int num_models, const float* gA, const float* gpi,
                const float* gB, float* gtk, float* gsk, float* gek, cudaStream_t st) {
  HMM_REQUIRE(tk && sk && ek && num_models > 0, "genepred_params_grad: null pointer");
  params_grad_kernel<<<1, 64, 0, st>>>(tk, sk, ek, num_models, gA, gpi, gB, gtk, gsk, gek);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

// ---- forward + posterior without any allocation: the workspace is a static device array ----------------------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every device
// that loads the library holds HMM_B200_NOALLOC_SCRATCH_BYTES of it until the process exits.  In host code the array's
// name is only a stand-in: its device address is looked up once per device (that device current) and handed to run().
}  // namespace gene15
}  // namespace hmm

static __device__ __align__(256) unsigned char g_noalloc_scratch[HMM_B200_NOALLOC_SCRATCH_BYTES];

namespace hmm {
namespace gene15 {

static int noalloc_scratch(void** out) {
  static void* ptrs[kMaxDevices] = {};
  int dev = -1;
  HMM_CUDA_OK(cudaGetDevice(&dev));
  HMM_REQUIRE(dev >= 0 && dev < kMaxDevices, "genepred_run_noalloc: device %d beyond the %d this library keeps state for", dev, kMaxDevices);
  std::lock_guard<std::mutex> lk(state_mutex());
  if (ptrs[dev] == nullptr) {
    void* p = nullptr;
    HMM_CUDA_OK(cudaGetSymbolAddress(&p, g_noalloc_scratch));   // the address on the CURRENT device
    HMM_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "genepred_run_noalloc: bad scratch address %p", p);
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

int run_noalloc(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
                const float* codon5, const float* hints, unsigned flags, float* out_post, float* out_alpha, float* out_beta,
                float* loglik, cudaStream_t st) {
  HMM_REQUIRE(num_models > 0 && batch > 0 && T > 0, "genepred_run_noalloc: num_models, batch and T must be positive (got %d,%d,%d)",
              num_models, batch, T);
  const size_t need = workspace_size(num_models, batch, T);
  if (need > (size_t)HMM_B200_NOALLOC_SCRATCH_BYTES) {
    set_error("genepred_run_noalloc: %d x %d sequences of %d positions need %zu bytes of scratch, the static array holds %zu "
              "(HMM_B200_NOALLOC_SCRATCH_BYTES); use hmm_b200_genepred_run with a caller-owned workspace",
              num_models, batch, T, need, (size_t)HMM_B200_NOALLOC_SCRATCH_BYTES);
    return HMM_B200_EWORKSPACE;
  }
  void* ws = nullptr;
  int rc = noalloc_scratch(&ws);
  if (rc) return rc;
  // The array keeps the last call's contents.  What must start at zero -- the per-sequence arrival counters of the
  // two-kernel route -- is cleared by run() on `st` in every call that uses them; the flags of the single-launch route
  // carry a per-call tag and need no clearing; everything else is written before it is read.
  return run(inputs, num_models, batch, T, A, pi, Bfull, codon5, hints, flags, out_post, out_alpha, out_beta, loglik, ws,
             (size_t)HMM_B200_NOALLOC_SCRATCH_BYTES, st);
}

}  // namespace gene15
}  // namespace hmm

extern "C" {

int hmm_b200_genepred_params(const float* transition_kernel, const float* start_kernel, const float* emission_kernel,
                             int num_models, float* A, float* pi, float* Bfull, void* stream) {
  return hmm::gene15::params(transition_kernel, start_kernel, emission_kernel, num_models, A, pi, Bfull,
                             static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_params_grad(const float* transition_kernel, const float* start_kernel, const float* emission_kernel,
                                  int num_models, const float* grad_A, const float* grad_pi, const float* grad_Bfull,
                                  float* grad_transition_kernel, float* grad_start_kernel, float* grad_emission_kernel,
                                  void* stream) {
  return hmm::gene15::params_grad(transition_kernel, start_kernel, emission_kernel, num_models, grad_A, grad_pi, grad_Bfull,
                                  grad_transition_kernel, grad_start_kernel, grad_emission_kernel,
                                  static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_run_compact(const unsigned short* class_probs_bf16, const unsigned char* nuc_codes, int num_models, int batch,
                                  int T, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                  const float* end_hints, unsigned flags, float* out_post, float* out_alpha, float* out_beta,
                                  float* loglik, void* workspace, size_t workspace_bytes, void* stream) {
  return hmm::gene15::run(nullptr, num_models, batch, T, A, pi, Bfull, codon5, end_hints, flags, out_post, out_alpha, out_beta,
                          loglik, workspace, workspace_bytes, static_cast<cudaStream_t>(stream), class_probs_bf16, nuc_codes);
}

int hmm_b200_genepred_run_host_compact(const unsigned short* host_class_probs_bf16, const unsigned char* host_nuc_codes,
                                       int num_models, int batch, int T, const float* A, const float* pi, const float* Bfull,
                                       const float* codon5, const float* host_end_hints, unsigned flags, float* host_out_post,
                                       float* host_out_alpha, float* host_out_beta, float* host_loglik, void* device_scratch,
                                       size_t device_scratch_bytes, void* stream) {
  return hmm::gene15::run_host(nullptr, num_models, batch, T, A, pi, Bfull, codon5, host_end_hints, flags, host_out_post,
                               host_out_alpha, host_out_beta, host_loglik, device_scratch, device_scratch_bytes,
                               static_cast<cudaStream_t>(stream), host_class_probs_bf16, host_nuc_codes);
}

size_t hmm_b200_genepred_viterbi_workspace_size(int num_models, int batch, int T) {
  return hmm::gene15::viterbi_workspace_size(num_models, batch, T);
}

int hmm_b200_genepred_viterbi(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                              const float* Bfull, const float* codon5, const float* end_hints, unsigned flags,
                              unsigned char* path, float* score, void* workspace, size_t workspace_bytes, void* stream) {
  return hmm::gene15::viterbi(inputs, num_models, batch, T, A, pi, Bfull, codon5, end_hints, flags, path, score, workspace,
                              workspace_bytes, static_cast<cudaStream_t>(stream));
}

size_t hmm_b200_genepred_grad_workspace_size(int num_models, int batch, int T) {
  return hmm::gene15::grad_workspace_size(num_models, batch, T);
}

int hmm_b200_genepred_grad(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                           const float* Bfull, const float* codon5, const float* end_hints, unsigned flags,
                           const float* log_post, const float* log_alpha_hat, const float* grad_loglik, float* grad_inputs,
                           float* grad_A, float* grad_pi, float* grad_Bfull, void* workspace, size_t workspace_bytes,
                           void* stream) {
  return hmm::gene15::grad(inputs, num_models, batch, T, A, pi, Bfull, codon5, end_hints, flags, log_post, log_alpha_hat,
                           grad_loglik, grad_inputs, grad_A, grad_pi, grad_Bfull, workspace, workspace_bytes,
                           static_cast<cudaStream_t>(stream));
}

size_t hmm_b200_genepred_workspace_size(int num_models, int batch, int T) {
  return hmm::gene15::workspace_size(num_models, batch, T);
}

size_t hmm_b200_genepred_host_scratch_size(int num_models, int batch, int T) {
  return hmm::gene15::host_scratch_size(num_models, batch, T);
}

int hmm_b200_genepred_run_host(const float* host_inputs, int num_models, int batch, int T, const float* A, const float* pi,
                               const float* Bfull, const float* codon5, const float* host_end_hints, unsigned flags,
                               float* host_out_post, float* host_out_alpha, float* host_out_beta, float* host_loglik,
                               void* device_scratch, size_t device_scratch_bytes, void* stream) {
  return hmm::gene15::run_host(host_inputs, num_models, batch, T, A, pi, Bfull, codon5, host_end_hints, flags, host_out_post,
                               host_out_alpha, host_out_beta, host_loglik, device_scratch, device_scratch_bytes,
                               static_cast<cudaStream_t>(stream));
}

int hmm_b200_set_fused_mode(int mode) { return hmm::gene15::set_fused_mode(mode); }
int hmm_b200_debug_phases(void* device_buffer) { return hmm::gene15::set_debug_buffer(device_buffer); }
int hmm_b200_profile_enable(int on) { return hmm::gene15::profile_enable(on); }
int hmm_b200_profile_read(float* ms_host, int n) { return hmm::gene15::profile_read(ms_host, n); }

int hmm_b200_set_tuning(int chunk_len, int finalize_threads, int reserved) {
  return hmm::gene15::set_tuning(chunk_len, finalize_threads, reserved);
}

int hmm_b200_genepred_prepare_codon(const float* codon64, float* codon5, void* stream) {
  return hmm::gene15::prepare_codon(codon64, codon5, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_run(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                          const float* Bfull, const float* codon5, const float* end_hints, unsigned flags,
                          float* out_post, float* out_alpha, float* out_beta, float* loglik, void* workspace,
                          size_t workspace_bytes, void* stream) {
  return hmm::gene15::run(inputs, num_models, batch, T, A, pi, Bfull, codon5, end_hints, flags, out_post, out_alpha,
                          out_beta, loglik, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_emission(const float* inputs, int num_models, int batch, int T, const float* Bfull,
                               const float* codon5, const float* end_hints, unsigned flags, float* out_E, void* stream) {
  return hmm::gene15::emission(inputs, num_models, batch, T, Bfull, codon5, end_hints, flags, out_E,
                               static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_shard_reduce(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                   const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                   const float* end_hints, unsigned flags, float* block_out, void* workspace,
                                   size_t workspace_bytes, void* stream) {
  return hmm::gene15::shard_reduce(inputs, num_models, batch, T_local, seq_off, seq_total, halo, A, pi, Bfull, codon5, end_hints,
                                   flags, block_out, workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_shard_finalize(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                     const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                     const float* end_hints, unsigned flags, const float* all_blocks, int rank, int world,
                                     float* out_post, float* out_alpha, float* out_beta, float* loglik, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  return hmm::gene15::shard_finalize(inputs, num_models, batch, T_local, seq_off, seq_total, halo, A, pi, Bfull, codon5, end_hints,
                                     flags, all_blocks, rank, world, out_post, out_alpha, out_beta, loglik, workspace,
                                     workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_shard_reduce_x(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                     const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                     const float* end_hints, unsigned flags, float* block_out,
                                     const unsigned long long* peer_block_tables, const unsigned long long* peer_flag_tables,
                                     int rank, int world, unsigned step_tag, void* workspace, size_t workspace_bytes, void* stream) {
  return hmm::gene15::shard_reduce_x(inputs, num_models, batch, T_local, seq_off, seq_total, halo, A, pi, Bfull, codon5, end_hints,
                                     flags, block_out, peer_block_tables, peer_flag_tables, rank, world, step_tag, workspace,
                                     workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_shard_finalize_x(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                       const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                       const float* end_hints, unsigned flags, const float* all_blocks, const unsigned* block_flags,
                                       int rank, int world, unsigned step_tag, float* out_post, float* out_alpha, float* out_beta,
                                       float* loglik, void* workspace, size_t workspace_bytes, void* stream) {
  return hmm::gene15::shard_finalize_x(inputs, num_models, batch, T_local, seq_off, seq_total, halo, A, pi, Bfull, codon5, end_hints,
                                       flags, all_blocks, block_flags, rank, world, step_tag, out_post, out_alpha, out_beta, loglik,
                                       workspace, workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_shard_fused_chunk(int num_models, int batch, int T_local, int world) {
  return hmm::gene15::shard_fused_chunk(num_models * batch, T_local, world);
}

int hmm_b200_genepred_shard_run_x(const float* inputs, int num_models, int batch, int T_local, int seq_off, int seq_total,
                                  const float* halo, const float* A, const float* pi, const float* Bfull, const float* codon5,
                                  const float* end_hints, unsigned flags, const unsigned long long* peer_block_tables,
                                  const unsigned long long* peer_flag_tables, int rank, int world, unsigned step_tag,
                                  float* out_post, float* loglik, void* workspace, size_t workspace_bytes, void* stream) {
  return hmm::gene15::shard_run_x(inputs, num_models, batch, T_local, seq_off, seq_total, halo, A, pi, Bfull, codon5, end_hints,
                                  flags, peer_block_tables, peer_flag_tables, rank, world, step_tag, out_post, loglik, workspace,
                                  workspace_bytes, static_cast<cudaStream_t>(stream));
}

int hmm_b200_genepred_run_noalloc(const float* inputs, int num_models, int batch, int T, const float* A, const float* pi,
                                  const float* Bfull, const float* codon5, const float* end_hints, unsigned flags,
                                  float* out_post, float* out_alpha, float* out_beta, float* loglik, void* stream) {
  return hmm::gene15::run_noalloc(inputs, num_models, batch, T, A, pi, Bfull, codon5, end_hints, flags, out_post, out_alpha,
                                  out_beta, loglik, static_cast<cudaStream_t>(stream));
}

}  // extern "C"

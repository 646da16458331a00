// Generic dense-transition HMM path for sm_100a: scaled forward / backward / posterior for any number of states q
// (profile HMMs q = 2L+3, multi-copy gene models q = 1+14k), from a materialised emission tensor.
//
// Reference semantics (MsaHmmCell.py:73-106 driven by BaseRNN.py:216-227, drivers MsaHMMLayer.py:227-521):
//   forward   R_0 = max(pi, eps);  R_t = max(a-hat_{t-1} . A, eps);  a_t = max(E_t, eps) * R_t;  S_t = sum_q a_t;
//             a-hat_t = a_t / S_t;  log alpha_t = log a-hat_t + sum_{tau<=t} log S_tau
//   backward  R_{T-1} = 1;  R_t = max(b-hat_{t+1} . A^T, eps);  b_t = max(E_t, eps) * R_t;  b-hat_t = b_t / S;
//             log beta_t = log R_t + sum_{tau>t} log S_tau
//   posterior log alpha + log beta - loglik
//
// One launch per time step and direction, stream-ordered: a shared-memory tiled fp32 GEMM
// state(B x q) . A(q x q) with the floor / emission / row-sum epilogue fused.  Normalisation is deferred to the consumer:
// the step writes a_t (un-normalised) and per-column-tile partial row sums; the next step divides by their sum
// (deterministic order), so no atomics and bit-reproducible results.  A tcgen05 version of the step GEMM for the
// learnMSA sizes (q >= 192) lives in dense_tc.cu and shares this file's scale bookkeeping and finish kernels.
#include "common.cuh"

namespace hmm {
namespace dense_tc {   // dense_tc.cu: tcgen05 step for profile-HMM sizes
bool supported(int k, int B, int q);
int slots(int q);
size_t extra_workspace(int k, int B, int q);
int run_steps(const float* E, const unsigned char* sym, const float* Bt, int nsym, const float* A, const float* pi, int k, int B, int T,
              int q, bool need_f, bool need_b, float* store_f, float* store_b, float* part[2][2], float* hist[2], void* extra,
              cudaStream_t st);
}  // namespace dense_tc
namespace dense {

// 0 = by size (tensor cores from kTcMinStates states), 1 = CUDA-core step always, 2 = tensor-core step always
static int g_path = 0;
constexpr int kTcMinStates = 192;
void set_path(int p) { g_path = (p == 1 || p == 2) ? p : 0; }
static bool use_tc(int k, int B, int q) {
  if (g_path == 1) return false;
  return dense_tc::supported(k, B, q) && (g_path == 2 || q >= kTcMinStates);
}

constexpr int BM = 64, BN = 64, BK = 16, TM = 4, TN = 4;   // 256 threads, 4x4 outputs each
constexpr int MAXP = 32;                                    // at most 32 column tiles (q <= 2048)

struct StepArgs {
  const float* state;    // (k, B, q) un-normalised previous vector (or the initial state when init)
  const float* part_in;  // (k, B, ntile) partial sums of `state` (ignored when init)
  const float* A;        // (k, q, q): forward A, backward A^T (row-major, so that U = state . A)
  const float* E;        // emission rows of this step: E + ((m*B + b)*T + t)*q
  float* state_out;      // (k, B, q)
  float* part_out;       // (k, B, ntile)
  float* store;          // optional (k, B, T, q) buffer; row t receives a_t (forward) / R_t (backward), un-normalised
  float* scale_hist;     // (k, B, T): S of the *input* state used at this step (1 when init) -> slot t
  int k, B, T, q, t, ntile;
  int init, reverse;
};

__global__ void __launch_bounds__(256) step_kernel(const StepArgs a) {
  __shared__ float As[BK][BM + 4];   // state tile, transposed: As[kk][row]
  __shared__ float Bs[BK][BN + 4];
  __shared__ float Sin[BM];          // 1 / S of the input rows
  __shared__ float red[BM][BN / TN + 1];
  const int m = blockIdx.z, row0 = blockIdx.y * BM, col0 = blockIdx.x * BN;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;   // thread tile: rows ty*4.., cols tx*4..
  const int q = a.q, B = a.B;
  const float* st = a.state + (size_t)m * B * q;
  const float* Am = a.A + (size_t)m * q * q;

  if (tid < BM) {
    float s = 1.f;
    const int r = row0 + tid;
    if (!a.init && r < B) {
      s = 0.f;
      const float* p = a.part_in + ((size_t)m * B + r) * a.ntile;
      for (int i = 0; i < a.ntile; ++i) s += p[i];
    }
    Sin[tid] = s;
    if (blockIdx.x == 0 && r < B && a.scale_hist) a.scale_hist[((size_t)m * B + r) * a.T + a.t] = s;
  }
  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  if (!a.init) {
    for (int k0 = 0; k0 < q; k0 += BK) {
      // state tile: BM x BK (256 threads x 4 elements)
#pragma unroll
      for (int l = 0; l < (BM * BK) / 256; ++l) {
        const int idx = tid + l * 256, r = idx / BK, kk = idx - r * BK;
        const int gr = row0 + r, gk = k0 + kk;
        As[kk][r] = (gr < B && gk < q) ? st[(size_t)gr * q + gk] : 0.f;
      }
#pragma unroll
      for (int l = 0; l < (BK * BN) / 256; ++l) {
        const int idx = tid + l * 256, kk = idx / BN, c = idx - kk * BN;
        const int gk = k0 + kk, gc = col0 + c;
        Bs[kk][c] = (gk < q && gc < q) ? Am[(size_t)gk * q + gc] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        float av[TM], bv[TN];
#pragma unroll
        for (int i = 0; i < TM; ++i) av[i] = As[kk][ty * TM + i];
#pragma unroll
        for (int j = 0; j < TN; ++j) bv[j] = Bs[kk][tx * TN + j];
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
          for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
      }
      __syncthreads();
    }
  } else {
    __syncthreads();
  }

  // epilogue: R = max(U / S_in, eps)  (init: U = the given state);  x = max(E, eps) * R
  float rsum[TM];
#pragma unroll
  for (int i = 0; i < TM; ++i) {
    rsum[i] = 0.f;
    const int r = row0 + ty * TM + i;
    if (r >= B) continue;
    const float inv = 1.0f / Sin[ty * TM + i];
    const float* Er = a.E + (((size_t)m * B + r) * a.T + a.t) * q;
    float* so = a.state_out + ((size_t)m * B + r) * q;
    float* sr = a.store ? a.store + (((size_t)m * B + r) * a.T + a.t) * q : nullptr;
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int c = col0 + tx * TN + j;
      if (c >= q) continue;
      const float U = a.init ? st[(size_t)r * q + c] : acc[i][j] * inv;
      const float R = fmaxf(U, kEps);
      const float x = fmaxf(Er[c], kEps) * R;
      so[c] = x;
      if (sr) sr[c] = a.reverse ? R : x;
      rsum[i] += x;
    }
  }
#pragma unroll
  for (int i = 0; i < TM; ++i) red[ty * TM + i][tx] = rsum[i];
  __syncthreads();
  if (tid < BM) {
    const int r = row0 + tid;
    if (r < B) {
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < BN / TN; ++j) s += red[tid][j];
      a.part_out[((size_t)m * B + r) * a.ntile + blockIdx.x] = s;
    }
  }
}

// Persistent recurrence for small state counts (q < 192: config 1's q = 67, multi-copy gene models): ONE launch runs all T
// steps of one sequence and direction per CTA — the transition matrix stays resident in shared memory, the state vector
// never leaves the SM, the next step's emission row is loaded while the current step computes, the row sum is a warp-shuffle
// reduction.  Same arithmetic and the same stored quantities as step_kernel (un-normalised a_t / R_t, scale history, final
// partial sums), so cum_kernel / finish_kernel are shared.  thread = state.
struct SeqArgs {
  const float* E; const float* A; const float* At; const float* pi;
  // one-hot form (E == NULL): E_t = Bt[m][sym_t], sym (k, B, T) bytes, Bt (k, nsym, q); the sequence's symbols are staged
  // in shared memory behind the transition matrix and the ring is filled from rows of the (L2-resident) table
  const unsigned char* sym; const float* Bt; int nsym;
  float* store[2];   // forward: a_t, backward: R_t  (k, B, T, q) or NULL when the direction is not wanted
  float* hist[2];    // (k, B, T)
  float* part_f;     // (k, B, ntile): partial sums of the last forward state (slot 0 gets the sum, the rest zeros)
  int k, B, T, q, ntile, dir0;   // dir0: direction of blockIdx.y == 0
};

// QR > 0: the thread keeps its column of A in QR registers (q <= QR) and reads the state vector as float4 broadcasts;
// QR == 0: the column is walked in shared memory (any q < 192)
constexpr int SEQ_PF = 8;          // emission rows per prefetch block
constexpr int SEQ_XS = 200;        // floats per state buffer: 192 state slots + 8 per-warp partial sums
constexpr int SEQ_HDR = 2 * SEQ_XS + 2 * SEQ_PF * 192;   // two state buffers + the two-block emission ring
// shared-memory accesses of the step loop through 32-bit shared addresses held in registers (a generic pointer makes the
// compiler rebuild the shared window base from SR_CgaCtaId inside the loop, on the step's dependency chain)
__device__ __forceinline__ float4 lds4(unsigned addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ float2 lds2(unsigned addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ float lds1(unsigned addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts1(unsigned addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}

template <int QR>
__global__ void __launch_bounds__(192) seq_kernel(const SeqArgs a) {
  extern __shared__ float sm[];
  // thread index and direction pinned in registers: left to itself the compiler re-reads SR_TID / SR_CTAID (20+ cycles
  // each, on the step's dependency chain) several times per step instead of keeping them
  int tid_ = threadIdx.x, d_ = a.dir0 + blockIdx.y;
  asm volatile("" : "+r"(tid_), "+r"(d_));
  const int q = a.q, T = a.T, tid = tid_, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
  // sm[0..400): two state buffers (state of step s is written to buffer (s+1)&1 — one barrier per step); each holds the
  // un-normalised state padded to 192 floats (16-byte aligned float4 reads) and the per-warp partial sums at +192
  float* ring = sm + 2 * SEQ_XS;  // [2][SEQ_PF][192] emission ring, a thread touches only its own column
  float* As = sm + SEQ_HDR;       // q x q, row-major: U_j = sum_i x_i A[i][j], thread j walks a column
  const int seq = blockIdx.x, m = seq / a.B, d = d_;
  const float* Am = (d == 0 ? a.A : a.At) + (size_t)m * q * q;
  for (int i = tid; i < q * q; i += nt) As[i] = __ldg(Am + i);
  for (int i = tid; i < 2 * SEQ_XS; i += nt) sm[i] = 0.f;
  const bool live = tid < q;
  const int dt = d == 0 ? 1 : -1;
  const int t_first = d == 0 ? 0 : T - 1;
  const ptrdiff_t dq = (ptrdiff_t)dt * q;
  const float* Eseq = a.E ? a.E + (size_t)seq * T * q + (size_t)t_first * q + tid : nullptr;
  unsigned char* syms = reinterpret_cast<unsigned char*>(As + (size_t)q * q);
  if (!a.E) {
    const unsigned char* src = a.sym + (size_t)seq * T;
    for (int i = tid; i < T; i += nt) syms[i] = min((int)__ldg(src + i), a.nsym - 1);
    __syncthreads();
  }
  const char* btm = reinterpret_cast<const char*>(a.Bt + (size_t)m * a.nsym * q + tid);
  const unsigned syms_s = (unsigned)__cvta_generic_to_shared(syms);
  int t_issue = t_first;
  float* store = (a.store[d] && live) ? a.store[d] + (size_t)seq * T * q + (size_t)t_first * q + tid : nullptr;
  float* hist = a.hist[d] + (size_t)seq * T + t_first;
  // emission rows arrive SEQ_PF steps ahead through cp.async into the ring (a step is shorter than a DRAM round trip);
  // the step loop itself stays rolled so that its body is fetched from the instruction cache
  // (running pointers made opaque to the compiler, which otherwise rebuilds them from the kernel parameters at every use)
  unsigned ring_s = (unsigned)__cvta_generic_to_shared(ring + tid), sm_s = (unsigned)__cvta_generic_to_shared(sm);
  long long dqb = (long long)dq * 4;
  const char* esrc = reinterpret_cast<const char*>(Eseq);
  asm volatile("" : "+r"(ring_s), "+r"(sm_s), "+l"(dqb), "+l"(esrc), "+l"(store), "+l"(hist), "+l"(btm));
  int e_left = live ? T : 0;
  const bool onehot = a.E == nullptr;
  const long long qb = (long long)q * 4;
  auto issue = [&](int blk) {
    const unsigned dst = ring_s + (unsigned)((blk & 1) * SEQ_PF * 192 * 4);
#pragma unroll
    for (int u = 0; u < SEQ_PF; ++u) {
      if (u < e_left) {
        const char* src = esrc;
        if (onehot) {
          unsigned sy;
          asm volatile("ld.shared.u8 %0, [%1];" : "=r"(sy) : "r"(syms_s + (unsigned)t_issue));
          src = btm + (long long)sy * qb;
        }
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst + u * 192 * 4), "l"(src) : "memory");
      }
      esrc += dqb;
      t_issue += dt;
    }
    e_left -= SEQ_PF;
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  issue(0);
  __syncthreads();
  float acol[QR > 0 ? QR : 1];
  if (QR > 0) {
#pragma unroll
    for (int i = 0; i < QR; ++i) acol[i] = (live && i < q) ? As[(size_t)i * q + tid] : 0.f;
  }
  float S_in = 1.f;
  const int nblk = (T + SEQ_PF - 1) / SEQ_PF;
#pragma unroll 1
  for (int blk = 0; blk < nblk; ++blk) {
    issue(blk + 1);
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    const int s_end = min(SEQ_PF, T - blk * SEQ_PF);
#pragma unroll 1
    for (int u = 0; u < s_end; ++u) {
      const int s = blk * SEQ_PF + u;
      const float e = live ? lds1(ring_s + (unsigned)(((blk & 1) * SEQ_PF + u) * 192 * 4)) : 0.f;
      const float* xr = sm + (s & 1) * SEQ_XS;          // state of step s-1
      const unsigned xr_s = sm_s + (unsigned)((s & 1) * SEQ_XS * 4), xw_s = sm_s + (unsigned)(((s & 1) ^ 1) * SEQ_XS * 4);
      float U;
      if (s == 0) {
        U = live ? (d == 0 ? __ldg(a.pi + (size_t)m * q + tid) : 1.f) : 0.f;
      } else {
        float u0 = 0.f, u1 = 0.f, u2 = 0.f, u3 = 0.f;
        if (QR > 0) {
#pragma unroll
          for (int i = 0; i < QR / 4; ++i) {
            const float4 v = lds4(xr_s + 16 * i);
            u0 = fmaf(v.x, acol[4 * i], u0); u1 = fmaf(v.y, acol[4 * i + 1], u1);
            u2 = fmaf(v.z, acol[4 * i + 2], u2); u3 = fmaf(v.w, acol[4 * i + 3], u3);
          }
        } else if (live) {
          const float* col = As + tid;
          int i = 0;
          for (; i + 3 < q; i += 4) {
            u0 = fmaf(xr[i], col[(size_t)i * q], u0);
            u1 = fmaf(xr[i + 1], col[(size_t)(i + 1) * q], u1);
            u2 = fmaf(xr[i + 2], col[(size_t)(i + 2) * q], u2);
            u3 = fmaf(xr[i + 3], col[(size_t)(i + 3) * q], u3);
          }
          for (; i < q; ++i) u0 = fmaf(xr[i], col[(size_t)i * q], u0);
        }
        // S_in: sum of the previous state over the (at most six) warps; unused slots hold zeros
        const float4 r4 = lds4(xr_s + 192 * 4);
        const float2 r2 = lds2(xr_s + 196 * 4);
        S_in = ((r4.x + r4.y) + (r4.z + r4.w)) + (r2.x + r2.y);
        U = ((u0 + u1) + (u2 + u3)) * (1.0f / S_in);
      }
      const float R = fmaxf(U, kEps);
      const float x = live ? fmaxf(e, kEps) * R : 0.f;
      if (store) { *store = d ? R : x; store = reinterpret_cast<float*>(reinterpret_cast<char*>(store) + dqb); }
      if (tid == 0) *hist = S_in;
      hist += dt;
      float v = x;
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (live) sts1(xw_s + 4 * tid, x);
      if (lane == 0) sts1(xw_s + 4 * (192 + warp), v);
      __syncthreads();
    }
  }
  if (d == 0 && tid < a.ntile) {
    const float* xr = sm + (T & 1) * SEQ_XS + 192;
    a.part_f[(size_t)seq * a.ntile + tid] = tid == 0 ? ((xr[0] + xr[1]) + (xr[2] + xr[3])) + (xr[4] + xr[5]) : 0.f;
  }
}

template <int QR>
static int seq_launch(dim3 grid, int nthreads, size_t smem, cudaStream_t st, const SeqArgs& a) {
  { int rcs = ensure_dynamic_smem(seq_kernel<QR>, 10 + QR / 16, smem); if (rcs) return rcs; }   // slots 11..22: one per build
  seq_kernel<QR><<<grid, nthreads, smem, st>>>(a);
  return 0;
}

// per sequence: cumulative log scales (double) and the log-likelihood.
//   forward:  hist[t] = S_{t-1} (1 at t = 0), final = partial sums of the last state  ->  cum_f[t] = sum_{tau<=t} log S_tau
//   backward: hist[t] = S_{t+1} (1 at t = T-1)                                        ->  cum_b[t] = sum_{tau>t} log S_tau
__global__ void __launch_bounds__(32) cum_kernel(const float* __restrict__ hist_f, const float* __restrict__ part_f,
                                                 const float* __restrict__ hist_b, double* __restrict__ cum_f,
                                                 double* __restrict__ cum_b, float* __restrict__ loglik,
                                                 double* __restrict__ loglik_d, int N, int T, int ntile) {
  // one warp per sequence: lane l owns the contiguous steps [l * seg, (l + 1) * seg); segment sums in double, a warp
  // prefix / suffix scan over the 32 sums, then the running sums inside each segment (same order of additions for every
  // launch: deterministic)
  const int n = blockIdx.x, lane = threadIdx.x;
  if (n >= N) return;
  const int seg = (T + 31) / 32, lo = min(T, lane * seg), hi = min(T, lo + seg);
  if (hist_f) {
    const float* h = hist_f + (size_t)n * T;
    float last = 0.f;
    if (hi == T && lo < T)
      for (int i = 0; i < ntile; ++i) last += part_f[(size_t)n * ntile + i];
    auto term = [&](int t) { return log((double)(t + 1 < T ? h[t + 1] : last)); };
    double mine = 0.0;
    for (int t = lo; t < hi; ++t) mine += term(t);
    double pre = mine;   // inclusive prefix over lanes
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double v = __shfl_up_sync(0xffffffffu, pre, o);
      if (lane >= o) pre += v;
    }
    const double total = __shfl_sync(0xffffffffu, pre, 31);
    if (cum_f) {
      double c = pre - mine;
      for (int t = lo; t < hi; ++t) { c += term(t); cum_f[(size_t)n * T + t] = c; }
    }
    if (loglik && lane == 0) loglik[n] = (float)total;
    if (loglik_d && lane == 0) loglik_d[n] = total;
  }
  if (hist_b && cum_b) {
    const float* h = hist_b + (size_t)n * T;
    auto term = [&](int t) { return t + 1 < T ? log((double)h[t]) : 0.0; };   // h[t] = S of b_{t+1}
    double mine = 0.0;
    for (int t = lo; t < hi; ++t) mine += term(t);
    double suf = mine;   // inclusive suffix over lanes
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double v = __shfl_down_sync(0xffffffffu, suf, o);
      if (lane + o < 32) suf += v;
    }
    double c = suf - mine;
    for (int t = hi - 1; t >= lo; --t) { c += term(t); cum_b[(size_t)n * T + t] = c; }
  }
}

// in place: stored a_t -> log alpha_t,  stored R_t -> log beta_t,  and the posterior
// keep_a / keep_b = 0: the buffer was only lent as scratch (HMM_B200_*_SCRATCH), nothing is written back to it.
// The cumulative log scales and the log-likelihood are doubles: the posterior's shift cum_f + cum_b - loglik is a small
// difference of sums that reach 1e3 .. 1e5, formed here in double and only then rounded (in fp32 the rows of exp(post)
// summed to 1 +- 1e-3 at |loglik| = 5000).
__global__ void finish_kernel(float* __restrict__ alpha, float* __restrict__ beta, float* __restrict__ post,
                              const float* __restrict__ hist_f, const float* __restrict__ part_f, const double* __restrict__ cum_f,
                              const double* __restrict__ cum_b, const double* __restrict__ loglik_d, int N, int T, int q, int ntile,
                              int no_loglik, int keep_a, int keep_b, int norm_a) {
  const size_t row = blockIdx.x;           // (n, t)
  const int n = (int)(row / T), t = (int)(row - (size_t)n * T);
  float S = 1.f;
  double cf = 0.0, cb = 0.0;
  if (alpha) {
    if (t + 1 < T) S = hist_f[(size_t)n * T + t + 1];
    else { S = 0.f; for (int i = 0; i < ntile; ++i) S += part_f[(size_t)n * ntile + i]; }
    cf = cum_f[(size_t)n * T + t];
  }
  if (beta) cb = cum_b[(size_t)n * T + t];
  const float shift = post ? (float)((cf + cb) - (no_loglik ? 0.0 : loglik_d[n])) : 0.f;
  const float cff = (float)cf, cbf = (float)cb;
  const float invS = 1.0f / S;
  for (int j = threadIdx.x; j < q; j += blockDim.x) {
    const size_t i = row * q + j;
    float la = 0.f, lb = 0.f;
    if (alpha) { la = logf(alpha[i] * invS); }
    if (beta) { lb = logf(beta[i]); }
    if (post) post[i] = (la + lb) + shift;
    if (alpha && keep_a) alpha[i] = norm_a ? la : la + cff;   // norm_a: log of the normalised forward vector (HMM_B200_ALPHA_NORMALIZED)
    if (beta && keep_b) beta[i] = lb + cbf;
  }
}

// One HmmCell.forward call (MsaHmmCell.py:73-106) for API parity: row = (model, sequence, conditional row).
//   E_t (k*n, q); scaled (k*n, qc*q); loglik (k*n, qc);  A (k, q, q) already oriented (A or A^T)
//   out (k*n, qc*q + qc): [log a-hat | loglik] (forward) or [log R | old loglik] (reverse)
__global__ void cell_step_kernel(const float* __restrict__ E, const float* __restrict__ scaled, const float* __restrict__ loglik,
                                 const float* __restrict__ A, float* __restrict__ out, float* __restrict__ new_scaled,
                                 float* __restrict__ new_loglik, int k, int n, int qc, int q, int reverse, int init) {
  extern __shared__ float sh[];   // q state values + 32 reduction slots
  float* sv = sh;
  float* red = sh + q;
  const size_t r = blockIdx.x;                 // over k*n*qc
  const size_t seq = r / qc;                   // (model, sequence)
  const int c = (int)(r - seq * qc);
  const int m = (int)(seq / n);
  const float* s = scaled + seq * (size_t)qc * q + (size_t)c * q;
  for (int i = threadIdx.x; i < q; i += blockDim.x) sv[i] = s[i];
  __syncthreads();
  const float* Am = A + (size_t)m * q * q;
  const float* Er = E + seq * (size_t)q;
  float local = 0.f;
  for (int j = threadIdx.x; j < q; j += blockDim.x) {
    float U;
    if (init) U = sv[j];
    else {
      U = 0.f;
      for (int i = 0; i < q; ++i) U = fmaf(sv[i], Am[(size_t)i * q + j], U);
    }
    const float R = fmaxf(U, kEps);
    const float x = fmaxf(Er[j], kEps) * R;
    new_scaled[seq * (size_t)qc * q + (size_t)c * q + j] = x;          // normalised below
    if (reverse) out[seq * ((size_t)qc * q + qc) + (size_t)c * q + j] = logf(R);
    local += x;
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = local;
  __syncthreads();
  float S = 0.f;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) S += red[w];
  const float old_ll = loglik[seq * qc + c];
  const float ll = old_ll + logf(S);
  const float inv = 1.0f / S;
  for (int j = threadIdx.x; j < q; j += blockDim.x) {
    const size_t si = seq * (size_t)qc * q + (size_t)c * q + j;
    const float ah = new_scaled[si] * inv;
    new_scaled[si] = ah;
    if (!reverse) out[seq * ((size_t)qc * q + qc) + (size_t)c * q + j] = logf(ah);
  }
  if (threadIdx.x == 0) {
    new_loglik[seq * qc + c] = ll;
    out[seq * ((size_t)qc * q + qc) + (size_t)qc * q + c] = reverse ? old_ll : ll;
  }
}

// fill (k, B, q) with pi broadcast (forward) or ones (backward)
__global__ void init_state_kernel(float* __restrict__ st, const float* __restrict__ pi, int k, int B, int q) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)k * B * q) return;
  const int j = (int)(i % q);
  const int m = (int)(i / ((size_t)B * q));
  st[i] = pi ? pi[(size_t)m * q + j] : 1.f;
}

// fallback of the one-hot form for the engines that read a materialised emission tensor (per-step CUDA-core engine,
// sequences too long for seq_kernel's symbol staging): E[n][t][j] = Bt[m][sym[n][t]][j], coalesced over j
__global__ void __launch_bounds__(256) onehot_expand_kernel(const unsigned char* __restrict__ sym, const float* __restrict__ Bt,
                                                            float* __restrict__ E, size_t rows, int B, int T, int q, int nsym) {
  for (size_t r = blockIdx.x; r < rows; r += gridDim.x) {
    const int m = (int)(r / ((size_t)B * T));
    const int sy = min((int)__ldg(sym + r), nsym - 1);
    const float* src = Bt + ((size_t)m * nsym + sy) * q;
    for (int j = threadIdx.x; j < q; j += blockDim.x) E[r * q + j] = __ldg(src + j);
  }
}

struct Layout {
  size_t st[2][2], part[2][2], hist[2], cum[2], ll, lld, extra, total;
};
static Layout layout(int k, int B, int T, int q) {
  Layout l;
  const int ntile = cdiv(q, BN) > dense_tc::slots(q) ? cdiv(q, BN) : dense_tc::slots(q);
  const size_t N = (size_t)k * B;
  size_t off = 0;
  for (int d = 0; d < 2; ++d)
    for (int p = 0; p < 2; ++p) { l.st[d][p] = off; off = align_up(off + N * q * sizeof(float), 256); }
  for (int d = 0; d < 2; ++d)
    for (int p = 0; p < 2; ++p) { l.part[d][p] = off; off = align_up(off + N * ntile * sizeof(float), 256); }
  for (int d = 0; d < 2; ++d) { l.hist[d] = off; off = align_up(off + N * T * sizeof(float), 256); }
  for (int d = 0; d < 2; ++d) { l.cum[d] = off; off = align_up(off + N * T * sizeof(double), 256); }
  l.ll = off; off = align_up(off + N * sizeof(float), 256);
  l.lld = off; off = align_up(off + N * sizeof(double), 256);
  l.extra = off;
  if (g_path != 1 && dense_tc::supported(k, B, q)) off += dense_tc::extra_workspace(k, B, q);
  l.total = off;
  return l;
}

size_t workspace_size(int k, int B, int T, int q) {
  if (k <= 0 || B <= 0 || T <= 0 || q <= 0) return 0;
  return layout(k, B, T, q).total;
}

// out_alpha / out_beta double as the per-step stores; out_post needs both (the caller passes scratch for the one it
// does not want back: see hmm_b200.h)
// symbols + table (one-hot form) need the expanded tensor only where no engine gathers: bytes of it behind the layout
static size_t onehot_expand_bytes(int k, int B, int T, int q) {
  const bool tc = use_tc(k, B, q);
  const size_t seq_smem = ((size_t)q * q + SEQ_HDR) * sizeof(float) + align_up((size_t)T, 16);
  const bool seq = !tc && g_path == 0 && q < kTcMinStates && seq_smem <= 200 * 1024;
  return (tc || seq) ? 0 : align_up((size_t)k * B * T * q * sizeof(float), 256);
}
size_t workspace_size_onehot(int k, int B, int T, int q) {
  if (k <= 0 || B <= 0 || T <= 0 || q <= 0) return 0;
  return layout(k, B, T, q).total + onehot_expand_bytes(k, B, T, q);
}

int run(const float* E, const float* A, const float* At, const float* pi, int k, int B, int T, int q, unsigned flags,
        float* out_post, float* out_alpha, float* out_beta, float* loglik, void* workspace, size_t workspace_bytes,
        cudaStream_t st, const unsigned char* sym = nullptr, const float* Bt = nullptr, int nsym = 0) {
  HMM_REQUIRE((E || (sym && Bt && nsym > 0)) && A && pi, "dense: null input pointer");
  if (!E) {
    const size_t xb = onehot_expand_bytes(k, B, T, q);
    if (xb) {   // no gathering engine for this shape: materialise E behind the regular workspace
      const size_t base = layout(k, B, T, q).total;
      if (!workspace || workspace_bytes < base + xb) {
        set_error("dense onehot: workspace too small (%zu < %zu bytes)", workspace_bytes, base + xb);
        return HMM_B200_EWORKSPACE;
      }
      float* Ex = reinterpret_cast<float*>(static_cast<unsigned char*>(workspace) + base);
      const size_t rows = (size_t)k * B * T;
      onehot_expand_kernel<<<(unsigned)(rows < 65535u * 16u ? rows : 65535u * 16u), 256, 0, st>>>(sym, Bt, Ex, rows, B, T, q, nsym);
      HMM_CUDA_OK(cudaGetLastError());
      E = Ex; sym = nullptr;
    }
  }
  HMM_REQUIRE(k > 0 && B > 0 && T > 0 && q > 0, "dense: k, B, T, q must be positive (got %d,%d,%d,%d)", k, B, T, q);
  HMM_REQUIRE(cdiv(q, BN) <= MAXP, "dense: q = %d exceeds the supported %d states", q, MAXP * BN);
  HMM_REQUIRE(cdiv(B, BM) <= 65535 && k <= 65535, "dense: batch / model count too large");
  HMM_REQUIRE(out_post || out_alpha || out_beta || loglik, "dense: no output requested");
  const bool need_f = out_post || out_alpha || loglik, need_b = out_post || out_beta;
  HMM_REQUIRE(!out_post || (out_alpha && out_beta), "dense: the posterior needs alpha and beta buffers (scratch is fine)");
  HMM_REQUIRE(!need_b || At || use_tc(k, B, q), "dense: the backward pass needs A^T");
  const Layout l = layout(k, B, T, q);
  if (!workspace || workspace_bytes < l.total) {
    set_error("dense: workspace too small (%zu < %zu bytes)", workspace_bytes, l.total);
    return HMM_B200_EWORKSPACE;
  }
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  auto F = [&](size_t o) { return reinterpret_cast<float*>(ws + o); };
  auto D = [&](size_t o) { return reinterpret_cast<double*>(ws + o); };
  const bool tc = use_tc(k, B, q);
  const int ntile = tc ? dense_tc::slots(q) : cdiv(q, BN);
  const size_t N = (size_t)k * B;
  const dim3 grid(ntile, cdiv(B, BM), k);
  const int ib = 256, ig = (int)((N * q + ib - 1) / ib);
  float* llbuf = loglik ? loglik : F(l.ll);

  if (tc) {
    float* part[2][2] = {{F(l.part[0][0]), F(l.part[0][1])}, {F(l.part[1][0]), F(l.part[1][1])}};
    float* hist[2] = {F(l.hist[0]), F(l.hist[1])};
    const int rc = dense_tc::run_steps(E, sym, Bt, nsym, A, pi, k, B, T, q, need_f, need_b, out_alpha, out_beta, part, hist, ws + l.extra, st);
    if (rc) return rc;
  }
  // small q: one persistent launch for all steps of both directions (seq_kernel); g_path == 1 keeps the per-step engine
  const size_t seq_smem = ((size_t)q * q + SEQ_HDR) * sizeof(float) + (E ? 0 : align_up((size_t)T, 16));
  const bool seq = !tc && g_path == 0 && q < kTcMinStates && seq_smem <= 200 * 1024 && N <= 2147483647ull;
  if (seq) {
    SeqArgs a;
    a.E = E; a.A = A; a.At = At; a.pi = pi;
    a.sym = sym; a.Bt = Bt; a.nsym = nsym;
    a.store[0] = need_f ? out_alpha : nullptr; a.store[1] = need_b ? out_beta : nullptr;
    a.hist[0] = F(l.hist[0]); a.hist[1] = F(l.hist[1]);
    a.part_f = F(l.part[0][T & 1]);
    a.k = k; a.B = B; a.T = T; a.q = q; a.ntile = ntile;
    a.dir0 = need_f ? 0 : 1;
    const int nthreads = ((q + 31) / 32) * 32;
    const dim3 sgrid((unsigned)N, (need_f && need_b) ? 2 : 1);
    // register column length in steps of 16 states (q = 67 runs 80 FMAs per step, not 96); the dropped terms are 0 * 0
    int rc = 0;
    switch ((q + 15) / 16) {
      case 1: rc = seq_launch<16>(sgrid, nthreads, seq_smem, st, a); break;
      case 2: rc = seq_launch<32>(sgrid, nthreads, seq_smem, st, a); break;
      case 3: rc = seq_launch<48>(sgrid, nthreads, seq_smem, st, a); break;
      case 4: rc = seq_launch<64>(sgrid, nthreads, seq_smem, st, a); break;
      case 5: rc = seq_launch<80>(sgrid, nthreads, seq_smem, st, a); break;
      case 6: rc = seq_launch<96>(sgrid, nthreads, seq_smem, st, a); break;
      case 7: rc = seq_launch<112>(sgrid, nthreads, seq_smem, st, a); break;
      case 8: rc = seq_launch<128>(sgrid, nthreads, seq_smem, st, a); break;
      case 9: rc = seq_launch<144>(sgrid, nthreads, seq_smem, st, a); break;
      case 10: rc = seq_launch<160>(sgrid, nthreads, seq_smem, st, a); break;
      case 11: rc = seq_launch<176>(sgrid, nthreads, seq_smem, st, a); break;
      default: rc = seq_launch<192>(sgrid, nthreads, seq_smem, st, a); break;
    }
    if (rc) return rc;
  }
  for (int d = 0; d < 2 && !tc && !seq; ++d) {
    if ((d == 0 && !need_f) || (d == 1 && !need_b)) continue;
    init_state_kernel<<<ig, ib, 0, st>>>(F(l.st[d][0]), d == 0 ? pi : nullptr, k, B, q);
    for (int s = 0; s < T; ++s) {
      StepArgs a;
      const int t = d == 0 ? s : T - 1 - s;
      a.state = F(l.st[d][s & 1]); a.part_in = F(l.part[d][s & 1]);
      a.state_out = F(l.st[d][(s + 1) & 1]); a.part_out = F(l.part[d][(s + 1) & 1]);
      a.A = d == 0 ? A : At; a.E = E;
      a.store = d == 0 ? out_alpha : out_beta;
      a.scale_hist = F(l.hist[d]);
      a.k = k; a.B = B; a.T = T; a.q = q; a.t = t; a.ntile = ntile; a.init = s == 0; a.reverse = d;
      step_kernel<<<grid, 256, 0, st>>>(a);
    }
  }
  HMM_CUDA_OK(cudaGetLastError());
  const float* part_f = F(l.part[0][tc ? ((T - 1) & 1) : (T & 1)]);   // slot written by the last forward step
  cum_kernel<<<(unsigned)N, 32, 0, st>>>(need_f ? F(l.hist[0]) : nullptr, part_f, need_b ? F(l.hist[1]) : nullptr,
                                                 (out_alpha || out_post) ? D(l.cum[0]) : nullptr,
                                                 need_b ? D(l.cum[1]) : nullptr, need_f ? llbuf : nullptr,
                                                 need_f ? D(l.lld) : nullptr, (int)N, T, ntile);
  if (out_post || out_alpha || out_beta) {
    HMM_REQUIRE(N * (size_t)T < 2147483647ull, "dense: k*B*T too large for one launch");
    finish_kernel<<<(unsigned)(N * T), 128, 0, st>>>(need_f ? out_alpha : nullptr, need_b ? out_beta : nullptr, out_post,
                                                     F(l.hist[0]), part_f, D(l.cum[0]), D(l.cum[1]), D(l.lld), (int)N, T, q, ntile,
                                                     (flags & HMM_B200_NO_LOGLIK) ? 1 : 0,
                                                     (flags & HMM_B200_ALPHA_SCRATCH) ? 0 : 1, (flags & HMM_B200_BETA_SCRATCH) ? 0 : 1,
                                                     (flags & HMM_B200_ALPHA_NORMALIZED) ? 1 : 0);
  }
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int cell_step(const float* E, const float* scaled, const float* loglik, const float* A, int k, int n, int qc, int q, int reverse,
              int init, float* out, float* new_scaled, float* new_loglik, cudaStream_t st) {
  HMM_REQUIRE(E && scaled && loglik && A && out && new_scaled && new_loglik, "cell_step: null pointer");
  HMM_REQUIRE(k > 0 && n > 0 && q > 0 && (qc == 1 || qc == q), "cell_step: bad shape (k=%d n=%d qc=%d q=%d)", k, n, qc, q);
  const size_t rows = (size_t)k * n * qc;
  HMM_REQUIRE(rows < 2147483647ull, "cell_step: too many rows");
  const int threads = q >= 128 ? 128 : (q >= 64 ? 64 : 32);
  cell_step_kernel<<<(unsigned)rows, threads, (q + 32) * sizeof(float), st>>>(E, scaled, loglik, A, out, new_scaled, new_loglik, k, n,
                                                                            qc, q, reverse, init);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace dense
}  // namespace hmm

extern "C" {

size_t hmm_b200_dense_workspace_size(int num_models, int batch, int T, int q) {
  return hmm::dense::workspace_size(num_models, batch, T, q);
}

int hmm_b200_dense_run(const float* E, const float* A, const float* A_transposed, const float* pi, int num_models, int batch, int T,
                       int q, unsigned flags, float* out_post, float* out_alpha, float* out_beta, float* loglik,
                       void* workspace, size_t workspace_bytes, void* stream) {
  return hmm::dense::run(E, A, A_transposed, pi, num_models, batch, T, q, flags, out_post, out_alpha, out_beta, loglik, workspace,
                         workspace_bytes, static_cast<cudaStream_t>(stream));
}

size_t hmm_b200_dense_workspace_size_onehot(int num_models, int batch, int T, int q) {
  return hmm::dense::workspace_size_onehot(num_models, batch, T, q);
}

int hmm_b200_dense_run_onehot(const unsigned char* symbols, const float* Bt, int num_symbols, const float* A,
                              const float* A_transposed, const float* pi, int num_models, int batch, int T, int q, unsigned flags,
                              float* out_post, float* out_alpha, float* out_beta, float* loglik, void* workspace,
                              size_t workspace_bytes, void* stream) {
  return hmm::dense::run(nullptr, A, A_transposed, pi, num_models, batch, T, q, flags, out_post, out_alpha, out_beta, loglik,
                         workspace, workspace_bytes, static_cast<cudaStream_t>(stream), symbols, Bt, num_symbols);
}

int hmm_b200_cell_step(const float* E_t, const float* scaled, const float* loglik, const float* A, int num_models, int n,
                       int qc, int q, int reverse, int init, float* out, float* new_scaled, float* new_loglik, void* stream) {
  return hmm::dense::cell_step(E_t, scaled, loglik, A, num_models, n, qc, q, reverse, init, out, new_scaled, new_loglik,
                               static_cast<cudaStream_t>(stream));
}

}  // extern "C"

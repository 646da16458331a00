// Device-side helpers shared by the gene-HMM kernels (gene15.cu: products / boundary / finalize / gradient / Viterbi;
// gene15_fused.cu: the single-launch forward+posterior kernel).  See gene15.cu for the algorithm description.
#pragma once
#include <stdlib.h>
#include "common.cuh"

namespace hmm {
namespace dense { void set_path(int p); }
namespace dense_tc { void set_debug(void* p); }
namespace gene15 {

constexpr int Q = 15;      // states
constexpr int QP = 16;     // padded row
constexpr int SIN = 20;    // input floats per position: 15 class probabilities + 5 nucleotide one-hot
constexpr int NEDGE = 23;
constexpr int MAT = QP * QP;  // floats per stored matrix block
constexpr int BST = 20;       // row stride of blocks staged for the scans (16-byte rows; <= 2-way bank conflicts)
constexpr int BSZ = QP * BST; // floats per staged block
constexpr int G = 32;         // chunks per tile
constexpr int HEAP = 2 * G;   // blocks per tile in heap order: [1] tile product, [2..G) inner nodes, [G..2G) chunks
constexpr int MAX_SCAN_TILES = 60;    // K1's last CTA of a sequence scans up to this many tile matrices itself
constexpr int K1_THREADS = G * 16;
constexpr int KST = 20;       // K1: floats between the coefficient rows of consecutive positions (80 B: the position-per-lane
                              // stores and loads of a warp then fall into 32 different bank groups; 64 B would be 4-way conflicts)
constexpr unsigned EMIT_LOG_FLOORED = 0x40000000u;   // internal flag of emission_kernel (not part of the C ABI)
#ifndef HMM_K1_BLOCKS
#ifndef HMM_K1_BLOCKS
#define HMM_K1_BLOCKS 2               // CTAs of K1 per SM the register budget is sized for (1: 128 regs, 2: 64 regs)
#endif
#endif
constexpr int MAX_L = 40;             // longest chunk (shared memory bound)
constexpr int DIV_SHIFT = 21;         // i / L == (i * mL) >> 21 for i < 4096, L <= 48
constexpr double LN2D = 0.69314718055994531;
constexpr int T5_FLOATS = 2 * 125 * 12;

// edges in the reference's kernel order (from, to)
__device__ constexpr int ESRC[NEDGE] = {0, 0, 14, 7, 5, 4, 4, 8, 1, 1, 11, 5, 5, 9, 2, 2, 12, 6, 6, 10, 3, 3, 13};
__device__ constexpr int EDST[NEDGE] = {0, 7, 0, 5, 14, 5, 8, 1, 1, 11, 4, 6, 9, 2, 2, 12, 5, 4, 10, 3, 3, 13, 6};
// core predecessor of core state j; predecessor / successor of pass state 7+p
__device__ constexpr int SRC1[7] = {0, 1, 2, 3, 6, 4, 5};
__device__ constexpr int PRED[8] = {0, 4, 5, 6, 1, 2, 3, 5};
__device__ constexpr int SUCC[8] = {5, 1, 2, 3, 4, 5, 6, 0};

struct Args {
  const float* inputs;
  const float* A;
  const float* pi;
  const float* Bfull;
  const float* codon5;  // [2][125][12] 5-symbol codon tables (prepare_codon)
  const float* hints;
  int num_models, batch, T;
  unsigned flags;
  int L;           // chunk length
  unsigned mL;     // (1 << DIV_SHIFT) / L + 1
  int C;           // chunks per sequence = ceil(T / L)
  int P;           // tiles per sequence  = ceil(C / G)
  int top_scan;    // who scans the tile matrices: 0 boundary_kernel, 1 K1's last CTA of the sequence, 2 every K2 CTA itself
  // sequence sharding (multi-GPU): this rank holds positions [seq_off, seq_off + T) of sequences of total length
  // seq_total; halo nucleotides and boundary vectors come from the neighbours.
  int seq_off, seq_total;
  const float* halo;   // [NB][4][5] nucleotide flags of positions -2,-1 (left) and T,T+1 (right) or NULL
  // compact wire format: when cls16 is set, `inputs` is NULL and K1 unpacks its tile itself
  const unsigned short* cls16;   // [NB][T][15] class probabilities as bf16 bit patterns
  const unsigned char* nuc8;     // [NB][T] nucleotide codes 0..3 = A,C,G,T, anything else = N
  float* Mh;   // [NB][P][HEAP][256] per tile: product tree in heap order (node k = node 2k . node 2k+1)
  float* Mt;   // [NB][P][256] tile matrices (= heap node 1), contiguous per sequence for the top-level scan
  float* bvf;  // [NB][P][16] forward boundary vector entering tile p, slot 15 = exponent (int bits)
  float* bvr;  // [NB][P][16] backward boundary vector leaving tile p, slot 15 = exponent (int bits)
  const float* in_f;  // [NB][16] boundary entering this shard, slot 15 = exponent bits (NULL: pi)     multi-GPU
  const float* in_r;  // [NB][16] boundary leaving this shard (NULL: ones)
  float* llws;        // [NB] log-likelihood
  float* Ews;         // [NB][P][G*CS] floored emission rows of every tile in K2's shared-memory layout (CS = 16 L + 4)
  unsigned* cnt;      // [NB] tiles of the sequence finished by K1 (zeroed per call)
  unsigned* flg;      // [NB][P] fused kernel: tile p of sequence n has published its tile matrix when flg == tag
  unsigned tag;       // per-call nonce of the fused kernel's flags (no reset between calls)
  int run;            // fused kernel: consecutive positions per emission thread
  int ll_external;    // the log-likelihood in llws comes from the cross-shard scan: boundary_kernel leaves it alone
  long long* dbg;     // optional phase timestamps (tools/phase_times.py), normally NULL
  float* out_post;
  float* out_alpha;
  float* out_beta;
  float* loglik;
};

// ------------------------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ long long clock_fenced() { long long t; asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) :: "memory"); return t; }
#define HMM_MARK(base, k) do { if (a.dbg && threadIdx.x == 0) a.dbg[(size_t)(base) * 16 + (k)] = clock_fenced(); } while (0)
// wall-clock stamps (globaltimer, ns; comparable across SMs) and the SM id: slot 14 at CTA start, 15 at CTA end, 13 = smid
__device__ __forceinline__ long long gtimer() { long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) :: "memory"); return t; }
__device__ __forceinline__ int smid() { int r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }
#define HMM_WALL(base, k) do { if (a.dbg && threadIdx.x == 0) { a.dbg[(size_t)(base) * 16 + (k)] = gtimer(); a.dbg[(size_t)(base) * 16 + 13] = smid(); } } while (0)

// programmatic dependent launch: the next kernel of the stream may start its CTAs (their prologue) while this grid
// drains; it touches nothing this grid wrote before griddep_wait()
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ int fexp(float x) { return ((__float_as_int(x) >> 23) & 0xff) - 127; }
__device__ __forceinline__ int clamp_exp(int e) { return max(-126, min(126, e)); }
// 2^(-e), e clamped to [-126, 126]
__device__ __forceinline__ float pow2neg(int e) { return __int_as_float((127 - clamp_exp(e)) << 23); }
// 2^(-e) with e = exponent of g, for a positive normal g < 2^127, straight from the exponent bits
__device__ __forceinline__ float recip_pow2(float g, int& e) {
  const int bits = __float_as_int(g);
  e = (bits >> 23) - 127;
  return __int_as_float(0x7F000000 - (bits & 0x7F800000));
}
// exponent of a positive float, 0 for non-positive input
__device__ __forceinline__ int pexp(float x) { return x > 0.f ? fexp(x) : 0; }

__device__ __forceinline__ int half_max_int(int v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o, 16));
  return v;
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
__device__ __forceinline__ float sum15p(const float* v) {
  return ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7])) + (((v[8] + v[9]) + (v[10] + v[11])) +
         ((v[12] + v[13]) + v[14]));
}
__device__ __forceinline__ float sum7(const float (&v)[7]) { return ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + v[6]); }

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

// ---- bulk asynchronous copies (TMA, 1-D) completing on an mbarrier; sizes and addresses are multiples of 16 bytes ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
// generic-proxy writes to shared memory made visible to the async proxy (before a bulk store reads them)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

// ------------------------------------------------------------------------------------------------------------
// emission (gene_pred_hmm_emitter.py:93-121 class part, :231-263 3-mer part, kmer.py:3-47 soft 3-mers)
// ------------------------------------------------------------------------------------------------------------
struct EmisSmem {
  float B[Q][QP];          // [state][class], column 15 = 0
  float T5[2][125][12];    // [left|right][5-symbol 3-mer][E2,START,EI0-2,IE0-2,STOP, pad]; symbol 4 = uniform
};

// Builds the 5-symbol tables from the reference's (2,9,64) codon tables: symbol 4 (N / outside the sequence) is the
// average over the four bases (kmer.py:21-33).  One CTA; run once per parameter set.
static __global__ void __launch_bounds__(256) prepare_codon_kernel(const float* __restrict__ codon, float* __restrict__ out) {
  __shared__ float T5[2][125][12];
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < T5_FLOATS; i += nt) (&T5[0][0][0])[i] = 0.f;
  __syncthreads();
  for (int i = tid; i < 2 * 9 * 64; i += nt) {
    const int side = i / 576, r = i - side * 576, s = r >> 6, k = r & 63;
    const int nr = k >> 4, fr = (k >> 2) & 3, pv = k & 3;
    T5[side][(nr * 5 + fr) * 5 + pv][s] = codon[i];
  }
  __syncthreads();
  for (int i = tid; i < 2 * 16 * 9; i += nt) {  // pivot = 4
    const int side = i / 144, r = i - side * 144, hi = r / 9, s = r - hi * 9;
    const int base = ((hi >> 2) * 5 + (hi & 3)) * 5;
    const float* t = &T5[side][base][s];
    T5[side][base + 4][s] = 0.25f * ((t[0] + t[12]) + (t[24] + t[36]));
  }
  __syncthreads();
  for (int i = tid; i < 2 * 20 * 9; i += nt) {  // far = 4
    const int side = i / 180, r = i - side * 180, hi = r / 9, s = r - hi * 9;
    const int nr = hi / 5, pv = hi - nr * 5;
    const float* t = &T5[side][nr * 25 + pv][s];
    T5[side][nr * 25 + 20 + pv][s] = 0.25f * ((t[0] + t[60]) + (t[120] + t[180]));
  }
  __syncthreads();
  for (int i = tid; i < 2 * 25 * 9; i += nt) {  // near = 4
    const int side = i / 225, r = i - side * 225, lo = r / 9, s = r - lo * 9;
    const float* t = &T5[side][lo][s];
    T5[side][100 + lo][s] = 0.25f * ((t[0] + t[300]) + (t[600] + t[900]));
  }
  __syncthreads();
  for (int i = tid; i < T5_FLOATS; i += nt) out[i] = (&T5[0][0][0])[i];
}

// B rows and the 5-symbol codon tables into shared memory (no barrier inside)
__device__ __forceinline__ void load_emis_params(EmisSmem& sp, const float* __restrict__ Bm, const float* __restrict__ codon5) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int i = tid; i < Q * QP; i += nt) {
    const int q = i >> 4, s = i & 15;
    sp.B[q][s] = (s < Q) ? __ldg(Bm + q * Q + s) : 0.f;
  }
  const float4* src = reinterpret_cast<const float4*>(codon5);
  float4* dst = reinterpret_cast<float4*>(&sp.T5[0][0][0]);
  for (int i = tid; i < T5_FLOATS / 4; i += nt) dst[i] = __ldg(src + i);
}

// pointer to the 5 nucleotide flags of (shard-local) position t, or nullptr when t lies outside the whole sequence
__device__ __forceinline__ const float* nuc_ptr(const Args& a, int n, int t) {
  if (t >= 0 && t < a.T) return a.inputs + ((size_t)n * a.T + t) * SIN + Q;
  if (a.halo == nullptr) return nullptr;
  const int gt = a.seq_off + t;
  if (gt < 0 || gt >= a.seq_total) return nullptr;
  if (t < 0 && t >= -2) return a.halo + ((size_t)n * 4 + (t + 2)) * 5;
  if (t >= a.T && t < a.T + 2) return a.halo + ((size_t)n * 4 + 2 + (t - a.T)) * 5;
  return nullptr;
}

// 0..3 exactly one-hot base, 4 uniform (N flag exactly 1 with no base set), -1 anything else
__device__ __forceinline__ int code_of(float x, float y, float z, float w, float f) {
  const int zeros = (x == 0.f) + (y == 0.f) + (z == 0.f) + (w == 0.f);
  if (f == 1.f) return zeros == 4 ? 4 : -1;
  const int ones = (x == 1.f) + (y == 1.f) + (z == 1.f) + (w == 1.f);
  if (ones != 1 || zeros != 3) return -1;
  return (y == 1.f) ? 1 : (z == 1.f) ? 2 : (w == 1.f) ? 3 : 0;
}
__device__ __forceinline__ int nuc_code(const Args& a, int n, int t) {
  const float* p = nuc_ptr(a, n, t);
  if (p == nullptr) return 4;   // outside the sequence: 1/4 each (kmer.py:27-33)
  return code_of(__ldg(p), __ldg(p + 1), __ldg(p + 2), __ldg(p + 3), __ldg(p + 4));
}

// distribution over A,C,G,T at position t (kmer.py:21-33)
__device__ __forceinline__ float4 nuc_dist(const Args& a, int n, int t) {
  const float* p = nuc_ptr(a, n, t);
  if (p == nullptr) return make_float4(.25f, .25f, .25f, .25f);
  const float add = (__ldg(p + 4) == 1.0f) ? 0.25f : 0.f;
  return make_float4(__ldg(p) + add, __ldg(p + 1) + add, __ldg(p + 2) + add, __ldg(p + 3) + add);
}
__device__ __forceinline__ float comp(const float4 p, int i) { return i == 0 ? p.x : i == 1 ? p.y : i == 2 ? p.z : p.w; }

// generic soft 3-mer contraction (rare: a nucleotide row that is neither one-hot nor N)
static __device__ __noinline__ void codon_side_soft(const EmisSmem& sp, int side, const float4 pv, const float4 nr, const float4 fr,
                                             float* out) {
  for (int s = 0; s < 9; ++s) out[s] = 0.f;
  for (int i = 0; i < 64; ++i) {
    const int a = i >> 4, b = (i >> 2) & 3, c = i & 3;
    const float w = comp(nr, a) * comp(fr, b) * comp(pv, c);
    if (w != 0.f) {
      const float* t = sp.T5[side][(a * 5 + b) * 5 + c];
      for (int s = 0; s < 9; ++s) out[s] = fmaf(w, t[s], out[s]);
    }
  }
}

// codon factors of the 9 constrained states at position t; cd points at the code of t (cd[-2..2] valid).
// The soft path (a nucleotide row that is neither one-hot nor N) works on its own scratch array, so that the arrays
// of the table path stay in registers.
__device__ __forceinline__ void codon_factors(const Args& a, const EmisSmem& sp, int n, int t, const signed char* cd, float (&cod)[9]) {
  float fl[9], fr[9];
  const int c0 = cd[0], l1 = cd[1], l2 = cd[2], r1 = cd[-1], r2 = cd[-2];
  if ((c0 | l1 | l2) >= 0) {
    const float4* t4 = reinterpret_cast<const float4*>(sp.T5[0][(l1 * 5 + l2) * 5 + c0]);
    const float4 u0 = t4[0], u1 = t4[1];
    fl[0] = u0.x; fl[1] = u0.y; fl[2] = u0.z; fl[3] = u0.w; fl[4] = u1.x; fl[5] = u1.y; fl[6] = u1.z; fl[7] = u1.w;
    fl[8] = sp.T5[0][(l1 * 5 + l2) * 5 + c0][8];
  } else {
    float tmp[12];
    codon_side_soft(sp, 0, nuc_dist(a, n, t), nuc_dist(a, n, t + 1), nuc_dist(a, n, t + 2), tmp);
#pragma unroll
    for (int s = 0; s < 9; ++s) fl[s] = tmp[s];
  }
  if ((c0 | r1 | r2) >= 0) {
    const float4* t4 = reinterpret_cast<const float4*>(sp.T5[1][(r1 * 5 + r2) * 5 + c0]);
    const float4 u0 = t4[0], u1 = t4[1];
    fr[0] = u0.x; fr[1] = u0.y; fr[2] = u0.z; fr[3] = u0.w; fr[4] = u1.x; fr[5] = u1.y; fr[6] = u1.z; fr[7] = u1.w;
    fr[8] = sp.T5[1][(r1 * 5 + r2) * 5 + c0][8];
  } else {
    float tmp[12];
    codon_side_soft(sp, 1, nuc_dist(a, n, t), nuc_dist(a, n, t - 1), nuc_dist(a, n, t - 2), tmp);
#pragma unroll
    for (int s = 0; s < 9; ++s) fr[s] = tmp[s];
  }
#pragma unroll
  for (int s = 0; s < 9; ++s) cod[s] = fl[s] * fr[s];
}

// Emission probabilities of NP (shard-local) positions t[0..NP) of sequence n, computed together so that every B row
// fetched from shared memory serves all of them.  FLOOR: max(e, eps) as the recursion applies it (MsaHmmCell.py:87).
// rows[k]: the 20 input floats of position t[k] (16-byte aligned; global memory, or the tile staged in shared memory)
template <bool FLOOR, int NP>
__device__ __forceinline__ void emit_x(const Args& a, const EmisSmem& sp, int n, const int (&t)[NP],
                                       const signed char* const (&cd)[NP], const float4 (&x)[NP][4], float (&e)[NP][Q]) {
  float cod[NP][9];
#pragma unroll
  for (int k = 0; k < NP; ++k) codon_factors(a, sp, n, t[k], cd[k], cod[k]);
  const float train = (a.flags & HMM_B200_TRAINING) ? 1e-7f : 0.f;
  const float free_cod = (1.0f / 4096.0f) + train;
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const float4* b = reinterpret_cast<const float4*>(sp.B[q]);
    const float4 b0 = b[0], b1 = b[1], b2 = b[2], b3 = b[3];   // b3.w == 0 removes x3.w (a nucleotide flag)
#pragma unroll
    for (int k = 0; k < NP; ++k) {
      const float s = dot4(x[k][3], b3, dot4(x[k][2], b2, dot4(x[k][1], b1, dot4(x[k][0], b0, 0.f))));
      e[k][q] = s * ((q < 6) ? free_cod : cod[k][q < 6 ? 0 : q - 6] + train);
    }
  }
  if (a.hints != nullptr) {
    const float* h = a.hints + (size_t)n * 2 * Q;
#pragma unroll
    for (int k = 0; k < NP; ++k) {
      const int g = a.seq_off + t[k];
      if (g == 0) {
#pragma unroll
        for (int q = 0; q < Q; ++q) e[k][q] *= __ldg(h + q);
      }
      if (g == a.seq_total - 1) {
#pragma unroll
        for (int q = 0; q < Q; ++q) e[k][q] *= __ldg(h + Q + q);
      }
    }
  }
  if (FLOOR) {
#pragma unroll
    for (int k = 0; k < NP; ++k) {
#pragma unroll
      for (int q = 0; q < Q; ++q) e[k][q] = fmaxf(e[k][q], kEps);
    }
  }
}

// the same with the rows read here
template <bool FLOOR, int NP>
__device__ __forceinline__ void emit_n(const Args& a, const EmisSmem& sp, int n, const int (&t)[NP],
                                       const signed char* const (&cd)[NP], const float* const (&rows)[NP], float (&e)[NP][Q]) {
  float4 x[NP][4];
#pragma unroll
  for (int k = 0; k < NP; ++k) {
    const float4* row = reinterpret_cast<const float4*>(rows[k]);
    x[k][0] = row[0]; x[k][1] = row[1]; x[k][2] = row[2]; x[k][3] = row[3];
  }
  emit_x<FLOOR, NP>(a, sp, n, t, cd, x, e);
}

// two positions i0 (valid) and i1 (valid iff two) of the tile
template <bool FLOOR>
__device__ __forceinline__ void emit_pair(const Args& a, const EmisSmem& sp, int n, int t0, int t1, bool two,
                                          const signed char* cd0, const signed char* cd1, float (&e0)[Q], float (&e1)[Q]) {
  if (!two) { t1 = t0; cd1 = cd0; }
  const int t[2] = {t0, t1};
  const signed char* const cd[2] = {cd0, cd1};
  const float* const rows[2] = {a.inputs + ((size_t)n * a.T + t0) * SIN, a.inputs + ((size_t)n * a.T + t1) * SIN};
  float e[2][Q];
  emit_n<FLOOR, 2>(a, sp, n, t, cd, rows, e);
#pragma unroll
  for (int q = 0; q < Q; ++q) { e0[q] = e[0][q]; e1[q] = e[1][q]; }
}

// codes of positions [t_lo - 2, t_lo + npos + 2) -> codes[0 .. npos + 4); up to 4 positions per thread with all
// loads in flight at once (npos <= G * MAX_L = 1536, blockDim >= 512 in the kernels that matter)
__device__ __forceinline__ void codes_stage(const Args& a, signed char* codes, int n, int t_lo, int npos) {
  const int nt = blockDim.x;
  for (int base = threadIdx.x; base < npos + 4; base += 4 * nt) {
    float4 u[4], v[4];
    int t[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      t[k] = t_lo - 2 + base + k * nt;
      if (base + k * nt < npos + 4 && t[k] >= 0 && t[k] < a.T) {
        const float* p = a.inputs + ((size_t)n * a.T + t[k]) * SIN + 12;
        u[k] = ldg4(p); v[k] = ldg4(p + 4);   // u.w = A flag, v = C,G,T,N flags
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = base + k * nt;
      if (i < npos + 4) {
        const int c = (t[k] >= 0 && t[k] < a.T) ? code_of(u[k].w, v[k].x, v[k].y, v[k].z, v[k].w) : nuc_code(a, n, t[k]);
        codes[i] = (signed char)c;
      }
    }
  }
}

// the same from the tile staged in shared memory: IN holds positions [t_in0, t_in0 + n_in) (20 floats each)
__device__ __forceinline__ void codes_from_smem(const Args& a, signed char* codes, const float* IN, int n, int t_lo, int npos,
                                                int t_in0, int n_in) {
  for (int i = threadIdx.x; i < npos + 4; i += blockDim.x) {
    const int t = t_lo - 2 + i, j = t - t_in0;
    int c;
    if (j >= 0 && j < n_in) {
      const float* p = IN + j * SIN + 12;
      const float4 v = *reinterpret_cast<const float4*>(p + 4);
      c = code_of(p[3], v.x, v.y, v.z, v.w);
    } else {
      c = nuc_code(a, n, t);
    }
    codes[i] = (signed char)c;
  }
}

__device__ __forceinline__ int divL(const Args& a, int i) { return (int)(((unsigned)i * a.mL) >> DIV_SHIFT); }
__device__ __forceinline__ int tile_row_offset(const Args& a, int i, int CS) {
  const int ci = divL(a, i);
  return ci * CS + (i - ci * a.L) * QP;
}

// ------------------------------------------------------------------------------------------------------------
// matrix blocks
// ------------------------------------------------------------------------------------------------------------

// Row 15 of a block whose rows 0-14 hold [M-hat | s'] (row stride QP); one half warp, lane hl.
__device__ __forceinline__ void finish_block(float* blk, int hl, float sp /* s'[hl], 0 for hl = 15 */, int E) {
  const int col = min(hl, Q - 1);
  float cs = 0.f;
#pragma unroll
  for (int i = 0; i < Q; ++i) cs = fmaf(__shfl_sync(0xffffffffu, sp, i, 16), blk[i * QP + col], cs);
  blk[15 * QP + hl] = (hl < Q) ? cs : __int_as_float(E);
}

// strip (row r, columns 4 jq .. 4 jq + 3) of C' = A-hat . diag(s'_B) B-hat, normalised by the row sum; the last
// strip's .w carries the un-normalised relative scale of the row, s'_A[r] * rowsum
__device__ __forceinline__ void mm_strip(const float* __restrict__ Ab, const float* __restrict__ Bb, float* __restrict__ Cb, int r, int jq) {
  float ar[QP];
  load_row16(Ab + r * QP, ar);   // ar[15] = s'_A[r]
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  float rs = 0.f;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    const float4 b = *reinterpret_cast<const float4*>(Bb + k * QP + 4 * jq);
    const float ak = ar[k] * Bb[k * QP + 15];
    rs += ak;
    acc.x = fmaf(ak, b.x, acc.x); acc.y = fmaf(ak, b.y, acc.y); acc.z = fmaf(ak, b.z, acc.z); acc.w = fmaf(ak, b.w, acc.w);
  }
  const float inv = __frcp_rn(fmaxf(rs, 1e-37f));
  float4 o = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
  if (jq == 3) o.w = ar[15] * rs;
  *reinterpret_cast<float4*>(Cb + r * QP + 4 * jq) = o;
}

// second half of a product, by one half warp (all 32 lanes of the warp must call): common power of two out of
// the row scales, column sums, exponent.  `write` = false computes but stores nothing.
__device__ __forceinline__ void mm_finish(const float* __restrict__ Ab, const float* __restrict__ Bb, float* __restrict__ Cb, int hl, bool write) {
  const float t = (hl < Q) ? Cb[hl * QP + 15] : 0.f;
  const int emax = half_max_int((hl < Q && t > 0.f) ? fexp(t) : -100000);
  const int e = emax == -100000 ? 0 : emax;
  const float sp = t * pow2neg(e);
  const int E = __float_as_int(Ab[255]) + __float_as_int(Bb[255]) + clamp_exp(e);
  const int col = min(hl, Q - 1);
  float cs = 0.f;
#pragma unroll
  for (int i = 0; i < Q; ++i) cs = fmaf(__shfl_sync(0xffffffffu, sp, i, 16), Cb[i * QP + col], cs);
  __syncwarp();
  if (write) {
    if (hl < Q) Cb[hl * QP + 15] = sp;
    Cb[15 * QP + hl] = (hl < Q) ? cs : __int_as_float(E);
  }
}

// One level of the product tree: dst[pr] = src[2 pr] . src[2 pr + 1] for pr < np, all threads of the CTA.
// Ends with a barrier.
__device__ __forceinline__ void tree_level(const float* __restrict__ src, float* __restrict__ dst, int np) {
  const int tid = threadIdx.x, nt = blockDim.x;
  for (int item = tid; item < np * 64; item += nt) {
    const int pr = item >> 6, r = (item >> 2) & 15, jq = item & 3;
    if (r < Q) mm_strip(src + (2 * pr) * MAT, src + (2 * pr + 1) * MAT, dst + pr * MAT, r, jq);
  }
  __syncthreads();
  const int hw = tid >> 4, hl = tid & 15, nhw = nt >> 4;
  for (int base = (hw & ~1); base < np; base += nhw) {
    const int pr = base + (hw & 1);
    const bool valid = pr < np;
    const int prc = valid ? pr : np - 1;
    mm_finish(src + (2 * prc) * MAT, src + (2 * prc + 1) * MAT, dst + prc * MAT, hl, valid);
  }
  __syncthreads();
}

// ---- the same product on the tensor cores: one warp per product, mma.sync m16n8k8 TF32 with the 3-pass split
// (hi.hi + lo.hi + hi.lo), which keeps fp32-level accuracy.  A-hat is the left operand as stored; the right operand
// is diag(s'_B) B-hat with column 15 = s'_B, so that column 15 of the product is the row sum (B-hat rows sum to 1).
__device__ __forceinline__ unsigned tf32_bits(float x) {
  unsigned r;
  asm("cvt.rna.tf32.f32 %0, %1;\n" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void split_tf32(float x, unsigned& hi, unsigned& lo) {
  hi = tf32_bits(x);
  lo = tf32_bits(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// C = A . B for stored blocks (row stride QP, shared memory), by one warp; same result format as mm_strip + mm_finish
__device__ __forceinline__ void warp_mm_tc(const float* __restrict__ Ab, const float* __restrict__ Bb, float* __restrict__ Cb, int lane) {
  const int g = lane >> 2, t = lane & 3;
  float c[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
  for (int kb = 0; kb < QP; kb += 8) {
    const int k0 = kb + t, k1 = kb + t + 4;
    unsigned ah[4], al[4];
    split_tf32(Ab[g * QP + k0], ah[0], al[0]);
    split_tf32(Ab[(g + 8) * QP + k0], ah[1], al[1]);
    split_tf32(Ab[g * QP + k1], ah[2], al[2]);
    split_tf32(Ab[(g + 8) * QP + k1], ah[3], al[3]);
    const float s0 = Bb[k0 * QP + 15];
    const float s1 = (k1 < Q) ? Bb[k1 * QP + 15] : 0.f;     // row 15 of the right operand is not a matrix row
#pragma unroll
    for (int nb = 0; nb < 2; ++nb) {
      const int nn = nb * 8 + g;
      float b0 = Bb[k0 * QP + nn], b1 = Bb[k1 * QP + nn];
      b0 = (nn == 15) ? s0 : b0 * s0;
      b1 = (k1 < Q) ? ((nn == 15) ? s1 : b1 * s1) : 0.f;
      unsigned bh[2], bl[2];
      split_tf32(b0, bh[0], bl[0]);
      split_tf32(b1, bh[1], bl[1]);
      mma_tf32(c[nb], ah, bh);
      mma_tf32(c[nb], al, bh);
      mma_tf32(c[nb], ah, bl);
    }
  }
  // this lane holds rows g (c[.][0..1]) and g+8 (c[.][2..3]), columns nb*8 + 2t, +1; row 15 is not a matrix row
  const bool r1ok = g < 7;
  const int src = (lane & ~3) | 3;
  const float rs0 = __shfl_sync(0xffffffffu, c[1][1], src);
  const float rs1 = __shfl_sync(0xffffffffu, c[1][3], src);
  const float inv0 = __frcp_rn(fmaxf(rs0, 1e-37f)), inv1 = __frcp_rn(fmaxf(rs1, 1e-37f));
  const float t0 = Ab[g * QP + 15] * rs0;
  const float t1 = r1ok ? Ab[(g + 8) * QP + 15] * rs1 : 0.f;
  int em = max(t0 > 0.f ? fexp(t0) : -100000, t1 > 0.f ? fexp(t1) : -100000);
#pragma unroll
  for (int o = 4; o < 32; o <<= 1) em = max(em, __shfl_xor_sync(0xffffffffu, em, o));
  const int e = em == -100000 ? 0 : em;
  const float ps = pow2neg(e);
  const float sp0 = t0 * ps, sp1 = t1 * ps;
  const int E = __float_as_int(Ab[MAT - 1]) + __float_as_int(Bb[MAT - 1]) + clamp_exp(e);
  float v[2][4];
#pragma unroll
  for (int nb = 0; nb < 2; ++nb) {
    v[nb][0] = c[nb][0] * inv0; v[nb][1] = c[nb][1] * inv0;
    v[nb][2] = r1ok ? c[nb][2] * inv1 : 0.f; v[nb][3] = r1ok ? c[nb][3] * inv1 : 0.f;
  }
  if (t == 3) { v[1][1] = sp0; v[1][3] = sp1; }      // column 15 = relative row scale
  // column sums of diag(s') C-hat over the 15 rows
  float cs[2][2];
#pragma unroll
  for (int nb = 0; nb < 2; ++nb) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float x = fmaf(sp0, v[nb][h], sp1 * v[nb][2 + h]);
#pragma unroll
      for (int o = 4; o < 32; o <<= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      cs[nb][h] = x;
    }
  }
#pragma unroll
  for (int nb = 0; nb < 2; ++nb) {
    *reinterpret_cast<float2*>(Cb + g * QP + nb * 8 + 2 * t) = make_float2(v[nb][0], v[nb][1]);
    if (r1ok) *reinterpret_cast<float2*>(Cb + (g + 8) * QP + nb * 8 + 2 * t) = make_float2(v[nb][2], v[nb][3]);
  }
  if (g == 7) {   // row 15: column sums, exponent in slot 15
    *reinterpret_cast<float2*>(Cb + 15 * QP + 2 * t) = make_float2(cs[0][0], cs[0][1]);
    *reinterpret_cast<float2*>(Cb + 15 * QP + 8 + 2 * t) = make_float2(cs[1][0], (t == 3) ? __int_as_float(E) : cs[1][1]);
  }
}

// One level of the product tree on the tensor cores: dst[pr] = src[2 pr] . src[2 pr + 1], one warp per product.
// Ends with a barrier.
__device__ __forceinline__ void tree_level_tc(const float* __restrict__ src, float* __restrict__ dst, int np) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int pr = warp; pr < np; pr += nw) warp_mm_tc(src + (2 * pr) * MAT, src + (2 * pr + 1) * MAT, dst + pr * MAT, lane);
  __syncthreads();
}

// Boundary vectors during scans: lane hl of a half warp holds v[hl]; e = power-of-two exponent (replicated):
// true vector = v * 2^e.
struct HV {
  float v;
  int e;
};

// One scan step for both half warps of a warp at once (no divergence: the shuffles are warp-wide).
//   bwd == 0 (row vector):    v <- v . M;   bwd == 1 (column vector):  r <- M . r,   M = 2^E diag(s') M-hat.
// blk: staged block (row stride BST).  ones: 16 floats of 1.0 in shared memory.
// `commit` == false leaves h untouched (the two half warps may need different trip counts).
__device__ __forceinline__ void matvec(const float* __restrict__ blk, const float* __restrict__ ones, int hl, int bwd,
                                       bool commit, HV& h) {
  const int row = min(hl, Q - 1);
  const float sp = blk[row * BST + 15];
  const int Eb = __float_as_int(blk[15 * BST + 15]);
  const float u = (hl < Q) ? (bwd ? h.v : h.v * sp) : 0.f;
  const float* mp = blk + (bwd ? row * BST : row);
  const int st = bwd ? 1 : BST;
  const float* dp = bwd ? blk + 15 * BST : ones;
  float acc0 = 0.f, acc1 = 0.f, s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int k = 0; k < Q; ++k) {
    const float uk = __shfl_sync(0xffffffffu, u, k, 16);
    const float c = mp[k * st];
    const float d = dp[k];
    if (k & 1) { acc1 = fmaf(uk, c, acc1); s1 = fmaf(uk, d, s1); } else { acc0 = fmaf(uk, c, acc0); s0 = fmaf(uk, d, s0); }
  }
  const int ex = pexp(s0 + s1);
  const float out = (bwd ? sp : 1.f) * (acc0 + acc1) * pow2neg(ex);
  if (commit) {
    h.v = (hl < Q) ? out : 0.f;
    h.e += clamp_exp(ex) + Eb;
  }
}

// The same step on a raw stored block (row stride QP, no padding; shared or global memory).  The forward half reads
// columns (conflict-free); the backward half walks its row along a diagonal (k = hl + step mod 16), so that the 16
// lanes hit 16 different banks and the vector entries arrive by a rotating shuffle.
__device__ __forceinline__ void matvec_raw(const float* __restrict__ blk, int hl, int bwd, bool commit, HV& h) {
  const int row = min(hl, Q - 1);
  const float sp = blk[row * QP + 15];
  const int Eb = __float_as_int(blk[MAT - 1]);
  const float u = (hl < Q) ? (bwd ? h.v : h.v * sp) : 0.f;
  float acc0 = 0.f, acc1 = 0.f, s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int k = 0; k < QP; ++k) {
    const int kk = bwd ? ((hl + k) & 15) : k;
    const float uk = __shfl_sync(0xffffffffu, u, kk, 16);
    const bool ok = kk < Q;                      // slot 15 holds scales / exponent bits, never a matrix entry
    float c = bwd ? blk[row * QP + kk] : blk[kk * QP + hl];
    float d = bwd ? blk[15 * QP + kk] : 1.f;
    c = ok ? c : 0.f;
    d = ok ? d : 0.f;
    if (k & 1) { acc1 = fmaf(uk, c, acc1); s1 = fmaf(uk, d, s1); } else { acc0 = fmaf(uk, c, acc0); s0 = fmaf(uk, d, s0); }
  }
  // The exponent taken out must be the SAME in the 16 lanes of a half warp (h.e is read from one lane, h.v from all).  The
  // backward half adds its terms in a lane-dependent (rotated) order, so its sums can differ in the last bit and, next to a
  // power of two, in the exponent: every lane uses lane 0's sum.
  const float ssum = __shfl_sync(0xffffffffu, s0 + s1, 0, 16);
  const int ex = pexp(ssum);
  const float out = (bwd ? sp : 1.f) * (acc0 + acc1) * pow2neg(ex);
  if (commit) {
    h.v = (hl < Q) ? out : 0.f;
    h.e += clamp_exp(ex) + Eb;
  }
}

// stage nblk stored blocks (row stride QP, global) into shared memory with row stride BST, asynchronously
__device__ __forceinline__ void stage_blocks_async(float* dst, const float* __restrict__ src, int nblk) {
  for (int i = threadIdx.x; i < nblk * (MAT / 4); i += blockDim.x) {
    const int b = i >> 6, r = (i >> 2) & 15, c = i & 3;
    cp_async16(dst + b * BSZ + r * BST + 4 * c, src + 4 * i);
  }
}


__device__ __forceinline__ void load_edges(const float* Aw, float (&w)[NEDGE]) {
#pragma unroll
  for (int e = 0; e < NEDGE; ++e) w[e] = Aw[e];
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

__device__ __forceinline__ void named_barrier_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// The top-level scan (TotalProbabilityCell.py:30-49 in scaled probability space) done by every K2 CTA for itself, by one
// warp, from the sequence's tile matrices in global memory (L2): lanes 0-15 run pi . M_0 ... M_{P-1} (snapshot on
// entering tile p; the end gives the log-likelihood), lanes 16-31 run M_{P-1} ... M_{p+1} . 1.  The next block's
// entries are fetched while the current one is used.  Returns the tile's boundary vector (entering for the forward half,
// leaving for the backward half); *ll is valid in lane 0.
constexpr int MAX_K2_SCAN = 16;       // sequences of up to this many tiles are scanned inside K2 (no counter, no extra kernel)
__device__ __forceinline__ HV scan_tiles_global(const float* Mt, int P, int p, const float* __restrict__ pi_m, int hl, int hi,
                                                float* ll) {
  const int row = min(hl, Q - 1);
  HV h, snap;
  h.v = (hl < Q) ? (hi == 0 ? __ldg(pi_m + hl) : 1.f) : 0.f;
  h.e = 0;
  snap = h;
  const int steps = hi == 0 ? P : P - 1 - p;
  float v[QP], sp, dcol;
  int Eb;
  auto fetch = [&](int it, float (&vv)[QP], float& spv, float& dv, int& ev) {
    const float* blk = Mt + (size_t)(hi == 0 ? it : P - 1 - it) * MAT;   // it < P: always inside the sequence's list
#pragma unroll
    for (int k = 0; k < QP; ++k) vv[k] = __ldcg(hi == 0 ? blk + k * QP + hl : blk + row * QP + k);
    spv = __ldcg(blk + row * QP + 15);
    dv = __ldcg(blk + 15 * QP + hl);
    ev = __float_as_int(__ldcg(blk + MAT - 1));
  };
  // both halves run P rounds (the shuffles are warp-wide); the backward half commits the first P - 1 - p of them, its
  // remaining rounds read blocks p .. 0 for nothing
  fetch(0, v, sp, dcol, Eb);
  for (int it = 0; it < P; ++it) {
    float vn[QP], spn = 0.f, dn = 0.f;
    int En = 0;
    const bool more = it + 1 < P;
    if (more) fetch(it + 1, vn, spn, dn, En);
    if (hi == 0 && it == p) snap = h;
    const float u = (hl < Q) ? (hi ? h.v : h.v * sp) : 0.f;
    float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
    for (int k = 0; k < Q; ++k) {
      const float uk = __shfl_sync(0xffffffffu, u, k, 16);
      if (k & 1) acc1 = fmaf(uk, v[k], acc1); else acc0 = fmaf(uk, v[k], acc0);
    }
    const float s = half_sum(hi ? ((hl < Q) ? u * dcol : 0.f) : u);
    const int ex = pexp(s);
    const float out = (hi ? sp : 1.f) * (acc0 + acc1) * pow2neg(ex);
    if (it < steps) {
      h.v = (hl < Q) ? out : 0.f;
      h.e += clamp_exp(ex) + Eb;
    }
    if (more) {
#pragma unroll
      for (int k = 0; k < QP; ++k) v[k] = vn[k];
      sp = spn; dcol = dn; Eb = En;
    }
  }
  const float tot = half_sum(h.v);
  *ll = (float)((double)h.e * LN2D + (double)logf(tot));   // meaningful in the forward half only
  return hi == 0 ? snap : h;
}

__device__ __forceinline__ void identity_block(float* blk, int lane) {
  for (int i = lane; i < MAT; i += 32) {
    const int r = i >> 4, c = i & 15;
    float v;
    if (r < Q) v = (c == r || c == 15) ? 1.f : 0.f;          // e_r | s' = 1
    else v = (c < Q) ? 1.f : __int_as_float(0);              // column sums 1 | E = 0
    blk[i] = v;
  }
}


// ---- sequence sharding: the block / flag tables of all ranks (peer memory) ----
constexpr int MAX_PEERS = 16;
struct Peers {
  float* blocks[MAX_PEERS];      // block table of every rank: [2][world][NB][256] (two step parities)
  unsigned* flags[MAX_PEERS];    // flag table of every rank:  [2][world][NB]
  int rank, world;
  unsigned tag;                  // step number (same on every rank); parity selects the table half
};

__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}


}  // namespace gene15
}  // namespace hmm

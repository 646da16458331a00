// 15-state gene-structure HMM: forward + posterior in ONE launch (the config-2 headline path).
//
// One 256-thread CTA per tile (G = 32 chunks x L positions of one sequence), two CTAs per SM, all tiles of the call
// co-resident, so a tile can wait for the other tiles of its sequence inside the kernel:
//
//   1 emission      a warp walks its block of positions in steps of 32 consecutive positions (lane = position), inputs
//                   read straight from HBM / L2 (the tile is prefetched towards L2 at kernel start, the next step's rows
//                   are in flight during the current one).  The class-probability contraction runs on packed FFMA2 (two
//                   states per instruction; B sits in shared memory as state pairs, the class probability is the scalar
//                   operand), the 3-mer factors are two table rows.  The floored emission row goes to the L2-resident
//                   workspace (phase 5 streams it back), the coefficient row of the two-step recurrence (gene15.cu, K1) to
//                   shared memory; neighbouring positions' nucleotide codes and pass-state emissions come by shuffles, so
//                   the whole phase needs ONE barrier (block boundaries are patched behind it).
//   2 rows          thread = two conditional rows of one chunk's transfer matrix, packed (FFMA2 with the coefficient as
//                   the scalar operand): 15 FFMA2 per position for two rows.
//   3 tree          the 5-level product tree of gene15.cu (mma.sync TF32 x 3), kept in shared memory.
//   4 exchange      the tile matrix goes to the workspace with a per-call tag; one warp polls the tags of the sequence's
//                   tiles (ld.acquire), the sequence's tile matrices are staged in shared memory and scanned by two warps,
//                   forward and backward (the role of TotalProbabilityCell, TotalProbabilityCell.py:30-49), and the tree is
//                   swept down to the chunk boundary vectors.
//   5 recursion     the exact recursions of MsaHmmCell.py:73-106 inside every chunk, both directions meeting in the middle
//                   (gene15.cu, K2); emission rows arrive through a 6-deep cp.async ring per thread, so that only the
//                   own rows / posterior products (Y) need shared memory: 110 KB per CTA instead of 153.
//   6 output        log posterior through a per-warp staging buffer, coalesced.
//
// Against the two-kernel path: no product tree / emission rows re-read through bulk copies, no second launch, no second
// wave (finalize_kernel fits one CTA per SM), half the FMA issue slots in phases 1 and 2.
// The same arithmetic as gene15.cu in the same order (floors, scales as exact powers of two): see the comments there.
// The launches are cooperative (launch_coresident): the in-kernel waits need the whole grid resident.
// Also in this file: the shard mode of the kernel (reducer CTAs: one launch per sequence-sharded step, template SHARD) and
// boundary_groups_kernel (boundary vectors of long tile lists behind products_kernel in one launch).
#include <atomic>
#include <chrono>
#include "gene15_dev.cuh"

namespace hmm {
namespace gene15 {

typedef unsigned long long u64;
constexpr int FT = 256;        // threads per CTA
constexpr int FRING = 6;       // emission rows in flight per recursion thread (an L2 round trip is ~4 recursion steps)
constexpr int FSCAN = 24;      // tile matrices of one sequence staged in shared memory for the top-level scan

__device__ __forceinline__ u64 pk2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ u64 bc2(float x) { u64 r; asm("mov.b64 %0, {%1, %1};" : "=l"(r) : "f"(x)); return r; }
__device__ __forceinline__ void upk2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

__device__ __forceinline__ void cp_async16m(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit_m() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait_m() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

// state held by slot s of a B row in pair order: pairs (0,1) (4,5) (6,7) (8,9) (10,11) (12,13) (14,-); I1, I2 share I0's row
__device__ constexpr int PSTATE[QP] = {0, 1, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, -1, -1, -1};

struct FSmem {
  float Aw[24];          // [0..6] A[src1(j)][j], [7..14] A[pred p][p], [15..22] A[p][succ p]
  float W[24];           // the 23 edge weights in the reference's kernel order
  float EF[G][QP];       // emission at every chunk's first transition step
  float EL[G][QP];       // emission at every chunk's last position
  int hef[HEAP], her[HEAP];
  unsigned long long bar;   // mbarrier: the codon tables have landed
  float ll;
  float pad_[1];
  int edge_hi[4];        // codes of positions T, T + 1 (the next shard's first nucleotides, 4 outside the sequence); 2 pads
  union {                // phase 1: last pass-state emissions of every run; phases 4-5: boundary vectors of the tree nodes
    float carry[FT / 32][8];
    struct { float hvf[HEAP][QP]; float hvr[HEAP][QP]; } v;
  };
  union {                // phase 1: B (as state pairs per class, in sp.B) and the codon tables; 4: the sequence's tile matrices;
    EmisSmem sp;         // 5: emission ring (16-byte piece c of lane l's row at piece c ^ ((l >> 1) & 3)); 6: output staging
    float tiles[FSCAN][MAT];
    float ring[2][FRING][32][QP];
    float stage[FT / 32][32 * Q];
  };
};

// Sequence-shard mode (multi-GPU): the launch carries extra "reducer" CTAs (blockIdx.x >= P) that do what
// shard_product_kernel + shard_scan_kernel do as kernels of their own (gene15.cu) while the tile CTAs wait.  One per group
// of G tiles: the group's product tree, then -- once the ranks' blocks are there -- the fold over ranks and groups and the
// down-sweep of its tree to the boundary vectors of its tiles.  One more per sequence: the product of the group matrices
// = the shard's block, stored into every rank's table over NVLink.  world == 0: not sharded (no reducers, the top-level
// scan of phase 4 instead).
struct FShard {
  int world, rank, PG;
  unsigned tag;                // step tag of the peer tables: the same on every rank (local flags carry Args::tag, per launch)
  float* Gt;                   // [NB][PG][256] group matrices
  unsigned* gflg;              // [NB][PG] group matrix published
  unsigned* bvflg;             // [NB][PG] boundary vectors of the group's tiles published
  const float* my_blocks;      // this rank's block table half of the step: [world][NB][256]
  const unsigned* my_flags;    // this rank's flag table half of the step:  [world][NB]
  Peers pr;                    // the tables of every rank (peer stores)
};

__host__ __device__ inline size_t fused_u_floats(int L) {
  size_t y = (size_t)G * (L * QP + 4), h = (size_t)HEAP * MAT, c = (size_t)G * L * QP;
  size_t m = y > h ? y : h;
  return m > c ? m : c;
}
__host__ __device__ inline size_t fused_smem_bytes(int L) { return ((sizeof(FSmem) + 127) / 128) * 128 + fused_u_floats(L) * sizeof(float); }

// coefficient rows: row i at 16 i floats, its 16-byte piece c at piece c ^ ((i >> 1) & 3): the row-per-lane stores of
// phase 1 are conflict-free at the 64-byte row stride (phase 2 reads a row as a broadcast)
__device__ __forceinline__ void coef_store(float* U, int i, const float (&k)[QP]) {
  float4* b = reinterpret_cast<float4*>(U + (size_t)i * QP);
  const int f = (i >> 1) & 3;
  b[0 ^ f] = make_float4(k[0], k[1], k[2], k[3]);
  b[1 ^ f] = make_float4(k[4], k[5], k[6], k[7]);
  b[2 ^ f] = make_float4(k[8], k[9], k[10], k[11]);
  b[3 ^ f] = make_float4(k[12], k[13], k[14], k[15]);
}
__device__ __forceinline__ float4 lds4f(unsigned addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
// ub: address of the coefficient rows in the shared window (128-byte aligned): one shift, one AND and four LOP3 per row
__device__ __forceinline__ void coef_load(unsigned ub, int i, float (&k)[QP]) {
  const unsigned rp = ub + ((unsigned)i << 6), f4 = ((unsigned)i << 3) & 0x30u;
  const float4 a0 = lds4f(rp | (0x00u ^ f4)), a1 = lds4f(rp | (0x10u ^ f4)), a2 = lds4f(rp | (0x20u ^ f4)), a3 = lds4f(rp | (0x30u ^ f4));
  k[0] = a0.x; k[1] = a0.y; k[2] = a0.z; k[3] = a0.w; k[4] = a1.x; k[5] = a1.y; k[6] = a1.z; k[7] = a1.w;
  k[8] = a2.x; k[9] = a2.y; k[10] = a2.z; k[11] = a2.w; k[12] = a3.x; k[13] = a3.y; k[14] = a3.z; k[15] = a3.w;
}
__device__ __forceinline__ float coef_at(const float* U, int i, int j) {
  return U[(size_t)i * QP + ((((j >> 2) ^ (i >> 1)) & 3) << 2) + (j & 3)];
}

// emission ring of one recursion thread: slot s at s * 32 * QP floats from the thread's base, 16-byte piece q at po[q]
constexpr int FRSL = 32 * QP;
__device__ __forceinline__ void ring_issue(float* ring, int slot, const int (&po)[4], const float* src) {
#pragma unroll
  for (int q = 0; q < 4; ++q) cp_async16m(ring + slot * FRSL + po[q], src + 4 * q);
}
__device__ __forceinline__ void ring_read(const float* ring, int slot, const int (&po)[4], float (&e)[QP]) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const float4 v = *reinterpret_cast<const float4*>(ring + slot * FRSL + po[q]);
    e[4 * q] = v.x; e[4 * q + 1] = v.y; e[4 * q + 2] = v.z; e[4 * q + 3] = v.w;
  }
}

__device__ __forceinline__ float lg2_fast(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }   // x normal

__device__ __forceinline__ u64 sum7p(const u64 (&v)[7]) { return add2(add2(add2(v[0], v[1]), add2(v[2], v[3])), add2(add2(v[4], v[5]), v[6])); }
__device__ __forceinline__ u64 sum15p2(const u64 (&v)[Q]) {
  return add2(add2(add2(add2(v[0], v[1]), add2(v[2], v[3])), add2(add2(v[4], v[5]), add2(v[6], v[7]))),
              add2(add2(add2(v[8], v[9]), add2(v[10], v[11])), add2(add2(v[12], v[13]), v[14])));
}

// codon factors of the 9 constrained states at position t from the nucleotide codes of t-2 .. t+2 (gene_pred_hmm_emitter.py:231-263)
__device__ __forceinline__ void codon_factors_reg(const Args& a, const EmisSmem& sp, int n, int t, int r2, int r1, int c0, int l1, int l2,
                                                  float (&cod)[9]) {
  float fl[9], fr[9];
  if ((c0 | l1 | l2) >= 0) {
    const float* tp = sp.T5[0][(l1 * 5 + l2) * 5 + c0];
    const float4 u0 = reinterpret_cast<const float4*>(tp)[0], u1 = reinterpret_cast<const float4*>(tp)[1];
    fl[0] = u0.x; fl[1] = u0.y; fl[2] = u0.z; fl[3] = u0.w; fl[4] = u1.x; fl[5] = u1.y; fl[6] = u1.z; fl[7] = u1.w; fl[8] = tp[8];
  } else {
    float tmp[12];
    codon_side_soft(sp, 0, nuc_dist(a, n, t), nuc_dist(a, n, t + 1), nuc_dist(a, n, t + 2), tmp);
#pragma unroll
    for (int s = 0; s < 9; ++s) fl[s] = tmp[s];
  }
  if ((c0 | r1 | r2) >= 0) {
    const float* tp = sp.T5[1][(r1 * 5 + r2) * 5 + c0];
    const float4 u0 = reinterpret_cast<const float4*>(tp)[0], u1 = reinterpret_cast<const float4*>(tp)[1];
    fr[0] = u0.x; fr[1] = u0.y; fr[2] = u0.z; fr[3] = u0.w; fr[4] = u1.x; fr[5] = u1.y; fr[6] = u1.z; fr[7] = u1.w; fr[8] = tp[8];
  } else {
    float tmp[12];
    codon_side_soft(sp, 1, nuc_dist(a, n, t), nuc_dist(a, n, t - 1), nuc_dist(a, n, t - 2), tmp);
#pragma unroll
    for (int s = 0; s < 9; ++s) fr[s] = tmp[s];
  }
#pragma unroll
  for (int s = 0; s < 9; ++s) cod[s] = fl[s] * fr[s];
}

// One step of the top-level scan with a whole warp on one direction: both half warps hold the vector (lane hl and lane
// 16 + hl the same value), each takes 8 of the 16 source slots and the halves are added by one exchange (matvec_raw,
// gene15_dev.cuh, is the one-half-warp form).  The exponent taken out comes from lane 0's sum, so that it is the same in every lane.
__device__ __forceinline__ void matvec_split(const float* __restrict__ blk, int hl, int hi, bool bwd, HV& h) {
  const int row = min(hl, Q - 1);
  const float sp = blk[row * QP + 15];
  const int Eb = __float_as_int(blk[MAT - 1]);
  const float u = (hl < Q) ? (bwd ? h.v : h.v * sp) : 0.f;
  float acc0 = 0.f, acc1 = 0.f, s0 = 0.f, s1 = 0.f;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int kq = 8 * hi + k;
    const int kk = bwd ? ((hl + kq) & 15) : kq;
    const float uk = __shfl_sync(0xffffffffu, u, kk, 16);
    const bool ok = kk < Q;                      // slot 15 holds scales / exponent bits, never a matrix entry
    float c = bwd ? blk[row * QP + kk] : blk[kk * QP + hl];
    float d = bwd ? blk[15 * QP + kk] : 1.f;
    c = ok ? c : 0.f;
    d = ok ? d : 0.f;
    if (k & 1) { acc1 = fmaf(uk, c, acc1); s1 = fmaf(uk, d, s1); } else { acc0 = fmaf(uk, c, acc0); s0 = fmaf(uk, d, s0); }
  }
  float acc = acc0 + acc1, ssum = s0 + s1;
  acc += __shfl_xor_sync(0xffffffffu, acc, 16);
  ssum += __shfl_xor_sync(0xffffffffu, ssum, 16);
  ssum = __shfl_sync(0xffffffffu, ssum, 0);
  const int ex = pexp(ssum);
  const float out = (bwd ? sp : 1.f) * acc * pow2neg(ex);
  h.v = (hl < Q) ? out : 0.f;
  h.e += clamp_exp(ex) + Eb;
}

// Down-sweep of a 32-leaf product tree in heap order (gene15.cu, finalize_kernel): from the vectors at the root (forward
// vector entering the tree's span in hvf[1] / hef[1], backward vector behind it in hvr[1] / her[1]) to those of every leaf.
// First half of the CTA's warps forward, second half backward; a half warp handles one node.
__device__ __forceinline__ void down_sweep(FSmem& S, const float* H, long long* dbg, int cta) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, hl = lane & 15, hi = lane >> 4;
  (void)dbg; (void)cta;
  {
    const int wpd = FT >> 6;
    const int dir = warp >= wpd ? 1 : 0;
    const int wd = warp - dir * wpd;
    const int row = min(hl, Q - 1);
    for (int lo = 1; lo < G; lo <<= 1) {
      for (int k0 = lo + 2 * wd; k0 < 2 * lo; k0 += 2 * wpd) {
        const int k = k0 + hi;
        const bool act = k < 2 * lo;
        const int kc = act ? k : k0;
        if (dir == 0) {
          const float v = S.v.hvf[kc][hl];
          const int e = S.hef[kc];
          const float* blk = H + (size_t)(2 * kc) * MAT;
          const float u = (hl < Q) ? v * blk[row * QP + 15] : 0.f;
          const int Eb = __float_as_int(blk[MAT - 1]);
          const float s = half_sum(u);
          float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
          for (int j = 0; j < Q; ++j) {
            const float uj = __shfl_sync(0xffffffffu, u, j, 16);
            const float cc = blk[j * QP + hl];
            if (j & 1) acc1 = fmaf(uj, cc, acc1); else acc0 = fmaf(uj, cc, acc0);
          }
          const int ex = pexp(s);
          if (act) {
            S.v.hvf[2 * k][hl] = v;
            S.v.hvf[2 * k + 1][hl] = (hl < Q) ? (acc0 + acc1) * pow2neg(ex) : 0.f;
            if (hl == 0) { S.hef[2 * k] = e; S.hef[2 * k + 1] = e + clamp_exp(ex) + Eb; }
          }
        } else {
          const float v = S.v.hvr[kc][hl];
          const int e = S.her[kc];
          const float* blk = H + (size_t)(2 * kc + 1) * MAT;
          const float sp = blk[15 + row * QP];
          const int Eb = __float_as_int(blk[MAT - 1]);
          const float u = (hl < Q) ? v : 0.f;
          const float s = half_sum((hl < Q) ? blk[15 * QP + hl] * v : 0.f);
          const float* rp = blk + row * QP;
          float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
          for (int j = 0; j < QP; ++j) {
            const int kk = (hl + j) & 15;
            const float uj = __shfl_sync(0xffffffffu, u, kk, 16);
            const float cc = rp[kk];
            if (j & 1) acc1 = fmaf(uj, cc, acc1); else acc0 = fmaf(uj, cc, acc0);
          }
          const int ex = pexp(s);
          if (act) {
            S.v.hvr[2 * k + 1][hl] = v;
            S.v.hvr[2 * k][hl] = (hl < Q) ? sp * (acc0 + acc1) * pow2neg(ex) : 0.f;
            if (hl == 0) { S.her[2 * k + 1] = e; S.her[2 * k] = e + clamp_exp(ex) + Eb; }
          }
        }
      }
      if (dir == 0) named_barrier_sync(2, wpd * 32); else named_barrier_sync(3, wpd * 32);
#ifdef HMM_FUSED_FINE_MARKS
      if (dbg && tid == 0) dbg[(size_t)(4096 + cta) * 16 + 31 - __clz(lo)] = clock_fenced();
#endif
    }
  }
}

__device__ __forceinline__ void shard_reducer(const Args& a, const FShard& sh);

// SHARD = false: the plain single-launch kernel (no reducer code, no halo look-ups, no shard branch in phase 4 in its image)
template <bool SHARD>
__global__ void __launch_bounds__(FT, 2) fused_kernel(const Args a, const FShard sh) {
  extern __shared__ __align__(128) unsigned char smem[];
  FSmem& S = *reinterpret_cast<FSmem*>(smem);
  float* U = reinterpret_cast<float*>(smem + ((sizeof(FSmem) + 127) / 128) * 128);
  const int L = a.L, T = a.T, CS = L * QP + 4, P = a.P;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, hl = lane & 15, hi = lane >> 4;
  if (SHARD && (int)blockIdx.x >= P) {      // shard mode: a reducer CTA
    shard_reducer(a, sh);
    return;
  }
  const bool seq_start = (a.seq_off == 0);   // position 0 of this launch is the start of the whole sequence
  const int p = blockIdx.x, n = blockIdx.y, m = n / a.batch;
  const int c0 = p * G, t_lo = c0 * L, npos = min(T, t_lo + G * L) - t_lo;
  const int nch = min(G, a.C - c0);
  const size_t tile = (size_t)n * P + p;
  const float* Am = a.A + (size_t)m * Q * Q;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  HMM_MARK(cta, 0);
  HMM_WALL(cta, 14);

  // ---- the first rows of every warp's block are requested before anything else (HBM latency runs under the parameter
  // loads).  A warp owns a block of 32 * run consecutive positions and walks it in `run` steps of 32 consecutive positions
  // (lane = position); a row = 4 pieces of class probabilities (the last holds the A flag) + 1 piece of C,G,T,N flags.
  // Lanes behind the tile's end recompute its last position (no divergence); their stores are switched off. ----
  const float* seq = a.inputs + (size_t)n * T * SIN;
  const int run = a.run, wb = warp * 32 * run;
  const bool wact = wb < npos;
  float4 xc[4], xn[4], f3 = make_float4(0.f, 0.f, 0.f, 0.f), f4 = f3;   // f3 / f4: pieces 3 / 4 of the row two steps ahead
  int code_prev = 4, code_cur = 4, code_next = 4;   // 4: outside the sequence, 1/4 each (kmer.py:27-33)
  // whether tile position i is a row of this sequence whose code a step of this warp needs
  auto row_of = [&](int i) { return reinterpret_cast<const float4*>(seq + (size_t)(t_lo + i) * SIN); };
  // code of a position outside this launch's rows: the neighbour shard's nucleotide (halo) or 4 outside the sequence
  auto edge = [&](int t) { return (SHARD && a.halo != nullptr && t >= -2 && t < T + 2) ? nuc_code(a, n, t) : 4; };
  if (wact) {
    {
      const float4* row = row_of(min(wb + lane, npos - 1));
#pragma unroll
      for (int q = 0; q < 4; ++q) xc[q] = __ldg(row + q);
      f4 = __ldg(row + 4);
    }
    if (t_lo + wb + lane < T) code_cur = -2;                  // marks: compute from xc / f4 once they have landed
    else code_cur = edge(t_lo + wb + lane);
    if (t_lo + wb + 32 + lane >= T) code_next = edge(t_lo + wb + 32 + lane);
    if (t_lo + wb + 32 + lane < T) {                          // step 1 (or, behind the block, the two positions the 3-mer windows reach)
      const float4* row = row_of(wb + 32 + lane);
      xn[3] = __ldg(row + 3);
      f3 = __ldg(row + 4);
      code_next = -2;
    }
    {
      const int t = t_lo + wb - 32 + lane;   // lanes 30, 31: the two positions in front of the block
      if (lane >= 30 && t >= 0) {
        const float* pp = seq + (size_t)t * SIN + 12;
        const float4 u = ldg4(pp), v = ldg4(pp + 4);
        code_prev = code_of(u.w, v.x, v.y, v.z, v.w);
      } else if (lane >= 30) {
        code_prev = edge(t);
      }
    }
  }
  // ---- parameters: the codon tables by one bulk copy (lands during the first class contraction), the rest gathered ----
  {
    if (tid == 0) {
      mbar_init(&S.bar, 1);
      mbar_init_fence();
      mbar_expect_tx(&S.bar, (unsigned)(T5_FLOATS * sizeof(float)));
      bulk_g2s(&S.sp.T5[0][0][0], a.codon5, (unsigned)(T5_FLOATS * sizeof(float)), &S.bar);
    }
    if (tid >= 32 && tid < 34) S.edge_hi[tid - 32] = edge(T + tid - 32);
    if (tid < 23) {
      float aw;
      if (tid < 7) aw = __ldg(Am + SRC1[tid] * Q + tid);
      else if (tid < 15) aw = __ldg(Am + PRED[tid - 7] * Q + tid);
      else aw = __ldg(Am + (tid - 8) * Q + SUCC[tid - 15]);
      S.Aw[tid] = aw;
      S.W[tid] = __ldg(Am + ESRC[tid] * Q + EDST[tid]);
    }
    {
      const int c = tid >> 4, slot = tid & 15;   // FT == Q * QP + 16: threads 0..239 fill B as pairs
      if (c < Q) {
        const int st = PSTATE[slot];
        S.sp.B[c][slot] = st >= 0 ? __ldg(a.Bfull + (size_t)m * Q * Q + st * Q + c) : 0.f;
      }
    }
  }
  __syncthreads();
  HMM_MARK(cta, 1);

  // ---- phase 1: emission + coefficient rows.  Every global and shared access of a step is contiguous across the warp,
  // the neighbours' nucleotide codes and the previous position's pass-state emissions come by shuffles; the class
  // probabilities of the next step and the flags of the step after it are in flight while the current step is computed ----
  float* Eg = a.Ews + tile * (size_t)G * CS;
  {
    float elast[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) elast[q] = 0.f;
    if (wact) {
      const float train = (a.flags & HMM_B200_TRAINING) ? 1e-7f : 0.f;
      const float free_cod = (1.0f / 4096.0f) + train;
      if (code_cur == -2) code_cur = code_of(xc[3].w, f4.x, f4.y, f4.z, f4.w);
      if (code_next == -2) code_next = code_of(xn[3].w, f3.x, f3.y, f3.z, f3.w);
      mbar_wait(&S.bar, 0);   // codon tables
#ifdef HMM_EMIT_UNROLL2
#pragma unroll 2
#else
#pragma unroll 1
#endif
      for (int k = 0; k < run; ++k) {
        const int i = wb + 32 * k + lane, ic = min(i, npos - 1), t = t_lo + ic;
        const bool active = i < npos;
        // in flight during this step: class probabilities of step k + 1, flags of step k + 2
        if (k + 1 < run) {
          const float4* row = row_of(min(i + 32, npos - 1));
#pragma unroll
          for (int q = 0; q < 3; ++q) xn[q] = __ldg(row + q);
        }
        const bool v2 = t_lo + i + 64 < T && (k + 2 < run || (k + 2 == run && lane < 2));
        if (v2) {
          const float4* row = row_of(i + 64);
          f3 = __ldg(row + 3);
          f4 = __ldg(row + 4);
        }
        // nucleotide codes of t-2 .. t+2: lanes l-2 .. l+2 of this step, across the step edges from the previous / next step
        int cm2, cm1, cp1, cp2;
        {
          const int s1 = (lane - 1) & 31, s2 = (lane - 2) & 31, s3 = (lane + 1) & 31, s4 = (lane + 2) & 31;
          const int a1 = __shfl_sync(0xffffffffu, code_cur, s1), b1 = __shfl_sync(0xffffffffu, code_prev, s1);
          const int a2 = __shfl_sync(0xffffffffu, code_cur, s2), b2 = __shfl_sync(0xffffffffu, code_prev, s2);
          const int a3 = __shfl_sync(0xffffffffu, code_cur, s3), b3 = __shfl_sync(0xffffffffu, code_next, s3);
          const int a4 = __shfl_sync(0xffffffffu, code_cur, s4), b4 = __shfl_sync(0xffffffffu, code_next, s4);
          cm1 = lane >= 1 ? a1 : b1;
          cm2 = lane >= 2 ? a2 : b2;
          cp1 = lane <= 30 ? a3 : b3;
          cp2 = lane <= 29 ? a4 : b4;
          if (!active) cm2 = cm1 = cp1 = cp2 = 4;   // lanes behind the tile's end: nobody reads their results
        }
        // class-probability contraction: two states per FFMA2, the class probability as the scalar operand
        float e[Q];
        {
          const float x[Q] = {xc[0].x, xc[0].y, xc[0].z, xc[0].w, xc[1].x, xc[1].y, xc[1].z, xc[1].w,
                              xc[2].x, xc[2].y, xc[2].z, xc[2].w, xc[3].x, xc[3].y, xc[3].z};
          u64 acc[7];
#pragma unroll
          for (int q = 0; q < 7; ++q) acc[q] = 0ull;
#pragma unroll
          for (int c = 0; c < Q; ++c) {
            const float4* bp = reinterpret_cast<const float4*>(S.sp.B[c]);
            const float4 b0 = bp[0], b1 = bp[1], b2 = bp[2], b3 = bp[3];
            const u64 xb = bc2(x[c]);
            acc[0] = fma2(xb, pk2(b0.x, b0.y), acc[0]);
            acc[1] = fma2(xb, pk2(b0.z, b0.w), acc[1]);
            acc[2] = fma2(xb, pk2(b1.x, b1.y), acc[2]);
            acc[3] = fma2(xb, pk2(b1.z, b1.w), acc[3]);
            acc[4] = fma2(xb, pk2(b2.x, b2.y), acc[4]);
            acc[5] = fma2(xb, pk2(b2.z, b2.w), acc[5]);
            acc[6] = fma2(xb, pk2(b3.x, b3.y), acc[6]);
          }
          float d15;
          upk2(acc[0], e[0], e[1]); e[2] = e[1]; e[3] = e[1];
          upk2(acc[1], e[4], e[5]); upk2(acc[2], e[6], e[7]); upk2(acc[3], e[8], e[9]);
          upk2(acc[4], e[10], e[11]); upk2(acc[5], e[12], e[13]); upk2(acc[6], e[14], d15);
        }
        {
          float cod[9];
          codon_factors_reg(a, S.sp, n, t, cm2, cm1, active ? code_cur : 4, cp1, cp2, cod);
#pragma unroll
          for (int q = 0; q < Q; ++q) e[q] *= ((q < 6) ? free_cod : cod[q < 6 ? 0 : q - 6] + train);
        }
        if (a.hints != nullptr) {
          const float* h = a.hints + (size_t)n * 2 * Q;
          if (a.seq_off + t == 0) {
#pragma unroll
            for (int q = 0; q < Q; ++q) e[q] *= __ldg(h + q);
          }
          if (a.seq_off + t == a.seq_total - 1) {
#pragma unroll
            for (int q = 0; q < Q; ++q) e[q] *= __ldg(h + Q + q);
          }
        }
#pragma unroll
        for (int q = 0; q < Q; ++q) e[q] = fmaxf(e[q], kEps);   // max(E, eps), MsaHmmCell.py:87
        // pass-state emissions of position t-1: the lane below, lane 0 from the previous step's last lane
        float ep[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float up = __shfl_up_sync(0xffffffffu, e[7 + q], 1);
          ep[q] = lane >= 1 ? up : elast[q];
        }
#pragma unroll
        for (int q = 0; q < 8; ++q) elast[q] = __shfl_sync(0xffffffffu, e[7 + q], 31);
        if (active) {
          const int ci = divL(a, i), c = c0 + ci;
          store_row16(Eg + (size_t)ci * CS + (i - ci * L) * QP, e, 0.f);
          {
            const int ts = c * L + ((c == 0 && seq_start) ? 1 : 0), t1 = min(T, (c + 1) * L) - 1;
            if (t == ts) store_row16(S.EF[ci], e, 0.f);
            if (t == t1) store_row16(S.EL[ci], e, 0.f);
          }
          float kc[QP];
          kc[15] = 0.f;
          const bool head = (k == 0 && lane == 0);   // first position of the block: its predecessor belongs to another warp
          if (t == 0 && seq_start) {
#pragma unroll
            for (int q = 0; q < Q; ++q) kc[q] = e[q];   // position 0 keeps its emission row (no transition into it)
          } else {
#pragma unroll
            for (int j = 0; j < 7; ++j) kc[j] = e[j] * S.Aw[j];
#pragma unroll
            for (int q = 0; q < 8; ++q) kc[7 + q] = e[SUCC[q]] * S.Aw[15 + q];
            if (!head || i == 0) {
              // (the tile's first position takes its own pass-state emissions: k2 is never read at a chunk's first position)
#pragma unroll
              for (int q = 0; q < 8; ++q) kc[7 + q] *= ((head ? e[7 + q] : ep[q]) * S.Aw[7 + q]);
            }
          }
          coef_store(U, i, kc);   // (a block's first row is completed behind the barrier with the previous warp's last emissions)
        }
        // rotate: step k + 1 becomes current; its flags (requested a step ago) give the code of the step after it
#pragma unroll
        for (int q = 0; q < 4; ++q) xc[q] = xn[q];
        xn[3] = f3;
        code_prev = code_cur;
        code_cur = code_next;
        code_next = v2 ? code_of(f3.w, f4.x, f4.y, f4.z, f4.w) : ((unsigned)(t_lo + i + 64 - T) < 2u ? S.edge_hi[t_lo + i + 64 - T] : 4);
      }
    }
    if (lane == 0) {
      *reinterpret_cast<float4*>(&S.carry[warp][0]) = make_float4(elast[0], elast[1], elast[2], elast[3]);
      *reinterpret_cast<float4*>(&S.carry[warp][4]) = make_float4(elast[4], elast[5], elast[6], elast[7]);
    }
    __syncthreads();
    if (lane == 0 && warp > 0 && wb < npos) {
      const float4 u0 = *reinterpret_cast<const float4*>(&S.carry[warp - 1][0]), u1 = *reinterpret_cast<const float4*>(&S.carry[warp - 1][4]);
      const float ep[8] = {u0.x, u0.y, u0.z, u0.w, u1.x, u1.y, u1.z, u1.w};
      float keep[QP];
      coef_load(smem_u32(U), wb, keep);
#pragma unroll
      for (int q = 0; q < 8; ++q) keep[7 + q] *= (ep[q] * S.Aw[7 + q]);
      coef_store(U, wb, keep);
    }
  }
  __syncthreads();
  HMM_MARK(cta, 2);

  // ---- phase 2: conditional rows, thread = (chunk, row pair), packed halves = rows ra | rb ----
  {
    const int ci = tid >> 3, rp = tid & 7, c = c0 + ci;
    const int ra = 2 * rp, rb = 2 * rp + 1;   // rb == 15: not a matrix row (stays zero)
    const unsigned ub = smem_u32(U);
    u64 vec[Q];
    float smA = 1.f, smB = 1.f;
    int esA = 0, esB = 0;
#pragma unroll
    for (int j = 0; j < Q; ++j) vec[j] = pk2(j == ra ? 1.f : 0.f, j == rb ? 1.f : 0.f);   // chunks beyond the sequence: identity
    if (c < a.C) {
      const bool first = (c == 0 && seq_start);
      const int t0 = c * L, t1 = min(T, t0 + L) - 1, ts = t0 + (first ? 1 : 0);
      const int ib = ci * L;
      float initA = 1.f, initB = 1.f;
      if (first) { initA = coef_at(U, ib, ra); initB = coef_at(U, ib, min(rb, Q - 1)); }   // E_0[row]
      if (rb >= Q) initB = 0.f;
      u64 xm[7], ym[8];
#pragma unroll
      for (int j = 0; j < 7; ++j) xm[j] = pk2(ra == j ? initA : 0.f, rb == j ? initB : 0.f);
#pragma unroll
      for (int q = 0; q < 8; ++q) ym[q] = pk2(ra == 7 + q ? initA : 0.f, rb == 7 + q ? initB : 0.f);
      if (ts > t1) {
#pragma unroll
        for (int j = 0; j < 7; ++j) vec[j] = xm[j];
#pragma unroll
        for (int q = 0; q < 8; ++q) vec[7 + q] = ym[q];
      } else {
        u64 x[7], xp[7];
        {
          float k1[QP];
          coef_load(ub, ib + (ts - t0), k1);
#pragma unroll
          for (int j = 0; j < 7; ++j) x[j] = mul2(bc2(k1[j]), xm[SRC1[j]]);
#pragma unroll
          for (int q = 0; q < 8; ++q) x[SUCC[q]] = fma2(bc2(S.EF[ci][SUCC[q]] * S.Aw[15 + q]), ym[q], x[SUCC[q]]);
#pragma unroll
          for (int j = 0; j < 7; ++j) xp[j] = xm[j];
        }
        int cnt = 0;
#pragma unroll 4
        for (int t = ts + 1; t <= t1; ++t) {
          float k[QP];
          coef_load(ub, ib + (t - t0), k);
          u64 xn[7];
          xn[0] = fma2(bc2(k[7 + 7]), xp[5], mul2(bc2(k[0]), x[0]));
          xn[1] = fma2(bc2(k[7 + 1]), xp[4], mul2(bc2(k[1]), x[1]));
          xn[2] = fma2(bc2(k[7 + 2]), xp[5], mul2(bc2(k[2]), x[2]));
          xn[3] = fma2(bc2(k[7 + 3]), xp[6], mul2(bc2(k[3]), x[3]));
          xn[4] = fma2(bc2(k[7 + 4]), xp[1], mul2(bc2(k[4]), x[6]));
          xn[5] = fma2(bc2(k[7 + 5]), xp[2], fma2(bc2(k[7 + 0]), xp[0], mul2(bc2(k[5]), x[4])));
          xn[6] = fma2(bc2(k[7 + 6]), xp[3], mul2(bc2(k[6]), x[5]));
#pragma unroll
          for (int j = 0; j < 7; ++j) { xp[j] = x[j]; x[j] = xn[j]; }
          if ((++cnt & 3) == 0) {
            float sA, sB;
            upk2(add2(sum7p(x), sum7p(xp)), sA, sB);
            const int exA = pexp(sA), exB = pexp(sB);
            const u64 sc = pk2(pow2neg(exA), pow2neg(exB));
#pragma unroll
            for (int j = 0; j < 7; ++j) { x[j] = mul2(x[j], sc); xp[j] = mul2(xp[j], sc); }
            esA += clamp_exp(exA);
            esB += clamp_exp(exB);
          }
        }
#pragma unroll
        for (int j = 0; j < 7; ++j) vec[j] = x[j];
#pragma unroll
        for (int q = 0; q < 8; ++q) vec[7 + q] = mul2(bc2(S.EL[ci][7 + q] * S.Aw[7 + q]), xp[PRED[q]]);
      }
      float sA, sB;
      upk2(sum15p2(vec), sA, sB);
      smA = fmaxf(sA, 1e-37f);
      smB = fmaxf(sB, 1e-37f);
      const u64 inv = pk2(__frcp_rn(smA), __frcp_rn(smB));
#pragma unroll
      for (int j = 0; j < Q; ++j) vec[j] = mul2(vec[j], inv);
    }
    // relative row scales: common power of two = the largest row exponent of the chunk (8 lanes hold its 15 rows)
    int emax = max(esA + fexp(smA), rb < Q ? esB + fexp(smB) : -100000);
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) emax = max(emax, __shfl_xor_sync(0xffffffffu, emax, o));
    const float spA = smA * pow2neg(emax - esA);
    const float spB = rb < Q ? smB * pow2neg(emax - esB) : 0.f;
    __syncthreads();   // every row is done with the coefficient rows: U becomes the product tree
    HMM_MARK(cta, 3);
    float* b = U + (size_t)(G + ci) * MAT;
    {
      float va[Q], vb[Q];
#pragma unroll
      for (int j = 0; j < Q; ++j) upk2(vec[j], va[j], vb[j]);
      store_row16(b + ra * QP, va, spA);
      if (rb < Q) store_row16(b + rb * QP, vb, spB);
    }
    __syncwarp();
    {
      // row 15: column sums of diag(s') M-hat (columns ra, rb of this thread), exponent in slot 15
      float cs0 = 0.f, cs1 = 0.f;
      const int colb = min(rb, Q - 1);
#pragma unroll
      for (int i = 0; i < Q; ++i) {
        const float si = b[i * QP + 15];
        cs0 = fmaf(si, b[i * QP + ra], cs0);
        cs1 = fmaf(si, b[i * QP + colb], cs1);
      }
      __syncwarp();
      b[15 * QP + ra] = cs0;
      b[15 * QP + rb] = rb < Q ? cs1 : __int_as_float(emax);
    }
  }
  __syncthreads();
  HMM_MARK(cta, 4);

  // ---- phase 3: product tree in heap order (node k = node 2k . node 2k+1, root at 1) ----
  float* H = U;
  for (int np = G / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  HMM_MARK(cta, 5);

  // ---- phase 4: publish the tile matrix, wait for the sequence's other tiles, top-level scan, down-sweep ----
  if (tid < MAT / 4) reinterpret_cast<float4*>(a.Mt + tile * MAT)[tid] = reinterpret_cast<const float4*>(H + MAT)[tid];
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    st_release_u32(a.flg + tile, a.tag);
    if (a.dbg) a.dbg[(size_t)cta * 16 + 10] = gtimer();
  }
  const bool add_ll = (a.flags & HMM_B200_NO_LOGLIK) != 0;
  const bool sharded = SHARD && sh.world > 0;
  if (sharded) {
    // shard mode: the group's reducer CTA publishes this tile's boundary vectors (cross-rank, cross-group, in-group folds)
    if (warp == 1) {
      const unsigned* f = sh.bvflg + (size_t)n * sh.PG + p / G;
      if (lane == 0) while (ld_acquire_u32(f) != a.tag) __nanosleep(64);
      if (a.dbg && lane == 0) a.dbg[(size_t)cta * 16 + 11] = gtimer();
      __syncwarp();
      const float* bv = (hi == 0 ? a.bvf : a.bvr) + tile * QP;
      const float bvv = __ldcg(bv + hl);
      const int bve = __float_as_int(__ldcg(bv + 15));
      if (hi == 0) { S.v.hvf[1][hl] = (hl < Q) ? bvv : 0.f; if (hl == 0) S.hef[1] = bve; }
      else         { S.v.hvr[1][hl] = (hl < Q) ? bvv : 0.f; if (hl == 0) S.her[1] = bve; }
      if (lane == 0) S.ll = __ldcg(a.llws + n);
    }
  }
  if (!sharded && warp == 1) {
    for (int q = lane; q < P; q += 32) {
      const unsigned* f = a.flg + (size_t)n * P + q;
      while (ld_acquire_u32(f) != a.tag) __nanosleep(64);
    }
    if (a.dbg && lane == 0) a.dbg[(size_t)cta * 16 + 11] = gtimer();
  }
  __syncthreads();   // every tile of the sequence has published its matrix
  const bool staged = !sharded && P <= FSCAN;
  if (staged) {
    // the sequence's tile matrices into shared memory with every load in flight at once: a scan step then costs a
    // shared-memory matvec instead of an L2 round trip
    const float4* src = reinterpret_cast<const float4*>(a.Mt + (size_t)n * P * MAT);
    float4* dst = reinterpret_cast<float4*>(&S.tiles[0][0]);
    for (int i = tid; i < P * (MAT / 4); i += FT) dst[i] = __ldcg(src + i);
    __syncthreads();
  }
  if (staged && (warp == 1 || warp == 2)) {
    // top-level scan (TotalProbabilityCell.py:30-49 in scaled probability space) on the staged matrices: warp 1 runs
    // pi . M_0 ... M_{P-1} (snapshot on entering tile p; the end gives the log-likelihood), warp 2 runs M_{P-1} ... M_{p+1} . 1;
    // in either warp the two half warps share every step (8 source states each, matvec_split)
    const bool bwd = warp == 2;
    HV h;
    h.v = (hl < Q) ? (bwd ? 1.f : __ldg(a.pi + (size_t)m * Q + hl)) : 0.f;
    h.e = 0;
    if (!bwd) {
      HV snap = h;
      for (int it = 0; it < P; ++it) {
        if (it == p) snap = h;
        matvec_split(S.tiles[it], hl, hi, false, h);
      }
      const float tot = half_sum(h.v);
      const float llv = (float)((double)h.e * LN2D + (double)logf(tot));
      if (lane == 0 && p == 0 && a.loglik) a.loglik[n] = llv;
      if (hi == 0) { S.v.hvf[1][hl] = (hl < Q) ? snap.v : 0.f; if (hl == 0) S.hef[1] = snap.e; }
      if (lane == 0) S.ll = llv;
    } else {
      for (int it = P - 1; it > p; --it) matvec_split(S.tiles[it], hl, hi, true, h);
      if (hi == 0) { S.v.hvr[1][hl] = (hl < Q) ? h.v : 0.f; if (hl == 0) S.her[1] = h.e; }
    }
  }
  if (!sharded && !staged && warp == 1) {
    // more tiles than the staging area holds: the same scan straight from L2, both directions in one warp
    float llv = 0.f;
    const HV bnd = scan_tiles_global(a.Mt + (size_t)n * P * MAT, P, p, a.pi + (size_t)m * Q, hl, hi, &llv);
    if (lane == 0 && p == 0 && a.loglik) a.loglik[n] = llv;
    if (hi == 0) { S.v.hvf[1][hl] = (hl < Q) ? bnd.v : 0.f; if (hl == 0) S.hef[1] = bnd.e; }
    else         { S.v.hvr[1][hl] = (hl < Q) ? bnd.v : 0.f; if (hl == 0) S.her[1] = bnd.e; }
    if (lane == 0) S.ll = llv;
  }
  __syncthreads();
  HMM_MARK(cta, 6);
  // recursion threads (warp 0 forward, warp 1 backward, lane = chunk): the first emission rows start their way back from the
  // workspace now, under the down-sweep
#ifdef HMM_FUSED_FINE_MARKS
  if (a.dbg && tid == 0) a.dbg[(size_t)(4096 + cta) * 16 + 8] = clock_fenced();
  if (a.dbg && tid == 0) a.dbg[(size_t)(4096 + cta) * 16 + 12] = clock_fenced();
#endif
  const int r_ci = lane, r_c = c0 + r_ci;
  const int r_len = (warp < 2 && r_ci < nch) ? min(L, T - r_c * L) : 0;
  const float* r_Ec = Eg + (size_t)r_ci * CS;
  float* r_ring = &S.ring[warp & 1][0][lane][0];
  const int r_po[4] = {(0 ^ ((lane >> 1) & 3)) << 2, (1 ^ ((lane >> 1) & 3)) << 2, (2 ^ ((lane >> 1) & 3)) << 2, (3 ^ ((lane >> 1) & 3)) << 2};
  if (warp < 2) {
#pragma unroll
    for (int k = 0; k < FRING; ++k) {
      if (k < r_len) ring_issue(r_ring, k, r_po, r_Ec + (size_t)(warp == 0 ? k : r_len - 1 - k) * QP);
      cp_async_commit_m();
    }
  }
#ifdef HMM_FUSED_FINE_MARKS
  if (a.dbg && tid == 0) a.dbg[(size_t)(4096 + cta) * 16 + 9] = clock_fenced();
#endif
  down_sweep(S, H, a.dbg, cta);
  __syncthreads();   // boundary vectors complete, tree dead: U becomes Y
  HMM_MARK(cta, 7);

  // ---- phase 5: exact recursions, warp 0 forward, warp 1 backward, lane = chunk; emission rows through the ring ----
  float* Y = U;
  if (warp < 2) {
    const int ci = r_ci, c = r_c;
    const int len = r_len, h = (len + 1) >> 1;
    const float* Ec = r_Ec;
    float* ring = r_ring;
    float w[NEDGE];
    load_edges(S.W, w);
    float* Yc = Y + (size_t)ci * CS;
    if (warp == 0) {
      float av[Q];
      {
        float tv[QP];
        load_row16(S.v.hvf[G + ci], tv);
#pragma unroll
        for (int j = 0; j < Q; ++j) av[j] = tv[j];
      }
      float kap = sum15(av);
      int nsum = 0;
      const bool first = (c == 0 && seq_start);
      int slot = 0;
      for (int ph = 0; ph < 2; ++ph) {
        const int u_lo = ph == 0 ? 0 : h, u_hi = ph == 0 ? h : len;
        for (int u = u_lo; u < u_hi; ++u) {
          float e[QP], R[Q];
          cp_async_wait_m<FRING - 1>();
          ring_read(ring, slot, r_po, e);
          if (u + FRING < len) ring_issue(ring, slot, r_po, Ec + (size_t)(u + FRING) * QP);
          cp_async_commit_m();
          slot = slot + 1 == FRING ? 0 : slot + 1;
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
          if (ph == 0) {
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
      float R[Q];
      {
        float tv[QP];
        load_row16(S.v.hvr[G + ci], tv);
#pragma unroll
        for (int j = 0; j < Q; ++j) R[j] = tv[j];
      }
      float kap = 1.f;
      int nsum = 0;
      int slot = 0;
      for (int ph = 0; ph < 2; ++ph) {
        const int u_hi = ph == 0 ? len : h, u_lo = ph == 0 ? h : 0;
        for (int u = u_hi - 1; u >= u_lo; --u) {
          float e[QP], Rf[Q], bq[Q];
          cp_async_wait_m<FRING - 1>();
          ring_read(ring, slot, r_po, e);
          if (u - FRING >= 0) ring_issue(ring, slot, r_po, Ec + (size_t)(u - FRING) * QP);
          cp_async_commit_m();
          slot = slot + 1 == FRING ? 0 : slot + 1;
          const float fl = kEps * kap;
#pragma unroll
          for (int j = 0; j < Q; ++j) Rf[j] = fmaxf(R[j], fl);
#pragma unroll
          for (int j = 0; j < Q; ++j) bq[j] = e[j] * Rf[j];
          if (ph == 0) {
            store_row16(Yc + u * QP, Rf, __int_as_float(nsum));
          } else {
            float o[QP];
            load_row16(Yc + u * QP, o);
            float gm[Q];
#pragma unroll
            for (int j = 0; j < Q; ++j) gm[j] = Rf[j] * o[j];
            store_row16(Yc + u * QP, gm, 0.f);
          }
          const float g = sum15(bq);
          int ex;
          const float sc = recip_pow2(g, ex);
          nsum += ex;
#pragma unroll
          for (int j = 0; j < Q; ++j) bq[j] *= sc;
          kap = g * sc;
          A_vec(bq, w, R);
        }
        if (ph == 0) named_barrier_sync(1, 64);
      }
    }
    cp_async_wait_m<0>();
  }
  __syncthreads();
  HMM_MARK(cta, 8);

  // ---- phase 6: log posterior = log g - log sum(g) (+ log-likelihood), warp = 32 consecutive positions, staged ----
  {
    const size_t obase = ((size_t)n * T + t_lo) * Q;
    const float ll_shift = add_ll ? S.ll : 0.f;
    float* stg = S.stage[warp];
    constexpr float LN2F = 0.69314718055994531f;
    for (int ib = warp * 32; ib < npos; ib += FT) {
      const int i = ib + lane;
      if (i < npos) {
        float g[QP];
        load_row16(Y + tile_row_offset(a, i, CS), g);
        float z = 0.f;
#pragma unroll
        for (int j = 0; j < Q; ++j) { g[j] = fmaxf(g[j], 1e-37f); z += g[j]; }
        const float lz = lg2_fast(z);   // every g >= 1e-37: normal numbers, no denormal branch needed
#pragma unroll
        for (int j = 0; j < Q; ++j) stg[lane * Q + j] = fmaf(lg2_fast(g[j]) - lz, LN2F, ll_shift);
      }
      __syncwarp();
      const int cnt = min(32, npos - ib) * Q;
      float* dst = a.out_post + obase + (size_t)ib * Q;
      for (int j = lane; j < cnt; j += 32) dst[j] = stg[j];
      __syncwarp();
    }
  }
  HMM_MARK(cta, 9);
  HMM_WALL(cta, 15);
}

// ------------------------------------------------------------------------------------------------------------
// shard mode: the reducer CTAs of sequence n (the work of shard_product_kernel / shard_scan_kernel, gene15.cu, driven by
// tagged flags instead of kernel boundaries).  blockIdx.x in [P, P + PG): the reducer of tile group g;  P + PG: the
// sequence's block reducer (product of the group matrices -> every rank's table).
// ------------------------------------------------------------------------------------------------------------
#define RMARK(k) do { if (a.dbg && tid == 0) a.dbg[(size_t)cta * 16 + (k)] = gtimer(); } while (0)

// Both reducers run their body twice: pass 0 on identity blocks, with no wait and nothing published, while the tiles are
// still at work -- it pulls the instructions of the pass that counts into the SM's instruction caches (a reducer is the
// only CTA of the launch that runs this code on its SM, and it runs it once: cold, the chain tiles -> vectors took about
// twice as long).

// product of the sequence's group matrices = the shard's block, stored with the step tag into the table of every rank
__device__ __forceinline__ void shard_block_reducer(const Args& a, const FShard& sh, float* H) {
  const int PG = sh.PG, NB = gridDim.y, n = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  RMARK(0);
  int W = 1;
  while (W < PG) W <<= 1;
  for (int pass = 0; pass < 2; ++pass) {
    const bool real = pass == 1;
    for (int b = (real ? PG : 0) + (tid >> 5); b < W; b += FT / 32) identity_block(H + (W + b) * MAT, lane);   // leaves [W, 2W), root at 1
    if (real) {
      if (tid < PG) {
        const unsigned* f = sh.gflg + (size_t)n * PG + tid;
        while (ld_acquire_u32(f) != a.tag) __nanosleep(40);
      }
      __syncthreads();
      RMARK(1);
      const float4* src = reinterpret_cast<const float4*>(sh.Gt + (size_t)n * PG * MAT);
      float4* dst = reinterpret_cast<float4*>(H + (size_t)W * MAT);
      for (int i = tid; i < PG * (MAT / 4); i += FT) dst[i] = __ldcg(src + i);
    }
    __syncthreads();
    for (int np = W / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  }
  RMARK(2);
  const float4* root = reinterpret_cast<const float4*>(H + MAT);
  const size_t half = (size_t)(sh.tag & 1u) * sh.world * NB;
  for (int i = tid; i < sh.world * (MAT / 4); i += FT) {
    const int r = i >> 6, j = i & 63;
    reinterpret_cast<float4*>(sh.pr.blocks[r] + (half + (size_t)sh.rank * NB + n) * MAT)[j] = root[j];
  }
  // the barrier puts every store of the CTA in front of the releasing threads' stores (a release is cumulative)
  __syncthreads();
  if (tid < sh.world) st_release_sys_u32(sh.pr.flags[tid] + half + (size_t)sh.rank * NB + n, sh.tag);
  RMARK(3);
}

__device__ __forceinline__ void shard_reducer(const Args& a, const FShard& sh) {
  extern __shared__ __align__(128) unsigned char smem[];
  FSmem& S = *reinterpret_cast<FSmem*>(smem);
  float* H = reinterpret_cast<float*>(smem + ((sizeof(FSmem) + 127) / 128) * 128);   // heap order: leaves [G, 2G), root at 1
  const int P = a.P, PG = sh.PG, NB = gridDim.y;
  const int g = (int)blockIdx.x - P, n = blockIdx.y, m = n / a.batch;
  if (g == PG) { shard_block_reducer(a, sh, H); return; }
  const int tid = threadIdx.x, lane = tid & 31, hl = lane & 15, hi = lane >> 4;
  const int nb = min(G, P - g * G);
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  RMARK(0);
  for (int pass = 0; pass < 2; ++pass) {
    const bool real = pass == 1;
    // 1. the group's tile matrices, as their CTAs publish them
    for (int b = (real ? nb : 0) + (tid >> 5); b < G; b += FT / 32) identity_block(H + (G + b) * MAT, lane);
    if (real) {
      // (a warp takes tiles w, w + 8, ...: one lane polls the tile's flag, the warp requests its matrix -- the copies run
      // while the slower tiles are still awaited)
      const float4* src = reinterpret_cast<const float4*>(a.Mt + ((size_t)n * P + (size_t)g * G) * MAT);
      float4* dst = reinterpret_cast<float4*>(H + G * MAT);
      for (int b = tid >> 5; b < nb; b += FT / 32) {
        if (lane == 0) {
          const unsigned* f = a.flg + (size_t)n * P + (size_t)g * G + b;
          while (ld_acquire_u32(f) != a.tag) __nanosleep(32);
        }
        __syncwarp();
        cp_async16m(dst + b * (MAT / 4) + lane, src + b * (MAT / 4) + lane);
        cp_async16m(dst + b * (MAT / 4) + 32 + lane, src + b * (MAT / 4) + 32 + lane);
      }
      cp_async_commit_m();
      cp_async_wait_m<0>();
      RMARK(1);
    }
    __syncthreads();
    // 2. the group's product tree (kept: the down-sweep of step 5 walks it) -> its root to Gt, tagged
    for (int np = G / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
    int nranks = 1, ngroups = 2;   // pass 0: three identity blocks
    if (real) {
      nranks = sh.world; ngroups = PG;
      if (tid < MAT / 4) reinterpret_cast<float4*>(sh.Gt + ((size_t)n * PG + g) * MAT)[tid] = reinterpret_cast<const float4*>(H + MAT)[tid];
      __syncthreads();
      if (tid == 0) st_release_u32(sh.gflg + (size_t)n * PG + g, a.tag);   // (cumulative over the CTA's stores: the barrier)
      RMARK(2);
      // 3. the other groups' matrices (S.tiles[world + s]) and the ranks' blocks (S.tiles[r]) as they arrive
      if (tid < PG && tid != g) {
        const unsigned* f = sh.gflg + (size_t)n * PG + tid;
        while (ld_acquire_u32(f) != a.tag) __nanosleep(40);
      }
      __syncthreads();
      RMARK(3);
      const float4* gs = reinterpret_cast<const float4*>(sh.Gt + (size_t)n * PG * MAT);
      float4* dst = reinterpret_cast<float4*>(&S.tiles[sh.world][0]);
      for (int i = tid; i < PG * (MAT / 4); i += FT)
        if ((i >> 6) != g) dst[i] = __ldcg(gs + i);
      if (tid < sh.world) {
        const unsigned* f = sh.my_flags + (size_t)tid * NB + n;
        while (ld_acquire_sys_u32(f) != sh.tag) __nanosleep(40);
      }
      __syncthreads();
      RMARK(4);
      for (int i = tid; i < sh.world * (MAT / 4); i += FT) {
        const int r = i >> 6, j = i & 63;
        reinterpret_cast<float4*>(&S.tiles[r][0])[j] = __ldcg(reinterpret_cast<const float4*>(sh.my_blocks + ((size_t)r * NB + n) * MAT) + j);
      }
    } else if (tid < 96) {
      identity_block(&S.tiles[tid >> 5][0], lane);
    }
    __syncthreads();
    // 4. folds (TotalProbabilityCell.py:30-49 across ranks, then across this rank's groups), ONE loop: lanes 0-15 forward,
    // lanes 16-31 backward; both halves run to the longer trip count and commit only their own steps
    if (tid < 32) {
      const int rank = real ? sh.rank : 0, gg = real ? g : 0;
      HV h;
      h.v = (hl < Q) ? (hi == 0 ? __ldg(a.pi + (size_t)m * Q + hl) : 1.f) : 0.f;
      h.e = 0;
      HV mine = h;
      auto ranks_done = [&]() {   // the forward half has crossed every rank: the log-likelihood; back to this rank's entry
        const float tot = half_sum(h.v);
        if (real && lane == 0) {   // every reducer of the sequence stores the same value: its own tiles read it behind its own flag
          const float ll = (float)((double)h.e * LN2D + (double)logf(tot));
          a.llws[n] = ll;
          if (g == 0 && a.loglik) a.loglik[n] = ll;
        }
        if (hi == 0) h = mine;
      };
      const int nf = nranks + ngroups - 1;
      for (int it = 0; it < nf; ++it) {
        int slot;
        bool commit;
        if (it < nranks) {
          if (hi == 0 && it == rank) mine = h;
          slot = hi == 0 ? it : nranks - 1 - it;
          commit = hi == 0 ? true : slot > rank;
        } else {
          if (it == nranks) ranks_done();
          const int s = hi == 0 ? it - nranks : ngroups - 1 - (it - nranks);   // forward: groups 0 .. g-1; backward: PG-1 .. g+1
          commit = hi == 0 ? s < gg : s > gg;
          slot = nranks + s;
        }
        matvec_raw(S.tiles[slot], hl, hi, commit, h);
      }
      if (ngroups == 1) ranks_done();
      if (hi == 0) { S.v.hvf[1][hl] = (hl < Q) ? h.v : 0.f; if (hl == 0) S.hef[1] = h.e; }
      else         { S.v.hvr[1][hl] = (hl < Q) ? h.v : 0.f; if (hl == 0) S.her[1] = h.e; }
    }
    __syncthreads();
    if (real) RMARK(5);
    // 5. down the group's tree to the vectors of its tiles
    down_sweep(S, H, nullptr, 0);
    __syncthreads();
  }
  for (int i = tid; i < nb * QP; i += FT) {
    const int b = i >> 4, c = i & 15;
    const size_t o = ((size_t)n * P + (size_t)g * G + b) * QP + c;
    a.bvf[o] = c < Q ? S.v.hvf[G + b][c] : __int_as_float(S.hef[G + b]);
    a.bvr[o] = c < Q ? S.v.hvr[G + b][c] : __int_as_float(S.her[G + b]);
  }
  __syncthreads();
  if (tid == 0) st_release_u32(sh.bvflg + (size_t)n * PG + g, a.tag);
  RMARK(6);
}
#undef RMARK

// ------------------------------------------------------------------------------------------------------------
// Boundary vectors of every tile for long tile lists (the two-kernel path behind products_kernel: config 5's 1 Mbp has 869
// tiles in 28 groups) in ONE launch, the reducer scheme of the shard mode without ranks: one CTA per group of 32 tile
// matrices builds the group's product tree and publishes its root; every CTA then folds the sequence's group matrices up to
// its own group from either end (two warps, matvec_split) and sweeps its tree down to the tiles' vectors.  Replaces
// group_reduce_kernel + boundary_kernel + group_down_kernel (gene15.cu: 43 us of the 233 at 1 Mbp).  All CTAs of the launch
// are resident at once (cooperative launch).
// ------------------------------------------------------------------------------------------------------------
constexpr int BG_MAX_GROUPS = 96;
__host__ __device__ inline size_t bg_smem_bytes(int PG) {
  return ((sizeof(FSmem) + 127) / 128) * 128 + ((size_t)HEAP + PG) * MAT * sizeof(float);
}
__global__ void __launch_bounds__(FT) boundary_groups_kernel(const Args a, float* __restrict__ Gt, unsigned* __restrict__ gflg,
                                                             unsigned tag) {
  extern __shared__ __align__(128) unsigned char smem[];
  FSmem& S = *reinterpret_cast<FSmem*>(smem);
  float* H = reinterpret_cast<float*>(smem + ((sizeof(FSmem) + 127) / 128) * 128);   // heap order: leaves [G, 2G), root at 1
  float* GM = H + (size_t)HEAP * MAT;                                                 // the sequence's group matrices
  const int g = blockIdx.x, n = blockIdx.y, PG = gridDim.x, P = a.P, m = n / a.batch;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, hl = lane & 15, hi = lane >> 4;
  const int nb = min(G, P - g * G);
  // 1. the group's product tree
  {
    const float4* src = reinterpret_cast<const float4*>(a.Mt + ((size_t)n * P + (size_t)g * G) * MAT);
    float4* dst = reinterpret_cast<float4*>(H + G * MAT);
    for (int i = tid; i < nb * (MAT / 4); i += FT) dst[i] = __ldg(src + i);
    for (int b = nb + warp; b < G; b += FT / 32) identity_block(H + (G + b) * MAT, lane);
  }
  __syncthreads();
  for (int np = G / 2; np >= 1; np >>= 1) tree_level_tc(H + 2 * np * MAT, H + np * MAT, np);
  // 2. publish the group matrix
  if (tid < MAT / 4) reinterpret_cast<float4*>(Gt + ((size_t)n * PG + g) * MAT)[tid] = reinterpret_cast<const float4*>(H + MAT)[tid];
  __syncthreads();
  if (tid == 0) st_release_u32(gflg + (size_t)n * PG + g, tag);
  // 3. the other groups' matrices
  for (int q = tid; q < PG; q += FT) {
    const unsigned* f = gflg + (size_t)n * PG + q;
    while (ld_acquire_u32(f) != tag) __nanosleep(64);
  }
  __syncthreads();
  {
    const float4* src = reinterpret_cast<const float4*>(Gt + (size_t)n * PG * MAT);
    float4* dst = reinterpret_cast<float4*>(GM);
    for (int i = tid; i < PG * (MAT / 4); i += FT) dst[i] = __ldcg(src + i);
  }
  __syncthreads();
  // 4. folds (TotalProbabilityCell.py:30-49 in scaled probability space): warp 1 from the left up to this group (group 0's
  // CTA to the end: the log-likelihood), warp 2 from the right down to the group behind this one
  if (warp == 1) {
    HV h;
    h.v = (hl < Q) ? __ldg(a.pi + (size_t)m * Q + hl) : 0.f;
    h.e = 0;
    const int steps = g == 0 ? PG : g;
    HV snap = h;
    for (int s = 0; s < steps; ++s) {
      if (s == g) snap = h;
      matvec_split(GM + (size_t)s * MAT, hl, hi, false, h);
    }
    if (g != 0) snap = h;
    if (g == 0 && !a.ll_external) {
      const float tot = half_sum(h.v);
      const float ll = (float)((double)h.e * LN2D + (double)logf(tot));
      if (lane == 0) { a.llws[n] = ll; if (a.loglik) a.loglik[n] = ll; }
    }
    if (hi == 0) { S.v.hvf[1][hl] = (hl < Q) ? snap.v : 0.f; if (hl == 0) S.hef[1] = snap.e; }
  } else if (warp == 2) {
    HV h;
    h.v = (hl < Q) ? 1.f : 0.f;
    h.e = 0;
    for (int s = PG - 1; s > g; --s) matvec_split(GM + (size_t)s * MAT, hl, hi, true, h);
    if (hi == 0) { S.v.hvr[1][hl] = (hl < Q) ? h.v : 0.f; if (hl == 0) S.her[1] = h.e; }
  }
  __syncthreads();
  // 5. down the group's tree to the vectors of its tiles
  down_sweep(S, H, nullptr, 0);
  __syncthreads();
  for (int i = tid; i < nb * QP; i += FT) {
    const int b = i >> 4, c = i & 15;
    const size_t o = ((size_t)n * P + (size_t)g * G + b) * QP + c;
    a.bvf[o] = c < Q ? S.v.hvf[G + b][c] : __int_as_float(S.hef[G + b]);
    a.bvr[o] = c < Q ? S.v.hvr[G + b][c] : __int_as_float(S.her[G + b]);
  }
}

// ------------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------------
enum { SLOT_FUSED = 23, SLOT_FUSED_SHARD = 28 };

// CTAs of the fused kernel that can be resident at once on the current device for chunk length L (0: does not fit)
int fused_capacity(int L) {
  static int cache[kMaxDevices][MAX_L + 1] = {};
  if (L < 1 || L > MAX_L) return 0;
  const int dev = current_device();
  {
    std::lock_guard<std::mutex> lk(state_mutex());
    if (cache[dev][L] != 0) return cache[dev][L] < 0 ? 0 : cache[dev][L];
  }
  const size_t smem = fused_smem_bytes(L);
  int cap = -1;
  if (smem <= (size_t)227 * 1024 && ensure_dynamic_smem(fused_kernel<false>, SLOT_FUSED, smem) == 0 &&
      ensure_dynamic_smem(fused_kernel<true>, SLOT_FUSED_SHARD, smem) == 0) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fused_kernel<true>, FT, smem) == cudaSuccess && per_sm > 0)
      cap = per_sm * device_sm_count();
    else
      (void)cudaGetLastError();
  }
  std::lock_guard<std::mutex> lk(state_mutex());
  cache[dev][L] = cap;
  return cap < 0 ? 0 : cap;
}

int fused_run_len(int L) { return cdiv(G * L, FT); }

static unsigned next_tag() {
  static std::atomic<unsigned> tag{(unsigned)std::chrono::steady_clock::now().time_since_epoch().count() | 1u};
  unsigned t = tag.fetch_add(2654435761u) + 2654435761u;
  return t ? t : 1u;
}

// Every CTA of these launches waits inside the kernel for others of the same grid, so the grid must be resident as a whole:
// a cooperative launch makes the driver guarantee it (or fail) also when other streams have kernels on the device (two
// plain launches from two streams could each get half of their CTAs placed and wait for ever).  No measurable cost: 61.7
// vs 61.8 us at config 2.
constexpr int FUSED_NOT_RESIDENT = 1001;   // internal: the cooperative launch was refused for size (not a C ABI code)
static int launch_coresident(const Args& a, const FShard& sh, dim3 grid, size_t smem, cudaStream_t st) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = dim3(FT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;
  at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  const char* refuse = getenv("HMM_B200_TEST_REFUSE_COOP");   // test hook: behave as if the device had refused the grid
  const cudaError_t e = (refuse && refuse[0] == '1') ? cudaErrorCooperativeLaunchTooLarge : (sh.world > 0 ? cudaLaunchKernelEx(&cfg, fused_kernel<true>, a, sh) : cudaLaunchKernelEx(&cfg, fused_kernel<false>, a, sh));
  if (e == cudaErrorCooperativeLaunchTooLarge) {
    // fewer CTAs fit than the occupancy query promised (a reduced SM share under MPS, a smaller shared-memory carve-out):
    // the caller takes the path that needs no co-residency
    (void)cudaGetLastError();
    set_error("genepred: %u x %u CTAs of the single-launch kernel cannot be resident at once on this device", grid.x, grid.y);
    return FUSED_NOT_RESIDENT;
  }
  HMM_CUDA_OK(e);
  HMM_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_fused(Args a, int NB, cudaStream_t st) {
  const size_t smem = fused_smem_bytes(a.L);
  int rc = ensure_dynamic_smem(fused_kernel<false>, SLOT_FUSED, smem);
  if (rc) return rc;
  a.run = fused_run_len(a.L);
  a.tag = next_tag();
  // Under CUDA graph capture the tag is baked into the graph and every replay would find its own flags of the previous
  // replay: a captured launch clears the tile flags first (a memset node in front of the kernel node).
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) { (void)cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
  if (cap == cudaStreamCaptureStatusActive) HMM_CUDA_OK(cudaMemsetAsync(a.flg, 0, (size_t)NB * a.P * sizeof(unsigned), st));
  FShard sh{};
  return launch_coresident(a, sh, dim3(a.P, NB), smem, st);
}

// shard mode: P tile CTAs + PG group reducers + 1 block reducer per sequence, all resident at once.  Local flags (tiles, groups, boundary
// vectors) carry a per-launch tag, the peer tables the step tag (the same on every rank); nothing is cleared between steps.
int launch_fused_shard(Args a, int NB, const unsigned long long* peer_blocks, const unsigned long long* peer_flags, int rank, int world,
                       unsigned tag, float* Gt, unsigned* gflg, unsigned* bvflg, cudaStream_t st) {
  const size_t smem = fused_smem_bytes(a.L);
  int rc = ensure_dynamic_smem(fused_kernel<true>, SLOT_FUSED_SHARD, smem);
  if (rc) return rc;
  a.run = fused_run_len(a.L);
  a.tag = next_tag();   // local flags (tiles, groups, boundary vectors): unique per launch whatever the workspace held before
  FShard sh{};
  sh.world = world; sh.rank = rank; sh.PG = cdiv(a.P, G); sh.tag = tag;
  HMM_REQUIRE(sh.PG <= G && world + sh.PG <= FSCAN, "genepred shard step: %d tile groups among %d ranks exceed the reducer's staging", sh.PG, world);
  sh.Gt = Gt; sh.gflg = gflg; sh.bvflg = bvflg;
  for (int r = 0; r < world; ++r) {
    sh.pr.blocks[r] = reinterpret_cast<float*>(static_cast<uintptr_t>(peer_blocks[r]));
    sh.pr.flags[r] = reinterpret_cast<unsigned*>(static_cast<uintptr_t>(peer_flags[r]));
  }
  sh.pr.rank = rank; sh.pr.world = world; sh.pr.tag = tag;
  const size_t half = (size_t)(tag & 1u) * world * NB;
  sh.my_blocks = sh.pr.blocks[rank] + half * MAT;
  sh.my_flags = sh.pr.flags[rank] + half;
  rc = launch_coresident(a, sh, dim3(a.P + sh.PG + 1, NB), smem, st);
  return rc == FUSED_NOT_RESIDENT ? HMM_B200_EINVAL : rc;   // (the message says why; the 4-launch route needs no co-residency)
}


// boundary vectors of long tile lists in one cooperative launch; FUSED_NOT_RESIDENT: does not apply (caller takes the three
// kernels of gene15.cu)
int launch_boundary_groups(const Args& a, int NB, float* Gt, unsigned* gflg, cudaStream_t st) {
  static const bool off = getenv("HMM_B200_BOUNDARY_GROUPS") && atoi(getenv("HMM_B200_BOUNDARY_GROUPS")) == 0;
  const int PG = cdiv(a.P, G);
  if (off || PG < 2 || PG > BG_MAX_GROUPS) return FUSED_NOT_RESIDENT;
  const size_t smem = bg_smem_bytes(PG);
  if (smem > (size_t)227 * 1024) return FUSED_NOT_RESIDENT;
  int rc = ensure_dynamic_smem(boundary_groups_kernel, 27, smem);
  if (rc) return rc;
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, boundary_groups_kernel, FT, smem) != cudaSuccess) { (void)cudaGetLastError(); return FUSED_NOT_RESIDENT; }
  if ((long long)NB * PG > (long long)per_sm * device_sm_count()) return FUSED_NOT_RESIDENT;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(PG, NB); cfg.blockDim = dim3(FT); cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;
  at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  const unsigned tag = next_tag();
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) { (void)cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
  if (cap == cudaStreamCaptureStatusActive) HMM_CUDA_OK(cudaMemsetAsync(gflg, 0, (size_t)NB * PG * sizeof(unsigned), st));   // (replays reuse the tag)
  const cudaError_t e = cudaLaunchKernelEx(&cfg, boundary_groups_kernel, a, Gt, gflg, tag);
  if (e == cudaErrorCooperativeLaunchTooLarge) { (void)cudaGetLastError(); return FUSED_NOT_RESIDENT; }
  HMM_CUDA_OK(e);
  return 0;
}

}  // namespace gene15
}  // namespace hmm

// tcgen05 (5th-generation tensor core) step for the generic dense-transition HMM path at profile-HMM sizes
// (learnMSA: q = 2L+3 ~ 1000 states, BASELINE config 4).  Same recurrence, same floors and the same deferred
// normalisation as the CUDA-core step in dense.cu (reference MsaHmmCell.py:73-106, Transitioner.py:320-331):
//
//     U = state . A        one 128 x 256 x q GEMM tile per CTA on the tensor cores, fp32 accumulation in TMEM
//     R = max(U / S_in, eps);  x = max(E_t, eps) * R;  partial row sums of x     fused epilogue
//
// fp32 accuracy from bf16 tensor cores: both operands are carried as bf16 pairs (hi = rn(x), lo = rn(x - hi)) and every
// k-block issues three MMAs (hi.hi + hi.lo + lo.hi) into the same accumulator; the dropped lo.lo term is 2^-16 relative.
// The state pair of step t+1 is written by the epilogue of step t, the transition pair once per call.
//
// One persistent launch runs every time step of both directions (forward t = s, backward t = T-1-s), see chain_kernel.
// Warp roles: warp 0 = TMA producer (one lane), warp 1 = TMEM allocation + MMA issue (one lane), warps 2..9 = epilogue
// (TMEM lane quarter = warp % 4, two warps per quarter).  Operand tiles are [rows][64] bf16 = 128-byte rows in the 128-byte-swizzled K-major
// layout, written by TMA and described to the MMA by shared-memory descriptors (SBO = 1024 B, +32 B per UMMA_K = 16).
#include <cuda.h>
#include <cuda_bf16.h>
#include "common.cuh"

namespace hmm {
namespace dense_tc {

constexpr int TBM = 128, TBN = 256, TBK = 64, STAGES = 2, UK = 16;
constexpr int A_TILE = TBM * TBK * 2;                    // 16 KB  (one bf16 tile of the state)
constexpr int B_TILE = TBN * TBK * 2;                    // 32 KB  (one bf16 tile of the transition matrix)
constexpr int STAGE_BYTES = 2 * A_TILE + 2 * B_TILE;     // hi + lo of both operands = 96 KB
constexpr int STG_PITCH = 33;
constexpr int STG_WARP = 32 * STG_PITCH;                 // floats per epilogue warp
constexpr int EPI_WARPS = 8;                             // two per TMEM lane quarter, each owns half of the tile's columns
constexpr int STG_BYTES = EPI_WARPS * STG_WARP * 4;
constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + STG_BYTES + 128;
constexpr int THREADS = 64 + EPI_WARPS * 32;
constexpr uint32_t TMEM_COLS = 512;                      // two 256-column accumulators
constexpr int MAXP = 32;
static_assert(SMEM_BYTES <= 232448, "shared memory budget");

struct Args {
  const float* E;        // (k, B, T, q), or NULL with the one-hot form:
  const unsigned char* sym;   // (k, B, T) symbol of every position
  const float* Bt;            // (k, nsym, q) emission table, Bt[m][x][j] = P(symbol x | state j): E_t = Bt[m][sym_t]
  int nsym;
  uint16_t* state;       // bf16 [dir][pp][hi|lo][k*Bpad][Npad]
  float* store[2];       // per direction (k, B, T, q) or null: a_t (forward) / R_t (backward), un-normalised
  float* part[2][2];     // per direction, ping-pong: (k, B, ntile) partial row sums
  float* hist[2];        // per direction (k, B, T)
  long long* dbg;        // optional phase timestamps of CTA (0,0): [role 0..2][item - DBG_IT0 (8)][8] globaltimer ns
  unsigned* cnt;         // [chain][dir] CTAs of the chain that finished a step (zeroed per call)
  int k, B, T, q, ntile, ntile_n, Bpad, Npad, nkb;   // ntile = partial-sum slots per row = 2 per column tile
  int dir0, ndir, chain0, nrb;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// spin with a bound: a protocol error traps instead of hanging the device
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  for (uint32_t it = 0; it < (1u << 28); ++it) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
  }
  __trap();
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major operand tile in the 128-byte swizzle: 8-row groups 1024 B apart, descriptor version 1 (sm_100), layout type 2
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr) {
  return (uint64_t)((addr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: D = f32, A = B = bf16, both K-major, M = 128, N = 256
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(TBN >> 3) << 17) | ((uint32_t)(TBM >> 4) << 24);

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
        "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
        "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
        "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}

__device__ __forceinline__ uint32_t pack_bf16(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

__device__ __forceinline__ size_t state_off(const Args& a, int dir, int pp, int hl) {
  return (size_t)((dir * 2 + pp) * 2 + hl) * a.k * a.Bpad * (size_t)a.Npad;
}

// ---- cross-CTA step counters (one per chain and direction, counts CTAs that finished a step) -------------------------
__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void wait_count(const unsigned* p, unsigned want) {
  for (uint32_t it = 0; it < (1u << 24); ++it) {
    if (ld_acquire(p) >= want) return;
    __nanosleep(64);
  }
  __trap();
}
constexpr int DBG_IT0 = 200;
__device__ __forceinline__ void dbg_stamp(const Args& a, int role, int it, int slot) {
  if (a.dbg && blockIdx.x == 0 && blockIdx.y == 0 && it >= DBG_IT0 && it < DBG_IT0 + 8) {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    a.dbg[(role * 8 + (it - DBG_IT0)) * 8 + slot] = t;
  }
}
__device__ __forceinline__ void proxy_fence_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }

// Persistent kernel: CTA (j, c) owns column tile j of chain c = (head, block of 128 sequences) for ALL time steps and both
// directions.  Its work items alternate forward step s / backward step s / forward step s+1 ...; item i accumulates into
// TMEM buffer i & 1, so the epilogue of one direction runs under the MMAs of the other.  The only cross-CTA dependency
// is inside a chain: step s+1 of a direction needs the state columns the chain's other column tiles wrote at step s;
// each CTA bumps cnt[chain][dir] when its epilogue is done and the TMA producer waits for ntile * s before it loads the
// state operand (the transition operand does not wait).  All CTAs of a launch are co-resident (cooperative launch).
__global__ void __launch_bounds__(THREADS, 1)
chain_kernel(const __grid_constant__ CUtensorMap tmS, const __grid_constant__ CUtensorMap tmA, const Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* stg_all = reinterpret_cast<float*>(smem + STAGES * STAGE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES + STG_BYTES);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
  const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + STAGES);
  const uint32_t tfull0 = smem_u32(bars + 2 * STAGES), tempty0 = smem_u32(bars + 2 * STAGES + 2);
  const uint32_t stage0 = smem_u32(smem);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int chain = a.chain0 + (int)blockIdx.y;
  const int m = chain / a.nrb, row0 = (chain % a.nrb) * TBM, n0 = blockIdx.x * TBN;
  const int nitems = (a.T - 1) * a.ndir;
  const unsigned* cnt_r = a.cnt + (size_t)chain * 2;

  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(tfull0 + 8 * i, 1); mbar_init(tempty0 + 8 * i, EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int kbg = 0;
      for (int it = 0; it < nitems; ++it) {
        const int s = 1 + it / a.ndir, dir = a.dir0 + it % a.ndir, pp = (s + 1) & 1;
        const int rS_hi = (int)(((size_t)((dir * 2 + pp) * 2 + 0) * a.k + m) * a.Bpad) + row0;
        const int rS_lo = (int)(((size_t)((dir * 2 + pp) * 2 + 1) * a.k + m) * a.Bpad) + row0;
        const int rA_hi = (int)(((size_t)(dir * 2 + 0) * a.k + m) * a.Npad) + n0;
        const int rA_lo = (int)(((size_t)(dir * 2 + 1) * a.k + m) * a.Npad) + n0;
        for (int kb = 0; kb < a.nkb; ++kb, ++kbg) {
          const int st = kbg % STAGES;
          const uint32_t ph = (uint32_t)(kbg / STAGES) & 1u;
          mbar_wait(empty0 + 8 * st, ph ^ 1u);
          const uint32_t fb = full0 + 8 * st, sb = stage0 + st * STAGE_BYTES;
          mbar_expect_tx(fb, STAGE_BYTES);
          tma_load_2d(sb + 2 * A_TILE, &tmA, fb, kb * TBK, rA_hi);
          tma_load_2d(sb + 2 * A_TILE + B_TILE, &tmA, fb, kb * TBK, rA_lo);
          if (kb == 0 && s >= 2) {   // the chain's column tiles must all have written step s-1 of this direction
            dbg_stamp(a, 0, it, 0);
            wait_count(cnt_r + dir, (unsigned)(a.ntile_n * (s - 1)));
            proxy_fence_global();
            dbg_stamp(a, 0, it, 1);
          }
          if (kb == a.nkb - 1) dbg_stamp(a, 0, it, 2);
          tma_load_2d(sb, &tmS, fb, kb * TBK, rS_hi);
          tma_load_2d(sb + A_TILE, &tmS, fb, kb * TBK, rS_lo);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int kbg = 0;
      for (int it = 0; it < nitems; ++it) {
        const int buf = it & 1;
        dbg_stamp(a, 1, it, 0);
        mbar_wait(tempty0 + 8 * buf, (((uint32_t)it >> 1) & 1u) ^ 1u);   // the epilogue drained this accumulator
        tc_fence_after();
        dbg_stamp(a, 1, it, 1);
        const uint32_t acc = tmem + (uint32_t)(buf * TBN);
        for (int kb = 0; kb < a.nkb; ++kb, ++kbg) {
          const int st = kbg % STAGES;
          const uint32_t ph = (uint32_t)(kbg / STAGES) & 1u;
          mbar_wait(full0 + 8 * st, ph);
          tc_fence_after();
          if (kb == 0) dbg_stamp(a, 1, it, 2);
          const uint32_t sb = stage0 + st * STAGE_BYTES;
#pragma unroll
          for (int k4 = 0; k4 < TBK / UK; ++k4) {
            const uint32_t ko = k4 * UK * 2;   // bytes along K inside the 128-byte swizzle atom
            const uint64_t s_hi = smem_desc(sb + ko), s_lo = smem_desc(sb + A_TILE + ko);
            const uint64_t a_hi = smem_desc(sb + 2 * A_TILE + ko), a_lo = smem_desc(sb + 2 * A_TILE + B_TILE + ko);
            tc_mma(acc, s_lo, a_hi, kIdesc, (kb | k4) ? 1u : 0u);
            tc_mma(acc, s_hi, a_lo, kIdesc, 1u);
            tc_mma(acc, s_hi, a_hi, kIdesc, 1u);
          }
          tc_commit(empty0 + 8 * st);
        }
        tc_commit(tfull0 + 8 * buf);
        dbg_stamp(a, 1, it, 3);
      }
    }
  } else {
    // ---- epilogue: thread = row of the tile (TMEM lane quarter = warp % 4), warp group ew / 4 = column half --------
    const int ew = warp - 2, qd = warp & 3, half = ew >> 2;
    constexpr int CH = TBN / 32 / (EPI_WARPS / 4);   // 32-column chunks per warp
    float* stg = stg_all + ew * STG_WARP;
    const int rl = qd * 32 + lane, r = row0 + rl;
    const bool valid = r < a.B;
    const size_t seq = (size_t)m * a.B + (valid ? r : 0);
    const size_t rowpitch = (size_t)a.T * a.q;
    const size_t rbase = ((size_t)m * a.B + row0 + qd * 32) * rowpitch;   // first row of this warp
    const int rows_here = min(32, a.B - (row0 + qd * 32));                // rows of this warp inside the batch (may be <= 0)
    const int cbase = n0 + half * (CH * 32);
    const int slot = blockIdx.x * (EPI_WARPS / 4) + half;                 // partial-sum slot of this warp group

    for (int it = 0; it < nitems; ++it) {
      const int s = 1 + it / a.ndir, dir = a.dir0 + it % a.ndir, pp = (s + 1) & 1, buf = it & 1;
      const int t = dir == 0 ? s : a.T - 1 - s;
      const float* Et = a.E ? a.E + (size_t)t * a.q : nullptr;
      // one-hot form: lane i holds the symbol of the warp's row i at this step; the row's emissions are a row of the
      // (L2-resident) table, fetched coalesced over the columns
      const float* Btm = a.Bt + (size_t)m * a.nsym * a.q;
      int mysym = 0;
      if (!Et && lane < rows_here) mysym = min((int)__ldg(a.sym + ((size_t)m * a.B + row0 + qd * 32 + lane) * a.T + t), a.nsym - 1);
      float* St = a.store[dir] ? a.store[dir] + (size_t)t * a.q : nullptr;
      const bool stamp = (ew == 0 && lane == 0);
      if (stamp) dbg_stamp(a, 2, it, 0);
      float ereg[32];
      {
        const int col = cbase + lane;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const int sy = __shfl_sync(0xffffffffu, mysym, i);
          ereg[i] = (i < rows_here && col < a.q) ? (Et ? __ldg(Et + rbase + (size_t)i * rowpitch + col) : __ldg(Btm + (size_t)sy * a.q + col)) : 0.f;
        }
      }
      if (s >= 2) {
        if (lane == 0) wait_count(cnt_r + dir, (unsigned)(a.ntile_n * (s - 1)));
        __syncwarp();
      }
      float Sin = 0.f;
      if (valid) {
        const float* p = a.part[dir][pp] + seq * a.ntile;
        for (int i = 0; i < a.ntile; ++i) Sin += __ldcg(p + i);
        if (blockIdx.x == 0 && half == 0) a.hist[dir][seq * a.T + t] = Sin;
      }
      const float inv = valid ? 1.0f / Sin : 0.f;
      uint16_t* o_hi = a.state + state_off(a, dir, pp ^ 1, 0) + ((size_t)m * a.Bpad + r) * a.Npad + cbase;
      uint16_t* o_lo = a.state + state_off(a, dir, pp ^ 1, 1) + ((size_t)m * a.Bpad + r) * a.Npad + cbase;
#pragma unroll
      for (int i = 0; i < 32; ++i) stg[i * STG_PITCH + lane] = ereg[i];
      __syncwarp();
      if (stamp) dbg_stamp(a, 2, it, 1);
      mbar_wait(tfull0 + 8 * buf, ((uint32_t)it >> 1) & 1u);
      tc_fence_after();
      if (stamp) dbg_stamp(a, 2, it, 2);

      float rsum = 0.f;
#pragma unroll 1
      for (int c = 0; c < CH; ++c) {
        const int col = cbase + c * 32 + lane;
        if (c + 1 < CH) {   // next chunk's emissions: in flight while this chunk is processed
          const int ncol = col + 32;
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int sy = __shfl_sync(0xffffffffu, mysym, i);
            ereg[i] = (i < rows_here && ncol < a.q) ? (Et ? __ldg(Et + rbase + (size_t)i * rowpitch + ncol) : __ldg(Btm + (size_t)sy * a.q + ncol)) : 0.f;
          }
        }
        uint32_t acc[32];
        tmem_ld32(tmem + ((uint32_t)(qd * 32) << 16) + (uint32_t)(buf * TBN + half * (CH * 32) + c * 32), acc);
        if (c + 1 == CH) {   // accumulator fully read: hand the buffer back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8 * buf) : "memory");
        }
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int j = 0; j < 32; j += 2) {
          float x2[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int cc = cbase + c * 32 + j + u;
            const float e = stg[lane * STG_PITCH + j + u];
            const float R = fmaxf(__uint_as_float(acc[j + u]) * inv, kEps);
            float x = fmaxf(e, kEps) * R;
            x = (valid && cc < a.q) ? x : 0.f;
            stg[lane * STG_PITCH + j + u] = dir ? R : x;
            rsum += x;
            x2[u] = x;
          }
          const __nv_bfloat16 h0 = __float2bfloat16_rn(x2[0]), h1 = __float2bfloat16_rn(x2[1]);
          const __nv_bfloat16 l0 = __float2bfloat16_rn(x2[0] - __bfloat162float(h0));
          const __nv_bfloat16 l1 = __float2bfloat16_rn(x2[1] - __bfloat162float(h1));
          hi[j >> 1] = pack_bf16(h0, h1);
          lo[j >> 1] = pack_bf16(l0, l1);
        }
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          reinterpret_cast<uint4*>(o_hi + c * 32)[v] = make_uint4(hi[4 * v], hi[4 * v + 1], hi[4 * v + 2], hi[4 * v + 3]);
          reinterpret_cast<uint4*>(o_lo + c * 32)[v] = make_uint4(lo[4 * v], lo[4 * v + 1], lo[4 * v + 2], lo[4 * v + 3]);
        }
        __syncwarp();
        if (St && col < a.q) {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (i < rows_here) St[rbase + (size_t)i * rowpitch + col] = stg[i * STG_PITCH + lane];
        }
        __syncwarp();
        if (c + 1 < CH) {
#pragma unroll
          for (int i = 0; i < 32; ++i) stg[i * STG_PITCH + lane] = ereg[i];
          __syncwarp();
        }
      }
      if (stamp) dbg_stamp(a, 2, it, 3);
      if (valid) a.part[dir][pp ^ 1][seq * a.ntile + slot] = rsum;
      // publish: every epilogue thread's global writes, then one counter bump per CTA
      proxy_fence_global();
      asm volatile("bar.sync 1, %0;" ::"n"(EPI_WARPS * 32) : "memory");
      if (threadIdx.x == 64) {
        __threadfence();
        atomicAdd(a.cnt + (size_t)chain * 2 + dir, 1u);
      }
      if (stamp) dbg_stamp(a, 2, it, 4);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS) : "memory");
  }
}

// First step of both directions (no transition, MsaHmmCell.py:77-78): U = pi (forward) / 1 (backward).
// grid (Bpad, k, ndir), one CTA per (padded) row; writes every column of the bf16 pair (zeros in the padding).
__global__ void __launch_bounds__(256) init_kernel(const Args a, const float* __restrict__ pi) {
  __shared__ float red[8];
  const int r = blockIdx.x, m = blockIdx.y, dir = a.dir0 + blockIdx.z;
  const int t = dir == 0 ? 0 : a.T - 1;
  const bool valid = r < a.B;
  const size_t seq = (size_t)m * a.B + (valid ? r : 0);
  const float* Er = a.E ? a.E + (seq * a.T + t) * a.q
                        : a.Bt + ((size_t)m * a.nsym + min((int)a.sym[seq * a.T + t], a.nsym - 1)) * a.q;
  float* Sr = a.store[dir] ? a.store[dir] + (seq * a.T + t) * a.q : nullptr;
  uint16_t* o_hi = a.state + state_off(a, dir, 0, 0) + ((size_t)m * a.Bpad + r) * a.Npad;
  uint16_t* o_lo = a.state + state_off(a, dir, 0, 1) + ((size_t)m * a.Bpad + r) * a.Npad;
  float local = 0.f;
  for (int c = threadIdx.x; c < a.Npad; c += blockDim.x) {
    float x = 0.f;
    if (valid && c < a.q) {
      const float U = dir == 0 ? pi[(size_t)m * a.q + c] : 1.f;
      const float R = fmaxf(U, kEps);
      x = fmaxf(Er[c], kEps) * R;
      if (Sr) Sr[c] = dir ? R : x;
    }
    const __nv_bfloat16 h = __float2bfloat16_rn(x);
    o_hi[c] = __bfloat16_as_ushort(h);
    o_lo[c] = __bfloat16_as_ushort(__float2bfloat16_rn(x - __bfloat162float(h)));
    local += x;
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0 && valid) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    float* p = a.part[dir][0] + seq * a.ntile;   // ping-pong slot 0 = the sums step 1 divides by
    p[0] = s;
    for (int i = 1; i < a.ntile; ++i) p[i] = 0.f;
    a.hist[dir][seq * a.T + t] = 1.f;
  }
}

// Transition operand, once per call: bf16 pair of A^T (forward: B[n][kk] = A[kk][n]) and of A (backward: B[n][kk] = A[n][kk]),
// zero padded to (Npad, Kpad).  out: [dir][hi|lo][k*Npad][Kpad]
__global__ void prep_kernel(const float* __restrict__ A, uint16_t* __restrict__ out, int k, int q, int Npad, int Kpad) {
  const size_t per = (size_t)k * Npad * Kpad;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= per) return;
  const int kk = (int)(i % Kpad);
  const int n = (int)((i / Kpad) % Npad);
  const int m = (int)(i / ((size_t)Kpad * Npad));
  float vf = 0.f, vb = 0.f;
  if (n < q && kk < q) {
    vf = A[((size_t)m * q + kk) * q + n];
    vb = A[((size_t)m * q + n) * q + kk];
  }
  const __nv_bfloat16 hf = __float2bfloat16_rn(vf), hb = __float2bfloat16_rn(vb);
  out[i] = __bfloat16_as_ushort(hf);
  out[per + i] = __bfloat16_as_ushort(__float2bfloat16_rn(vf - __bfloat162float(hf)));
  out[2 * per + i] = __bfloat16_as_ushort(hb);
  out[3 * per + i] = __bfloat16_as_ushort(__float2bfloat16_rn(vb - __bfloat162float(hb)));
}

// ---- host ----------------------------------------------------------------------------------------------------------
int slots(int q);
typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiled encode_fn() {
  static EncodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiled>(p);
  }
  return fn;
}

// rows x cols bf16, row-major, box = box_rows x 64 columns, 128-byte swizzle
static int make_map(CUtensorMap* map, void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiled enc = encode_fn();
  if (!enc) {
    set_error("dense_tc: cuTensorMapEncodeTiled is not available from this driver");
    return HMM_B200_ECUDA;
  }
  const cuuint64_t dims[2] = {cols, rows};
  const cuuint64_t strides[1] = {cols * 2};
  const cuuint32_t box[2] = {(cuuint32_t)TBK, box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult rc = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("dense_tc: cuTensorMapEncodeTiled failed (%d) for a %llu x %llu tensor", (int)rc, (unsigned long long)rows,
              (unsigned long long)cols);
    return HMM_B200_ECUDA;
  }
  return 0;
}

static long long* g_dbg = nullptr;
void set_debug(void* p) { g_dbg = static_cast<long long*>(p); }   // tools/c4_phases.py

struct Dims {
  int Bpad, Npad, Kpad, ntile_n, nrb, nchain;
  size_t state_bytes, aop_bytes, cnt_bytes;
};
static Dims dims(int k, int B, int q) {
  Dims d;
  d.Bpad = cdiv(B, TBM) * TBM;
  d.Npad = cdiv(q, TBN) * TBN;
  d.Kpad = cdiv(q, TBK) * TBK;
  d.ntile_n = d.Npad / TBN;
  d.nrb = d.Bpad / TBM;
  d.nchain = k * d.nrb;
  d.state_bytes = (size_t)8 * k * d.Bpad * d.Npad * 2;
  d.aop_bytes = (size_t)4 * k * d.Npad * d.Kpad * 2;
  d.cnt_bytes = (size_t)d.nchain * 2 * sizeof(unsigned);
  return d;
}

static int sm_count() { return device_sm_count(); }

bool supported(int k, int B, int q) {
  const Dims d = dims(k, B, q);
  return slots(q) <= MAXP && (size_t)8 * k * d.Bpad < 2147483647ull && (size_t)4 * k * d.Npad < 2147483647ull &&
         d.ntile_n <= 128;
}
// partial-sum slots per row: one per epilogue warp group (column half) of every column tile
int slots(int q) { return cdiv(q, TBN) * (EPI_WARPS / 4); }

// extra workspace (beyond the part / hist / cum / ll slots shared with dense.cu): bf16 state pairs, transition pairs, counters
size_t extra_workspace(int k, int B, int q) {
  const Dims d = dims(k, B, q);
  return align_up(d.state_bytes, 1024) + align_up(d.aop_bytes, 1024) + align_up(d.cnt_bytes, 1024) + 1024;
}

// Runs every step of the requested directions.  part[d][p], hist[d]: the slots of dense.cu's layout with slots(q) per row.
int run_steps(const float* E, const unsigned char* sym, const float* Bt, int nsym, const float* A, const float* pi, int k, int B, int T,
              int q, bool need_f, bool need_b, float* store_f, float* store_b, float* part[2][2], float* hist[2], void* extra,
              cudaStream_t st) {
  const Dims d = dims(k, B, q);
  { int rcs = ensure_dynamic_smem(chain_kernel, 9, SMEM_BYTES); if (rcs) return rcs; }   // slot 9: dense_tc chain
  const int sms = sm_count();
  HMM_REQUIRE(sms >= d.ntile_n, "dense_tc: %d column tiles need as many SMs (device has %d)", d.ntile_n, sms);
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(extra) + 1023) & ~(uintptr_t)1023);
  uint16_t* state = reinterpret_cast<uint16_t*>(base);
  uint16_t* aop = reinterpret_cast<uint16_t*>(base + align_up(d.state_bytes, 1024));
  unsigned* cnt = reinterpret_cast<unsigned*>(base + align_up(d.state_bytes, 1024) + align_up(d.aop_bytes, 1024));
  CUtensorMap tmS, tmA;
  int rc = make_map(&tmS, state, (uint64_t)8 * k * d.Bpad, (uint64_t)d.Npad, TBM);
  if (rc) return rc;
  rc = make_map(&tmA, aop, (uint64_t)4 * k * d.Npad, (uint64_t)d.Kpad, TBN);
  if (rc) return rc;

  Args a;
  a.E = E; a.sym = sym; a.Bt = Bt; a.nsym = nsym; a.state = state; a.cnt = cnt; a.dbg = g_dbg;
  a.store[0] = store_f; a.store[1] = store_b;
  for (int dd = 0; dd < 2; ++dd) {
    a.part[dd][0] = part[dd][0]; a.part[dd][1] = part[dd][1];
    a.hist[dd] = hist[dd];
  }
  a.k = k; a.B = B; a.T = T; a.q = q; a.ntile = slots(q); a.ntile_n = d.ntile_n; a.Bpad = d.Bpad; a.Npad = d.Npad;
  a.nkb = d.Kpad / TBK;
  a.dir0 = need_f ? 0 : 1;
  a.ndir = (need_f ? 1 : 0) + (need_b ? 1 : 0);
  a.chain0 = 0; a.nrb = d.nrb;

  const size_t per = (size_t)k * d.Npad * d.Kpad;
  HMM_CUDA_OK(cudaMemsetAsync(cnt, 0, d.cnt_bytes, st));
  prep_kernel<<<(unsigned)((per + 255) / 256), 256, 0, st>>>(A, aop, k, q, d.Npad, d.Kpad);
  init_kernel<<<dim3(d.Bpad, k, a.ndir), 256, 0, st>>>(a, pi);
  HMM_CUDA_OK(cudaGetLastError());
  if (T < 2) return 0;
  // chains are independent: as many per launch as are co-resident (one CTA per SM), the rest in further launches
  const int per_launch = sms / d.ntile_n;
  for (int c0 = 0; c0 < d.nchain; c0 += per_launch) {
    Args b = a;
    b.chain0 = c0;
    const int nc = d.nchain - c0 < per_launch ? d.nchain - c0 : per_launch;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(d.ntile_n, nc, 1);
    cfg.blockDim = dim3(THREADS, 1, 1);
    cfg.dynamicSmemBytes = SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    HMM_CUDA_OK(cudaLaunchKernelEx(&cfg, chain_kernel, tmS, tmA, b));
  }
  return 0;
}

}  // namespace dense_tc
}  // namespace hmm

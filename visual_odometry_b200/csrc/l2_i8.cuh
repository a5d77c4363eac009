// l2_i8.cuh — K6i: SIFT / NORM_L2 matching as an exact INTEGER contraction (tcgen05 kind::i8).
//
// cv2-SIFT descriptors are integers in [0, 255] (SURVEY 8a A3'), i.e. exactly uint8: the dot
// products of d^2(i,j) = |a_i|^2 + |b_j|^2 - 2 a_i.b_j run as u8 x u8 -> s32 MMAs at twice the
// bf16 rate (K = 32 per instruction), exact by construction.  Used when every value of the frame
// set is an integer in [0, 255] and |d|^2 <= 1 000 250 (SIFT: 262 144); anything else keeps the
// bf16 kernel of l2.cu (decided on the device, flags[2]).
//   * a fifth K = 32 step in S8 x S8 (its own instruction descriptor, same accumulator) multiplies
//     a "ones" group of the row operand (127 x 31, 1) with -floor(|b_j|^2 / 2) written as 32 signed
//     bytes, so the accumulator holds  acc = a.b - floor(|b|^2/2);  2 acc - (|b|^2 & 1) orders the
//     columns of a row by distance.  Padded columns get -128 everywhere: below every real value.
//   * epilogue: key = acc * 512 + cc_j with the per-column constant cc_j = 256 (1 - |b_j|^2 & 1) +
//     (255 - j mod 256) (1 KB per column tile, through its own small smem ring): ONE IMAD + half a
//     3-input max per element; largest key = nearest neighbour, lowest column on ties.
//   * the cross-check direction is a second MMA pass with the operands swapped, as in K1t: at
//     640 tensor cycles per 128 x 256 accumulator the epilogue, not the MMA, is what is scarce.
// Kernel skeleton = hamming_tc.cuh (persistent, warp-specialised, two 256-column accumulators that
// double-buffer each other, work item = (pair, direction, 256-row block)); the eight epilogue warps
// work as two independent groups, one per accumulator (as in hamming_f4.cuh).  Measured
// (profiles/r2_summary.md): with the whole epilogue removed the MMA + bulk-copy side alone needs
// 1263 clocks per accumulator (640 of MMA): at 41 KB per 1280 tensor clocks the four smem stages
// do not cover the L2 latency (see profiles/r2_summary.md for the latency model and the variants).
#pragma once
#include "hamming_tc.cuh"

namespace vo {
namespace l8 {

constexpr int ROWS = 256;
constexpr int KG_BYTES = ROWS * 16;                 // 4096
constexpr int KG_DATA = 8;                          // 128 uint8
constexpr int TILE_BYTES = 12 * KG_BYTES + 1024;    // data 8, ones 2, ext 2, + cc[256] int32 = 50176
constexpr int A_BYTES = 10 * KG_BYTES;              // data + ones
constexpr int B_BYTES = 10 * KG_BYTES;              // data + ext
constexpr uint32_t LBO = KG_BYTES, SBO = 128;
constexpr int NSTAGE = 4;
constexpr int NCC = 4;                              // ring of per-tile column constants
constexpr int EPI_WARPS = 8;                     // warp & 3 = TMEM lane quarter, warp >> 2 = accumulator
constexpr int THREADS = (EPI_WARPS + 2) * 32;
constexpr int H_MAX = 127 * 127 * 31 + 126;         // largest floor(|b|^2/2) the ext group can hold
constexpr int KEY_NONE_BELOW = -(H_MAX + 1) * 512;  // keys of padded columns lie below this
constexpr int MININT = (int)0x80000000;

__host__ __device__ constexpr uint32_t idesc_i8(int a_signed, int b_signed, int M, int N) {
  return (2u << 4) | ((uint32_t)a_signed << 7) | ((uint32_t)b_signed << 10) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}

// ------------------------------------------------------------------ prep: fp32 -> uint8 operand tiles
// flags[2] |= 1 when the frame set is not eligible (non-integer, outside [0,255], norm too large)
static __global__ void __launch_bounds__(256)
tile_kernel(const float* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap, int ntile,
            const float* __restrict__ norms /*[F][norm_stride]*/, int norm_stride,
            uint8_t* __restrict__ tiles, int32_t* __restrict__ flags) {
  const int f = blockIdx.y, tl = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = min(nkp[f], n_cap);
  uint8_t* tile = tiles + ((size_t)f * ntile + tl) * TILE_BYTES;
  bool bad = false;
  for (int rl = warp; rl < ROWS; rl += 8) {
    const int r = tl * ROWS + rl;
    const bool ok = r < n;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ok) v = __ldg(reinterpret_cast<const float4*>(desc + ((size_t)f * n_cap + r) * 128) + lane);
    bad |= (truncf(v.x) != v.x) | (truncf(v.y) != v.y) | (truncf(v.z) != v.z) | (truncf(v.w) != v.w) |
           !(v.x >= 0.f && v.x <= 255.f) | !(v.y >= 0.f && v.y <= 255.f) | !(v.z >= 0.f && v.z <= 255.f) |
           !(v.w >= 0.f && v.w <= 255.f);
    const uint32_t pk = ((uint32_t)(int)v.x & 255u) | (((uint32_t)(int)v.y & 255u) << 8) |
                        (((uint32_t)(int)v.z & 255u) << 16) | (((uint32_t)(int)v.w & 255u) << 24);
    const int k = 4 * lane;
    *reinterpret_cast<uint32_t*>(tile + (size_t)(k >> 4) * KG_BYTES + rl * 16 + (k & 15)) = pk;
    if (lane < 4) {
      // lane 0, 1: ones group (kg 8, 9); lane 2, 3: ext group (kg 10, 11)
      uint4 o = make_uint4(0u, 0u, 0u, 0u);
      if (lane == 0) o = make_uint4(0x7f7f7f7fu, 0x7f7f7f7fu, 0x7f7f7f7fu, 0x7f7f7f7fu);
      if (lane == 1) o = make_uint4(0x7f7f7f7fu, 0x7f7f7f7fu, 0x7f7f7f7fu, 0x017f7f7fu);   // slots 16..30 = 127, slot 31 = 1
      if (lane >= 2) {
        uint32_t wds[4];
        if (!ok) {
          wds[0] = wds[1] = wds[2] = wds[3] = 0x80808080u;        // -128 everywhere: never wins
        } else {
          const float nbf = norms[(size_t)f * norm_stride + r];
          const bool fits = nbf <= (float)(2 * H_MAX);
          bad |= !fits;
          const int nb = fits ? (int)nbf : 0;
          const int h = nb >> 1;
          int m = h / 127;                  // spread over slots 0..30, at most 127 each
          const int rem = h - 127 * m;      // slot 31 (times 1)
#pragma unroll
          for (int wd = 0; wd < 4; ++wd) {
            uint32_t x = 0u;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              const int slot = (lane - 2) * 16 + wd * 4 + b;
              int c;
              if (slot == 31) c = rem;
              else { const int left = m - 127 * slot; c = left <= 0 ? 0 : (left > 127 ? 127 : left); }
              x |= ((uint32_t)(-c) & 255u) << (8 * b);
            }
            wds[wd] = x;
          }
        }
        o = make_uint4(wds[0], wds[1], wds[2], wds[3]);
      }
      *reinterpret_cast<uint4*>(tile + (size_t)(8 + lane) * KG_BYTES + rl * 16) = o;
    }
    if (lane == 4) {
      int e = 0;
      if (ok) {
        const float nbf = norms[(size_t)f * norm_stride + r];
        e = (nbf <= (float)(2 * H_MAX)) ? ((int)nbf & 1) : 0;
      }
      reinterpret_cast<int32_t*>(tile + 12 * KG_BYTES)[rl] = 256 * (1 - e) + (255 - rl);
    }
  }
  if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(&flags[2], 1);
}

struct __align__(1024) Smem {
  uint8_t A[A_BYTES];
  uint8_t B[NSTAGE][B_BYTES];
  int32_t cc[NCC][ROWS];
  uint64_t full_a, empty_a;
  uint64_t full_b[NSTAGE], empty_b[NSTAGE];
  uint64_t full_c[NCC], empty_c[NCC];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

struct Top { int k1, k2, t1, t2; };
// fold the two best keys of column tile t (h >= l) into the running state; keys of different tiles
// compare by (acc, parity) = key >> 8 only, and the earlier tile keeps its place on ties
template <bool TOP2>
__device__ __forceinline__ void fold_tile(Top& s, int h, int l, int t) {
  if (!TOP2) {
    const bool p = (h >> 8) > (s.k1 >> 8);
    s.k1 = p ? h : s.k1;
    s.t1 = p ? t : s.t1;
  } else {
    const bool a = (h >> 8) > (s.k1 >> 8);
    const bool b = a ? ((l >> 8) > (s.k1 >> 8)) : ((h >> 8) > (s.k2 >> 8));
    const int n2 = a ? (b ? l : s.k1) : (b ? h : s.k2);
    const int nt2 = a ? (b ? t : s.t1) : (b ? t : s.t2);
    s.k2 = n2; s.t2 = nt2;
    s.k1 = a ? h : s.k1;
    s.t1 = a ? t : s.t1;
  }
}
// 16-byte shared-memory load with an explicit state space; volatile without a memory clobber:
// ordered against the barrier / TMEM asm statements around it, free against the arithmetic
__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr));
  return v;
}
__device__ __forceinline__ int imad_i(int a, int b, int c) {
  int d;
  asm("mad.lo.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// rowbest layout of l2.cu: float4 (d2_1, idx_1 bits, d2_2, idx_2 bits) at
// [((pair * 2 + dir) * n_splits + split) * row_stride + row]
// VARIANT (compile time: a RUN-time switch around the TMEM loads costs 15-25 % by itself, measured):
// 0 = the per-column constants through the generic load the compiler emits;
// 1, 2 = explicit ld.shared for them (the default).
template <bool TOP2, int VARIANT>
__global__ void __launch_bounds__(THREADS, 1)
match_kernel(const uint8_t* __restrict__ tiles, const int32_t* __restrict__ nkp,
             const int32_t* __restrict__ pairs, int n_cap, int ntile_cap, int ndir, long n_items,
             const float* __restrict__ norms, int norm_stride, int n_splits, int row_stride,
             const int32_t* __restrict__ flags, float4* __restrict__ rowbest) {
  if (flags[2] != 0) return;     // not uint8 data: the bf16 kernel of l2.cu runs instead
  extern __shared__ uint8_t smem_raw[];
  Smem& S = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nrbp = ntile_cap;

  if (threadIdx.x == 0) {
    mbar_init(&S.full_a, 1);
    mbar_init(&S.empty_a, 1);
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(&S.full_b[s], 1); mbar_init(&S.empty_b[s], 1); }
    for (int s = 0; s < NCC; ++s) { mbar_init(&S.full_c[s], 1); mbar_init(&S.empty_c[s], EPI_WARPS); }
    for (int s = 0; s < 2; ++s) { mbar_init(&S.tmem_full[s], 1); mbar_init(&S.tmem_empty[s], EPI_WARPS / 2); }
    fence_barrier_init();
  }
  if (warp == EPI_WARPS) tc::tmem_alloc(&S.tmem_base, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = S.tmem_base;

  if (warp == EPI_WARPS + 1) {
    // ===== bulk-copy producer =====
    if (lane == 0) {
      uint32_t g = 0, n_it = 0;
      for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const htc::Item it = htc::get_item(item, nkp, pairs, n_cap, nrbp, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.empty_a, (n_it & 1) ^ 1);
        mbar_expect_tx(&S.full_a, A_BYTES);
        bulk_g2s(S.A, tiles + ((size_t)it.fq * ntile_cap + it.rbp) * TILE_BYTES, A_BYTES, &S.full_a);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const uint8_t* src = tiles + ((size_t)it.ft * ntile_cap + t) * TILE_BYTES;
          const int sc = g % NCC;
          mbar_wait(&S.empty_c[sc], ((g / NCC) & 1) ^ 1);
          mbar_expect_tx(&S.full_c[sc], ROWS * 4);
          bulk_g2s(S.cc[sc], src + 12 * KG_BYTES, ROWS * 4, &S.full_c[sc]);
          const int st = g % NSTAGE;
          mbar_wait(&S.empty_b[st], ((g / NSTAGE) & 1) ^ 1);
          mbar_expect_tx(&S.full_b[st], B_BYTES);
          bulk_g2s(S.B[st], src, KG_DATA * KG_BYTES, &S.full_b[st]);                                   // uint8 data
          bulk_g2s(S.B[st] + KG_DATA * KG_BYTES, src + 10 * KG_BYTES, 2 * KG_BYTES, &S.full_b[st]);    // ext group
        }
        ++n_it;
      }
    }
  } else if (warp == EPI_WARPS) {
    // ===== MMA issuer: 4 x (u8 x u8) + 1 x (s8 x s8) of 128 x 256 x 32 per row block and tile =====
    if (lane == 0) {
      const uint32_t id_data = idesc_i8(0, 0, 128, ROWS), id_ext = idesc_i8(1, 1, 128, ROWS);
      uint32_t g = 0, n_it = 0, use[2] = {0u, 0u};
      for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const htc::Item it = htc::get_item(item, nkp, pairs, n_cap, nrbp, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.full_a, n_it & 1);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % NSTAGE;
          mbar_wait(&S.full_b[st], (g / NSTAGE) & 1);
          const uint32_t b_addr = smem_u32(S.B[st]);
#pragma unroll
          for (int rb = 0; rb < 2; ++rb) {
            if (rb >= it.nrb) break;
            mbar_wait(&S.tmem_empty[rb], (use[rb] & 1) ^ 1);
            tc::fence_after_sync();
            const uint32_t a_addr = smem_u32(S.A) + rb * 2048;
            const uint32_t d_addr = tmem + rb * ROWS;
#pragma unroll
            for (int k = 0; k < 5; ++k)
              htc::mma_i8(d_addr, tc::smem_desc(a_addr + k * 2 * LBO, LBO, SBO),
                          tc::smem_desc(b_addr + k * 2 * LBO, LBO, SBO), k < 4 ? id_data : id_ext, k > 0);
            tc::commit(&S.tmem_full[rb]);
            ++use[rb];
          }
          tc::commit(&S.empty_b[st]);
        }
        tc::commit(&S.empty_a);
        ++n_it;
      }
    }
  } else {
    // ===== epilogue warps: thread = TMEM lane = row; warps 0..3 drain accumulator 0 (row block 0),
    // warps 4..7 accumulator 1, independently: the barrier / TMEM-load / ranking chain of one
    // accumulator overlaps the other's, and nothing is merged through shared memory =====
    const int quarter = warp & 3, rb = warp >> 2;
    const int rl = quarter * 32 + lane;
    const int mul = 512 + (int)(n_items < 0);      // = 512, opaque: keeps the IMAD form
    uint32_t use = 0, g = 0;
    for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
      const htc::Item it = htc::get_item(item, nkp, pairs, n_cap, nrbp, ndir);
      if (it.ntile == 0) continue;
      Top st;
      st.k1 = st.k2 = MININT; st.t1 = st.t2 = -1;
      for (int t = 0; t < it.ntile; ++t, ++g) {
        const int sc = g % NCC;
        mbar_wait(&S.full_c[sc], (g / NCC) & 1);
        if (rb < it.nrb) {
          const uint4* ccp = reinterpret_cast<const uint4*>(&S.cc[sc][0]);
          const uint32_t ccs = smem_u32(&S.cc[sc][0]);
          mbar_wait(&S.tmem_full[rb], use & 1);
          ++use;
          tc::fence_after_sync();
          const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + rb * ROWS;
          uint32_t v[2][32];
          int h = MININT, l = MININT;
          auto rank = [&](uint32_t (&x)[32], int chunk) {
#pragma unroll
            for (int c4 = 0; c4 < 8; ++c4) {
              // broadcast shared-memory read of four column constants
              const uint4 cc = VARIANT >= 1 ? lds128(ccs + (chunk * 8 + c4) * 16) : ccp[chunk * 8 + c4];
              x[4 * c4 + 0] = (uint32_t)imad_i((int)x[4 * c4 + 0], mul, (int)cc.x);
              x[4 * c4 + 1] = (uint32_t)imad_i((int)x[4 * c4 + 1], mul, (int)cc.y);
              x[4 * c4 + 2] = (uint32_t)imad_i((int)x[4 * c4 + 2], mul, (int)cc.z);
              x[4 * c4 + 3] = (uint32_t)imad_i((int)x[4 * c4 + 3], mul, (int)cc.w);
            }
            if (TOP2) {
              int ch, cl;
              htc::top2_32(x, ch, cl);
              htc::merge2(h, l, ch, cl);
            } else {
              h = max(h, htc::max32(x));
            }
          };
          tc::ld_32x32_issue(taddr, v[0]);
          tc::ld_wait(v[0]);
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            if (c + 1 < 8) tc::ld_32x32_issue(taddr + 32 * (c + 1), v[(c + 1) & 1]);
            if (c == 7) {             // every load has been waited for: the accumulator is in registers
              tc::fence_before_sync();
              __syncwarp();
              if (lane == 0) tc::mbar_arrive(&S.tmem_empty[rb]);
            }
            rank(v[c & 1], c);
            if (c + 1 < 8) tc::ld_wait(v[(c + 1) & 1]);
          }
          fold_tile<TOP2>(st, h, l, t);
        }
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&S.empty_c[sc]);     // this warp is done with the tile's constants
      }
      // ---- convert to squared distances, write ----
      if (rb < it.nrb) {
        Top m = st;
        if (m.t1 < 0 && m.t2 >= 0) { m.k1 = m.k2; m.t1 = m.t2; m.t2 = -1; }
        const int row = it.rbp * ROWS + rb * 128 + rl;
        if (row < it.nq) {
          const int na = (int)norms[(size_t)it.fq * norm_stride + row];
          const float INF = __int_as_float(0x7f800000);
          auto decode = [&](int key, int tile, float& d2, int& col) {
            col = -1; d2 = INF;
            if (tile < 0 || key < KEY_NONE_BELOW) return;
            col = tile * ROWS + 255 - (key & 255);
            if (col >= it.nt) { col = -1; return; }
            const int acc = key >> 9, e = 1 - ((key >> 8) & 1);
            d2 = (float)max(na - 2 * acc + e, 0);
          };
          float4 r;
          int c1, c2 = -1;
          float d1, d2v = INF;
          decode(m.k1, m.t1, d1, c1);
          if (TOP2) decode(m.k2, m.t2, d2v, c2);
          r.x = d1; r.y = __int_as_float(c1); r.z = d2v; r.w = __int_as_float(c2);
          float4* dst = rowbest + (((size_t)it.pair * 2 + it.dir) * n_splits) * row_stride + row;
          dst[0] = r;
          for (int s = 1; s < n_splits; ++s)      // the merge in l2_compact_kernel reads every split
            dst[(size_t)s * row_stride] = make_float4(INF, __int_as_float(-1), INF, __int_as_float(-1));
        }
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == EPI_WARPS) tc::tmem_dealloc(tmem, 512);
}

}  // namespace l8
}  // namespace vo

// l2.cu — K6: SIFT / NORM_L2 descriptor matching on the tcgen05 tensor cores.
//
// Replaces cv2.BFMatcher(NORM_L2, crossCheck=True).match (reference src/visual_slam.py:19,
// src/image_pair.py:234) and knn(k=2)+ratio (src/feature_detection.py:20-26).
//   d^2(i,j) = |a_i|^2 + |b_j|^2 - 2 a_i.b_j : the a.b term is the one dense contraction of the
// hot path, so it runs as a bf16 GEMM (tcgen05.mma, fp32 accumulate in TMEM).  cv2-SIFT
// descriptors are integers in [0,255]: exact in bf16, dot products < 2^24 exact in fp32, so the
// GEMM is EXACT for real SIFT data; other float descriptors get an fp32 rerank of the candidates.
//
// The GEMM has only K = 128, so one 128x128 accumulator tile is 576 tensor-pipe cycles of MMA
// against 16384 values to rank: the epilogue, not the MMA, is the scarce resource.  Design:
//   * the accumulator already holds an ORDERED INTEGER: a ninth K-step adds
//       -s^2|b_j|^2/2 + MAGIC   (MAGIC = 1.5*2^22, s = a power-of-two scale)
//     so the fp32 bit pattern of every element is  const + 2*s^2*(a.b - |b|^2/2): ranking is
//     integer max on the raw bits, the column index rides in 7 low bits of a packed key
//     (one IMAD on the FMA pipe + half a 3-input VIMNMX3 per element, no branches, no
//     divergence; ties resolve to the lowest column exactly like OpenCV).
//   * persistent CTAs (one per SM) walk a list of work items (pair, direction, 256-row block,
//     column split) so TMEM allocation, barrier set-up and pipeline fill are paid once per SM.
//   * every 40 KB B tile streamed from L2 feeds TWO 128-row A blocks (18 MMAs), halving the
//     L2 -> SM traffic that capped the one-block version at ~60 % of the tensor rate.
//
// Pipeline per launch:
//   l2_norm_kernel   exact fp32 squared norms + the largest norm (fixes the scale s).
//   l2_tile_kernel   fp32 [F,n,128] -> bf16 tiles already in the UMMA canonical K-major
//                    no-swizzle layout (one 128x160 tile = one contiguous 40 KB block = ONE 1-D
//                    TMA bulk copy, no tensor map).
//   l2_top2_tc_kernel  warp-specialised: warp 17 = TMA producer, warp 16 = single-thread MMA
//                    issuer, warps 0-15 = epilogue (thread = TMEM lane = query row).
//   l2_rerank_kernel / l2_compact_kernel  fp32 rerank when the data is not exactly
//                    representable; cross-check / ratio test, sqrtf, stable compaction, gather.
#include <cuda_bf16.h>

#include "tc.cuh"
#include "l2_i8.cuh"

namespace vo {

constexpr int L2_DIM = 128;
constexpr int L2_BLK = 128;                       // rows per tile (UMMA M and N)
// a tile = 20 K-groups (kg) of 8 bf16 columns x 128 rows: kg 0..15 the descriptor, kg 16 = [1,1,1,0..]
// (A-side extras), kg 17 = 0, kg 18 = (-s^2|b|^2/2 + MAGIC) split into three bf16 (B-side
// extras), kg 19 = 0.  A ninth K=16 MMA step over (kg16,kg17) x (kg18,kg19) adds that term.
constexpr int L2_KG = 20;
constexpr int L2_TILE_ELEMS = L2_BLK * 8 * L2_KG;  // 20480 bf16 = 40 KB
constexpr int L2_TILE_BYTES = L2_TILE_ELEMS * 2;
constexpr uint32_t L2_LBO = 2048;                 // between the two K core matrices of a K=16 slice
constexpr uint32_t L2_SBO = 128;                  // between 8-row groups
constexpr int L2_NSTAGE = 3;
constexpr int L2_RBK = 2;                         // 128-row A blocks per work item
constexpr int L2_EPI_WARPS = 16;  // 4 per TMEM lane quarter, 32 columns each
constexpr int L2_THREADS = (L2_EPI_WARPS + 2) * 32;  // + MMA warp + TMA warp
// fp32(MAGIC + x) has bit pattern bits(MAGIC) + 2x for every multiple x of 0.5 with |x| < 2^21
#define L2_MAGIC 6291456.0f
constexpr uint32_t L2_MAGIC_BITS = 0x4AC00000u;
constexpr uint32_t L2_QBIAS = L2_MAGIC_BITS & 0x1ffffffu;   // what (key >> 7) reads for x = 0
constexpr float L2_PAD_SCORE = 4194304.0f;        // MAGIC - 2^21: below every real score
// column keys (single-pass mutual mode): ((bits - s^2|a_i|^2 - MAGIC_BITS + 2^23) << 8) | (255 - r)
// = (2^23 - s^2 d^2(i,j)) << 8 | (255 - row inside the 256-row block): unsigned max over a column
// = nearest query row, lowest row on ties.  Padded rows carry |a|^2 := 2^22 + 2^21 and always lose.
constexpr uint32_t L2_COL_OFFSET = L2_MAGIC_BITS - (1u << 23);
constexpr uint32_t L2_PAD_NORM = (1u << 22) + (1u << 21);
enum { L2_MODE_TOP1 = 0, L2_MODE_TOP2 = 1, L2_MODE_TOP1_COLMAX = 2, L2_MODE_TOP1_COLMAX_FRAG = 3 };

struct L2Ws {
  __nv_bfloat16* tiles;   // [F][nblk][20480]
  float* norms;           // [F][nblk*128]   (+inf on padded rows)
  int32_t* flags;         // [0] inexact data, [1] bits of the largest squared norm, [2] not uint8 data
  uint8_t* tiles8;        // [F][ntile256][l8::TILE_BYTES] uint8 operand tiles of the integer path (l2_i8.cuh)
  float4* rowbest;        // [P][2][S][nblk*128]  (d2_1, idx_1, d2_2, idx_2) idx as int bits
  uint32_t* colpart;      // [P][nrbp*4][nblk*128] best query row of every train column among the 64
                          // rows one epilogue warp owns (256-row block x TMEM lane quarter)
  size_t bytes;
};

static inline size_t l2_align(size_t x) { return (x + 1023) & ~(size_t)1023; }

static L2Ws l2_carve(void* base, int n_frames, int n_pairs, int n_cap, int splits) {
  L2Ws w;
  const size_t nblk = (n_cap + L2_BLK - 1) / L2_BLK;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += l2_align(bytes);
    return p;
  };
  w.tiles = (__nv_bfloat16*)take((size_t)n_frames * nblk * L2_TILE_BYTES);
  w.norms = (float*)take((size_t)n_frames * nblk * L2_BLK * sizeof(float));
  w.flags = (int32_t*)take(256);
  w.rowbest = (float4*)take((size_t)n_pairs * 2 * splits * nblk * L2_BLK * sizeof(float4));
  w.colpart = (uint32_t*)take((size_t)n_pairs * ((nblk + 1) / 2) * 4 * nblk * L2_BLK * sizeof(uint32_t));
  w.tiles8 = (uint8_t*)take((size_t)n_frames * ((n_cap + l8::ROWS - 1) / l8::ROWS) * l8::TILE_BYTES);
  w.bytes = off;
  return w;
}

// bit 0: producer loads each stage only once (no per-tile TMA); bit 1: epilogue skips the scan;
// bit 2: epilogue skips the TMEM load too (diagnostics only: results are wrong when set);
// bit 3: MAGIC is added by the epilogue (FADD) instead of inside the ninth MMA step.
__device__ int g_l2_debug = 0;

// scale s = 2^k with s^2 * max|d|^2 <= 2^20: every score  s^2 (a.b - |b|^2/2)  then lies inside
// the +-2^21 window around MAGIC, and is a multiple of 0.5 (= exact) for integer descriptors.
__device__ __forceinline__ float l2_scale(const int32_t* flags) {
  const float maxn2 = __int_as_float(flags[1]);
  if (!(maxn2 > 0.f) || !(maxn2 < 3.0e38f)) return 1.f;
  int e;
  frexpf(maxn2, &e);             // maxn2 = m * 2^e, m in [0.5, 1)  =>  maxn2 <= 2^e
  const int k2 = 20 - e;         // s^2 <= 2^k2
  const int k = (k2 >= 0) ? (k2 >> 1) : -((1 - k2) >> 1);
  return ldexpf(1.f, k);
}

// ------------------------------------------------------------------ prep 1: norms + largest norm
__global__ void __launch_bounds__(256)
l2_norm_kernel(const float* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
               int nblk, L2Ws w) {
  const int f = blockIdx.y, blk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = nkp[f];
  float mx = 0.f;
  for (int rl = warp; rl < L2_BLK; rl += 8) {
    const int r = blk * L2_BLK + rl;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < n) v = __ldg(reinterpret_cast<const float4*>(desc + ((size_t)f * n_cap + r) * L2_DIM) + lane);
    float ss = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0)
      w.norms[((size_t)f * nblk + blk) * L2_BLK + rl] = (r < n) ? ss : __int_as_float(0x7f800000);
    if (r < n) mx = fmaxf(mx, ss);
  }
  // non-negative floats order like their bit patterns
  if (lane == 0 && mx > 0.f) atomicMax(&w.flags[1], __float_as_int(mx));
}

// ------------------------------------------------------------------ prep 2: fp32 -> bf16 tiles
// tile layout (bytes): kg*2048 + rg*128 + rr*16 + kk*2  with r = 8*rg+rr, k = 8*kg+kk
__global__ void __launch_bounds__(256)
l2_tile_kernel(const float* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
               int nblk, int skip_if_i8, L2Ws w) {
  if (skip_if_i8 && w.flags[2] == 0) return;     // uint8 data: the integer kernel (l2_i8.cuh) runs
  const int f = blockIdx.y, blk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = nkp[f];
  const float sc = l2_scale(w.flags);
  const bool fold_magic = (g_l2_debug & 8) == 0;
  __nv_bfloat16* tile = w.tiles + ((size_t)f * nblk + blk) * L2_TILE_ELEMS;
  bool bad = false;
  for (int rl = warp; rl < L2_BLK; rl += 8) {
    const int r = blk * L2_BLK + rl;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < n) v = __ldg(reinterpret_cast<const float4*>(desc + ((size_t)f * n_cap + r) * L2_DIM) + lane);
    v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;      // power of two: exact
    const __nv_bfloat16 b0 = __float2bfloat16_rn(v.x), b1 = __float2bfloat16_rn(v.y),
                        b2 = __float2bfloat16_rn(v.z), b3 = __float2bfloat16_rn(v.w);
    // exact path needs: representable in bf16 AND integer after scaling (scores multiples of 1)
    bad |= (__bfloat162float(b0) != v.x) | (__bfloat162float(b1) != v.y) |
           (__bfloat162float(b2) != v.z) | (__bfloat162float(b3) != v.w) |
           (truncf(v.x) != v.x) | (truncf(v.y) != v.y) | (truncf(v.z) != v.z) | (truncf(v.w) != v.w);
    const int k = 4 * lane;
    const size_t byte_off = (size_t)(k >> 3) * 2048 + (size_t)(rl >> 3) * 128 + (rl & 7) * 16 + (k & 7) * 2;
    __nv_bfloat162 lo, hi;
    lo.x = b0; lo.y = b1; hi.x = b2; hi.y = b3;
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(reinterpret_cast<char*>(tile) + byte_off) = pk;
    if (lane < 4) {  // the four extra K-groups of this row (16 bytes each)
      uint4 e = make_uint4(0, 0, 0, 0);
      if (lane == 0) {
        e.x = 0x3f803f80u;  // bf16 (1, 1)
        e.y = 0x00003f80u;  // bf16 (1, 0)
      } else if (lane == 2) {
        const float base = fold_magic ? L2_MAGIC : 0.f;
        float x = fold_magic ? L2_PAD_SCORE : -2097152.0f;   // padded column never wins
        if (r < n) {
          const float ss = w.norms[((size_t)f * nblk + blk) * L2_BLK + rl];
          x = fmaf(-0.5f * sc * sc, ss, base);
        }
        const __nv_bfloat16 h = __float2bfloat16_rn(x);
        const float r1 = x - __bfloat162float(h);
        const __nv_bfloat16 m = __float2bfloat16_rn(r1);
        const __nv_bfloat16 l = __float2bfloat16_rn(r1 - __bfloat162float(m));
        e.x = (uint32_t)__bfloat16_as_ushort(h) | ((uint32_t)__bfloat16_as_ushort(m) << 16);
        e.y = (uint32_t)__bfloat16_as_ushort(l);
      }
      *reinterpret_cast<uint4*>(reinterpret_cast<char*>(tile) + (size_t)(16 + lane) * 2048 +
                                (size_t)(rl >> 3) * 128 + (rl & 7) * 16) = e;
    }
  }
  if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(&w.flags[0], 1);
}

// ------------------------------------------------------------------ tensor-core top-1 / top-2
struct __align__(1024) L2Smem {
  uint8_t A[L2_RBK][L2_TILE_BYTES];
  uint8_t B[L2_NSTAGE][L2_TILE_BYTES];
  uint4 part[3][L2_RBK][L2_BLK];  // (key1, tile1, key2, tile2) of column quarters 1..3
  uint64_t full_a, empty_a;
  uint64_t full_b[L2_NSTAGE], empty_b[L2_NSTAGE];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

struct L2Item {
  int pair, dir, split, rbp;   // rbp = index of the 256-row block
  int fq, ft, nq, nt;
  int t_lo, ntile, nrb;        // column tiles [t_lo, t_lo+ntile), valid 128-row blocks (0..2)
};

__device__ __forceinline__ L2Item l2_item(int item, const int32_t* __restrict__ nkp,
                                          const int32_t* __restrict__ pairs, int nblk, int nrbp,
                                          int n_splits, int ndir) {
  L2Item it;
  it.rbp = item % nrbp;
  int x = item / nrbp;
  it.split = x % n_splits;
  x /= n_splits;
  it.dir = x % ndir;
  it.pair = x / ndir;
  it.fq = pairs[2 * it.pair + it.dir];
  it.ft = pairs[2 * it.pair + 1 - it.dir];
  it.nq = nkp[it.fq];
  it.nt = nkp[it.ft];
  const int ntile_all = (it.nt + L2_BLK - 1) / L2_BLK;
  const int per = (ntile_all + n_splits - 1) / n_splits;
  it.t_lo = it.split * per;
  it.ntile = max(min(ntile_all, it.t_lo + per) - it.t_lo, 0);
  const int r0 = it.rbp * L2_RBK * L2_BLK;
  it.nrb = r0 >= it.nq ? 0 : (r0 + L2_BLK < it.nq ? 2 : 1);
  if (it.nrb == 0) it.ntile = 0;
  return it;
}

// a packed key = (score bits << 7) | (31 - column inside the 32-column chunk): unsigned max =
// largest score, lowest column on ties.  t1/t2 hold the chunk number (tile * 4 + quarter).  Keys
// of different chunks compare by score only (key >> 7) and an earlier chunk keeps its place on
// ties, so the global order is "lowest column index wins".
struct L2Top {
  uint32_t k1, k2;
  int t1, t2;
};

__device__ __forceinline__ uint32_t umax3(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u32(a, b, c); }

// fold one tile's (h1 >= h2, both of tile t) into the running top-2
template <bool TOP2>
__device__ __forceinline__ void l2_fold_tile(L2Top& s, uint32_t h1, uint32_t h2, int t) {
  if (!TOP2) {
    const bool p = (h1 >> 7) > (s.k1 >> 7);
    s.k1 = p ? h1 : s.k1;
    s.t1 = p ? t : s.t1;
  } else {
    const bool a = (h1 >> 7) > (s.k1 >> 7);
    // a: new best = h1; second = h2 if it beats the old best (strictly), else the old best
    // !a: best stays; second = h1 if it beats the old second (strictly)
    const bool b = a ? ((h2 >> 7) > (s.k1 >> 7)) : ((h1 >> 7) > (s.k2 >> 7));
    const uint32_t n2 = a ? (b ? h2 : s.k1) : (b ? h1 : s.k2);
    const int nt2 = a ? (b ? t : s.t1) : (b ? t : s.t2);
    s.k2 = n2;
    s.t2 = nt2;
    s.k1 = a ? h1 : s.k1;
    s.t1 = a ? t : s.t1;
  }
}

template <bool ADDM, int C>
struct L2Pack {
  static __device__ __forceinline__ void run(uint32_t (&k)[32], const uint32_t (&v)[32], uint32_t mul) {
    uint32_t bits = v[C];
    if (ADDM) bits = __float_as_uint(__fadd_rn(__uint_as_float(bits), L2_MAGIC));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(k[C]) : "r"(bits), "r"(mul), "n"(31 - C));
    L2Pack<ADDM, C + 1>::run(k, v, mul);
  }
};
template <bool ADDM>
struct L2Pack<ADDM, 32> {
  static __device__ __forceinline__ void run(uint32_t (&)[32], const uint32_t (&)[32], uint32_t) {}
};

// rank 32 accumulator values (one TMEM chunk of this thread's row) into the running state.
// key = bits * 128 + (31 - c): `mul` holds 128 in a register the compiler cannot see through, so
// the index stays an immediate addend and the whole packing is ONE IMAD on the FMA pipe.
template <bool TOP2, bool ADDM>
__device__ __forceinline__ void l2_rank_chunk(L2Top& s, const uint32_t (&v)[32], uint32_t mul,
                                              int t) {
  uint32_t k[32];
  L2Pack<ADDM, 0>::run(k, v, mul);
  if (!TOP2) {
    // 3-input max tree: 32 -> 11 -> 4 -> 2 -> 1  (16 VIMNMX3/VIMNMX, depth 4)
    uint32_t a[11];
#pragma unroll
    for (int c = 0; c < 10; ++c) a[c] = umax3(k[3 * c], k[3 * c + 1], k[3 * c + 2]);
    a[10] = max(k[30], k[31]);
    const uint32_t b0 = umax3(a[0], a[1], a[2]), b1 = umax3(a[3], a[4], a[5]),
                   b2 = umax3(a[6], a[7], a[8]), b3 = max(a[9], a[10]);
    const uint32_t m = max(umax3(b0, b1, b2), b3);
    l2_fold_tile<false>(s, m, 0u, t);
  } else {
    // tournament of sorted pairs: merge (a1>=a2),(b1>=b2) -> (max(a1,b1), max3(min(a1,b1),a2,b2))
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      hi[c] = max(k[2 * c], k[2 * c + 1]);
      lo[c] = min(k[2 * c], k[2 * c + 1]);
    }
#pragma unroll
    for (int w = 8; w >= 1; w >>= 1) {
#pragma unroll
      for (int c = 0; c < w; ++c) {
        const uint32_t a1 = hi[c], b1 = hi[c + w];
        lo[c] = umax3(min(a1, b1), lo[c], lo[c + w]);
        hi[c] = max(a1, b1);
      }
    }
    l2_fold_tile<true>(s, hi[0], lo[0], t);
  }
}

// single-pass mutual mode: column maxima of one 32-column chunk over BOTH row blocks.  The two
// rows a thread owns are folded first (VIMNMX); the 32 rows of the warp are then folded by a
// transposing butterfly: at offset OFF a lane keeps the half of its columns selected by its lane
// bit and receives the partner's values for that half (31 SHFL + 31 max + 62 selects per chunk;
// one warp REDUX per column measured ~4x slower).  Lane c ends up with the best of column c.
template <int OFF>
__device__ __forceinline__ void l2_butterfly(uint32_t (&k)[32], int lane) {
  const bool hi = (lane & OFF) != 0;
#pragma unroll
  for (int j = 0; j < OFF; ++j) {
    const uint32_t send = hi ? k[j] : k[j + OFF];
    const uint32_t keep = hi ? k[j + OFF] : k[j];
    k[j] = max(keep, __shfl_xor_sync(0xffffffffu, send, OFF));
  }
}
template <bool ADDM, bool TWO>
__device__ __forceinline__ uint32_t l2_col_chunk_redux(const uint32_t (&v0)[32],
                                                       const uint32_t (&v1)[32], uint32_t mul2,
                                                       uint32_t cadd0, uint32_t cadd1, int lane) {
  uint32_t mine = 0u;
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    uint32_t b0 = v0[c], b1 = v1[c];
    if (ADDM) {
      b0 = __float_as_uint(__fadd_rn(__uint_as_float(b0), L2_MAGIC));
      b1 = __float_as_uint(__fadd_rn(__uint_as_float(b1), L2_MAGIC));
    }
    uint32_t k = b0 * mul2 + cadd0;
    if (TWO) k = max(k, b1 * mul2 + cadd1);
    const uint32_t r = __reduce_max_sync(0xffffffffu, k);
    mine = (lane == c) ? r : mine;
  }
  return mine;
}
template <bool ADDM, bool TWO>
__device__ __forceinline__ uint32_t l2_col_chunk(const uint32_t (&v0)[32], const uint32_t (&v1)[32],
                                                 uint32_t mul2, uint32_t cadd0, uint32_t cadd1,
                                                 int lane) {
  uint32_t k[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) {
    uint32_t b0 = v0[c], b1 = v1[c];
    if (ADDM) {
      b0 = __float_as_uint(__fadd_rn(__uint_as_float(b0), L2_MAGIC));
      b1 = __float_as_uint(__fadd_rn(__uint_as_float(b1), L2_MAGIC));
    }
    k[c] = b0 * mul2 + cadd0;
    if (TWO) k[c] = max(k[c], b1 * mul2 + cadd1);
  }
  l2_butterfly<16>(k, lane);
  l2_butterfly<8>(k, lane);
  l2_butterfly<4>(k, lane);
  l2_butterfly<2>(k, lane);
  l2_butterfly<1>(k, lane);
  return k[0];
}

// merge another warp's partial (other column quarter): order = score desc, global column asc
__device__ __forceinline__ bool l2_before(uint32_t ka, int ta, uint32_t kb, int tb) {
  const uint32_t sa = ka >> 7, sb = kb >> 7;
  if (sa != sb) return sa > sb;
  const int ca = ta * 32 + 31 - (int)(ka & 127u), cb = tb * 32 + 31 - (int)(kb & 127u);
  return ca < cb;
}
__device__ __forceinline__ void l2_insert(L2Top& s, uint32_t k, int t) {
  if (t < 0) return;
  if (s.t1 < 0 || l2_before(k, t, s.k1, s.t1)) {
    s.k2 = s.k1; s.t2 = s.t1; s.k1 = k; s.t1 = t;
  } else if (s.t2 < 0 || l2_before(k, t, s.k2, s.t2)) {
    s.k2 = k; s.t2 = t;
  }
}

// IMAD forms the compiler must not take apart (it would split the addend off or fuse it into a
// VIADDMNMX chain, which defeats the 3-input max trees): d = a * b + c
__device__ __forceinline__ uint32_t imad_r(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
// row keys of one fragment row p of a 16-lane half (8 columns 8 i + 2 j4 + e, i = 0..3, e = 0..1);
// radd[2 i + e] = 31 - (8 i + 2 j4 + e)
template <int P>
__device__ __forceinline__ uint32_t l2_frag_rowmax(const uint32_t (&v)[16], uint32_t mul, const uint32_t (&radd)[8]) {
  const uint32_t k0 = imad_r(v[0 + 2 * P], mul, radd[0]), k1 = imad_r(v[1 + 2 * P], mul, radd[1]);
  const uint32_t k2 = imad_r(v[4 + 2 * P], mul, radd[2]), k3 = imad_r(v[5 + 2 * P], mul, radd[3]);
  const uint32_t k4 = imad_r(v[8 + 2 * P], mul, radd[4]), k5 = imad_r(v[9 + 2 * P], mul, radd[5]);
  const uint32_t k6 = imad_r(v[12 + 2 * P], mul, radd[6]), k7 = imad_r(v[13 + 2 * P], mul, radd[7]);
  return max(umax3(umax3(k0, k1, k2), umax3(k3, k4, k5), k6), k7);
}

// Column half of the fragment epilogue: cmax[c] = best key of column 8 (c >> 1) + 2 j4 + (c & 1)
// among the thread's 8 rows.  A three-stage transposing butterfly over g8 leaves ONE column per
// thread; one 128-byte store per warp.
__device__ __forceinline__ void l2_frag_colstore(const L2Ws& w, const L2Item& it, const uint32_t (&cmax)[8],
                                                 int quarter, int cq, int g8, int j4, int t, int nblk,
                                                 int nrbp) {
  const bool b0 = (g8 & 1) != 0, b1 = (g8 & 2) != 0, b2 = (g8 & 4) != 0;
  uint32_t c1[4], c2[2];
#pragma unroll
  for (int i = 0; i < 4; ++i) {          // exchange over e
    const uint32_t send = b0 ? cmax[2 * i] : cmax[2 * i + 1], keep = b0 ? cmax[2 * i + 1] : cmax[2 * i];
    c1[i] = max(keep, __shfl_xor_sync(0xffffffffu, send, 4));
  }
#pragma unroll
  for (int x = 0; x < 2; ++x) {          // exchange over the low bit of i
    const uint32_t send = b1 ? c1[2 * x] : c1[2 * x + 1], keep = b1 ? c1[2 * x + 1] : c1[2 * x];
    c2[x] = max(keep, __shfl_xor_sync(0xffffffffu, send, 8));
  }
  const uint32_t send = b2 ? c2[0] : c2[1], keep = b2 ? c2[1] : c2[0];
  const uint32_t cm = max(keep, __shfl_xor_sync(0xffffffffu, send, 16));
  const int col = 8 * (2 * ((g8 >> 2) & 1) + ((g8 >> 1) & 1)) + 2 * j4 + (g8 & 1);
  const size_t ncol = (size_t)nblk * L2_BLK;
  w.colpart[(((size_t)it.pair * nrbp + it.rbp) * 4 + quarter) * ncol + (size_t)(it.t_lo + t) * L2_BLK +
            cq * 32 + col] = cm;
}

// Row butterfly of one row block: rmax[h * 2 + p] -> the best key of ONE row per thread
__device__ __forceinline__ uint32_t l2_frag_rowbest(const uint32_t (&rmax)[4], bool jb0, bool jb1) {
  uint32_t r1[2];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const uint32_t send = jb0 ? rmax[2 * h] : rmax[2 * h + 1], keep = jb0 ? rmax[2 * h + 1] : rmax[2 * h];
    r1[h] = max(keep, __shfl_xor_sync(0xffffffffu, send, 1));
  }
  const uint32_t send = jb1 ? r1[0] : r1[1], keep = jb1 ? r1[1] : r1[0];
  return max(keep, __shfl_xor_sync(0xffffffffu, send, 2));
}

// Key construction of one 16-lane half of a row block (the IMAD half of the epilogue): r0 / r1 =
// best row key among the thread's 8 columns of rows g8 / g8 + 8, cmax[c] folds the two rows in
// (column keys: bits * 256 + cadd(row); mul2 = 0 switches a row block off)
template <bool ADDM, bool FIRST>
__device__ __forceinline__ void l2_frag_keys(uint32_t (&v)[16], uint32_t cadd_p0, uint32_t cadd_p1, uint32_t mul,
                                             uint32_t mul2, const uint32_t (&radd)[8], uint32_t& r0,
                                             uint32_t& r1, uint32_t (&cmax)[8]) {
  if (ADDM) {
#pragma unroll
    for (int k = 0; k < 16; ++k) v[k] = __float_as_uint(__fadd_rn(__uint_as_float(v[k]), L2_MAGIC));
  }
  r0 = l2_frag_rowmax<0>(v, mul, radd);
  r1 = l2_frag_rowmax<1>(v, mul, radd);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const uint32_t k0 = imad_r(v[4 * (c >> 1) + (c & 1)], mul2, cadd_p0);
    const uint32_t k1 = imad_r(v[4 * (c >> 1) + 2 + (c & 1)], mul2, cadd_p1);
    cmax[c] = FIRST ? max(k0, k1) : umax3(cmax[c], k0, k1);
  }
}

// One B tile of the single-pass mutual mode with fragment-shaped TMEM loads (16x256b.x4: a thread
// holds rows {g8, g8 + 8} x columns {8i + 2 j4 + e} of a 16-lane half; two halves, two row blocks
// -> 8 rows x 8 columns per thread).  Rows: max over the thread's 8 columns (4 min/max ops), then a
// two-stage transposing butterfly over j4 leaves ONE row per row block in every thread.  Columns:
// max over the thread's 8 rows, then a three-stage butterfly over g8 leaves ONE column per thread.
// Software-pipelined so that the select/shuffle work (ALU pipe) has key construction (IMAD, FMA
// pipe) or a TMEM load to run beside: one 16-register load is in flight while the previous one is
// turned into keys, the column butterfly of tile t - 1 runs under the first load of tile t, the
// row butterfly of block 0 beside the keys of block 1.
template <bool ADDM>
__device__ __forceinline__ void l2_frag_tile(L2Smem& S, const L2Ws& w, const L2Item& it, L2Top (&st)[L2_RBK],
                                             const uint32_t (&cadd8)[8], const uint32_t (&radd)[8],
                                             uint32_t (&cpend)[8], uint32_t tmem, int as, int quarter, int cq,
                                             int g8, int j4, int lane, int t, int tcol, uint32_t mul,
                                             uint32_t mul2, int nblk, int nrbp, int dbg) {
  const uint32_t tbase = tmem + ((uint32_t)(quarter * 32) << 16) + as * (L2_RBK * L2_BLK) + cq * 32;
  const bool two = it.nrb == 2;
  const uint32_t mul2b = two ? mul2 : 0u;  // a missing second row block never wins a column
  const bool jb0 = (j4 & 1) != 0, jb1 = (j4 & 2) != 0;
  uint32_t va[16], vb[16];                 // [4 i + 2 p + e] of one 16-lane half
  uint32_t ra[4], rb4[4];
#ifdef L2_FRAG_TIMING                     // timing experiments only: the branches split the schedule
  const bool noload = (dbg & 4) != 0, norank = (dbg & 2) != 0;
#else
  constexpr bool noload = false, norank = false;
#endif
  if (noload) {
#pragma unroll
    for (int k = 0; k < 16; ++k) va[k] = vb[k] = L2_MAGIC_BITS + (uint32_t)(k + lane);
  }
  if (!noload) tc::ld_16x256b_x4_issue(tbase, va);                                  // block 0, lanes 0..15
  if (t > 0 && !norank) l2_frag_colstore(w, it, cpend, quarter, cq, g8, j4, t - 1, nblk, nrbp);
  if (!noload) { tc::ld_wait16(va); tc::ld_16x256b_x4_issue(tbase + (16u << 16), vb); }   // block 0, 16..31
  if (!norank) l2_frag_keys<ADDM, true>(va, cadd8[0], cadd8[1], mul, mul2, radd, ra[0], ra[1], cpend);
  if (!noload) { tc::ld_wait16(vb); tc::ld_16x256b_x4_issue(tbase + L2_BLK, va); }        // block 1, 0..15
  if (!norank) l2_frag_keys<ADDM, false>(vb, cadd8[2], cadd8[3], mul, mul2, radd, ra[2], ra[3], cpend);
  if (!noload) { tc::ld_wait16(va); tc::ld_16x256b_x4_issue(tbase + L2_BLK + (16u << 16), vb); }
  if (!norank) {
    l2_fold_tile<false>(st[0], l2_frag_rowbest(ra, jb0, jb1), 0u, tcol);
    l2_frag_keys<ADDM, false>(va, cadd8[4], cadd8[5], mul, mul2b, radd, rb4[0], rb4[1], cpend);
  }
  if (!noload) tc::ld_wait16(vb);
  tc::fence_before_sync();
  __syncwarp();
  if (lane == 0) tc::mbar_arrive(&S.tmem_empty[as]);   // the accumulators are in registers
  if (norank) return;
  l2_frag_keys<ADDM, false>(vb, cadd8[6], cadd8[7], mul, mul2b, radd, rb4[2], rb4[3], cpend);
  const uint32_t best1 = l2_frag_rowbest(rb4, jb0, jb1);
  if (two) l2_fold_tile<false>(st[1], best1, 0u, tcol);
  if (t == it.ntile - 1) l2_frag_colstore(w, it, cpend, quarter, cq, g8, j4, t, nblk, nrbp);
}

// MODE: TOP1 = nearest neighbour of every row (cross-check from two directions, fallback for
// inexact data); TOP2 = two nearest (knn + ratio); TOP1_COLMAX = nearest neighbour of every row
// AND of every column from the same accumulator (cross-check in ONE pass, exact data only).
// run_if: -1 always, 0 only when the data is exact (flags[0] == 0), 1 only when it is not.
// Persistent: CTA c handles work items c, c + gridDim.x, ...
template <int MODE, bool ADDM>
__global__ void __launch_bounds__(L2_THREADS, 1)
l2_top2_tc_kernel(const int32_t* __restrict__ nkp, const int32_t* __restrict__ pairs, int nblk,
                  int n_splits, int ndir, int n_items, int run_if, L2Ws w) {
  if (run_if >= 8) {                              // bit 3: skip when the integer path applies
    if (w.flags[2] == 0) return;
    run_if -= 9;                                  // 8 -> -1, 9 -> 0, 10 -> 1
  }
  constexpr bool TOP2 = (MODE == L2_MODE_TOP2);
  constexpr bool FRAG = (MODE == L2_MODE_TOP1_COLMAX_FRAG);   // fragment-shaped TMEM loads
  constexpr bool COLMAX = (MODE == L2_MODE_TOP1_COLMAX) || FRAG;
  if (run_if >= 0 && (w.flags[0] != 0) != (run_if != 0)) return;
  extern __shared__ uint8_t smem_raw[];
  L2Smem& S = *reinterpret_cast<L2Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nrbp = (nblk + L2_RBK - 1) / L2_RBK;

  if (threadIdx.x == 0) {
    mbar_init(&S.full_a, 1);
    mbar_init(&S.empty_a, 1);
    for (int s = 0; s < L2_NSTAGE; ++s) { mbar_init(&S.full_b[s], 1); mbar_init(&S.empty_b[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&S.tmem_full[s], 1); mbar_init(&S.tmem_empty[s], L2_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == L2_EPI_WARPS) tc::tmem_alloc(&S.tmem_base, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = S.tmem_base;
  // (the round-1 timing switches read g_l2_debug here at run time; branches around the TMEM loads
  // cost time by themselves, measured on K6i in round 2, so they are compiled out: the measurements
  // they produced are in profiles/r1_summary.md)
  constexpr int dbg = 0;

  if (warp == L2_EPI_WARPS + 1) {
    // ===== TMA producer: A blocks once per item, one 40 KB bulk copy per B tile =====
    if (lane == 0) {
      uint32_t g = 0, n_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const L2Item it = l2_item(item, nkp, pairs, nblk, nrbp, n_splits, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.empty_a, (n_it & 1) ^ 1);     // MMAs of the previous item are done with A
        mbar_expect_tx(&S.full_a, (uint32_t)it.nrb * L2_TILE_BYTES);
        for (int k = 0; k < it.nrb; ++k)
          bulk_g2s(S.A[k], w.tiles + ((size_t)it.fq * nblk + it.rbp * L2_RBK + k) * L2_TILE_ELEMS,
                   L2_TILE_BYTES, &S.full_a);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % L2_NSTAGE;
          mbar_wait(&S.empty_b[st], ((g / L2_NSTAGE) & 1) ^ 1);
          if ((dbg & 1) && g >= L2_NSTAGE) { tc::mbar_arrive(&S.full_b[st]); continue; }
          mbar_expect_tx(&S.full_b[st], L2_TILE_BYTES);
          bulk_g2s(S.B[st], w.tiles + ((size_t)it.ft * nblk + it.t_lo + t) * L2_TILE_ELEMS,
                   L2_TILE_BYTES, &S.full_b[st]);
        }
        ++n_it;
      }
    }
  } else if (warp == L2_EPI_WARPS) {
    // ===== MMA issuer: one thread, nrb x 9 x (128x128x16) per B tile, accumulators double-buffered
    if (lane == 0) {
      const uint32_t idesc = tc::idesc_bf16_f32(L2_BLK, L2_BLK);
      uint32_t g = 0, n_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const L2Item it = l2_item(item, nkp, pairs, nblk, nrbp, n_splits, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.full_a, n_it & 1);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % L2_NSTAGE, as = g & 1;
          mbar_wait(&S.full_b[st], (g / L2_NSTAGE) & 1);
          mbar_wait(&S.tmem_empty[as], ((g >> 1) & 1) ^ 1);
          tc::fence_after_sync();
          const uint32_t b_addr = smem_u32(S.B[st]);
          for (int rb = 0; rb < it.nrb; ++rb) {
            const uint32_t a_addr = smem_u32(S.A[rb]);
            const uint32_t d_addr = tmem + as * (L2_RBK * L2_BLK) + rb * L2_BLK;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
              // steps 0..7: descriptor columns; step 8: A extras (kg16,17) x B extras (kg18,19)
              const uint32_t ao = (k < 8) ? k * 2 * L2_LBO : 16 * L2_LBO;
              const uint32_t bo = (k < 8) ? k * 2 * L2_LBO : 18 * L2_LBO;
              tc::mma_bf16(d_addr, tc::smem_desc(a_addr + ao, L2_LBO, L2_SBO),
                           tc::smem_desc(b_addr + bo, L2_LBO, L2_SBO), idesc, k > 0);
            }
          }
          tc::commit(&S.empty_b[st]);     // smem stage reusable once these MMAs retire
          tc::commit(&S.tmem_full[as]);   // accumulators ready for the epilogue
        }
        tc::commit(&S.empty_a);           // A blocks reusable
        ++n_it;
      }
    }
  } else {
    // ===== epilogue warps 0..15: thread = TMEM lane = query row, warp/4 = 32-column quarter ====
    const int quarter = warp & 3, cq = warp >> 2;
    // FRAG: tcgen05.ld 16x256b hands a thread 2 rows x 8 columns of every 16-lane half; after the
    // row butterfly thread (g8, j4) owns row 16 (j4 >> 1) + 8 (j4 & 1) + g8 of its warp's 32
    const int g8 = lane >> 2, j4 = lane & 3;
    const int rl = FRAG ? quarter * 32 + 16 * (j4 >> 1) + 8 * (j4 & 1) + g8 : quarter * 32 + lane;
    const uint32_t mul = 128u + (uint32_t)(n_items < 0);   // = 128, opaque to the compiler
    const uint32_t mul2 = 256u + (uint32_t)(n_items < 0);  // = 256 (column keys)
    const float inv_s2 = 1.f / (l2_scale(w.flags) * l2_scale(w.flags));
    uint32_t radd[8];                      // FRAG: 31 - column of the thread's 8 fragment columns
#pragma unroll
    for (int k = 0; k < 8; ++k) radd[k] = 31u - (uint32_t)(8 * (k >> 1) + 2 * j4 + (k & 1));
    uint32_t g = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const L2Item it = l2_item(item, nkp, pairs, nblk, nrbp, n_splits, ndir);
      if (it.nrb == 0) continue;
      L2Top st[L2_RBK];
#pragma unroll
      for (int rb = 0; rb < L2_RBK; ++rb) { st[rb].k1 = st[rb].k2 = 0u; st[rb].t1 = st[rb].t2 = -1; }
      uint32_t cadd[L2_RBK] = {0u, 0u};
      uint32_t cpend[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // FRAG: column keys of the previous tile
      uint32_t cadd8[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};   // FRAG: the 8 rows this thread sees, slot = rb * 4 + h * 2 + p
      if (COLMAX && !FRAG) {
        const float s2 = l2_scale(w.flags) * l2_scale(w.flags);
#pragma unroll
        for (int rb = 0; rb < L2_RBK; ++rb) {
          const int row = (it.rbp * L2_RBK + rb) * L2_BLK + rl;
          uint32_t na = L2_PAD_NORM;
          if (rb < it.nrb && row < it.nq)
            na = (uint32_t)__float2int_rn(s2 * w.norms[(size_t)it.fq * nblk * L2_BLK + row]);
          cadd[rb] = (255u - (uint32_t)(rb * L2_BLK + rl)) - ((na + L2_COL_OFFSET) << 8);
        }
      }
      if constexpr (FRAG) {
        // the 8 norm loads are unconditional (clamped index) so that they are all in flight at once
        const float s2 = l2_scale(w.flags) * l2_scale(w.flags);
        float nv[8];
#pragma unroll
        for (int sl = 0; sl < 8; ++sl) {
          const int rb = sl >> 2, rloc = quarter * 32 + 16 * ((sl >> 1) & 1) + 8 * (sl & 1) + g8;
          const int row = min((it.rbp * L2_RBK + rb) * L2_BLK + rloc, nblk * L2_BLK - 1);
          nv[sl] = w.norms[(size_t)it.fq * nblk * L2_BLK + row];
        }
#pragma unroll
        for (int sl = 0; sl < 8; ++sl) {
          const int rb = sl >> 2, rloc = quarter * 32 + 16 * ((sl >> 1) & 1) + 8 * (sl & 1) + g8;
          const int row = (it.rbp * L2_RBK + rb) * L2_BLK + rloc;
          const uint32_t na = (rb < it.nrb && row < it.nq) ? (uint32_t)__float2int_rn(s2 * nv[sl]) : L2_PAD_NORM;
          cadd8[sl] = (255u - (uint32_t)(rb * L2_BLK + rloc)) - ((na + L2_COL_OFFSET) << 8);
        }
      }
      for (int t = 0; t < it.ntile; ++t, ++g) {
        const int as = g & 1;
        mbar_wait(&S.tmem_full[as], (g >> 1) & 1);
        tc::fence_after_sync();
        const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + as * (L2_RBK * L2_BLK) + cq * 32;
        const int tcol = (it.t_lo + t) * 4 + cq;   // chunk number
        if constexpr (FRAG) {
          l2_frag_tile<ADDM>(S, w, it, st, cadd8, radd, cpend, tmem, as, quarter, cq, g8, j4, lane, t, tcol, mul, mul2,
                             nblk, nrbp, dbg);
          continue;
        }
        uint32_t v0[32], v1[32];
        if (dbg & 4) {
#pragma unroll
          for (int c = 0; c < 32; ++c) v0[c] = v1[c] = L2_MAGIC_BITS;
          tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(&S.tmem_empty[as]);
          if (dbg & 2) continue;
          l2_rank_chunk<TOP2, ADDM>(st[0], v0, mul, tcol);
          l2_rank_chunk<TOP2, ADDM>(st[1], v1, mul, tcol);
          continue;
        }
        // block 0 is ranked while the load of block 1 is in flight (block 1's TMEM columns always
        // exist; their content is ignored when the item has a single row block)
        tc::ld_32x32_issue(taddr, v0);
        tc::ld_wait(v0);
        tc::ld_32x32_issue(taddr + L2_BLK, v1);
        if (!(dbg & 2)) l2_rank_chunk<TOP2, ADDM>(st[0], v0, mul, tcol);
        tc::ld_wait(v1);
        tc::fence_before_sync();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&S.tmem_empty[as]);   // both accumulators are in registers
        if (!(dbg & 2) && it.nrb == 2) l2_rank_chunk<TOP2, ADDM>(st[1], v1, mul, tcol);
        if (COLMAX && !(dbg & 2)) {
          uint32_t cm;
          if (dbg & 32)
            cm = (it.nrb == 2) ? l2_col_chunk_redux<ADDM, true>(v0, v1, mul2, cadd[0], cadd[1], lane)
                               : l2_col_chunk_redux<ADDM, false>(v0, v1, mul2, cadd[0], cadd[1], lane);
          else
            cm = (it.nrb == 2) ? l2_col_chunk<ADDM, true>(v0, v1, mul2, cadd[0], cadd[1], lane)
                               : l2_col_chunk<ADDM, false>(v0, v1, mul2, cadd[0], cadd[1], lane);
          // one coalesced 128 B store per warp: best row of 32 columns among this warp's 64 rows
          const size_t ncol = (size_t)nblk * L2_BLK;
          w.colpart[(((size_t)it.pair * nrbp + it.rbp) * 4 + quarter) * ncol +
                    (size_t)(it.t_lo + t) * L2_BLK + cq * 32 + lane] = cm;
        }
      }
      // ---- merge the four column quarters of every row, convert, write ----
      if (cq > 0) {
#pragma unroll
        for (int rb = 0; rb < L2_RBK; ++rb)
          S.part[cq - 1][rb][rl] = make_uint4(st[rb].k1, (uint32_t)st[rb].t1, st[rb].k2, (uint32_t)st[rb].t2);
      }
      asm volatile("bar.sync 1, 512;" ::: "memory");   // the 16 epilogue warps only
      if (cq == 0) {
#pragma unroll
        for (int rb = 0; rb < L2_RBK; ++rb) {
          if (rb >= it.nrb) break;
          L2Top m = st[rb];
          if (m.t2 >= 0 && m.t1 < 0) { m.k1 = m.k2; m.t1 = m.t2; m.t2 = -1; }
#pragma unroll
          for (int p = 0; p < 3; ++p) {
            const uint4 o = S.part[p][rb][rl];
            l2_insert(m, o.x, (int)o.y);
            if (TOP2) l2_insert(m, o.z, (int)o.w);
          }
          const int row = (it.rbp * L2_RBK + rb) * L2_BLK + rl;
          if (row < it.nq) {
            // d^2 = |a|^2 - 2 (a.b - |b|^2/2); exact for integer descriptors
            const float na = w.norms[(size_t)it.fq * nblk * L2_BLK + row];
            int c1 = m.t1 >= 0 ? m.t1 * 32 + 31 - (int)(m.k1 & 127u) : -1;
            int c2 = (TOP2 && m.t2 >= 0) ? m.t2 * 32 + 31 - (int)(m.k2 & 127u) : -1;
            if (c1 >= it.nt) c1 = -1;     // padded column (only when fewer real columns exist)
            if (c2 >= it.nt) c2 = -1;
            const float x1 = (float)((int)(m.k1 >> 7) - (int)L2_QBIAS);   // 2 s^2 (a.b - |b|^2/2)
            const float x2 = (float)((int)(m.k2 >> 7) - (int)L2_QBIAS);
            float4 r;
            r.x = c1 >= 0 ? fmaxf(fmaf(-x1, inv_s2, na), 0.f) : __int_as_float(0x7f800000);
            r.y = __int_as_float(c1);
            r.z = c2 >= 0 ? fmaxf(fmaf(-x2, inv_s2, na), 0.f) : __int_as_float(0x7f800000);
            r.w = __int_as_float(c2);
            w.rowbest[(((size_t)it.pair * 2 + it.dir) * n_splits + it.split) * nblk * L2_BLK + row] = r;
          }
        }
      }
      asm volatile("bar.sync 1, 512;" ::: "memory");   // part[] may be overwritten by the next item
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == L2_EPI_WARPS) tc::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------ fp32 rerank (non-integer data)
// When some descriptor value is not exactly representable in bf16 the GEMM distances are
// approximate: recompute the two candidates of every row exactly in fp32 (same formula as cv2:
// sum of squared differences) and re-order them.  One warp per row.
__global__ void __launch_bounds__(256)
l2_rerank_kernel(const float* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
                 const int32_t* __restrict__ pairs, int nblk, int n_splits, int ndir, L2Ws w) {
  if (w.flags[0] == 0) return;
  const int pair = blockIdx.y / ndir, dir = blockIdx.y % ndir;
  const int fq = pairs[2 * pair + dir], ft = pairs[2 * pair + 1 - dir];
  const int nq = nkp[fq];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + warp;
  if (row >= nq) return;
  const float4 a = __ldg(reinterpret_cast<const float4*>(desc + ((size_t)fq * n_cap + row) * L2_DIM) + lane);
  for (int s = 0; s < n_splits; ++s) {
    float4* slot = w.rowbest + (((size_t)pair * 2 + dir) * n_splits + s) * nblk * L2_BLK + row;
    float4 rbv = *slot;
    int idx[2] = {__float_as_int(rbv.y), __float_as_int(rbv.w)};
    float d[2];
    for (int c = 0; c < 2; ++c) {
      float ss = __int_as_float(0x7f800000);
      if (idx[c] >= 0) {
        const float4 b = __ldg(reinterpret_cast<const float4*>(desc + ((size_t)ft * n_cap + idx[c]) * L2_DIM) + lane);
        const float dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z, dw = a.w - b.w;
        ss = dx * dx + dy * dy + dz * dz + dw * dw;
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
      }
      d[c] = ss;
    }
    if (lane == 0) {
      const bool swap = (d[1] < d[0]) || (d[1] == d[0] && idx[1] >= 0 && idx[1] < idx[0]);
      *slot = swap ? make_float4(d[1], __int_as_float(idx[1]), d[0], __int_as_float(idx[0]))
                   : make_float4(d[0], __int_as_float(idx[0]), d[1], __int_as_float(idx[1]));
    }
  }
}

// ------------------------------------------------------------------ compaction
__device__ __forceinline__ void merge_top2(float4 v, float& s1, int& i1, float& s2, int& i2) {
  const int a = __float_as_int(v.y), b = __float_as_int(v.w);
  if (a >= 0 && (v.x < s2 || (v.x == s2 && (i2 < 0 || a < i2)))) {
    if (v.x < s1 || (v.x == s1 && (i1 < 0 || a < i1))) { s2 = s1; i2 = i1; s1 = v.x; i1 = a; }
    else { s2 = v.x; i2 = a; }
  }
  if (b >= 0 && (v.z < s2 || (v.z == s2 && (i2 < 0 || b < i2)))) {
    if (v.z < s1 || (v.z == s1 && (i1 < 0 || b < i1))) { s2 = s1; i2 = i1; s1 = v.z; i1 = b; }
    else { s2 = v.z; i2 = b; }
  }
}

constexpr int L2C_THREADS = 1024;
__global__ void __launch_bounds__(L2C_THREADS)
l2_compact_kernel(const int32_t* __restrict__ nkp, const float* __restrict__ kp, int n_cap,
                  const int32_t* __restrict__ pairs, int nblk, int n_splits, int mode, double ratio,
                  int colmode, L2Ws w, int32_t* __restrict__ q, int32_t* __restrict__ t,
                  float* __restrict__ dist, int32_t* __restrict__ t2, float* __restrict__ dist2,
                  double* __restrict__ pts, int32_t* __restrict__ n_out) {
  __shared__ int warp_tot[L2C_THREADS / 32];
  __shared__ int base_s;
  const int pair = blockIdx.x;
  const int f1 = pairs[2 * pair], f2 = pairs[2 * pair + 1];
  const int n1 = nkp[f1], n2 = nkp[f2];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const size_t stride = (size_t)nblk * L2_BLK;
  const float4* fwd = w.rowbest + ((size_t)pair * 2 + 0) * n_splits * stride;
  const float4* bwd = w.rowbest + ((size_t)pair * 2 + 1) * n_splits * stride;
  const int nrbp = (nblk + L2_RBK - 1) / L2_RBK;
  const uint32_t* colp = w.colpart + (size_t)pair * nrbp * 4 * stride;
  // colmode bit 1: the integer path may have run (two directions in rowbest, no column partials)
  const bool use_col = (colmode & 1) != 0 && w.flags[0] == 0 && !((colmode & 2) != 0 && w.flags[2] == 0);
  const size_t ob = (size_t)pair * n_cap;
  const float INF = __int_as_float(0x7f800000);
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n1; i0 += L2C_THREADS) {
    const int i = i0 + tid;
    float s1 = INF, s2 = INF;
    int i1 = -1, i2 = -1;
    bool keep = false;
    if (i < n1 && n2 > 0) {
      for (int s = 0; s < n_splits; ++s) merge_top2(fwd[s * stride + i], s1, i1, s2, i2);
      if (mode == VO_MATCH_MUTUAL && use_col) {
        // single-pass path: best query row of column i1, folded over the 256-row blocks
        // (ascending, strict >: the lowest row wins ties, as in OpenCV)
        int best_row = -1;
        if (i1 >= 0) {
          const int nrbp_q = (n1 + L2_RBK * L2_BLK - 1) / (L2_RBK * L2_BLK);
          uint32_t best = 0u;
          for (int r = 0; r < nrbp_q * 4; ++r) {   // ascending 256-row blocks; quarters by key
            const uint32_t k = colp[(size_t)r * stride + i1];
            const int row = (r >> 2) * L2_RBK * L2_BLK + 255 - (int)(k & 255u);
            if ((k >> 8) > best || ((k >> 8) == best && row < best_row)) { best = k >> 8; best_row = row; }
          }
        }
        keep = (best_row == i);
      } else if (mode == VO_MATCH_MUTUAL) {
        float b1 = INF, b2 = INF;
        int j1 = -1, j2 = -1;
        if (i1 >= 0)
          for (int s = 0; s < n_splits; ++s) merge_top2(bwd[s * stride + i1], b1, j1, b2, j2);
        keep = (j1 == i);
      } else {
        const float d1 = sqrtf(s1), d2 = sqrtf(s2);
        keep = i1 >= 0 && (ratio < 0.0 ? true : (i2 >= 0 && (double)d1 < ratio * (double)d2));
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_tot[wid] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int wv = 0; wv < wid; ++wv) off += warp_tot[wv];
    off += __popc(bal & ((1u << lane) - 1u));
    if (keep) {
      q[ob + off] = i;
      t[ob + off] = i1;
      dist[ob + off] = sqrtf(s1);
      if (t2) t2[ob + off] = i2;
      if (dist2) dist2[ob + off] = i2 >= 0 ? sqrtf(s2) : 0.f;
      if (pts && kp) {
        const float2 p1 = reinterpret_cast<const float2*>(kp)[(size_t)f1 * n_cap + i];
        const float2 p2 = reinterpret_cast<const float2*>(kp)[(size_t)f2 * n_cap + i1];
        double2* o = reinterpret_cast<double2*>(pts) + 2 * (ob + off);
        o[0] = make_double2((double)p1.x, (double)p1.y);
        o[1] = make_double2((double)p2.x, (double)p2.y);
      }
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int wv = 0; wv < L2C_THREADS / 32; ++wv) tot += warp_tot[wv];
      base_s += tot;
    }
    __syncthreads();
  }
  if (tid == 0) n_out[pair] = base_s;
}

}  // namespace vo

using namespace vo;

static int l2_num_sms() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!sms[dev] &&
      (cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms[dev] <= 0))
    sms[dev] = 148;
  return sms[dev];
}

// column splits only when there are too few (pair, direction, 256-row block) items to give every
// SM at least one
static int l2_pick_splits(int n_pairs, int ndir, int nblk) {
  const long items = (long)n_pairs * ndir * ((nblk + L2_RBK - 1) / L2_RBK);
  int s = (int)((l2_num_sms() + items - 1) / items);
  if (s > nblk) s = nblk;
  if (s > 8) s = 8;
  return s < 1 ? 1 : s;
}

static int g_l2_debug_host = 0;
extern "C" int vo_debug_l2(int flags) {
  g_l2_debug_host = flags;
  return cudaMemcpyToSymbol(g_l2_debug, &flags, sizeof(int)) == cudaSuccess ? VO_OK : VO_ERR_CUDA;
}

extern "C" size_t vo_match_l2_workspace_bytes(int n_frames, int n_pairs, int n_cap) {
  if (n_frames <= 0 || n_pairs <= 0 || n_cap <= 0) return 0;
  return l2_carve(nullptr, n_frames, n_pairs, n_cap, 8).bytes;
}

extern "C" int vo_match_l2_batch(const float* desc, const int32_t* nkp, const float* kp, int n_cap,
                                 int dim, int n_frames, const int32_t* pairs, int n_pairs, int mode,
                                 double ratio, int32_t* q, int32_t* t, float* dist, int32_t* t2,
                                 float* dist2, double* pts, int32_t* n_out, void* ws,
                                 size_t ws_bytes, void* stream) {
  VO_REQUIRE(desc && nkp && pairs && q && t && dist && n_out && ws, VO_ERR_INVALID_ARG,
             "vo_match_l2_batch: null pointer");
  VO_REQUIRE(dim == L2_DIM, VO_ERR_INVALID_ARG, "vo_match_l2_batch: dim must be 128 (SIFT), got %d", dim);
  VO_REQUIRE(n_cap > 0 && n_frames > 0 && n_pairs >= 0 && n_pairs <= 32767, VO_ERR_INVALID_ARG,
             "vo_match_l2_batch: bad sizes");
  VO_REQUIRE(mode == VO_MATCH_MUTUAL || mode == VO_MATCH_KNN2_RATIO, VO_ERR_INVALID_ARG,
             "vo_match_l2_batch: bad mode %d", mode);
  VO_REQUIRE(ws_bytes >= vo_match_l2_workspace_bytes(n_frames, n_pairs, n_cap), VO_ERR_WORKSPACE,
             "vo_match_l2_batch: workspace too small");
  if (n_pairs == 0) return VO_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int nblk = (n_cap + L2_BLK - 1) / L2_BLK;
  const int ndir = (mode == VO_MATCH_MUTUAL) ? 2 : 1;   // rerank grid: directions that may exist
  const int splits = l2_pick_splits(n_pairs, 1, nblk);
  L2Ws w = l2_carve(ws, n_frames, n_pairs, n_cap, splits);
  const bool use_i8 = (g_l2_debug_host & 128) == 0;    // debug bit 7: bf16 kernels only
  const int ntile8 = (n_cap + l8::ROWS - 1) / l8::ROWS;
  cudaMemsetAsync(w.flags, 0, 4 * sizeof(int32_t), st);
  l2_norm_kernel<<<dim3(nblk, n_frames), 256, 0, st>>>(desc, nkp, n_cap, nblk, w);
  if (use_i8)
    l8::tile_kernel<<<dim3(ntile8, n_frames), 256, 0, st>>>(desc, nkp, n_cap, ntile8, w.norms, nblk * L2_BLK,
                                                           w.tiles8, w.flags);
  l2_tile_kernel<<<dim3(nblk, n_frames), 256, 0, st>>>(desc, nkp, n_cap, nblk, use_i8 ? 1 : 0, w);
  typedef void (*L2Kernel)(const int32_t*, const int32_t*, int, int, int, int, int, L2Ws);
  static const L2Kernel kernels[8] = {
      l2_top2_tc_kernel<L2_MODE_TOP1, false>, l2_top2_tc_kernel<L2_MODE_TOP2, false>,
      l2_top2_tc_kernel<L2_MODE_TOP1_COLMAX, false>, l2_top2_tc_kernel<L2_MODE_TOP1_COLMAX_FRAG, false>,
      l2_top2_tc_kernel<L2_MODE_TOP1, true>, l2_top2_tc_kernel<L2_MODE_TOP2, true>,
      l2_top2_tc_kernel<L2_MODE_TOP1_COLMAX, true>, l2_top2_tc_kernel<L2_MODE_TOP1_COLMAX_FRAG, true>};
  static bool attr_set[64] = {false};     // cudaFuncSetAttribute is per device
  const int smem = (int)sizeof(L2Smem) + 1024;
  const int smem8 = (int)sizeof(l8::Smem) + 1024;
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!attr_set[dev]) {
    for (int k = 0; k < 8; ++k)
      if (cudaFuncSetAttribute(kernels[k], cudaFuncAttributeMaxDynamicSharedMemorySize, smem) !=
          cudaSuccess)
        return check_launch("vo_match_l2_batch(smem attribute)");
    typedef void (*K8)(const uint8_t*, const int32_t*, const int32_t*, int, int, int, long, const float*, int,
                       int, int, const int32_t*, float4*);
    const K8 all8[6] = {l8::match_kernel<false, 0>, l8::match_kernel<false, 1>, l8::match_kernel<false, 2>,
                        l8::match_kernel<true, 0>,  l8::match_kernel<true, 1>,  l8::match_kernel<true, 2>};
    for (int k = 0; k < 6; ++k)
      if (cudaFuncSetAttribute(all8[k], cudaFuncAttributeMaxDynamicSharedMemorySize, smem8) != cudaSuccess)
        return check_launch("vo_match_l2_batch(smem attribute, integer kernel)");
    attr_set[dev] = true;
  }
  const int addm = (g_l2_debug_host & 8) ? 4 : 0;
  // experimental: fragment-shaped TMEM loads (fewer instructions, measured 3 % slower: profiles/r1_summary.md)
  const bool frag_layout = (g_l2_debug_host & 64) != 0;
  const bool two_pass_only = (g_l2_debug_host & 16) != 0;   // diagnostics: old mutual path
  {
    ProfScope prof(2, st);
    const int nrbp = (nblk + L2_RBK - 1) / L2_RBK;
    if (use_i8) {   // uint8 data (cv2 SIFT): both directions as integer MMA passes; exits at once otherwise
      const int nd = (mode == VO_MATCH_MUTUAL) ? 2 : 1;
      const long n_items = (long)n_pairs * nd * ntile8;
      const int grid = (int)(n_items < l2_num_sms() ? n_items : l2_num_sms());
      // epilogue variant of the integer kernel (compile-time instantiations): default 1 (explicit
      // ld.shared for the column constants: 0.260 vs 0.302 ms mutual at 200 x 2000^2); debug bits
      // 8..9 = 1 selects variant 0, = 2 variant 2 (early release: equal to 1 within noise)
      static const int vmap[4] = {1, 0, 2, 1};
      const int variant = vmap[(g_l2_debug_host >> 8) & 3];
      typedef void (*K8)(const uint8_t*, const int32_t*, const int32_t*, int, int, int, long, const float*, int,
                         int, int, const int32_t*, float4*);
      const K8 tab[6] = {l8::match_kernel<false, 0>, l8::match_kernel<false, 1>, l8::match_kernel<false, 2>,
                         l8::match_kernel<true, 0>,  l8::match_kernel<true, 1>,  l8::match_kernel<true, 2>};
      tab[(mode == VO_MATCH_MUTUAL ? 0 : 3) + variant]<<<grid, l8::THREADS, smem8, st>>>(
          w.tiles8, nkp, pairs, n_cap, ntile8, nd, n_items, w.norms, nblk * L2_BLK, splits, nblk * L2_BLK,
          w.flags, w.rowbest);
    }
    auto launch = [&](int mode_k, int nd, int run_if) {
      const long n_items = (long)n_pairs * nd * splits * nrbp;
      const int grid = (int)(n_items < l2_num_sms() ? n_items : l2_num_sms());
      kernels[mode_k + addm]<<<grid, L2_THREADS, smem, st>>>(nkp, pairs, nblk, splits, nd,
                                                            (int)n_items, use_i8 ? run_if + 9 : run_if, w);
    };
    if (mode != VO_MATCH_MUTUAL) {
      launch(L2_MODE_TOP2, 1, -1);
    } else if (two_pass_only) {
      launch(L2_MODE_TOP1, 2, -1);
    } else {
      // exact (integer) data: rows and columns ranked from ONE accumulator pass; anything else:
      // both directions + fp32 rerank.  The prep kernels decide on the device (flags[0]); the
      // launch that does not apply exits at once.
      launch(frag_layout ? L2_MODE_TOP1_COLMAX_FRAG : L2_MODE_TOP1_COLMAX, 1, 0);
      // inexact data: TWO candidates per row in both directions, so that the fp32 rerank can
      // undo a bf16 rank flip between the best two (a true neighbour the bf16 ranking puts third
      // or lower is still missed: approximate for non-integer descriptors, see vo_b200.h)
      launch(L2_MODE_TOP2, 2, 1);
    }
  }
  l2_rerank_kernel<<<dim3((n_cap + 7) / 8, n_pairs * ndir), 256, 0, st>>>(desc, nkp, n_cap, pairs,
                                                                        nblk, splits, ndir, w);
  l2_compact_kernel<<<n_pairs, L2C_THREADS, 0, st>>>(nkp, kp, n_cap, pairs, nblk, splits, mode, ratio,
                                                    (two_pass_only ? 0 : 1) | (use_i8 ? 2 : 0), w, q, t, dist, t2, dist2, pts, n_out);
  return check_launch("vo_match_l2_batch");
}

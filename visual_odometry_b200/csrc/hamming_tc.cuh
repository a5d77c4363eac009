// hamming_tc.cuh — K1t: 256-bit Hamming matching as an exact int8 contraction on tcgen05.
//
// Same reference call as hamming.cu (cv2.BFMatcher(NORM_HAMMING, crossCheck=True).match,
// reference src/image_pair.py:234 with the matcher of src/image_and_keypoints.py:9; knn(k=2) +
// ratio of src/feature_detection.py:20-26), for large batches of pairs.
//
// With the bits of a descriptor written as +-16 (int8), the dot product of two descriptors is
// 256 * (256 - 2 * Hamming): an s32 tensor-core GEMM with K = 256 ranks the neighbours EXACTLY
// (kind::i8, s32 accumulators in TMEM; no rounding anywhere).  A ninth K = 32 step multiplies a
// "ones" group of the row operand (16, 1, 64 x 9) with an "index" group of the column operand
// ((255 - j) >> 4, (255 - j) & 15, and -128 x 9 on padded rows), so the accumulator ALREADY holds
//     key(i, j) = 256 * dot(i, j) + (255 - j mod 256)          (padded columns: < -65536)
// and the whole epilogue is one signed max per element: the largest key of a row is its nearest
// neighbour, lowest column on ties = OpenCV's order.  The cross-check needs the nearest query of
// every train descriptor as well; a transposed reduction over TMEM lanes costs ~4 issue slots per
// element (measured on K6, profiles/r1_summary.md), so the second direction is simply a second
// MMA pass with the operands swapped (the operand form is the same for both sides): the tensor
// pipe has the headroom, the integer pipes do not.
//
// Operand form ("tile"): 256 rows x 320 bytes in the UMMA canonical K-major no-swizzle layout,
//   byte(r, kg, kk) = kg * 4096 + r * 16 + kk,   kg = 16-byte K group:
//   kg 0..15 = the 256 bits as +-16, kg 16..17 = ones group, kg 18..19 = index group.
// A row operand is kg 0..17 (one 72 KB bulk copy), a column operand kg 0..15 + 18..19 (two).
// Kernel: persistent, warp-specialised (warp 9 = bulk-copy producer, warp 8 = one-thread MMA
// issuer, warps 0..7 = epilogue); a work item = (pair, direction, 256-row block); every 72 KB
// column tile feeds two M = 128 row blocks with N = 256 (18 x tcgen05.mma 128x256x32), the two
// 256-column accumulators double-buffer each other.
#pragma once
#include "tc.cuh"

namespace vo {
namespace htc {

constexpr int ROWS = 256;                       // rows per operand tile
constexpr int KG_BYTES = ROWS * 16;             // one 16-byte K group of all rows = 4096
constexpr int TILE_BYTES = 20 * KG_BYTES;       // 81920 in global memory
constexpr int OPER_BYTES = 18 * KG_BYTES;       // 73728 in shared memory
constexpr uint32_t LBO = KG_BYTES, SBO = 128;
constexpr int NSTAGE = 2;
constexpr int EPI_WARPS = 8;                    // warp & 3 = TMEM lane quarter, warp >> 2 = column half
constexpr int THREADS = (EPI_WARPS + 2) * 32;
constexpr int KEY_MININT = (int)0x80000000;


// kind::i8: A, B signed 8 bit, K-major, D = s32, M x N
__host__ __device__ constexpr uint32_t idesc_s8_s32(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc,
                                       uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// ------------------------------------------------------------------ prep: bits -> int8 operand tiles
// four bits -> four bytes, 0x10 (+16) for a set bit and 0xF0 (-16) for a clear one
__device__ __forceinline__ uint32_t spread4(uint32_t b4) {
  const uint32_t x = (b4 * 0x00204081u) & 0x01010101u;
  return 0xF0F0F0F0u ^ (x * 0xE0u);
}

static __global__ void __launch_bounds__(ROWS)
expand_kernel(const uint8_t* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
              int ntile, uint8_t* __restrict__ tiles) {
  const int f = blockIdx.y, tl = blockIdx.x, r = threadIdx.x;
  const int row = tl * ROWS + r;
  const bool ok = row < min(nkp[f], n_cap);
  uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  if (ok) {
    const uint4* p = reinterpret_cast<const uint4*>(desc + ((size_t)f * n_cap + row) * 32);
    const uint4 a = __ldg(p), b = __ldg(p + 1);
    w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
  }
  uint8_t* tile = tiles + ((size_t)f * ntile + tl) * TILE_BYTES + (size_t)r * 16;
#pragma unroll
  for (int kg = 0; kg < 16; ++kg) {
    const uint32_t bits = (w[kg >> 1] >> ((kg & 1) * 16)) & 0xffffu;
    uint4 o = make_uint4(0u, 0u, 0u, 0u);   // padded rows: dot = 0 with everything
    if (ok) o = make_uint4(spread4(bits & 15u), spread4((bits >> 4) & 15u),
                           spread4((bits >> 8) & 15u), spread4(bits >> 12));
    *reinterpret_cast<uint4*>(tile + (size_t)kg * KG_BYTES) = o;
  }
  // ones group: bytes (16, 1, 64 x 9, 0 ...)
  *reinterpret_cast<uint4*>(tile + 16 * KG_BYTES) = make_uint4(0x40400110u, 0x40404040u, 0x00404040u, 0u);
  *reinterpret_cast<uint4*>(tile + 17 * KG_BYTES) = make_uint4(0u, 0u, 0u, 0u);
  // index group: real row ((255 - r) >> 4, (255 - r) & 15, 0 ...): 16 * hi + lo = 255 - r;
  // padded row (0, 0, -128 x 9): 9 * 64 * -128 = -73728 < every real key (>= -65536)
  const uint32_t x = 255u - (uint32_t)r;
  const uint4 idx = ok ? make_uint4((x >> 4) | ((x & 15u) << 8), 0u, 0u, 0u)
                       : make_uint4(0x80800000u, 0x80808080u, 0x00808080u, 0u);
  *reinterpret_cast<uint4*>(tile + 18 * KG_BYTES) = idx;
  *reinterpret_cast<uint4*>(tile + 19 * KG_BYTES) = make_uint4(0u, 0u, 0u, 0u);
}

// ------------------------------------------------------------------ the contraction + ranking
struct __align__(1024) Smem {
  uint8_t A[OPER_BYTES];
  uint8_t B[NSTAGE][OPER_BYTES];
  int2 part[2][2][128];          // [item parity][row block][row]: (best, second) of column half 1
  uint64_t full_a, empty_a;
  uint64_t full_b[NSTAGE], empty_b[NSTAGE];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

struct Item {
  int pair, dir, rbp;      // rbp = 256-row block of the row-side frame
  int fq, ft, nq, nt;      // row-side / column-side frame and their key-point counts
  int ntile, nrb;          // column tiles, valid 128-row blocks (0..2)
};

__device__ __forceinline__ Item get_item(long item, const int32_t* __restrict__ nkp,
                                         const int32_t* __restrict__ pairs, int n_cap, int nrbp,
                                         int ndir) {
  Item it;
  it.rbp = (int)(item % nrbp);
  const long x = item / nrbp;
  it.dir = (int)(x % ndir);
  it.pair = (int)(x / ndir);
  it.fq = pairs[2 * it.pair + it.dir];
  it.ft = pairs[2 * it.pair + 1 - it.dir];
  it.nq = min(nkp[it.fq], n_cap);
  it.nt = min(nkp[it.ft], n_cap);
  it.ntile = (it.nt + ROWS - 1) / ROWS;
  const int r0 = it.rbp * ROWS;
  it.nrb = r0 >= it.nq ? 0 : (r0 + 128 < it.nq ? 2 : 1);
  if (it.nrb == 0) it.ntile = 0;
  if (it.ntile == 0) it.nrb = 0;
  return it;
}

__device__ __forceinline__ int smax3(int a, int b, int c) { return __vimax3_s32(a, b, c); }

// largest of 32 keys: 3-input max tree (16 instructions)
__device__ __forceinline__ int max32(const uint32_t (&v)[32]) {
  int a[11];
#pragma unroll
  for (int c = 0; c < 10; ++c) a[c] = smax3((int)v[3 * c], (int)v[3 * c + 1], (int)v[3 * c + 2]);
  a[10] = max((int)v[30], (int)v[31]);
  const int b0 = smax3(a[0], a[1], a[2]), b1 = smax3(a[3], a[4], a[5]), b2 = smax3(a[6], a[7], a[8]),
            b3 = max(a[9], a[10]);
  return max(smax3(b0, b1, b2), b3);
}
// two largest of 32 keys (all keys of a row are distinct): tournament of sorted pairs
__device__ __forceinline__ void top2_32(const uint32_t (&v)[32], int& h, int& l) {
  int hi[16], lo[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    hi[c] = max((int)v[2 * c], (int)v[2 * c + 1]);
    lo[c] = min((int)v[2 * c], (int)v[2 * c + 1]);
  }
#pragma unroll
  for (int w = 8; w >= 1; w >>= 1) {
#pragma unroll
    for (int c = 0; c < w; ++c) {
      const int a1 = hi[c], b1 = hi[c + w];
      lo[c] = smax3(min(a1, b1), lo[c], lo[c + w]);
      hi[c] = max(a1, b1);
    }
  }
  h = hi[0];
  l = lo[0];
}
// merge two sorted pairs (h1 >= l1), (h2 >= l2) into the first
__device__ __forceinline__ void merge2(int& h1, int& l1, int h2, int l2) {
  l1 = smax3(min(h1, h2), l1, l2);
  h1 = max(h1, h2);
}
// tile-local key 256 * dot + (255 - j mod 256)  ->  ((dot + 256) << 16) | (65535 - j): signed max
// over all tiles = smallest distance, lowest column.  Padded columns stay negative.
__device__ __forceinline__ int to_global(int m, int cbase) {
  const int kp = max(m, -(1 << 22)) + 65536;   // (clamp: an empty slot must not overflow below)
  return (kp >> 8) * 65536 + (kp & 255) + cbase;
}
// -> (distance << 16) | index, the key format of match_compact_kernel; 0xffffffff = none
__device__ __forceinline__ uint32_t out_key(int g) {
  if (g < 0) return 0xffffffffu;
  const uint32_t h = (512u - ((uint32_t)g >> 16)) >> 1;
  return (h << 16) | (65535u - ((uint32_t)g & 65535u));
}

// DBG (compile time; a run-time switch around the TMEM loads would cost time by itself): timing
// experiments whose results are wrong: 2 = the epilogue skips the ranking, 6 = and the TMEM loads.
template <bool TOP2, int DBG = 0>
__global__ void __launch_bounds__(THREADS, 1)
match_kernel(const uint8_t* __restrict__ tiles, const int32_t* __restrict__ nkp,
             const int32_t* __restrict__ pairs, int n_cap, int ntile_cap, int ndir, long n_items,
             uint32_t* __restrict__ rowbest /*[P][n_cap][2]*/, uint32_t* __restrict__ colmin /*[P][n_cap]*/) {
  extern __shared__ uint8_t smem_raw[];
  Smem& S = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nrbp = ntile_cap;

  if (threadIdx.x == 0) {
    mbar_init(&S.full_a, 1);
    mbar_init(&S.empty_a, 1);
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(&S.full_b[s], 1); mbar_init(&S.empty_b[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&S.tmem_full[s], 1); mbar_init(&S.tmem_empty[s], EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == EPI_WARPS) tc::tmem_alloc(&S.tmem_base, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = S.tmem_base;
  constexpr int dbg = DBG;

  if (warp == EPI_WARPS + 1) {
    // ===== bulk-copy producer: the row operand once per item, 72 KB per column tile =====
    if (lane == 0) {
      uint32_t g = 0, n_it = 0;
      for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.empty_a, (n_it & 1) ^ 1);       // MMAs of the previous item are done with A
        mbar_expect_tx(&S.full_a, OPER_BYTES);
        bulk_g2s(S.A, tiles + ((size_t)it.fq * ntile_cap + it.rbp) * TILE_BYTES, OPER_BYTES, &S.full_a);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % NSTAGE;
          mbar_wait(&S.empty_b[st], ((g / NSTAGE) & 1) ^ 1);
          mbar_expect_tx(&S.full_b[st], OPER_BYTES);
          const uint8_t* src = tiles + ((size_t)it.ft * ntile_cap + t) * TILE_BYTES;
          bulk_g2s(S.B[st], src, 16 * KG_BYTES, &S.full_b[st]);                           // bits
          bulk_g2s(S.B[st] + 16 * KG_BYTES, src + 18 * KG_BYTES, 2 * KG_BYTES, &S.full_b[st]);  // index group
        }
        ++n_it;
      }
    }
  } else if (warp == EPI_WARPS) {
    // ===== MMA issuer: one thread, 9 x (128 x 256 x 32) per row block and column tile =====
    if (lane == 0) {
      const uint32_t idesc = idesc_s8_s32(128, ROWS);
      uint32_t g = 0, n_it = 0, use[2] = {0u, 0u};
      for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.full_a, n_it & 1);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % NSTAGE;
          mbar_wait(&S.full_b[st], (g / NSTAGE) & 1);
          const uint32_t b_addr = smem_u32(S.B[st]);
#pragma unroll
          for (int rb = 0; rb < 2; ++rb) {
            if (rb >= it.nrb) break;
            mbar_wait(&S.tmem_empty[rb], (use[rb] & 1) ^ 1);   // the epilogue has drained accumulator rb
            tc::fence_after_sync();
            const uint32_t a_addr = smem_u32(S.A) + rb * 2048;   // rows 128..255 = row groups 16..31
            const uint32_t d_addr = tmem + rb * ROWS;
#pragma unroll
            for (int k = 0; k < 9; ++k)   // k < 8: the bits; k = 8: ones group x index group
              mma_i8(d_addr, tc::smem_desc(a_addr + k * 2 * LBO, LBO, SBO),
                     tc::smem_desc(b_addr + k * 2 * LBO, LBO, SBO), idesc, k > 0);
            tc::commit(&S.tmem_full[rb]);
            ++use[rb];
          }
          tc::commit(&S.empty_b[st]);       // stage reusable once these MMAs retire
        }
        tc::commit(&S.empty_a);
        ++n_it;
      }
    }
  } else {
    // ===== epilogue warps: thread = TMEM lane = row, 128 of the 256 columns per warp =====
    const int quarter = warp & 3, half = warp >> 2;
    const int rl = quarter * 32 + lane;
    uint32_t use[2] = {0u, 0u}, n_it = 0;
    for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
      const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
      if (it.ntile == 0) continue;
      int b1[2] = {KEY_MININT, KEY_MININT}, b2[2] = {KEY_MININT, KEY_MININT};
      for (int t = 0; t < it.ntile; ++t) {
        const int cbase = 65535 - 255 - t * ROWS;
#pragma unroll
        for (int rb = 0; rb < 2; ++rb) {
          if (rb >= it.nrb) break;
          mbar_wait(&S.tmem_full[rb], use[rb] & 1);
          ++use[rb];
          tc::fence_after_sync();
          const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + rb * ROWS + half * 128;
          uint32_t va[32], vb[32];
          int h = KEY_MININT, l = KEY_MININT;
          auto rank = [&](const uint32_t (&v)[32]) {
            if constexpr ((dbg & 2) != 0) return;
            if (TOP2) {
              int ch, cl;
              top2_32(v, ch, cl);
              merge2(h, l, ch, cl);
            } else {
              h = max(h, max32(v));
            }
          };
          if constexpr ((dbg & 4) != 0) {
#pragma unroll
            for (int c = 0; c < 32; ++c) va[c] = vb[c] = (uint32_t)(c + lane);
            rank(va);
            rank(vb);
            rank(va);
          } else {
            tc::ld_32x32_issue(taddr, va);
            tc::ld_wait(va);
            tc::ld_32x32_issue(taddr + 32, vb);
            rank(va);
            tc::ld_wait(vb);
            tc::ld_32x32_issue(taddr + 64, va);
            rank(vb);
            tc::ld_wait(va);
            tc::ld_32x32_issue(taddr + 96, vb);
            rank(va);
            tc::ld_wait(vb);
          }
          tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(&S.tmem_empty[rb]);   // accumulator rb is in registers
          rank(vb);
          if (TOP2) merge2(b1[rb], b2[rb], to_global(h, cbase), to_global(l, cbase));
          else b1[rb] = max(b1[rb], to_global(h, cbase));
        }
      }
      // ---- merge the two column halves of every row, write ----
      if (half == 1) {
#pragma unroll
        for (int rb = 0; rb < 2; ++rb) S.part[n_it & 1][rb][rl] = make_int2(b1[rb], b2[rb]);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");   // the 8 epilogue warps only
      if (half == 0) {
#pragma unroll
        for (int rb = 0; rb < 2; ++rb) {
          if (rb >= it.nrb) break;
          const int2 o = S.part[n_it & 1][rb][rl];
          int h = b1[rb], l = b2[rb];
          merge2(h, l, o.x, o.y);
          const int row = it.rbp * ROWS + rb * 128 + rl;
          if (row < it.nq) {
            if (it.dir == 0)
              reinterpret_cast<uint2*>(rowbest)[(size_t)it.pair * n_cap + row] =
                  make_uint2(out_key(h), TOP2 ? out_key(l) : 0xffffffffu);
            else
              colmin[(size_t)it.pair * n_cap + row] = out_key(h);
          }
        }
      }
      ++n_it;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == EPI_WARPS) tc::tmem_dealloc(tmem, 512);
}

struct Ws {
  uint8_t* tiles;      // [F][ntile][TILE_BYTES]
  uint32_t* rowbest;   // [P][n_cap][2]
  uint32_t* colmin;    // [P][n_cap]
  size_t bytes;
};
static inline Ws carve(void* base, int n_frames, int n_pairs, int n_cap) {
  Ws w;
  const size_t ntile = (n_cap + ROWS - 1) / ROWS;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += (bytes + 1023) & ~(size_t)1023;
    return p;
  };
  w.tiles = (uint8_t*)take((size_t)n_frames * ntile * TILE_BYTES);
  w.rowbest = (uint32_t*)take((size_t)n_pairs * n_cap * 2 * sizeof(uint32_t));
  w.colmin = (uint32_t*)take((size_t)n_pairs * n_cap * sizeof(uint32_t));
  w.bytes = off;
  return w;
}

}  // namespace htc
}  // namespace vo

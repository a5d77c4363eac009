// hamming_f4.cuh — K1f: 256-bit Hamming matching as an exact fp4 (e2m1) contraction on tcgen05.
//
// Same reference call as hamming.cu / hamming_tc.cuh (cv2.BFMatcher(NORM_HAMMING,
// crossCheck=True).match, reference src/image_pair.py:234 with the matcher of
// src/image_and_keypoints.py:9; knn(k=2) + ratio of src/feature_detection.py:20-26).
//
// The bits of a descriptor written as +-6 (e2m1 0x7 / 0xF): the dot product of two descriptors
// is 36 * (256 - 2 * Hamming), every partial sum is an integer below 2^24 and therefore exact in
// the fp32 accumulator, and kind::mxf4 moves twice the K per tensor clock of kind::i8.  The block
// scales of kind::mxf4 are all 1.0 (ue8m0 0x7F): the scale-factor columns of TMEM are filled once
// per CTA with tcgen05.st, so no scale-factor layout is involved.
//
// A fifth K = 64 step multiplies a "ones" group of the row operand with an "index" group of the
// column operand (four base-4 digits of x = 207 - j as e2m1 values): the accumulator holds
//     key(i, j) = 36 * dot(i, j) + 0.25 * (207 - j mod 208)
// (one Hamming step = 72 > 0.25 * 207), so the epilogue is one fp32 max per element (FMNMX3): the
// largest key of a row is its nearest neighbour, lowest column on ties = OpenCV's order.  Padded
// columns (beyond the frame's key points) are all-zero operands (key 0); the epilogue masks them
// in the last column tile only.
//
// The ROW operand lives in TMEM (tcgen05.mma [d], [a_tmem], b_desc): it is constant over a work
// item, so the bulk-copy producer never waits for it at an item boundary and every instruction reads
// only the column tile (N x 32 bytes) from shared memory.  Measured per MMA with the epilogue
// removed: 156 clk with both operands in shared memory (N = 240), 139 clk with this form (floor
// 104 clk); of the rest ~80 clk per accumulator are the hand-off between the MMA issuer and the
// epilogue warps that two accumulators do not cover (profiles/r2t_k1f_one_pass_experiment_negative.log).
//
// TMEM (512 columns): two 128 x 208 fp32 accumulators (0..207, 208..415), the row operand of the
// two row blocks (416..455, 456..495: 40 columns = 320 fp4 values per row) and 16 columns of scale
// factors (496..511); hence column tiles of 208 rows.
// Operand forms:
//   rows    : [F][rows][128 bytes] row-major, 2 bits-as-nibbles per byte in descriptor bit order;
//             a loader thread (= TMEM lane) copies its row into TMEM with tcgen05.st
//   columns : 208-row tiles, UMMA canonical K-major no-swizzle layout, 16-byte K groups (32 values):
//             byte(r, kg, kk) = kg * 3328 + r * 16 + kk, kg 0..7 bits, 8..9 index group
// Kernel: persistent, warp-specialised: warps 0..7 epilogue, warp 8 MMA issuer (one thread),
// warp 9 bulk-copy producer, warps 10..13 row-operand loaders; a work item = (pair, direction,
// 256-row block); every column tile feeds two M = 128 row blocks (5 x tcgen05.mma 128x208x64).
#pragma once
#include "tc.cuh"

namespace vo {
namespace hf4 {

constexpr int ROWS_A = 256, ROWS_B = 208;
constexpr int A_ROW_BYTES = 128;                            // row operand in global memory
constexpr int KG_B = ROWS_B * 16;                           // one 16-byte K group of all rows
constexpr int TILE_B = 10 * KG_B;                           // 33280 bytes
constexpr int NSTAGE = 5;
constexpr int EPI_WARPS = 8;                                // warp & 3 = TMEM lane quarter, warp >> 2 = accumulator
constexpr int THREADS = (EPI_WARPS + 2 + 4) * 32;
constexpr int A_COL = 416, A_COLS = 40, SFA_COL = 496, SFB_COL = 504;
constexpr int KEY_MININT = (int)0x80000000;
#define HF4_NEG (-3.0e38f)

// kind::mxf4: A, B e2m1 (K-major), scales ue8m0, D = fp32, M x N, K = 64 per instruction
__host__ __device__ constexpr uint32_t idesc_mxf4(int M, int N) {
  return (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (1u << 23) | ((uint32_t)(M >> 4) << 24);
}
// A from TMEM (lane = row, 8 fp4 values per 32-bit column), B from shared memory
__device__ __forceinline__ void mma_mxf4(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc,
                                         uint32_t idesc, uint32_t accumulate, uint32_t sfa, uint32_t sfb) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], [%1], %2, %3, [%5], [%6], p;\n"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(sfa), "r"(sfb)
      : "memory");
}
// 32 TMEM lanes x 16 columns <- one value (the scale factors: every byte 0x7F = 2^0)
__device__ __forceinline__ void st_32x16_fill(uint32_t taddr, uint32_t v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};\n"
      "tcgen05.wait::st.sync.aligned;" ::"r"(taddr), "r"(v)
      : "memory");
}
__device__ __forceinline__ void ld_32x16_issue(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
// 32 TMEM lanes x 32 columns <- 32 registers per thread (thread = lane)
__device__ __forceinline__ void st_32x32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};" ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
               : "memory");
}
__device__ __forceinline__ void st_32x8(uint32_t taddr, uint32_t r0) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %2, %2, %2, %2, %2, %2};" ::"r"(taddr),
               "r"(r0), "r"(0u)
               : "memory");
}
__device__ __forceinline__ void st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ------------------------------------------------------------------ prep: bits -> e2m1 operand tiles
// eight bits -> eight nibbles, 0x7 (+6) for a set bit and 0xF (-6) for a clear one
__device__ __forceinline__ uint32_t spread8(uint32_t b8) {
  uint32_t x = b8;
  x = (x | (x << 12)) & 0x000F000Fu;
  x = (x | (x << 6)) & 0x03030303u;
  x = (x | (x << 3)) & 0x11111111u;
  return 0xFFFFFFFFu ^ (x << 3);
}

// thread = (row, 32-bit word of the descriptor): 32 rows per block; every store of a warp covers
// whole 128-byte lines (row operand) or 64 consecutive bytes per K group (column tiles)
static __global__ void __launch_bounds__(256)
expand_kernel(const uint8_t* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
              int rows_a, int ntile_b, uint8_t* __restrict__ a_rows, uint8_t* __restrict__ tiles_b) {
  const int f = blockIdx.y;
  const int row = blockIdx.x * 32 + (threadIdx.x >> 3), kg = threadIdx.x & 7;
  const bool ok = row < min(nkp[f], n_cap);
  const bool in_a = row < rows_a, in_b = row < ntile_b * ROWS_B;
  uint4 o = make_uint4(0u, 0u, 0u, 0u);   // padded rows: dot = 0 with everything
  if (ok) {
    const uint32_t w = __ldg(reinterpret_cast<const uint32_t*>(desc + ((size_t)f * n_cap + row) * 32) + kg);
    o = make_uint4(spread8(w & 255u), spread8((w >> 8) & 255u), spread8((w >> 16) & 255u), spread8(w >> 24));
  }
  uint8_t* tb = tiles_b + ((size_t)f * ntile_b + row / ROWS_B) * TILE_B + (size_t)(row % ROWS_B) * 16;
  if (in_a) *reinterpret_cast<uint4*>(a_rows + ((size_t)f * rows_a + row) * A_ROW_BYTES + kg * 16) = o;
  if (in_b) *reinterpret_cast<uint4*>(tb + (size_t)kg * KG_B) = o;
  if (in_b && kg < 2) {
    // index group (K groups 8, 9): x = 207 - r = 64 d3 + 16 d2 + 4 d1 + d0; slots 0..3 = d3, 4 = d2,
    // 5 = d1 (values 0, 1, 2, 3 = e2m1 codes 0, 2, 4, 5), slot 6 = d0 / 2 (code d0); against the
    // ones group (ONES_WORD: slots 0..4 = 4.0, slot 5 = 1.0, slot 6 = 0.5) of the row operand:
    //   4 * (4 * d3) + 4 * d2 + 1 * d1 + 0.5 * (d0 / 2) = x / 4
    const uint32_t x = (uint32_t)(ROWS_B - 1 - row % ROWS_B);
    const uint32_t d3 = x >> 6, d2 = (x >> 4) & 3u, d1 = (x >> 2) & 3u, d0 = x & 3u;
    const uint32_t c3 = 2u * d3 - (d3 == 3u), c2 = 2u * d2 - (d2 == 3u), c1 = 2u * d1 - (d1 == 3u);
    const uint32_t word = (ok && kg == 0) ? (c3 * 0x1111u) | (c2 << 16) | (c1 << 20) | (d0 << 24) : 0u;
    *reinterpret_cast<uint4*>(tb + (size_t)(8 + kg) * KG_B) = make_uint4(word, 0u, 0u, 0u);
  }
}
constexpr uint32_t ONES_WORD = 0x01266666u;

// ------------------------------------------------------------------ the contraction + ranking
struct __align__(1024) Smem {
  uint8_t B[NSTAGE][TILE_B];
  uint64_t full_a, empty_a;
  uint64_t full_b[NSTAGE], empty_b[NSTAGE];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

template <bool V> struct MaskedTag { static constexpr bool value = V; };

struct Item {
  int pair, dir, rbp;      // rbp = 256-row block of the row-side frame
  int fq, ft, nq, nt;      // row-side / column-side frame and their key-point counts
  int ntile, nrb;          // column tiles (of 208), valid 128-row blocks (0..2)
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
  it.ntile = (it.nt + ROWS_B - 1) / ROWS_B;
  const int r0 = it.rbp * ROWS_A;
  it.nrb = r0 >= it.nq ? 0 : (r0 + 128 < it.nq ? 2 : 1);
  if (it.nrb == 0) it.ntile = 0;
  if (it.ntile == 0) it.nrb = 0;
  return it;
}

__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float r;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
  return r;
}
__device__ __forceinline__ int imax3(int a, int b, int c) { return __vimax3_s32(a, b, c); }

// largest of 32 keys: 3-input max tree (16 instructions)
__device__ __forceinline__ float max32(const uint32_t (&u)[32]) {
  float a[11];
#pragma unroll
  for (int c = 0; c < 10; ++c)
    a[c] = fmax3(__uint_as_float(u[3 * c]), __uint_as_float(u[3 * c + 1]), __uint_as_float(u[3 * c + 2]));
  a[10] = fmaxf(__uint_as_float(u[30]), __uint_as_float(u[31]));
  const float b0 = fmax3(a[0], a[1], a[2]), b1 = fmax3(a[3], a[4], a[5]), b2 = fmax3(a[6], a[7], a[8]),
              b3 = fmaxf(a[9], a[10]);
  return fmaxf(fmax3(b0, b1, b2), b3);
}
// two largest of 32 keys (the real keys of a row are distinct): tournament of sorted pairs
__device__ __forceinline__ void top2_32(const uint32_t (&u)[32], float& h, float& l) {
  float hi[16], lo[16];
#pragma unroll
  for (int c = 0; c < 16; ++c) {
    const float x = __uint_as_float(u[2 * c]), y = __uint_as_float(u[2 * c + 1]);
    hi[c] = fmaxf(x, y);
    lo[c] = fminf(x, y);
  }
#pragma unroll
  for (int w = 8; w >= 1; w >>= 1) {
#pragma unroll
    for (int c = 0; c < w; ++c) {
      const float a1 = hi[c], b1 = hi[c + w];
      lo[c] = fmax3(fminf(a1, b1), lo[c], lo[c + w]);
      hi[c] = fmaxf(a1, b1);
    }
  }
  h = hi[0];
  l = lo[0];
}
__device__ __forceinline__ void merge2f(float& h1, float& l1, float h2, float l2) {
  l1 = fmax3(fminf(h1, h2), l1, l2);
  h1 = fmaxf(h1, h2);
}
__device__ __forceinline__ void merge2i(int& h1, int& l1, int h2, int l2) {
  l1 = imax3(min(h1, h2), l1, l2);
  h1 = max(h1, h2);
}
// tile-local key 36 * dot + (207 - j) / 4  ->  ((256 - H) << 16) | (65535 - column): signed max over
// all tiles = smallest distance, lowest column.  4 * key + 36864 = 288 * (256 - H) + (207 - j).
__device__ __forceinline__ int to_global(float m, int tile) {
  if (m < -1.0e6f) return KEY_MININT;          // empty slot / masked column
  const int y = (int)(m * 4.0f) + 36864;
  const int u = y / 288, x = y - 288 * u;
  return (u << 16) | (65535 - (tile * ROWS_B + (ROWS_B - 1 - x)));
}
// -> (distance << 16) | index, the key format of match_compact_kernel; 0xffffffff = none
__device__ __forceinline__ uint32_t out_key(int g) {
  if (g < 0) return 0xffffffffu;
  const uint32_t h = 256u - ((uint32_t)g >> 16);
  return (h << 16) | (65535u - ((uint32_t)g & 65535u));
}

// PROBE: one item, column tile 0, row block 0: the raw accumulator goes to `acc` [128][208].
// DBG (compile time, timing experiments whose results are wrong): 1 = no TMEM loads, no ranking;
// 2 = the loads, no ranking
template <bool TOP2, bool PROBE = false, int DBG = 0>
__global__ void __launch_bounds__(THREADS, 1)
match_kernel(const uint8_t* __restrict__ a_rows, const uint8_t* __restrict__ tiles_b,
             const int32_t* __restrict__ nkp, const int32_t* __restrict__ pairs, int n_cap,
             int rows_a, int ntile_b, int ndir, long n_items,
             uint32_t* __restrict__ rowbest /*[P][n_cap][2]*/, uint32_t* __restrict__ colmin /*[P][n_cap]*/,
             float* __restrict__ acc) {
  extern __shared__ uint8_t smem_raw[];
  Smem& S = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nrbp = rows_a / ROWS_A;

  if (threadIdx.x == 0) {
    mbar_init(&S.full_a, 4);       // the four loader warps
    mbar_init(&S.empty_a, 1);
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(&S.full_b[s], 1); mbar_init(&S.empty_b[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&S.tmem_full[s], 1); mbar_init(&S.tmem_empty[s], EPI_WARPS / 2); }
    fence_barrier_init();
  }
  if (warp == EPI_WARPS) tc::tmem_alloc(&S.tmem_base, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = S.tmem_base;
  // scale factors: every byte of columns 496..511 on all 128 lanes = 0x7F (ue8m0 2^0)
  if (warp < 4) st_32x16_fill(tmem + ((uint32_t)(warp * 32) << 16) + SFA_COL, 0x7F7F7F7Fu);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();

  if (warp >= EPI_WARPS + 2) {
    // ===== row-operand loaders: thread = TMEM lane = row; 128 bytes of bits + the ones group =====
    const int quarter = warp & 3;
    const uint32_t lanes = (uint32_t)(quarter * 32) << 16;
    uint32_t n_it = 0;
    for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
      const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
      if (it.ntile == 0) continue;
      uint32_t w[2][32];
#pragma unroll
      for (int rb = 0; rb < 2; ++rb) {
        const uint4* src = reinterpret_cast<const uint4*>(
            a_rows + ((size_t)it.fq * rows_a + it.rbp * ROWS_A + rb * 128 + quarter * 32 + lane) * A_ROW_BYTES);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 v = __ldg(src + c);
          w[rb][4 * c] = v.x; w[rb][4 * c + 1] = v.y; w[rb][4 * c + 2] = v.z; w[rb][4 * c + 3] = v.w;
        }
      }
      mbar_wait(&S.empty_a, (n_it & 1) ^ 1);       // the MMAs of the previous item are done with it
      tc::fence_after_sync();
#pragma unroll
      for (int rb = 0; rb < 2; ++rb) {
        st_32x32(tmem + lanes + A_COL + rb * A_COLS, w[rb]);
        st_32x8(tmem + lanes + A_COL + rb * A_COLS + 32, ONES_WORD);
      }
      st_wait();
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&S.full_a);
      ++n_it;
    }
  } else if (warp == EPI_WARPS + 1) {
    // ===== bulk-copy producer: 32.5 KB per column tile =====
    if (lane == 0) {
      uint32_t g = 0;
      for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % NSTAGE;
          mbar_wait(&S.empty_b[st], ((g / NSTAGE) & 1) ^ 1);
          mbar_expect_tx(&S.full_b[st], TILE_B);
          bulk_g2s(S.B[st], tiles_b + ((size_t)it.ft * ntile_b + t) * TILE_B, TILE_B, &S.full_b[st]);
        }
      }
    }
  } else if (warp == EPI_WARPS) {
    // ===== MMA issuer: one thread, 5 x (128 x 208 x 64) per row block and column tile =====
    if (lane == 0) {
      const uint32_t idesc = idesc_mxf4(128, ROWS_B);
      const uint32_t sfa = tmem + SFA_COL, sfb = tmem + SFB_COL;
      uint32_t g = 0, n_it = 0, use[2] = {0u, 0u};
      for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
        if (it.ntile == 0) continue;
        mbar_wait(&S.full_a, n_it & 1);
        tc::fence_after_sync();
        for (int t = 0; t < it.ntile; ++t, ++g) {
          const int st = g % NSTAGE;
          mbar_wait(&S.full_b[st], (g / NSTAGE) & 1);
          const uint32_t b_addr = smem_u32(S.B[st]);
#pragma unroll
          for (int rb = 0; rb < 2; ++rb) {
            if (rb >= it.nrb) break;
            mbar_wait(&S.tmem_empty[rb], (use[rb] & 1) ^ 1);   // the epilogue has drained accumulator rb
            tc::fence_after_sync();
            const uint32_t a_addr = tmem + A_COL + rb * A_COLS;
            const uint32_t d_addr = tmem + rb * ROWS_B;
#pragma unroll
            for (int k = 0; k < 5; ++k)   // k < 4: the bits; k = 4: ones group x index group
              mma_mxf4(d_addr, a_addr + k * 8, tc::smem_desc(b_addr + k * 2 * KG_B, KG_B, 128), idesc, k > 0,
                       sfa, sfb);
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
    // ===== epilogue warps: thread = TMEM lane = row; warps 0..3 drain accumulator 0 (row block 0),
    // warps 4..7 accumulator 1: the two groups run independently, so the latency chain of one
    // accumulator (barrier, TMEM loads, ranking) overlaps the other's =====
    const int quarter = warp & 3, rb = warp >> 2;
    const int rl = quarter * 32 + lane;
    constexpr int NFULL = ROWS_B / 32, REM = ROWS_B % 32, NCH = NFULL + (REM ? 1 : 0);
    static_assert(REM == 0 || REM == 16, "column tile = 32-column chunks + an optional 16-column one");
    uint32_t use = 0;
    for (long item = blockIdx.x; item < n_items; item += gridDim.x) {
      const Item it = get_item(item, nkp, pairs, n_cap, nrbp, ndir);
      if (rb >= it.nrb) continue;
      int b1 = KEY_MININT, b2 = KEY_MININT;
      for (int t = 0; t < it.ntile; ++t) {
        const int nvalid = min(it.nt - t * ROWS_B, ROWS_B);   // valid columns of this tile
        mbar_wait(&S.tmem_full[rb], use & 1);
        ++use;
        tc::fence_after_sync();
        const uint32_t taddr = tmem + ((uint32_t)(quarter * 32) << 16) + rb * ROWS_B;
        float h = HF4_NEG, l = HF4_NEG;
        // MASKED: the last column tile of a frame whose key-point count is no multiple of the tile
        // (a separate instantiation: the selects must not run on the other tiles)
        auto drain = [&](auto masked_tag) {
          constexpr bool MASKED = decltype(masked_tag)::value;
          uint32_t v[2][32];
          auto rank = [&](uint32_t (&x)[32], int c0) {
            if constexpr (DBG != 0) return;
            if constexpr (MASKED) {
#pragma unroll
              for (int c = 0; c < 32; ++c)
                if (c0 + c >= nvalid) x[c] = __float_as_uint(HF4_NEG);
            }
            if constexpr (PROBE) {
              if (rb == 0 && t == 0) {
#pragma unroll
                for (int c = 0; c < 32; ++c)
                  if (c0 + c < ROWS_B) acc[rl * ROWS_B + c0 + c] = __uint_as_float(x[c]);
              }
            }
            if (TOP2) {
              float ch, cl;
              top2_32(x, ch, cl);
              merge2f(h, l, ch, cl);
            } else {
              h = fmaxf(h, max32(x));
            }
          };
          if constexpr (DBG != 1) {
            tc::ld_32x32_issue(taddr, v[0]);
            tc::ld_wait(v[0]);
          }
#pragma unroll
          for (int c = 0; c < NCH; ++c) {
            if constexpr (DBG != 1) {
              if (c + 1 < NFULL) tc::ld_32x32_issue(taddr + 32 * (c + 1), v[(c + 1) & 1]);
              else if (c + 1 < NCH) ld_32x16_issue(taddr + 32 * (c + 1), v[(c + 1) & 1]);
            }
            if (c == NCH - 1) {       // every load has been waited for: the accumulator is in registers
              tc::fence_before_sync();
              __syncwarp();
              if (lane == 0) tc::mbar_arrive(&S.tmem_empty[rb]);
            }
            rank(v[c & 1], 32 * c);
            if constexpr (DBG != 1) {
              if (c + 1 < NCH) {
                tc::ld_wait(v[(c + 1) & 1]);
                if (c + 1 >= NFULL) {
#pragma unroll
                  for (int e = 16; e < 32; ++e) v[(c + 1) & 1][e] = __float_as_uint(HF4_NEG);   // (constants: folded)
                }
              }
            }
          }
        };
        if (nvalid < ROWS_B) drain(MaskedTag<true>{});
        else drain(MaskedTag<false>{});
        if (TOP2) merge2i(b1, b2, to_global(h, t), to_global(l, t));
        else b1 = max(b1, to_global(h, t));
      }
      const int row = it.rbp * ROWS_A + rb * 128 + rl;
      if (row < it.nq) {
        if (it.dir == 0)
          reinterpret_cast<uint2*>(rowbest)[(size_t)it.pair * n_cap + row] =
              make_uint2(out_key(b1), TOP2 ? out_key(b2) : 0xffffffffu);
        else
          colmin[(size_t)it.pair * n_cap + row] = out_key(b1);
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == EPI_WARPS) tc::tmem_dealloc(tmem, 512);
}

struct Ws {
  uint8_t* a_rows;     // [F][rows_a][128]
  uint8_t* tiles_b;    // [F][ntile_b][TILE_B]
  uint32_t* rowbest;   // [P][n_cap][2]
  uint32_t* colmin;    // [P][n_cap]
  size_t bytes;
};
static inline Ws carve(void* base, int n_frames, int n_pairs, int n_cap) {
  Ws w;
  const size_t rows_a = (size_t)((n_cap + ROWS_A - 1) / ROWS_A) * ROWS_A, ntb = (n_cap + ROWS_B - 1) / ROWS_B;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? (char*)base + off : nullptr;
    off += (bytes + 1023) & ~(size_t)1023;
    return p;
  };
  w.a_rows = (uint8_t*)take((size_t)n_frames * rows_a * A_ROW_BYTES);
  w.tiles_b = (uint8_t*)take((size_t)n_frames * ntb * TILE_B);
  w.rowbest = (uint32_t*)take((size_t)n_pairs * n_cap * 2 * sizeof(uint32_t));
  w.colmin = (uint32_t*)take((size_t)n_pairs * n_cap * sizeof(uint32_t));
  w.bytes = off;
  return w;
}

}  // namespace hf4
}  // namespace vo

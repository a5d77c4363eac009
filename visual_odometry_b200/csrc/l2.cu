// l2.cu — K6: SIFT / NORM_L2 descriptor matching on the tcgen05 tensor cores.
//
// Replaces cv2.BFMatcher(NORM_L2, crossCheck=True).match (reference src/visual_slam.py:19,
// src/image_pair.py:234) and knn(k=2)+ratio (src/feature_detection.py:20-26).
//   d^2(i,j) = |a_i|^2 + |b_j|^2 - 2 a_i.b_j : the a.b term is the one dense contraction of the
// hot path, so it runs as a bf16 GEMM (tcgen05.mma, fp32 accumulate in TMEM).  cv2-SIFT
// descriptors are integers in [0,255]: exact in bf16, dot products < 2^24 exact in fp32, so the
// GEMM is EXACT for real SIFT data; other float descriptors get an fp32 rerank of the candidates.
//
// Pipeline per launch:
//   l2_prep_kernel   fp32 [F,n,128] -> bf16 tiles already in the UMMA canonical K-major
//                    no-swizzle layout (one 128x128 tile = one contiguous 32 KB block, so a tile
//                    is ONE 1-D TMA bulk copy, no tensor map) + exact fp32 squared norms.
//   l2_top2_tc_kernel  warp-specialised: warp 5 = TMA producer, warp 4 = single-thread MMA
//                    issuer, warps 0-3 = epilogue (thread = TMEM lane = query row) keeping a
//                    running top-2 of (d^2, index) straight out of TMEM.  Runs in both directions
//                    (query->train, train->query) so the cross-check needs no column reduction.
//   l2_compact_kernel  merges split partials, cross-check / ratio test, sqrtf, stable compaction,
//                    keypoint gather.
#include <cuda_bf16.h>

#include "tc.cuh"

namespace vo {

constexpr int L2_DIM = 128;
constexpr int L2_BLK = 128;                       // rows per tile (UMMA M and N)
// a tile = 20 K-groups (kg) of 8 bf16 columns x 128 rows: kg 0..15 the descriptor, kg 16 = [1,1,1,0..]
// (A-side extras), kg 17 = 0, kg 18 = -|b|^2/2 split into three bf16 (B-side extras), kg 19 = 0.
// A ninth K=16 MMA step over (kg16,kg17) x (kg18,kg19) adds -|b_j|^2/2 to the accumulator, so the
// epilogue ranks  a.b - |b|^2/2  directly: no norm loads, no FMA per element.
constexpr int L2_KG = 20;
constexpr int L2_TILE_ELEMS = L2_BLK * 8 * L2_KG;  // 20480 bf16 = 40 KB
constexpr int L2_TILE_BYTES = L2_TILE_ELEMS * 2;
constexpr uint32_t L2_LBO = 2048;                 // between the two K core matrices of a K=16 slice
constexpr uint32_t L2_SBO = 128;                  // between 8-row groups
constexpr int L2_NSTAGE = 3;
constexpr int L2_EPI_WARPS = 16;  // 4 per TMEM lane quarter, 32 columns each
constexpr int L2_THREADS = (L2_EPI_WARPS + 2) * 32;  // + MMA warp + TMA warp

struct L2Ws {
  __nv_bfloat16* tiles;   // [F][nblk][16384]
  float* norms;           // [F][nblk*128]   (+inf on padded rows)
  int32_t* inexact;       // [1] set when some value is not exactly representable in bf16
  float4* rowbest;        // [P][2][S][nblk*128]  (d2_1, idx_1, d2_2, idx_2) idx as int bits
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
  w.inexact = (int32_t*)take(256);
  w.rowbest = (float4*)take((size_t)n_pairs * 2 * splits * nblk * L2_BLK * sizeof(float4));
  w.bytes = off;
  return w;
}

// ------------------------------------------------------------------ prep: fp32 -> bf16 tiles
// tile layout (bytes): kg*2048 + rg*128 + rr*16 + kk*2  with r = 8*rg+rr, k = 8*kg+kk
__global__ void __launch_bounds__(256)
l2_prep_kernel(const float* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
               int nblk, L2Ws w) {
  const int f = blockIdx.y, blk = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = nkp[f];
  __nv_bfloat16* tile = w.tiles + ((size_t)f * nblk + blk) * L2_TILE_ELEMS;
  bool bad = false;
  for (int rl = warp; rl < L2_BLK; rl += 8) {
    const int r = blk * L2_BLK + rl;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < n) v = __ldg(reinterpret_cast<const float4*>(desc + ((size_t)f * n_cap + r) * L2_DIM) + lane);
    const __nv_bfloat16 b0 = __float2bfloat16_rn(v.x), b1 = __float2bfloat16_rn(v.y),
                        b2 = __float2bfloat16_rn(v.z), b3 = __float2bfloat16_rn(v.w);
    bad |= (__bfloat162float(b0) != v.x) | (__bfloat162float(b1) != v.y) |
           (__bfloat162float(b2) != v.z) | (__bfloat162float(b3) != v.w);
    const int k = 4 * lane;
    const size_t byte_off = (size_t)(k >> 3) * 2048 + (size_t)(rl >> 3) * 128 + (rl & 7) * 16 + (k & 7) * 2;
    __nv_bfloat162 lo, hi;
    lo.x = b0; lo.y = b1; hi.x = b2; hi.y = b3;
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(reinterpret_cast<char*>(tile) + byte_off) = pk;
    float ss = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    if (lane == 0)
      w.norms[((size_t)f * nblk + blk) * L2_BLK + rl] = (r < n) ? ss : __int_as_float(0x7f800000);
    if (lane < 4) {  // the four extra K-groups of this row (16 bytes each)
      uint4 e = make_uint4(0, 0, 0, 0);
      if (lane == 0) {
        e.x = 0x3f803f80u;  // bf16 (1, 1)
        e.y = 0x00003f80u;  // bf16 (1, 0)
      } else if (lane == 2) {
        __nv_bfloat16 h, m, l;
        if (r < n) {
          const float x = -0.5f * ss;
          h = __float2bfloat16_rn(x);
          const float r1 = x - __bfloat162float(h);
          m = __float2bfloat16_rn(r1);
          l = __float2bfloat16_rn(r1 - __bfloat162float(m));
        } else {
          h = __float2bfloat16_rn(__int_as_float(0xff800000));  // -inf: padded column never wins
          m = l = __float2bfloat16_rn(0.f);
        }
        e.x = (uint32_t)__bfloat16_as_ushort(h) | ((uint32_t)__bfloat16_as_ushort(m) << 16);
        e.y = (uint32_t)__bfloat16_as_ushort(l);
      }
      *reinterpret_cast<uint4*>(reinterpret_cast<char*>(tile) + (size_t)(16 + lane) * 2048 +
                                (size_t)(rl >> 3) * 128 + (rl & 7) * 16) = e;
    }
  }
  if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(w.inexact, 1);
}

// ------------------------------------------------------------------ tensor-core top-2
// bit 0: producer loads each stage only once (no per-tile TMA); bit 1: epilogue skips the scan;
// bit 2: epilogue skips the TMEM load too.  Diagnostics only (results are wrong when set).
__device__ int g_l2_debug = 0;

struct __align__(1024) L2Smem {
  uint8_t A[L2_TILE_BYTES];
  uint8_t B[L2_NSTAGE][L2_TILE_BYTES];
  float4 part[3][L2_BLK];  // top-2 of column quarters 1..3 of each row, merged by quarter 0
  uint64_t full_a;
  uint64_t full_b[L2_NSTAGE], empty_b[L2_NSTAGE];
  uint64_t tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

// running top-2 of the LARGEST scores; ties keep the earlier (lower) column
__device__ __forceinline__ void top2_insert(float s, int c, float& s1, int& i1, float& s2, int& i2) {
  if (s > s1) { s2 = s1; i2 = i1; s1 = s; i1 = c; }
  else { s2 = s; i2 = c; }
}

// TOP2 = false: cross-check matching needs only the nearest neighbour of every row (in both
// directions), which halves the number of threshold hits and shortens the hit path.
template <bool TOP2>
__global__ void __launch_bounds__(L2_THREADS, 1)
l2_top2_tc_kernel(const int32_t* __restrict__ nkp, const int32_t* __restrict__ pairs, int nblk,
                  int n_splits, L2Ws w) {
  extern __shared__ uint8_t smem_raw[];
  L2Smem& S = *reinterpret_cast<L2Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int pair = blockIdx.y >> 1, dir = blockIdx.y & 1, split = blockIdx.z;
  const int fq = pairs[2 * pair + dir], ft = pairs[2 * pair + 1 - dir];
  const int nq = nkp[fq], nt = nkp[ft];
  const int rb = blockIdx.x;
  if (rb * L2_BLK >= nq) return;
  const int ntile_all = (nt + L2_BLK - 1) / L2_BLK;
  const int per = (ntile_all + n_splits - 1) / n_splits;
  const int t_lo = split * per, t_hi = min(ntile_all, t_lo + per);
  const int ntile = max(t_hi - t_lo, 0);

  if (threadIdx.x == 0) {
    mbar_init(&S.full_a, 1);
    for (int s = 0; s < L2_NSTAGE; ++s) { mbar_init(&S.full_b[s], 1); mbar_init(&S.empty_b[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&S.tmem_full[s], 1); mbar_init(&S.tmem_empty[s], L2_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == L2_EPI_WARPS) tc::tmem_alloc(&S.tmem_base, 256);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = S.tmem_base;
  const float NEG_INF = __int_as_float(0xff800000);
  const int dbg = g_l2_debug;

  if (warp == L2_EPI_WARPS + 1) {
    // ===== TMA producer: one 40 KB bulk copy per tile =====
    if (lane == 0 && ntile > 0) {
      mbar_expect_tx(&S.full_a, L2_TILE_BYTES);
      bulk_g2s(S.A, w.tiles + ((size_t)fq * nblk + rb) * L2_TILE_ELEMS, L2_TILE_BYTES, &S.full_a);
      for (int t = 0; t < ntile; ++t) {
        const int st = t % L2_NSTAGE;
        mbar_wait(&S.empty_b[st], ((t / L2_NSTAGE) & 1) ^ 1);
        if ((dbg & 1) && t >= L2_NSTAGE) { tc::mbar_arrive(&S.full_b[st]); continue; }
        mbar_expect_tx(&S.full_b[st], L2_TILE_BYTES);
        bulk_g2s(S.B[st], w.tiles + ((size_t)ft * nblk + t_lo + t) * L2_TILE_ELEMS, L2_TILE_BYTES,
                 &S.full_b[st]);
      }
    }
  } else if (warp == L2_EPI_WARPS) {
    // ===== MMA issuer: one thread, 9 x (128x128x16) per tile, accumulator double-buffered =====
    if (lane == 0 && ntile > 0) {
      const uint32_t idesc = tc::idesc_bf16_f32(L2_BLK, L2_BLK);
      mbar_wait(&S.full_a, 0);
      const uint32_t a_addr = smem_u32(S.A);
      for (int t = 0; t < ntile; ++t) {
        const int st = t % L2_NSTAGE, as = t & 1;
        mbar_wait(&S.full_b[st], (t / L2_NSTAGE) & 1);
        mbar_wait(&S.tmem_empty[as], ((t >> 1) & 1) ^ 1);
        tc::fence_after_sync();
        const uint32_t b_addr = smem_u32(S.B[st]);
#pragma unroll
        for (int k = 0; k < 9; ++k) {
          // steps 0..7: descriptor columns; step 8: A extras (kg16,17) x B extras (kg18,19)
          const uint32_t ao = (k < 8) ? k * 2 * L2_LBO : 16 * L2_LBO;
          const uint32_t bo = (k < 8) ? k * 2 * L2_LBO : 18 * L2_LBO;
          tc::mma_bf16(tmem + as * L2_BLK, tc::smem_desc(a_addr + ao, L2_LBO, L2_SBO),
                       tc::smem_desc(b_addr + bo, L2_LBO, L2_SBO), idesc, k > 0);
        }
        tc::commit(&S.empty_b[st]);     // smem stage reusable once these MMAs retire
        tc::commit(&S.tmem_full[as]);   // accumulator ready for the epilogue
      }
    }
  } else {
    // ===== epilogue warps 0..7: thread = TMEM lane = query row, warp/4 = column half ==========
    // running top-2 of  a.b - |b|^2/2  (largest = nearest), 64 columns per TMEM load
    const int quarter = warp & 3, cq = warp >> 2;   // TMEM lane quarter, column quarter
    const int rl = quarter * 32 + lane;
    const int row = rb * L2_BLK + rl;
    float s1 = NEG_INF, s2 = NEG_INF;
    int i1 = -1, i2 = -1;
    for (int t = 0; t < ntile; ++t) {
      const int as = t & 1;
      mbar_wait(&S.tmem_full[as], (t >> 1) & 1);
      tc::fence_after_sync();
      const int c0 = (t_lo + t) * L2_BLK + cq * 32;
      float v[32];
      if (dbg & 4) {
#pragma unroll
        for (int c = 0; c < 32; ++c) v[c] = NEG_INF;
      } else {
        tc::ld_32x32(tmem + ((uint32_t)(quarter * 32) << 16) + as * L2_BLK + cq * 32, v);
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&S.tmem_empty[as]);   // values are in registers now
      if (dbg & 2) continue;
      // group-of-8 prefilter (7 FMNMX + 1 branch).  With 32 independent rows per warp some lane
      // almost always hits, so the hit path is kept short: locate the first occurrence of the
      // group maximum, insert it, knock it out, re-test (a second hit in one group is rare).
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float* x = &v[8 * g];
        float m = fmaxf(fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3])),
                        fmaxf(fmaxf(x[4], x[5]), fmaxf(x[6], x[7])));
        if (TOP2) {
          while (m > s2) {
            int pos = 7;
#pragma unroll
            for (int k = 6; k >= 0; --k) pos = (x[k] == m) ? k : pos;
            top2_insert(m, c0 + 8 * g + pos, s1, i1, s2, i2);
#pragma unroll
            for (int k = 0; k < 8; ++k) x[k] = (k == pos) ? NEG_INF : x[k];
            m = fmaxf(fmaxf(fmaxf(x[0], x[1]), fmaxf(x[2], x[3])),
                      fmaxf(fmaxf(x[4], x[5]), fmaxf(x[6], x[7])));
          }
        } else if (m > s1) {
          int pos = 7;
#pragma unroll
          for (int k = 6; k >= 0; --k) pos = (x[k] == m) ? k : pos;
          s1 = m;
          i1 = c0 + 8 * g + pos;
        }
      }
    }
    if (cq > 0) S.part[cq - 1][rl] = make_float4(s1, __int_as_float(i1), s2, __int_as_float(i2));
    asm volatile("bar.sync 1, 512;" ::: "memory");   // the 16 epilogue warps only
    if (cq == 0) {
      // merge the other column quarters; equal scores keep the lower column index
#pragma unroll
      for (int p = 0; p < 3; ++p) {
        const float4 o = S.part[p][rl];
        const int a = __float_as_int(o.y), b = __float_as_int(o.w);
        if (a >= 0 && (o.x > s2 || (o.x == s2 && (i2 < 0 || a < i2)))) {
          if (o.x > s1 || (o.x == s1 && (i1 < 0 || a < i1))) { s2 = s1; i2 = i1; s1 = o.x; i1 = a; }
          else { s2 = o.x; i2 = a; }
        }
        if (b >= 0 && (o.z > s2 || (o.z == s2 && (i2 < 0 || b < i2)))) {
          if (o.z > s1 || (o.z == s1 && (i1 < 0 || b < i1))) { s2 = s1; i2 = i1; s1 = o.z; i1 = b; }
          else { s2 = o.z; i2 = b; }
        }
      }
      if (row < nq) {
        // d^2 = |a|^2 - 2 (a.b - |b|^2/2); exact for integer descriptors
        const float na = w.norms[(size_t)fq * nblk * L2_BLK + row];
        float4 r;
        r.x = i1 >= 0 ? fmaxf(fmaf(-2.f, s1, na), 0.f) : __int_as_float(0x7f800000);
        r.y = __int_as_float(i1);
        r.z = i2 >= 0 ? fmaxf(fmaf(-2.f, s2, na), 0.f) : __int_as_float(0x7f800000);
        r.w = __int_as_float(i2);
        w.rowbest[(((size_t)pair * 2 + dir) * n_splits + split) * nblk * L2_BLK + row] = r;
      }
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == L2_EPI_WARPS) tc::tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------ fp32 rerank (non-integer data)
// When some descriptor value is not exactly representable in bf16 the GEMM distances are
// approximate: recompute the two candidates of every row exactly in fp32 (same formula as cv2:
// sum of squared differences) and re-order them.  One warp per row.
__global__ void __launch_bounds__(256)
l2_rerank_kernel(const float* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
                 const int32_t* __restrict__ pairs, int nblk, int n_splits, L2Ws w) {
  if (*w.inexact == 0) return;
  const int pair = blockIdx.y >> 1, dir = blockIdx.y & 1;
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
                  L2Ws w, int32_t* __restrict__ q, int32_t* __restrict__ t,
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
      if (mode == VO_MATCH_MUTUAL) {
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

static int l2_pick_splits(int n_pairs, int nblk) {
  const long ctas = (long)n_pairs * 2 * nblk;
  int s = (int)((148L * 2 + ctas - 1) / ctas);
  if (s > nblk) s = nblk;
  if (s > 8) s = 8;
  return s < 1 ? 1 : s;
}

extern "C" int vo_debug_l2(int flags) {
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
  const int splits = l2_pick_splits(n_pairs, nblk);
  L2Ws w = l2_carve(ws, n_frames, n_pairs, n_cap, splits);
  cudaMemsetAsync(w.inexact, 0, sizeof(int32_t), st);
  l2_prep_kernel<<<dim3(nblk, n_frames), 256, 0, st>>>(desc, nkp, n_cap, nblk, w);
  static bool attr_set = false;
  const int smem = (int)sizeof(L2Smem) + 1024;
  if (!attr_set) {
    if (cudaFuncSetAttribute(l2_top2_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             smem) != cudaSuccess ||
        cudaFuncSetAttribute(l2_top2_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             smem) != cudaSuccess)
      return check_launch("vo_match_l2_batch(smem attribute)");
    attr_set = true;
  }
  {
    ProfScope prof(2, st);
    const dim3 grid(nblk, n_pairs * 2, splits);
    if (mode == VO_MATCH_MUTUAL)
      l2_top2_tc_kernel<false><<<grid, L2_THREADS, smem, st>>>(nkp, pairs, nblk, splits, w);
    else
      l2_top2_tc_kernel<true><<<grid, L2_THREADS, smem, st>>>(nkp, pairs, nblk, splits, w);
  }
  l2_rerank_kernel<<<dim3((n_cap + 7) / 8, n_pairs * 2), 256, 0, st>>>(desc, nkp, n_cap, pairs, nblk,
                                                                     splits, w);
  l2_compact_kernel<<<n_pairs, L2C_THREADS, 0, st>>>(nkp, kp, n_cap, pairs, nblk, splits, mode, ratio,
                                                    w, q, t, dist, t2, dist2, pts, n_out);
  return check_launch("vo_match_l2_batch");
}

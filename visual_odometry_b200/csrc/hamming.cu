// hamming.cu — K1: 256-bit ORB descriptor matching, one N1 x N2 popc pass per frame pair.
//
// Replaces cv2.BFMatcher(NORM_HAMMING, crossCheck=True).match (reference src/image_pair.py:234,
// matcher built at src/image_and_keypoints.py:9 / src/visual_slam.py:18) and the brute-force
// knnMatch(k=2)+ratio of src/feature_detection.py:20-26.
//
// Layout: each thread keeps RQ query descriptors (8 x u32 each) in registers; train descriptors
// stream through shared memory in TT-descriptor tiles staged by 1-D TMA bulk copies
// (cp.async.bulk + mbarrier, NSTAGE deep).  Every lane reads the same train descriptor
// (broadcast LDS.128), so one smem read feeds 32*RQ distances.  Row (query) top-2 lives in
// registers as packed keys (dist<<16 | trainIdx): min() on the key is "smallest distance, lowest
// index on ties" = OpenCV's order.  Column (train) minima use keys (dist<<16 | queryIdx):
// per-thread min over RQ -> REDUX warp min -> shared atomicMin -> one global atomicMin per tile
// column, so forward NN, second NN and backward NN all come out of ONE pass (cv2 does two).
#include "common.cuh"
#include "hamming_tc.cuh"
#include "hamming_f4.cuh"

namespace vo {

constexpr int HM_THREADS = 128;
constexpr int HM_TT = 128;     // train descriptors per smem tile (4 KB)
constexpr int HM_NSTAGE = 4;
constexpr uint32_t KEY_INF = 0xffffffffu;


// 256-bit Hamming distance with a carry-save adder tree (Harley-Seal): the popc unit (XU pipe,
// 16 lanes/clk/SM) is the scarce resource, LOP3 (ALU pipe, 64 lanes/clk/SM) is not.  Three
// full adders compress seven of the eight XOR words into ones / twos words, so the distance is
//   popc(ones) + popc(x7) + 2*(popc(c1) + popc(c2) + popc(c3))        = 5 popc + 6 LOP3
// instead of 8 popc.  (A fourth adder over c1..c3 would give 4 popc + 8 LOP3, which moves the
// bottleneck to the ALU pipe.)  Exact: a full adder preserves  a + b + c = sum + 2*carry  per bit.
// (inline PTX so that ptxas keeps exactly one LOP3 per sum and per carry: left to itself the
// compiler re-associates the XORs into the adders and ends up with 18 LOP3 instead of 14.)
__device__ __forceinline__ uint32_t csa_sum(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));  // a ^ b ^ c
  return r;
}
__device__ __forceinline__ uint32_t csa_carry(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));  // majority
  return r;
}
// FOUR = false: 5 POPC + 14 LOP3;  FOUR = true: a fourth adder over the carries gives
// popc(ones) + popc(x7) + 2 popc(twos) + 4 popc(fours) = 4 POPC + 16 LOP3.  The kernel alternates
// the two forms over its query rows so that the XU (POPC) and ALU (LOP3) pipes finish together.
// The weighted sum of the popcounts runs on the FMA pipe (IMAD with multipliers the compiler
// cannot see through: `one` = 1, `two` = 2), which is otherwise idle; the ALU pipe keeps LOP3/min.
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
  return r;
}
template <bool FOUR>
__device__ __forceinline__ uint32_t hamming256(const uint32_t (&q)[8], const uint4& ta,
                                               const uint4& tb, uint32_t one, uint32_t two) {
  const uint32_t x0 = q[0] ^ ta.x, x1 = q[1] ^ ta.y, x2 = q[2] ^ ta.z, x3 = q[3] ^ ta.w;
  const uint32_t x4 = q[4] ^ tb.x, x5 = q[5] ^ tb.y, x6 = q[6] ^ tb.z, x7 = q[7] ^ tb.w;
  const uint32_t s1 = csa_sum(x0, x1, x2), c1 = csa_carry(x0, x1, x2);
  const uint32_t s2 = csa_sum(x3, x4, x5), c2 = csa_carry(x3, x4, x5);
  const uint32_t ones = csa_sum(s1, s2, x6), c3 = csa_carry(s1, s2, x6);
  const uint32_t p1 = imad(__popc(ones), one, __popc(x7));
  if (FOUR) {
    const uint32_t twos = csa_sum(c1, c2, c3), fours = csa_carry(c1, c2, c3);
    // p1 + 2 * (popc(twos) + 2 * popc(fours))
    return imad(imad(__popc(fours), two, __popc(twos)), two, p1);
  }
  // p1 + 2 * (popc(c1) + popc(c2) + popc(c3))
  return imad(imad(__popc(c1), one, imad(__popc(c2), one, __popc(c3))), two, p1);
}

template <int RQ>
__global__ void __launch_bounds__(HM_THREADS)
hamming_top2_kernel(const uint8_t* __restrict__ desc, const int32_t* __restrict__ nkp, int n_cap,
                    const int32_t* __restrict__ pairs, const uint8_t* __restrict__ desc2_direct,
                    int n1_direct, int n2_direct, int n_splits,
                    uint32_t* __restrict__ rowbest /*[P][S][n_cap][2]*/,
                    uint32_t* __restrict__ colmin /*[P][n_cap]*/) {
  __shared__ __align__(128) uint4 tile[HM_NSTAGE][HM_TT * 2];
  __shared__ uint32_t cmin[HM_NSTAGE][HM_TT];
  __shared__ __align__(8) uint64_t full[HM_NSTAGE];

  const int pair = blockIdx.y;
  const int split = blockIdx.z;
  // pairs == nullptr: single pair given directly as (desc, desc2_direct)
  const int f1 = pairs ? pairs[2 * pair] : 0, f2 = pairs ? pairs[2 * pair + 1] : 0;
  const int n1 = pairs ? nkp[f1] : n1_direct, n2 = pairs ? nkp[f2] : n2_direct;
  const int q0 = blockIdx.x * (HM_THREADS * RQ);
  if (q0 >= n1) return;

  // train range of this split, in whole tiles
  const int n_tiles_all = (n2 + HM_TT - 1) / HM_TT;
  const int tiles_per_split = (n_tiles_all + n_splits - 1) / n_splits;
  const int tile_lo = split * tiles_per_split;
  const int tile_hi = min(n_tiles_all, tile_lo + tiles_per_split);
  const int n_tiles = tile_hi - tile_lo;

  const int tid = threadIdx.x;
  const uint8_t* d2 = pairs ? desc + (size_t)f2 * n_cap * 32 : desc2_direct;

  if (tid == 0) {
    for (int s = 0; s < HM_NSTAGE; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  for (int k = tid; k < HM_NSTAGE * HM_TT; k += HM_THREADS) (&cmin[0][0])[k] = KEY_INF;
  __syncthreads();

  auto issue = [&](int k) {  // tile k of this split -> stage k % NSTAGE   (thread 0 only)
    const int j0 = (tile_lo + k) * HM_TT;
    const uint32_t bytes = (uint32_t)min(HM_TT, n2 - j0) * 32u;
    const int s = k % HM_NSTAGE;
    mbar_expect_tx(&full[s], bytes);
    bulk_g2s(&tile[s][0], d2 + (size_t)j0 * 32, bytes, &full[s]);
  };
  if (tid == 0)
    for (int k = 0; k < min(HM_NSTAGE, n_tiles); ++k) issue(k);

  // query descriptors -> registers
  uint32_t qd[RQ][8];
  uint32_t qkey[RQ];
  uint32_t m1[RQ], m2[RQ];
  const uint4* d1 = reinterpret_cast<const uint4*>(desc + (size_t)f1 * n_cap * 32);
#pragma unroll
  for (int r = 0; r < RQ; ++r) {
    const int i = q0 + r * HM_THREADS + tid;
    const bool ok = i < n1;
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (ok) {
      a = __ldg(d1 + 2 * i);
      b = __ldg(d1 + 2 * i + 1);
    }
    qd[r][0] = a.x; qd[r][1] = a.y; qd[r][2] = a.z; qd[r][3] = a.w;
    qd[r][4] = b.x; qd[r][5] = b.y; qd[r][6] = b.z; qd[r][7] = b.w;
    // invalid queries get the largest query index bits + a distance bias that can never win
    qkey[r] = ok ? (uint32_t)i : 0x7fff0000u;
    m1[r] = KEY_INF;
    m2[r] = KEY_INF;
  }

  const uint32_t sh16 = 65536u + (uint32_t)(n_splits < 0);   // = 65536, opaque: keeps the IMAD
  const uint32_t one = 1u + (uint32_t)(n_splits < 0), two = 2u + (uint32_t)(n_splits < 0);
  for (int k = 0; k < n_tiles; ++k) {
    const int s = k % HM_NSTAGE;
    const int j0 = (tile_lo + k) * HM_TT;
    const int nj = min(HM_TT, n2 - j0);
    mbar_wait(&full[s], (k / HM_NSTAGE) & 1);
    const uint4* tp = &tile[s][0];
    // two train descriptors per step: their keys are ordered once (lo, hi) and folded into the
    // running top-2 with 5 min/max instead of 6; key packing runs on the FMA pipe (IMAD)
    int jj = 0;
    for (; jj + 1 < nj; jj += 2) {
      const uint4 ta0 = tp[2 * jj], tb0 = tp[2 * jj + 1];
      const uint4 ta1 = tp[2 * jj + 2], tb1 = tp[2 * jj + 3];
      const uint32_t jkey = (uint32_t)(j0 + jj);
      uint32_t ck0 = KEY_INF, ck1 = KEY_INF;
#pragma unroll
      for (int r = 0; r < RQ; ++r) {
        const uint32_t d0 = (r & 1) ? hamming256<true>(qd[r], ta0, tb0, one, two) : hamming256<false>(qd[r], ta0, tb0, one, two);
        const uint32_t d1 = (r & 1) ? hamming256<true>(qd[r], ta1, tb1, one, two) : hamming256<false>(qd[r], ta1, tb1, one, two);
        uint32_t rk0, rk1, c0, c1;
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(rk0) : "r"(d0), "r"(sh16), "r"(jkey));
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(rk1) : "r"(d1), "r"(sh16), "r"(jkey + 1u));
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c0) : "r"(d0), "r"(sh16), "r"(qkey[r]));
        asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(c1) : "r"(d1), "r"(sh16), "r"(qkey[r]));
        const uint32_t lo = min(rk0, rk1), hi = max(rk0, rk1);
        m2[r] = __vimin3_u32(m2[r], max(m1[r], lo), hi);
        m1[r] = min(m1[r], lo);
        ck0 = min(ck0, c0);
        ck1 = min(ck1, c1);
      }
      ck0 = warp_min_u32(ck0);
      ck1 = warp_min_u32(ck1);
      if ((tid & 31) == 0) {
        atomicMin(&cmin[s][jj], ck0);
        atomicMin(&cmin[s][jj + 1], ck1);
      }
    }
    if (jj < nj) {   // odd tail of the tile
      const uint4 ta = tp[2 * jj], tb = tp[2 * jj + 1];
      const uint32_t jkey = (uint32_t)(j0 + jj);
      uint32_t ck = KEY_INF;
#pragma unroll
      for (int r = 0; r < RQ; ++r) {
        const uint32_t d = hamming256<false>(qd[r], ta, tb, one, two);
        const uint32_t rk = (d << 16) | jkey;
        m2[r] = min(m2[r], max(m1[r], rk));
        m1[r] = min(m1[r], rk);
        ck = min(ck, (d << 16) + qkey[r]);
      }
      ck = warp_min_u32(ck);
      if ((tid & 31) == 0) atomicMin(&cmin[s][jj], ck);
    }
    __syncthreads();  // all warps done with tile[s] and cmin[s]
    if (tid < nj) {
      const uint32_t v = cmin[s][tid];
      if (v < 0x7fff0000u) atomicMin(colmin + (size_t)pair * n_cap + j0 + tid, v);
    }
    if (tid < HM_TT) cmin[s][tid] = KEY_INF;
    if (tid == 0 && k + HM_NSTAGE < n_tiles) {
      fence_proxy_async();
      issue(k + HM_NSTAGE);
    }
  }

  uint2* rb = reinterpret_cast<uint2*>(rowbest) + ((size_t)pair * n_splits + split) * n_cap;
#pragma unroll
  for (int r = 0; r < RQ; ++r) {
    const int i = q0 + r * HM_THREADS + tid;
    if (i < n1) rb[i] = make_uint2(m1[r], m2[r]);
  }
}

// K1b: per pair, merge split partials, apply cross-check or ratio test, stable-compact by
// queryIdx, gather KeyPoint.pt of both sides into float64 (image_pair.py:294-299).
constexpr int CP_THREADS = 1024;
__global__ void __launch_bounds__(CP_THREADS)
match_compact_kernel(const uint32_t* __restrict__ rowbest, const uint32_t* __restrict__ colmin,
                     const int32_t* __restrict__ nkp, const float* __restrict__ kp, int n_cap,
                     const int32_t* __restrict__ pairs, int n1_direct, int n2_direct,
                     int n_splits, int mode, double ratio, int32_t* __restrict__ q, int32_t* __restrict__ t, float* __restrict__ dist,
                     int32_t* __restrict__ t2, float* __restrict__ dist2,
                     double* __restrict__ pts, int32_t* __restrict__ n_out) {
  __shared__ int warp_tot[CP_THREADS / 32];
  __shared__ int base_s;
  const int pair = blockIdx.x;
  const int f1 = pairs ? pairs[2 * pair] : 0, f2 = pairs ? pairs[2 * pair + 1] : 0;
  const int n1 = pairs ? nkp[f1] : n1_direct, n2 = pairs ? nkp[f2] : n2_direct;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const uint2* rb = reinterpret_cast<const uint2*>(rowbest) + (size_t)pair * n_splits * n_cap;
  const uint32_t* cm = colmin + (size_t)pair * n_cap;
  const size_t ob = (size_t)pair * n_cap;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < n1; i0 += CP_THREADS) {
    const int i = i0 + tid;
    uint32_t a = KEY_INF, b = KEY_INF;
    bool keep = false;
    if (i < n1 && n2 > 0) {
      for (int s = 0; s < n_splits; ++s) {
        const uint2 v = rb[(size_t)s * n_cap + i];
        // merge two sorted pairs
        const uint32_t lo = min(a, v.x), hi = max(a, v.x);
        b = min(hi, min(b, v.y));
        a = lo;
      }
      if (mode == VO_MATCH_MUTUAL) {
        keep = (cm[a & 0xffffu] & 0xffffu) == (uint32_t)i;
      } else {
        // Lowe ratio as written in feature_detection.py:24: m.distance < ratio * n.distance
        // ratio < 0: plain nearest neighbour for every query (no test)
        // (evaluated in double like the Python expression it restates)
        keep = ratio < 0.0 ? true
                           : (b != KEY_INF) && ((double)(a >> 16) < ratio * (double)(b >> 16));
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_tot[wid] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int w = 0; w < wid; ++w) off += warp_tot[w];
    off += __popc(bal & ((1u << lane) - 1u));
    if (keep) {
      const int j = a & 0xffffu;
      q[ob + off] = i;
      t[ob + off] = j;
      dist[ob + off] = (float)(a >> 16);
      if (t2) t2[ob + off] = (b == KEY_INF) ? -1 : (int)(b & 0xffffu);
      if (dist2) dist2[ob + off] = (b == KEY_INF) ? 0.f : (float)(b >> 16);
      if (pts && kp) {
        const float2 p1 = reinterpret_cast<const float2*>(kp)[(size_t)f1 * n_cap + i];
        const float2 p2 = reinterpret_cast<const float2*>(kp)[(size_t)f2 * n_cap + j];
        double4 o = make_double4((double)p1.x, (double)p1.y, (double)p2.x, (double)p2.y);
        reinterpret_cast<double4*>(pts)[ob + off] = o;
      }
    }
    __syncthreads();
    if (tid == 0) {
      int tot = 0;
      for (int w = 0; w < CP_THREADS / 32; ++w) tot += warp_tot[w];
      base_s += tot;
    }
    __syncthreads();
  }
  if (tid == 0) n_out[pair] = base_s;
}

}  // namespace vo

using namespace vo;

static size_t popc_workspace_bytes(int n_pairs, int n_cap) {
  // rowbest for up to 16 splits + colmin
  return (size_t)n_pairs * n_cap * (16 * 2 + 1) * sizeof(uint32_t);
}

// bits 0..1: 0 = choose by batch size, 1 = always the popc kernel, 2 / 3 = always the int8 / fp4
// tensor-core kernel (need n_frames > 0); bits 3..5: timing-experiment instantiations of those
static int g_hamming_debug = 0;
static int g_hamming_tensor_default = 3;   // 2 = K1t (int8), 3 = K1f (fp4)
extern "C" int vo_debug_hamming(int flags) {
  g_hamming_debug = flags;
  return VO_OK;
}

static int device_sms() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!sms[dev] &&
      (cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms[dev] <= 0))
    sms[dev] = 148;
  return sms[dev];
}

// The tensor-core path pays a per-frame operand expansion and works in (pair, direction, 256-row
// block) items of ~10 us each: worth it once the items give every SM something to do.
static int use_tensor_path(int n_frames, int n_pairs, int n_cap, int mode) {
  const int path = g_hamming_debug & 3;
  if (n_frames <= 0 || path == 1) return 0;
  if (path >= 2) return path;
  const long items = (long)n_pairs * (mode == VO_MATCH_MUTUAL ? 2 : 1) * ((n_cap + htc::ROWS - 1) / htc::ROWS);
  return (n_cap >= 512 && items >= device_sms() / 2) ? g_hamming_tensor_default : 0;
}

extern "C" size_t vo_match_workspace_bytes(int n_frames, int n_pairs, int n_cap) {
  if (n_pairs <= 0 || n_cap <= 0) return 0;
  const size_t a = popc_workspace_bytes(n_pairs, n_cap);
  const size_t b = n_frames > 0 ? htc::carve(nullptr, n_frames, n_pairs, n_cap).bytes : 0;
  const size_t c = n_frames > 0 ? hf4::carve(nullptr, n_frames, n_pairs, n_cap).bytes : 0;
  return a > b ? (a > c ? a : c) : (b > c ? b : c);
}

// Train-range splits serve two purposes: filling the SMs when there are few (pair, query tile)
// CTAs, and avoiding a mostly empty last wave when there are many: the split count with the best
// wave efficiency  waves / ceil(waves)  wins (each extra split costs ~1.5 % in partial merges).
static int pick_splits(int n_pairs, int qtiles, int n_cap, int resident_ctas) {
  const long ctas = (long)n_pairs * qtiles;
  const int max_s = (n_cap + HM_TT - 1) / HM_TT;
  if (ctas * 4 <= resident_ctas) {           // tiny batch: spread over the SMs
    int s = (int)((resident_ctas + ctas - 1) / ctas);
    if (s > 16) s = 16;
    if (s > max_s) s = max_s;
    return s < 1 ? 1 : s;
  }
  int best = 1;
  double best_eff = -1.0;
  for (int s = 1; s <= 4 && s <= max_s; ++s) {
    const double waves = (double)(ctas * s) / resident_ctas;
    const double full = (double)((ctas * s + resident_ctas - 1) / resident_ctas);
    const double eff = waves / full - 0.015 * (s - 1);
    if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
  }
  return best;
}

template <int RQ>
static int hamming_resident_ctas() {
  static int cached = 0;
  if (!cached) {
    int dev = 0, sms = 148, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, hamming_top2_kernel<RQ>, HM_THREADS,
                                                      0) != cudaSuccess || per_sm <= 0)
      per_sm = 4;
    cached = sms * per_sm;
  }
  return cached;
}

static int launch_hamming(const uint8_t* desc, const uint8_t* desc2, int n1d, int n2d,
                          const int32_t* nkp, const float* kp, int n_cap, const int32_t* pairs,
                          int n_pairs, int mode, double ratio, int32_t* q, int32_t* t, float* dist,
                          int32_t* t2, float* dist2, double* pts, int32_t* n_out, void* ws,
                          cudaStream_t st) {
  // small batches: 128-query tiles + train splits to fill 148 SMs; big batches: 512-query tiles
  const bool small = (long)n_pairs * ((n_cap + 511) / 512) < 148L * 4;
  const int rq = small ? 1 : 4;
  const int qtiles = (n_cap + HM_THREADS * rq - 1) / (HM_THREADS * rq);
  const int splits = pick_splits(n_pairs, qtiles, n_cap, rq == 1 ? hamming_resident_ctas<1>()
                                                                 : hamming_resident_ctas<4>());
  uint32_t* rowbest = (uint32_t*)ws;  // 8-byte aligned (uint2 access); colmin sits behind it
  uint32_t* colmin = rowbest + (size_t)n_pairs * n_cap * 32;
  cudaMemsetAsync(colmin, 0xff, (size_t)n_pairs * n_cap * sizeof(uint32_t), st);
  for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
    const int np = n_pairs - p0 < 65535 ? n_pairs - p0 : 65535;
    dim3 grid(qtiles, np, splits);
    uint32_t* rb = rowbest + (size_t)p0 * splits * n_cap * 2;
    uint32_t* cm = colmin + (size_t)p0 * n_cap;
    const int32_t* pp = pairs ? pairs + 2 * p0 : nullptr;
    {
      ProfScope prof(0, st);
      if (rq == 1)
      hamming_top2_kernel<1><<<grid, HM_THREADS, 0, st>>>(desc, nkp, n_cap, pp, desc2, n1d, n2d,
                                                         splits, rb, cm);
    else
      hamming_top2_kernel<4><<<grid, HM_THREADS, 0, st>>>(desc, nkp, n_cap, pp, desc2, n1d, n2d,
                                                         splits, rb, cm);
    }
    const size_t ob = (size_t)p0 * n_cap;
    ProfScope prof_c(8, st);
    match_compact_kernel<<<np, CP_THREADS, 0, st>>>(
        rb, cm, nkp, kp, n_cap, pp, n1d, n2d, splits, mode, ratio, q + ob, t + ob, dist + ob,
        t2 ? t2 + ob : nullptr, dist2 ? dist2 + ob : nullptr, pts ? pts + ob * 4 : nullptr,
        n_out + p0);
  }
  return check_launch("vo_match_hamming");
}

// tensor-core path: expand every frame once, both directions as MMA passes, the same compaction
static int launch_hamming_tc(const uint8_t* desc, const int32_t* nkp, const float* kp, int n_cap,
                             int n_frames, const int32_t* pairs, int n_pairs, int mode, double ratio,
                             int32_t* q, int32_t* t, float* dist, int32_t* t2, float* dist2,
                             double* pts, int32_t* n_out, void* ws, cudaStream_t st) {
  const int ntile = (n_cap + htc::ROWS - 1) / htc::ROWS;
  const htc::Ws w = htc::carve(ws, n_frames, n_pairs, n_cap);
  const int smem = (int)sizeof(htc::Smem) + 1024;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!attr_set[dev]) {
    if (cudaFuncSetAttribute(htc::match_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(htc::match_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(htc::match_kernel<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(htc::match_kernel<false, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
      return check_launch("vo_match_hamming_batch(smem attribute)");
    attr_set[dev] = true;
  }
  {
    ProfScope prof(4, st);
    htc::expand_kernel<<<dim3(ntile, n_frames), htc::ROWS, 0, st>>>(desc, nkp, n_cap, ntile, w.tiles);
  }
  const int ndir = mode == VO_MATCH_MUTUAL ? 2 : 1;
  const long n_items = (long)n_pairs * ndir * ntile;
  const int grid = (int)(n_items < device_sms() ? n_items : device_sms());
  {
    ProfScope prof(0, st);
    const int kd = (g_hamming_debug >> 2) & 6;     // timing experiments: compile-time instantiations
    if (mode == VO_MATCH_MUTUAL && kd == 2)
      htc::match_kernel<false, 2><<<grid, htc::THREADS, smem, st>>>(w.tiles, nkp, pairs, n_cap, ntile, ndir,
                                                                   n_items, w.rowbest, w.colmin);
    else if (mode == VO_MATCH_MUTUAL && kd == 6)
      htc::match_kernel<false, 6><<<grid, htc::THREADS, smem, st>>>(w.tiles, nkp, pairs, n_cap, ntile, ndir,
                                                                   n_items, w.rowbest, w.colmin);
    else if (mode == VO_MATCH_MUTUAL)
      htc::match_kernel<false><<<grid, htc::THREADS, smem, st>>>(w.tiles, nkp, pairs, n_cap, ntile, ndir,
                                                                n_items, w.rowbest, w.colmin);
    else
      htc::match_kernel<true><<<grid, htc::THREADS, smem, st>>>(w.tiles, nkp, pairs, n_cap, ntile, ndir,
                                                               n_items, w.rowbest, w.colmin);
  }
  for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
    const int np = n_pairs - p0 < 65535 ? n_pairs - p0 : 65535;
    const size_t ob = (size_t)p0 * n_cap;
    ProfScope prof_c(8, st);
    match_compact_kernel<<<np, CP_THREADS, 0, st>>>(
        w.rowbest + ob * 2, w.colmin + ob, nkp, kp, n_cap, pairs + 2 * p0, 0, 0, 1, mode, ratio, q + ob,
        t + ob, dist + ob, t2 ? t2 + ob : nullptr, dist2 ? dist2 + ob : nullptr,
        pts ? pts + ob * 4 : nullptr, n_out + p0);
  }
  return check_launch("vo_match_hamming_batch(tensor)");
}

// fp4 tensor-core path (K1f): both operand forms of every frame once, then as K1t
static int launch_hamming_f4(const uint8_t* desc, const int32_t* nkp, const float* kp, int n_cap,
                             int n_frames, const int32_t* pairs, int n_pairs, int mode, double ratio,
                             int32_t* q, int32_t* t, float* dist, int32_t* t2, float* dist2,
                             double* pts, int32_t* n_out, void* ws, cudaStream_t st, float* probe_acc) {
  const int nta = (n_cap + hf4::ROWS_A - 1) / hf4::ROWS_A, ntb = (n_cap + hf4::ROWS_B - 1) / hf4::ROWS_B;
  const int rows_a = nta * hf4::ROWS_A;
  const hf4::Ws w = hf4::carve(ws, n_frames, n_pairs, n_cap);
  const int smem = (int)sizeof(hf4::Smem) + 1024;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  dev &= 63;
  if (!attr_set[dev]) {
    if (cudaFuncSetAttribute(hf4::match_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(hf4::match_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(hf4::match_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(hf4::match_kernel<false, false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess ||
        cudaFuncSetAttribute(hf4::match_kernel<false, false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess)
      return check_launch("vo_match_hamming_batch(smem attribute, fp4)");
    attr_set[dev] = true;
  }
  {
    ProfScope prof(4, st);
    const int rows = rows_a > ntb * hf4::ROWS_B ? rows_a : ntb * hf4::ROWS_B;
    hf4::expand_kernel<<<dim3((rows + 31) / 32, n_frames), 256, 0, st>>>(desc, nkp, n_cap, rows_a, ntb,
                                                                           w.a_rows, w.tiles_b);
  }
  const int ndir = mode == VO_MATCH_MUTUAL ? 2 : 1;
  const long n_items = (long)n_pairs * ndir * nta;
  const int grid = (int)(n_items < device_sms() ? n_items : device_sms());
  {
    ProfScope prof(0, st);
    const int kd = (g_hamming_debug >> 3) & 7;     // timing experiments: compile-time instantiations
#define VO_F4_ARGS w.a_rows, w.tiles_b, nkp, pairs, n_cap, rows_a, ntb
    if (probe_acc)
      hf4::match_kernel<false, true><<<1, hf4::THREADS, smem, st>>>(VO_F4_ARGS, 1, 1, w.rowbest, w.colmin, probe_acc);
    else if (mode == VO_MATCH_MUTUAL && kd == 1)
      hf4::match_kernel<false, false, 1><<<grid, hf4::THREADS, smem, st>>>(VO_F4_ARGS, ndir, n_items, w.rowbest, w.colmin, nullptr);
    else if (mode == VO_MATCH_MUTUAL && kd == 2)
      hf4::match_kernel<false, false, 2><<<grid, hf4::THREADS, smem, st>>>(VO_F4_ARGS, ndir, n_items, w.rowbest, w.colmin, nullptr);
    else if (mode == VO_MATCH_MUTUAL)
      hf4::match_kernel<false><<<grid, hf4::THREADS, smem, st>>>(VO_F4_ARGS, ndir, n_items, w.rowbest, w.colmin, nullptr);
    else
      hf4::match_kernel<true><<<grid, hf4::THREADS, smem, st>>>(VO_F4_ARGS, ndir, n_items, w.rowbest, w.colmin, nullptr);
#undef VO_F4_ARGS
  }
  if (probe_acc) return check_launch("vo_debug_hamming_f4_acc");
  for (int p0 = 0; p0 < n_pairs; p0 += 65535) {
    const int np = n_pairs - p0 < 65535 ? n_pairs - p0 : 65535;
    const size_t ob = (size_t)p0 * n_cap;
    ProfScope prof_c(8, st);
    match_compact_kernel<<<np, CP_THREADS, 0, st>>>(
        w.rowbest + ob * 2, w.colmin + ob, nkp, kp, n_cap, pairs + 2 * p0, 0, 0, 1, mode, ratio, q + ob,
        t + ob, dist + ob, t2 ? t2 + ob : nullptr, dist2 ? dist2 + ob : nullptr,
        pts ? pts + ob * 4 : nullptr, n_out + p0);
  }
  return check_launch("vo_match_hamming_batch(fp4 tensor)");
}

// debug: the raw fp32 accumulator of K1f for frames (0, 1) of `desc`, rows 0..127 x columns 0..207:
// acc[i][j] = 36 * (256 - 2 * Hamming(i, j)) + (207 - j) / 4 when the contraction is right
extern "C" int vo_debug_hamming_f4_acc(const uint8_t* desc, const int32_t* nkp, int n_cap,
                                       const int32_t* pairs, float* acc, void* ws, size_t ws_bytes,
                                       void* stream) {
  VO_REQUIRE(desc && nkp && pairs && acc && ws, VO_ERR_INVALID_ARG, "vo_debug_hamming_f4_acc: null pointer");
  VO_REQUIRE(ws_bytes >= vo_match_workspace_bytes(2, 1, n_cap), VO_ERR_WORKSPACE,
             "vo_debug_hamming_f4_acc: workspace too small");
  return launch_hamming_f4(desc, nkp, nullptr, n_cap, 2, pairs, 1, VO_MATCH_MUTUAL, 0.0, nullptr, nullptr,
                           nullptr, nullptr, nullptr, nullptr, nullptr, ws, (cudaStream_t)stream, acc);
}

extern "C" int vo_match_hamming_batch(const uint8_t* desc, const int32_t* nkp, const float* kp,
                                      int n_cap, int n_frames, const int32_t* pairs, int n_pairs,
                                      int mode, double ratio, int32_t* q, int32_t* t, float* dist,
                                      int32_t* t2, float* dist2, double* pts, int32_t* n_out,
                                      void* ws, size_t ws_bytes, void* stream) {
  VO_REQUIRE(desc && nkp && pairs && q && t && dist && n_out && ws, VO_ERR_INVALID_ARG,
             "vo_match_hamming_batch: null pointer");
  VO_REQUIRE(n_cap > 0 && n_cap <= 65535 && n_pairs >= 0 && n_frames >= 0, VO_ERR_INVALID_ARG,
             "vo_match_hamming_batch: n_cap must be in 1..65535 (got %d)", n_cap);
  VO_REQUIRE(mode == VO_MATCH_MUTUAL || mode == VO_MATCH_KNN2_RATIO, VO_ERR_INVALID_ARG,
             "vo_match_hamming_batch: bad mode %d", mode);
  VO_REQUIRE(ws_bytes >= vo_match_workspace_bytes(n_frames, n_pairs, n_cap), VO_ERR_WORKSPACE,
             "vo_match_hamming_batch: workspace too small");
  if (n_pairs == 0) return VO_OK;
  const int path = use_tensor_path(n_frames, n_pairs, n_cap, mode);
  if (path == 3)
    return launch_hamming_f4(desc, nkp, kp, n_cap, n_frames, pairs, n_pairs, mode, ratio, q, t, dist,
                             t2, dist2, pts, n_out, ws, (cudaStream_t)stream, nullptr);
  if (path == 2)
    return launch_hamming_tc(desc, nkp, kp, n_cap, n_frames, pairs, n_pairs, mode, ratio, q, t, dist,
                             t2, dist2, pts, n_out, ws, (cudaStream_t)stream);
  return launch_hamming(desc, nullptr, 0, 0, nkp, kp, n_cap, pairs, n_pairs, mode, ratio, q, t,
                        dist, t2, dist2, pts, n_out, ws, (cudaStream_t)stream);
}

extern "C" int vo_match_hamming(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int mode,
                                double ratio, int32_t* q, int32_t* t, float* dist, int32_t* t2,
                                float* dist2, int32_t* n_out, void* ws, size_t ws_bytes,
                                void* stream) {
  VO_REQUIRE(q && t && dist && n_out && ws, VO_ERR_INVALID_ARG, "vo_match_hamming: null pointer");
  VO_REQUIRE(n1 >= 0 && n2 >= 0 && n1 <= 65535 && n2 <= 65535, VO_ERR_INVALID_ARG,
             "vo_match_hamming: descriptor counts must be in 0..65535");
  VO_REQUIRE(mode == VO_MATCH_MUTUAL || mode == VO_MATCH_KNN2_RATIO, VO_ERR_INVALID_ARG,
             "vo_match_hamming: bad mode %d", mode);
  const int n_cap = n1 > n2 ? n1 : n2;
  if (n1 == 0 || n2 == 0) {
    cudaMemsetAsync(n_out, 0, sizeof(int32_t), (cudaStream_t)stream);
    return check_launch("vo_match_hamming");
  }
  VO_REQUIRE(d1 && d2, VO_ERR_INVALID_ARG, "vo_match_hamming: null descriptors");
  VO_REQUIRE(ws_bytes >= vo_match_workspace_bytes(0, 1, n_cap), VO_ERR_WORKSPACE,
             "vo_match_hamming: workspace too small");
  return launch_hamming(d1, d2, n1, n2, nullptr, nullptr, n_cap, nullptr, 1, mode, ratio, q, t,
                        dist, t2, dist2, nullptr, n_out, ws, (cudaStream_t)stream);
}

// ---------------------------------------------------------------- no-allocation variant
// vo_match_hamming with its scratch (row top-2 keys per train split + column minima, 132 bytes per
// key point) in a file-scope device array instead of a caller workspace: for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  No
// counterpart in the reference.
//   * limit: max(n1, n2) <= VO_STATIC_MAX_KP = 8192 key points (1 081 344 bytes of device memory,
//     held by every GPU that loads the library until the process exits);
//   * the array's name is only a host-side stand-in in host code: its device address is taken
//     once per device with cudaGetSymbolAddress (that device current) and passed to the kernels;
//   * every call shares the array and finds the previous call's contents in it: the column minima
//     are re-initialised by the launch group itself (cudaMemsetAsync on the caller's stream) and
//     every row key that is read is written first, so back-to-back calls on ONE stream are
//     independent; calls that overlap on the GPU (two streams, two host threads) RACE on it.
// Calls no cudaMalloc / cudaFree / allocating library routine.
constexpr int VO_STATIC_MAX_KP = 8192;
static __device__ __align__(256) uint32_t g_static_scratch[(size_t)VO_STATIC_MAX_KP * (16 * 2 + 1)];

static int static_scratch_address(void** out) {
  static void* addr[64] = {nullptr};
  static bool known[64] = {false};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("vo_match_hamming_static: cannot identify the current device");
    return VO_ERR_CUDA;
  }
  if (!known[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      set_error("vo_match_hamming_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return VO_ERR_CUDA;
    }
    addr[dev] = p;
    known[dev] = true;
  }
  *out = addr[dev];
  return VO_OK;
}

extern "C" int vo_match_hamming_static_max_keypoints(void) { return VO_STATIC_MAX_KP; }

extern "C" int vo_match_hamming_static(const uint8_t* d1, int n1, const uint8_t* d2, int n2, int mode,
                                       double ratio, int32_t* q, int32_t* t, float* dist, int32_t* t2,
                                       float* dist2, int32_t* n_out, void* stream) {
  VO_REQUIRE(q && t && dist && n_out, VO_ERR_INVALID_ARG, "vo_match_hamming_static: null pointer");
  VO_REQUIRE(n1 >= 0 && n2 >= 0, VO_ERR_INVALID_ARG, "vo_match_hamming_static: negative descriptor count");
  VO_REQUIRE(n1 <= VO_STATIC_MAX_KP && n2 <= VO_STATIC_MAX_KP, VO_ERR_WORKSPACE,
             "vo_match_hamming_static: %d x %d descriptors exceed the static scratch (at most %d per "
             "side); use vo_match_hamming with a caller workspace", n1, n2, VO_STATIC_MAX_KP);
  VO_REQUIRE(mode == VO_MATCH_MUTUAL || mode == VO_MATCH_KNN2_RATIO, VO_ERR_INVALID_ARG,
             "vo_match_hamming_static: bad mode %d", mode);
  const int n_cap = n1 > n2 ? n1 : n2;
  if (n1 == 0 || n2 == 0) {
    cudaMemsetAsync(n_out, 0, sizeof(int32_t), (cudaStream_t)stream);
    return check_launch("vo_match_hamming_static");
  }
  VO_REQUIRE(d1 && d2, VO_ERR_INVALID_ARG, "vo_match_hamming_static: null descriptors");
  void* ws = nullptr;
  const int rc = static_scratch_address(&ws);
  if (rc != VO_OK) return rc;
  return launch_hamming(d1, d2, n1, n2, nullptr, nullptr, n_cap, nullptr, 1, mode, ratio, q, t,
                        dist, t2, dist2, nullptr, n_out, ws, (cudaStream_t)stream);
}

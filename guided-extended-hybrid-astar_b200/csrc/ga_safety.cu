// ga_safety.cu -- occupancy pipeline: patch insertion (K7) and ONE fused kernel (K1 + K2):
// threshold -> bit planes -> disk dilation -> byte expansion -> 4x4 max-pool.
//
// Replaces CollisionChecker::passLocalMapData (collision_checking.cpp:269-313),
// CollisionChecker::processSafetyPatch (collision_checking.cpp:193-224) and Pooling::execute
// (LIB/src/cuda_lib/max_pool.cu:21-79).
//
// Design: thresholded cells take the values {FREE 0, UNKNOWN 1, OCC 2} and cv::dilate is a max
// filter, so the whole pipeline is exact on two 1-bit planes: P2 = (cell == OCC) and
// P1 = (cell != FREE), 32 cells per word; max over the disk = P1' + P2' where ' is binary
// dilation.  The elliptical structuring element is read by columns: a cell at horizontal distance w
// of the centre is covered by the rows |dy| <= hmax(w).  A CTA owns a tile of 256 x 64 cells:
//   0  reads the raw bytes of the tile + halo (16-byte loads, all in flight before the first is used),
//      thresholds them (SWAR compare + DP4A bit gather) and writes 16-bit halves of the two bit planes
//      to shared memory -- the planes never exist in global memory; of the halo words left and right
//      only the inner 16 cells are read (r <= 16);
//   V  cumulative vertical ORs V_h(y) = OR of the plane rows y-h .. y+h, two words per thread (64-bit
//      shared accesses), stored at the distinct hmax values (7 for the 21 x 21 disk);
//   B  dilated(y) = OR_w shift(+-w) of V_hmax(w)(y) by funnel shifts, two output words per thread;
//      the rows of a warp are 4 apart so that its shared accesses (row pitch 10 words) are conflict free;
//   C  bits -> bytes through a 256-entry table (4 cells per lookup), 16-byte coalesced stores;
//   P  4x4 max-pool of the tile from the same shared words (8 pooled cells per thread).
// HBM traffic is the compulsory 1 B/cell read + 1 B/cell write + 1/16 B/cell pooled write; the halo
// re-reads (1.5x) hit L2.
#include <cuda.h>  // CUtensorMap (types only: the encoder is looked up through the runtime, libcuda is not linked)

#include "ga_internal.cuh"

#define DT_TW 8            // tile width in words (256 cells)
#define DT_TH 64           // tile height in rows
#define DT_SW (DT_TW + 2)  // staged plane words per row (one halo word each side); also the pitch of the output planes
#define DT_SH (2 * DT_SW)  // the same in 16-cell halves
#define DT_NH (DT_SH - 2)  // halves that are read: the outer half of each halo word is never needed (r <= 16)
#define DT_THREADS 256
static_assert(GA_MAX_DISK_R <= 16, "phase 0 reads only the inner half of the halo words");

// SWAR byte compare: bit 7 of byte k of the result = (byte k of x > c), any c in 0..255.
// k7 = 0x01010101 * (127 - (c & 127)): adding it to the low 7 bits carries into bit 7 iff they exceed c & 127;
// sel = all ones when c >= 128 (then bit 7 of x must be set as well), 0 otherwise (then bit 7 of x suffices).
// Three logic operations and one add, no branch on the threshold.
__device__ __forceinline__ uint32_t bytes_gt(uint32_t x, uint32_t k7, uint32_t sel)
{
  const uint32_t m = (x & 0x7F7F7F7Fu) + k7;
  return ((m & x) | (~sel & (m | x))) & 0x80808080u;
}
// 16 compared bytes (0x80 where the compare passed) -> 16 bits.  A byte-wise dot product with the weights 1, 2, 4, 8
// (16 .. 128 for the next word) adds up 128 x the bit pattern: one DP4A per word.
__device__ __forceinline__ uint32_t gather16(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3)
{
  const uint32_t lo = __dp4a(c1, 0x80402010u, __dp4a(c0, 0x08040201u, 0u));  // 128 * bits 0..7
  const uint32_t hi = __dp4a(c3, 0x80402010u, __dp4a(c2, 0x08040201u, 0u));  // 128 * bits 8..15
  return (lo >> 7) + (hi << 1);
}

// The shipped footprint (default vehicle, SAFETY_DISTANCE_M 0.1) gives the 21 x 21 disk: radius 10, half widths
// 10,10,10,10,9,9,8,7,6,4,0 for |dy| = 0..10.  Its column view as compile-time tables lets phases V and B unroll
// completely (SPEC10); any other element runs the same kernel with the tables read from `se`.
__device__ constexpr int K10_R = 10, K10_NLVL = 7;
__device__ constexpr int K10_LVL_H[K10_NLVL] = { 3, 5, 6, 7, 8, 9, 10 };
__device__ constexpr int K10_LVL_OF_W[K10_R + 1] = { 6, 5, 5, 5, 5, 4, 4, 3, 2, 1, 0 };

// LOADW: widest aligned access the row pitch allows -- 16 (dim % 16 == 0), 4 (dim % 4 == 0) or 1.
// TMA: the raw bytes of the tile + halo arrive as ONE cp.async.bulk.tensor.2d (needs LOADW == 16: the row pitch must be a
// multiple of 16 bytes); otherwise by per-thread cp.async (LOADW 16 / 4) or byte loads (LOADW 1).
template <bool SPEC10, int LOADW, bool TMA>
__global__ void __launch_bounds__(DT_THREADS, SPEC10 ? 7 : 1)
safety_kernel(const uint8_t* __restrict__ patch, int dim, uint8_t* __restrict__ safety, uint8_t* __restrict__ astar,
              int astar_dim, const ga_se se, const uint32_t lo, const uint4 thr, const __grid_constant__ CUtensorMap tmap)
{
  extern __shared__ __align__(128) uint32_t smem[];
  __shared__ uint32_t s_lut[256];   // (P1 nibble | P2 nibble << 4) -> 4 cell bytes
  __shared__ __align__(8) unsigned long long s_mbar;
  const int r = SPEC10 ? K10_R : se.r;
  const int n_lvl = SPEC10 ? K10_NLVL : se.n_lvl;
  const int rows_in = DT_TH + 2 * r;
  uint32_t* sP = smem;                          // [plane][rows_in][DT_SW]   thresholded bit planes
  uint32_t* sV = sP + ((2 * rows_in * DT_SW + 31) & ~31);  // (128-byte aligned: the bulk copy's destination) [level][DT_TH][DT_SW]  vertical ORs over |dy| <= hmax(level) of ONE plane at a time
  uint32_t* sOut = sV + n_lvl * DT_TH * DT_SW;  // [plane][DT_TH][DT_SW]  dilated planes
  constexpr int lvl_stride = DT_TH * DT_SW;
  const int w0 = blockIdx.x * DT_TW, y0 = blockIdx.y * DT_TH;
  const int tid = threadIdx.x;

  {
    // table: byte k = ((idx >> k) & 1) + ((idx >> (4 + k)) & 1)
    const uint32_t a = tid & 15, b = tid >> 4;
    s_lut[tid] = (((a & 0xFu) * 0x00204081u) & 0x01010101u) + (((b & 0xFu) * 0x00204081u) & 0x01010101u);
  }
  // ---- phase 0: threshold + pack.  One item = 16 cells (half a plane word), stored as 16 bits.  Cells outside the
  //      image are 0 in both planes (OpenCV ignores them).  Items are numbered row by row over the DT_NH halves a row
  //      needs; a thread's next item is DT_THREADS further: (row, half) advance by constants, no division in the loop.
  uint16_t* sP16 = reinterpret_cast<uint16_t*>(sP);
  constexpr int P0_N = ((DT_TH + 2 * (SPEC10 ? K10_R : GA_MAX_DISK_R)) * DT_NH + DT_THREADS - 1) / DT_THREADS;  // items per thread
  constexpr int STEP_R = DT_THREADS / DT_NH, STEP_C = DT_THREADS % DT_NH;
  // compare constants of bytes_gt for MAX_THRESH and MIN_THRESH - 1, made on the host: as opaque kernel arguments they
  // stay operands of one three-input logic operation (computed here, the compiler turns `sel` back into a predicate
  // and a second, predicated operation per compare)
  const uint32_t hi_k7 = thr.x, hi_sel = thr.y, lo_k7 = thr.z, lo_sel = thr.w;
  const int xh0 = (w0 - 1) * 32 + 16;  // first cell of the first half that is read (inner half of the left halo word)
  // the outer halves of the halo words: never shifted into the tile, zeroed so that phase V works on defined words
  for (int i = tid; i < 4 * rows_in; i += DT_THREADS) sP16[(i >> 1) * DT_SH + (i & 1) * (DT_SH - 1)] = 0;
  {
    // the raw bytes are staged in the (still unused) area of the vertical ORs by asynchronous copies: every item of the
    // tile is in flight at once (one memory round trip per tile, no registers held), rows / columns outside the image
    // are zero-filled (src-size 0).  A thread thresholds the items it copied itself: no barrier in between.
    uint8_t* sRaw = reinterpret_cast<uint8_t*>(sV);
    const uint32_t raw0 = (uint32_t)__cvta_generic_to_shared(sRaw) + tid * 16;
    int rl = tid / DT_NH, c = tid - rl * DT_NH;
    if (TMA)
    {
      // one bulk tensor copy per tile: box = DT_NH * 4 words x rows_in rows of the patch seen as a 2-D array of 32-bit
      // words, rows and columns outside the image are zero-filled by the copy engine; completion on an mbarrier
      const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
      if (tid == 0)
      {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar) : "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(rows_in * DT_NH * 16) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(sRaw)), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(mbar), "r"(xh0 / 4),
                     "r"(y0 - r)
                     : "memory");
      }
      __syncthreads();  // the barrier object exists before anybody waits on it
      uint32_t done = 0;
      for (int spin = 0; !done; ++spin)
      {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(mbar) : "memory");
        if (spin > (1 << 22)) __trap();  // a copy that never completes is an error, not a hang
      }
    }
    else
    {
#pragma unroll
    for (int j = 0; j < P0_N; ++j)
    {
      if (rl < rows_in)
      {
        const int y = y0 - r + rl, x = xh0 + c * 16;
        const bool ok = (unsigned)y < (unsigned)dim && x >= 0 && x < dim;
        const uint8_t* p = patch + (ok ? (uint32_t)(y * dim + x) : 0u);  // patch_dim <= 32768: the offset fits 32 bits
        const uint32_t dst = raw0 + j * (DT_THREADS * 16);
        if (LOADW == 16)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(p), "r"(ok ? 16 : 0) : "memory");
        else if (LOADW == 4)
        {
#pragma unroll
          for (int k = 0; k < 4; ++k)
          {
            const bool okk = ok && x + 4 * k < dim;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst + 4 * k), "l"(okk ? p + 4 * k : patch), "r"(okk ? 4 : 0) : "memory");
          }
        }
        else
        {
          uint32_t w[4] = { 0, 0, 0, 0 };
          if (ok)
          {
#pragma unroll
            for (int k = 0; k < 16; ++k)
              if (x + k < dim) w[k >> 2] |= (uint32_t)__ldg(p + k) << (8 * (k & 3));
          }
          *reinterpret_cast<uint4*>(sRaw + (tid + j * DT_THREADS) * 16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
      rl += STEP_R; c += STEP_C;
      if (c >= DT_NH) { c -= DT_NH; ++rl; }
    }
    if (LOADW >= 4) asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    }
    rl = tid / DT_NH; c = tid - rl * DT_NH;
#pragma unroll
    for (int j = 0; j < P0_N; ++j)
    {
      if (rl < rows_in)
      {
        const uint4 q = *reinterpret_cast<const uint4*>(sRaw + (tid + j * DT_THREADS) * 16);
        const uint32_t v[4] = { q.x, q.y, q.z, q.w };
        // v > MAX_THRESH -> OCC  (collision_checking.cpp:214,220).  Cells outside the image were zero-filled: never OCC
        const uint32_t b_occ = gather16(bytes_gt(v[0], hi_k7, hi_sel), bytes_gt(v[1], hi_k7, hi_sel),
                                        bytes_gt(v[2], hi_k7, hi_sel), bytes_gt(v[3], hi_k7, hi_sel));
        // !(v < MIN_THRESH) -> not FREE  (:213,219)
        uint32_t b_nf = gather16(bytes_gt(v[0], lo_k7, lo_sel), bytes_gt(v[1], lo_k7, lo_sel),
                                 bytes_gt(v[2], lo_k7, lo_sel), bytes_gt(v[3], lo_k7, lo_sel));
        if (lo == 0u)
        {
          // nothing is below MIN_THRESH = 0: every cell inside the image is at least UNKNOWN (the zero fill does not tell)
          const int y = y0 - r + rl, x = xh0 + c * 16;
          b_nf = ((unsigned)y < (unsigned)dim && x >= 0 && x < dim) ? 0xFFFFu : 0u;
          if (LOADW != 16 && x + 16 > dim && x < dim) b_nf &= (1u << (dim - x)) - 1u;
        }
        b_nf |= b_occ;  // setTo(OCC) runs after setTo(FREE): OCC wins
        sP16[rl * DT_SH + c + 1] = (uint16_t)b_occ;
        sP16[(rows_in + rl) * DT_SH + c + 1] = (uint16_t)b_nf;
      }
      rl += STEP_R; c += STEP_C;
      if (c >= DT_NH) { c -= DT_NH; ++rl; }
    }
  }
  __syncthreads();

  // ---- phase V: vertical ORs.  A cell at horizontal distance w of the centre column is covered by the SE rows
  //      |dy| <= hmax(w), so  dilated(y) = OR_w shift(+-w) of V_hmax(w)(y),  V_h(y) = OR of the plane rows y-h .. y+h,
  //      stored at the distinct hmax values (7 for the 21 x 21 disk).  The horizontal work then runs on the 64 output
  //      rows only.  One item = two staged words (64-bit accesses) of TWO neighbouring output rows y, y+1: they share
  //      the core C_h = OR of the rows y-h+1 .. y+h,  V_h(y) = C_h | row(y-h),  V_h(y+1) = C_h | row(y+1+h), and the
  //      two rows just used extend the core to C_(h+1): 11 shared-memory loads per output row instead of 21.
  //      Lanes run along the rows first (8 row pairs x pitch 20 words visit every fourth bank pair): conflict free.
  //      The two planes go through V and B one after the other: half the shared memory, 7 CTAs per SM -- the 1024
  //      tiles of a 4096^2 patch are resident at once (148 x 7 = 1036) instead of running as 1.4 waves.
  constexpr int RP = DT_SW / 2;  // 64-bit pairs per row
  static_assert(DT_TW == 8 && DT_TH == 64 && DT_THREADS == 256, "item decoding of phases V and B");
  for (int pl = 0; pl < 2; ++pl)
  {
    if (tid < (DT_TH / 2) * RP)
    {
      const int t = tid >> 3, cp = t % RP, y = 2 * ((t / RP) * 8 + (tid & 7));
      const uint2* src = reinterpret_cast<const uint2*>(sP) + (pl * rows_in + r + y) * RP + cp;
      uint2* dst = reinterpret_cast<uint2*>(sV) + y * RP + cp;
      const uint2 r0 = src[0], r1 = src[RP];
      uint2 core = make_uint2(r0.x | r1.x, r0.y | r1.y);  // C_1
      int lv = 0;
      if (!SPEC10 && se.lvl_h[0] == 0)
      {
        dst[0] = r0;
        dst[RP] = r1;
        lv = 1;
      }
      if (SPEC10)
      {
#pragma unroll
        for (int h = 1; h <= K10_R; ++h)
        {
          const uint2 u = src[-h * RP], d = src[(h + 1) * RP];
          if (lv < K10_NLVL && K10_LVL_H[lv] == h)  // compile-time after unrolling
          {
            dst[lv * (lvl_stride / 2)] = make_uint2(core.x | u.x, core.y | u.y);
            dst[lv * (lvl_stride / 2) + RP] = make_uint2(core.x | d.x, core.y | d.y);
            ++lv;
          }
          core.x |= u.x | d.x;
          core.y |= u.y | d.y;
        }
      }
      else
      {
        for (int h = 1; h <= r; ++h)
        {
          const uint2 u = src[-h * RP], d = src[(h + 1) * RP];
          if (lv < n_lvl && se.lvl_h[lv] == h)
          {
            dst[lv * (lvl_stride / 2)] = make_uint2(core.x | u.x, core.y | u.y);
            dst[lv * (lvl_stride / 2) + RP] = make_uint2(core.x | d.x, core.y | d.y);
            ++lv;
          }
          core.x |= u.x | d.x;
          core.y |= u.y | d.y;
        }
      }
    }
    __syncthreads();

  // ---- phase B: horizontal shifts of the vertical ORs.  One item = two neighbouring output words (staged words
  //      2p+1, 2p+2) from the aligned pairs p and p+1 of every level.  The 8 rows a warp works on are 4 apart:
  //      4 rows x pitch 10 words tile the 32 banks exactly (10 * 4 = 40 = 8 mod 32), loads and stores are conflict free.
  {
    const uint2* sV64 = reinterpret_cast<const uint2*>(sV);
    uint2* sOut64 = reinterpret_cast<uint2*>(sOut);
    const int warp = tid >> 5, lane = tid & 31;
    const int rl = (warp >> 2) * 32 + (warp & 3) + 4 * (lane >> 2), p = lane & 3;
    {
      const uint2* base = sV64 + rl * RP + p;
      uint32_t out0, out1;
      if (SPEC10)
      {
        uint2 A[K10_NLVL], B[K10_NLVL];
#pragma unroll
        for (int l = 0; l < K10_NLVL; ++l)
        {
          A[l] = base[l * (lvl_stride / 2)];
          B[l] = base[l * (lvl_stride / 2) + 1];
        }
        out0 = A[K10_LVL_OF_W[0]].y;
        out1 = B[K10_LVL_OF_W[0]].x;
#pragma unroll
        for (int w = 1; w <= K10_R; ++w)
        {
          const int l = K10_LVL_OF_W[w];  // cells at distance exactly w
          out0 |= __funnelshift_l(A[l].x, A[l].y, w) | __funnelshift_r(A[l].y, B[l].x, w);
          out1 |= __funnelshift_l(A[l].y, B[l].x, w) | __funnelshift_r(B[l].x, B[l].y, w);
        }
      }
      else
      {
        int cur = se.lvl_of_w[0];
        uint2 A = base[cur * (lvl_stride / 2)], B = base[cur * (lvl_stride / 2) + 1];
        out0 = A.y;
        out1 = B.x;
        for (int w = 1; w <= r; ++w)
        {
          const int lv = se.lvl_of_w[w];
          if (lv != cur)
          {
            cur = lv;
            A = base[cur * (lvl_stride / 2)];
            B = base[cur * (lvl_stride / 2) + 1];
          }
          out0 |= __funnelshift_l(A.x, A.y, w) | __funnelshift_r(A.y, B.x, w);
          out1 |= __funnelshift_l(A.y, B.x, w) | __funnelshift_r(B.x, B.y, w);
        }
      }
      sOut64[(pl * DT_TH + rl) * RP + p] = make_uint2(out0, out1);
    }
  }
  __syncthreads();
  }

  // ---- phase C: bits -> bytes, coalesced 16-byte stores (a warp writes 2 x 256 consecutive cells).  The four rows of
  //      a thread are independent and branch free up to the store: their shared-memory loads overlap.
  {
    constexpr int NJ = DT_TH * DT_TW * 2 / DT_THREADS, RSTEP = DT_THREADS / 16;
    const int sub = tid & 1, wl = (tid >> 1) & (DT_TW - 1), rl0 = tid >> 4;
    const int x = (w0 + wl) * 32 + sub * 16;
    uint8_t* dst0 = safety + (uint32_t)((y0 + rl0) * dim + x);
    uint32_t p1[NJ], p2[NJ];
#pragma unroll
    for (int j = 0; j < NJ; ++j)
    {
      p2[j] = sOut[(rl0 + j * RSTEP) * DT_SW + wl] >> (sub * 16);
      p1[j] = sOut[(DT_TH + rl0 + j * RSTEP) * DT_SW + wl] >> (sub * 16);
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j)
    {
      uint32_t o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k] = s_lut[((p1[j] >> (4 * k)) & 0xFu) | (((p2[j] >> (4 * k)) & 0xFu) << 4)];
      uint8_t* dst = dst0 + (uint32_t)(j * RSTEP * dim);
      if (y0 + rl0 + j * RSTEP < dim && x < dim)
      {
        if (LOADW == 16)
          *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
        else if (LOADW == 4)
        {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (x + 4 * k < dim) reinterpret_cast<uint32_t*>(dst)[k] = o[k];
        }
        else
        {
          const int n = min(16, dim - x);
          for (int k = 0; k < n; ++k) dst[k] = (uint8_t)(o[k >> 2] >> (8 * (k & 3)));
        }
      }
    }
  }

  // ---- P: 4x4 max-pool of the tile (tile origin is a multiple of 4 in both axes; trailing
  //      patch_dim mod 4 rows / columns are ignored, max_pool.cu:33-36); one thread = 8 pooled cells
  for (int it = tid; it < (DT_TH / 4) * DT_TW; it += DT_THREADS)
  {
    const int wl = it & (DT_TW - 1), yl = it >> 3;
    const int X = (w0 + wl) * 8, Y = y0 / 4 + yl;
    if (X >= astar_dim || Y >= astar_dim) continue;
    uint32_t o2 = 0, o1 = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
      o2 |= sOut[(yl * 4 + j) * DT_SW + wl];
      o1 |= sOut[(DT_TH + yl * 4 + j) * DT_SW + wl];
    }
    // nibble k of n = any bit of nibble k set, per plane; cell = n1 + n2
    o1 |= o1 >> 1; o1 |= o1 >> 2; o1 &= 0x11111111u;
    o2 |= o2 >> 1; o2 |= o2 >> 2; o2 &= 0x11111111u;
    const uint32_t s = o1 + o2;  // 8 nibbles, each 0..2
    uint32_t lo8 = s & 0xFFFFu, hi8 = s >> 16;
    lo8 = (lo8 | (lo8 << 8)) & 0x00FF00FFu; lo8 = (lo8 | (lo8 << 4)) & 0x0F0F0F0Fu;
    hi8 = (hi8 | (hi8 << 8)) & 0x00FF00FFu; hi8 = (hi8 | (hi8 << 4)) & 0x0F0F0F0Fu;
    uint8_t* dst = astar + (size_t)Y * astar_dim + X;
    if ((astar_dim & 7) == 0)
      *reinterpret_cast<uint2*>(dst) = make_uint2(lo8, hi8);
    else
    {
      const int n = min(8, astar_dim - X);
      for (int k = 0; k < n; ++k) dst[k] = (uint8_t)((k < 4 ? lo8 : hi8) >> (8 * (k & 3)));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host entry points
// ---------------------------------------------------------------------------------------------

extern "C" int ga_patch_insert(ga_ctx* c, const uint8_t* window, int ox, int oy, int dim)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, window && dim > 0, GA_ERR_ARG, "bad window");
  GA_REQUIRE(c, c->patch_dim > 0, GA_ERR_STATE, "no patch (call ga_patch_reset)");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int pd = c->patch_dim;
  // collision_checking.cpp:287-290: the whole window is dropped when it overhangs in x
  if (ox < 0 || ox + dim > pd) return GA_OK;
  // :297-301: rows outside the patch are skipped
  int r0 = 0, r1 = dim;  // [r0, r1) rows of the window that land inside
  if (oy < 0) r0 = -oy;
  if (oy + dim > pd) r1 = pd - oy;
  if (r1 <= r0) return GA_OK;
  GA_CUDA(c, cudaMemcpy2DAsync(c->d_patch + (size_t)(oy + r0) * pd + ox, (size_t)pd, window + (size_t)r0 * dim,
                               (size_t)dim, (size_t)dim, (size_t)(r1 - r0), cudaMemcpyHostToDevice, c->stream));
  return GA_OK;
}

// The patch as a 2-D tensor of 32-bit words for cp.async.bulk.tensor: box = the DT_NH halves (x 4 words) a tile row
// needs x rows_in rows, zero fill outside.  cuTensorMapEncodeTiled is a driver entry point; it is looked up through
// the runtime so that the library keeps linking against cudart only.  false = not available: the kernel then stages
// with per-thread cp.async (same result).
static bool patch_tensor_map(ga_ctx* c, int dim, int rows_in)
{
  typedef CUresult (*encode_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                               const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static encode_t encode = nullptr;
  static bool looked = false;
  if (!looked)
  {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      encode = reinterpret_cast<encode_t>(fn);
    else
      (void)cudaGetLastError();
    looked = true;
  }
  c->tm_ptr = c->d_patch;
  c->tm_dim = dim;
  c->tm_rows = rows_in;
  if (!encode) return false;
  CUtensorMap tm;
  const cuuint64_t gdim[2] = { (cuuint64_t)(dim / 4), (cuuint64_t)dim };
  const cuuint64_t gstride[1] = { (cuuint64_t)dim };  // bytes between rows: a multiple of 16 (LOADW == 16)
  const cuuint32_t box[2] = { (cuuint32_t)(DT_NH * 4), (cuuint32_t)rows_in };
  const cuuint32_t estr[2] = { 1, 1 };
  if (encode(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, c->d_patch, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  static_assert(sizeof(tm) <= sizeof(c->tm_patch), "tensor map storage");
  memcpy(c->tm_patch, &tm, sizeof(tm));
  return true;
}

extern "C" int ga_set_safety_staging(ga_ctx* c, int use_tma)
{
  if (!c) return GA_ERR_ARG;
  c->use_tma = use_tma != 0;
  return GA_OK;
}

extern "C" int ga_build_safety(ga_ctx* c)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp, GA_ERR_STATE, "no patch / params");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int dim = c->patch_dim;
  // min_thresh_ / max_thresh_ are uint8_t (collision_checking.hpp:51-52)
  const uint32_t lo = (uint8_t)c->cp.min_thresh, hi = (uint8_t)c->cp.max_thresh;
  {
    ga_launch_scope ls(c, GA_K_DILATE);
    const int rows_in = DT_TH + 2 * c->se.r;
    // thresholded planes + the vertical ORs of every level (one plane at a time) + the dilated planes
    const size_t stage = (size_t)rows_in * DT_NH * 16;  // raw bytes of the tile + halo, staged where the vertical ORs and the dilated planes go later
    const size_t smem = sizeof(uint32_t) * (size_t)((2 * rows_in * DT_SW + 31) & ~31) + std::max(stage, sizeof(uint32_t) * (size_t)(c->se.n_lvl + 2) * DT_TH * DT_SW);
    const int words = (dim + 31) / 32;
    dim3 grid((words + DT_TW - 1) / DT_TW, (dim + DT_TH - 1) / DT_TH);
    // the shipped 21 x 21 disk has compile-time tables
    bool spec10 = c->se.r == K10_R && c->se.n_lvl == K10_NLVL;
    for (int l = 0; spec10 && l < K10_NLVL; ++l) spec10 = c->se.lvl_h[l] == K10_LVL_H[l];
    for (int w = 0; spec10 && w <= K10_R; ++w) spec10 = c->se.lvl_of_w[w] == K10_LVL_OF_W[w];
    const int loadw = (dim & 15) == 0 ? 16 : (dim & 3) == 0 ? 4 : 1;
    // v > MAX_THRESH;  !(v < MIN_THRESH) <=> v > MIN_THRESH - 1 (MIN_THRESH = 0: nothing is below, handled in the kernel)
    const uint32_t lm = (lo - 1u) & 0xFFu;
    const uint4 thr = make_uint4(0x01010101u * (127u - (hi & 127u)), hi >= 128u ? 0xFFFFFFFFu : 0u,
                                 0x01010101u * (127u - (lm & 127u)), lm >= 128u ? 0xFFFFFFFFu : 0u);
    // 16-byte row pitch: the tile + halo comes in as one bulk tensor copy (TMA), described once per patch
    CUtensorMap tm;
    memset(&tm, 0, sizeof(tm));
    bool tma = false;
    if (loadw == 16 && c->use_tma)
    {
      if (c->tm_ptr != c->d_patch || c->tm_dim != dim || c->tm_rows != rows_in) c->tm_ok = patch_tensor_map(c, dim, rows_in);
      tma = c->tm_ok;
      if (tma) memcpy(&tm, c->tm_patch, sizeof(tm));
    }
    typedef void (*kern_t)(const uint8_t*, int, uint8_t*, uint8_t*, int, const ga_se, const uint32_t, const uint4, const CUtensorMap);
    const kern_t kern = tma      ? (spec10 ? safety_kernel<true, 16, true> : safety_kernel<false, 16, true>)
                        : spec10 ? (loadw == 16 ? safety_kernel<true, 16, false> : loadw == 4 ? safety_kernel<true, 4, false> : safety_kernel<true, 1, false>)
                                 : (loadw == 16 ? safety_kernel<false, 16, false> : loadw == 4 ? safety_kernel<false, 4, false> : safety_kernel<false, 1, false>);
    GA_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    GA_CUDA(c, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    kern<<<grid, DT_THREADS, smem, c->stream>>>(c->d_patch, dim, c->d_safety, c->d_astar, c->astar_dim, c->se, lo, thr, tm);
  }
  GA_CUDA(c, cudaGetLastError());
  return GA_OK;
}

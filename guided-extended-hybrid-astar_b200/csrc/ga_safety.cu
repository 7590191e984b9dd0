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
// dilation.  The elliptical structuring element is a stack of horizontal spans (one half width per
// row dy).  A CTA owns a tile of 256 x 64 cells:
//   0  reads the raw bytes of the tile + halo (16-byte loads), thresholds and packs them into the
//      two bit planes in shared memory -- the planes never exist in global memory;
//   A  per plane word, the span dilations D_w = D_(w-1) | cells << w | cells >> w for w = 1..r by
//      32-bit funnel shifts; the distinct half widths the SE uses are kept;
//   B  vertical OR over the 2r+1 SE rows, four words per thread (128-bit shared loads);
//   C  bits -> bytes through a 256-entry table (4 cells per lookup), 16-byte coalesced stores;
//   P  4x4 max-pool of the tile from the same shared words (8 pooled cells per thread).
// HBM traffic is the compulsory 1 B/cell read + 1 B/cell write + 1/16 B/cell pooled write; the halo
// re-reads (1.6x) hit L2.
#include "ga_internal.cuh"

#define DT_TW 8          // tile width in words (256 cells)
#define DT_TH 64         // tile height in rows
#define DT_SW (DT_TW + 2)  // staged plane words per row (one halo word each side)
#define DT_THREADS 256

// SWAR byte compare: bit 7 of byte k of the result = (byte k of x > c), any c in 0..255.
// k7 = 0x01010101 * (127 - (c & 127)): adding it to the low 7 bits carries into bit 7 iff they exceed c & 127.
__device__ __forceinline__ uint32_t bytes_gt(uint32_t x, uint32_t k7, bool c_high)
{
  const uint32_t m = (x & 0x7F7F7F7Fu) + k7;
  return (c_high ? (m & x) : (m | x)) & 0x80808080u;
}
// bits 7, 15, 23, 31 -> bits 0..3
__device__ __forceinline__ uint32_t nibble_of_high_bits(uint32_t m) { return ((m >> 7) * 0x01020408u) >> 24; }

// The shipped footprint (default vehicle, SAFETY_DISTANCE_M 0.1) gives the 21 x 21 disk: radius 10, half widths
// 10,10,10,10,9,9,8,7,6,4,0 for |dy| = 0..10.  Its column view as compile-time tables lets phases V and B unroll
// completely (SPEC10); any other element runs the same kernel with the tables read from `se`.
__device__ constexpr int K10_R = 10, K10_NLVL = 7;
__device__ constexpr int K10_LVL_H[K10_NLVL] = { 3, 5, 6, 7, 8, 9, 10 };
__device__ constexpr int K10_LVL_OF_W[K10_R + 1] = { 6, 5, 5, 5, 5, 4, 4, 3, 2, 1, 0 };

template <bool SPEC10>
__global__ void __launch_bounds__(DT_THREADS)
safety_kernel(const uint8_t* __restrict__ patch, int dim, uint8_t* __restrict__ safety, uint8_t* __restrict__ astar,
              int astar_dim, const ga_se se, uint32_t lo4, uint32_t hi4)
{
  extern __shared__ __align__(16) uint32_t smem[];
  __shared__ uint32_t s_lut[256];   // (P1 nibble | P2 nibble << 4) -> 4 cell bytes
  const int r = SPEC10 ? K10_R : se.r;
  const int n_lvl = SPEC10 ? K10_NLVL : se.n_lvl;
  const int rows_in = DT_TH + 2 * r;
  uint32_t* sP = smem;                                   // [plane][rows_in][DT_SW]   thresholded bit planes
  uint32_t* sV = sP + 2 * rows_in * DT_SW;               // [plane][level][DT_TH][DT_SW]  vertical ORs over |dy| <= hmax(level)
  uint32_t* sOut = sV + 2 * n_lvl * DT_TH * DT_SW;       // [plane][DT_TH][DT_TW]   dilated planes
  constexpr int lvl_stride = DT_TH * DT_SW;
  const int plane_stride = n_lvl * lvl_stride;
  const int w0 = blockIdx.x * DT_TW, y0 = blockIdx.y * DT_TH;
  const int tid = threadIdx.x;
  const bool vec = (dim & 15) == 0;
  const uint32_t lo = lo4 & 0xFFu, hi = hi4 & 0xFFu;
  const uint32_t hi_k7 = 0x01010101u * (127u - (hi & 127u)), lo_k7 = 0x01010101u * (127u - ((lo - 1u) & 127u));
  const bool hi_high = hi >= 128u, lo_high = (lo - 1u) >= 128u && lo != 0u;

  {
    // table: byte k = ((idx >> k) & 1) + ((idx >> (4 + k)) & 1)
    const uint32_t a = tid & 15, b = tid >> 4;
    s_lut[tid] = (((a & 0xFu) * 0x00204081u) & 0x01010101u) + (((b & 0xFu) * 0x00204081u) & 0x01010101u);
  }

  // ---- phase 0: threshold + pack.  One item = 16 cells (half a plane word); two neighbouring
  //      lanes make a word.  Cells outside the image are 0 in both planes (OpenCV ignores them).
  const int n_items = rows_in * DT_SW * 2;
  constexpr int P0_MAX = ((DT_TH + 2 * GA_MAX_DISK_R) * DT_SW * 2 + DT_THREADS - 1) / DT_THREADS;  // items per thread, largest SE
  if (vec)
  {
    // dim % 16 == 0: an item is 16 aligned bytes, fully inside or fully outside the image.  All loads
    // of a thread are issued before the first one is used (one memory round trip per tile).
    uint4 v[P0_MAX];
#pragma unroll
    for (int j = 0; j < P0_MAX; ++j)
    {
      const int it = tid + j * DT_THREADS;
      const int half = it & 1, wi = (it >> 1) % DT_SW, rl = (it >> 1) / DT_SW;
      const int y = y0 - r + rl, x = (w0 - 1 + wi) * 32 + half * 16;
      v[j] = make_uint4(0, 0, 0, 0);  // below any MIN_THRESH > 0: FREE, the neutral element of the max filter
      if (it < n_items && y >= 0 && y < dim && x >= 0 && x < dim) v[j] = __ldg(reinterpret_cast<const uint4*>(patch + (size_t)y * dim + x));
    }
#pragma unroll
    for (int j = 0; j < P0_MAX; ++j)
    {
      const int it = tid + j * DT_THREADS;
      if (it - (tid & 31) >= n_items) break;  // whole warps leave together: the shuffle below needs all lanes
      const int half = it & 1, wi = (it >> 1) % DT_SW, rl = (it >> 1) / DT_SW;
      const int y = y0 - r + rl, x = (w0 - 1 + wi) * 32 + half * 16;
      const bool inside = it < n_items && y >= 0 && y < dim && x >= 0 && x < dim;
      const uint32_t w[4] = { v[j].x, v[j].y, v[j].z, v[j].w };
      // 16 byte compares -> 16 bits: the compare leaves 0x80 in every byte that passes; a byte-wise dot product with the
      // weights 1, 2, 4, 8 (16 .. 128 for the next word) adds up 128 x the bit pattern -- one DP4A per word instead of
      // shift / multiply / shift per word and the shift-ors that assemble the nibbles
      const uint32_t o_lo = __dp4a(bytes_gt(w[1], hi_k7, hi_high), 0x80402010u, __dp4a(bytes_gt(w[0], hi_k7, hi_high), 0x08040201u, 0u));
      const uint32_t o_hi = __dp4a(bytes_gt(w[3], hi_k7, hi_high), 0x80402010u, __dp4a(bytes_gt(w[2], hi_k7, hi_high), 0x08040201u, 0u));
      const uint32_t n_lo = __dp4a(bytes_gt(w[1], lo_k7, lo_high), 0x80402010u, __dp4a(bytes_gt(w[0], lo_k7, lo_high), 0x08040201u, 0u));
      const uint32_t n_hi = __dp4a(bytes_gt(w[3], lo_k7, lo_high), 0x80402010u, __dp4a(bytes_gt(w[2], lo_k7, lo_high), 0x08040201u, 0u));
      uint32_t b_occ = (o_lo >> 7) | ((o_hi >> 7) << 8);                                  // v > MAX_THRESH -> OCC  (collision_checking.cpp:214,220)
      uint32_t b_nf = (lo == 0) ? 0xFFFFu : ((n_lo >> 7) | ((n_hi >> 7) << 8));           // !(v < MIN_THRESH), i.e. v > MIN_THRESH - 1  (:213,219); nothing is below 0
      if (!inside) b_occ = b_nf = 0;
      b_nf |= b_occ;  // setTo(OCC) runs after setTo(FREE): OCC wins
      // lanes 2k, 2k+1 hold the halves of one word
      const uint32_t o_occ = __shfl_xor_sync(0xFFFFFFFFu, b_occ, 1), o_nf = __shfl_xor_sync(0xFFFFFFFFu, b_nf, 1);
      if (!half && it < n_items)
      {
        sP[rl * DT_SW + wi] = b_occ | (o_occ << 16);
        sP[(rows_in + rl) * DT_SW + wi] = b_nf | (o_nf << 16);
      }
    }
  }
  else
  {
    for (int it = tid; it < n_items / 2; it += DT_THREADS)  // one item = one plane word, byte loads
    {
      const int wi = it % DT_SW, rl = it / DT_SW;
      const int y = y0 - r + rl, x = (w0 - 1 + wi) * 32;
      uint32_t b_occ = 0, b_nf = 0;
      if (y >= 0 && y < dim && x >= 0 && x < dim)
      {
        const uint8_t* p = patch + (size_t)y * dim + x;
        const int n = min(32, dim - x);
        for (int k = 0; k < n; ++k)
        {
          const uint32_t c = __ldg(p + k);
          b_occ |= (c > hi ? 1u : 0u) << k;
          b_nf |= (c >= lo ? 1u : 0u) << k;
        }
      }
      sP[rl * DT_SW + wi] = b_occ;
      sP[(rows_in + rl) * DT_SW + wi] = b_nf | b_occ;
    }
  }
  __syncthreads();

  // ---- phase V: vertical ORs.  A cell at horizontal distance w of the centre column is covered by the SE rows
  //      |dy| <= hmax(w), so  dilated(y) = OR_w shift(+-w) of V_hmax(w)(y),  V_h(y) = OR of the plane rows y-h .. y+h.
  //      One item = one staged word of one output row: the cumulative OR over h = 0 .. r, stored at the distinct hmax
  //      values (7 for the 21 x 21 disk).  The horizontal work then runs on the 64 output rows only and needs no
  //      per-input-row span chains: 40 + 55 instructions per output word against 260 for spans-then-OR.
  for (int it = tid; it < 2 * DT_TH * DT_SW; it += DT_THREADS)
  {
    const int pl = it >= DT_TH * DT_SW, rem = it - pl * (DT_TH * DT_SW);  // rem = rl * DT_SW + wi: rows are contiguous
    const uint32_t* src = sP + pl * rows_in * DT_SW + r * DT_SW + rem;
    uint32_t* dst = sV + pl * plane_stride + rem;
    uint32_t acc = src[0];
    if (SPEC10)
    {
      int lv = 0;
#pragma unroll
      for (int h = 1; h <= K10_R; ++h)
      {
        acc |= src[h * DT_SW] | src[-h * DT_SW];
        if (lv < K10_NLVL && K10_LVL_H[lv] == h)  // compile-time after unrolling
        {
          dst[lv * lvl_stride] = acc;
          ++lv;
        }
      }
    }
    else
    {
      int lv = 0;
      if (se.lvl_h[0] == 0)
      {
        dst[0] = acc;
        lv = 1;
      }
      for (int h = 1; h <= r; ++h)
      {
        acc |= src[h * DT_SW] | src[-h * DT_SW];
        if (lv < n_lvl && se.lvl_h[lv] == h)
        {
          dst[lv * lvl_stride] = acc;
          ++lv;
        }
      }
    }
  }
  __syncthreads();

  // ---- phase B: horizontal shifts of the vertical ORs, one item = one output word
  for (int it = tid; it < 2 * DT_TH * DT_TW; it += DT_THREADS)
  {
    const int wl = it & (DT_TW - 1), rl = (it >> 3) & (DT_TH - 1), pl = it >> 9;
    static_assert(DT_TW == 8 && DT_TH == 64, "item decoding of phase B");
    const uint32_t* base = sV + pl * plane_stride + rl * DT_SW + wl;
    uint32_t out;
    if (SPEC10)
    {
      uint32_t a[K10_NLVL], b[K10_NLVL], c[K10_NLVL];
#pragma unroll
      for (int l = 0; l < K10_NLVL; ++l)
      {
        a[l] = base[l * lvl_stride];
        b[l] = base[l * lvl_stride + 1];
        c[l] = base[l * lvl_stride + 2];
      }
      out = b[K10_LVL_OF_W[0]];
#pragma unroll
      for (int w = 1; w <= K10_R; ++w)
      {
        constexpr int dummy = 0;
        (void)dummy;
        const int l = K10_LVL_OF_W[w];
        out |= __funnelshift_l(a[l], b[l], w) | __funnelshift_r(b[l], c[l], w);  // cells at distance exactly w
      }
    }
    else
    {
      int cur = se.lvl_of_w[0];
      uint32_t a = base[cur * lvl_stride], b = base[cur * lvl_stride + 1], c = base[cur * lvl_stride + 2];
      out = b;
      for (int w = 1; w <= r; ++w)
      {
        const int lv = se.lvl_of_w[w];
        if (lv != cur)
        {
          cur = lv;
          a = base[cur * lvl_stride];
          b = base[cur * lvl_stride + 1];
          c = base[cur * lvl_stride + 2];
        }
        out |= __funnelshift_l(a, b, w) | __funnelshift_r(b, c, w);
      }
    }
    sOut[(pl * DT_TH + rl) * DT_TW + wl] = out;
  }
  __syncthreads();

  // ---- phase C: bits -> bytes, coalesced 16-byte stores (a warp writes 512 consecutive cells)
  for (int it = tid; it < DT_TH * DT_TW * 2; it += DT_THREADS)
  {
    const int sub = it & 1, wl = (it >> 1) & (DT_TW - 1), rl = it >> 4;
    const int y = y0 + rl;
    const int x = (w0 + wl) * 32 + sub * 16;
    if (y >= dim || x >= dim) continue;
    const uint32_t p2 = sOut[rl * DT_TW + wl] >> (sub * 16);
    const uint32_t p1 = sOut[(DT_TH + rl) * DT_TW + wl] >> (sub * 16);
    uint8_t* dst = safety + (size_t)y * dim + x;
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = s_lut[((p1 >> (4 * k)) & 0xFu) | (((p2 >> (4 * k)) & 0xFu) << 4)];
    if (vec)
      *reinterpret_cast<uint4*>(dst) = make_uint4(o[0], o[1], o[2], o[3]);
    else
    {
      const int n = min(16, dim - x);
      for (int k = 0; k < n; ++k) dst[k] = (uint8_t)(o[k >> 2] >> (8 * (k & 3)));
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
      o2 |= sOut[(yl * 4 + j) * DT_TW + wl];
      o1 |= sOut[(DT_TH + yl * 4 + j) * DT_TW + wl];
    }
    // nibble k of n = any bit of nibble k set, per plane; cell = n1 + n2
    o1 |= o1 >> 1; o1 |= o1 >> 2; o1 &= 0x11111111u;
    o2 |= o2 >> 1; o2 |= o2 >> 2; o2 &= 0x11111111u;
    const uint32_t s = o1 + o2;  // 8 nibbles, each 0..2
    uint32_t lo = s & 0xFFFFu, hi = s >> 16;
    lo = (lo | (lo << 8)) & 0x00FF00FFu; lo = (lo | (lo << 4)) & 0x0F0F0F0Fu;
    hi = (hi | (hi << 8)) & 0x00FF00FFu; hi = (hi | (hi << 4)) & 0x0F0F0F0Fu;
    uint8_t* dst = astar + (size_t)Y * astar_dim + X;
    if ((astar_dim & 7) == 0)
      *reinterpret_cast<uint2*>(dst) = make_uint2(lo, hi);
    else
    {
      const int n = min(8, astar_dim - X);
      for (int k = 0; k < n; ++k) dst[k] = (uint8_t)((k < 4 ? lo : hi) >> (8 * (k & 3)));
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

extern "C" int ga_build_safety(ga_ctx* c)
{
  if (!c) return GA_ERR_ARG;
  GA_REQUIRE(c, c->patch_dim > 0 && c->have_cp, GA_ERR_STATE, "no patch / params");
  GA_CUDA(c, cudaSetDevice(c->device));
  const int dim = c->patch_dim;
  // min_thresh_ / max_thresh_ are uint8_t (collision_checking.hpp:51-52)
  const uint32_t lo = (uint8_t)c->cp.min_thresh, hi = (uint8_t)c->cp.max_thresh;
  const uint32_t lo4 = lo * 0x01010101u, hi4 = hi * 0x01010101u;
  {
    ga_launch_scope ls(c, GA_K_DILATE);
    const int rows_in = DT_TH + 2 * c->se.r;
    const size_t smem = sizeof(uint32_t) * ((size_t)2 * rows_in * DT_SW + (size_t)2 * c->se.n_lvl * DT_TH * DT_SW + 2 * DT_TH * DT_TW);
    const int words = (dim + 31) / 32;
    dim3 grid((words + DT_TW - 1) / DT_TW, (dim + DT_TH - 1) / DT_TH);
    // the shipped 21 x 21 disk has compile-time tables
    bool spec10 = c->se.r == K10_R && c->se.n_lvl == K10_NLVL;
    for (int l = 0; spec10 && l < K10_NLVL; ++l) spec10 = c->se.lvl_h[l] == K10_LVL_H[l];
    for (int w = 0; spec10 && w <= K10_R; ++w) spec10 = c->se.lvl_of_w[w] == K10_LVL_OF_W[w];
    if (spec10)
    {
      GA_CUDA(c, cudaFuncSetAttribute(safety_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      safety_kernel<true><<<grid, DT_THREADS, smem, c->stream>>>(c->d_patch, dim, c->d_safety, c->d_astar, c->astar_dim, c->se, lo4, hi4);
    }
    else
    {
      GA_CUDA(c, cudaFuncSetAttribute(safety_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      safety_kernel<false><<<grid, DT_THREADS, smem, c->stream>>>(c->d_patch, dim, c->d_safety, c->d_astar, c->astar_dim, c->se, lo4, hi4);
    }
  }
  GA_CUDA(c, cudaGetLastError());
  return GA_OK;
}

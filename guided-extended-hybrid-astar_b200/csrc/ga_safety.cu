// ga_safety.cu -- occupancy pipeline: patch insertion (K7), threshold + bit-plane pack (K1a),
// disk dilation + byte expansion + 4x4 max-pool (K1b/K2).
//
// Replaces CollisionChecker::passLocalMapData (collision_checking.cpp:269-313),
// CollisionChecker::processSafetyPatch (collision_checking.cpp:193-224) and Pooling::execute
// (LIB/src/cuda_lib/max_pool.cu:21-79).
//
// Design: thresholded cells take the values {FREE 0, UNKNOWN 1, OCC 2} and cv::dilate is a max
// filter, so the whole pipeline is exact on two 1-bit planes: P2 = (cell == OCC) and
// P1 = (cell != FREE), 32 cells per word.  max over the disk = P1' + P2' where ' is binary
// dilation.  The elliptical structuring element is a stack of horizontal spans (one half width
// per row dy); a horizontal span dilation of a bit row is a log-step shift/OR on a 64-bit
// window, the vertical combination an OR of (2r+1) staged words.  HBM traffic is the compulsory
// 1 B/cell read (pack) + 1 B/cell write + 1/16 B/cell pooled write; the 2 x 1/8 B/cell planes
// live in L2.
#include "ga_internal.cuh"

// ---------------------------------------------------------------------------------------------
// K1a: threshold + pack.  CPT cells per thread (16: uint4 loads, 4: u32 loads, 1: byte + ballot).
// ---------------------------------------------------------------------------------------------

__device__ __forceinline__ uint32_t nibble_of_mask(uint32_t m)
{
  // m has 0xFF / 0x00 per byte; returns bit k = byte k set
  return ((m & 0x08040201u) * 0x01010101u) >> 24;
}

template <int CPT>
__global__ void __launch_bounds__(256)
pack_planes_kernel(const uint8_t* __restrict__ patch, int dim, int pitch, uint32_t* __restrict__ occ,
                   uint32_t* __restrict__ nf, uint32_t lo4, uint32_t hi4)
{
  // one warp covers 32 * CPT consecutive cells of one row
  constexpr int LPW = 32 / CPT;  // lanes per 32-cell word
  const int lane = threadIdx.x & 31;
  const int warps_per_row = (pitch * LPW + 31) / 32;
  const long long warp_id = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const int y = (int)(warp_id / warps_per_row);
  if (y >= dim) return;
  const int wr = (int)(warp_id % warps_per_row);
  const int x0 = (wr * 32 + lane) * CPT;  // first cell of this thread
  const uint8_t* row = patch + (size_t)y * dim;

  uint32_t b_occ = 0, b_nf = 0;  // CPT bits each
  if (CPT == 16)
  {
    if (x0 < dim)  // dim % 16 == 0: fully inside
    {
      const uint4 v = *reinterpret_cast<const uint4*>(row + x0);
      const uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
      for (int k = 0; k < 4; ++k)
      {
        b_occ |= nibble_of_mask(__vcmpgtu4(w[k], hi4)) << (4 * k);
        b_nf |= nibble_of_mask(__vcmpgeu4(w[k], lo4)) << (4 * k);
      }
    }
  }
  else if (CPT == 4)
  {
    if (x0 < dim)  // dim % 4 == 0
    {
      const uint32_t w = *reinterpret_cast<const uint32_t*>(row + x0);
      b_occ = nibble_of_mask(__vcmpgtu4(w, hi4));
      b_nf = nibble_of_mask(__vcmpgeu4(w, lo4));
    }
  }
  else
  {
    if (x0 < dim)
    {
      const uint32_t v = row[x0];
      b_occ = v > (hi4 & 0xFF);
      b_nf = v >= (lo4 & 0xFF);
    }
  }
  b_nf |= b_occ;  // setTo(OCC) runs after setTo(FREE) (collision_checking.cpp:219-220): OCC wins
  // gather LPW lanes into one word
  const int sub = lane % LPW;
  uint32_t w_occ = b_occ << (sub * CPT);
  uint32_t w_nf = b_nf << (sub * CPT);
#pragma unroll
  for (int s = 1; s < LPW; s <<= 1)
  {
    w_occ |= __shfl_xor_sync(0xFFFFFFFFu, w_occ, s);
    w_nf |= __shfl_xor_sync(0xFFFFFFFFu, w_nf, s);
  }
  const int word = (wr * 32 + lane) / LPW;
  if (sub == 0 && word < pitch)
  {
    occ[(size_t)y * pitch + word] = w_occ;
    nf[(size_t)y * pitch + word] = w_nf;
  }
}

// ---------------------------------------------------------------------------------------------
// K1b + K2: dilate the two planes with the disk, expand to bytes, pool 4x4.
// CTA tile: TW words (TW*32 cells) x TH rows.
// ---------------------------------------------------------------------------------------------

#define DT_TW 8
#define DT_TH 64
#define DT_THREADS 256

// horizontal span dilation of a 64-bit window: bit j of the result = OR of window bits
// [j + r - w, j + r + w]  (window bit i = cell 32*word - r + i)
__device__ __forceinline__ uint32_t hspan(const uint64_t s1, const uint64_t s2, const uint64_t s4, const uint64_t s8,
                                          const uint64_t s16, const uint64_t s32, int r, int w)
{
  const int L = 2 * w + 1;
  uint64_t sp;
  if (L >= 32) sp = s32 | (s32 >> (L - 32));
  else if (L >= 16) sp = s16 | (s16 >> (L - 16));
  else if (L >= 8) sp = s8 | (s8 >> (L - 8));
  else if (L >= 4) sp = s4 | (s4 >> (L - 4));
  else if (L >= 2) sp = s2 | (s2 >> (L - 2));
  else sp = s1;
  return (uint32_t)(sp >> (r - w));
}

// spread 4 bits to 4 bytes (0/1 each)
__device__ __forceinline__ uint32_t spread4(uint32_t nib)
{
  return ((nib & 0xFu) * 0x00204081u) & 0x01010101u;
}

template <int CPS>  // cells per store: 16 (uint4), 4 (u32), 1 (byte)
__global__ void __launch_bounds__(DT_THREADS)
dilate_pool_kernel(const uint32_t* __restrict__ occ, const uint32_t* __restrict__ nf, int dim, int pitch,
                   uint8_t* __restrict__ safety, uint8_t* __restrict__ astar, int astar_dim, const ga_se se)
{
  extern __shared__ uint32_t smem[];
  const int r = se.r;
  const int nv = se.n_var;
  const int rows_in = DT_TH + 2 * r;
  // sH[plane][variant][row][word]
  uint32_t* sH = smem;
  const int plane_stride = nv * rows_in * DT_TW;
  uint32_t* sOut = smem + 2 * plane_stride;  // [plane][DT_TH][DT_TW]

  const int w0 = blockIdx.x * DT_TW;
  const int y0 = blockIdx.y * DT_TH;

  // ---- phase A: horizontal spans of every staged row
  for (int it = threadIdx.x; it < 2 * rows_in * DT_TW; it += DT_THREADS)
  {
    const int wl = it % DT_TW;
    const int rl = (it / DT_TW) % rows_in;
    const int pl = it / (DT_TW * rows_in);
    const int y = y0 - r + rl;
    const int w = w0 + wl;
    uint32_t a = 0, b = 0, c = 0;
    if (y >= 0 && y < dim)
    {
      const uint32_t* rowp = (pl ? nf : occ) + (size_t)y * pitch;
      if (w - 1 >= 0 && w - 1 < pitch) a = __ldg(rowp + w - 1);
      if (w < pitch) b = __ldg(rowp + w);
      if (w + 1 < pitch) c = __ldg(rowp + w + 1);
    }
    uint32_t* dst = sH + pl * plane_stride + rl * DT_TW + wl;
    if ((a | b | c) == 0)
    {
      for (int v = 0; v < nv; ++v) dst[v * rows_in * DT_TW] = 0;
      continue;
    }
    // window bit i = cell 32*w - r + i, i in [0, 32 + 2r)
    uint64_t win = ((uint64_t)b << r) | ((uint64_t)c << (32 + r));
    if (r) win |= (uint64_t)(a >> (32 - r));
    const uint64_t s1 = win;
    const uint64_t s2 = s1 | (s1 >> 1);
    const uint64_t s4 = s2 | (s2 >> 2);
    const uint64_t s8 = s4 | (s4 >> 4);
    const uint64_t s16 = s8 | (s8 >> 8);
    const uint64_t s32 = s16 | (s16 >> 16);
    for (int v = 0; v < nv; ++v) dst[v * rows_in * DT_TW] = hspan(s1, s2, s4, s8, s16, s32, r, se.var_w[v]);
  }
  __syncthreads();

  // ---- phase B: vertical OR over the SE rows
  for (int it = threadIdx.x; it < 2 * DT_TH * DT_TW; it += DT_THREADS)
  {
    const int wl = it % DT_TW;
    const int rl = (it / DT_TW) % DT_TH;
    const int pl = it / (DT_TW * DT_TH);
    const uint32_t* base = sH + pl * plane_stride + wl;
    uint32_t acc = 0;
    for (int i = 0; i <= 2 * r; ++i) acc |= base[(se.row_var[i] * rows_in + rl + i) * DT_TW];
    sOut[(pl * DT_TH + rl) * DT_TW + wl] = acc;
  }
  __syncthreads();

  // ---- phase C: expand to bytes, coalesced stores of CPS cells per thread
  {
    constexpr int SPW = 32 / CPS;  // stores per word
    for (int it = threadIdx.x; it < DT_TH * DT_TW * SPW; it += DT_THREADS)
    {
      const int sl = it % (DT_TW * SPW);
      const int rl = it / (DT_TW * SPW);
      const int wl = sl / SPW;
      const int sub = sl % SPW;
      const int y = y0 + rl;
      const int x = (w0 + wl) * 32 + sub * CPS;
      if (y >= dim || x >= dim) continue;
      const uint32_t p2 = sOut[rl * DT_TW + wl] >> (sub * CPS);
      const uint32_t p1 = sOut[(DT_TH + rl) * DT_TW + wl] >> (sub * CPS);
      uint8_t* dst = safety + (size_t)y * dim + x;
      if (CPS == 16)
      {
        uint4 o;
        o.x = spread4(p1) + spread4(p2);
        o.y = spread4(p1 >> 4) + spread4(p2 >> 4);
        o.z = spread4(p1 >> 8) + spread4(p2 >> 8);
        o.w = spread4(p1 >> 12) + spread4(p2 >> 12);
        *reinterpret_cast<uint4*>(dst) = o;
      }
      else if (CPS == 4)
      {
        *reinterpret_cast<uint32_t*>(dst) = spread4(p1) + spread4(p2);
      }
      else
      {
        *dst = (uint8_t)((p1 & 1u) + (p2 & 1u));
      }
    }
  }

  // ---- K2: 4x4 max-pool of the tile (tile origin is a multiple of 4 in both axes)
  for (int it = threadIdx.x; it < (DT_TH / 4) * DT_TW * 8; it += DT_THREADS)
  {
    const int xl = it % (DT_TW * 8);
    const int yl = it / (DT_TW * 8);
    const int X = w0 * 8 + xl;
    const int Y = y0 / 4 + yl;
    if (X >= astar_dim || Y >= astar_dim) continue;
    const int wl = xl / 8;
    const int sh = (xl % 8) * 4;
    uint32_t o2 = 0, o1 = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j)
    {
      o2 |= sOut[(yl * 4 + j) * DT_TW + wl];
      o1 |= sOut[(DT_TH + yl * 4 + j) * DT_TW + wl];
    }
    const uint32_t v = (((o1 >> sh) & 0xFu) ? 1u : 0u) + (((o2 >> sh) & 0xFu) ? 1u : 0u);
    astar[(size_t)Y * astar_dim + X] = (uint8_t)v;
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
  const int pitch = c->plane_pitch;
  // min_thresh_ / max_thresh_ are uint8_t (collision_checking.hpp:51-52)
  const uint32_t lo = (uint8_t)c->cp.min_thresh, hi = (uint8_t)c->cp.max_thresh;
  const uint32_t lo4 = lo * 0x01010101u, hi4 = hi * 0x01010101u;
  {
    ga_launch_scope ls(c, GA_K_PACK);
    const int cpt = (dim % 16 == 0) ? 16 : (dim % 4 == 0) ? 4 : 1;
    const int lpw = 32 / cpt;
    const long long warps = (long long)dim * ((pitch * lpw + 31) / 32);
    const unsigned blocks = (unsigned)((warps * 32 + 255) / 256);
    if (cpt == 16)
      pack_planes_kernel<16><<<blocks, 256, 0, c->stream>>>(c->d_patch, dim, pitch, c->d_plane_occ, c->d_plane_nf, lo4, hi4);
    else if (cpt == 4)
      pack_planes_kernel<4><<<blocks, 256, 0, c->stream>>>(c->d_patch, dim, pitch, c->d_plane_occ, c->d_plane_nf, lo4, hi4);
    else
      pack_planes_kernel<1><<<blocks, 256, 0, c->stream>>>(c->d_patch, dim, pitch, c->d_plane_occ, c->d_plane_nf, lo4, hi4);
  }
  GA_CUDA(c, cudaGetLastError());
  {
    ga_launch_scope ls(c, GA_K_DILATE);
    const int rows_in = DT_TH + 2 * c->se.r;
    const size_t smem = sizeof(uint32_t) * ((size_t)2 * c->se.n_var * rows_in * DT_TW + 2 * DT_TH * DT_TW);
    dim3 grid((pitch + DT_TW - 1) / DT_TW, (dim + DT_TH - 1) / DT_TH);
    const int cps = (dim % 16 == 0) ? 16 : (dim % 4 == 0) ? 4 : 1;
    if (cps == 16)
    {
      GA_CUDA(c, cudaFuncSetAttribute(dilate_pool_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      dilate_pool_kernel<16><<<grid, DT_THREADS, smem, c->stream>>>(c->d_plane_occ, c->d_plane_nf, dim, pitch, c->d_safety,
                                                                     c->d_astar, c->astar_dim, c->se);
    }
    else if (cps == 4)
    {
      GA_CUDA(c, cudaFuncSetAttribute(dilate_pool_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      dilate_pool_kernel<4><<<grid, DT_THREADS, smem, c->stream>>>(c->d_plane_occ, c->d_plane_nf, dim, pitch, c->d_safety,
                                                                    c->d_astar, c->astar_dim, c->se);
    }
    else
    {
      GA_CUDA(c, cudaFuncSetAttribute(dilate_pool_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      dilate_pool_kernel<1><<<grid, DT_THREADS, smem, c->stream>>>(c->d_plane_occ, c->d_plane_nf, dim, pitch, c->d_safety,
                                                                    c->d_astar, c->astar_dim, c->se);
    }
  }
  GA_CUDA(c, cudaGetLastError());
  return GA_OK;
}

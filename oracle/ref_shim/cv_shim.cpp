// TEST INFRASTRUCTURE ONLY -- definitions behind opencv2/imgproc.hpp (the declaration shim).
// Third-party arithmetic the reference gets from OpenCV 4.8.0 is restated here with the
// semantics pinned against cv2 4.13 in tests/golden/cv_golden.npz:
//   getStructuringElement(MORPH_ELLIPSE) and distanceTransform(DIST_L2, DIST_MASK_5) come from
//   the plain-C oracle (gao_ellipse_rows / gao_distance_transform), dilate is a max over the
//   kernel's in-image support written out below (row runs of the kernel, border ignored).
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <mutex>
#include <new>

#include "ga_oracle.h"
#include "opencv2/imgproc.hpp"

namespace cv
{
int shim_dt_variant = 0;

// Matrix storage: uninitialised, and recycled through a small free list -- cv::Mat gets the same from malloc, which
// hands a just-freed block back without touching the kernel.  (A value-initialised std::vector per temporary made the
// element-wise ops of processSafetyPatch cost more in page faults than in arithmetic.)
namespace
{
struct pool_t
{
  std::vector<std::pair<size_t, uchar*>> free_list;
};
pool_t& pool()
{
  // never destroyed: the reference keeps cv::Mat objects in static members (CollisionChecker::dil_kernel_), whose
  // storage comes back here after every other destructor has run
  static pool_t* p = new pool_t();
  return *p;
}
}  // namespace
static std::mutex g_pool_mutex;
std::shared_ptr<uchar> shim_alloc(size_t bytes)
{
  pool_t& p = pool();
  uchar* buf = nullptr;
  std::unique_lock<std::mutex> lock(g_pool_mutex);
  for (size_t i = 0; i < p.free_list.size(); ++i)
    if (p.free_list[i].first == bytes)
    {
      buf = p.free_list[i].second;
      p.free_list.erase(p.free_list.begin() + (long)i);
      break;
    }
  lock.unlock();
  if (!buf) buf = static_cast<uchar*>(std::malloc(bytes ? bytes : 1));
  if (!buf) throw std::bad_alloc();
  return std::shared_ptr<uchar>(buf, [bytes](uchar* q) {
    pool_t& pp = pool();
    std::lock_guard<std::mutex> g(g_pool_mutex);
    if (pp.free_list.size() < 24)
      pp.free_list.emplace_back(bytes, q);
    else
      std::free(q);
  });
}


static void need_u8(const Mat& m, const char* who)
{
  if (m.type() != CV_8UC1) throw std::runtime_error(std::string("cv shim: ") + who + " wants CV_8UC1");
}

Mat Mat::clone() const
{
  Mat m(rows, cols, type_);
  for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
  return m;
}

Mat& Mat::setTo(double value, const Mat& mask)
{
  need_u8(*this, "setTo");
  const uchar v = (uchar)value;
  for (int y = 0; y < rows; ++y)
  {
    uchar* d = ptr<uchar>(y);
    const uchar* m = mask.ptr<uchar>(y);
    for (int x = 0; x < cols; ++x) d[x] = m[x] ? v : d[x];  // select, not a branch: vectorises like OpenCV's copyMask
  }
  return *this;
}

Mat& Mat::operator*=(double s)
{
  if (type_ != CV_32FC1) throw std::runtime_error("cv shim: operator*= wants CV_32FC1");
  // OpenCV: dst = saturate_cast<float>(src * alpha) evaluated in double for a 32F matrix
  for (int y = 0; y < rows; ++y)
  {
    float* d = ptr<float>(y);
    for (int x = 0; x < cols; ++x) d[x] = (float)((double)d[x] * s);
  }
  return *this;
}

Mat operator==(const Mat& m, double v)
{
  Mat out(m.rows, m.cols, CV_8UC1);
  for (int y = 0; y < m.rows; ++y)
  {
    uchar* o = out.ptr<uchar>(y);
    if (m.type() == CV_8UC1)
    {
      const uchar* s = m.ptr<uchar>(y);
      for (int x = 0; x < m.cols; ++x) o[x] = ((double)s[x] == v) ? 255 : 0;
    }
    else
    {
      const float* s = m.ptr<float>(y);
      for (int x = 0; x < m.cols; ++x) o[x] = ((double)s[x] == v) ? 255 : 0;
    }
  }
  return out;
}
Mat operator==(double v, const Mat& m) { return m == v; }

Mat operator-(double s, const Mat& m)
{
  need_u8(m, "scalar - Mat");
  Mat out(m.rows, m.cols, CV_8UC1);
  for (int y = 0; y < m.rows; ++y)
  {
    uchar* o = out.ptr<uchar>(y);
    const uchar* p = m.ptr<uchar>(y);
    if (s == std::floor(s) && s >= 0 && s <= 255)
    {
      const int si = (int)s;  // saturate_cast<uchar>(s - p) for an integral scalar: plain integer arithmetic (SIMD)
      for (int x = 0; x < m.cols; ++x) o[x] = (uchar)(si > p[x] ? si - p[x] : 0);
    }
    else
      for (int x = 0; x < m.cols; ++x) o[x] = (uchar)std::clamp(std::nearbyint(s - (double)p[x]), 0.0, 255.0);
  }
  return out;
}

// cv::compare(u8 Mat, scalar): the operator is chosen once, the row loops are byte compares (SIMD), as in OpenCV.
// A u8 value compared with a scalar: v < t  <=>  v < ceil(t), v > t  <=>  v > floor(t) -- integer thresholds.
void compare(const Mat& src, double value, Mat& dst, int cmpop)
{
  need_u8(src, "compare");
  dst.create(src.rows, src.cols, CV_8UC1);
  // reduce every operator to "lo <= v <= hi" on bytes (possibly negated): two byte compares per cell
  const double fl = std::floor(value), ce = std::ceil(value);
  int lo = 0, hi = 255;
  bool negate = false;
  switch (cmpop)
  {
    case CMP_EQ: lo = hi = (fl == value && value >= 0 && value <= 255) ? (int)value : -1; break;
    case CMP_GT: lo = (int)std::clamp(fl + 1, 0.0, 256.0); break;
    case CMP_GE: lo = (int)std::clamp(ce, 0.0, 256.0); break;
    case CMP_LT: hi = (int)std::clamp(ce - 1, -1.0, 255.0); break;
    case CMP_LE: hi = (int)std::clamp(fl, -1.0, 255.0); break;
    default: lo = hi = (fl == value && value >= 0 && value <= 255) ? (int)value : -1; negate = true; break;
  }
  const bool none = lo > 255 || hi < 0 || lo > hi;
  const uchar l8 = (uchar)std::clamp(lo, 0, 255), h8 = (uchar)std::clamp(hi, 0, 255);
  const uchar yes = negate ? 0 : 255, no = negate ? 255 : 0;
  for (int y = 0; y < src.rows; ++y)
  {
    const uchar* __restrict s = src.ptr<uchar>(y);
    uchar* __restrict d = dst.ptr<uchar>(y);
    if (none)
      std::memset(d, no, (size_t)src.cols);
    else
      for (int x = 0; x < src.cols; ++x) d[x] = (s[x] >= l8 && s[x] <= h8) ? yes : no;
  }
}

void subtract(double s, const Mat& src, Mat& dst)
{
  Mat r = s - src;
  dst = r;
}

void bitwise_and(const Mat& a, const Mat& b, Mat& dst)
{
  need_u8(a, "bitwise_and");
  dst.create(a.rows, a.cols, CV_8UC1);
  for (int y = 0; y < a.rows; ++y)
  {
    const uchar *p = a.ptr<uchar>(y), *q = b.ptr<uchar>(y);
    uchar* d = dst.ptr<uchar>(y);
    for (int x = 0; x < a.cols; ++x) d[x] = p[x] & q[x];
  }
}

Mat getStructuringElement(int shape, Size ksize, Point anchor)
{
  if (shape != MORPH_ELLIPSE || ksize.width != ksize.height || (ksize.width & 1) == 0)
    throw std::runtime_error("cv shim: only odd square MORPH_ELLIPSE kernels");
  const int r = ksize.width / 2;
  if (anchor.x != -1 && (anchor.x != r || anchor.y != r)) throw std::runtime_error("cv shim: anchor must be the centre");
  Mat k(ksize.height, ksize.width, CV_8UC1);
  std::vector<int> hw(ksize.width);
  gao_ellipse_rows(r, hw.data(), k.data);
  return k;
}

// cv::dilate for a kernel whose rows are single centred runs (every MORPH_ELLIPSE kernel is): out-of-image taps
// are ignored, i.e. the image is padded with 0.  In OpenCV's complexity class so that timing the reference through
// this shim is fair: one horizontal running maximum per DISTINCT run width -- 7 for the 21 x 21 disk -- into a plane
// each (built by window doubling, SIMD byte maxima, see hmax_rows), then one vertical pass that takes
// the maximum over the kernel rows of the plane belonging to each row's width (byte maxima, auto-vectorised).
// Kernels with other row shapes fall back to a plain scan (not used by the reference).
// Horizontal running maxima of one row for several half widths at once.  Windows are built by doubling:
// M_1 = row, M_2[i] = max(M_1[i], M_1[i+1]), M_4[i] = max(M_2[i], M_2[i+2]) ... and a window of k cells starting at i
// is max(M_p[i], M_p[i + k - p]) with p the largest power of two <= k.  Every pass is an element-wise byte maximum of
// two shifted arrays, which the compiler turns into SIMD (pmaxub) -- the same cost class as OpenCV's row filters.
struct hmax_scratch
{
  std::vector<uchar> m[6];  // M_1 .. M_32, padded
};
static void hmax_rows(const uchar* in, int W, const std::vector<int>& half_widths, std::vector<uchar*>& out_rows, hmax_scratch& sc)
{
  const int P = 64;               // zero padding on both sides (>= the largest half width, keeps every index in range)
  const int L = W + 2 * P;
  for (auto& v : sc.m)
    if ((int)v.size() != L) v.assign((size_t)L, 0);
  uchar* __restrict m1 = sc.m[0].data();
  std::memset(m1, 0, (size_t)P);
  std::memcpy(m1 + P, in, (size_t)W);
  std::memset(m1 + P + W, 0, (size_t)P);
  int top = 0;                    // highest level built
  int kmax = 1;
  for (int hw : half_widths) kmax = std::max(kmax, 2 * hw + 1);
  for (int lv = 1, p = 2; p <= kmax && lv < 6; ++lv, p <<= 1)
  {
    const uchar* __restrict a = sc.m[lv - 1].data();
    uchar* __restrict d = sc.m[lv].data();
    const int sh = p >> 1, n = L - p + 1;
    for (int i = 0; i < n; ++i) d[i] = std::max(a[i], a[i + sh]);
    for (int i = n; i < L; ++i) d[i] = 0;
    top = lv;
  }
  for (size_t v = 0; v < half_widths.size(); ++v)
  {
    const int hw = half_widths[v], k = 2 * hw + 1;
    int lv = 0, p = 1;
    while (lv < top && p * 2 <= k)
    {
      ++lv;
      p <<= 1;
    }
    const uchar* __restrict a = sc.m[lv].data() + (P - hw);
    const uchar* __restrict b = a + (k - p);
    uchar* __restrict d = out_rows[v];
    for (int x = 0; x < W; ++x) d[x] = std::max(a[x], b[x]);
  }
}

void dilate(const Mat& src, Mat& dst, const Mat& kernel)
{
  need_u8(src, "dilate");
  const int kh = kernel.rows, kw = kernel.cols, ay = kh / 2, ax = kw / 2;
  // half width of the centred run of every kernel row (-1: empty row); anything else -> generic scan
  std::vector<int> row_hw((size_t)kh, -1);
  bool centred = true;
  for (int j = 0; j < kh && centred; ++j)
  {
    const uchar* k = kernel.ptr<uchar>(j);
    int first = -1, last = -1, cnt = 0;
    for (int i = 0; i < kw; ++i)
      if (k[i])
      {
        if (first < 0) first = i;
        last = i;
        ++cnt;
      }
    if (cnt == 0) continue;
    if (cnt != last - first + 1 || ax - first != last - ax) centred = false;
    row_hw[(size_t)j] = last - ax;
  }
  const Mat in = src.clone();  // the reference dilates in place (collision_checking.cpp:223)
  dst.create(src.rows, src.cols, CV_8UC1);
  const int H = in.rows, W = in.cols;
  if (!centred)
  {
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x)
      {
        uchar m = 0;
        for (int j = 0; j < kh; ++j)
          for (int i = 0; i < kw; ++i)
          {
            const int yy = y + j - ay, xx = x + i - ax;
            if (kernel.ptr<uchar>(j)[i] && yy >= 0 && yy < H && xx >= 0 && xx < W) m = std::max(m, in.ptr<uchar>(yy)[xx]);
          }
        dst.ptr<uchar>(y)[x] = m;
      }
    return;
  }
  // distinct widths -> planes of horizontal maxima
  std::vector<int> widths;
  for (int hw : row_hw)
    if (hw >= 0 && std::find(widths.begin(), widths.end(), hw) == widths.end()) widths.push_back(hw);
  std::vector<std::shared_ptr<uchar>> plane_mem;
  std::vector<uchar*> plane;
  for (size_t v = 0; v < widths.size(); ++v)
  {
    plane_mem.push_back(shim_alloc((size_t)H * W));
    plane.push_back(plane_mem.back().get());
  }
  hmax_scratch sc;
  std::vector<uchar*> rows(widths.size());
  for (int y = 0; y < H; ++y)
  {
    for (size_t v = 0; v < widths.size(); ++v) rows[v] = plane[v] + (size_t)y * W;
    hmax_rows(in.ptr<uchar>(y), W, widths, rows, sc);
  }
  std::vector<int> row_plane((size_t)kh, -1);
  for (int j = 0; j < kh; ++j)
    if (row_hw[(size_t)j] >= 0)
      row_plane[(size_t)j] = (int)(std::find(widths.begin(), widths.end(), row_hw[(size_t)j]) - widths.begin());
  for (int y = 0; y < H; ++y)
  {
    uchar* __restrict d = dst.ptr<uchar>(y);
    std::memset(d, 0, (size_t)W);
    for (int j = 0; j < kh; ++j)
    {
      const int yy = y + j - ay;
      if (row_plane[(size_t)j] < 0 || yy < 0 || yy >= H) continue;
      const uchar* __restrict p = plane[(size_t)row_plane[(size_t)j]] + (size_t)yy * W;
      for (int x = 0; x < W; ++x) d[x] = std::max(d[x], p[x]);
    }
  }
}

void distanceTransform(const Mat& src, Mat& dst, int distanceType, int maskSize)
{
  need_u8(src, "distanceTransform");
  if (distanceType != DIST_L2 || maskSize != DIST_MASK_5) throw std::runtime_error("cv shim: DIST_L2 / DIST_MASK_5 only");
  dst.create(src.rows, src.cols, CV_32FC1);
  if (dst.step != (size_t)dst.cols * 4) throw std::runtime_error("cv shim: distanceTransform wants a dense destination");
  gao_distance_transform(src.data, (int)src.step, src.cols, src.rows, reinterpret_cast<float*>(dst.data), shim_dt_variant);
}

void findNonZero(const Mat& src, std::vector<Point>& idx)
{
  need_u8(src, "findNonZero");
  idx.clear();
  for (int y = 0; y < src.rows; ++y)
  {
    const uchar* s = src.ptr<uchar>(y);
    for (int x = 0; x < src.cols; ++x)
      if (s[x]) idx.emplace_back(x, y);
  }
}
}  // namespace cv

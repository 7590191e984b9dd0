// TEST INFRASTRUCTURE ONLY -- definitions behind opencv2/imgproc.hpp (the declaration shim).
// Third-party arithmetic the reference gets from OpenCV 4.8.0 is restated here with the
// semantics pinned against cv2 4.13 in tests/golden/cv_golden.npz:
//   getStructuringElement(MORPH_ELLIPSE) and distanceTransform(DIST_L2, DIST_MASK_5) come from
//   the plain-C oracle (gao_ellipse_rows / gao_distance_transform), dilate is a max over the
//   kernel's in-image support written out below (row runs of the kernel, border ignored).
#include <algorithm>

#include "ga_oracle.h"
#include "opencv2/imgproc.hpp"

namespace cv
{
int shim_dt_variant = 0;

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
    for (int x = 0; x < cols; ++x)
      if (m[x]) d[x] = v;
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
    for (int x = 0; x < m.cols; ++x) o[x] = (uchar)std::clamp(s - (double)p[x], 0.0, 255.0);
  }
  return out;
}

void compare(const Mat& src, double value, Mat& dst, int cmpop)
{
  need_u8(src, "compare");
  dst.create(src.rows, src.cols, CV_8UC1);
  for (int y = 0; y < src.rows; ++y)
  {
    const uchar* s = src.ptr<uchar>(y);
    uchar* d = dst.ptr<uchar>(y);
    for (int x = 0; x < src.cols; ++x)
    {
      const double v = s[x];
      bool r = false;
      switch (cmpop)
      {
        case CMP_EQ: r = v == value; break;
        case CMP_GT: r = v > value; break;
        case CMP_GE: r = v >= value; break;
        case CMP_LT: r = v < value; break;
        case CMP_LE: r = v <= value; break;
        default: r = v != value; break;
      }
      d[x] = r ? 255 : 0;
    }
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
// this shim is fair: one horizontal running maximum (van Herk / Gil-Werman, ~3 comparisons per pixel whatever the
// width) per DISTINCT run width -- 7 for the 21 x 21 disk -- into a plane each, then one vertical pass that takes
// the maximum over the kernel rows of the plane belonging to each row's width (byte maxima, auto-vectorised).
// Kernels with other row shapes fall back to a plain scan (not used by the reference).
static void hmax_row(const uchar* in, uchar* out, int W, int hw, std::vector<uchar>& fwd, std::vector<uchar>& bwd)
{
  if (hw == 0)
  {
    std::memcpy(out, in, (size_t)W);
    return;
  }
  const int k = 2 * hw + 1, P = hw;                 // window, zero padding on both sides
  const int n = ((W + 2 * P + k - 1) / k) * k;      // padded length, a whole number of blocks
  fwd.assign((size_t)n, 0);
  bwd.assign((size_t)n, 0);
  // padded signal p[i] = in[i - P]
  for (int b0 = 0; b0 < n; b0 += k)
  {
    uchar m = 0;
    for (int i = b0; i < b0 + k; ++i)
    {
      const int x = i - P;
      const uchar v = (x >= 0 && x < W) ? in[x] : 0;
      m = std::max(m, v);
      fwd[(size_t)i] = m;
    }
    m = 0;
    for (int i = b0 + k - 1; i >= b0; --i)
    {
      const int x = i - P;
      const uchar v = (x >= 0 && x < W) ? in[x] : 0;
      m = std::max(m, v);
      bwd[(size_t)i] = m;
    }
  }
  // window of output x covers padded [x, x + k - 1]
  for (int x = 0; x < W; ++x) out[x] = std::max(bwd[(size_t)x], fwd[(size_t)(x + k - 1)]);
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
  std::vector<std::vector<uchar>> plane(widths.size(), std::vector<uchar>((size_t)H * W));
  std::vector<uchar> fwd, bwd;
  for (size_t v = 0; v < widths.size(); ++v)
    for (int y = 0; y < H; ++y) hmax_row(in.ptr<uchar>(y), plane[v].data() + (size_t)y * W, W, widths[v], fwd, bwd);
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
      const uchar* __restrict p = plane[(size_t)row_plane[(size_t)j]].data() + (size_t)yy * W;
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

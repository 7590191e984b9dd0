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

void dilate(const Mat& src, Mat& dst, const Mat& kernel)
{
  need_u8(src, "dilate");
  const int kh = kernel.rows, kw = kernel.cols, ay = kh / 2, ax = kw / 2;
  // runs of set taps per kernel row
  struct run { int dy, x0, x1; };
  std::vector<run> runs;
  for (int j = 0; j < kh; ++j)
  {
    const uchar* k = kernel.ptr<uchar>(j);
    for (int i = 0; i < kw;)
    {
      if (!k[i]) { ++i; continue; }
      int e = i;
      while (e + 1 < kw && k[e + 1]) ++e;
      runs.push_back({ j - ay, i - ax, e - ax });
      i = e + 1;
    }
  }
  const Mat in = src.clone();  // the reference dilates in place (collision_checking.cpp:223)
  dst.create(src.rows, src.cols, CV_8UC1);
  const int H = in.rows, W = in.cols;
  // horizontal running max per distinct run width would be faster; values are tiny (0..2) so an
  // early-out at the image maximum keeps the plain scan acceptable for a checker
  uchar vmax = 0;
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) vmax = std::max(vmax, in.ptr<uchar>(y)[x]);
  for (int y = 0; y < H; ++y)
  {
    uchar* d = dst.ptr<uchar>(y);
    for (int x = 0; x < W; ++x)
    {
      uchar m = 0;
      for (const run& r : runs)
      {
        const int yy = y + r.dy;
        if (yy < 0 || yy >= H) continue;
        const int x0 = std::max(x + r.x0, 0), x1 = std::min(x + r.x1, W - 1);
        const uchar* row = in.ptr<uchar>(yy);
        for (int xx = x0; xx <= x1; ++xx) m = std::max(m, row[xx]);
        if (m == vmax) break;
      }
      d[x] = m;
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

// TEST INFRASTRUCTURE ONLY -- declaration shim that lets the reference's translation units
// (collision_checking.cpp, a_star.cpp, VoronoiDiagramGenerator.cpp ...) compile where they lie,
// without OpenCV.  Only the handful of cv:: names those files use exist here; semantics follow
// OpenCV 4.x for exactly the call patterns at collision_checking.cpp:106-108,197-223 and
// a_star.cpp:426-473.  Definitions: ../cv_shim.cpp.  Not part of the product library.
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

typedef int64_t int64;
typedef unsigned char uchar;

#define CV_8UC1 0
#define CV_32FC1 5

namespace cv
{
struct Point
{
  int x = 0, y = 0;
  Point() = default;
  Point(int x_, int y_) : x(x_), y(y_) {}
};
struct Size
{
  int width = 0, height = 0;
  Size() = default;
  Size(int w, int h) : width(w), height(h) {}
};
struct Range
{
  int start = 0, end = 0;
  Range() = default;
  Range(int s, int e) : start(s), end(e) {}
};

enum { CMP_EQ = 0, CMP_GT = 1, CMP_GE = 2, CMP_LT = 3, CMP_LE = 4, CMP_NE = 5 };
enum { MORPH_RECT = 0, MORPH_CROSS = 1, MORPH_ELLIPSE = 2 };
enum { MORPH_ERODE = 0, MORPH_DILATE = 1 };
enum { DIST_L1 = 1, DIST_L2 = 2 };
enum { DIST_MASK_3 = 3, DIST_MASK_5 = 5 };

template <typename T>
using Ptr = std::shared_ptr<T>;

std::shared_ptr<uchar> shim_alloc(size_t bytes);

class Mat
{
public:
  int rows = 0, cols = 0, type_ = CV_8UC1;
  size_t step = 0;  // bytes per row
  uchar* data = nullptr;
  std::shared_ptr<uchar> own;  // pooled, uninitialised storage (shim_alloc): temporaries are recycled as malloc does for OpenCV

  Mat() = default;
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(int r, int c, int type, void* ext) : rows(r), cols(c), type_(type), step((size_t)c * esz(type)), data((uchar*)ext) {}
  static size_t esz(int type) { return type == CV_32FC1 ? 4 : 1; }
  size_t elemSize() const { return esz(type_); }
  int type() const { return type_; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  void create(int r, int c, int type)
  {
    if (data && rows == r && cols == c && type_ == type) return;  // cv::Mat::create keeps a matching buffer
    own = shim_alloc((size_t)r * c * esz(type));
    rows = r;
    cols = c;
    type_ = type;
    step = (size_t)c * esz(type);
    data = own.get();
  }
  template <typename T>
  T* ptr(int y) { return reinterpret_cast<T*>(data + (size_t)y * step); }
  template <typename T>
  const T* ptr(int y) const { return reinterpret_cast<const T*>(data + (size_t)y * step); }
  Mat operator()(const Range& ry, const Range& rx) const
  {
    if (ry.start < 0 || rx.start < 0 || ry.end > rows || rx.end > cols || ry.end < ry.start || rx.end < rx.start)
      throw std::runtime_error("cv shim: ROI outside the matrix (OpenCV would assert)");
    Mat m;
    m.rows = ry.end - ry.start;
    m.cols = rx.end - rx.start;
    m.type_ = type_;
    m.step = step;
    m.data = data + (size_t)ry.start * step + (size_t)rx.start * elemSize();
    m.own = own;
    return m;
  }
  Mat clone() const;
  Mat& setTo(double value, const Mat& mask);
  Mat& operator*=(double s);
};

Mat operator==(const Mat& m, double v);
Mat operator==(double v, const Mat& m);
Mat operator-(double s, const Mat& m);  // saturating, u8 only (a_star.cpp:429)

void compare(const Mat& src, double value, Mat& dst, int cmpop);
void subtract(double s, const Mat& src, Mat& dst);
void bitwise_and(const Mat& a, const Mat& b, Mat& dst);
Mat getStructuringElement(int shape, Size ksize, Point anchor = Point(-1, -1));
void dilate(const Mat& src, Mat& dst, const Mat& kernel);
void distanceTransform(const Mat& src, Mat& dst, int distanceType, int maskSize);
void findNonZero(const Mat& src, std::vector<Point>& idx);

// which of OpenCV's two DIST_L2 / DIST_MASK_5 arithmetic variants distanceTransform uses
// (SURVEY.md H2): 0 = 16.16 fixed point (IPP off), 1 = float32 raster chamfer (IPP on)
extern int shim_dt_variant;

// ---- cv::cuda: only referenced from the compiled-out USE_GPU_MASK_DILATE branch
namespace cuda
{
class GpuMat
{
public:
  void upload(const Mat&) { throw std::runtime_error("cv shim: no cv::cuda"); }
  void download(Mat&) const { throw std::runtime_error("cv shim: no cv::cuda"); }
  GpuMat& setTo(double, const GpuMat&) { throw std::runtime_error("cv shim: no cv::cuda"); }
};
class Filter
{
public:
  virtual ~Filter() = default;
  virtual void apply(const GpuMat&, GpuMat&) { throw std::runtime_error("cv shim: no cv::cuda"); }
};
inline Ptr<Filter> createMorphologyFilter(int, int, const Mat&) { return std::make_shared<Filter>(); }
inline void compare(const GpuMat&, double, GpuMat&, int) { throw std::runtime_error("cv shim: no cv::cuda"); }
inline void subtract(double, const GpuMat&, GpuMat&) { throw std::runtime_error("cv shim: no cv::cuda"); }
inline void bitwise_and(const GpuMat&, const GpuMat&, GpuMat&) { throw std::runtime_error("cv shim: no cv::cuda"); }
}  // namespace cuda
}  // namespace cv

// oracle/stubs_cv/opencv2/opencv.hpp — TEST INFRASTRUCTURE ONLY.
//
// A FUNCTIONAL miniature of the OpenCV 2.4 types that /root/reference/src/color_coding.cpp and src/IO_flow.cpp use, so
// that those two reference sources compile UNMODIFIED where they lie (oracle/Makefile, target _ref/libdis_refviz.so)
// and pin the colour-coding and .flo oracles against the reference itself.  Containers only: all
// the arithmetic stays in the reference's own code.  (The engine sources use the inert stub in oracle/stubs/.)
#ifndef DIS_ORACLE_STUB_OPENCV_FUNCTIONAL
#define DIS_ORACLE_STUB_OPENCV_FUNCTIONAL
#include <math.h>   // as OpenCV 2.4 core/types_c.h does: with libstdc++ (and MSVC) this makes the float overloads of sqrt / atan2 / fabs visible unqualified
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_32F 5
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_32FC4 CV_MAKETYPE(CV_32F, 4)
typedef unsigned char uchar;
inline int cvIsNaN(double value) { return std::isnan(value) ? 1 : 0; }   // OpenCV 2.4 core/types_c.h: takes a double

namespace cv {
using std::max;
using std::min;
template <typename T, int N>
struct Vec {
  T val[N];
  Vec() { for (int i = 0; i < N; ++i) val[i] = T(0); }
  Vec(T a, T b, T c) { val[0] = a; val[1] = b; val[2] = c; }
  T& operator[](int i) { return val[i]; }
  const T& operator[](int i) const { return val[i]; }
};
typedef Vec<uchar, 3> Vec3b;
typedef Vec<int, 3> Vec3i;
typedef Vec<float, 2> Vec2f;
typedef Vec<float, 4> Vec4f;
struct Point2f { float x, y; };
struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Scalar { double val[4]; static Scalar all(double v) { Scalar s; for (int i = 0; i < 4; ++i) s.val[i] = v; return s; } };

class Mat {
 public:
  int rows, cols, type_;
  uchar* data;
  std::vector<uchar> own;
  Mat() : rows(0), cols(0), type_(0), data(nullptr) {}
  Mat(int r, int c, int type, void* ext) : rows(r), cols(c), type_(type), data(static_cast<uchar*>(ext)) {}
  int channels() const { return (type_ >> 3) + 1; }
  size_t elemSize() const { return (size_t)channels() * ((type_ & 7) == CV_32F ? 4 : 1); }
  Size size() const { return Size(cols, rows); }
  void create(Size s, int type) {
    rows = s.height; cols = s.width; type_ = type;
    own.assign((size_t)rows * cols * elemSize(), 0);
    data = own.data();
  }
  void setTo(const Scalar& s) {   // only ever called with Scalar::all(0) on 8-bit images here
    const uchar v = (uchar)s.val[0];
    std::memset(data, v, (size_t)rows * cols * elemSize());
  }
  template <typename T> T& at(int y, int x) { return reinterpret_cast<T*>(data)[(size_t)y * cols + x]; }
  template <typename T> const T& at(int y, int x) const { return reinterpret_cast<const T*>(data)[(size_t)y * cols + x]; }
};
template <typename T>
class Mat_ : public Mat {
 public:
  Mat_(int r, int c, T* ext) : Mat(r, c, CV_32FC2, ext) {}
  const T& operator()(int y, int x) const { return reinterpret_cast<const T*>(data)[(size_t)y * cols + x]; }
};
}  // namespace cv
#endif

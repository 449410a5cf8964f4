// oracle/stubs/opencv2/opencv.hpp — TEST INFRASTRUCTURE ONLY (see oracle/README.md).
//
// The reference engine includes OpenCV only for its GUI "draw_grid" branch
// (/root/reference/src/optical_flow.cpp:93-145), which the oracle never enters
// (draw_grid = false). OpenCV's C++ headers are not installed here, so the names that
// branch mentions are declared as inert placeholders to let the three engine sources
// compile unmodified. No arithmetic on the measured path goes through this file.
#ifndef DIS_ORACLE_STUB_OPENCV
#define DIS_ORACLE_STUB_OPENCV
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

enum { CV_8UC1 = 0, CV_32FC1 = 5, CV_GRAY2RGB = 8 };

namespace cv {
enum { INTER_NEAREST = 0, WINDOW_AUTOSIZE = 1 };
struct Size { Size() {} Size(int, int) {} };
struct Rect { Rect(int, int, int, int) {} };
struct Point { template <typename A, typename B> Point(A, B) {} };
struct Scalar { Scalar(double = 0, double = 0, double = 0, double = 0) {} };
struct Mat {
  Mat() {}
  Mat(int, int, int, void*) {}
  Mat operator()(const Rect&) const { return Mat(); }
  void convertTo(Mat&, int) const {}
};
inline void cvtColor(const Mat&, Mat&, int) {}
inline void resize(const Mat&, Mat&, Size, double, double, int) {}
inline void line(Mat&, Point, Point, Scalar, int = 1) {}
inline void namedWindow(const char*, int) {}
inline void imshow(const char*, const Mat&) {}
inline int waitKey(int) { return 0; }
}  // namespace cv
#endif

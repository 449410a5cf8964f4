// oracle/ref_viz_driver.cpp — TEST INFRASTRUCTURE ONLY.  C entry points around the reference's own colour coding
// (/root/reference/src/color_coding.cpp:83-117, draw_optical_flow) and .flo I/O (/root/reference/src/IO_flow.cpp:10-98),
// compiled unmodified where they lie against oracle/stubs_cv/ into oracle/_ref/libdis_refviz.so.
#include <cstdint>
#include <opencv2/opencv.hpp>
#include <color_coding.hpp>
#include <IO_flow.hpp>

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API void ref_color_code(const float* flow, int w, int h, float maxmotion, uint8_t* bgr) {
  cv::Mat_<cv::Point2f> f(h, w, reinterpret_cast<cv::Point2f*>(const_cast<float*>(flow)));
  cv::Mat dst;
  draw_optical_flow(f, dst, maxmotion);
  std::memcpy(bgr, dst.data, (size_t)3 * w * h);
}

REF_API void ref_save_flo(const float* flow, int w, int h, const char* path) {
  cv::Mat m(h, w, CV_32FC2, const_cast<float*>(flow));
  SaveFlowFile(m, path);
}

REF_API void ref_read_flo(const char* path, float* flow, int w, int h) {
  cv::Mat m(h, w, CV_32FC2, flow);
  ReadFlowFile(m, path);
}

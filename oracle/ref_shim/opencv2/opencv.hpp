// TEST INFRASTRUCTURE ONLY.  Stand-in for <opencv2/opencv.hpp> so that the reference's detector
// templates (external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.{hpp,inl}) compile
// UNMODIFIED with plain g++ in a container that has no OpenCV headers.  Only what those two files
// name is declared: a dense float matrix with rows / cols / at<float>(i, j) / clone / zeros, and the
// filter2D signature Cen2018 mentions (declared, never called by the oracle driver).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <utility>
#include <vector>

#define CV_32F 5

namespace cv {

struct Point { int x, y; Point(int x_ = 0, int y_ = 0) : x(x_), y(y_) {} };
enum { BORDER_REFLECT101 = 4 };

class Mat {
 public:
  int rows = 0, cols = 0;
  Mat() = default;
  Mat(int r, int c) : rows(r), cols(c), data_(std::make_shared<std::vector<float>>((size_t)r * c, 0.f)) {}
  static Mat zeros(int r, int c, int /*type*/) { return Mat(r, c); }
  Mat clone() const { Mat m(rows, cols); if (data_) *m.data_ = *data_; return m; }
  Mat& operator/=(float s) { if (data_) for (float& v : *data_) v /= s; return *this; }
  template <typename T> T& at(int i, int j) { return (*data_)[(size_t)i * cols + j]; }
  template <typename T> const T& at(int i, int j) const { return (*data_)[(size_t)i * cols + j]; }
 private:
  std::shared_ptr<std::vector<float>> data_;
};

inline void filter2D(const Mat&, Mat&, int, const Mat&, Point, double, int) {
  throw std::runtime_error("cv::filter2D is not part of the oracle shim");
}

}  // namespace cv

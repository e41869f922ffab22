// TEST INFRASTRUCTURE ONLY.  Stand-in for vtr_radar/data_types/point.hpp (which needs PCL): the
// container interface the detector templates use (clear / push_back / insert / begin / end) and a
// point type with the reference's field types (float x, y, z, rho, theta, phi; int64 timestamp --
// external/vtr3/main/src/vtr_radar/include/vtr_radar/data_types/point.hpp:29-47,100-130).
#pragma once
#include <cstdint>
#include <vector>

namespace pcl {
template <class PointT>
class PointCloud : public std::vector<PointT> {};
}  // namespace pcl

namespace vtr {
namespace radar {
struct OraclePoint {
  float x = 0, y = 0, z = 0;
  float rho = 0, theta = 0, phi = 0;
  int64_t timestamp = 0;
};
}  // namespace radar
}  // namespace vtr

// TEST INFRASTRUCTURE ONLY.  Runs the reference's UNMODIFIED detector templates
//   external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.inl:39-85   KStrongest<PointT>::run
//   external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.inl:345-400 ModifiedCACFAR<PointT>::run
// (compiled from where they lie under /root/reference against the shim headers in oracle/ref_shim/) on
// scans read from a file, and writes the polar points they return.  Built by oracle/Makefile into
// oracle/_ref/ (git-ignored); used by oracle/make_golden_kstrong.py to produce tests/golden/kstrong_ref.npz.
//
// input  (little endian): int32 detector (0 k-strongest, 1 modified CA-CFAR), int32 rows, int32 cols,
//                          float32 res, float64 p[8] (k-strongest: k, threshold2, threshold3, minr, maxr,
//                          range_offset; CFAR: width, guard, threshold, threshold2, threshold3, minr, maxr,
//                          range_offset), float64 azimuth[rows], int64 time[rows], float32 scan[rows*cols]
// output: int32 count, then per point float32 rho, float32 phi, int64 timestamp
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "vtr_radar/detector/detector.hpp"

using vtr::radar::OraclePoint;

template <typename T>
static void rd(FILE* f, T* p, size_t n) {
  if (fread(p, sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
}

int main(int argc, char** argv) {
  if (argc != 3) { fprintf(stderr, "usage: %s in.bin out.bin\n", argv[0]); return 2; }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 2;
  int32_t det, rows, cols; float res; double p[8];
  rd(f, &det, 1); rd(f, &rows, 1); rd(f, &cols, 1); rd(f, &res, 1); rd(f, p, 8);
  std::vector<double> az(rows); std::vector<int64_t> tm(rows); std::vector<float> scan((size_t)rows * cols);
  rd(f, az.data(), rows); rd(f, tm.data(), rows); rd(f, scan.data(), scan.size());
  fclose(f);
  cv::Mat m(rows, cols);
  for (int i = 0; i < rows; ++i)
    for (int j = 0; j < cols; ++j) m.at<float>(i, j) = scan[(size_t)i * cols + j];
  pcl::PointCloud<OraclePoint> cloud;
  if (det == 0) {
    vtr::radar::KStrongest<OraclePoint> d((int)p[0], p[1], p[2], p[3], p[4], p[5]);
    d.run(m, res, tm, az, cloud);
  } else {
    vtr::radar::ModifiedCACFAR<OraclePoint> d((int)p[0], (int)p[1], p[2], p[3], p[4], p[5], p[6], p[7]);
    d.run(m, res, tm, az, cloud);
  }
  FILE* g = fopen(argv[2], "wb");
  if (!g) return 2;
  const int32_t count = (int32_t)cloud.size();
  fwrite(&count, 4, 1, g);
  for (const auto& q : cloud) { fwrite(&q.rho, 4, 1, g); fwrite(&q.phi, 4, 1, g); fwrite(&q.timestamp, 8, 1, g); }
  fclose(g);
  return 0;
}

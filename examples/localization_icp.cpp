// ROS-free C++ consumer of include/mm_masking_b200.h (the reachable part).
//
// What it stands in for: the production localisation ICP of vtr3,
//   external/vtr3/main/src/vtr_radar_lidar/src/modules/localization/localization_icp_module.cpp:276-420
// -- per step: 1-NN of every (masked) radar point in the lidar submap (kd-tree, :284-293), a distance gate
// (:296-306), a Cauchy-robust point-to-point Gauss-Newton step (:327-366, `huber_delta`/cauchy k of
// gen_data/config/radar_lidar_config.yaml:176) -- with the per-point weights a trained mask would supply through
// the TorchModule hook (external/vtr3/main/src/vtr_torch/include/vtr_torch/modules/torch_module.hpp:30-90).
// The module itself needs ROS 2, PCL and STEAM, none of which exist in this image ("unavailable on this box");
// this translation unit shows that a plain C++ caller can drive the same computation through the C ABI: it
// includes the header, links libmm_masking_b200.so, owns every buffer (cudaMalloc), and runs point-to-point ICP
// with a hard distance gate, Cauchy weights and per-point prior weights on one scan/map pair.
//
//   localization_icp <in.bin> [max_iterations] [gate_m] [cauchy_k]
//   in.bin (little endian): int32 n, int32 m, float32 source[n*3], float32 target[m*3], float32 weight[n],
//                           float32 T_init[16]
//   stdout: "T" then the 4x4 pose (row-major, %.9g), "iterations <k>", "converged <0|1>", "matched_ratio <r>"
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../include/mm_masking_b200.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 3; } } while (0)

template <typename T>
static bool rd(FILE* f, T* p, size_t count) { return fread(p, sizeof(T), count, f) == count; }

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s in.bin [max_iterations] [gate_m] [cauchy_k]\n", argv[0]); return 2; }
  const int max_it = argc > 2 ? atoi(argv[2]) : 30;
  const double gate = argc > 3 ? atof(argv[3]) : 5.0, cauchy_k = argc > 4 ? atof(argv[4]) : 1.0;
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 2; }
  int32_t n = 0, m = 0;
  if (!rd(f, &n, 1) || !rd(f, &m, 1) || n <= 0 || m <= 0) return 2;
  std::vector<float> src((size_t)n * 3), tgt((size_t)m * 3), w(n), T0(16);
  if (!rd(f, src.data(), src.size()) || !rd(f, tgt.data(), tgt.size()) || !rd(f, w.data(), w.size()) || !rd(f, T0.data(), 16)) return 2;
  fclose(f);

  if (mmb_abi_version() != MMB_ABI_VERSION) { fprintf(stderr, "ABI mismatch\n"); return 3; }
  mmb_icp_params p = {};
  p.N = 1; p.n = n; p.m = m; p.target_cols = 3;
  p.dtype = MMB_F32; p.icp_type = MMB_PT2PT; p.dim = 3; p.max_iterations = max_it;
  p.differentiable = 0;                    // hard gate / plain Cauchy: the deployed (non-differentiable) form
  p.const_iter = 0; p.use_trim = 1; p.trim_dist = gate;
  p.loss_kind = MMB_LOSS_CAUCHY; p.loss_metric = cauchy_k;
  p.has_weight = 1; p.save_for_backward = 0; p.nn_slices = 0; p.nn_backend = MMB_NN_AUTO;
  p.tolerance = 1e-6; p.tanh_steepness = 5.0; p.match_ratio_thresh = 0.1;

  const size_t ws_bytes = mmb_icp_workspace_bytes(&p);
  if (!ws_bytes) { fprintf(stderr, "parameters rejected\n"); return 3; }
  float *d_src, *d_tgt, *d_w, *d_T0, *d_T, *d_pc, *d_del, *d_wh, *d_cost, *d_it, *d_mr;
  uint8_t* d_conv; int32_t* d_done; void* d_ws;
  const size_t R = (size_t)3 * n;
  CK(cudaMalloc(&d_src, src.size() * 4)); CK(cudaMalloc(&d_tgt, tgt.size() * 4)); CK(cudaMalloc(&d_w, n * 4));
  CK(cudaMalloc(&d_T0, 64)); CK(cudaMalloc(&d_T, 64)); CK(cudaMalloc(&d_pc, src.size() * 4));
  CK(cudaMalloc(&d_del, (size_t)max_it * 6 * 4)); CK(cudaMalloc(&d_wh, (size_t)max_it * R * 4));
  CK(cudaMalloc(&d_cost, (size_t)max_it * 4)); CK(cudaMalloc(&d_conv, 1)); CK(cudaMalloc(&d_it, 4)); CK(cudaMalloc(&d_mr, 4));
  CK(cudaMalloc(&d_done, 4)); CK(cudaMalloc(&d_ws, ws_bytes));
  CK(cudaMemcpy(d_src, src.data(), src.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_tgt, tgt.data(), tgt.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_w, w.data(), n * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_T0, T0.data(), 64, cudaMemcpyHostToDevice));
  CK(cudaMemset(d_del, 0, (size_t)max_it * 6 * 4)); CK(cudaMemset(d_wh, 0, (size_t)max_it * R * 4));
  CK(cudaMemset(d_cost, 0, (size_t)max_it * 4)); CK(cudaMemset(d_conv, 0, 1)); CK(cudaMemset(d_it, 0, 4));
  CK(cudaMemset(d_mr, 0, 4)); CK(cudaMemset(d_done, 0, 4));

  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  const int rc = mmb_icp_forward(&p, d_src, d_tgt, d_T0, d_w, d_T, d_pc, d_del, d_wh, d_cost, d_conv, d_it, d_mr, d_done,
                                 nullptr, d_ws, ws_bytes, st);
  if (rc != MMB_OK) { fprintf(stderr, "mmb_icp_forward: %s %s\n", mmb_status_string(rc), mmb_last_cuda_error()); return 3; }
  CK(cudaStreamSynchronize(st));
  float T[16], it = 0, mr = 0; uint8_t conv = 0;
  CK(cudaMemcpy(T, d_T, 64, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&it, d_it, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&mr, d_mr, 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&conv, d_conv, 1, cudaMemcpyDeviceToHost));
  printf("T\n");
  for (int i = 0; i < 4; ++i) printf("%.9g %.9g %.9g %.9g\n", T[4 * i], T[4 * i + 1], T[4 * i + 2], T[4 * i + 3]);
  printf("iterations %d\nconverged %d\nmatched_ratio %.9g\n", (int)it, (int)conv, mr);
  return 0;
}

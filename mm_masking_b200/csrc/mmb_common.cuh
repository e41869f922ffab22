// Shared device helpers for the mm_masking_b200 kernels (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <limits.h>

#include <atomic>

#include "../../include/mm_masking_b200.h"

namespace mmb {

// ---- error plumbing -------------------------------------------------------------------------
void set_last_cuda_error(cudaError_t e);
extern std::atomic<long long> g_launch_count;
// called once after every kernel launch of the library: error check + launch count
inline int check_launch() {
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_last_cuda_error(e); return MMB_ERR_CUDA; }
  return MMB_OK;
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Bump allocator over the caller's workspace; every block is 256-byte aligned so bulk
// (TMA) copies and vector loads from it are legal.
struct Carver {
  char* base; size_t off;
  explicit Carver(void* p) : base(static_cast<char*>(p)), off(0) {}
  template <typename T> T* take(size_t count) {
    T* r = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off = align_up(off + count * sizeof(T), 256);
    return r;
  }
};

// ---- packed fp32x2 arithmetic (FADD2 / FMUL2 / FFMA2 on sm_100a) and 3-input min (FMNMX3) ---
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
  f32x2 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
  asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
  f32x2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
  f32x2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
  f32x2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;
}
__device__ __forceinline__ float min3(float a, float b, float c) {
  float r; asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r;
}

// ---- mbarrier + 1-D bulk async copy (TMA engine, UBLKCP in SASS) -------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// ---- reductions ----------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Deterministic block sum of K doubles per thread.  `scratch` holds K * (blockDim/32 + 1)
// doubles.  After the call every thread can read the K totals from scratch + K * (blockDim/32).
template <int K>
__device__ __forceinline__ const double* block_sum(double (&v)[K], double* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = warp_sum(v[k]);
  __syncthreads();                          // scratch may still be read from a previous round
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) scratch[k * nwarp + warp] = v[k];
  }
  __syncthreads();
  double* totals = scratch + K * nwarp;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    double s = 0.0;
    for (int w = 0; w < nwarp; ++w) s += scratch[k * nwarp + w];
    totals[k] = s;
  }
  __syncthreads();
  return totals;
}

// ---- small fp64 linear algebra (one thread) ---------------------------------------------------
// Solve A x = b for symmetric positive definite A (D <= 6) by Cholesky; falls back to Gaussian
// elimination with partial pivoting when a pivot is not positive (the reference inverts with LU,
// ICP.py:191).  A is row-major D x D and is not modified.
template <int D>
__device__ inline void solve_spd(const double* A, const double* b, double* x) {
  double L[D * D];
  bool ok = true;
  for (int i = 0; i < D && ok; ++i) {
    for (int j = 0; j <= i; ++j) {
      double s = A[i * D + j];
      for (int k = 0; k < j; ++k) s -= L[i * D + k] * L[j * D + k];
      if (i == j) {
        if (!(s > 0.0)) { ok = false; break; }
        L[i * D + i] = sqrt(s);
      } else {
        L[i * D + j] = s / L[j * D + j];
      }
    }
  }
  if (ok) {
    double y[D];
    for (int i = 0; i < D; ++i) {
      double s = b[i];
      for (int k = 0; k < i; ++k) s -= L[i * D + k] * y[k];
      y[i] = s / L[i * D + i];
    }
    for (int i = D - 1; i >= 0; --i) {
      double s = y[i];
      for (int k = i + 1; k < D; ++k) s -= L[k * D + i] * x[k];
      x[i] = s / L[i * D + i];
    }
    return;
  }
  double M[D * (D + 1)];
  for (int i = 0; i < D; ++i) {
    for (int j = 0; j < D; ++j) M[i * (D + 1) + j] = A[i * D + j];
    M[i * (D + 1) + D] = b[i];
  }
  for (int c = 0; c < D; ++c) {
    int piv = c;
    for (int r2 = c + 1; r2 < D; ++r2)
      if (fabs(M[r2 * (D + 1) + c]) > fabs(M[piv * (D + 1) + c])) piv = r2;
    if (piv != c)
      for (int j = 0; j <= D; ++j) {
        double t = M[c * (D + 1) + j]; M[c * (D + 1) + j] = M[piv * (D + 1) + j]; M[piv * (D + 1) + j] = t;
      }
    double pv = M[c * (D + 1) + c];
    if (pv == 0.0) pv = 1e-300;
    for (int r2 = c + 1; r2 < D; ++r2) {
      double f = M[r2 * (D + 1) + c] / pv;
      for (int j = c; j <= D; ++j) M[r2 * (D + 1) + j] -= f * M[c * (D + 1) + j];
    }
  }
  for (int i = D - 1; i >= 0; --i) {
    double s = M[i * (D + 1) + D];
    for (int k = i + 1; k < D; ++k) s -= M[i * (D + 1) + k] * x[k];
    double pv = M[i * (D + 1) + i];
    if (pv == 0.0) pv = 1e-300;
    x[i] = s / pv;
  }
}

// R = exp(phi^) (Rodrigues; series near 0).  Row-major 3x3.
__device__ inline void so3_exp(const double* phi, double* R) {
  const double x = phi[0], y = phi[1], z = phi[2];
  const double t2 = x * x + y * y + z * z;
  double a, b;                         // R = I + a K + b K^2
  if (t2 < 1e-12) {
    a = 1.0 - t2 / 6.0 + t2 * t2 / 120.0;
    b = 0.5 - t2 / 24.0 + t2 * t2 / 720.0;
  } else {
    const double t = sqrt(t2);
    a = sin(t) / t;
    b = (1.0 - cos(t)) / t2;
  }
  R[0] = 1.0 - b * (y * y + z * z); R[1] = -a * z + b * x * y;         R[2] = a * y + b * x * z;
  R[3] = a * z + b * x * y;         R[4] = 1.0 - b * (x * x + z * z); R[5] = -a * x + b * y * z;
  R[6] = -a * y + b * x * z;        R[7] = a * x + b * y * z;         R[8] = 1.0 - b * (x * x + y * y);
}

// out = J_l(phi)^T s, with J_l the left Jacobian of SO(3): d exp(phi^) = (J_l dphi)^ exp(phi^).
__device__ inline void so3_left_jacobian_T_apply(const double* phi, const double* s, double* out) {
  const double x = phi[0], y = phi[1], z = phi[2];
  const double t2 = x * x + y * y + z * z;
  double b, c;                         // J_l = I + b K + c K^2 ; J_l^T = I - b K + c K^2
  if (t2 < 1e-12) {
    b = 0.5 - t2 / 24.0;
    c = 1.0 / 6.0 - t2 / 120.0;
  } else {
    const double t = sqrt(t2);
    b = (1.0 - cos(t)) / t2;
    c = (t - sin(t)) / (t2 * t);
  }
  // K s = phi x s ; K^2 s = phi x (phi x s)
  const double k1x = y * s[2] - z * s[1], k1y = z * s[0] - x * s[2], k1z = x * s[1] - y * s[0];
  const double k2x = y * k1z - z * k1y, k2y = z * k1x - x * k1z, k2z = x * k1y - y * k1x;
  out[0] = s[0] - b * k1x + c * k2x;
  out[1] = s[1] - b * k1y + c * k2y;
  out[2] = s[2] - b * k1z + c * k2z;
}

// Rotate a source point: a = C p, evaluated with a fixed operation order so that every kernel
// that needs the transformed point produces identical bits.
__device__ __forceinline__ void rotate_point(const double* C, double px, double py, double pz, double& ax,
                                             double& ay, double& az) {
  ax = fma(C[2], pz, fma(C[1], py, __dmul_rn(C[0], px)));
  ay = fma(C[5], pz, fma(C[4], py, __dmul_rn(C[3], px)));
  az = fma(C[8], pz, fma(C[7], py, __dmul_rn(C[6], px)));
}

template <typename T> __device__ __forceinline__ T from_double(double v);
template <> __device__ __forceinline__ float from_double<float>(double v) { return __double2float_rn(v); }
template <> __device__ __forceinline__ double from_double<double>(double v) { return v; }

}  // namespace mmb

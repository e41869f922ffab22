// Radar preprocessing kernels for sm_100a: per-point bilinear weight sampling (K6, forward and
// adjoint), polar -> Cartesian resampling (K1), GO-CFAR mask + peak extraction (K2a) and
// K-strongest extraction (K2b).
//
// What is computed follows mm_masking/radar_utils.py of the reference (and, for K-strongest,
// external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.inl:39-85); see
// include/mm_masking_b200.h for the per-function mapping.  How it is computed does not: the
// reference builds these from dozens of ATen launches (two Python loops of 1 198 `torch.sum`
// calls for the CFAR windows, a (B,640,400) expanded searchsorted, per-sample `nonzero`), here each
// stage is one pass over the scan in HBM order with the row staged in shared memory.
#include "mmb_common.cuh"

#include <type_traits>

#include <cub/device/device_radix_sort.cuh>

namespace mmb {

// All of these kernels are bound by HBM traffic; grids are sized in whole rows / pixels.
constexpr int ROW_THREADS = 256;

// Inputs come as fp32, fp64 or raw Navtech bytes (radar_utils.py:20-27: 11 header bytes per row,
// value = byte / 255 in fp32).
enum { IN_F32 = 0, IN_F64 = 1, IN_U8 = 2 };

template <typename T> struct Scan {
  const void* base; int kind; size_t row_stride;      // elements (bytes for u8) between azimuth rows
  __device__ __forceinline__ T at(size_t row, int col) const {
    if (kind == IN_F32) return (T) static_cast<const float*>(base)[row * row_stride + col];
    if (kind == IN_F64) return (T) static_cast<const double*>(base)[row * row_stride + col];
    return (T)__fdiv_rn((float)static_cast<const uint8_t*>(base)[row * row_stride + 11 + col], 255.0f);
  }
};

template <typename T> __device__ __forceinline__ T t_floor(T v);
template <> __device__ __forceinline__ float t_floor<float>(float v) { return floorf(v); }
template <> __device__ __forceinline__ double t_floor<double>(double v) { return floor(v); }
template <typename T> __device__ __forceinline__ T t_tanh(T v);
template <> __device__ __forceinline__ float t_tanh<float>(float v) { return tanhf(v); }
template <> __device__ __forceinline__ double t_tanh<double>(double v) { return tanh(v); }

// Unfused arithmetic in the working type, so that products and sums round where ATen's do.
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }

// ---------------------------------------------------------------------------------------------
// bilinear taps of `grid_sample(mode='bilinear', padding_mode='zeros', align_corners=True)`
// (ATen CPU kernel: ix = (gx + 1) * ((W - 1) / 2); weights nw = (1-fy)(1-fx) ...; taps outside
// the image contribute 0)
// ---------------------------------------------------------------------------------------------
template <typename T> struct Taps { int x0, y0; T nw, ne, sw, se; };

template <typename T>
__device__ __forceinline__ Taps<T> bilinear_taps(T gx, T gy, int W, int H) {
  const T ix = mul_rn(add_rn(gx, (T)1), (T)(W - 1) / (T)2);
  const T iy = mul_rn(add_rn(gy, (T)1), (T)(H - 1) / (T)2);
  const T xw = t_floor(ix), yn = t_floor(iy);
  const T w = add_rn(ix, -xw), e = add_rn((T)1, -w);
  const T n = add_rn(iy, -yn), s = add_rn((T)1, -n);
  Taps<T> t;
  // keep far-out-of-range coordinates (the -100 of padded points) inside int range
  t.x0 = (int)fmax(fmin((double)xw, 2.0e9), -2.0e9);
  t.y0 = (int)fmax(fmin((double)yn, 2.0e9), -2.0e9);
  t.nw = mul_rn(s, e); t.ne = mul_rn(s, w); t.sw = mul_rn(n, e); t.se = mul_rn(n, w);
  return t;
}

__device__ __forceinline__ bool inside(int x, int y, int W, int H) { return x >= 0 && x < W && y >= 0 && y < H; }

// ---------------------------------------------------------------------------------------------
// K6: per-point weight sampling (radar_utils.py:108-126) and its adjoint
// ---------------------------------------------------------------------------------------------
// grid = (v, u) * 2 / (W - 1) with u = -x / cart_res, v = y / cart_res (radar_utils.py:374-391);
// points with x == 0 and y == 0 are batch padding and sample at (-100, -100) -> 0 (:118-122).
template <typename T>
__device__ __forceinline__ void point_grid(const T* __restrict__ p, T cart_res, int W, T& gx, T& gy) {
  const T x = p[0], y = p[1];
  if (x == (T)0 && y == (T)0) { gx = (T)-100; gy = (T)-100; return; }
  const T u = div_rn(-x, cart_res), v = div_rn(y, cart_res);
  gx = mul_rn(div_rn(v, (T)(W - 1)), (T)2);
  gy = mul_rn(div_rn(u, (T)(W - 1)), (T)2);
}

template <typename T>
__global__ void sample_weights_kernel(const T* __restrict__ mask, const T* __restrict__ pts, int pt_cols, int N, int n,
                                      int H, int W, T cart_res, T* __restrict__ out) {
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)N * n) return;
  const int b = (int)(g / n);
  T gx, gy;
  point_grid<T>(pts + g * pt_cols, cart_res, W, gx, gy);
  const Taps<T> t = bilinear_taps<T>(gx, gy, W, H);
  const T* img = mask + (size_t)b * H * W;
  T acc = (T)0;
  if (inside(t.x0, t.y0, W, H)) acc = mul_rn(img[(size_t)t.y0 * W + t.x0], t.nw);
  if (inside(t.x0 + 1, t.y0, W, H)) acc = add_rn(acc, mul_rn(img[(size_t)t.y0 * W + t.x0 + 1], t.ne));
  if (inside(t.x0, t.y0 + 1, W, H)) acc = add_rn(acc, mul_rn(img[(size_t)(t.y0 + 1) * W + t.x0], t.sw));
  if (inside(t.x0 + 1, t.y0 + 1, W, H)) acc = add_rn(acc, mul_rn(img[(size_t)(t.y0 + 1) * W + t.x0 + 1], t.se));
  out[g] = acc;
}

// grad_mask must be zero-filled by the caller; taps are accumulated with atomics (as ATen's own
// grid_sample backward does), so colliding taps sum in an unspecified order.
template <typename T>
__global__ void sample_weights_bwd_kernel(const T* __restrict__ grad_w, const T* __restrict__ pts, int pt_cols, int N,
                                          int n, int H, int W, T cart_res, T* __restrict__ grad_mask) {
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)N * n) return;
  const int b = (int)(g / n);
  T gx, gy;
  point_grid<T>(pts + g * pt_cols, cart_res, W, gx, gy);
  const Taps<T> t = bilinear_taps<T>(gx, gy, W, H);
  T* img = grad_mask + (size_t)b * H * W;
  const T go = grad_w[g];
  if (inside(t.x0, t.y0, W, H)) atomicAdd(img + (size_t)t.y0 * W + t.x0, mul_rn(go, t.nw));
  if (inside(t.x0 + 1, t.y0, W, H)) atomicAdd(img + (size_t)t.y0 * W + t.x0 + 1, mul_rn(go, t.ne));
  if (inside(t.x0, t.y0 + 1, W, H)) atomicAdd(img + (size_t)(t.y0 + 1) * W + t.x0, mul_rn(go, t.sw));
  if (inside(t.x0 + 1, t.y0 + 1, W, H)) atomicAdd(img + (size_t)(t.y0 + 1) * W + t.x0 + 1, mul_rn(go, t.se));
}

// Deterministic form of the adjoint: every tap becomes a (pixel key, contribution) pair, the pairs
// are sorted by key with a STABLE radix sort (CUB, toolkit library; equal keys keep their tap order), and the
// first element of every run of equal keys adds the run in that order and stores the sum -- no floating-point
// atomics, the same bits on every run.  Taps outside the image get the key `total_pixels` and sort to the end.
template <typename T>
__global__ void sample_weights_taps_kernel(const T* __restrict__ grad_w, const T* __restrict__ pts, int pt_cols, int N,
                                           int n, int H, int W, T cart_res, unsigned long long* __restrict__ keys,
                                           T* __restrict__ vals) {
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)N * n) return;
  const int b = (int)(g / n);
  T gx, gy;
  point_grid<T>(pts + g * pt_cols, cart_res, W, gx, gy);
  const Taps<T> t = bilinear_taps<T>(gx, gy, W, H);
  const unsigned long long base = (unsigned long long)b * H * W, none = (unsigned long long)N * H * W;
  const T go = grad_w[g];
  const int xs[4] = {t.x0, t.x0 + 1, t.x0, t.x0 + 1}, ys[4] = {t.y0, t.y0, t.y0 + 1, t.y0 + 1};
  const T ws[4] = {t.nw, t.ne, t.sw, t.se};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const bool in = inside(xs[k], ys[k], W, H);
    keys[4 * g + k] = in ? base + (unsigned long long)ys[k] * W + xs[k] : none;
    vals[4 * g + k] = in ? mul_rn(go, ws[k]) : (T)0;
  }
}

template <typename T>
__global__ void sample_weights_runs_kernel(const unsigned long long* __restrict__ keys, const T* __restrict__ vals,
                                           size_t count, unsigned long long none, T* __restrict__ grad_mask) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const unsigned long long k = keys[i];
  if (k >= none || (i > 0 && keys[i - 1] == k)) return;       // not the head of a run
  T acc = vals[i];
  for (size_t j = i + 1; j < count && keys[j] == k; ++j) acc = add_rn(acc, vals[j]);
  grad_mask[k] = acc;
}

struct SampleBwdWorkspace {
  unsigned long long *keys_in, *keys_out;
  void *vals_in, *vals_out, *temp;
  size_t temp_bytes, bytes;
};

static int key_bits(unsigned long long none) {
  int bits = 1;
  while (bits < 64 && (none >> bits)) ++bits;
  return bits;
}

template <typename T>
static SampleBwdWorkspace carve_sample_bwd(void* base, int N, int n, int H, int W) {
  Carver c(base);
  SampleBwdWorkspace w;
  const size_t count = (size_t)N * n * 4;
  w.keys_in = c.take<unsigned long long>(count); w.keys_out = c.take<unsigned long long>(count);
  w.vals_in = c.take<T>(count); w.vals_out = c.take<T>(count);
  w.temp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, w.temp_bytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                  (const T*)nullptr, (T*)nullptr, count, 0, key_bits((unsigned long long)N * H * W));
  w.temp = c.take<char>(w.temp_bytes);
  w.bytes = c.off;
  return w;
}

template <typename T>
static int run_sample_bwd_det(int N, int n, int pt_cols, int H, int W, double cart_res, const void* grad_w, const void* pts,
                              void* grad_mask, void* workspace, cudaStream_t st) {
  SampleBwdWorkspace w = carve_sample_bwd<T>(workspace, N, n, H, W);
  const size_t count = (size_t)N * n * 4;
  const unsigned long long none = (unsigned long long)N * H * W;
  sample_weights_taps_kernel<T><<<(int)(((size_t)N * n + 255) / 256), 256, 0, st>>>(
      static_cast<const T*>(grad_w), static_cast<const T*>(pts), pt_cols, N, n, H, W, (T)cart_res, w.keys_in,
      static_cast<T*>(w.vals_in));
  int rc = check_launch();
  if (rc) return rc;
  if (cub::DeviceRadixSort::SortPairs(w.temp, w.temp_bytes, w.keys_in, w.keys_out, static_cast<const T*>(w.vals_in),
                                      static_cast<T*>(w.vals_out), count, 0, key_bits(none), st) != cudaSuccess)
    return check_launch();
  sample_weights_runs_kernel<T><<<(int)((count + 255) / 256), 256, 0, st>>>(w.keys_out, static_cast<const T*>(w.vals_out),
                                                                           count, none, static_cast<T*>(grad_mask));
  return check_launch();
}

// ---------------------------------------------------------------------------------------------
// K1: polar -> Cartesian (radar_utils.py:258-336)
// ---------------------------------------------------------------------------------------------
// One thread per output pixel.  range_tab / angle_tab are the (width x width) pixel-centre tables of
// form_cart_range_angle_grid (radar_utils.py:399-419); they depend only on the image width and the
// dtype and are computed once on the host with the reference's own torch calls, so that the only
// arithmetic done here is the per-scan part.  The scan's azimuth table sits in shared memory for the
// searchsorted of the "wobble fix" (:295-309).
struct PolarArgs {
  const void* scan; int scan_kind; size_t scan_row_stride; size_t scan_batch_stride;   // (B,A,R)
  const void* azimuths;          // (B,A) working dtype
  const void* range_tab; const void* angle_tab;   // (width*width)
  void* out;                     // (B,width,width)
  const void* grad_out; void* grad_scan;           // adjoint: (B,width,width) in, (B,A,R) accumulated
  int B, A, R, width;
  int interpolate_crossover, fix_wobble;
  double radar_res;
};

// BWD = true: the adjoint with respect to the scan (what autograd does for the grid_sample of
// radar_utils.py:334): the same taps, grad_out[pixel] * weight accumulated into grad_scan.
template <typename T, bool BWD>
__global__ void __launch_bounds__(ROW_THREADS) polar_to_cart_kernel(PolarArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* saz = reinterpret_cast<T*>(smem_raw);
  const int b = blockIdx.y;
  const T* az = static_cast<const T*>(a.azimuths) + (size_t)b * a.A;
  for (int i = threadIdx.x; i < a.A; i += blockDim.x) saz[i] = az[i];
  __syncthreads();

  // 16 x 16 pixel tile per block (a warp covers 16 x 2 pixels): neighbouring pixels sample
  // neighbouring polar bins, so a compact tile touches few 32-byte sectors of the scan
  const int npix = a.width * a.width;
  const int tiles_x = (a.width + 15) >> 4;
  const int px = (blockIdx.x % tiles_x) * 16 + (threadIdx.x & 15);
  const int py = (blockIdx.x / tiles_x) * 16 + (threadIdx.x >> 4);
  if (px >= a.width || py >= a.width) return;
  const int pix = py * a.width + px;
  const T range = static_cast<const T*>(a.range_tab)[pix];
  const T angle = static_cast<const T*>(a.angle_tab)[pix];
  const T res = (T)a.radar_res, half = (T)(a.radar_res / 2);

  T u = div_rn(add_rn(range, -half), res);                                   // :288
  T v;
  if (a.fix_wobble) {
    int lo = 0, hi = a.A;                    // searchsorted, side = left: first i with az[i] >= angle
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (saz[mid] < angle) lo = mid + 1; else hi = mid;
    }
    int c3 = lo;
    if (c3 == a.A) c3 -= 1;                                                  // :300
    int c2 = c3 - 1;
    if (c2 < 0) c2 += 1;                                                     // :302
    const T a3 = saz[c3], a2 = saz[c2];
    const T diff = add_rn(angle, -a3);
    T num = mul_rn(diff, diff < (T)0 ? (T)1 : (T)0);
    num = mul_rn(num, c3 > 0 ? (T)1 : (T)0);
    const T delta = div_rn(num, add_rn(add_rn(a3, -a2), (T)1e-14));          // :308
    v = add_rn((T)c3, delta);
  } else {
    const T step = div_rn(add_rn(saz[a.A - 1], -saz[0]), (T)(a.A - 1));      // :286
    v = div_rn(add_rn(angle, -saz[0]), step);                                // :290
  }
  if (u < (T)0) u = (T)0;                                                    // :314
  int H = a.A;
  if (a.interpolate_crossover) { H = a.A + 2; v = add_rn(v, (T)1); }         // :317-319
  const T gx = add_rn(mul_rn(div_rn(u, (T)(a.R - 1)), (T)2), (T)-1);         // :322
  const T gy = add_rn(mul_rn(div_rn(v, (T)(H - 1)), (T)2), (T)-1);           // :323
  const Taps<T> t = bilinear_taps<T>(gx, gy, a.R, H);

  if (BWD) {
    const T go = static_cast<const T*>(a.grad_out)[(size_t)b * npix + pix];
    T* gs = static_cast<T*>(a.grad_scan) + (size_t)b * a.A * a.R;
    auto put = [&](int x, int y, T w) {
      if (!inside(x, y, a.R, H)) return;
      int r = y;
      if (a.interpolate_crossover) r = y == 0 ? a.A - 1 : (y == a.A + 1 ? 0 : y - 1);
      atomicAdd(gs + (size_t)r * a.R + x, mul_rn(go, w));
    };
    put(t.x0, t.y0, t.nw); put(t.x0 + 1, t.y0, t.ne); put(t.x0, t.y0 + 1, t.sw); put(t.x0 + 1, t.y0 + 1, t.se);
    return;
  }
  Scan<T> sc; sc.base = a.scan; sc.kind = a.scan_kind; sc.row_stride = a.scan_row_stride;
  const size_t row0 = (size_t)b * a.A;
  auto fetch = [&](int x, int y) -> T {
    if (!inside(x, y, a.R, H)) return (T)0;
    int r = y;
    if (a.interpolate_crossover) r = y == 0 ? a.A - 1 : (y == a.A + 1 ? 0 : y - 1);   // wrapped pad rows
    return sc.at(row0 + r, x);
  };
  T acc = mul_rn(fetch(t.x0, t.y0), t.nw);
  acc = add_rn(acc, mul_rn(fetch(t.x0 + 1, t.y0), t.ne));
  acc = add_rn(acc, mul_rn(fetch(t.x0, t.y0 + 1), t.sw));
  acc = add_rn(acc, mul_rn(fetch(t.x0 + 1, t.y0 + 1), t.se));
  static_cast<T*>(a.out)[(size_t)b * npix + pix] = acc;
}

// ---------------------------------------------------------------------------------------------
// block-wide exclusive scan of one value per thread (blockDim.x <= 1024, multiple of 32)
// ---------------------------------------------------------------------------------------------
template <typename V>
__device__ __forceinline__ V block_exclusive_scan(V x, V* warp_tot /* >= 33 entries */, V& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  V inc = x;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const V y = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += y;
  }
  __syncthreads();                              // warp_tot may still be read from a previous call
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    V w = lane < nwarp ? warp_tot[lane] : V(0);
    V winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const V y = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += y;
    }
    if (lane < nwarp) warp_tot[lane] = winc - w;   // exclusive prefix of the warp totals
    if (lane == 31) warp_tot[32] = winc;           // block total (lanes >= nwarp carry the last sum)
  }
  __syncthreads();
  total = warp_tot[32];
  return warp_tot[warp] + inc - x;
}

// ---------------------------------------------------------------------------------------------
// K2a part 1: GO-CFAR mask (radar_utils.py:29-69)
// ---------------------------------------------------------------------------------------------
// One block per azimuth row.  The window sums of :46-53 are differences of an fp64 prefix sum of
// the row staged in shared memory: exact for uint8-derived scans whatever the order (a sum of 50
// such values needs < 40 bits), so the mask does not depend on how a host library orders a
// float32 sum.  thres = a * max(left, right) / w2 + b inside [mincol, maxcol), 1000 outside (:55-60).
struct CfarArgs {
  const void* scan; int scan_kind; size_t scan_row_stride;
  void* mask;                    // (rows, R): float when !diff, working dtype when diff
  int rows, R;
  int w2, guard, mincol, maxcol; // radar_utils.py:34-40
  int lo, hi;                    // columns [lo, hi) cover every window: staged in shared memory
  double a_thresh, b_thresh, steep;
  int diff;
};

// byte -> fp32 value table of load_radar (radar_utils.py:20-27: byte / 255 in fp32), built once per block so that
// the decode is one shared-memory load instead of an IEEE division per element
#define MMB_L1(i) ((float)(i) / 255.0f)
#define MMB_L4(i) MMB_L1(i), MMB_L1((i) + 1), MMB_L1((i) + 2), MMB_L1((i) + 3)
#define MMB_L16(i) MMB_L4(i), MMB_L4((i) + 4), MMB_L4((i) + 8), MMB_L4((i) + 12)
#define MMB_L64(i) MMB_L16(i), MMB_L16((i) + 16), MMB_L16((i) + 32), MMB_L16((i) + 48)
// fl32(i / 255), the division done once by the compiler in IEEE single precision
__device__ const float g_byte_lut[256] = {MMB_L64(0), MMB_L64(64), MMB_L64(128), MMB_L64(192)};
__device__ __forceinline__ void build_byte_lut(float* lut) {
  for (int i = threadIdx.x; i < 256; i += blockDim.x) lut[i] = g_byte_lut[i];
}
// element (row, col) of the scan; u8 rows go through the table
template <typename T>
__device__ __forceinline__ T scan_value(const Scan<T>& sc, const float* lut, size_t row, int col) {
  if (sc.kind == IN_U8) return (T)lut[static_cast<const uint8_t*>(sc.base)[row * sc.row_stride + 11 + col]];
  return sc.at(row, col);
}

// The prefix array lives in shared memory with one pad double after every 8 (logical i -> i + i / 8): the staging
// threads own 8 consecutive elements each (stride 9 doubles across the lanes: every bank once per half-warp), and
// the window lookups of consecutive columns stay within three wavefronts.  Unpadded, the 64-byte lane stride of
// the staging and the 32-byte stride of a four-column-per-thread lookup saturated the shared-memory pipe
// (profiles/r02_radar_ncu_summary.md: 3.8x the ideal wavefronts, LSU data pipe 95 % busy).
__host__ __device__ __forceinline__ int pidx(int i) { return i + (i >> 3); }
__host__ __device__ __forceinline__ size_t prefix_bytes(int entries) { return (size_t)(pidx(entries) + 2) * sizeof(double); }

// Stage x[lo, hi) as an inclusive fp64 prefix sum: pre[0] = 0, pre[i + 1] = x[lo] + ... + x[lo + i] (logical indices).
// Every thread takes 8 consecutive elements, scans them in registers, and one block scan of the thread totals
// supplies the offsets.  For uint8-derived scans every partial sum is exact in fp64 (the values are multiples of
// 2^-32 below 1), so the prefix does not depend on this blocking.
// `pad` entries after pre[len] repeat the total, so that a window that runs past hi (a slice truncated at the row
// end in the reference) needs no index clamp.
template <typename T>
__device__ __forceinline__ void stage_prefix(const Scan<T>& sc, const float* lut, size_t row, int lo, int hi, double* pre,
                                             double* wt, int pad = 0) {
  const int len = hi - lo;
  double carry = 0.0;
  const uint8_t* pb = static_cast<const uint8_t*>(sc.base) + row * sc.row_stride + 11 + lo;
  const float* pf = static_cast<const float*>(sc.base) + row * sc.row_stride + lo;
  const double* pd = static_cast<const double*>(sc.base) + row * sc.row_stride + lo;
  for (int base = 0; base < len; base += 8 * (int)blockDim.x) {
    const int s = base + 8 * (int)threadIdx.x;
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.0;
    if (s < len) {
      if (sc.kind == IN_U8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) if (s + k < len) v[k] = (double)lut[pb[s + k]];
      } else if (sc.kind == IN_F32) {
#pragma unroll
        for (int k = 0; k < 8; ++k) if (s + k < len) v[k] = (double)pf[s + k];
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) if (s + k < len) v[k] = pd[s + k];
      }
#pragma unroll
      for (int k = 1; k < 8; ++k) v[k] += v[k - 1];
    }
    double total;
    const double off = block_exclusive_scan<double>(v[7], wt, total) + carry;
    if (s < len) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (s + k < len) pre[pidx(s + k + 1)] = v[k] + off;
    }
    carry += total;
  }
  if (threadIdx.x == 0) pre[0] = 0.0;
  for (int i = threadIdx.x; i < pad; i += blockDim.x) pre[pidx(len + 1 + i)] = carry;
  __syncthreads();
}

// Forward, the shipped form: one thread per output pixel does the scans of a batch group one after the other.
// What depends only on the pixel -- its range / angle, u = (range - res/2) / res and the normalised gx (:288,:322) --
// is evaluated once per thread instead of once per scan; the searchsorted of the wobble fix starts from the index a
// uniform azimuth spacing would give and walks to the exact answer (one or two comparisons for encoder jitter,
// correct for any sorted table); raw Navtech bytes are decoded through a 256-entry table in shared memory instead of
// an IEEE division per tap.  Every operation that produces a value is the same rounded operation, in the same
// order, as in the one-scan kernel above (which the adjoint still uses), so the outputs are bit-identical.
constexpr int POLAR_BGROUP = 8;
template <typename T>
__global__ void __launch_bounds__(ROW_THREADS) polar_to_cart_fwd_kernel(PolarArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* saz_all = reinterpret_cast<T*>(smem_raw);                 // (nb, A)
  __shared__ float lut[256];
  const int b0 = blockIdx.y * POLAR_BGROUP, nb = min(POLAR_BGROUP, a.B - b0);
  const T* az = static_cast<const T*>(a.azimuths) + (size_t)b0 * a.A;
  for (int i = threadIdx.x; i < nb * a.A; i += blockDim.x) saz_all[i] = az[i];
  if (a.scan_kind == IN_U8) build_byte_lut(lut);
  __syncthreads();

  const int npix = a.width * a.width;
  const int tiles_x = (a.width + 15) >> 4;
  const int px = (blockIdx.x % tiles_x) * 16 + (threadIdx.x & 15);
  const int py = (blockIdx.x / tiles_x) * 16 + (threadIdx.x >> 4);
  if (px >= a.width || py >= a.width) return;
  const int pix = py * a.width + px;
  const T range = static_cast<const T*>(a.range_tab)[pix];
  const T angle = static_cast<const T*>(a.angle_tab)[pix];
  const T res = (T)a.radar_res, half = (T)(a.radar_res / 2);
  T u = div_rn(add_rn(range, -half), res);                                   // :288
  if (u < (T)0) u = (T)0;                                                    // :314
  const T gx = add_rn(mul_rn(div_rn(u, (T)(a.R - 1)), (T)2), (T)-1);         // :322
  const int H = a.interpolate_crossover ? a.A + 2 : a.A;
  // the x half of bilinear_taps (same operations): column, its two weights, which of the two columns exist
  const T ix = mul_rn(add_rn(gx, (T)1), (T)(a.R - 1) / (T)2);
  const T xw = t_floor(ix);
  const T wx = add_rn(ix, -xw), ex = add_rn((T)1, -wx);
  const int x0 = (int)fmax(fmin((double)xw, 2.0e9), -2.0e9);
  const bool x0ok = x0 >= 0 && x0 < a.R, x1ok = x0 + 1 >= 0 && x0 + 1 < a.R;
  const T hy = (T)(H - 1) / (T)2;
  int guess = (int)((float)angle * ((float)a.A * 0.15915494f));             // angle / (2 pi / A): a starting point only
  guess = min(max(guess, 0), a.A);
  const bool u8 = a.scan_kind == IN_U8;
  const size_t scan_stride = (size_t)a.A * a.scan_row_stride;
  const uint8_t* sb = static_cast<const uint8_t*>(a.scan) + (size_t)b0 * scan_stride + 11;
  const T* st = static_cast<const T*>(a.scan) + (size_t)b0 * scan_stride;
  const int rs = (int)a.scan_row_stride;

  for (int bb = 0; bb < nb; ++bb, sb += scan_stride, st += scan_stride) {
    const T* saz = saz_all + bb * a.A;
    T v;
    if (a.fix_wobble) {
      int c3 = guess;                        // searchsorted, side = left: first i with az[i] >= angle
      while (c3 > 0 && !(saz[c3 - 1] < angle)) --c3;
      while (c3 < a.A && saz[c3] < angle) ++c3;
      if (c3 == a.A) c3 -= 1;                                                // :300
      int c2 = c3 - 1;
      if (c2 < 0) c2 += 1;                                                   // :302
      const T a3 = saz[c3], a2 = saz[c2];
      const T diff = add_rn(angle, -a3);
      T num = mul_rn(diff, diff < (T)0 ? (T)1 : (T)0);
      num = mul_rn(num, c3 > 0 ? (T)1 : (T)0);
      const T delta = div_rn(num, add_rn(add_rn(a3, -a2), (T)1e-14));        // :308
      v = add_rn((T)c3, delta);
    } else {
      const T step = div_rn(add_rn(saz[a.A - 1], -saz[0]), (T)(a.A - 1));    // :286
      v = div_rn(add_rn(angle, -saz[0]), step);                              // :290
    }
    if (a.interpolate_crossover) v = add_rn(v, (T)1);                        // :317-319
    const T gy = add_rn(mul_rn(div_rn(v, (T)(H - 1)), (T)2), (T)-1);         // :323
    // the y half of bilinear_taps
    const T iy = mul_rn(add_rn(gy, (T)1), hy);
    const T yn = t_floor(iy);
    const T ny = add_rn(iy, -yn), sy = add_rn((T)1, -ny);
    const int y0 = (int)fmax(fmin((double)yn, 2.0e9), -2.0e9);
    // rows of the (wrap-padded) scan behind y0 and y0 + 1, or -1
    auto scan_row = [&](int y) -> int {
      if (y < 0 || y >= H) return -1;
      if (!a.interpolate_crossover) return y;
      return y == 0 ? a.A - 1 : (y == a.A + 1 ? 0 : y - 1);                  // wrapped pad rows
    };
    const int r0 = scan_row(y0), r1 = scan_row(y0 + 1);
    auto tap = [&](int r, bool xok, int x) -> T {
      if (r < 0 || !xok) return (T)0;
      const int o = r * rs + x;
      return u8 ? (T)lut[sb[o]] : st[o];
    };
    T acc = mul_rn(tap(r0, x0ok, x0), mul_rn(sy, ex));
    acc = add_rn(acc, mul_rn(tap(r0, x1ok, x0 + 1), mul_rn(sy, wx)));
    acc = add_rn(acc, mul_rn(tap(r1, x0ok, x0), mul_rn(ny, ex)));
    acc = add_rn(acc, mul_rn(tap(r1, x1ok, x0 + 1), mul_rn(ny, wx)));
    static_cast<T*>(a.out)[(size_t)(b0 + bb) * npix + pix] = acc;
  }
}

__device__ __forceinline__ double cfar_threshold_at(const CfarArgs& a, const double* pre, int j) {
  if (j < a.mincol || j >= a.maxcol) return 1000.0;
  auto P = [&](int k) { return pre[pidx(min(k, a.hi) - a.lo)]; };     // slices past the row end are truncated
  const double left = P(j - a.guard) - P(j - a.w2 - a.guard);
  const double right = P(j + a.w2 + a.guard + 1) - P(j + a.guard + 1);
  return a.a_thresh * (fmax(left, right) / (double)a.w2) + a.b_thresh;
}

template <typename T>
__device__ __forceinline__ T cfar_mask_value(const CfarArgs& a, T x, double thres) {
  if (!a.diff) return (double)x > thres ? (T)1 : (T)0;                                   // :67
  const T soft = add_rn(mul_rn((T)0.5, t_tanh<T>(add_rn(mul_rn((T)a.steep, add_rn(x, -(T)thres)), (T)2.5))), (T)0.5);   // :64
  return (soft > (T)0.99 || soft < (T)-0.99) ? soft : (T)0;                              // hardshrink :65
}

// Non-differentiable mask bit x > thres (radar_utils.py:67).  The window sums come from the prefix; the threshold is
// first formed without the fp64 division (one fma with a / w2), and only when x lies within 1e-12 (1 + |t|) of it --
// far outside the rounding of either form -- is the reference's own expression evaluated: same bits, no division on
// the common path.
__device__ __forceinline__ bool cfar_hard_bit(const CfarArgs& a, const double* pre, int j, double x, double a_over_w2) {
  auto P = [&](int k) { return pre[pidx(min(k, a.hi) - a.lo)]; };
  const double left = P(j - a.guard) - P(j - a.w2 - a.guard);
  const double right = P(j + a.w2 + a.guard + 1) - P(j + a.guard + 1);
  const double t = fma(fmax(left, right), a_over_w2, a.b_thresh);
  if (fabs(x - t) > 1.0e-12 * (1.0 + fabs(t))) return x > t;
  return x > a.a_thresh * (fmax(left, right) / (double)a.w2) + a.b_thresh;
}

// !diff: float mask, four columns per thread, 16-byte stores; columns outside [mincol, maxcol) are zero and read
// nothing.  The prefix is padded past hi (see stage_prefix), so the window ends need no clamp.
template <typename T>
__global__ void __launch_bounds__(ROW_THREADS, 8) cfar_mask_hard_kernel(CfarArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* pre = reinterpret_cast<double*>(smem_raw);
  __shared__ double wt[33];
  __shared__ float lut[256];
  const size_t row = blockIdx.x;
  Scan<T> sc; sc.base = a.scan; sc.kind = a.scan_kind; sc.row_stride = a.scan_row_stride;
  if (a.scan_kind == IN_U8) { build_byte_lut(lut); __syncthreads(); }
  const int pad = a.w2 + a.guard + 2;
  stage_prefix<T>(sc, lut, row, a.lo, a.hi, pre, wt, pad);
  const double aw = a.a_thresh / (double)a.w2, bt = a.b_thresh;
  const uint8_t* pb = static_cast<const uint8_t*>(sc.base) + row * sc.row_stride + 11;
  const T* pt = static_cast<const T*>(sc.base) + row * sc.row_stride;
  const bool u8 = a.scan_kind == IN_U8;
  // window ends relative to column j, as indices into pre (which starts at column lo)
  const int oA = -a.guard - a.lo, oB = -a.w2 - a.guard - a.lo, oC = a.w2 + a.guard + 1 - a.lo, oD = a.guard + 1 - a.lo;
  float* out = static_cast<float*>(a.mask) + row * a.R;
  // zeros outside [mincol, maxcol): 16-byte stores where the row allows it
  if ((a.R & 3) == 0) {
    const int z0 = (a.mincol >> 2) << 2, z1 = ((a.maxcol + 3) >> 2) << 2;      // quads fully outside the window
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = 4 * threadIdx.x; j < z0; j += 4 * blockDim.x) *reinterpret_cast<float4*>(out + j) = zero;
    for (int j = z1 + 4 * threadIdx.x; j < a.R; j += 4 * blockDim.x) *reinterpret_cast<float4*>(out + j) = zero;
    for (int j = z0 + threadIdx.x; j < a.mincol; j += blockDim.x) out[j] = 0.f;
    for (int j = a.maxcol + threadIdx.x; j < min(z1, a.R); j += blockDim.x) out[j] = 0.f;
  } else {
    for (int j = threadIdx.x; j < a.mincol; j += blockDim.x) out[j] = 0.f;
    for (int j = a.maxcol + threadIdx.x; j < a.R; j += blockDim.x) out[j] = 0.f;
  }
  // one column per lane: consecutive lanes read consecutive prefix entries and write consecutive floats
  for (int j = a.mincol + threadIdx.x; j < a.maxcol; j += blockDim.x) {
    const double left = pre[pidx(j + oA)] - pre[pidx(j + oB)], right = pre[pidx(j + oC)] - pre[pidx(j + oD)];
    const double mx = fmax(left, right);
    const double t = fma(mx, aw, bt);
    const double x = u8 ? (double)lut[pb[j]] : (double)pt[j];
    bool one = x > t;
    if (fabs(x - t) <= 1.0e-12 * (1.0 + fabs(t))) one = x > a.a_thresh * (mx / (double)a.w2) + bt;   // the reference's own form
    out[j] = one ? 1.f : 0.f;
  }
}

template <typename T, bool DIFF>
__global__ void __launch_bounds__(ROW_THREADS) cfar_mask_kernel(CfarArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* pre = reinterpret_cast<double*>(smem_raw);
  __shared__ double wt[33];
  const size_t row = blockIdx.x;
  __shared__ float lut[256];
  Scan<T> sc; sc.base = a.scan; sc.kind = a.scan_kind; sc.row_stride = a.scan_row_stride;
  if (a.scan_kind == IN_U8) { build_byte_lut(lut); __syncthreads(); }
  stage_prefix<T>(sc, lut, row, a.lo, a.hi, pre, wt);
  typedef typename std::conditional<DIFF, T, float>::type OT;
  OT* out = static_cast<OT*>(a.mask) + row * a.R;
  for (int j = threadIdx.x; j < a.R; j += blockDim.x) {
    OT m = (OT)0;
    if (j >= a.mincol && j < a.maxcol) m = (OT)cfar_mask_value<T>(a, sc.at(row, j), cfar_threshold_at(a, pre, j));
    out[j] = m;
  }
}

// ---------------------------------------------------------------------------------------------
// K2a part 2: peak extraction (radar_utils.py:71-106, 167-195)
// ---------------------------------------------------------------------------------------------
// arr[k] = (res * k) * mask[k]; the "peak" array is
//     v[k] = arr[k] * z[k+1] + arr[k+1] * z[k]   (k < R-1),  v[R-1] = 0,
// z = (arr == 0) (non-diff) or 1 - tanh(steep * arr) (diff).  The non-zero v of the flattened
// scan, in order, are paired (2i, 2i+1) and averaged into (rho, phi) -> (rho cos phi, rho sin phi, 0).
// Three kernels:
//   rows:    one block per azimuth row, arr staged in shared memory: the non-zero v of the row, in range
//            order, go to the row's slot of a staging area (rowcap values per row; only the entries
//            that exist are written) and their number to counts (B*A).  The row is evaluated ONCE.
//   offsets: exclusive scan of the A counts per scan -> offsets (B, A+1)      [one block per scan]
//   points:  one thread per point: entries (2i, 2i+1) of the flattened scan are located through the
//            offsets (binary search over the rows) and read from the staging area.
// The row source is either a mask tensor (the reference's extract_pc signature) or the CFAR mask
// evaluated on the fly from the scan (fused cfar + extraction: the mask is never written).
struct PeakArgs {
  // mask source
  const void* mask; size_t mask_row_stride;        // (rows, R) working dtype, or nullptr for the fused source
  CfarArgs cfar;                                    // fused source (non-diff or diff)
  int rows, A, R;
  float res;
  int diff; double steep;
  int* counts;                  // (rows)
  void* stage; int rowcap;      // (rows, rowcap) working dtype: the row's non-zero v in range order
};

// columns [c0, c1) that can hold a non-zero v: with the fused source the mask is zero outside
// [mincol, maxcol), so entries can only sit in [mincol-1, maxcol)
__host__ __device__ __forceinline__ void peak_span(const PeakArgs& a, bool fused, int& c0, int& c1) {
  c0 = fused ? (a.cfar.mincol - 1 > 0 ? a.cfar.mincol - 1 : 0) : 0;
  c1 = fused ? (a.cfar.maxcol < a.R ? a.cfar.maxcol : a.R) : a.R;
  if (c1 < c0) c1 = c0;
}

template <typename T, bool FUSED>
__device__ __forceinline__ void stage_arr(const PeakArgs& a, size_t row, T* arr, double* pre, double* wt) {
  if (FUSED) {
    __shared__ float lut[256];
    Scan<T> sc; sc.base = a.cfar.scan; sc.kind = a.cfar.scan_kind; sc.row_stride = a.cfar.scan_row_stride;
    if (a.cfar.scan_kind == IN_U8) { build_byte_lut(lut); __syncthreads(); }
    stage_prefix<T>(sc, lut, row, a.cfar.lo, a.cfar.hi, pre, wt);
    const double aw = a.cfar.a_thresh / (double)a.cfar.w2;
    const int k0 = max(a.cfar.mincol - 1, 0), k1 = min(a.cfar.maxcol + 1, a.R);   // arr is read on [mincol-1, maxcol]
    for (int k = k0 + threadIdx.x; k < k1; k += blockDim.x) {
      T m = (T)0;
      if (k >= a.cfar.mincol && k < a.cfar.maxcol) {
        const T x = scan_value<T>(sc, lut, row, k);
        if (a.cfar.diff) m = cfar_mask_value<T>(a.cfar, x, cfar_threshold_at(a.cfar, pre, k));
        else m = cfar_hard_bit(a.cfar, pre, k, (double)x, aw) ? (T)1 : (T)0;
      }
      arr[k] = mul_rn((T)__fmul_rn(a.res, (float)k), m);
    }
  } else {
    const T* mrow = static_cast<const T*>(a.mask) + row * a.mask_row_stride;
    for (int k = threadIdx.x; k < a.R; k += blockDim.x) arr[k] = mul_rn((T)__fmul_rn(a.res, (float)k), mrow[k]);
  }
  __syncthreads();
}

template <typename T>
__device__ __forceinline__ T peak_value(const PeakArgs& a, const T* arr, int k) {
  if (k >= a.R - 1) return (T)0;
  const T x0 = arr[k], x1 = arr[k + 1];
  T z0, z1;
  if (a.diff) {
    z0 = add_rn((T)1, -t_tanh<T>(mul_rn((T)a.steep, x0)));
    z1 = add_rn((T)1, -t_tanh<T>(mul_rn((T)a.steep, x1)));
  } else {
    z0 = x0 == (T)0 ? (T)1 : (T)0;
    z1 = x1 == (T)0 ? (T)1 : (T)0;
  }
  return add_rn(mul_rn(x0, z1), mul_rn(x1, z0));
}

template <typename T, bool FUSED>
__global__ void __launch_bounds__(ROW_THREADS) peak_rows_kernel(PeakArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* pre = reinterpret_cast<double*>(smem_raw);                       // fused source only
  T* arr = reinterpret_cast<T*>(smem_raw + (FUSED ? prefix_bytes(a.cfar.hi - a.cfar.lo + 1) : 0));
  __shared__ double wt[33];
  __shared__ int wi[33];
  const size_t row = blockIdx.x;
  stage_arr<T, FUSED>(a, row, arr, pre, wt);

  // each thread owns a contiguous chunk so that ordinals follow the range order
  int c0, c1;
  peak_span(a, FUSED, c0, c1);
  const int chunk = (c1 - c0 + blockDim.x - 1) / blockDim.x;
  const int s = c0 + threadIdx.x * chunk, e = min(s + chunk, c1);
  int cnt = 0;
  for (int k = s; k < e; ++k) cnt += peak_value<T>(a, arr, k) != (T)0 ? 1 : 0;
  int total;
  int pos = block_exclusive_scan<int>(cnt, wi, total);
  if (threadIdx.x == 0) a.counts[row] = total;
  if (cnt == 0) return;
  T* sv = static_cast<T*>(a.stage) + row * (size_t)a.rowcap;               // total <= c1 - c0 = rowcap
  for (int k = s; k < e; ++k) {
    const T v = peak_value<T>(a, arr, k);
    if (v != (T)0) sv[pos++] = v;
  }
}

// offsets (B, A+1): exclusive scan of each scan's row counts; offsets[b][A] = number of entries.
__global__ void row_offsets_kernel(const int* __restrict__ counts, int A, int* __restrict__ offsets,
                                   int* __restrict__ totals) {
  __shared__ int wi[33];
  const int b = blockIdx.x;
  int carry = 0;
  for (int base = 0; base < A; base += blockDim.x) {
    const int r = base + threadIdx.x;
    const int c = r < A ? counts[(size_t)b * A + r] : 0;
    int total;
    const int ex = block_exclusive_scan<int>(c, wi, total);
    if (r < A) offsets[(size_t)b * (A + 1) + r] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    offsets[(size_t)b * (A + 1) + A] = carry;
    if (totals) totals[b] = carry;
  }
}

// One thread per point: entries (2i, 2i+1) -> (rho, phi) -> xyz (radar_utils.py:96-104, 187-195).
// The two entries are consecutive in the FLATTENED scan, so the second may sit in a later row.
template <typename T>
__global__ void peak_points_kernel(const T* __restrict__ stage, int rowcap, const int* __restrict__ offsets,
                                   const T* __restrict__ azimuths, const T* __restrict__ T_ab, int A, int cap,
                                   int out_cap, T* __restrict__ ent_val, int* __restrict__ ent_row,
                                   T* __restrict__ out) {
  const int b = blockIdx.y;
  const int* off = offsets + (size_t)b * (A + 1);
  const int npts = off[A] / 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npts || i >= out_cap || 2 * i + 1 >= cap) return;
  const int e0 = 2 * i, e1 = e0 + 1;
  int r0 = 0, hi = A;                            // off[r0] <= e0 < off[hi]; empty rows share an offset,
  while (hi - r0 > 1) {                          // so the LAST row with off <= e0 is the one holding e0
    const int mid = (r0 + hi) >> 1;
    if (off[mid] <= e0) r0 = mid; else hi = mid;
  }
  int r1 = r0;
  while (off[r1 + 1] <= e1) ++r1;                // e1 < off[A]: terminates
  const T* sb = stage + (size_t)b * A * rowcap;
  const T v0 = sb[(size_t)r0 * rowcap + (e0 - off[r0])], v1 = sb[(size_t)r1 * rowcap + (e1 - off[r1])];
  T* ev = ent_val + (size_t)b * cap;
  int* er = ent_row + (size_t)b * cap;
  ev[e0] = v0; ev[e1] = v1; er[e0] = r0; er[e1] = r1;
  const T rho = div_rn(add_rn(v1, v0), (T)2);
  const T phi = div_rn(add_rn(azimuths[(size_t)b * A + r1], azimuths[(size_t)b * A + r0]), (T)2);
  // cos/sin: correctly rounded (evaluated in fp64, rounded once), then the reference's product;
  // the host libraries the reference runs on differ from each other by an ulp here.
  T x = mul_rn(rho, (T)cos((double)phi)), y = mul_rn(rho, (T)sin((double)phi)), z = (T)0;
  if (T_ab) {
    const T* M = T_ab + (size_t)b * 16;
    const T xx = add_rn(add_rn(mul_rn(M[0], x), mul_rn(M[1], y)), M[3]);
    const T yy = add_rn(add_rn(mul_rn(M[4], x), mul_rn(M[5], y)), M[7]);
    const T zz = add_rn(add_rn(mul_rn(M[8], x), mul_rn(M[9], y)), M[11]);
    x = xx; y = yy; z = zz;
  }
  T* o = out + ((size_t)b * out_cap + i) * 3;
  o[0] = x; o[1] = y; o[2] = z;
}

// ---------------------------------------------------------------------------------------------
// K2b: K-strongest per azimuth (vtr_radar detector.inl:39-85, KStrongest<PointT>::run)
// ---------------------------------------------------------------------------------------------
// Per row: mean over bins [j0, j1) accumulated in double and divided by the DOUBLE N = maxcol - mincol
// (not the bin count, detector.inl:47-61); thres = mean * t2 + t3; among bins with value >= thres
// take up to k by descending value, ties to the lower bin (the C++ comparator is ill-typed and its
// order implementation-defined, SURVEY.md 8a note; this is the intended order); rho = float(j) * res
// + range_offset; (x, y) = rho (cos phi, sin phi) (navtech_extraction_module.cpp:33-39).
struct KStrongArgs {
  const void* scan; int scan_kind; size_t scan_row_stride;
  int rows, A, R, j0, j1, k;
  double N, t2, t3;
  int* sel_bin;                 // (rows, k)
  int* counts;                  // (rows)
};

__device__ __forceinline__ unsigned int ordered_bits(float v) {
  const unsigned int u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// One WARP per azimuth row, KSTRONG_WARPS rows per block, no block-wide synchronisation inside the selection.
//   pass 1: the lanes stride over the row's bins (bytes decoded through the shared table), summing in fp64 -- the
//           mean of detector.inl:57-61, exact for uint8-derived scans whatever the order;
//   pass 2: the bins at or above the threshold are compacted (ballot) into the warp's list of 64-bit keys
//           (ordered value bits || inverted bin index), in bin order, up to KSTRONG_LIST entries;
//   pass 3: k rounds of a warp-wide maximum over the list (two 32-bit REDUX), the winner is cleared.
//   A row with more candidates than the list holds (a saturated row) is selected without storage instead: round r
//   takes the largest key below the key of round r - 1, re-reading the row (L1-resident) every round.
// Order within a row: descending value, ties to the lower bin -- the choice the literal reference leaves open
// (tests/kstrong_pin.py).
constexpr int KSTRONG_WARPS = 8;
constexpr int KSTRONG_LIST = 256;
__global__ void __launch_bounds__(KSTRONG_WARPS * 32) kstrong_rows_kernel(KStrongArgs a) {
  __shared__ unsigned long long keys_all[KSTRONG_WARPS][KSTRONG_LIST];
  __shared__ float lut[256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int len = a.j1 - a.j0;
  unsigned long long* keys = keys_all[warp];
  if (a.scan_kind == IN_U8) build_byte_lut(lut);
  __syncthreads();
  const size_t row = (size_t)blockIdx.x * KSTRONG_WARPS + warp;
  if (row >= (size_t)a.rows) return;                        // whole warp
  const bool u8 = a.scan_kind == IN_U8;
  const uint8_t* pb = static_cast<const uint8_t*>(a.scan) + row * a.scan_row_stride + 11 + a.j0;
  const float* pf = static_cast<const float*>(a.scan) + row * a.scan_row_stride + a.j0;
  auto value = [&](int i) -> float { return u8 ? lut[pb[i]] : pf[i]; };
  auto key_of = [&](float v, int i) -> unsigned long long {
    return ((unsigned long long)ordered_bits(v) << 32) | (unsigned int)(0x7fffffff - i);
  };
  double sum = 0.0;
  for (int i = lane; i < len; i += 32) sum += (double)value(i);
  sum = warp_sum(sum);
  const double thres = (sum / a.N) * a.t2 + a.t3;
  int cnt = 0;
  for (int base = 0; base < len; base += 32) {
    const int i = base + lane;
    const float v = i < len ? value(i) : 0.f;
    const bool pass = i < len && (double)v >= thres;
    const unsigned m = __ballot_sync(0xffffffffu, pass);
    const int at = cnt + __popc(m & ((1u << lane) - 1u));
    if (pass && at < KSTRONG_LIST) keys[at] = key_of(v, i);
    cnt += __popc(m);
  }
  __syncwarp();
  int taken = 0;
  if (cnt <= KSTRONG_LIST) {
    for (int round = 0; round < a.k && round < cnt; ++round) {
      unsigned long long best = 0ull; int where = -1;
      for (int i = lane; i < cnt; i += 32) {
        const unsigned long long key = keys[i];
        if (key > best) { best = key; where = i; }
      }
      const unsigned hi = (unsigned)(best >> 32), lo = (unsigned)best;
      const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
      const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
      if (hi == mhi && lo == mlo && where >= 0) {            // keys are unique: exactly one lane
        keys[where] = 0ull;
        a.sel_bin[row * a.k + round] = a.j0 + (0x7fffffff - (int)mlo);
      }
      ++taken;
      __syncwarp();
    }
  } else {
    unsigned long long prev = ~0ull;
    for (int round = 0; round < a.k && round < cnt; ++round) {
      unsigned long long best = 0ull;
      for (int i = lane; i < len; i += 32) {
        const float v = value(i);
        if ((double)v >= thres) {
          const unsigned long long key = key_of(v, i);
          if (key < prev && key > best) best = key;
        }
      }
      const unsigned hi = (unsigned)(best >> 32), lo = (unsigned)best;
      const unsigned mhi = __reduce_max_sync(0xffffffffu, hi);
      const unsigned mlo = __reduce_max_sync(0xffffffffu, hi == mhi ? lo : 0u);
      prev = ((unsigned long long)mhi << 32) | mlo;
      if (lane == 0) a.sel_bin[row * a.k + round] = a.j0 + (0x7fffffff - (int)mlo);
      ++taken;
    }
  }
  if (lane == 0) a.counts[row] = taken;
}

template <typename T>
__global__ void kstrong_points_kernel(const int* __restrict__ sel_bin, const int* __restrict__ counts,
                                      const int* __restrict__ offsets, const T* __restrict__ azimuths, int A, int k,
                                      float res, double range_offset, int out_cap, T* __restrict__ out,
                                      int* __restrict__ out_row, int* __restrict__ out_bin) {
  const int b = blockIdx.y;
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= A * k) return;
  const int r = g / k, slot = g % k;
  const size_t row = (size_t)b * A + r;
  if (slot >= counts[row]) return;
  const int pos = offsets[(size_t)b * (A + 1) + r] + slot;
  if (pos >= out_cap) return;
  const int j = sel_bin[row * k + slot];
  const float rho = (float)((double)__fmul_rn((float)j, res) + range_offset);
  const T phi = azimuths[row];
  const size_t o = (size_t)b * out_cap + pos;
  out[o * 3 + 0] = mul_rn((T)rho, (T)cos((double)phi));
  out[o * 3 + 1] = mul_rn((T)rho, (T)sin((double)phi));
  out[o * 3 + 2] = (T)0;
  if (out_row) out_row[o] = r;
  if (out_bin) out_bin[o] = j;
}

// ---------------------------------------------------------------------------------------------
// host-side dispatch
// ---------------------------------------------------------------------------------------------
static int scan_kind_of(int32_t dtype, int32_t scan_u8) { return scan_u8 ? IN_U8 : (dtype == MMB_F64 ? IN_F64 : IN_F32); }

static int fill_cfar(const mmb_cfar_params* p, const void* scan, CfarArgs& c) {
  if (!p) return MMB_ERR_NULL_POINTER;
  if (p->B <= 0 || p->A <= 0 || p->R <= 1) return MMB_ERR_INVALID_SHAPE;
  if (p->dtype != MMB_F32 && p->dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (p->scan_u8 && p->dtype != MMB_F32) return MMB_ERR_UNSUPPORTED;
  if (!(p->res > 0.0)) return MMB_ERR_INVALID_SHAPE;
  int width = p->width;
  if (width % 2 == 0) width += 1;                                            // radar_utils.py:34
  c.w2 = width / 2; c.guard = p->guard;
  if (c.w2 <= 0 || c.guard < 0) return MMB_ERR_INVALID_SHAPE;
  const double lo = p->minr / p->res + c.w2 + c.guard + 1, hi = p->maxr / p->res - c.w2 - c.guard;
  c.mincol = lo < 0 ? 0 : (int)lo;                                           // max(0, int(...)) :38
  c.maxcol = (int)hi;
  if (c.maxcol > p->R) c.maxcol = p->R;                                      // min(R, int(...)) :39
  if (c.maxcol < c.mincol) c.maxcol = c.mincol;
  c.lo = c.mincol - c.w2 - c.guard;
  if (c.lo < 0) return MMB_ERR_UNSUPPORTED;        // a negative slice start wraps around in the reference
  c.hi = c.maxcol + c.w2 + c.guard + 1;
  if (c.hi > p->R) c.hi = p->R;
  if (c.hi < c.lo) c.hi = c.lo;
  c.scan = scan; c.scan_kind = scan_kind_of(p->dtype, p->scan_u8);
  c.scan_row_stride = p->scan_u8 ? (size_t)p->R + 11 : (size_t)p->R;
  c.rows = p->B * p->A; c.R = p->R;
  c.a_thresh = p->a_thresh; c.b_thresh = p->b_thresh; c.steep = p->steep_fact; c.diff = p->diff ? 1 : 0;
  c.mask = nullptr;
  return MMB_OK;
}

template <typename K> static int opt_in_smem(K kernel, size_t bytes) {
  if (bytes > 227 * 1024) return MMB_ERR_UNSUPPORTED;
  if (bytes > 48 * 1024 &&
      cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) {
    set_last_cuda_error(cudaGetLastError());
    return MMB_ERR_CUDA;
  }
  return MMB_OK;
}

struct ExtractWs { int* counts; int* offsets; void* stage; size_t bytes; };
static ExtractWs carve_extract(void* base, int B, int A, int rowcap, size_t elt) {
  Carver c(base);
  ExtractWs w;
  w.counts = c.take<int>((size_t)B * A);
  w.offsets = c.take<int>((size_t)B * (A + 1));
  w.stage = c.take<unsigned char>((size_t)B * A * rowcap * elt);
  w.bytes = c.off;
  return w;
}

template <typename T, bool FUSED>
static int launch_peak_rows(const PeakArgs& a, cudaStream_t st) {
  size_t smem = (size_t)a.R * sizeof(T) + 16;
  if (FUSED) smem += prefix_bytes(a.cfar.hi - a.cfar.lo + 1);
  int rc = opt_in_smem(peak_rows_kernel<T, FUSED>, smem);
  if (rc) return rc;
  peak_rows_kernel<T, FUSED><<<a.rows, ROW_THREADS, smem, st>>>(a);
  return check_launch();
}

static int fill_peak(const mmb_extract_params* p, const void* src, PeakArgs& a, bool need_src = true) {
  if (!p || (need_src && !src)) return MMB_ERR_NULL_POINTER;
  if (p->B <= 0 || p->A <= 0 || p->R <= 1) return MMB_ERR_INVALID_SHAPE;
  if (p->dtype != MMB_F32 && p->dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  a.mask = nullptr; a.mask_row_stride = p->R;
  a.rows = p->B * p->A; a.A = p->A; a.R = p->R;
  a.res = (float)p->res; a.diff = p->diff ? 1 : 0; a.steep = p->steep_fact;
  a.counts = nullptr; a.stage = nullptr; a.rowcap = p->R;
  if (p->fused) {
    if (p->cfar.B != p->B || p->cfar.A != p->A || p->cfar.R != p->R || p->cfar.dtype != p->dtype)
      return MMB_ERR_INVALID_SHAPE;
    const int rc = fill_cfar(&p->cfar, src, a.cfar);
    if (rc) return rc;
  } else {
    a.mask = src;
  }
  int c0, c1;
  peak_span(a, p->fused != 0, c0, c1);
  a.rowcap = c1 - c0 > 0 ? c1 - c0 : 1;
  return MMB_OK;
}

// ---------------------------------------------------------------------------------------------
// N2: occupancy BEV of a point cloud (radar_utils.py:142-165, `extract_bev_from_pts`)
// ---------------------------------------------------------------------------------------------
// idx = (-x / res + half, y / res + half) (point_to_cart_idx, :374-397; half = 320, its DEFAULT width / 2,
// whatever W is, :144); a coordinate < 0 or > W-1
// is replaced by W//2 (:147-148, per coordinate); the four floor/ceil pixels get 1 (:156-159); the
// centre pixel, which collects the out-of-range and padding points, ends as 0 (:163).  Writes are
// idempotent, so no atomics; the caller zero-fills `bev`.
template <typename T>
__global__ void bev_from_points_kernel(const T* __restrict__ pts, int pt_cols, int N, int n, int W, T cart_res,
                                       T half, T* __restrict__ bev) {
  const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= (size_t)N * n) return;
  const int b = (int)(g / n);
  const T* p = pts + g * pt_cols;
  T u = add_rn(div_rn(-p[0], cart_res), half), v = add_rn(div_rn(p[1], cart_res), half);
  const T mid = (T)(W / 2);
  if (u < (T)0) u = mid;
  if (u > (T)(W - 1)) u = mid;
  if (v < (T)0) v = mid;
  if (v > (T)(W - 1)) v = mid;
  if (!(u >= (T)0 && u <= (T)(W - 1) && v >= (T)0 && v <= (T)(W - 1))) return;      // NaN coordinates
  const int uf = (int)t_floor(u), uc = (int)ceil((double)u), vf = (int)t_floor(v), vc = (int)ceil((double)v);
  T* img = bev + (size_t)b * W * W;
  const int c = W / 2;
  auto put = [&](int r, int col) { if (!(r == c && col == c)) img[(size_t)r * W + col] = (T)1; };
  put(uc, vf); put(uc, vc); put(uf, vf); put(uf, vc);
}

// ---------------------------------------------------------------------------------------------
// N3: Cartesian -> polar resampling (radar_utils.py:338-372, `radar_cartesian_to_polar`) + adjoint
// ---------------------------------------------------------------------------------------------
// Output pixel (azimuth i, range bin j) samples the Cartesian image at X = sin(az_i) r_j,
// Y = cos(az_i) r_j: u = X / res, v = -Y / res, normalised by (W-1)/2 resp. (H-1)/2, bilinear,
// zeros outside, align_corners.  r_j comes from the caller (`torch.linspace`, form_polar_range_grid
// :421-438).  One block per (range segment, azimuth, scan); sin/cos once per block.
struct CartPolarArgs {
  const void* cart;              // (B,H,W)
  const void* azimuths;          // (B,A)
  const void* range_tab;         // (R)
  void* out;                     // (B,A,R)
  const void* grad_out; void* grad_cart;
  int B, A, R, H, W;
  double cart_res;
};

template <typename T, bool BWD>
__global__ void __launch_bounds__(ROW_THREADS) cart_to_polar_kernel(CartPolarArgs a) {
  const int i = blockIdx.y, b = blockIdx.z;
  __shared__ T s_sc[2];
  if (threadIdx.x == 0) {
    const T az = static_cast<const T*>(a.azimuths)[(size_t)b * a.A + i];
    s_sc[0] = (T)sin((double)az); s_sc[1] = (T)cos((double)az);
    if (sizeof(T) == 4) { s_sc[0] = (T)sinf((float)az); s_sc[1] = (T)cosf((float)az); }
  }
  __syncthreads();
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= a.R) return;
  const T r = static_cast<const T*>(a.range_tab)[j];
  const T X = mul_rn(s_sc[0], r), Y = mul_rn(s_sc[1], r);                       // :347-348
  const T res = (T)a.cart_res;
  T u = div_rn(X, res), v = div_rn(-Y, res);                                   // :354-355
  u = mul_rn(div_rn(u, (T)(a.W - 1)), (T)2);                                   // :358-359
  v = mul_rn(div_rn(v, (T)(a.H - 1)), (T)2);
  const Taps<T> t = bilinear_taps<T>(u, v, a.W, a.H);
  const size_t o = ((size_t)b * a.A + i) * a.R + j;
  if (BWD) {
    const T go = static_cast<const T*>(a.grad_out)[o];
    T* gc = static_cast<T*>(a.grad_cart) + (size_t)b * a.H * a.W;
    if (inside(t.x0, t.y0, a.W, a.H)) atomicAdd(gc + (size_t)t.y0 * a.W + t.x0, mul_rn(go, t.nw));
    if (inside(t.x0 + 1, t.y0, a.W, a.H)) atomicAdd(gc + (size_t)t.y0 * a.W + t.x0 + 1, mul_rn(go, t.ne));
    if (inside(t.x0, t.y0 + 1, a.W, a.H)) atomicAdd(gc + (size_t)(t.y0 + 1) * a.W + t.x0, mul_rn(go, t.sw));
    if (inside(t.x0 + 1, t.y0 + 1, a.W, a.H)) atomicAdd(gc + (size_t)(t.y0 + 1) * a.W + t.x0 + 1, mul_rn(go, t.se));
    return;
  }
  const T* img = static_cast<const T*>(a.cart) + (size_t)b * a.H * a.W;
  T acc = (T)0;
  if (inside(t.x0, t.y0, a.W, a.H)) acc = mul_rn(img[(size_t)t.y0 * a.W + t.x0], t.nw);
  if (inside(t.x0 + 1, t.y0, a.W, a.H)) acc = add_rn(acc, mul_rn(img[(size_t)t.y0 * a.W + t.x0 + 1], t.ne));
  if (inside(t.x0, t.y0 + 1, a.W, a.H)) acc = add_rn(acc, mul_rn(img[(size_t)(t.y0 + 1) * a.W + t.x0], t.sw));
  if (inside(t.x0 + 1, t.y0 + 1, a.W, a.H)) acc = add_rn(acc, mul_rn(img[(size_t)(t.y0 + 1) * a.W + t.x0 + 1], t.se));
  static_cast<T*>(a.out)[o] = acc;
}

}  // namespace mmb

using namespace mmb;

extern "C" {

int mmb_sample_weights_fwd(int32_t N, int32_t n, int32_t pt_cols, int32_t H, int32_t W, int32_t dtype,
                           double cart_res, const void* mask, const void* pts, void* w_out, void* stream) {
  if (N <= 0 || n <= 0 || H <= 1 || W <= 1 || pt_cols < 2) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!mask || !pts || !w_out) return MMB_ERR_NULL_POINTER;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = (int)(((size_t)N * n + 255) / 256);
  if (dtype == MMB_F32)
    sample_weights_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(mask), static_cast<const float*>(pts),
                                                         pt_cols, N, n, H, W, (float)cart_res, static_cast<float*>(w_out));
  else
    sample_weights_kernel<double><<<blocks, 256, 0, st>>>(static_cast<const double*>(mask),
                                                          static_cast<const double*>(pts), pt_cols, N, n, H, W, cart_res,
                                                          static_cast<double*>(w_out));
  return check_launch();
}

int mmb_sample_weights_bwd(int32_t N, int32_t n, int32_t pt_cols, int32_t H, int32_t W, int32_t dtype,
                           double cart_res, const void* grad_w, const void* pts, void* grad_mask, void* stream) {
  if (N <= 0 || n <= 0 || H <= 1 || W <= 1 || pt_cols < 2) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!grad_w || !pts || !grad_mask) return MMB_ERR_NULL_POINTER;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int blocks = (int)(((size_t)N * n + 255) / 256);
  if (dtype == MMB_F32)
    sample_weights_bwd_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(grad_w),
                                                             static_cast<const float*>(pts), pt_cols, N, n, H, W,
                                                             (float)cart_res, static_cast<float*>(grad_mask));
  else
    sample_weights_bwd_kernel<double><<<blocks, 256, 0, st>>>(static_cast<const double*>(grad_w),
                                                              static_cast<const double*>(pts), pt_cols, N, n, H, W,
                                                              cart_res, static_cast<double*>(grad_mask));
  return check_launch();
}

size_t mmb_sample_weights_bwd_det_workspace_bytes(int32_t N, int32_t n, int32_t H, int32_t W, int32_t dtype) {
  if (N <= 0 || n <= 0 || H <= 1 || W <= 1) return 0;
  return dtype == MMB_F64 ? carve_sample_bwd<double>(nullptr, N, n, H, W).bytes : carve_sample_bwd<float>(nullptr, N, n, H, W).bytes;
}

int mmb_sample_weights_bwd_det(int32_t N, int32_t n, int32_t pt_cols, int32_t H, int32_t W, int32_t dtype, double cart_res,
                               const void* grad_w, const void* pts, void* grad_mask, void* workspace,
                               size_t workspace_bytes, void* stream) {
  if (N <= 0 || n <= 0 || H <= 1 || W <= 1 || pt_cols < 2) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!grad_w || !pts || !grad_mask || !workspace) return MMB_ERR_NULL_POINTER;
  if (reinterpret_cast<uintptr_t>(workspace) % 256) return MMB_ERR_MISALIGNED;
  if (workspace_bytes < mmb_sample_weights_bwd_det_workspace_bytes(N, n, H, W, dtype)) return MMB_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return dtype == MMB_F32 ? run_sample_bwd_det<float>(N, n, pt_cols, H, W, cart_res, grad_w, pts, grad_mask, workspace, st)
                          : run_sample_bwd_det<double>(N, n, pt_cols, H, W, cart_res, grad_w, pts, grad_mask, workspace, st);
}

int mmb_polar_to_cart(int32_t B, int32_t A, int32_t R, int32_t width, int32_t dtype, int32_t scan_u8,
                      const void* scan, const void* azimuths, const void* range_tab, const void* angle_tab,
                      double radar_res, int32_t interpolate_crossover, int32_t fix_wobble, void* out, void* stream) {
  if (B <= 0 || A <= 1 || R <= 1 || width <= 0) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (scan_u8 && dtype != MMB_F32) return MMB_ERR_UNSUPPORTED;
  if (!scan || !azimuths || !range_tab || !angle_tab || !out) return MMB_ERR_NULL_POINTER;
  PolarArgs a;
  a.scan = scan; a.scan_kind = scan_kind_of(dtype, scan_u8);
  a.scan_row_stride = scan_u8 ? (size_t)R + 11 : (size_t)R; a.scan_batch_stride = 0;
  a.azimuths = azimuths; a.range_tab = range_tab; a.angle_tab = angle_tab; a.out = out;
  a.grad_out = nullptr; a.grad_scan = nullptr;
  a.B = B; a.A = A; a.R = R; a.width = width;
  a.interpolate_crossover = interpolate_crossover ? 1 : 0; a.fix_wobble = fix_wobble ? 1 : 0;
  a.radar_res = radar_res;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int tiles = (width + 15) / 16;
  dim3 grid(tiles * tiles, (B + POLAR_BGROUP - 1) / POLAR_BGROUP);
  const int nbmax = B < POLAR_BGROUP ? B : POLAR_BGROUP;
  int rc;
  if (dtype == MMB_F32) {
    const size_t smem = (size_t)nbmax * A * sizeof(float);
    if ((rc = opt_in_smem(polar_to_cart_fwd_kernel<float>, smem))) return rc;
    polar_to_cart_fwd_kernel<float><<<grid, ROW_THREADS, smem, st>>>(a);
  } else {
    const size_t smem = (size_t)nbmax * A * sizeof(double);
    if ((rc = opt_in_smem(polar_to_cart_fwd_kernel<double>, smem))) return rc;
    polar_to_cart_fwd_kernel<double><<<grid, ROW_THREADS, smem, st>>>(a);
  }
  return check_launch();
}

int mmb_cfar_columns(const mmb_cfar_params* p, int32_t* w2, int32_t* mincol, int32_t* maxcol) {
  CfarArgs c;
  const int rc = fill_cfar(p, nullptr, c);
  if (rc) return rc;
  if (w2) *w2 = c.w2;
  if (mincol) *mincol = c.mincol;
  if (maxcol) *maxcol = c.maxcol;
  return MMB_OK;
}

int mmb_cfar_mask(const mmb_cfar_params* p, const void* scan, void* mask, void* stream) {
  CfarArgs c;
  int rc = fill_cfar(p, scan, c);
  if (rc) return rc;
  if (!scan || !mask) return MMB_ERR_NULL_POINTER;
  c.mask = mask;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t smem = prefix_bytes(c.hi - c.lo + 1);
#define MMB_CFAR_LAUNCH(T, D)                                                  \
  do {                                                                         \
    if ((rc = opt_in_smem(cfar_mask_kernel<T, D>, smem))) return rc;           \
    cfar_mask_kernel<T, D><<<c.rows, ROW_THREADS, smem, st>>>(c);              \
  } while (0)
#define MMB_CFAR_HARD(T)                                                       \
  do {                                                                         \
    const size_t smem_h = prefix_bytes(c.hi - c.lo + 1 + c.w2 + c.guard + 2);   \
    if ((rc = opt_in_smem(cfar_mask_hard_kernel<T>, smem_h))) return rc;       \
    cfar_mask_hard_kernel<T><<<c.rows, ROW_THREADS, smem_h, st>>>(c);          \
  } while (0)
  if (p->dtype == MMB_F32) { if (c.diff) MMB_CFAR_LAUNCH(float, true); else MMB_CFAR_HARD(float); }
  else { if (c.diff) MMB_CFAR_LAUNCH(double, true); else MMB_CFAR_HARD(double); }
#undef MMB_CFAR_HARD
#undef MMB_CFAR_LAUNCH
  return check_launch();
}

size_t mmb_extract_workspace_bytes(const mmb_extract_params* p) {
  PeakArgs a;
  if (fill_peak(p, nullptr, a, false)) return 0;
  return carve_extract(nullptr, p->B, p->A, a.rowcap, (size_t)(p->dtype == MMB_F64 ? 8 : 4)).bytes;
}

int mmb_extract_count(const mmb_extract_params* p, const void* mask_or_scan, void* workspace, size_t workspace_bytes,
                      int32_t* totals, void* stream) {
  PeakArgs a;
  int rc = fill_peak(p, mask_or_scan, a);
  if (rc) return rc;
  if (!workspace || !totals) return MMB_ERR_NULL_POINTER;
  if (reinterpret_cast<uintptr_t>(workspace) % 256) return MMB_ERR_MISALIGNED;
  ExtractWs w = carve_extract(workspace, p->B, p->A, a.rowcap, (size_t)(p->dtype == MMB_F64 ? 8 : 4));
  if (workspace_bytes < w.bytes) return MMB_ERR_WORKSPACE_TOO_SMALL;
  a.counts = w.counts; a.stage = w.stage;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p->dtype == MMB_F32) rc = p->fused ? launch_peak_rows<float, true>(a, st) : launch_peak_rows<float, false>(a, st);
  else rc = p->fused ? launch_peak_rows<double, true>(a, st) : launch_peak_rows<double, false>(a, st);
  if (rc) return rc;
  row_offsets_kernel<<<p->B, 256, 0, st>>>(w.counts, p->A, w.offsets, totals);
  return check_launch();
}

int mmb_extract_write(const mmb_extract_params* p, const void* azimuths, const void* T_ab,
                      void* workspace, size_t workspace_bytes, int32_t entry_cap, void* ent_val, int32_t* ent_row,
                      int32_t out_cap, void* points_out, void* stream) {
  PeakArgs a;
  int rc = fill_peak(p, nullptr, a, false);
  if (rc) return rc;
  if (!workspace || !azimuths || !ent_val || !ent_row || !points_out) return MMB_ERR_NULL_POINTER;
  if (entry_cap <= 0 || out_cap <= 0) return MMB_ERR_INVALID_SHAPE;
  if (reinterpret_cast<uintptr_t>(workspace) % 256) return MMB_ERR_MISALIGNED;
  ExtractWs w = carve_extract(workspace, p->B, p->A, a.rowcap, (size_t)(p->dtype == MMB_F64 ? 8 : 4));
  if (workspace_bytes < w.bytes) return MMB_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid((out_cap + 255) / 256, p->B);
  if (p->dtype == MMB_F32)
    peak_points_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(w.stage), a.rowcap, w.offsets,
                                                    static_cast<const float*>(azimuths), static_cast<const float*>(T_ab),
                                                    p->A, entry_cap, out_cap, static_cast<float*>(ent_val), ent_row,
                                                    static_cast<float*>(points_out));
  else
    peak_points_kernel<double><<<grid, 256, 0, st>>>(static_cast<const double*>(w.stage), a.rowcap, w.offsets,
                                                     static_cast<const double*>(azimuths),
                                                     static_cast<const double*>(T_ab), p->A, entry_cap, out_cap,
                                                     static_cast<double*>(ent_val), ent_row,
                                                     static_cast<double*>(points_out));
  return check_launch();
}

// ---- K-strongest ---------------------------------------------------------------------------------
struct KStrongWs { int* sel_bin; int* counts; int* offsets; size_t bytes; };
static KStrongWs carve_kstrong(void* base, int B, int A, int k) {
  Carver c(base);
  KStrongWs w;
  w.sel_bin = c.take<int>((size_t)B * A * k);
  w.counts = c.take<int>((size_t)B * A);
  w.offsets = c.take<int>((size_t)B * (A + 1));
  w.bytes = c.off;
  return w;
}

static int validate_kstrong(const mmb_kstrong_params* p) {
  if (!p) return MMB_ERR_NULL_POINTER;
  if (p->B <= 0 || p->A <= 0 || p->R <= 0 || p->k <= 0 || p->k > 1024) return MMB_ERR_INVALID_SHAPE;
  if (p->dtype != MMB_F32 && p->dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!(p->res > 0.0)) return MMB_ERR_INVALID_SHAPE;
  return MMB_OK;
}

size_t mmb_kstrong_workspace_bytes(const mmb_kstrong_params* p) {
  if (validate_kstrong(p)) return 0;
  return carve_kstrong(nullptr, p->B, p->A, p->k).bytes;
}

int mmb_kstrong_select(const mmb_kstrong_params* p, const void* scan, void* workspace, size_t workspace_bytes,
                       int32_t* totals, void* stream) {
  int rc = validate_kstrong(p);
  if (rc) return rc;
  if (!scan || !workspace || !totals) return MMB_ERR_NULL_POINTER;
  if (reinterpret_cast<uintptr_t>(workspace) % 256) return MMB_ERR_MISALIGNED;
  KStrongWs w = carve_kstrong(workspace, p->B, p->A, p->k);
  if (workspace_bytes < w.bytes) return MMB_ERR_WORKSPACE_TOO_SMALL;
  // detector.inl:45-51: `res` is a float there, the column bounds are doubles
  const double resf = (double)(float)p->res;
  double mincol = p->minr / resf;
  if (mincol > p->R || mincol < 0) mincol = 0;
  double maxcol = p->maxr / resf;
  if (maxcol > p->R || maxcol < 0) maxcol = p->R;
  KStrongArgs a;
  a.scan = scan;
  a.scan_kind = p->scan_u8 ? IN_U8 : IN_F32;
  a.scan_row_stride = p->scan_u8 ? (size_t)p->R + 11 : (size_t)p->R;
  a.rows = p->B * p->A; a.A = p->A; a.R = p->R; a.k = p->k;
  a.j0 = (int)mincol;
  int j1 = (int)maxcol;
  if ((double)j1 < maxcol) j1 += 1;                  // `j < maxcol` with a double bound
  if (j1 > p->R) j1 = p->R;
  a.j1 = j1 < a.j0 ? a.j0 : j1;
  a.N = maxcol - mincol; a.t2 = p->threshold2; a.t3 = p->threshold3;
  a.sel_bin = w.sel_bin; a.counts = w.counts;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  kstrong_rows_kernel<<<(a.rows + KSTRONG_WARPS - 1) / KSTRONG_WARPS, KSTRONG_WARPS * 32, 0, st>>>(a);
  if ((rc = check_launch())) return rc;
  row_offsets_kernel<<<p->B, 256, 0, st>>>(w.counts, p->A, w.offsets, totals);
  return check_launch();
}

int mmb_kstrong_write(const mmb_kstrong_params* p, const void* azimuths, void* workspace, size_t workspace_bytes,
                      int32_t out_cap, void* points_out, int32_t* rows_out, int32_t* bins_out, void* stream) {
  int rc = validate_kstrong(p);
  if (rc) return rc;
  if (!azimuths || !workspace || !points_out) return MMB_ERR_NULL_POINTER;
  if (out_cap <= 0) return MMB_ERR_INVALID_SHAPE;
  KStrongWs w = carve_kstrong(workspace, p->B, p->A, p->k);
  if (workspace_bytes < w.bytes) return MMB_ERR_WORKSPACE_TOO_SMALL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  dim3 grid((p->A * p->k + 255) / 256, p->B);
  if (p->dtype == MMB_F32)
    kstrong_points_kernel<float><<<grid, 256, 0, st>>>(w.sel_bin, w.counts, w.offsets, static_cast<const float*>(azimuths),
                                                       p->A, p->k, (float)p->res, p->range_offset, out_cap,
                                                       static_cast<float*>(points_out), rows_out, bins_out);
  else
    kstrong_points_kernel<double><<<grid, 256, 0, st>>>(w.sel_bin, w.counts, w.offsets,
                                                        static_cast<const double*>(azimuths), p->A, p->k, (float)p->res,
                                                        p->range_offset, out_cap, static_cast<double*>(points_out),
                                                        rows_out, bins_out);
  return check_launch();
}

int mmb_polar_to_cart_bwd(int32_t B, int32_t A, int32_t R, int32_t width, int32_t dtype, const void* grad_out,
                          const void* azimuths, const void* range_tab, const void* angle_tab, double radar_res,
                          int32_t interpolate_crossover, int32_t fix_wobble, void* grad_scan, void* stream) {
  if (B <= 0 || A <= 1 || R <= 1 || width <= 0) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!grad_out || !azimuths || !range_tab || !angle_tab || !grad_scan) return MMB_ERR_NULL_POINTER;
  PolarArgs a;
  a.scan = nullptr; a.scan_kind = IN_F32; a.scan_row_stride = (size_t)R; a.scan_batch_stride = 0;
  a.azimuths = azimuths; a.range_tab = range_tab; a.angle_tab = angle_tab; a.out = nullptr;
  a.grad_out = grad_out; a.grad_scan = grad_scan;
  a.B = B; a.A = A; a.R = R; a.width = width;
  a.interpolate_crossover = interpolate_crossover ? 1 : 0; a.fix_wobble = fix_wobble ? 1 : 0;
  a.radar_res = radar_res;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int tiles = (width + 15) / 16;
  dim3 grid(tiles * tiles, B);
  int rc;
  if (dtype == MMB_F32) {
    const size_t smem = (size_t)A * sizeof(float);
    if ((rc = opt_in_smem(polar_to_cart_kernel<float, true>, smem))) return rc;
    polar_to_cart_kernel<float, true><<<grid, ROW_THREADS, smem, st>>>(a);
  } else {
    const size_t smem = (size_t)A * sizeof(double);
    if ((rc = opt_in_smem(polar_to_cart_kernel<double, true>, smem))) return rc;
    polar_to_cart_kernel<double, true><<<grid, ROW_THREADS, smem, st>>>(a);
  }
  return check_launch();
}

int mmb_bev_from_points(int32_t N, int32_t n, int32_t pt_cols, int32_t width, int32_t dtype, double cart_res,
                        double center_offset, const void* pts, void* bev, void* stream) {
  if (N <= 0 || n <= 0 || pt_cols < 2 || width <= 0) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!pts || !bev) return MMB_ERR_NULL_POINTER;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t esz = dtype == MMB_F32 ? 4 : 8;
  if (cudaMemsetAsync(bev, 0, (size_t)N * width * width * esz, st) != cudaSuccess) return check_launch();
  const size_t total = (size_t)N * n;
  const int blocks = (int)((total + 255) / 256);
  if (dtype == MMB_F32)
    bev_from_points_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(pts), pt_cols, N, n, width,
                                                          (float)cart_res, (float)center_offset, static_cast<float*>(bev));
  else
    bev_from_points_kernel<double><<<blocks, 256, 0, st>>>(static_cast<const double*>(pts), pt_cols, N, n, width,
                                                           cart_res, center_offset, static_cast<double*>(bev));
  return check_launch();
}

static int cart_polar_launch(int bwd, int32_t B, int32_t A, int32_t R, int32_t H, int32_t W, int32_t dtype,
                             double cart_res, const void* cart, const void* azimuths, const void* range_tab, void* out,
                             const void* grad_out, void* grad_cart, void* stream) {
  if (B <= 0 || A <= 0 || R <= 0 || H <= 1 || W <= 1) return MMB_ERR_INVALID_SHAPE;
  if (dtype != MMB_F32 && dtype != MMB_F64) return MMB_ERR_UNSUPPORTED;
  if (!azimuths || !range_tab) return MMB_ERR_NULL_POINTER;
  if (bwd ? (!grad_out || !grad_cart) : (!cart || !out)) return MMB_ERR_NULL_POINTER;
  CartPolarArgs a;
  a.cart = cart; a.azimuths = azimuths; a.range_tab = range_tab; a.out = out;
  a.grad_out = grad_out; a.grad_cart = grad_cart;
  a.B = B; a.A = A; a.R = R; a.H = H; a.W = W; a.cart_res = cart_res;
  dim3 grid((R + ROW_THREADS - 1) / ROW_THREADS, A, B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == MMB_F32) {
    if (bwd) cart_to_polar_kernel<float, true><<<grid, ROW_THREADS, 0, st>>>(a);
    else cart_to_polar_kernel<float, false><<<grid, ROW_THREADS, 0, st>>>(a);
  } else {
    if (bwd) cart_to_polar_kernel<double, true><<<grid, ROW_THREADS, 0, st>>>(a);
    else cart_to_polar_kernel<double, false><<<grid, ROW_THREADS, 0, st>>>(a);
  }
  return check_launch();
}

int mmb_cart_to_polar(int32_t B, int32_t A, int32_t R, int32_t H, int32_t W, int32_t dtype, double cart_res,
                      const void* cart, const void* azimuths, const void* range_tab, void* out, void* stream) {
  return cart_polar_launch(0, B, A, R, H, W, dtype, cart_res, cart, azimuths, range_tab, out, nullptr, nullptr, stream);
}

int mmb_cart_to_polar_bwd(int32_t B, int32_t A, int32_t R, int32_t H, int32_t W, int32_t dtype, double cart_res,
                          const void* grad_out, const void* azimuths, const void* range_tab, void* grad_cart,
                          void* stream) {
  return cart_polar_launch(1, B, A, R, H, W, dtype, cart_res, nullptr, azimuths, range_tab, nullptr, grad_out, grad_cart,
                           stream);
}

}  // extern "C"

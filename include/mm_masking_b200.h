/*
 * mm_masking_b200 -- C ABI of the B200 (sm_100a) hot path of utiasASRL/mm_masking.
 *
 * Every entry point takes plain device pointers, sizes and a CUDA stream handle
 * (`void*` == cudaStream_t); no C++ or torch types cross this boundary.  The caller
 * (PyTorch's allocator, through mm_masking_b200/_capi.py) owns every buffer; the library
 * allocates nothing and keeps no state between calls.  All functions enqueue work on
 * `stream` and return without synchronising.  Return value: 0 on success, otherwise an
 * mmb_status code (see mmb_status_string).
 *
 * Each function names the reference interface it replaces (paths relative to the
 * reference checkout).
 */
#ifndef MM_MASKING_B200_H
#define MM_MASKING_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MMB_ABI_VERSION 5

typedef enum mmb_status {
  MMB_OK = 0,
  MMB_ERR_NULL_POINTER = 1,
  MMB_ERR_INVALID_SHAPE = 2,
  MMB_ERR_UNSUPPORTED = 3,
  MMB_ERR_WORKSPACE_TOO_SMALL = 4,
  MMB_ERR_MISALIGNED = 5,
  MMB_ERR_CUDA = 6,
  MMB_ERR_SCRATCH_LIMIT = 7   /* mmb_nn_search_static: the problem needs more than MMB_NN_STATIC_SCRATCH_BYTES */
} mmb_status;

enum { MMB_F32 = 0, MMB_F64 = 1 };
enum { MMB_PT2PL = 0, MMB_PT2PT = 1 };
enum { MMB_LOSS_NONE = 0, MMB_LOSS_HUBER = 1, MMB_LOSS_CAUCHY = 2 };
/* Nearest-neighbour backend.  Both return the same indices (exact squared differences, lowest
 * index on ties); BRUTE evaluates all n*m pairs like nn.py:32, GRID visits only the cells that can
 * hold the neighbour (Morton-ordered cells, 3 x 3 blocks of a quadtree level, separate step kernel), ROWS
 * covers the disc of the current bound row by row on a row-major cell grid and evaluates the residual rows,
 * the normal equations and the solve in the same launch (one launch per iteration).  AUTO: ROWS when
 * m >= 8192, BRUTE below. */
enum { MMB_NN_AUTO = 0, MMB_NN_BRUTE = 1, MMB_NN_GRID = 2, MMB_NN_ROWS = 3 };
#define MMB_PROFILE_CLASSES 8

int mmb_abi_version(void);
const char* mmb_status_string(int status);
/* cudaGetLastError text of the most recent MMB_ERR_CUDA on this thread ("" if none). */
const char* mmb_last_cuda_error(void);
/* Number of kernels this library has launched in the calling process so far (every launch site counts once).
 * Measurement aid: bench.py differences it around its timed regions for `gpu_launches`. */
long long mmb_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Weighted differentiable ICP.
 * Replaces external/dICP/dICP/ICP.py:45-293 (`ICP.icp` / `ICP.dICP`) on dense inputs; the
 * ragged-list normalisation of ICP.py:295-497 stays on the host (mm_masking_b200/dICP/ICP.py).
 * ---------------------------------------------------------------------------------------- */
typedef struct mmb_icp_params {
  int32_t N;               /* pairs in the batch */
  int32_t n;               /* source rows per pair (after padding) */
  int32_t m;               /* target rows per pair (after padding) */
  int32_t target_cols;     /* 3 (xyz) or 6 (xyz + normal); pt2pl needs 6 (ICP.py:101) */
  int32_t dtype;           /* MMB_F32 / MMB_F64: source, target, T_init, weight and all outputs */
  int32_t icp_type;        /* MMB_PT2PL / MMB_PT2PT (ICP.py:28) */
  int32_t dim;             /* 3, or 2 = optimise yaw,x,y only (ICP.py:105-114,176-179) */
  int32_t max_iterations;  /* ICP.py:29 */
  int32_t differentiable;  /* soft tanh trim / pseudo-Huber vs hard variants (loss.py:21-56) */
  int32_t const_iter;      /* config const_iter (dICP_config.yaml:7) */
  int32_t use_trim;        /* trim_dist is not None and >= 0 (ICP.py:152) */
  int32_t loss_kind;       /* MMB_LOSS_* (ICP.py:156-159) */
  int32_t has_weight;      /* prior per-point weights given (ICP.py:310-316) */
  int32_t save_for_backward;
  int32_t nn_slices;       /* 0 = choose automatically; else target slices per query tile */
  int32_t nn_backend;      /* MMB_NN_* */
  double tolerance;        /* ICP.py:30 */
  double trim_dist;
  double loss_metric;
  double tanh_steepness;   /* dICP_config.yaml:3 */
  double match_ratio_thresh; /* dICP_config.yaml:19 */
} mmb_icp_params;

size_t mmb_icp_workspace_bytes(const mmb_icp_params* p);
/* bytes of the opaque `saved` buffer mmb_icp_backward needs (0 if !save_for_backward) */
size_t mmb_icp_saved_bytes(const mmb_icp_params* p);

/*
 * Forward.  R = n (pt2pl) or 3n (pt2pt).  Histories are laid out for max_iterations;
 * only the first *iters_done entries of each are meaningful (ICP.py:273-275).
 *   source (N,n,3)  target (N,m,target_cols)  T_init (N,4,4)  weight (N,n) or NULL
 *   T_out (N,4,4)  pc_out (N,n,3)  deltas (N,max_it,6)  weights_hist (N,max_it,R)
 *   costs (N,max_it)  converged (N) u8  iterations (N)  matched_ratio (N)  -- dtype as inputs
 *   iters_done (1) int32   saved: mmb_icp_saved_bytes or NULL
 */
int mmb_icp_forward(const mmb_icp_params* p, const void* source, const void* target,
                    const void* T_init, const void* weight, void* T_out, void* pc_out,
                    void* deltas, void* weights_hist, void* costs, uint8_t* converged,
                    void* iterations, void* matched_ratio, int32_t* iters_done, void* saved,
                    void* workspace, size_t workspace_bytes, void* stream);

/*
 * mmb_icp_forward plus per-kernel-class device timing, for bench.py's roofline figures.  CUDA
 * events are recorded on `stream` around every launch; the call synchronises on the last one
 * and writes to HOST memory, MMB_PROFILE_CLASSES entries each: host_ms[c] = summed durations (ms)
 * of class c = 0 nearest-neighbour launches, 1 step (residual/normal-equation/solve) launches,
 * 2 the rest (init, target packing, result assembly), 3 the cold-start bound pass (subsample pack +
 * search, brute-force backend), 4 the cell-grid build (grid backend); host_launches[c] = number
 * of timed intervals in class c.
 */
int mmb_icp_forward_profiled(const mmb_icp_params* p, const void* source, const void* target,
                             const void* T_init, const void* weight, void* T_out, void* pc_out,
                             void* deltas, void* weights_hist, void* costs, uint8_t* converged,
                             void* iterations, void* matched_ratio, int32_t* iters_done,
                             void* saved, void* workspace, size_t workspace_bytes, void* stream,
                             float* host_ms, int32_t* host_launches);

/*
 * Backward of mmb_icp_forward: the hand-derived adjoint of the unrolled iterations (what
 * autograd computes for ICP.py:130-250; correspondence indices are constants).
 *   grad_T (N,4,4) or NULL, grad_pc (N,n,3) or NULL: incoming gradients.
 *   grad_source (N,n,3), grad_T_init (N,4,4), grad_weight (N,n): written; any may be NULL.
 *   grad_target (N,m,target_cols): ACCUMULATED into with atomics, caller zero-fills; may be NULL.
 */
int mmb_icp_backward(const mmb_icp_params* p, const void* source, const void* target,
                     const void* weight, const void* saved, const void* grad_T,
                     const void* grad_pc, void* grad_source, void* grad_target,
                     void* grad_T_init, void* grad_weight, void* workspace,
                     size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------
 * Exact 1-NN (brute force or cell grid, MMB_NN_*).  Replaces external/dICP/dICP/nn.py:23-40 / :72-92
 * (`torch.cdist` -> `argmin` -> `gather`): index of the nearest target row for every query
 * row, squared differences evaluated exactly, lowest index on ties.  dim == 2 ignores z.
 *   query (N,n,3)  target (N,m,target_cols)  idx_out (N,n) int32
 * ---------------------------------------------------------------------------------------- */
size_t mmb_nn_workspace_bytes(int32_t N, int32_t n, int32_t m, int32_t dtype, int32_t nn_slices,
                              int32_t nn_backend);
int mmb_nn_search(int32_t N, int32_t n, int32_t m, int32_t target_cols, int32_t dtype,
                  int32_t dim, int32_t nn_slices, int32_t nn_backend, const void* query,
                  const void* target, int32_t* idx_out, void* workspace, size_t workspace_bytes,
                  void* stream);

/* The same search for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Arguments, results and error codes are those of mmb_nn_search without the workspace: the scratch memory is a
 * file-scope `static __device__` array of MMB_NN_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the
 * library holds that much until the process exits).  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine) and launches the same kernels, so idx_out is bit-identical to mmb_nn_search's.
 *   - a problem whose mmb_nn_workspace_bytes(...) exceeds MMB_NN_STATIC_SCRATCH_BYTES is rejected with
 *     MMB_ERR_SCRATCH_LIMIT before anything is launched (16 MiB holds e.g. 4 x (2 000 queries, 50 000 targets) fp32);
 *   - every call on a device shares the one array and finds the previous call's contents in it; the call clears
 *     what must start at zero itself, on `stream`.  Calls that overlap on the GPU (two streams, two host threads,
 *     two graphs replayed concurrently) RACE on the array: serialise them, or use mmb_nn_search with one workspace
 *     per stream. */
#define MMB_NN_STATIC_SCRATCH_BYTES (16u << 20)
int mmb_nn_search_static(int32_t N, int32_t n, int32_t m, int32_t target_cols, int32_t dtype,
                         int32_t dim, int32_t nn_slices, int32_t nn_backend, const void* query,
                         const void* target, int32_t* idx_out, void* stream);


/* ------------------------------------------------------------------------------------------
 * Radar preprocessing (mm_masking/radar_utils.py).  dtype = MMB_F32 / MMB_F64 is the working type
 * of every floating-point buffer of a call.  `scan_u8` != 0: the scan is raw Navtech rows,
 * (B,A,R+11) uint8 with 11 header bytes per row and value = byte / 255 (radar_utils.py:20-27,
 * `load_radar`), MMB_F32 only; otherwise (B,A,R) in the working type.
 * ---------------------------------------------------------------------------------------- */

/* Per-point bilinear sample of the weight mask.  Replaces radar_utils.py:108-126
 * (`extract_weights` = `point_to_cart_idx(min_to_plus_1=True)` :374-391 + `F.grid_sample(bilinear,
 * zeros, align_corners=True)`); points with x == 0 and y == 0 are batch padding and get weight 0.
 *   mask (N,H,W)   pts (N,n,pt_cols) [x,y in columns 0,1]   w_out (N,n) */
int mmb_sample_weights_fwd(int32_t N, int32_t n, int32_t pt_cols, int32_t H, int32_t W,
                           int32_t dtype, double cart_res, const void* mask, const void* pts,
                           void* w_out, void* stream);
/* Adjoint of the above with respect to the mask (what autograd does for radar_utils.py:126).
 *   grad_w (N,n)   grad_mask (N,H,W): ACCUMULATED into with atomics, caller zero-fills. */
int mmb_sample_weights_bwd(int32_t N, int32_t n, int32_t pt_cols, int32_t H, int32_t W,
                           int32_t dtype, double cart_res, const void* grad_w, const void* pts,
                           void* grad_mask, void* stream);
/* The same adjoint without floating-point atomics: taps sorted by pixel with a stable radix sort,
 * every run of equal pixels summed in tap order -- bit-identical from run to run.  grad_mask zero-filled by the
 * caller; workspace: mmb_sample_weights_bwd_det_workspace_bytes, 256-byte aligned. */
size_t mmb_sample_weights_bwd_det_workspace_bytes(int32_t N, int32_t n, int32_t H, int32_t W, int32_t dtype);
int mmb_sample_weights_bwd_det(int32_t N, int32_t n, int32_t pt_cols, int32_t H, int32_t W,
                               int32_t dtype, double cart_res, const void* grad_w, const void* pts,
                               void* grad_mask, void* workspace, size_t workspace_bytes, void* stream);

/* Polar -> Cartesian resampling.  Replaces radar_utils.py:258-336
 * (`radar_polar_to_cartesian_diff`).  range_tab / angle_tab: the (width*width) pixel-centre tables
 * of `form_cart_range_angle_grid` (:399-419), batch-invariant, supplied by the caller.
 *   scan (B,A,R)   azimuths (B,A)   out (B,width,width) */
int mmb_polar_to_cart(int32_t B, int32_t A, int32_t R, int32_t width, int32_t dtype,
                      int32_t scan_u8, const void* scan, const void* azimuths,
                      const void* range_tab, const void* angle_tab, double radar_res,
                      int32_t interpolate_crossover, int32_t fix_wobble, void* out, void* stream);

/* Adjoint of mmb_polar_to_cart with respect to the scan (what autograd computes through the
 * `grid_sample` of radar_utils.py:334; the wrap-padded azimuth rows of :317-319 fold back onto rows
 * A-1 and 0).   grad_out (B,width,width)   grad_scan (B,A,R): ACCUMULATED into, caller zero-fills. */
int mmb_polar_to_cart_bwd(int32_t B, int32_t A, int32_t R, int32_t width, int32_t dtype,
                          const void* grad_out, const void* azimuths, const void* range_tab,
                          const void* angle_tab, double radar_res, int32_t interpolate_crossover,
                          int32_t fix_wobble, void* grad_scan, void* stream);

/* Cartesian -> polar resampling.  Replaces radar_utils.py:338-372 (`radar_cartesian_to_polar`).
 * range_tab (R): the range of every polar bin (`form_polar_range_grid` :421-438), from the caller.
 *   cart (B,H,W)   azimuths (B,A)   out (B,A,R) */
int mmb_cart_to_polar(int32_t B, int32_t A, int32_t R, int32_t H, int32_t W, int32_t dtype,
                      double cart_res, const void* cart, const void* azimuths,
                      const void* range_tab, void* out, void* stream);
/*   grad_out (B,A,R)   grad_cart (B,H,W): ACCUMULATED into with atomics, caller zero-fills. */
int mmb_cart_to_polar_bwd(int32_t B, int32_t A, int32_t R, int32_t H, int32_t W, int32_t dtype,
                          double cart_res, const void* grad_out, const void* azimuths,
                          const void* range_tab, void* grad_cart, void* stream);

/* Occupancy BEV of a point cloud.  Replaces radar_utils.py:142-165 (`extract_bev_from_pts`): the four
 * floor/ceil pixels of every point are 1, out-of-range coordinates go to the centre pixel, which
 * ends as 0.  center_offset: the pixel offset `point_to_cart_idx` adds (320 = its default width / 2,
 * which the reference uses whatever `cart_pixel_width` is, :144).
 *   pts (N,n,pt_cols)   bev (N,width,width): fully written (zero-filled here). */
int mmb_bev_from_points(int32_t N, int32_t n, int32_t pt_cols, int32_t width, int32_t dtype,
                        double cart_res, double center_offset, const void* pts, void* bev,
                        void* stream);

/* GO-CFAR mask.  Replaces radar_utils.py:29-69 (`cfar_mask`). */
typedef struct mmb_cfar_params {
  int32_t B, A, R;         /* scans, azimuths per scan, range bins */
  int32_t dtype;
  int32_t scan_u8;
  int32_t width;           /* window width (made odd like radar_utils.py:34) */
  int32_t guard;
  int32_t diff;            /* soft tanh mask + hardshrink(0.99) (:64-65) vs x > thres (:67) */
  double res, minr, maxr, a_thresh, b_thresh, steep_fact;
} mmb_cfar_params;
/* the half width and the [mincol, maxcol) column range of radar_utils.py:35-39 (host only) */
int mmb_cfar_columns(const mmb_cfar_params* p, int32_t* w2, int32_t* mincol, int32_t* maxcol);
/*   scan (B,A,R)   mask (B,A,R): float32 when !diff (as `torch.where(.., 1.0, 0.0)`), else dtype */
int mmb_cfar_mask(const mmb_cfar_params* p, const void* scan, void* mask, void* stream);

/* Peak extraction.  Replaces radar_utils.py:71-106 (`extract_pc`, with
 * `mean_peaks_parallel_fast` :167-185 and `pol_2_cart` :187-195).  Two phases because the
 * number of points is data dependent: count -> (caller reads totals, allocates) -> write.  The
 * count call evaluates every azimuth row once and parks its peak entries in the workspace; the
 * write call only assembles the points from them, so the workspace must not be touched in between.
 * fused != 0: the first buffer is the SCAN and the CFAR mask of `cfar` is evaluated on the fly. */
typedef struct mmb_extract_params {
  int32_t B, A, R;
  int32_t dtype;
  int32_t diff;
  int32_t fused;
  double res, steep_fact;
  mmb_cfar_params cfar;    /* used when fused */
} mmb_extract_params;
size_t mmb_extract_workspace_bytes(const mmb_extract_params* p);
/*   totals (B) int32 device: number of non-zero peak entries per scan (= 2 x points) */
int mmb_extract_count(const mmb_extract_params* p, const void* mask_or_scan, void* workspace,
                      size_t workspace_bytes, int32_t* totals, void* stream);
/*   same workspace as the count call, contents intact.  ent_val (B,entry_cap) dtype, ent_row
 *   (B,entry_cap) int32: the peak entries (value, azimuth row) in flattened order, for inspection;
 *   azimuths (B,A); T_ab (B,4,4) or NULL;
 *   points_out (B,out_cap,3): scan b holds totals[b]/2 points, azimuth-major, range ascending. */
int mmb_extract_write(const mmb_extract_params* p, const void* azimuths,
                      const void* T_ab, void* workspace, size_t workspace_bytes,
                      int32_t entry_cap, void* ent_val, int32_t* ent_row, int32_t out_cap,
                      void* points_out, void* stream);

/* K-strongest per azimuth.  Replaces
 * external/vtr3/main/src/vtr_radar/include/vtr_radar/detector/detector.inl:39-85
 * (`KStrongest<PointT>::run`) + navtech_extraction_module.cpp:33-39 (`pol2Cart2D`); order within
 * a row: descending intensity, ties to the lower bin.  Scan is float32 or raw bytes. */
typedef struct mmb_kstrong_params {
  int32_t B, A, R;
  int32_t dtype;           /* azimuths / points dtype */
  int32_t scan_u8;
  int32_t k;
  double res, minr, maxr, threshold2, threshold3, range_offset;
} mmb_kstrong_params;
size_t mmb_kstrong_workspace_bytes(const mmb_kstrong_params* p);
int mmb_kstrong_select(const mmb_kstrong_params* p, const void* scan, void* workspace,
                       size_t workspace_bytes, int32_t* totals, void* stream);
/*   points_out (B,out_cap,3); rows_out / bins_out (B,out_cap) int32 or NULL */
int mmb_kstrong_write(const mmb_kstrong_params* p, const void* azimuths, void* workspace,
                      size_t workspace_bytes, int32_t out_cap, void* points_out, int32_t* rows_out,
                      int32_t* bins_out, void* stream);

/* FP32 FMA-pipe probe used by bench.py for the roofline denominator: runs `iters` dependent
 * rounds of 8 independent FFMA (packed=0) or FFMA2 (packed=1) chains per thread on
 * blocks x 256 threads; flops = blocks*256*iters*8*2*(packed?2:1). */
int mmb_fp32_probe(int32_t blocks, int32_t iters, int32_t packed, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MM_MASKING_B200_H */

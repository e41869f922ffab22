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

#define MMB_ABI_VERSION 1

typedef enum mmb_status {
  MMB_OK = 0,
  MMB_ERR_NULL_POINTER = 1,
  MMB_ERR_INVALID_SHAPE = 2,
  MMB_ERR_UNSUPPORTED = 3,
  MMB_ERR_WORKSPACE_TOO_SMALL = 4,
  MMB_ERR_MISALIGNED = 5,
  MMB_ERR_CUDA = 6
} mmb_status;

enum { MMB_F32 = 0, MMB_F64 = 1 };
enum { MMB_PT2PL = 0, MMB_PT2PT = 1 };
enum { MMB_LOSS_NONE = 0, MMB_LOSS_HUBER = 1, MMB_LOSS_CAUCHY = 2 };

int mmb_abi_version(void);
const char* mmb_status_string(int status);
/* cudaGetLastError text of the most recent MMB_ERR_CUDA on this thread ("" if none). */
const char* mmb_last_cuda_error(void);

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
  int32_t reserved;
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
 * and writes to HOST memory: host_ms[0..3] = summed durations (ms) of the nearest-neighbour
 * launches, the step (residual/normal-equation/solve) launches, the rest (init, target
 * packing, result assembly) and the cold-start bound pass (subsample pack + search);
 * host_launches[0..3] = number of timed intervals in each class.
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
 * Brute-force 1-NN.  Replaces external/dICP/dICP/nn.py:23-40 / :72-92
 * (`torch.cdist` -> `argmin` -> `gather`): index of the nearest target row for every query
 * row, squared differences evaluated exactly, lowest index on ties.  dim == 2 ignores z.
 *   query (N,n,3)  target (N,m,target_cols)  idx_out (N,n) int32
 * ---------------------------------------------------------------------------------------- */
size_t mmb_nn_workspace_bytes(int32_t N, int32_t n, int32_t m, int32_t nn_slices);
int mmb_nn_search(int32_t N, int32_t n, int32_t m, int32_t target_cols, int32_t dtype,
                  int32_t dim, int32_t nn_slices, const void* query, const void* target,
                  int32_t* idx_out, void* workspace, size_t workspace_bytes, void* stream);

/* FP32 FMA-pipe probe used by bench.py for the roofline denominator: runs `iters` dependent
 * rounds of 8 independent FFMA (packed=0) or FFMA2 (packed=1) chains per thread on
 * blocks x 256 threads; flops = blocks*256*iters*8*2*(packed?2:1). */
int mmb_fp32_probe(int32_t blocks, int32_t iters, int32_t packed, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MM_MASKING_B200_H */

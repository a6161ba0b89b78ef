// Host-side state of one solver handle (one process, one GPU, one point shard).
#ifndef CB200_SOLVER_STATE_H_
#define CB200_SOLVER_STATE_H_

#include <cuda_runtime.h>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "cg_kernels.cuh"
#include "explicit_plan.h"
#include "flatten.h"
#include "multi_gpu.h"

namespace cb200 {

// Minimal NCCL binding resolved with dlopen at first use, so single-GPU
// processes never load (or need) libnccl.
struct UniqueId { char internal[128]; };  // ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES 128)
struct Nccl {
  void* lib = nullptr;
  void* comm = nullptr;
  int (*GetUniqueId)(UniqueId*) = nullptr;
  int (*CommInitRank)(void**, int, UniqueId, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};

}  // namespace cb200

struct cb200_solver {
  cb200_options opt{};
  // global sizes
  int P = 0, C = 0, N = 0;
  bool canonical = false;
  // shard
  int p_begin = 0, p_end = 0, r_begin = 0, r_end = 0;
  int P_loc = 0, N_loc = 0;
  int nvec = 0;           // 9C
  int vec_blocks = 0;     // CTAs of the camera-vector kernels
  int num_tiles = 0, num_long = 0, num_empty = 0, num_chunks = 0, num_dup_pairs = 0;
  int nrep = 0, rep_stride = 0;
  int num_sms = 0;        // persistent grids are sized in multiples of it
  bool has_D = false, values_resident = false, b_resident = false, isc_ready = false;
  bool sj_ready = false;
  bool fused_step_ok = false;  // cg_step_kernel fits the device as one co-resident grid

  cudaStream_t stream = nullptr;
  cudaEvent_t ev[8] = {};
  // camera-sized start-up work of a solve runs beside the preconditioner build (single GPU)
  cudaStream_t side_stream = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;

  // device: structure
  int *d_cam = nullptr, *d_row_start = nullptr, *d_long_points = nullptr, *d_empty_points = nullptr;
  cb200::TileMeta* d_tiles = nullptr;
  int *d_cam_perm = nullptr, *d_chunk_begin = nullptr, *d_cam_chunk_start = nullptr, *d_dup_pairs = nullptr;
  int *d_e_pos = nullptr, *d_f_pos = nullptr;
  // device: values
  double *d_E = nullptr, *d_F = nullptr, *d_b = nullptr, *d_De = nullptr, *d_Df = nullptr;
  double *d_values_staging = nullptr;  // non-canonical layouts only
  size_t staging_len = 0;
  // device: derived
  double *d_ete_inv = nullptr, *d_weights = nullptr, *d_chunk_partial = nullptr, *d_upper = nullptr;
  double *d_dup_blocks = nullptr, *d_sj = nullptr, *d_ftf_inv = nullptr;
  double *d_yrep = nullptr, *d_y_e = nullptr, *d_ete9 = nullptr;
  double* d_model = nullptr;  // when set, back substitution also accumulates the model cost change here
  double* d_x10 = nullptr;  // the vector being multiplied, padded to 10 doubles per camera (16-byte gathers)
  // device: camera vectors
  double *d_rhs = nullptr, *d_x = nullptr, *d_r = nullptr, *d_p = nullptr, *d_z = nullptr, *d_q = nullptr;
  double *d_t1 = nullptr, *d_t2 = nullptr, *d_rows = nullptr, *d_pts = nullptr;
  cb200::CgState* d_cg = nullptr;
  cb200::CgState* h_cg = nullptr;  // pinned; only the scalar head is copied back
  double* d_trace = nullptr;
  double* h_trace = nullptr;
  int max_trace = 0;

  // widened boundary: device-resident bundle adjustment problem (c_api_lm.inc)
  bool bal_attached = false, bal_params = false;
  double *d_obs = nullptr, *d_cameras = nullptr, *d_points = nullptr, *d_cand_cameras = nullptr,
         *d_cand_points = nullptr;
  int* d_row_point = nullptr;
  double *d_scale = nullptr;       // [3 P_loc + 9C] Jacobi scaling (points, then cameras)
  double *d_norm2 = nullptr;       // [3 P_loc + 9C] squared column norms of the unscaled Jacobian
  double *d_grad = nullptr;        // [3 P_loc + 9C] J'r
  double *d_cam_acc = nullptr;     // privatised camera accumulators of the evaluator
  double *d_cost_partial = nullptr;
  double *d_lm_scalars = nullptr;  // cost, model, |x|^2, |delta|^2, gradient max, gradient norm^2
  double *h_lm_scalars = nullptr;  // pinned
  int eval_blocks = 0;

  // explicit reduced camera matrix (use_explicit_schur_complement; explicit_kernels.cuh)
  int ex_cells = 0, ex_items = 0;
  int *d_ex_row_ptr = nullptr, *d_ex_col = nullptr, *d_ex_t_ptr = nullptr, *d_ex_t_cell = nullptr,
      *d_ex_t_row = nullptr, *d_ex_srow = nullptr, *d_ex_counter = nullptr;
  long long* d_ex_pair_start = nullptr;
  unsigned int* d_ex_pair_cell = nullptr;
  cb200::FormItem* d_ex_items = nullptr;
  double* d_S = nullptr;  // ex_cells x 81
  std::vector<int> ex_row_ptr, ex_col;  // host copy of the block structure (cb200_explicit_structure)
  bool ex_ready = false;                // d_S holds the matrix of the resident values / D

  cb200::Nccl nccl;
  // sums over the GPUs through NVLink peer memory (cg_step.cuh); peer_allreduce false = NCCL path
  cb200::PeerChannel peer[2];
  long long global_dup_pairs = 0;             // duplicate (point, camera) pairs over all ranks
  bool peer_allreduce = false;
  // multi-GPU handle (cb200_options::num_gpus > 1): `multi` on the handle the caller holds, `group` on its shards
  cb200::MultiGpu* multi = nullptr;
  cb200::InProcGroup* group = nullptr;
  bool writes_replicated = true;              // this shard writes camera-sized (replicated) results to host memory
  // back substitution from the per-iteration history (stream_kernels.cuh, StreamArgs::v_store)
  double *d_v_rhs = nullptr, *d_v_hist = nullptr;
  bool use_history = false;
  // deterministic mode: per-chunk partial sums of the camera gather
  double* d_chunk_partial9 = nullptr;
  // chunked upload of cb200_solve: tile / row ranges of the chunks, copy stream, one event per chunk
  std::vector<int> chunk_tile, chunk_row;     // h2d_chunks + 1 entries each
  cudaStream_t copy_stream = nullptr;
  std::vector<cudaEvent_t> chunk_event;
  cudaEvent_t ev_copy_start = nullptr;
  // CUDA graphs of the launch sequences of a solve, by sequence key; any change of a launch parameter
  // (diagonal on / off, trace, model pointer ...) bumps graph_config and drops them
  bool graphs_ok = false;
  int last_iterations = 0;                    // PCG iterations of the previous solve (length of the next first batch)
  long long graph_sig = 0;
  std::map<long long, cudaGraphExec_t> graphs;
  std::map<long long, int> graph_launches;    // kernels inside each graph
  std::set<long long> graph_seen;             // sequences that already ran once un-captured (warm-up)
  cb200::CgState* h_cg_init = nullptr;        // pinned: the head a solve starts from (source of the graph's H2D node)
  int stream_ctas_per_sm[4][4] = {};          // occupancy of point_stream_kernel<MODE, BUILD, DET>, filled on first use
  int launches = 0;
  std::vector<void*> allocations;
};

#endif  // CB200_SOLVER_STATE_H_

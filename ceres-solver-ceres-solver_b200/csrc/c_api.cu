// libceres_b200_schur.so — C ABI (include/ceres_b200_schur.h) over the CUDA
// kernels. Host orchestration of IterativeSchurComplementSolver::SolveImpl
// (iterative_schur_complement_solver.cc:64-157) with every vector, scalar and
// convergence test resident on the device.
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <chrono>
#include <limits>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

#include "camera_kernels.cuh"
#include "cg_kernels.cuh"
#include "cg_step.cuh"
#include "explicit_kernels.cuh"
#include "kernels.cuh"
#include "lm_kernels.cuh"
#include "solver_state.h"
#include "stream_kernels.cuh"

using namespace cb200;

namespace {

thread_local std::string g_error;

int Fail(int code, const std::string& msg) {
  g_error = msg;
  return code;
}

#define CUDA_OK(expr)                                                                  \
  do {                                                                                 \
    cudaError_t e_ = (expr);                                                           \
    if (e_ != cudaSuccess)                                                             \
      return Fail(CB200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

#define CHECK_OK(expr)        \
  do {                        \
    int rc_ = (expr);         \
    if (rc_ != CB200_OK) return rc_; \
  } while (0)

template <typename T>
int DeviceAlloc(cb200_solver* s, T** ptr, size_t count) {
  *ptr = nullptr;
  if (count == 0) count = 1;
  CUDA_OK(cudaMalloc(reinterpret_cast<void**>(ptr), count * sizeof(T)));
  s->allocations.push_back(*ptr);
  return CB200_OK;
}

template <typename T>
int DeviceUpload(cb200_solver* s, T** ptr, const std::vector<T>& host) {
  CHECK_OK(DeviceAlloc(s, ptr, host.size()));
  if (!host.empty())
    CUDA_OK(cudaMemcpy(*ptr, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice));
  return CB200_OK;
}

int LoadNccl(Nccl* n) {
  if (n->lib != nullptr) return CB200_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* name : names) {
    n->lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
    if (n->lib) break;
  }
  if (!n->lib) return Fail(CB200_ERR_NCCL, std::string("dlopen libnccl: ") + dlerror());
  n->GetUniqueId = reinterpret_cast<decltype(n->GetUniqueId)>(dlsym(n->lib, "ncclGetUniqueId"));
  n->CommInitRank = reinterpret_cast<decltype(n->CommInitRank)>(dlsym(n->lib, "ncclCommInitRank"));
  n->AllReduce = reinterpret_cast<decltype(n->AllReduce)>(dlsym(n->lib, "ncclAllReduce"));
  n->AllGather = reinterpret_cast<decltype(n->AllGather)>(dlsym(n->lib, "ncclAllGather"));
  n->CommDestroy = reinterpret_cast<decltype(n->CommDestroy)>(dlsym(n->lib, "ncclCommDestroy"));
  n->GetErrorString = reinterpret_cast<decltype(n->GetErrorString)>(dlsym(n->lib, "ncclGetErrorString"));
  if (!n->GetUniqueId || !n->CommInitRank || !n->AllReduce || !n->CommDestroy)
    return Fail(CB200_ERR_NCCL, "libnccl lacks a required symbol");
  return CB200_OK;
}

// Sum a device buffer over all ranks (camera-sized data only).
// an environment switch is on when it is set to anything but "" or "0"
bool EnvFlag(const char* name) {
  const char* e = std::getenv(name);
  return e != nullptr && e[0] != '\0' && !(e[0] == '0' && e[1] == '\0');
}

int NcclAllReduce(cb200_solver* s, double* buf, size_t count) {
  if (s->opt.world_size <= 1) return CB200_OK;
  const int rc = s->nccl.AllReduce(buf, buf, count, /*ncclDouble*/ 8, /*ncclSum*/ 0, s->nccl.comm, s->stream);
  if (rc != 0)
    return Fail(CB200_ERR_NCCL, std::string("ncclAllReduce: ") +
                                    (s->nccl.GetErrorString ? s->nccl.GetErrorString(rc) : "error"));
  return CB200_OK;
}

void FillPeerRefs(cb200_solver* s, int channel, PeerRefs* p) {
  const PeerChannel& ch = s->peer[channel];
  p->world = s->opt.world_size; p->rank = s->opt.rank;
  p->channel = channel;
  p->slot_stride = ch.slot;
  p->epoch = &s->d_cg->peer_epoch[channel];
  for (int r = 0; r < p->world; ++r) {
    p->base[r] = reinterpret_cast<double*>(ch.base[r]);
    p->flag[r] = reinterpret_cast<unsigned int*>(reinterpret_cast<char*>(ch.base[r]) + 2 * ch.slot * sizeof(double));
  }
}

// Sum (or maximum) over the ranks, in place. Peer-memory path (rank-ordered, deterministic, one small
// kernel) when the buffers are mapped and the vector fits a slot, NCCL otherwise. `channel` 1 is the
// side stream's set of buffers (the caller must be launching into the side stream).
int AllReduce(cb200_solver* s, double* buf, size_t count, int channel = 0, bool max_op = false) {
  if (s->opt.world_size <= 1) return CB200_OK;
  if (!s->peer_allreduce || count > s->peer[channel].slot) {
    if (s->nccl.comm == nullptr) return Fail(CB200_ERR_NCCL, "no NCCL communicator for a sum that does not fit the peer buffers");
    const int rc = s->nccl.AllReduce(buf, buf, count, /*ncclDouble*/ 8, max_op ? /*ncclMax*/ 2 : /*ncclSum*/ 0,
                                     s->nccl.comm, s->stream);
    if (rc != 0)
      return Fail(CB200_ERR_NCCL, std::string("ncclAllReduce: ") +
                                      (s->nccl.GetErrorString ? s->nccl.GetErrorString(rc) : "error"));
    return CB200_OK;
  }
  PeerRefs p;
  FillPeerRefs(s, channel, &p);
  const int grid = (int)std::max<size_t>(1, std::min<size_t>(64, (count + 255) / 256));
  if (max_op) peer_allreduce_kernel<true><<<grid, 256, 0, s->stream>>>(p, buf, count, s->d_cg);
  else peer_allreduce_kernel<false><<<grid, 256, 0, s->stream>>>(p, buf, count, s->d_cg);
  ++s->launches;
  if (cudaGetLastError() != cudaSuccess) return Fail(CB200_ERR_CUDA, "peer all-reduce launch failed");
  return CB200_OK;
}

// ---- multi-GPU handle: one worker thread per shard ------------------------------------------------
void WorkerLoop(MultiGpu* mg, int rank) {
  unsigned seen = 0;
  for (;;) {
    // a short spin first: calls on a solver come back to back, and a futex wake-up costs more than a PCG iteration
    const auto t0 = std::chrono::steady_clock::now();
    while (mg->posted.load(std::memory_order_acquire) == seen &&
           std::chrono::steady_clock::now() - t0 < std::chrono::microseconds(200)) {}
    std::function<int(cb200_solver*, int)> task;
    {
      std::unique_lock<std::mutex> lock(mg->m);
      mg->cv_work.wait(lock, [&] { return mg->stop || mg->posted.load(std::memory_order_relaxed) != seen; });
      if (mg->stop) return;
      seen = mg->posted.load(std::memory_order_relaxed);
      task = mg->task;
    }
    g_error.clear();
    const int rc = task(rank < (int)mg->shards.size() ? mg->shards[rank] : nullptr, rank);
    {
      std::lock_guard<std::mutex> lock(mg->m);
      mg->rc[rank] = rc;
      mg->error[rank] = g_error;
      if (--mg->pending == 0) mg->cv_done.notify_all();
    }
  }
}

// Run fn(shard, rank) on every shard's worker thread and wait; the first failure is the result.
int Dispatch(cb200_solver* s, const std::function<int(cb200_solver*, int)>& fn) {
  MultiGpu* mg = s->multi;
  {
    std::lock_guard<std::mutex> lock(mg->m);
    mg->task = fn;
    mg->pending = (int)mg->threads.size();
    mg->posted.fetch_add(1, std::memory_order_release);
  }
  mg->cv_work.notify_all();
  {
    std::unique_lock<std::mutex> lock(mg->m);
    mg->cv_done.wait(lock, [&] { return mg->pending == 0; });
  }
  for (size_t r = 0; r < mg->rc.size(); ++r)
    if (mg->rc[r] != CB200_OK) return Fail(mg->rc[r], "GPU shard " + std::to_string(r) + ": " + mg->error[r]);
  return CB200_OK;
}

inline int Blocks(long long n, int threads) { return (int)((n + threads - 1) / threads); }

}  // namespace

extern "C" {

const char* cb200_last_error(void) { return g_error.c_str(); }
int cb200_version(void) { return 100; }

int cb200_nccl_unique_id(char* out128) {
  if (!out128) return Fail(CB200_ERR_INVALID_ARGUMENT, "null output");
  Nccl n;
  CHECK_OK(LoadNccl(&n));
  UniqueId id;
  const int rc = n.GetUniqueId(&id);
  if (rc != 0) return Fail(CB200_ERR_NCCL, "ncclGetUniqueId failed");
  std::memcpy(out128, id.internal, 128);
  return CB200_OK;
}

}  // extern "C" (the shard constructor below is internal)

namespace {

void DestroyShard(cb200_solver* s) {
  if (!s) return;
  if (s->opt.device_id >= 0) cudaSetDevice(s->opt.device_id);
  if (s->stream) cudaStreamSynchronize(s->stream);
  if (s->side_stream) cudaStreamSynchronize(s->side_stream);
  if (s->copy_stream) cudaStreamSynchronize(s->copy_stream);
  bool release_peer_memory = true;
  if (s->group != nullptr) {
    // shards of one handle leave together: nobody frees a buffer a slower shard may still be summing
    s->group->Barrier();
  } else if (s->peer_allreduce && s->h_cg && !s->h_cg->peer_timeout && s->d_cg && s->nccl.comm) {
    // One process per GPU: one NCCL all-reduce is the rendezvous (it completes here only after every rank
    // has reached it, i.e. is past all its peer reads). If a peer never shows up, give up after a few
    // seconds and leak the buffers rather than free memory somebody may still read.
    release_peer_memory = false;
    double* scratch = &s->d_cg->partial[5][0];
    if (s->nccl.AllReduce(scratch, scratch, 1, /*ncclDouble*/ 8, /*ncclSum*/ 0, s->nccl.comm, s->stream) == 0) {
      const auto t0 = std::chrono::steady_clock::now();
      while (std::chrono::steady_clock::now() - t0 < std::chrono::seconds(5)) {
        if (cudaStreamQuery(s->stream) != cudaErrorNotReady) { release_peer_memory = true; break; }
      }
    }
    if (!release_peer_memory) {
      cudaGetLastError();
      delete s;  // the stream still holds the pending collective: leave everything else to process exit
      return;
    }
  }
  for (auto& kv : s->graphs) cudaGraphExecDestroy(kv.second);
  for (PeerChannel& ch : s->peer) {
    for (int r = 0; r < 16 && ch.ipc; ++r)
      if (ch.base[r] && ch.base[r] != ch.d_buf) cudaIpcCloseMemHandle(ch.base[r]);
    if (ch.d_buf) cudaFree(ch.d_buf);
  }
  if (s->nccl.comm && s->nccl.CommDestroy) s->nccl.CommDestroy(s->nccl.comm);
  for (void* p : s->allocations) cudaFree(p);
  if (s->h_cg) cudaFreeHost(s->h_cg);
  if (s->h_cg_init) cudaFreeHost(s->h_cg_init);
  if (s->h_lm_scalars) cudaFreeHost(s->h_lm_scalars);
  for (auto& e : s->ev)
    if (e) cudaEventDestroy(e);
  for (auto& e : s->chunk_event)
    if (e) cudaEventDestroy(e);
  if (s->ev_copy_start) cudaEventDestroy(s->ev_copy_start);
  if (s->ev_fork) cudaEventDestroy(s->ev_fork);
  if (s->ev_join) cudaEventDestroy(s->ev_join);
  if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
  if (s->side_stream) cudaStreamDestroy(s->side_stream);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
}

int ValidateOptions(const cb200_options& o) {
  if (o.preconditioner_type < CB200_IDENTITY || o.preconditioner_type > CB200_SCHUR_POWER_SERIES_EXPANSION)
    return Fail(CB200_ERR_INVALID_ARGUMENT,
                "unsupported preconditioner_type (CLUSTER_JACOBI / CLUSTER_TRIDIAGONAL are not "
                "implemented on the device and there is no CPU fallback)");
  if (o.world_size < 1 || o.rank < 0 || o.rank >= o.world_size)
    return Fail(CB200_ERR_INVALID_ARGUMENT, "bad rank / world_size");
  if (o.num_gpus < 0 || o.num_gpus > kMaxPeers) return Fail(CB200_ERR_INVALID_ARGUMENT, "bad num_gpus");
  if (o.num_gpus > 1 && o.world_size > 1)
    return Fail(CB200_ERR_INVALID_ARGUMENT, "num_gpus > 1 (one handle, several GPUs) excludes world_size > 1 (one process per GPU)");
  if (o.world_size > 1 && !o.nccl_unique_id)
    return Fail(CB200_ERR_INVALID_ARGUMENT, "world_size > 1 needs nccl_unique_id");
  if (o.residual_reset_period <= 0 || o.max_num_iterations < 0 || o.min_num_iterations < 0)
    return Fail(CB200_ERR_INVALID_ARGUMENT, "bad iteration options");
  const bool several = o.world_size > 1 || o.num_gpus > 1;
  if (o.explicit_switch_iterations < 0 ||
      (o.explicit_switch_iterations > 0 &&
       (several || o.preconditioner_type == CB200_SCHUR_POWER_SERIES_EXPANSION || o.use_spse_initialization ||
        o.deterministic)))
    return Fail(CB200_ERR_INVALID_ARGUMENT,
                "explicit_switch_iterations needs one GPU, a block-diagonal preconditioner, no SPSE initialisation "
                "and no deterministic mode");
  if (o.use_explicit_schur_complement) {  // solver.cc:277-290
    if (o.preconditioner_type != CB200_SCHUR_JACOBI)
      return Fail(CB200_ERR_INVALID_ARGUMENT,
                  "use_explicit_schur_complement only supports SCHUR_JACOBI as the preconditioner.");
    if (o.use_spse_initialization)
      return Fail(CB200_ERR_INVALID_ARGUMENT,
                  "use_explicit_schur_complement does not support use_spse_initialization.");
    if (o.deterministic)
      return Fail(CB200_ERR_INVALID_ARGUMENT, "deterministic mode covers the implicit operator only");
  }
  if (o.h2d_chunks < 0 || o.h2d_chunks > 256) return Fail(CB200_ERR_INVALID_ARGUMENT, "bad h2d_chunks");
  return CB200_OK;
}

// One shard context: one GPU, one contiguous point range. `group` non-null: a shard of a multi-GPU handle
// (same process), created on its worker thread.
int CreateShard(const cb200_structure* structure, const cb200_options& o, InProcGroup* group, cb200_solver** out) {
  *out = nullptr;
  FlatStructure flat;
  std::string err;
  const int rc = Flatten(*structure, o.num_eliminate_blocks, &flat, &err);
  if (rc != CB200_OK) return Fail(rc, err);
  if (flat.num_obs == 0 || flat.num_cameras == 0 || flat.num_points == 0)
    return Fail(CB200_ERR_UNSUPPORTED_STRUCTURE, "empty problem");

  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return Fail(CB200_ERR_CUDA, "no CUDA device (this library has no CPU fallback)");
  if (o.device_id < 0 || o.device_id >= ndev) return Fail(CB200_ERR_INVALID_ARGUMENT, "bad device_id");
  CUDA_OK(cudaSetDevice(o.device_id));

  cb200_solver* s = new cb200_solver;
  s->opt = o;
  s->opt.nccl_unique_id = nullptr;
  s->group = group;
  s->writes_replicated = group == nullptr || o.rank == 0;
  s->P = flat.num_points; s->C = flat.num_cameras; s->N = flat.num_obs;
  s->canonical = flat.canonical_layout;
  const std::vector<int> bounds = ShardPoints(flat.point_row_start, o.world_size);
  s->p_begin = bounds[o.rank]; s->p_end = bounds[o.rank + 1];
  s->r_begin = flat.point_row_start[s->p_begin]; s->r_end = flat.point_row_start[s->p_end];
  s->P_loc = s->p_end - s->p_begin; s->N_loc = s->r_end - s->r_begin;
  s->nvec = 9 * s->C;
  s->vec_blocks = std::min(kMaxVecBlocks, Blocks(s->nvec, kVecThreads));
  s->nrep = o.y_replicas > 0 ? o.y_replicas : 8;
  s->rep_stride = (10 * s->C + 31) / 32 * 32;  // accumulators are padded to 10 doubles per camera

  auto cleanup_fail = [&](int code) {
    const std::string keep = g_error;
    if (s->group != nullptr) { s->group->Abort(); s->group = nullptr; }  // nobody waits for this shard any more
    DestroyShard(s);
    g_error = keep;
    return code;
  };
#define TRY(expr) do { int rc2_ = (expr); if (rc2_ != CB200_OK) return cleanup_fail(rc2_); } while (0)
#define TRY_CUDA(expr) do { cudaError_t e2_ = (expr); if (e2_ != cudaSuccess) { Fail(CB200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e2_)); return cleanup_fail(CB200_ERR_CUDA); } } while (0)

  TRY_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  for (auto& e : s->ev) TRY_CUDA(cudaEventCreate(&e));
  TRY_CUDA(cudaStreamCreateWithFlags(&s->side_stream, cudaStreamNonBlocking));
  TRY_CUDA(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
  TRY_CUDA(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
  TRY_CUDA(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
  TRY_CUDA(cudaEventCreate(&s->ev_copy_start));

  // local structure arrays
  std::vector<int> row_start(s->P_loc + 1), cam(s->N_loc);
  for (int p = 0; p <= s->P_loc; ++p) row_start[p] = flat.point_row_start[s->p_begin + p] - s->r_begin;
  std::copy(flat.cam_idx.begin() + s->r_begin, flat.cam_idx.begin() + s->r_end, cam.begin());
  const TilePlan plan = PlanTiles(row_start.data(), cam.data(), s->P_loc);
  s->num_tiles = (int)plan.tiles.size();
  s->num_long = (int)plan.long_points.size();
  s->num_empty = (int)plan.empty_points.size();
  TRY_CUDA(cudaDeviceGetAttribute(&s->num_sms, cudaDevAttrMultiProcessorCount, o.device_id));

  // chunks of the overlapped upload (cb200_solve): contiguous tile ranges that start at a tile-group start,
  // and the row ranges they cover (rows of long / empty points in between ride with the chunk before them)
  {
    int K = o.h2d_chunks > 0 ? o.h2d_chunks : 16;
    if (s->num_tiles < 4 * K || !s->canonical) K = 1;
    s->chunk_tile.assign(K + 1, s->num_tiles);
    s->chunk_row.assign(K + 1, s->N_loc);
    s->chunk_tile[0] = 0; s->chunk_row[0] = 0;
    for (int k = 1; k < K; ++k) {
      int t = (int)((long long)s->num_tiles * k / K);
      while (t < s->num_tiles && (plan.tiles[t].np & kTileContinues)) ++t;
      t = std::max(t, s->chunk_tile[k - 1]);
      s->chunk_tile[k] = t;
      s->chunk_row[k] = t < s->num_tiles ? plan.tiles[t].r0 : s->N_loc;
    }
    s->chunk_event.assign(K, nullptr);
    for (auto& e : s->chunk_event) TRY_CUDA(cudaEventCreate(&e));
  }

  // camera-major order of the local rows, cut into chunks of one camera each
  std::vector<int> cam_count(s->C + 1, 0);
  for (int r = 0; r < s->N_loc; ++r) ++cam_count[cam[r] + 1];
  std::partial_sum(cam_count.begin(), cam_count.end(), cam_count.begin());
  std::vector<int> cam_perm(s->N_loc), cursor(cam_count.begin(), cam_count.end() - 1);
  for (int r = 0; r < s->N_loc; ++r) cam_perm[cursor[cam[r]]++] = r;
  std::vector<int> chunk_begin, cam_chunk_start(s->C + 1, 0);
  for (int c = 0; c < s->C; ++c) {
    cam_chunk_start[c] = (int)chunk_begin.size();
    for (int b = cam_count[c]; b < cam_count[c + 1]; b += kChunkObs) chunk_begin.push_back(b);
  }
  cam_chunk_start[s->C] = (int)chunk_begin.size();
  s->num_chunks = (int)chunk_begin.size();
  chunk_begin.push_back(s->N_loc);
  // a chunk must end where its camera ends: fix the per-chunk end by storing
  // explicit begins only (end = next begin) -- true because chunks are emitted
  // camera by camera in camera-major order.

  // duplicate (point, camera) pairs
  std::vector<int> dup_pairs;
  {
    std::vector<std::pair<int, int>> cr;
    for (int p = 0; p < s->P_loc; ++p) {
      cr.clear();
      for (int r = row_start[p]; r < row_start[p + 1]; ++r) cr.emplace_back(cam[r], r);
      std::sort(cr.begin(), cr.end());
      for (size_t i = 0; i < cr.size(); ++i)
        for (size_t j = i + 1; j < cr.size() && cr[j].first == cr[i].first; ++j) {
          dup_pairs.push_back(p); dup_pairs.push_back(cr[i].second); dup_pairs.push_back(cr[j].second);
        }
    }
  }
  s->num_dup_pairs = (int)dup_pairs.size() / 3;

  TRY(DeviceUpload(s, &s->d_cam, cam));
  TRY(DeviceUpload(s, &s->d_row_start, row_start));
  TRY(DeviceUpload(s, &s->d_tiles, plan.tiles));
  TRY(DeviceUpload(s, &s->d_long_points, plan.long_points));
  TRY(DeviceUpload(s, &s->d_empty_points, plan.empty_points));
  TRY(DeviceUpload(s, &s->d_cam_perm, cam_perm));
  TRY(DeviceUpload(s, &s->d_chunk_begin, chunk_begin));
  TRY(DeviceUpload(s, &s->d_cam_chunk_start, cam_chunk_start));
  TRY(DeviceUpload(s, &s->d_dup_pairs, dup_pairs));
  if (!s->canonical) {
    std::vector<int> e_pos(flat.e_pos.begin() + s->r_begin, flat.e_pos.begin() + s->r_end);
    std::vector<int> f_pos(flat.f_pos.begin() + s->r_begin, flat.f_pos.begin() + s->r_end);
    TRY(DeviceUpload(s, &s->d_e_pos, e_pos));
    TRY(DeviceUpload(s, &s->d_f_pos, f_pos));
    long long max_pos = 0;
    for (int r = 0; r < s->N; ++r)
      max_pos = std::max<long long>(max_pos, std::max((long long)flat.e_pos[r] + 6, (long long)flat.f_pos[r] + 18));
    s->staging_len = (size_t)max_pos;
    TRY(DeviceAlloc(s, &s->d_values_staging, s->staging_len));
  }
  const size_t Nl = s->N_loc, Pl = s->P_loc, nv = s->nvec;
  TRY(DeviceAlloc(s, &s->d_E, Nl * 6));
  TRY(DeviceAlloc(s, &s->d_F, Nl * 18));
  TRY(DeviceAlloc(s, &s->d_b, Nl * 2));
  TRY(DeviceAlloc(s, &s->d_De, Pl * 3));
  TRY(DeviceAlloc(s, &s->d_Df, nv));
  TRY(DeviceAlloc(s, &s->d_ete_inv, Pl * 6));
  TRY(DeviceAlloc(s, &s->d_ete9, Pl * 9));
  TRY(DeviceAlloc(s, &s->d_weights, Nl * kWeightStride));
  TRY(DeviceAlloc(s, &s->d_chunk_partial, (size_t)s->num_chunks * 45));
  TRY(DeviceAlloc(s, &s->d_upper, (size_t)s->C * 45));
  TRY(DeviceAlloc(s, &s->d_dup_blocks, (size_t)s->C * 81));
  TRY(DeviceAlloc(s, &s->d_sj, (size_t)s->C * 81));
  TRY(DeviceAlloc(s, &s->d_ftf_inv, (size_t)s->C * 81));
  TRY(DeviceAlloc(s, &s->d_yrep, (size_t)s->nrep * s->rep_stride));
  TRY(DeviceAlloc(s, &s->d_y_e, Pl * 3));
  TRY(DeviceAlloc(s, &s->d_x10, (size_t)10 * s->C));
  TRY_CUDA(cudaMemset(s->d_x10, 0, (size_t)10 * s->C * sizeof(double)));
  double** vecs[] = {&s->d_rhs, &s->d_x, &s->d_r, &s->d_p, &s->d_z, &s->d_q, &s->d_t1, &s->d_t2};
  for (double** v : vecs) TRY(DeviceAlloc(s, v, nv));
  TRY(DeviceAlloc(s, &s->d_rows, Nl * 2));
  TRY(DeviceAlloc(s, &s->d_pts, Pl * 3));
  TRY(DeviceAlloc(s, &s->d_cg, 1));
  TRY_CUDA(cudaMemset(s->d_cg, 0, sizeof(CgState)));
  TRY_CUDA(cudaMemset(s->d_yrep, 0, (size_t)s->nrep * s->rep_stride * sizeof(double)));
  TRY_CUDA(cudaMemset(s->d_x, 0, nv * sizeof(double)));
  TRY_CUDA(cudaMallocHost(reinterpret_cast<void**>(&s->h_cg), sizeof(CgState)));
  std::memset(s->h_cg, 0, sizeof(CgState));
  TRY_CUDA(cudaMallocHost(reinterpret_cast<void**>(&s->h_cg_init), sizeof(CgState)));
  std::memset(s->h_cg_init, 0, sizeof(CgState));
  // back substitution from the per-iteration history: the implicit operator from a zero start only
  s->use_history = !o.use_explicit_schur_complement && o.explicit_switch_iterations == 0 &&
                   !o.use_spse_initialization && !EnvFlag("CB200_NO_BACKSUB_HISTORY");
  if (s->use_history) {
    TRY(DeviceAlloc(s, &s->d_v_rhs, Pl * 3));
    TRY(DeviceAlloc(s, &s->d_v_hist, (size_t)kHistCap * Pl * 3));
    TRY_CUDA(cudaMemset(s->d_v_rhs, 0, std::max<size_t>(1, Pl * 3) * sizeof(double)));  // points without rows stay 0
    TRY_CUDA(cudaMemset(s->d_v_hist, 0, std::max<size_t>(1, (size_t)kHistCap * Pl * 3) * sizeof(double)));
  }
  if (o.deterministic) TRY(DeviceAlloc(s, &s->d_chunk_partial9, (size_t)s->num_chunks * 9));
  s->graphs_ok = o.cuda_graphs >= 0 && o.explicit_switch_iterations == 0 && !EnvFlag("CB200_NO_GRAPHS");

#define SET_SMEM2(MODE, BUILD, DET)                                                        \
  TRY_CUDA(cudaFuncSetAttribute(point_stream_kernel<MODE, BUILD, DET>,                     \
                                cudaFuncAttributeMaxDynamicSharedMemorySize,               \
                                StreamLayout<MODE, BUILD>::kSmemBytes))
#define SET_SMEM(MODE, BUILD) SET_SMEM2(MODE, BUILD, false); SET_SMEM2(MODE, BUILD, true)
  SET_SMEM(kModeSx, false); SET_SMEM(kModePseZ, false); SET_SMEM(kModeRhs, false);
  SET_SMEM(kModeBack, false); SET_SMEM(kModeRhs, true);
#undef SET_SMEM
#undef SET_SMEM2
  TRY_CUDA(cudaFuncSetAttribute(camera_block_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kChunkSmemBytes));
  TRY_CUDA(cudaFuncSetAttribute(camera_block_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kChunkSmemBytes));

  {
    // the fused PCG step needs all its CTAs co-resident (grid barrier)
    int coop = 0, per_sm = 0;
    TRY_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, o.device_id));
    TRY_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cg_step_kernel<kStepNormal, 1>, kStepThreads, 0));
    const int ctas = Blocks(s->C, kStepCams);
    s->fused_step_ok = coop != 0 && ctas <= per_sm * s->num_sms && ctas <= kMaxStepBlocks &&
                       !EnvFlag("CB200_NO_FUSED_STEP");
  }
  const int world = o.world_size;
  const bool want_peer = world > 1 && world <= kMaxPeers && s->fused_step_ok && !EnvFlag("CB200_NO_PEER_ALLREDUCE");
  // NCCL: always with one process per GPU (handle exchange, fallback sums, the explicit matrix); shards of
  // one handle need it only for the explicit matrix or when peer access is unavailable (decided below)
  auto init_nccl = [&](const char* id128) -> int {
    CHECK_OK(LoadNccl(&s->nccl));
    UniqueId id;
    std::memcpy(id.internal, id128, 128);
    const int nrc = s->nccl.CommInitRank(&s->nccl.comm, world, id, o.rank);
    if (nrc != 0)
      return Fail(CB200_ERR_NCCL, std::string("ncclCommInitRank: ") +
                                      (s->nccl.GetErrorString ? s->nccl.GetErrorString(nrc) : "error"));
    return CB200_OK;
  };
  if (world > 1 && group == nullptr) TRY(init_nccl(o.nccl_unique_id));
  if (world > 1) {
    s->peer[0].slot = (size_t)81 * s->C;   // the largest per-solve sum: the cross-term blocks
    s->peer[1].slot = (size_t)9 * s->C;    // the reduced rhs on the side stream
    double ok = want_peer ? 1.0 : 0.0;
    for (PeerChannel& ch : s->peer) {
      if (ok == 0.0) break;
      const size_t bytes = 2 * ch.slot * sizeof(double) + 256;
      if (cudaMalloc(&ch.d_buf, bytes) != cudaSuccess || cudaMemset(ch.d_buf, 0, bytes) != cudaSuccess) {
        ok = 0.0;
        cudaGetLastError();
      }
    }
    if (group != nullptr) {
      // one process: peer access between the devices, addresses through host memory
      for (int r = 0; r < world && ok != 0.0; ++r) {
        const int dev = o.device_ids[r];
        if (dev == o.device_id) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, o.device_id, dev) != cudaSuccess || !can) { ok = 0.0; break; }
        const cudaError_t e = cudaDeviceEnablePeerAccess(dev, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = 0.0;
        cudaGetLastError();
      }
      for (int c = 0; c < 2; ++c) group->peer_buf[c][o.rank] = s->peer[c].d_buf;
      const double all_ok = group->Sum(o.rank, ok);  // (also the barrier that publishes the addresses)
      if (all_ok < 0.0) { Fail(CB200_ERR_STATE, "another GPU shard failed to initialise"); return cleanup_fail(CB200_ERR_STATE); }
      s->peer_allreduce = all_ok == (double)world;
      for (int c = 0; c < 2; ++c)
        for (int r = 0; r < world; ++r) s->peer[c].base[r] = group->peer_buf[c][r];
      const bool need_nccl = !s->peer_allreduce || o.use_explicit_schur_complement;
      if (need_nccl) {
        if (o.rank == 0) {
          if (cb200_nccl_unique_id(group->nccl_id) != CB200_OK) std::memset(group->nccl_id, 0, 128);
        }
        if (!group->Barrier()) { Fail(CB200_ERR_STATE, "another GPU shard failed to initialise"); return cleanup_fail(CB200_ERR_STATE); }
        const int nrc = init_nccl(group->nccl_id);
        const double bad = group->Sum(o.rank, nrc == CB200_OK ? 0.0 : 1.0);
        if (bad != 0.0) {
          if (nrc == CB200_OK) Fail(CB200_ERR_NCCL, "NCCL initialisation failed on another shard");
          return cleanup_fail(CB200_ERR_NCCL);
        }
      }
    } else if (s->nccl.AllGather != nullptr) {
      // one process per GPU: CUDA IPC handles travel through the NCCL communicator, so the C ABI needs
      // nothing new. All ranks must take the same decision: a sum of success flags settles it.
      struct Handles { cudaIpcMemHandle_t h[2]; };
      Handles mine;
      std::memset(&mine, 0, sizeof(mine));
      for (int c = 0; c < 2 && ok != 0.0; ++c)
        if (cudaIpcGetMemHandle(&mine.h[c], s->peer[c].d_buf) != cudaSuccess) { ok = 0.0; cudaGetLastError(); }
      std::vector<Handles> all(world);
      Handles* d_handles = nullptr;
      TRY(DeviceAlloc(s, &d_handles, (size_t)world));
      TRY_CUDA(cudaMemcpy(d_handles + o.rank, &mine, sizeof(mine), cudaMemcpyHostToDevice));
      if (s->nccl.AllGather(d_handles + o.rank, d_handles, sizeof(mine), /*ncclChar*/ 0, s->nccl.comm, s->stream) != 0) {
        Fail(CB200_ERR_NCCL, "ncclAllGather (IPC handles) failed");
        return cleanup_fail(CB200_ERR_NCCL);
      }
      TRY_CUDA(cudaStreamSynchronize(s->stream));
      TRY_CUDA(cudaMemcpy(all.data(), d_handles, all.size() * sizeof(mine), cudaMemcpyDeviceToHost));
      // every rank must have produced handles before anybody opens one
      double* d_ok = nullptr;
      TRY(DeviceAlloc(s, &d_ok, 1));
      TRY_CUDA(cudaMemcpy(d_ok, &ok, sizeof(double), cudaMemcpyHostToDevice));
      TRY(NcclAllReduce(s, d_ok, 1));
      TRY_CUDA(cudaStreamSynchronize(s->stream));
      double have = 0.0;
      TRY_CUDA(cudaMemcpy(&have, d_ok, sizeof(double), cudaMemcpyDeviceToHost));
      ok = (have == (double)world) ? 1.0 : 0.0;
      for (int c = 0; c < 2 && ok != 0.0; ++c) {
        s->peer[c].ipc = true;
        for (int r = 0; r < world && ok != 0.0; ++r) {
          if (r == o.rank) { s->peer[c].base[r] = s->peer[c].d_buf; continue; }
          if (cudaIpcOpenMemHandle(&s->peer[c].base[r], all[r].h[c], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
            s->peer[c].base[r] = nullptr;
            ok = 0.0;
            cudaGetLastError();
          }
        }
      }
      TRY_CUDA(cudaMemcpy(d_ok, &ok, sizeof(double), cudaMemcpyHostToDevice));
      TRY(NcclAllReduce(s, d_ok, 1));
      TRY_CUDA(cudaStreamSynchronize(s->stream));
      TRY_CUDA(cudaMemcpy(&ok, d_ok, sizeof(double), cudaMemcpyDeviceToHost));
      s->peer_allreduce = (ok == (double)world);
    }
  }
  if (o.use_explicit_schur_complement || o.explicit_switch_iterations > 0) {
    // block structure of S from the GLOBAL problem (identical on every rank), destinations of
    // this shard's outer products
    const ExplicitStructure es = PlanExplicitStructure(flat);
    const ExplicitFormPlan fp = PlanExplicitForm(es, row_start.data(), cam.data(), s->P_loc);
    s->ex_cells = (int)es.col.size();
    s->ex_items = (int)fp.items.size();
    s->ex_row_ptr = es.row_ptr;
    s->ex_col = es.col;
    TRY(DeviceUpload(s, &s->d_ex_row_ptr, es.row_ptr));
    TRY(DeviceUpload(s, &s->d_ex_col, es.col));
    TRY(DeviceUpload(s, &s->d_ex_t_ptr, es.t_ptr));
    TRY(DeviceUpload(s, &s->d_ex_t_cell, es.t_cell));
    TRY(DeviceUpload(s, &s->d_ex_t_row, es.t_row));
    TRY(DeviceUpload(s, &s->d_ex_srow, fp.srow));
    TRY(DeviceUpload(s, &s->d_ex_pair_start, fp.pair_start));
    TRY(DeviceUpload(s, &s->d_ex_pair_cell, fp.pair_cell));
    TRY(DeviceUpload(s, &s->d_ex_items, fp.items));
    TRY(DeviceAlloc(s, &s->d_ex_counter, 1));
    TRY(DeviceAlloc(s, &s->d_S, (size_t)s->ex_cells * 81));
  }
  s->global_dup_pairs = s->num_dup_pairs;
  if (world > 1) {
    // does any rank hold duplicate (point, camera) pairs? (decides whether the cross-term blocks are summed)
    double dups = (double)s->num_dup_pairs;
    if (group != nullptr) {
      dups = group->Sum(o.rank, dups);
      if (dups < 0.0) { Fail(CB200_ERR_STATE, "another GPU shard failed to initialise"); return cleanup_fail(CB200_ERR_STATE); }
    } else {
      double* d_dups = nullptr;
      TRY(DeviceAlloc(s, &d_dups, 1));
      TRY_CUDA(cudaMemcpy(d_dups, &dups, sizeof(double), cudaMemcpyHostToDevice));
      TRY(NcclAllReduce(s, d_dups, 1));
      TRY_CUDA(cudaStreamSynchronize(s->stream));
      TRY_CUDA(cudaMemcpy(&dups, d_dups, sizeof(double), cudaMemcpyDeviceToHost));
    }
    s->global_dup_pairs = (long long)dups;
  }
#undef TRY
#undef TRY_CUDA
  *out = s;
  return CB200_OK;
}

}  // namespace

extern "C" {

void cb200_destroy(cb200_solver* s) {
  if (!s) return;
  if (s->multi == nullptr) { DestroyShard(s); return; }
  MultiGpu* mg = s->multi;
  if (!mg->shards.empty())
    Dispatch(s, [mg](cb200_solver*, int rank) { DestroyShard(mg->shards[rank]); mg->shards[rank] = nullptr; return (int)CB200_OK; });
  {
    std::lock_guard<std::mutex> lock(mg->m);
    mg->stop = true;
  }
  mg->cv_work.notify_all();
  for (std::thread& t : mg->threads) t.join();
  delete mg;
  delete s;
}

int cb200_create(const cb200_structure* structure, const cb200_options* options, cb200_solver** out) {
  if (!structure || !options || !out) return Fail(CB200_ERR_INVALID_ARGUMENT, "null argument");
  *out = nullptr;
  CHECK_OK(ValidateOptions(*options));
  if (options->num_gpus <= 1) return CreateShard(structure, *options, nullptr, out);

  // ---- one handle, several GPUs: a shard per device, each on its own worker thread ----
  const int n = options->num_gpus;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return Fail(CB200_ERR_CUDA, "no CUDA device (this library has no CPU fallback)");
  cb200_options base = *options;
  bool all_zero = true;
  for (int g = 0; g < n; ++g) all_zero = all_zero && base.device_ids[g] == 0;
  for (int g = 0; g < n; ++g) {
    if (all_zero) base.device_ids[g] = g;
    if (base.device_ids[g] < 0 || base.device_ids[g] >= ndev)
      return Fail(CB200_ERR_INVALID_ARGUMENT, "num_gpus / device_ids name a device this machine does not have");
    for (int h = 0; h < g; ++h)
      if (base.device_ids[h] == base.device_ids[g]) return Fail(CB200_ERR_INVALID_ARGUMENT, "device_ids repeats a device");
  }
  // structure errors are reported once, from here, before any thread starts
  {
    FlatStructure flat;
    std::string err;
    const int rc = Flatten(*structure, base.num_eliminate_blocks, &flat, &err);
    if (rc != CB200_OK) return Fail(rc, err);
    if (flat.num_obs == 0 || flat.num_cameras == 0 || flat.num_points == 0)
      return Fail(CB200_ERR_UNSUPPORTED_STRUCTURE, "empty problem");
  }
  cb200_solver* parent = new cb200_solver;
  parent->opt = base;
  parent->opt.nccl_unique_id = nullptr;
  MultiGpu* mg = new MultiGpu;
  parent->multi = mg;
  mg->group.world = n;
  mg->shards.assign(n, nullptr);
  mg->rc.assign(n, CB200_OK);
  mg->error.assign(n, std::string());
  for (int g = 0; g < n; ++g) mg->threads.emplace_back(WorkerLoop, mg, g);
  const int rc = Dispatch(parent, [&](cb200_solver*, int rank) {
    cb200_options o = base;
    o.num_gpus = 1;
    o.rank = rank;
    o.world_size = n;
    o.device_id = base.device_ids[rank];
    const int r = CreateShard(structure, o, &mg->group, &mg->shards[rank]);
    // every shard must exist before any of them is used (or torn down): agree on the outcome
    // (a shard that failed early never reached the rendezvous inside CreateShard: it must not block here)
    return r;
  });
  if (rc != CB200_OK) {
    const std::string keep = g_error;
    for (cb200_solver*& c : mg->shards)
      if (c != nullptr) c->group = nullptr;  // no rendezvous: some shards do not exist
    cb200_destroy(parent);
    g_error = keep;
    return rc;
  }
  cb200_solver* s0 = mg->shards[0];
  parent->P = s0->P; parent->C = s0->C; parent->N = s0->N; parent->nvec = s0->nvec;
  parent->canonical = s0->canonical;
  parent->p_begin = 0; parent->p_end = s0->P; parent->r_begin = 0; parent->r_end = s0->N;
  parent->P_loc = s0->P; parent->N_loc = s0->N;
  parent->ex_cells = s0->ex_cells;
  *out = parent;
  return CB200_OK;
}

int cb200_shard_plan(const cb200_structure* structure, int num_eliminate_blocks, int world_size,
                     int* point_bounds, int* row_bounds) {
  if (!structure || world_size < 1 || !point_bounds || !row_bounds)
    return Fail(CB200_ERR_INVALID_ARGUMENT, "null argument or bad world_size");
  FlatStructure flat;
  std::string err;
  const int rc = Flatten(*structure, num_eliminate_blocks, &flat, &err);
  if (rc != CB200_OK) return Fail(rc, err);
  const std::vector<int> bounds = ShardPoints(flat.point_row_start, world_size);
  for (int g = 0; g <= world_size; ++g) {
    point_bounds[g] = bounds[g];
    row_bounds[g] = flat.point_row_start[bounds[g]];
  }
  return CB200_OK;
}

int cb200_explicit_plan(const cb200_structure* structure, int num_eliminate_blocks, int* num_cells,
                        int* row_ptr, int* cols) {
  if (!structure) return Fail(CB200_ERR_INVALID_ARGUMENT, "null structure");
  FlatStructure flat;
  std::string err;
  const int rc = Flatten(*structure, num_eliminate_blocks, &flat, &err);
  if (rc != CB200_OK) return Fail(rc, err);
  const ExplicitStructure es = PlanExplicitStructure(flat);
  if (num_cells) *num_cells = (int)es.col.size();
  if (row_ptr) std::copy(es.row_ptr.begin(), es.row_ptr.end(), row_ptr);
  if (cols) std::copy(es.col.begin(), es.col.end(), cols);
  return CB200_OK;
}

int cb200_num_rows(const cb200_solver* s) { return s ? 2 * s->N : 0; }
int cb200_num_cols(const cb200_solver* s) { return s ? 3 * s->P + 9 * s->C : 0; }
int cb200_num_cols_e(const cb200_solver* s) { return s ? 3 * s->P : 0; }
int cb200_num_cols_f(const cb200_solver* s) { return s ? 9 * s->C : 0; }
int cb200_shard(const cb200_solver* s, int* pb, int* pe, int* rb, int* re) {
  if (!s) return Fail(CB200_ERR_INVALID_ARGUMENT, "null solver");
  if (pb) *pb = s->p_begin;
  if (pe) *pe = s->p_end;
  if (rb) *rb = s->r_begin;
  if (re) *re = s->r_end;
  return CB200_OK;
}

}  // extern "C"

#include "c_api_solve.inc"
#include "c_api_explicit.inc"
#include "c_api_entry.inc"
#include "c_api_lm.inc"

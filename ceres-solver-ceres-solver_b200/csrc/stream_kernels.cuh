// The hot kernels: one persistent, TMA-pipelined pass over the Jacobian of a
// point shard that performs a whole ImplicitSchurComplement product.
//
// Reference (what one launch replaces):
//   ImplicitSchurComplement::RightMultiplyAndAccumulate  implicit_schur_complement.cc:106-144
//     = RightMultiplyAndAccumulateF, LeftMultiplyAndAccumulateE, block-diagonal
//       (E'E)^-1 multiply, RightMultiplyAndAccumulateE, LeftMultiplyAndAccumulateF
//       (partitioned_matrix_view_impl.h:112-375, block_sparse_matrix.cc:239-274)
//       and the ParallelSetZero / ParallelAssign passes over the 2N / 3P temporaries
//   ImplicitSchurComplement::UpdateRhs                   :251-276
//   ...::InversePowerSeriesOperatorRightMultiplyAccumulate :146-174
//   ImplicitSchurComplement::BackSubstitute              :208-243
//   (BUILD) UpdateBlockDiagonalEtE + AddDiagonalAndInvert
//           partitioned_matrix_view_impl.h:446-523, implicit_schur_complement.cc:179-204
//
// Structure: persistent CTAs of kStreamWarps autonomous warps. A warp owns a
// contiguous range of tiles (whole points, <= 32 rows: one lane per row) and a
// private ring of shared-memory stages. Its elected lane streams the tile's E
// rows, F rows, static side-band (TileMeta), (E'E)^-1 blocks and b rows into a
// stage with cp.async.bulk (TMA engine) completing on an mbarrier, always
// kStages-1 tiles ahead; the warp pulls its blocks into registers, refills the
// stage at once, and does the per-point segmented reduction with warp
// shuffles. No CTA-wide synchronisation exists after start-up, so an SM keeps
// ~12 independent tile pipelines in flight. Every Jacobian block crosses HBM
// exactly once per product.
#ifndef CB200_STREAM_KERNELS_CUH_
#define CB200_STREAM_KERNELS_CUH_

#include "cg_kernels.cuh"
#include "kernels.cuh"

namespace cb200 {

// pipeline depth and minimum resident CTAs per SM of the E'E / rhs (BUILD) pass and of the S*x pass (the other
// modes keep 2 / 3); macros so that tools/build_variant.py can A/B them
#ifndef CB200_STAGES_BUILD
#define CB200_STAGES_BUILD 1
#endif
#ifndef CB200_MINCTAS_BUILD
#define CB200_MINCTAS_BUILD 4
#endif
#ifndef CB200_STAGES_SX
#define CB200_STAGES_SX 1
#endif
// S*x with ONE data stage and the tile side-band (TileMeta) in its own ring of two slots: the camera ids of the
// NEXT tile are there one tile ahead (the x gather needs them), while E / F / (E'E)^-1 use a single stage that is
// refilled as soon as the lanes hold their blocks in registers: 10 KB instead of 17 KB per warp -> 16 warps per SM
#ifndef CB200_SX_META_RING
#define CB200_SX_META_RING 1
#endif
constexpr int kStreamWarps = 4;
constexpr int kStreamThreads = kStreamWarps * 32;
constexpr int kMetaBytes = (int)sizeof(TileMeta);
static_assert(kTileObs == 32, "one lane per row");

enum PointMode {
  kModeSx = 0,    // y += F'(t - E v),  t = F x          S*x without the D_f^2 term
  kModePseZ = 1,  // y += F'(E v),      t = F x          F'E (E'E)^-1 E'F x
  kModeRhs = 2,   // y += F'(t - E v),  t = b            reduced right-hand side
  kModeBack = 3   // y_e = v,           t = b - F x      back substitution
};                // with u = sum_rows E' t and v = (E'E + D_e^2)^-1 u per point

struct StreamArgs {
  const double* E;          // [N x 6]
  const double* F;          // [N x 18]
  const double* b;          // [2N] (kModeRhs, kModeBack); may be null in BUILD (t = 0)
  const TileMeta* tiles;
  int num_tiles;
  double* ete_inv;          // [P x 6] upper triangle of (E'E + D_e^2)^-1; written when BUILD
  const double* De;         // [3P] or null (BUILD)
  double* weights;          // [N x 3] or null (BUILD): W_r = I - E_r inverse(ete) E_r'
  const double* x;          // [10C] camera vector padded to 10 per camera (unused for kModeRhs)
  double* yrep;             // nrep privatised camera accumulators (10 per camera), rep_stride apart
  int nrep;
  int rep_stride;
  double* y_e;              // [3P] (kModeBack)
  const int* done;          // device convergence flag (or null): the kernel is a no-op when it is set ...
  int run_when_done;        // ... or, if this is non-zero, when it is NOT set (speculative back substitution)
  double* model;            // (kModeBack) or null: accumulates sum_rows m.(b - m/2) with m = E y_e + F x, the
                            // model cost change of the step -y (trust_region_minimizer.cc:430-438)
  // Back substitution without a pass over the Jacobian: y_e = (E'E+D^2)^-1 E'(b - F x) with x = sum_k alpha_k p_k
  // equals v_rhs - sum_k alpha_k v_k, v_k being exactly the per-point intermediate of the k-th S*p pass. The rhs
  // pass stores v_rhs, the first kHistCap S*p passes store their v (24 bytes per point each), and a solve that
  // ends within kHistCap iterations combines them (point_combine_kernel) instead of streaming J once more.
  double* v_store;          // (kModeRhs) [3P] or null: v of every point
  double* v_hist;           // (kModeSx) [kHistCap x v_hist_stride] or null: v of every point of PCG iteration cg->iter
  size_t v_hist_stride;
  const CgState* cg;        // with v_hist, or (kModeBack) to skip the pass when the history covers the solve
  double* rows_out;         // (DET) [2N]: w_r of every row instead of the atomic scatter of F_r' w_r
  unsigned long long* stamp;  // or null: CTA 0 writes %globaltimer here at entry (start of a solve)
  int zero_points;          // (kModeBack) y_e = 0 instead of the back-substituted value: the step [0; x] of an explicit
                            // Schur solve that ended without SUCCESS (schur_complement_solver.cc:135-160), whose model
                            // cost change the trust-region loop still needs (trust_region_minimizer.cc:430-447)
};

template <int MODE, bool BUILD>
struct StreamLayout {
  static constexpr bool kUsesB = (MODE == kModeRhs || MODE == kModeBack);
  static constexpr bool kLoadsInv = !BUILD;
  // modes that take the meta ring + single data stage: S*x and the power-series product (CB200_SX_META_RING >= 1),
  // back substitution too (>= 2)
  static constexpr bool kRingMode = !BUILD && ((CB200_SX_META_RING >= 1 && (MODE == 0 || MODE == 1)) ||
                                               (CB200_SX_META_RING >= 2 && MODE == 3));
  static constexpr int kStages = BUILD ? CB200_STAGES_BUILD : (kRingMode ? CB200_STAGES_SX : 2);
  // per stage (every offset a multiple of 16 bytes)
  static constexpr int kOffE = 0;
  static constexpr int kOffF = kOffE + kTileObs * 48;
  static constexpr int kOffMeta = kOffF + kTileObs * 144;
  static constexpr bool kMetaRing = kRingMode;
  static constexpr int kMetaSlots = kMetaRing ? 2 : 0;
  static constexpr int kOffInv = kOffMeta + (kMetaRing ? 0 : kMetaBytes);
  static constexpr int kOffB = kOffInv + (kLoadsInv ? kTilePoints * 48 : 0);
  static constexpr int kStageBytes = kOffB + (kUsesB ? kTileObs * 16 : 0);
  // per warp: stages, (the meta ring,) then the 32 x 10 transposition scratch, then the barriers
  static constexpr int kOffMetaRing = kStages * kStageBytes;
  static constexpr int kOffX = kOffMetaRing + kMetaSlots * kMetaBytes;
  static constexpr int kOffBars = kOffX + 320 * 8;
  static constexpr int kNumBars = kStages + kMetaSlots;
  static constexpr int kOffCarry = kOffBars + 16 * ((kNumBars * 8 + 15) / 16);  // 18 doubles, long points only
  static constexpr int kWarpBytes = kOffCarry + 144;
  static constexpr int kSmemBytes = kStreamWarps * kWarpBytes;
};

template <typename L>
__device__ __forceinline__ void IssueTile(const StreamArgs& a, int t, const int4 h /* r0, n, p0, np */,
                                          unsigned char* st, uint64_t* bar, uint64_t stream_policy) {
  const bool use_b = L::kUsesB && a.b != nullptr;
  const uint32_t bytes = (uint32_t)h.y * 192u + (L::kMetaRing ? 0u : (uint32_t)kMetaBytes) +
                         (L::kLoadsInv ? (uint32_t)(h.w & 255) * 48u : 0u) + (use_b ? (uint32_t)h.y * 16u : 0u);
  mbar_expect_tx(bar, bytes);
  if (!L::kMetaRing) bulk_g2s(st + L::kOffMeta, a.tiles + t, (uint32_t)kMetaBytes, bar);
  // Jacobian blocks and b are read once per pass: first in line for eviction
  bulk_g2s_hint(st + L::kOffE, a.E + (size_t)h.x * 6, (uint32_t)h.y * 48u, bar, stream_policy);
  bulk_g2s_hint(st + L::kOffF, a.F + (size_t)h.x * 18, (uint32_t)h.y * 144u, bar, stream_policy);
  if (L::kLoadsInv) bulk_g2s(st + L::kOffInv, a.ete_inv + (size_t)h.z * 6, (uint32_t)(h.w & 255) * 48u, bar);
  if (use_b) bulk_g2s_hint(st + L::kOffB, a.b + (size_t)h.x * 2, (uint32_t)h.y * 16u, bar, stream_policy);
}

// (meta ring) the side-band record of tile t into a ring slot
__device__ __forceinline__ void IssueMeta(const StreamArgs& a, int t, unsigned char* slot, uint64_t* bar) {
  mbar_expect_tx(bar, (uint32_t)kMetaBytes);
  bulk_g2s(slot, a.tiles + t, (uint32_t)kMetaBytes, bar);
}

// Tile header (r0, n, p0, np). volatile asm: the compiler must keep the value in
// registers across the loop iteration instead of re-issuing the load where it is used.
__device__ __forceinline__ int4 LoadHeader(const TileMeta* t) {
  int4 h;
  asm volatile("ld.global.nc.v4.s32 {%0, %1, %2, %3}, [%4];"
               : "=r"(h.x), "=r"(h.y), "=r"(h.z), "=r"(h.w) : "l"(t));
  return h;
}

// Inclusive segmented sum along the lanes; segments are lane ranges [head, tail].
// mask[s] is 1.0 where lane - 2^s >= head, else 0.0: one DFMA per step and value.
struct SegMask { double m[5]; };
__device__ __forceinline__ SegMask MakeSegMask(int lane, int head) {
  SegMask k;
#pragma unroll
  for (int s = 0; s < 5; ++s) k.m[s] = (lane - (1 << s) >= head) ? 1.0 : 0.0;
  return k;
}
__device__ __forceinline__ double SegScan(double v, const SegMask& k) {
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    const double up = __shfl_up_sync(0xffffffffu, v, 1 << s);
    v = fma(up, k.m[s], v);
  }
  return v;
}

template <int MODE, bool BUILD, bool DET = false>
__global__ void __launch_bounds__(kStreamThreads, BUILD ? CB200_MINCTAS_BUILD : 3)
point_stream_kernel(const StreamArgs a) {
  using L = StreamLayout<MODE, BUILD>;
  static_assert(!BUILD || MODE == kModeRhs, "the E'E build rides on the rhs pass");
  if (a.done != nullptr && ((*a.done != 0) != (a.run_when_done != 0))) return;
  if (MODE == kModeBack && a.cg != nullptr && a.cg->hist_valid && a.cg->iter <= kHistCap) return;  // point_combine_kernel does it
  if (a.stamp != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *a.stamp = GlobalTimerNs();
  // (kModeSx) slot of this product's v in the history, or -1
  int hist_slot = -1;
  if (MODE == kModeSx && a.v_hist != nullptr) {
    const int it = a.cg->iter;
    if (a.cg->hist_valid && it >= 1 && it <= kHistCap) hist_slot = it - 1;
  }
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  unsigned char* wsm = smem + warp * L::kWarpBytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(wsm + L::kOffBars);
  double* sXw = reinterpret_cast<double*>(wsm + L::kOffX);
  // contiguous tile range of this warp
  const int gw = blockIdx.x * kStreamWarps + warp;
  const int GW = gridDim.x * kStreamWarps;
  int t_begin = (int)((long long)a.num_tiles * gw / GW);
  int t_end = (int)((long long)a.num_tiles * (gw + 1) / GW);
  // a group of chunk tiles (one long point) stays with one warp: ranges start at group starts
  while (t_begin < a.num_tiles && (LoadHeader(a.tiles + t_begin).w & kTileContinues)) ++t_begin;
  while (t_end < a.num_tiles && (LoadHeader(a.tiles + t_end).w & kTileContinues)) ++t_end;
  if (t_begin >= t_end) return;
  // header of the tile to issue next, loaded one iteration before it is needed
  int4 hnext = make_int4(0, 0, 0, 0);
  const uint64_t stream_policy = policy_evict_first();
  if (lane == 0) {
    for (int s = 0; s < L::kNumBars; ++s) mbar_init(full + s, 1);
    for (int s = 0; s < L::kStages && t_begin + s < t_end; ++s)
      IssueTile<L>(a, t_begin + s, LoadHeader(a.tiles + t_begin + s),
                   wsm + s * L::kStageBytes, full + s, stream_policy);
    for (int k = 0; k < L::kMetaSlots && t_begin + k < t_end; ++k)
      IssueMeta(a, t_begin + k, wsm + L::kOffMetaRing + k * kMetaBytes, full + L::kStages + k);
    if (t_begin + L::kStages < t_end) hnext = LoadHeader(a.tiles + t_begin + L::kStages);
  }
  __syncwarp();
  double* yacc = a.yrep + (size_t)(gw % a.nrep) * a.rep_stride;

  // x is gathered one tile ahead: the loads of tile i+1's camera 9-vectors are in
  // flight while tile i is processed. Camera vectors are padded to 10 doubles, so
  // the warp's 32 x 10 values are 160 aligned 16-byte pieces: lane l loads pieces
  // j*32 + l (consecutive lanes read consecutive pieces of a camera).
  double2 xq[5];
  auto gather_x = [&](const TileMeta* m) {
    const int c_own = m->cam[lane];  // 10 * camera; 0 for idle lanes (no select may sit on the loads)
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int idx = j * 32 + lane;
      const int o = idx / 5;
      const int h = idx - 5 * o;
      const int c = __shfl_sync(0xffffffffu, c_own, o);
      xq[j] = __ldg(reinterpret_cast<const double2*>(a.x + c) + h);
    }
  };
  if (MODE != kModeRhs) {
    if (L::kMetaRing) {
      mbar_wait(full + L::kStages, 0);
      gather_x(reinterpret_cast<const TileMeta*>(wsm + L::kOffMetaRing));
    } else {
      mbar_wait(full + 0, 0);
      gather_x(reinterpret_cast<const TileMeta*>(wsm + L::kOffMeta));
    }
  }

  // carry of a long point across its chunk tiles: kept in shared memory (touched by chunk
  // tiles only, so it costs the common path no registers); every lane reads / writes the same
  double* cu = reinterpret_cast<double*>(wsm + L::kOffCarry);  // [3] running sums of u
  double* cv = cu + 3;                                          // [3] v of the finished point
  double* cq = cu + 6;                                          // [6] running sums of E'E
  double* cc = cu + 12;                                         // [6] adjugate inverse of the finished point
  if (lane < 18) cu[lane] = 0.0;
  __syncwarp();

  double model_acc = 0.0;  // (kModeBack) this lane's share of the model cost change
  for (int t = t_begin, i = 0; t < t_end; ++t, ++i) {
    const int s = i % L::kStages;
    unsigned char* st = wsm + s * L::kStageBytes;
    mbar_wait(full + s, (i / L::kStages) & 1);
    if (L::kMetaRing) mbar_wait(full + L::kStages + (i & 1), (i >> 1) & 1);
    const TileMeta* m = L::kMetaRing
                            ? reinterpret_cast<const TileMeta*>(wsm + L::kOffMetaRing + (i & 1) * kMetaBytes)
                            : reinterpret_cast<const TileMeta*>(st + L::kOffMeta);
    const int r0 = m->r0, n = m->n, p0 = m->p0;
    const int flags = m->np;
    const int kind = (flags >> kTileKindShift) & 3;  // 0 whole points, 1 / 2 chunk of a long point
    // ---- everything this lane needs from the stage goes to registers ----
    double Ei[6], Fi[18];
    double t0 = 0.0, t1 = 0.0;
    double fz0 = 0.0, fz1 = 0.0;  // (kModeBack) F x of this row
    const int cam = m->cam[lane];
    const int seg = m->seg[lane];
    const int pt = seg & 255, head = (seg >> 8) & 255, tail = (seg >> 16) & 255;
    const bool active = lane < n;
    const bool is_tail = active && lane == tail;
    {
      const double2* e2 = reinterpret_cast<const double2*>(st + L::kOffE) + lane * 3;
      const double2* f2 = reinterpret_cast<const double2*>(st + L::kOffF) + lane * 9;
#pragma unroll
      // idle lanes (lane >= n) read stale stage bytes: each is its own segment, and
      // nothing they compute is stored or scattered
      for (int k = 0; k < 3; ++k) { const double2 v = e2[k]; Ei[2 * k] = v.x; Ei[2 * k + 1] = v.y; }
#pragma unroll
      for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; Fi[2 * k] = v.x; Fi[2 * k + 1] = v.y; }
      if (L::kUsesB && a.b != nullptr) {
        const double2 bb = reinterpret_cast<const double2*>(st + L::kOffB)[lane];
        t0 = bb.x; t1 = bb.y;
      }
    }
    double inv[6];
    double de0 = 0.0, de1 = 0.0, de2 = 0.0;
    if (is_tail) {
      if (L::kLoadsInv) {
        const double2* g = reinterpret_cast<const double2*>(st + L::kOffInv) + pt * 3;
#pragma unroll
        for (int k = 0; k < 3; ++k) { const double2 v = g[k]; inv[2 * k] = v.x; inv[2 * k + 1] = v.y; }
      }
      if (BUILD && a.De != nullptr) {
        const double* d = a.De + 3 * (size_t)(p0 + pt);
        de0 = __ldg(d); de1 = __ldg(d + 1); de2 = __ldg(d + 2);
      }
    }
    if (MODE != kModeRhs) {
#pragma unroll
      for (int j = 0; j < 5; ++j) reinterpret_cast<double2*>(sXw)[j * 32 + lane] = xq[j];  // park this tile's x
    }
    __syncwarp();
    // the stage is free: refill it with the tile kStages ahead
    if (lane == 0 && t + L::kStages < t_end) {
      IssueTile<L>(a, t + L::kStages, hnext, st, full + s, stream_policy);
      if (t + L::kStages + 1 < t_end) hnext = LoadHeader(a.tiles + t + L::kStages + 1);
    }
    // (meta ring) this tile's record is in registers: its slot takes the record two tiles ahead
    if (L::kMetaRing && lane == 0 && t + 2 < t_end)
      IssueMeta(a, t + 2, wsm + L::kOffMetaRing + (i & 1) * kMetaBytes, full + L::kStages + (i & 1));

    // ---- t = F x; start the gather of the next tile ----
    if (MODE != kModeRhs) {
      if (t + 1 < t_end) {
        if (L::kMetaRing) {
          mbar_wait(full + L::kStages + ((i + 1) & 1), ((i + 1) >> 1) & 1);
          gather_x(reinterpret_cast<const TileMeta*>(wsm + L::kOffMetaRing + ((i + 1) & 1) * kMetaBytes));
        } else {
          const int s1 = (i + 1) % L::kStages;
          mbar_wait(full + s1, ((i + 1) / L::kStages) & 1);
          gather_x(reinterpret_cast<const TileMeta*>(wsm + s1 * L::kStageBytes + L::kOffMeta));
        }
      }
      double xr[10];
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        const double2 v = reinterpret_cast<const double2*>(sXw)[lane * 5 + k];
        xr[2 * k] = v.x; xr[2 * k + 1] = v.y;
      }
      // two chains per output row halve the dependent-DFMA depth
      double f0 = Fi[0] * xr[0], f0b = Fi[1] * xr[1], f1 = Fi[9] * xr[0], f1b = Fi[10] * xr[1];
#pragma unroll
      for (int k = 2; k < 8; k += 2) {
        f0 = fma(Fi[k], xr[k], f0); f0b = fma(Fi[k + 1], xr[k + 1], f0b);
        f1 = fma(Fi[9 + k], xr[k], f1); f1b = fma(Fi[10 + k], xr[k + 1], f1b);
      }
      f0 = fma(Fi[8], xr[8], f0) + f0b;
      f1 = fma(Fi[17], xr[8], f1) + f1b;
      if (MODE == kModeBack) { t0 -= f0; t1 -= f1; fz0 = f0; fz1 = f1; } else { t0 = f0; t1 = f1; }
      __syncwarp();  // sXw is reused by the scatter below
    }
    // ---- u = sum over the point's rows of E_r' t_r (and of E_r'E_r) ----
    const SegMask km = MakeSegMask(lane, head);
    double u0 = SegScan(Ei[0] * t0 + Ei[3] * t1, km);
    double u1 = SegScan(Ei[1] * t0 + Ei[4] * t1, km);
    double u2 = SegScan(Ei[2] * t0 + Ei[5] * t1, km);
    double c6[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};  // adjugate inverse (Schur-Jacobi weights)
    double q[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (BUILD) {
      q[0] = SegScan(Ei[0] * Ei[0] + Ei[3] * Ei[3], km);
      q[1] = SegScan(Ei[0] * Ei[1] + Ei[3] * Ei[4], km);
      q[2] = SegScan(Ei[0] * Ei[2] + Ei[3] * Ei[5], km);
      q[3] = SegScan(Ei[1] * Ei[1] + Ei[4] * Ei[4], km);
      q[4] = SegScan(Ei[1] * Ei[2] + Ei[4] * Ei[5], km);
      q[5] = SegScan(Ei[2] * Ei[2] + Ei[5] * Ei[5], km);
    }
    if (kind == 1) {
      // first-pass chunk of a long point: add the chunk's sums (at lane n-1) to the carry
      u0 = cu[0] + __shfl_sync(0xffffffffu, u0, n - 1);
      u1 = cu[1] + __shfl_sync(0xffffffffu, u1, n - 1);
      u2 = cu[2] + __shfl_sync(0xffffffffu, u2, n - 1);
      if (BUILD) {
#pragma unroll
        for (int k = 0; k < 6; ++k) q[k] = cq[k] + __shfl_sync(0xffffffffu, q[k], n - 1);
      }
      const bool last = (flags & kTileLastChunk) != 0;
      __syncwarp();
      if (lane == 0) {
        cu[0] = last ? 0.0 : u0; cu[1] = last ? 0.0 : u1; cu[2] = last ? 0.0 : u2;
        if (BUILD) {
#pragma unroll
          for (int k = 0; k < 6; ++k) cq[k] = last ? 0.0 : q[k];
        }
      }
      __syncwarp();
      if (!last) continue;
    }
    if (BUILD && is_tail && kind != 2) {
      double mm[9] = {q[0], q[1], q[2], q[1], q[3], q[4], q[2], q[4], q[5]};
      mm[0] += de0 * de0; mm[4] += de1 * de1; mm[8] += de2 * de2;
      if (a.weights != nullptr) {
        double c[9];
        Inverse3x3(mm, c);
        c6[0] = c[0]; c6[1] = c[1]; c6[2] = c[2]; c6[3] = c[4]; c6[4] = c[5]; c6[5] = c[8];
      }
      LltInverseUpper<3>(mm);
      inv[0] = mm[0]; inv[1] = mm[1]; inv[2] = mm[2]; inv[3] = mm[4]; inv[4] = mm[5]; inv[5] = mm[8];
      double2* g = reinterpret_cast<double2*>(a.ete_inv + (size_t)(p0 + pt) * 6);
      g[0] = make_double2(inv[0], inv[1]);
      g[1] = make_double2(inv[2], inv[3]);
      g[2] = make_double2(inv[4], inv[5]);
    }
    // ---- v = inverse * u at the point's last row, then to all its rows ----
    double v0 = 0.0, v1 = 0.0, v2 = 0.0;
    if (is_tail && kind != 2) {
      v0 = inv[0] * u0 + inv[1] * u1 + inv[2] * u2;
      v1 = inv[1] * u0 + inv[3] * u1 + inv[4] * u2;
      v2 = inv[2] * u0 + inv[4] * u1 + inv[5] * u2;
      if (MODE == kModeBack && a.zero_points) { v0 = 0.0; v1 = 0.0; v2 = 0.0; }
      if (MODE == kModeBack) {
        double* y = a.y_e + (size_t)(p0 + pt) * 3;
        y[0] = v0; y[1] = v1; y[2] = v2;
      }
      if (MODE == kModeRhs && a.v_store != nullptr) {
        double* y = a.v_store + (size_t)(p0 + pt) * 3;
        y[0] = v0; y[1] = v1; y[2] = v2;
      }
      if (MODE == kModeSx && hist_slot >= 0) {
        double* y = a.v_hist + (size_t)hist_slot * a.v_hist_stride + (size_t)(p0 + pt) * 3;
        y[0] = v0; y[1] = v1; y[2] = v2;
      }
    }
    if (MODE == kModeBack && a.model == nullptr) continue;
    // (a chunk's carry must be the same in every lane, idle ones included: they are live rows later)
    const int src = (kind == 1) ? n - 1 : tail;
    v0 = __shfl_sync(0xffffffffu, v0, src);
    v1 = __shfl_sync(0xffffffffu, v1, src);
    v2 = __shfl_sync(0xffffffffu, v2, src);
    if (BUILD && a.weights != nullptr) {
#pragma unroll
      for (int k = 0; k < 6; ++k) c6[k] = __shfl_sync(0xffffffffu, c6[k], src);
    }
    if (kind == 1) {
      // the point's sums are complete: keep v (and the adjugate inverse) for its second-pass chunks
      if (lane == 0) {
        cv[0] = v0; cv[1] = v1; cv[2] = v2;
#pragma unroll
        for (int k = 0; k < 6; ++k) cc[k] = c6[k];
      }
      __syncwarp();
      continue;
    }
    if (kind == 2) {
      v0 = cv[0]; v1 = cv[1]; v2 = cv[2];
#pragma unroll
      for (int k = 0; k < 6; ++k) c6[k] = cc[k];
    }
    if (MODE == kModeBack) {
      // model cost change of the step -y: m = J y for this row, b = t + F x
      const double m0 = Ei[0] * v0 + Ei[1] * v1 + Ei[2] * v2 + fz0;
      const double m1 = Ei[3] * v0 + Ei[4] * v1 + Ei[5] * v2 + fz1;
      if (active) model_acc += m0 * (t0 + fz0 - 0.5 * m0) + m1 * (t1 + fz1 - 0.5 * m1);
      continue;
    }
    // ---- w_r = t_r - E_r v (or E_r v); F_r' w_r goes to the camera accumulator ----
    {
      const double ev0 = Ei[0] * v0 + Ei[1] * v1 + Ei[2] * v2;
      const double ev1 = Ei[3] * v0 + Ei[4] * v1 + Ei[5] * v2;
      const double w0 = (MODE == kModePseZ) ? ev0 : t0 - ev0;
      const double w1 = (MODE == kModePseZ) ? ev1 : t1 - ev1;
      if (DET) {
        // deterministic mode: the row's w leaves as is; camera_gather_kernel sums F_r' w_r per camera in a fixed order
        if (active && !(BUILD && a.b == nullptr))
          reinterpret_cast<double2*>(a.rows_out)[(size_t)r0 + lane] = make_double2(w0, w1);
      } else {
#pragma unroll
        for (int k = 0; k < 9; ++k) sXw[lane * 9 + k] = Fi[k] * w0 + Fi[9 + k] * w1;
      }
    }
    if (BUILD && a.weights != nullptr) {
      // W_r = I - E_r c E_r' with c the adjugate inverse (InvertPSDMatrix<3>)
      const double g00 = Ei[0] * c6[0] + Ei[1] * c6[1] + Ei[2] * c6[2];
      const double g01 = Ei[0] * c6[1] + Ei[1] * c6[3] + Ei[2] * c6[4];
      const double g02 = Ei[0] * c6[2] + Ei[1] * c6[4] + Ei[2] * c6[5];
      const double g10 = Ei[3] * c6[0] + Ei[4] * c6[1] + Ei[5] * c6[2];
      const double g11 = Ei[3] * c6[1] + Ei[4] * c6[3] + Ei[5] * c6[4];
      const double g12 = Ei[3] * c6[2] + Ei[4] * c6[4] + Ei[5] * c6[5];
      if (active) {
        double* w = a.weights + (size_t)(r0 + lane) * kWeightStride;
        const double w00 = 1.0 - (g00 * Ei[0] + g01 * Ei[1] + g02 * Ei[2]);
        const double w01 = -(g00 * Ei[3] + g01 * Ei[4] + g02 * Ei[5]);
        const double w11 = 1.0 - (g10 * Ei[3] + g11 * Ei[4] + g12 * Ei[5]);
        if (kWeightStride == 4) {
          reinterpret_cast<double2*>(w)[0] = make_double2(w00, w01);
          reinterpret_cast<double2*>(w)[1] = make_double2(w11, 0.0);
        } else {
          w[0] = w00; w[1] = w01; w[2] = w11;
        }
      }
    }
    __syncwarp();
    if (!DET && !(BUILD && a.b == nullptr)) {
#pragma unroll
      for (int j = 0; j < 9; ++j) {
        const int idx = j * 32 + lane;
        const int o = idx / 9;
        const int k = idx - 9 * o;
        const int c = __shfl_sync(0xffffffffu, cam, o);
        if (o < n) atomicAdd(yacc + c + k, sXw[idx]);
      }
    }
    __syncwarp();
  }
  if (MODE == kModeBack && a.model != nullptr) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) model_acc += __shfl_xor_sync(0xffffffffu, model_acc, off);
    if (lane == 0) atomicAdd(a.model, model_acc);  // one atomic per warp, not per tile
  }
}

// Points without rows (legal in the block structure, absent from real bundle
// adjustment programs): (E'E + D^2)^-1 = D^-2 and a zero back-substituted step.
__global__ void empty_points_kernel(const int* __restrict__ empty_points, int count,
                                    const double* __restrict__ De, double* __restrict__ ete_inv,
                                    double* __restrict__ y_e) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const int p = empty_points[i];
  if (ete_inv != nullptr) {
    double mm[9] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
    if (De != nullptr)
      for (int k = 0; k < 3; ++k) mm[4 * k] = De[3 * (size_t)p + k] * De[3 * (size_t)p + k];
    LltInverseUpper<3>(mm);
    double* g = ete_inv + (size_t)p * 6;
    g[0] = mm[0]; g[1] = mm[1]; g[2] = mm[2]; g[3] = mm[4]; g[4] = mm[5]; g[5] = mm[8];
  }
  if (y_e != nullptr) { y_e[3 * (size_t)p] = 0.0; y_e[3 * (size_t)p + 1] = 0.0; y_e[3 * (size_t)p + 2] = 0.0; }
}

// A point with more rows than a tile: one CTA walks its rows twice (the second
// pass hits L2). grid = number of long points. Same modes as above.
template <int MODE, bool BUILD>
__global__ void __launch_bounds__(256)
point_long_kernel(const StreamArgs a, const int* __restrict__ long_points,
                  const int* __restrict__ row_start, const int* __restrict__ cam) {
  if (a.done != nullptr && ((*a.done != 0) != (a.run_when_done != 0))) return;
  if (MODE == kModeBack && a.cg != nullptr && a.cg->hist_valid && a.cg->iter <= kHistCap) return;
  __shared__ double sbuf[8];
  const int p = long_points[blockIdx.x];
  const int r0 = row_start[p], r1 = row_start[p + 1];
  const bool use_b = (MODE == kModeRhs || MODE == kModeBack) && a.b != nullptr;
  double u[3] = {0.0, 0.0, 0.0};
  double q[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
  for (int r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
    const double* Ei = a.E + (size_t)r * 6;
    const double* Fi = a.F + (size_t)r * 18;
    double t0 = 0.0, t1 = 0.0;
    if (use_b) { t0 = a.b[2 * (size_t)r]; t1 = a.b[2 * (size_t)r + 1]; }
    if (MODE != kModeRhs) {
      const double* x = a.x + 10 * (size_t)cam[r];
      double f0 = 0.0, f1 = 0.0;
      for (int k = 0; k < 9; ++k) { f0 += Fi[k] * x[k]; f1 += Fi[9 + k] * x[k]; }
      if (MODE == kModeBack) { t0 -= f0; t1 -= f1; } else { t0 = f0; t1 = f1; }
    }
    for (int k = 0; k < 3; ++k) u[k] += Ei[k] * t0 + Ei[3 + k] * t1;
    if (BUILD) {
      q[0] += Ei[0] * Ei[0] + Ei[3] * Ei[3];
      q[1] += Ei[0] * Ei[1] + Ei[3] * Ei[4];
      q[2] += Ei[0] * Ei[2] + Ei[3] * Ei[5];
      q[3] += Ei[1] * Ei[1] + Ei[4] * Ei[4];
      q[4] += Ei[1] * Ei[2] + Ei[4] * Ei[5];
      q[5] += Ei[2] * Ei[2] + Ei[5] * Ei[5];
    }
  }
  for (int k = 0; k < 3; ++k) u[k] = BlockSum(u[k], sbuf);
  double inv[6], c[9];
  if (BUILD) {
    for (int k = 0; k < 6; ++k) q[k] = BlockSum(q[k], sbuf);
    double mm[9] = {q[0], q[1], q[2], q[1], q[3], q[4], q[2], q[4], q[5]};
    if (a.De != nullptr) {
      const double* d = a.De + 3 * (size_t)p;
      mm[0] += d[0] * d[0]; mm[4] += d[1] * d[1]; mm[8] += d[2] * d[2];
    }
    Inverse3x3(mm, c);
    LltInverseUpper<3>(mm);
    inv[0] = mm[0]; inv[1] = mm[1]; inv[2] = mm[2]; inv[3] = mm[4]; inv[4] = mm[5]; inv[5] = mm[8];
    if (threadIdx.x == 0)
      for (int k = 0; k < 6; ++k) a.ete_inv[(size_t)p * 6 + k] = inv[k];
  } else {
    for (int k = 0; k < 6; ++k) inv[k] = a.ete_inv[(size_t)p * 6 + k];
  }
  const bool zero = MODE == kModeBack && a.zero_points;
  const double v0 = zero ? 0.0 : inv[0] * u[0] + inv[1] * u[1] + inv[2] * u[2];
  const double v1 = zero ? 0.0 : inv[1] * u[0] + inv[3] * u[1] + inv[4] * u[2];
  const double v2 = zero ? 0.0 : inv[2] * u[0] + inv[4] * u[1] + inv[5] * u[2];
  if (threadIdx.x == 0) {
    double* vs = nullptr;
    if (MODE == kModeRhs) vs = a.v_store;
    if (MODE == kModeSx && a.v_hist != nullptr) {
      const int it = a.cg->iter;
      if (a.cg->hist_valid && it >= 1 && it <= kHistCap) vs = a.v_hist + (size_t)(it - 1) * a.v_hist_stride;
    }
    if (vs != nullptr) { vs[3 * (size_t)p] = v0; vs[3 * (size_t)p + 1] = v1; vs[3 * (size_t)p + 2] = v2; }
  }
  if (MODE == kModeBack) {
    if (threadIdx.x == 0) { a.y_e[3 * (size_t)p] = v0; a.y_e[3 * (size_t)p + 1] = v1; a.y_e[3 * (size_t)p + 2] = v2; }
    if (a.model != nullptr) {
      double val = 0.0;
      for (int r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
        const double* Ei = a.E + (size_t)r * 6;
        const double* Fi = a.F + (size_t)r * 18;
        const double* x = a.x + 10 * (size_t)cam[r];
        double m0 = Ei[0] * v0 + Ei[1] * v1 + Ei[2] * v2, m1 = Ei[3] * v0 + Ei[4] * v1 + Ei[5] * v2;
        for (int k = 0; k < 9; ++k) { m0 += Fi[k] * x[k]; m1 += Fi[9 + k] * x[k]; }
        val += m0 * (a.b[2 * (size_t)r] - 0.5 * m0) + m1 * (a.b[2 * (size_t)r + 1] - 0.5 * m1);
      }
      val = BlockSum(val, sbuf);
      if (threadIdx.x == 0) atomicAdd(a.model, val);
    }
    return;
  }
  double* y = a.yrep + (size_t)(blockIdx.x % a.nrep) * a.rep_stride;
  for (int r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
    const double* Ei = a.E + (size_t)r * 6;
    const double* Fi = a.F + (size_t)r * 18;
    if (BUILD && a.weights != nullptr) {
      double g0[3], g1[3];
      for (int j = 0; j < 3; ++j) {
        g0[j] = Ei[0] * c[j] + Ei[1] * c[3 + j] + Ei[2] * c[6 + j];
        g1[j] = Ei[3] * c[j] + Ei[4] * c[3 + j] + Ei[5] * c[6 + j];
      }
      double* w = a.weights + (size_t)r * kWeightStride;
      w[0] = 1.0 - (g0[0] * Ei[0] + g0[1] * Ei[1] + g0[2] * Ei[2]);
      w[1] = -(g0[0] * Ei[3] + g0[1] * Ei[4] + g0[2] * Ei[5]);
      w[2] = 1.0 - (g1[0] * Ei[3] + g1[1] * Ei[4] + g1[2] * Ei[5]);
    }
    if (BUILD && a.b == nullptr) continue;
    double t0 = 0.0, t1 = 0.0;
    if (MODE == kModeRhs) { t0 = a.b[2 * (size_t)r]; t1 = a.b[2 * (size_t)r + 1]; }
    else {
      const double* x = a.x + 10 * (size_t)cam[r];
      for (int k = 0; k < 9; ++k) { t0 += Fi[k] * x[k]; t1 += Fi[9 + k] * x[k]; }
    }
    const double ev0 = Ei[0] * v0 + Ei[1] * v1 + Ei[2] * v2;
    const double ev1 = Ei[3] * v0 + Ei[4] * v1 + Ei[5] * v2;
    const double w0 = (MODE == kModePseZ) ? ev0 : t0 - ev0;
    const double w1 = (MODE == kModePseZ) ? ev1 : t1 - ev1;
    if (a.rows_out != nullptr) {  // deterministic mode
      a.rows_out[2 * (size_t)r] = w0; a.rows_out[2 * (size_t)r + 1] = w1;
      continue;
    }
    double* yc = y + 10 * (size_t)cam[r];
    for (int k = 0; k < 9; ++k) atomicAdd(yc + k, Fi[k] * w0 + Fi[9 + k] * w1);
  }
}

// Back substitution from the history (see StreamArgs::v_store): y_e = v_rhs - sum_{k < iter} alpha_k v_k.
// Runs behind a finished solve (done set) that stayed within kHistCap iterations; otherwise the
// kModeBack pass does the work. One thread per entry of y_e; fixed order => deterministic.
__global__ void __launch_bounds__(256)
point_combine_kernel(const CgState* __restrict__ st, const double* __restrict__ v_rhs,
                     const double* __restrict__ v_hist, size_t stride, size_t n, double* __restrict__ y_e) {
  if (!st->done || !st->hist_valid || st->iter > kHistCap) return;
  const int iters = st->iter;
  double alpha[kHistCap];
#pragma unroll
  for (int k = 0; k < kHistCap; ++k) alpha[k] = st->alpha_hist[k];
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double y = v_rhs[i];
#pragma unroll
    for (int k = 0; k < kHistCap; ++k)
      if (k < iters && alpha[k] != 0.0) y -= alpha[k] * v_hist[(size_t)k * stride + i];
    y_e[i] = y;
  }
}

// [P x 6] upper triangles -> [P x 9] full symmetric row-major blocks (seam output only)
__global__ void expand_symmetric3_kernel(const double* __restrict__ in6, int P, double* __restrict__ out9) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const double* a = in6 + (size_t)p * 6;
  double* o = out9 + (size_t)p * 9;
  o[0] = a[0]; o[1] = a[1]; o[2] = a[2];
  o[3] = a[1]; o[4] = a[3]; o[5] = a[4];
  o[6] = a[2]; o[7] = a[4]; o[8] = a[5];
}

}  // namespace cb200
#endif  // CB200_STREAM_KERNELS_CUH_

// kernels_tile.cuh -- pass B of the normal-equation build, fast path: Schur elimination of a TILE of
// tracks whose free poses fall inside one window of consecutive pose blocks.
//
// The host sorts tracks by their first pose block and cuts the order into tiles (solver.cuh).  Inside a
// tile every track's local system lives in the same small column space
//     [ 7 * Wt window pose columns | B border columns ]   (Lt columns, padded to L8 = multiple of 8),
// so the tile's whole contribution  - sum_tracks (G Hinv) G'  is a dense symmetric Lt x Lt matrix.  It is
// kept in REGISTERS, spread over the CTA as 4x4 (or 8x4) patches (thread = patch), and goes to the reduced system
// with one atomicAdd per entry per tile instead of one per entry per track.  Per round of <= K tracks:
//   phase b1  one THREAD per residual block: residual, analytic dr/dXc (A) and dr/dX (E), Cauchy
//             correction, then the three 3x3 matrices  M = (sr A)'E, N = [Z]x M, Nrc = [Xc - t_rc]x M.
//             Every pose / rig / offset column of F'E is  coef' {M | N | Nrc}  with a per-image
//             3-vector coef (CamCoef, common.cuh), so no Jacobian column is formed per observation.
//   phase b2  one warp per track: E'E, E'r over its blocks, Hinv, G = F'E into shared memory
//             (dense-in-window rows, zero where the track has no observation), Y = G Hinv
//   phase c   all threads: rank-3K update of the register patches from the staged tracks
//             (24 LDS per 48 DFMA), plus rhs -= G Hinv E'r
//   flush     once per tile
// Tracks that do not fit a window go through the generic kernel (kernels_build.cuh).
#pragma once
#include "kernels_backsub.cuh"

namespace mvba {

struct TileArgs {
  BuildArgs b;
  const CamCoef* coef;
  const int32_t* tile_trk0;   // ntiles + 1: range into trk_order
  const int32_t* tile_w0;     // first pose block of the window
  const int32_t* tile_wn;     // number of pose blocks in the window
  const int32_t* trk_order;   // track ids, tile-major
  int ntiles;
  int K;                      // tracks staged per round (<= 32)
  int icap;                   // residual blocks staged per round
  int nx;                     // max free non-pose columns of one block (focal, centre, distortion, depth)
  int L8max;                  // padded local width the shared buffers are sized for
};


// What phase b1 leaves per residual block.  lb / le / lr / lo: LOCAL column of the first beg-pose, end-pose and
// rig-transform column and of the offset column in the tile's column space (-1: constant or absent).
struct TRec {
  double M[9], N[9], Nrc[9], H[6], g[3];
  int cid, lb, le, lr, lo, nx, fcols, pad;  // nx: staged non-pose entries; fcols: 0 for the landmark priors
};
static_assert(sizeof(TRec) == 320, "TRec layout");

struct TTrack { int p, start[5], count[5], q0; };  // CSR ranges of the five block kinds, first staged slot

// Shared memory: Gs[K][L8max][3] | Ys[K][L8max][3] | ves[K][4] | Xs[icap][4 nx] | rec[icap] | trk[K+1]
__host__ __device__ inline size_t tile_smem(int K, int icap, int nx, int L8max) {
  return ((size_t)K * L8max * 6 + (size_t)K * 4 + (size_t)icap * 4 * nx) * 8 + (size_t)icap * sizeof(TRec) + (size_t)(K + 1) * sizeof(TTrack) + 16;
}

// Phase b1: one thread evaluates residual block `j` of staged track `tk`.  Nothing here is Jacobi-scaled on
// the f side (tile_fold scales whole columns once); E carries the landmark scaling.
__device__ inline double tile_eval(const View& v, const BuildArgs& a, const Reduced& S, const TTrack& tk, int j, int w0, int Wt,
                                   TRec& rec, double* X, double* jd) {
  Item it;
  const int p = tk.p;
  int kind = 0;
  while (j >= tk.count[kind]) { j -= tk.count[kind]; kind++; }
  const int64_t idx = (int64_t)tk.start[kind] + j;
  if (kind == 0) eval_cam<true>(v, a.x, a.tab, idx, it, jd, jd + MVBA_MAXDZ);
  else if (kind == 1) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, idx, it);
  else if (kind == 2) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, idx, it);
  else eval_prior(v, a.x, idx, it);
  double s = 0;
  for (int k = 0; k < it.nr; k++) s += it.r[k] * it.r[k];
  double hr, sr;
  cauchy(it.loss_a, s, &hr, &sr);
  double se[3] = {1, 1, 1};
  if (a.jscale_e) { se[0] = a.jscale_e[3 * (size_t)p]; se[1] = a.jscale_e[3 * (size_t)p + 1]; se[2] = a.jscale_e[3 * (size_t)p + 2]; }
  double Es[9];
#pragma unroll
  for (int k = 0; k < 9; k++) Es[k] = (k / 3 < it.nr) ? sr * it.E[k] * se[k % 3] : 0.0;
  double H[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const double* e = Es + 3 * r;
    const double rs = r < it.nr ? sr * it.r[r] : 0.0;
    H[0] += e[0] * e[0]; H[1] += e[0] * e[1]; H[2] += e[0] * e[2]; H[3] += e[1] * e[1]; H[4] += e[1] * e[2]; H[5] += e[2] * e[2];
    g[0] += e[0] * rs; g[1] += e[1] * rs; g[2] += e[2] * rs;
  }
#pragma unroll
  for (int k = 0; k < 6; k++) rec.H[k] = H[k];
#pragma unroll
  for (int k = 0; k < 3; k++) rec.g[k] = g[k];
  rec.cid = it.cid; rec.lb = -1; rec.le = -1; rec.lr = -1; rec.lo = -1; rec.nx = 0; rec.fcols = 0;
  if (kind >= 3) return hr;  // priors: no f columns
  rec.fcols = 1;
  const SensorDev& sd = v.sens[it.sensor];
  {
    const int p0 = v.cam_p0[it.cid], p1 = v.cam_p1[it.cid];
    const int cb = v.pose_col[p0], ce = p1 >= 0 ? v.pose_col[p1] : -1;
    rec.lb = cb < 0 ? -1 : cb - 7 * w0;
    rec.le = ce < 0 ? -1 : ce - 7 * w0;
    rec.lr = sd.c_rig < 0 ? -1 : 7 * Wt + (sd.c_rig - S.npc);
    rec.lo = sd.c_off < 0 ? -1 : 7 * Wt + (sd.c_off - S.npc);
  }
  // M[a][b] = sum_r sr A[r][a] Es[r][b];  N = [Z]x M;  Nrc = [Xc - t_rc]x M
  const double* trc = a.tab[it.cid].trc;
  const double y[3] = {it.Xc[0] - __ldg(trc), it.Xc[1] - __ldg(trc + 1), it.Xc[2] - __ldg(trc + 2)};
#pragma unroll
  for (int b = 0; b < 3; b++) {
    double m[3];
#pragma unroll
    for (int aa = 0; aa < 3; aa++) {
      double t = 0;
      for (int r = 0; r < it.nr; r++) t += it.A[3 * r + aa] * Es[3 * r + b];
      m[aa] = sr * t;
    }
    rec.M[b] = m[0]; rec.M[3 + b] = m[1]; rec.M[6 + b] = m[2];
    rec.N[b] = it.Z[1] * m[2] - it.Z[2] * m[1]; rec.N[3 + b] = it.Z[2] * m[0] - it.Z[0] * m[2]; rec.N[6 + b] = it.Z[0] * m[1] - it.Z[1] * m[0];
    rec.Nrc[b] = y[1] * m[2] - y[2] * m[1]; rec.Nrc[3 + b] = y[2] * m[0] - y[0] * m[2]; rec.Nrc[6 + b] = y[0] * m[1] - y[1] * m[0];
  }
  // free non-pose columns (focal, centre, distortion | depth transform, scale): direct J'E with the target column
  if (sd.c_focal >= 0 || sd.c_center >= 0 || sd.c_dist >= 0 || sd.c_d2i >= 0 || sd.c_scale >= 0) {
    const int vw = item_vwidth(v, it);
    int n = 0;
    for (int vc = MVBA_VPOSE; vc < vw; vc++) {
      const int gc = item_gcol(v, it, vc);
      if (gc < 0) continue;
      double col[3] = {0, 0, 0};
      item_jcol(v, it, a.tab, a.dtab, vc, jd, jd + MVBA_MAXDZ, col);
#pragma unroll
      for (int b = 0; b < 3; b++) {
        double x = 0;
        for (int r = 0; r < it.nr; r++) x += col[r] * Es[3 * r + b];
        X[4 * n + b] = sr * x;
      }
      X[4 * n + 3] = (double)(7 * Wt + (gc - S.npc));
      n++;
    }
    rec.nx = n;
  }
  return hr;
}

// Phase b2: one warp folds the staged blocks of one track into G (dense-in-window rows), Hinv, Y.
// Lane = coefficient row of CamCoef (23 rows); blocks are walked four at a time so that the
// coefficient loads (L2) of a group are in flight together.
__device__ inline void tile_fold(const TileArgs& ta, const Reduced& S, const TTrack& tk, const TRec* rec, const double* Xs,
                                 int w0, int Wt, int Lt, int L8, double* G, double* Y, double* ve_out, int lane) {
  const BuildArgs& a = ta.b;
  const int p = tk.p;
  const int n = tk.count[0] + tk.count[1] + tk.count[2] + tk.count[3] + tk.count[4];
  for (int li = lane; li < 3 * L8; li += 32) G[li] = 0;
  double H[6] = {0, 0, 0, 0, 0, 0}, gE[3] = {0, 0, 0};
  for (int i = lane; i < n; i += 32) {
    const TRec& r = rec[tk.q0 + i];
#pragma unroll
    for (int k = 0; k < 6; k++) H[k] += r.H[k];
#pragma unroll
    for (int k = 0; k < 3; k++) gE[k] += r.g[k];
  }
#pragma unroll
  for (int k = 0; k < 6; k++) H[k] = warp_sum(H[k]);
#pragma unroll
  for (int k = 0; k < 3; k++) gE[k] = warp_sum(gE[k]);
  __syncwarp();
  const int row = lane;
  const int msel = (row >= 17 && row <= 20) ? 2 : ((row >= 3 && row <= 6) || (row >= 10 && row <= 13) || row == 21) ? 1 : 0;
  const int role = row < 7 ? 0 : row < 14 ? 1 : row < 21 ? 2 : 3;
  const int roff = row < 7 ? row : row < 14 ? row - 7 : row < 21 ? row - 14 : 0;
  for (int i0 = 0; i0 < n; i0 += 4) {
    int li[4];
    double c[4][3];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      li[u] = -1;
      if (i0 + u < n && row < 23) {
        const TRec& r = rec[tk.q0 + i0 + u];
        const int base = role == 0 ? r.lb : role == 1 ? r.le : role == 2 ? r.lr : r.lo;
        if (r.fcols && base >= 0) {
          li[u] = base + roff;
          const double* cf = ta.coef[r.cid].c[row];
          c[u][0] = __ldg(cf); c[u][1] = __ldg(cf + 1); c[u][2] = __ldg(cf + 2);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (i0 + u >= n) break;
      const TRec& r = rec[tk.q0 + i0 + u];
      double out[3] = {0, 0, 0};
      if (li[u] >= 0) {
        const double* m = msel == 0 ? r.M : msel == 1 ? r.N : r.Nrc;
#pragma unroll
        for (int b = 0; b < 3; b++) out[b] = c[u][0] * m[b] + c[u][1] * m[3 + b] + c[u][2] * m[6 + b];
      }
      // the offset column has two rows (21: wa' N, 22: dt' M)
      if (r.lo >= 0) {
#pragma unroll
        for (int b = 0; b < 3; b++) { const double o22 = __shfl_sync(0xffffffffu, out[b], 22); if (row == 21) out[b] += o22; }
      }
      if (row < 22 && li[u] >= 0) { G[3 * li[u]] += out[0]; G[3 * li[u] + 1] += out[1]; G[3 * li[u] + 2] += out[2]; }
      if (r.nx > 0) {  // non-pose columns staged by phase b1 (distinct columns, disjoint from the pose part)
        const double* X = Xs + (size_t)(tk.q0 + i0 + u) * 4 * ta.nx;
        for (int k = lane; k < r.nx; k += 32) {
          const int l2 = (int)X[4 * k + 3];
          G[3 * l2] += X[4 * k]; G[3 * l2 + 1] += X[4 * k + 1]; G[3 * l2 + 2] += X[4 * k + 2];
        }
      }
      __syncwarp();
    }
  }
  H[0] += fmin(fmax(H[0], a.min_diag), a.max_diag) / a.radius;
  H[3] += fmin(fmax(H[3], a.min_diag), a.max_diag) / a.radius;
  H[5] += fmin(fmax(H[5], a.min_diag), a.max_diag) / a.radius;
  double Hi[6];
  inv_sym3(H, Hi);
  if (lane < 6) a.Hinv[6 * (size_t)p + lane] = Hi[lane];
  if (lane < 3) a.ge[3 * (size_t)p + lane] = gE[lane];
  if (lane == 0) {
    ve_out[0] = Hi[0] * gE[0] + Hi[1] * gE[1] + Hi[2] * gE[2];
    ve_out[1] = Hi[1] * gE[0] + Hi[3] * gE[1] + Hi[4] * gE[2];
    ve_out[2] = Hi[2] * gE[0] + Hi[4] * gE[1] + Hi[5] * gE[2];
  }
  // Jacobi column scaling once per local column, then Y = G Hinv
  for (int l = lane; l < L8; l += 32) {
    double sc = 1.0;
    if (a.jscale_f && l < Lt) sc = a.jscale_f[l < 7 * Wt ? 7 * w0 + l : S.npc + (l - 7 * Wt)];
    const double g0 = G[3 * l] * sc, g1 = G[3 * l + 1] * sc, g2 = G[3 * l + 2] * sc;
    G[3 * l] = g0; G[3 * l + 1] = g1; G[3 * l + 2] = g2;
    Y[3 * l] = Hi[0] * g0 + Hi[1] * g1 + Hi[2] * g2;
    Y[3 * l + 1] = Hi[1] * g0 + Hi[3] * g1 + Hi[4] * g2;
    Y[3 * l + 2] = Hi[2] * g0 + Hi[4] * g1 + Hi[5] * g2;
  }
}

// NT threads; register patches of PR x 4 (PR = 4 or 8), one per thread, on or above the diagonal.
template <int NT, int PR>
__global__ void __launch_bounds__(NT, 1) tile_kernel(View v, TileArgs a, Reduced S) {
  extern __shared__ double smem_d[];
  constexpr int NW = NT / 32, STEP = PR / 4;
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  const int K = a.K;
  double* Gs = smem_d;
  double* Ys = Gs + (size_t)K * a.L8max * 3;
  double* ves = Ys + (size_t)K * a.L8max * 3;
  double* Xs = ves + (size_t)K * 4;
  TRec* rec = reinterpret_cast<TRec*>(Xs + (size_t)a.icap * 4 * a.nx);
  TTrack* trk = reinterpret_cast<TTrack*>(rec + a.icap);
  __shared__ int s_kc, s_nit;
  const int B = S.B;
  double cost_acc = 0;
  double jd[2 * MVBA_MAXDZ];

  for (int t = blockIdx.x; t < a.ntiles; t += gridDim.x) {
    const int t0 = a.tile_trk0[t], t1 = a.tile_trk0[t + 1];
    const int w0 = a.tile_w0[t], Wt = a.tile_wn[t];
    const int Lt = 7 * Wt + B, L8 = (Lt + 7) & ~7;
    const int nb4 = L8 >> 2, nbr = L8 / PR;
    const int npatch = nbr * nb4 - STEP * (nbr * (nbr - 1) / 2);  // sum_i (nb4 - STEP i)
    const bool worker = tid < npatch;
    int bi = 0, bj = 0;
    if (worker) {
      int rem = tid;
      while (rem >= nb4 - STEP * bi) { rem -= nb4 - STEP * bi; bi++; }
      bj = STEP * bi + rem;
    }
    double acc[PR][4];
#pragma unroll
    for (int i = 0; i < PR; i++)
#pragma unroll
      for (int j = 0; j < 4; j++) acc[i][j] = 0;
    double rhs_acc = 0;

    int r0 = t0;
    while (r0 < t1) {
      // ---- round set-up: as many tracks as fit K slots and icap residual blocks (warp 0) ----
      if (wib == 0) {
        int cnt = 0;
        TTrack tk{};
        if (lane < K && r0 + lane < t1) {
          const int p = a.trk_order[r0 + lane];
          tk.p = p;
          tk.start[0] = (int)v.trk_obs0[p]; tk.count[0] = (int)(v.trk_obs0[p + 1] - v.trk_obs0[p]);
          tk.start[1] = (int)v.trk_dtri0[p]; tk.count[1] = (int)(v.trk_dtri0[p + 1] - v.trk_dtri0[p]);
          tk.start[2] = (int)v.trk_dmesh0[p]; tk.count[2] = (int)(v.trk_dmesh0[p + 1] - v.trk_dmesh0[p]);
          tk.start[3] = (int)v.trk_pm0[p]; tk.count[3] = (int)(v.trk_pm0[p + 1] - v.trk_pm0[p]);
          tk.start[4] = (int)v.trk_pt0[p]; tk.count[4] = (int)(v.trk_pt0[p + 1] - v.trk_pt0[p]);
          cnt = tk.count[0] + tk.count[1] + tk.count[2] + tk.count[3] + tk.count[4];
        }
        int incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
        const bool fits = cnt > 0 && incl <= a.icap;
        const unsigned mk = __ballot_sync(0xffffffffu, fits);
        const int kc = mk == 0xffffffffu ? 32 : __ffs(~mk) - 1;  // in order: the first track that does not fit ends the round
        tk.q0 = incl - cnt;
        if (lane < kc) trk[lane] = tk;
        if (lane == kc - 1) s_nit = incl;
        if (lane == 0) s_kc = kc;
      }
      __syncthreads();
      const int kc = s_kc, nit = s_nit;
      // ---- phase b1: one thread per residual block ----
      for (int q = tid; q < nit; q += NT) {
        int k = 0;
        while (k + 1 < kc && trk[k + 1].q0 <= q) k++;
        cost_acc += tile_eval(v, a.b, S, trk[k], q - trk[k].q0, w0, Wt, rec[q], Xs + (size_t)q * 4 * a.nx, jd);
      }
      __syncthreads();
      // ---- phase b2: one warp per track ----
      for (int k = wib; k < kc; k += NW)
        tile_fold(a, S, trk[k], rec, Xs, w0, Wt, Lt, L8, Gs + (size_t)k * a.L8max * 3, Ys + (size_t)k * a.L8max * 3, ves + 4 * k, lane);
      __syncthreads();
      // ---- phase c: register patches += Y G' over the staged tracks ----
      if (worker) {
        for (int k = 0; k < kc; k++) {
          const double2* yp = reinterpret_cast<const double2*>(Ys + ((size_t)k * a.L8max + PR * bi) * 3);
          const double2* gp = reinterpret_cast<const double2*>(Gs + ((size_t)k * a.L8max + 4 * bj) * 3);
          double y[3 * PR], g[12];
#pragma unroll
          for (int i = 0; i < 3 * PR / 2; i++) { double2 tmp = yp[i]; y[2 * i] = tmp.x; y[2 * i + 1] = tmp.y; }
#pragma unroll
          for (int i = 0; i < 6; i++) { double2 tmp = gp[i]; g[2 * i] = tmp.x; g[2 * i + 1] = tmp.y; }
#pragma unroll
          for (int i = 0; i < PR; i++)
#pragma unroll
            for (int j = 0; j < 4; j++)
              acc[i][j] = fma(y[3 * i], g[3 * j], fma(y[3 * i + 1], g[3 * j + 1], fma(y[3 * i + 2], g[3 * j + 2], acc[i][j])));
        }
      }
      if (tid < L8) {
        for (int k = 0; k < kc; k++) {
          const double* gp = Gs + ((size_t)k * a.L8max + tid) * 3;
          const double* ve = ves + 4 * k;
          rhs_acc = fma(gp[0], ve[0], fma(gp[1], ve[1], fma(gp[2], ve[2], rhs_acc)));
        }
      }
      __syncthreads();
      r0 += kc;
    }
    // ---- flush: local column -> reduced column; window poses are consecutive pose blocks ----
    if (worker) {
#pragma unroll
      for (int i = 0; i < PR; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int li = PR * bi + i, lj = 4 * bj + j;
          const double val = acc[i][j];
          if (li > lj || lj >= Lt || val == 0.0) continue;
          const int gi = li < 7 * Wt ? 7 * w0 + li : S.npc + (li - 7 * Wt);
          const int gj = lj < 7 * Wt ? 7 * w0 + lj : S.npc + (lj - 7 * Wt);
          addS(S, gi, gj, -val);
        }
    }
    if (tid < Lt && rhs_acc != 0.0) {
      const int gi = tid < 7 * Wt ? 7 * w0 + tid : S.npc + (tid - 7 * Wt);
      atomicAdd(S.rhs + gi, -rhs_acc);
    }
  }
  cost_acc = warp_sum(cost_acc);
  if (lane == 0 && cost_acc != 0.0) atomicAdd(a.b.cost, cost_acc);
}

}  // namespace mvba

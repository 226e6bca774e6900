// kernels_tile.cuh -- pass B of the normal-equation build, fast path: Schur elimination of a TILE of
// tracks whose free poses fall inside one window of consecutive pose blocks.
//
// The host sorts tracks by their first pose block and cuts the order into tiles, tiles into ROUNDS of <= K
// tracks / <= icap residual blocks, and writes STATIC tables for every round (solver.cuh): which residual
// block each staged slot evaluates, and for every (track, touched parameter block) GROUP the list of staged
// slots that contribute to it.  Inside a tile every track's local system lives in the same small column space
//     [ 7 * Wt window pose columns | B border columns ]   (Lt columns, padded to L8 = multiple of 8),
// so the tile's whole contribution  - sum_tracks (G Hinv) G'  is a dense symmetric Lt x Lt matrix.  It is
// kept in REGISTERS, spread over the CTA as 4x4 (or 8x4) patches (thread = patch), and goes to the reduced system
// with one atomicAdd per entry per tile instead of one per entry per track.  Per round:
//   phase b1  one THREAD per residual block: residual, analytic dr/dXc (A) and dr/dX (E), Cauchy
//             correction, M = (sr A)'E, and the block's share of E'E, E'r.
//   phase b2  one THREAD per (track, group, row): row of G = F'E as a GATHER over the group's contributors:
//             every pose / rig / offset row of one block's F'E is  d' M  with  d = coef, coef x Z or
//             coef x (Xc - t_rc)  for a per-image 3-vector coef (CamCoef, common.cuh), so no Jacobian column is
//             formed per observation, nothing is accumulated with atomics and no warp waits on another.
//             Threads k < kc meanwhile sum E'E, E'r of track k and invert.
//   phase b3  all threads: Y = G Hinv (dense-in-window rows)
//   phase c   all threads: rank-3K update of the register patches from the staged tracks
//             (24 LDS per 48 DFMA), plus rhs -= G Hinv E'r
//   flush     once per tile
// Tracks that do not fit a window go through the generic kernel (kernels_build.cuh).
#pragma once
#include "kernels_backsub.cuh"

namespace mvba {

// Static round tables (host-built, solver.cuh).  A slot is the position of a residual block in the round's staging
// buffer; a track's blocks occupy consecutive slots in the order CAM, DEPTH_TRI, DEPTH_MESH, MESH_TRI prior, TRI prior.
struct TBlk { int32_t idx; int32_t kind; };                       // index inside its kind's array
struct TTrk { int32_t p; int16_t q0, n; };                        // landmark, first slot, number of slots
struct TGrp { uint32_t con0; uint16_t lcol; uint8_t k, ncon; };   // contributors, first local column, track slot; ncon bit 7: offset column (1 row)
static_assert(sizeof(TBlk) == 8 && sizeof(TTrk) == 8 && sizeof(TGrp) == 8, "tile table layout");
// contributor: (slot << 2) | role, role 0 beg pose (coef rows 0-6), 1 end pose (7-13), 2 rig transform (14-20), 3 offset (21, 22)

struct TileArgs {
  BuildArgs b;
  const CamCoef* coef;
  const int32_t* tile_w0;     // first pose block of the window
  const int32_t* tile_wn;     // number of pose blocks in the window
  const int32_t* tile_round0; // ntiles + 1: range of rounds
  const int32_t* round_trk0;  // nrounds + 1: range into ttrk
  const int32_t* round_blk0;  // nrounds + 1: range into blk
  const int32_t* round_grp0;  // nrounds + 1: range into grp
  const TBlk* blk;
  const TTrk* ttrk;
  const TGrp* grp;
  const uint16_t* con;
  int ntiles;
  int K;                      // tracks staged per round (<= 32)
  int icap;                   // residual blocks staged per round
  int nx;                     // max free non-pose columns of one block (focal, centre, distortion, depth)
  int L8max;                  // padded local width the shared buffers are sized for
};

// What phase b1 leaves per residual block.
struct TRec {
  double M[9], Z[3], y[3], H[6], g[3];
  int cid, nx;  // cid < 0: no f columns (landmark priors); nx: staged non-pose entries
};
static_assert(sizeof(TRec) == 200, "TRec layout");

// Shared memory: Gs[K][L8max][3] | Ys[K][L8max][3] | ves[K][4] | His[K][6] | cs[L8max] | Xs[icap][4 nx] | rec[icap]
__host__ __device__ inline size_t tile_smem(int K, int icap, int nx, int L8max) {
  return ((size_t)K * L8max * 6 + (size_t)K * 10 + (size_t)L8max + (size_t)icap * 4 * nx) * 8 + (size_t)icap * sizeof(TRec) + 16;
}

// Phase b1: one thread evaluates one residual block.  Nothing here is Jacobi-scaled on the f side (phase b2 scales
// whole columns once); E carries the landmark scaling.
__device__ inline double tile_eval(const View& v, const BuildArgs& a, const Reduced& S, const TBlk bk, int Wt,
                                   TRec& rec, double* X, double* jd) {
  Item it;
  const int kind = bk.kind;
  const int64_t idx = bk.idx;
  if (kind == 0) eval_cam<true>(v, a.x, a.tab, idx, it, jd, jd + MVBA_MAXDZ);
  else if (kind == 1) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, idx, it);
  else if (kind == 2) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, idx, it);
  else eval_prior(v, a.x, idx, it);
  const int p = it.pid;
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
  rec.cid = -1; rec.nx = 0;
  if (kind >= 3) return hr;  // priors: no f columns
  rec.cid = it.cid;
  const SensorDev& sd = v.sens[it.sensor];
  // M[a][b] = sum_r sr A[r][a] Es[r][b];  the rotation rows use  coef'[Z]x M = (coef x Z)' M,  so Z and y = Xc - t_rc are kept
  const double* trc = a.tab[it.cid].trc;
#pragma unroll
  for (int k = 0; k < 3; k++) { rec.Z[k] = it.Z[k]; rec.y[k] = it.Xc[k] - __ldg(trc + k); }
#pragma unroll
  for (int b = 0; b < 3; b++)
#pragma unroll
    for (int aa = 0; aa < 3; aa++) {
      double t = 0;
      for (int r = 0; r < it.nr; r++) t += it.A[3 * r + aa] * Es[3 * r + b];
      rec.M[3 * aa + b] = sr * t;
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

// Phase b2: row `row` of group `gr` of G = F'E for one staged track, gathered over the group's contributors and
// Jacobi-scaled.  G points at the track's dense-in-window rows.
__device__ __forceinline__ void tile_gather(const TileArgs& ta, const TGrp gr, int row, const TRec* __restrict__ rec,
                                            const double* __restrict__ cs, double* __restrict__ G) {
  const bool single = (gr.ncon & 0x80) != 0;
  const int ncon = gr.ncon & 0x7f;
  if (row >= (single ? 1 : 7)) return;
  double o0 = 0, o1 = 0, o2 = 0;
  for (int c = 0; c < ncon; c++) {
    const unsigned cb = __ldg(ta.con + gr.con0 + c);
    const TRec& r = rec[cb >> 2];
    const int role = cb & 3;
    double d0, d1, d2;
    if (role < 3) {
      const double* cf = ta.coef[r.cid].c[7 * role + row];
      const double c0 = __ldg(cf), c1 = __ldg(cf + 1), c2 = __ldg(cf + 2);
      if (row < 3) { d0 = c0; d1 = c1; d2 = c2; }  // translations: coef' M
      else {                                        // rotations: coef' [w]x M = (coef x w)' M, w = Z or Xc - t_rc
        const double* w = role == 2 ? r.y : r.Z;
        d0 = c1 * w[2] - c2 * w[1]; d1 = c2 * w[0] - c0 * w[2]; d2 = c0 * w[1] - c1 * w[0];
      }
    } else {  // time offset: dalpha (wa x Z + dt)
      const double* cf = ta.coef[r.cid].c[21];
      const double c0 = __ldg(cf), c1 = __ldg(cf + 1), c2 = __ldg(cf + 2);
      d0 = c1 * r.Z[2] - c2 * r.Z[1] + __ldg(cf + 3); d1 = c2 * r.Z[0] - c0 * r.Z[2] + __ldg(cf + 4); d2 = c0 * r.Z[1] - c1 * r.Z[0] + __ldg(cf + 5);
    }
    o0 = fma(d0, r.M[0], fma(d1, r.M[3], fma(d2, r.M[6], o0)));
    o1 = fma(d0, r.M[1], fma(d1, r.M[4], fma(d2, r.M[7], o1)));
    o2 = fma(d0, r.M[2], fma(d1, r.M[5], fma(d2, r.M[8], o2)));
  }
  const int l = gr.lcol + row;
  const double sc = cs[l];
  G[3 * l] = o0 * sc; G[3 * l + 1] = o1 * sc; G[3 * l + 2] = o2 * sc;
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
  double* His = ves + (size_t)K * 4;
  double* cs = His + (size_t)K * 6;
  double* Xs = cs + a.L8max;
  TRec* rec = reinterpret_cast<TRec*>(Xs + (size_t)a.icap * 4 * a.nx);
  const int B = S.B;
  const size_t gstride = (size_t)a.L8max * 3;
  double cost_acc = 0;
  double jd[2 * MVBA_MAXDZ];

  for (int t = blockIdx.x; t < a.ntiles; t += gridDim.x) {
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
    // Jacobi scale of every local column (the previous tile's readers are behind its last barrier)
    for (int l = tid; l < a.L8max; l += NT)
      cs[l] = (a.b.jscale_f && l < Lt) ? a.b.jscale_f[l < 7 * Wt ? 7 * w0 + l : S.npc + (l - 7 * Wt)] : 1.0;

    const int rd0 = a.tile_round0[t], rd1 = a.tile_round0[t + 1];
    for (int rd = rd0; rd < rd1; rd++) {
      const int trk0 = a.round_trk0[rd], kc = a.round_trk0[rd + 1] - trk0;
      const int blk0 = a.round_blk0[rd], nit = a.round_blk0[rd + 1] - blk0;
      const int grp0 = a.round_grp0[rd], ngrp = a.round_grp0[rd + 1] - grp0;
      // ---- phase b1: one thread per residual block; everyone clears G ----
      for (int q = tid; q < nit; q += NT)
        cost_acc += tile_eval(v, a.b, S, a.blk[blk0 + q], Wt, rec[q], Xs + (size_t)q * 4 * a.nx, jd);
      {
        double2* gz = reinterpret_cast<double2*>(Gs);
        const int nz = (int)((size_t)kc * gstride / 2);
        for (int i = tid; i < nz; i += NT) gz[i] = make_double2(0.0, 0.0);
      }
      __syncthreads();
      // ---- phase b2: one thread per (group, row) gathers a row of G; threads k < kc reduce and invert E'E ----
      for (int i = tid; i < 8 * ngrp; i += NT) {
        const TGrp gr = a.grp[grp0 + (i >> 3)];
        tile_gather(a, gr, i & 7, rec, cs, Gs + (size_t)gr.k * gstride);
      }
      if (tid < kc) {
        const TTrk tk = a.ttrk[trk0 + tid];
        double H[6] = {0, 0, 0, 0, 0, 0}, gE[3] = {0, 0, 0};
        for (int i = 0; i < tk.n; i++) {
          const TRec& r = rec[tk.q0 + i];
#pragma unroll
          for (int k = 0; k < 6; k++) H[k] += r.H[k];
#pragma unroll
          for (int k = 0; k < 3; k++) gE[k] += r.g[k];
        }
        H[0] += fmin(fmax(H[0], a.b.min_diag), a.b.max_diag) / a.b.radius;
        H[3] += fmin(fmax(H[3], a.b.min_diag), a.b.max_diag) / a.b.radius;
        H[5] += fmin(fmax(H[5], a.b.min_diag), a.b.max_diag) / a.b.radius;
        double Hi[6];
        inv_sym3(H, Hi);
#pragma unroll
        for (int k = 0; k < 6; k++) { a.b.Hinv[6 * (size_t)tk.p + k] = Hi[k]; His[6 * tid + k] = Hi[k]; }
#pragma unroll
        for (int k = 0; k < 3; k++) a.b.ge[3 * (size_t)tk.p + k] = gE[k];
        ves[4 * tid] = Hi[0] * gE[0] + Hi[1] * gE[1] + Hi[2] * gE[2];
        ves[4 * tid + 1] = Hi[1] * gE[0] + Hi[3] * gE[1] + Hi[4] * gE[2];
        ves[4 * tid + 2] = Hi[2] * gE[0] + Hi[4] * gE[1] + Hi[5] * gE[2];
      }
      if (a.nx > 0) {  // free intrinsics / depth columns staged by phase b1: a warp per track, disjoint from the gathered columns
        for (int k = wib; k < kc; k += NW) {
          const TTrk tk = a.ttrk[trk0 + k];
          double* G = Gs + (size_t)k * gstride;
          for (int i = 0; i < tk.n; i++) {
            const int nxi = rec[tk.q0 + i].nx;
            const double* X = Xs + (size_t)(tk.q0 + i) * 4 * a.nx;
            for (int kk = lane; kk < nxi; kk += 32) {
              const int l2 = (int)X[4 * kk + 3];
              const double sc = cs[l2];
              G[3 * l2] += X[4 * kk] * sc; G[3 * l2 + 1] += X[4 * kk + 1] * sc; G[3 * l2 + 2] += X[4 * kk + 2] * sc;
            }
            __syncwarp();
          }
        }
      }
      __syncthreads();
      // ---- phase b3: Y = G Hinv ----
      for (int i = tid; i < kc * L8; i += NT) {
        const int k = i / L8, l = i - k * L8;
        const double* g = Gs + (size_t)k * gstride + 3 * l;
        const double* Hi = His + 6 * k;
        double* y = Ys + (size_t)k * gstride + 3 * l;
        const double g0 = g[0], g1 = g[1], g2 = g[2];
        y[0] = Hi[0] * g0 + Hi[1] * g1 + Hi[2] * g2;
        y[1] = Hi[1] * g0 + Hi[3] * g1 + Hi[4] * g2;
        y[2] = Hi[2] * g0 + Hi[4] * g1 + Hi[5] * g2;
      }
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

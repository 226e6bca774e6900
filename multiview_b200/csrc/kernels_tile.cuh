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
// gather group: first local column, track slot, contributor count (bit 7: offset column, 1 row), the first four
// contributors inline, the rest at con[over0 ...]
struct alignas(16) TGrp { uint32_t over0; uint16_t lcol; uint8_t k, ncon; uint16_t c[4]; };
static_assert(sizeof(TBlk) == 8 && sizeof(TTrk) == 8 && sizeof(TGrp) == 16, "tile table layout");
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
  int direct;                 // free non-pose columns go straight from phase b1 into Gt (no Xs staging, one round at a time)
  int nmono;                  // direct mode: monomials of the widest floating RPC model (0: none); phase b2 forms those columns
};
#define MVBA_XR 14            // per-block RPC record of direct mode: ux uy 1/den_x 1/den_y -ox/den_x -oy/den_y  sr Es[0..5]  Jf[0..1]

// What phase b1 leaves per residual block.
struct TRec {
  double M[9], Z[3], y[3], H[6], g[3];
  int cid, nx;  // cid < 0: no f columns (landmark priors); nx: staged non-pose entries
};
static_assert(sizeof(TRec) == 200, "TRec layout");

// Row length of the K-major staging matrices: 3 K rounded up so that KK = 4 (mod 16) doubles, which makes the
// DMMA fragment loads (8 rows x 4 consecutive doubles per half warp) bank-conflict free.
__host__ __device__ inline int tile_kk(int K) { int kk = 3 * K; while ((kk & 15) != 4) kk++; return kk; }

// Shared memory: Gt[L8max][KK] | Yt[L8max][KK] | ves[KK] | His[K][6] | cs[L8max] | Xs[2 icap][4 nx] | rec[2 icap]
// (icap = residual blocks of ONE round; phase b1 evaluates two rounds at a time)
// Gt[l][3 k + b] = row l of G = F'E of staged track k (dense in the window, zero where the track has no block),
// Yt = G Hinv in the same layout: phase c is the GEMM  C += Yt Gt'  with the staged tracks as the K dimension.
// direct: no Xs and records of one round only (the free non-pose columns of a block are added to Gt by the thread that evaluates it:
// with an RPC distortion block floating, Xs was 1.2 KB per residual block and held the round to six tracks)
__host__ __device__ inline size_t tile_smem(int K, int icap, int nx, int L8max, bool direct = false) {
  const size_t KK = tile_kk(K);
  if (direct) return (2 * (size_t)L8max * KK + KK + (size_t)K * 6 + (size_t)L8max + (size_t)icap * MVBA_XR) * 8 + (size_t)icap * sizeof(TRec) + 16;
  return (2 * (size_t)L8max * KK + KK + (size_t)K * 6 + (size_t)L8max + (size_t)2 * icap * 4 * nx) * 8 + (size_t)2 * icap * sizeof(TRec) + 16;
}

// Phase b1: one thread evaluates one residual block.  Nothing here is Jacobi-scaled on the f side (phase b2 scales
// whole columns once); E carries the landmark scaling.
// HAS_X = false (no free intrinsics / depth columns anywhere): no distortion Jacobian, no dynamically indexed
// column loop, so the block's Item stays in registers.
// Gk != nullptr (direct mode): the free non-pose columns are added to the track's three columns of Gt (Gk = Gt + 3 k, row stride KK)
// with shared-memory atomics, already Jacobi-scaled (cs); the column loop starts at `rot` so that the blocks of one track, evaluated
// by neighbouring lanes, are on different rows at any one time.
template <bool HAS_X>
__device__ __forceinline__ double tile_eval(const View& v, const BuildArgs& a, const Reduced& S, const TBlk bk, int Wt,
                                            TRec& rec, double* X, double* jd, double* Gk = nullptr, const double* cs = nullptr,
                                            int KK = 0, int rot = 0, double* xr = nullptr) {
  Item it;
  const int kind = bk.kind & 0xff;
  const int64_t idx = bk.idx;
  // direct mode, RPC sensor: the 38+ distortion columns are formed on the fly below, no dr/d(distortion) table (1 KB of local memory
  // per thread, which the 29 KB of L1 left beside this kernel's shared memory cannot hold)
  const bool rpc_fly = HAS_X && Gk != nullptr && kind == 0 && v.sens[v.cam_sensor[v.cid[idx]]].rpc_deg > 0;
  if (kind == 0) eval_cam<true>(v, a.x, a.tab, idx, it, HAS_X && !rpc_fly ? jd : nullptr, HAS_X && !rpc_fly ? jd + MVBA_MAXDZ : nullptr);
  else if (kind == 1) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, idx, it);
  else if (kind == 2) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, idx, it);
  else eval_prior(v, a.x, idx, it);
  const int p = it.pid;
  double s = it.r[0] * it.r[0] + it.r[1] * it.r[1];
  if (it.nr == 3) s += it.r[2] * it.r[2];
  double hr = 0, sr;
  if (a.want_cost) cauchy(it.loss_a, s, &hr, &sr);
  else sr = cauchy_sr(it.loss_a, s);
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
      double t = it.A[aa] * Es[b] + it.A[3 + aa] * Es[3 + b];  // every block kind has at least two rows
      if (it.nr == 3) t += it.A[6 + aa] * Es[6 + b];
      rec.M[3 * aa + b] = sr * t;
    }
  // free non-pose columns (focal, centre, distortion | depth transform, scale): direct J'E with the target column
  if (HAS_X && (sd.c_focal >= 0 || sd.c_center >= 0 || sd.c_dist >= 0 || sd.c_d2i >= 0 || sd.c_scale >= 0)) {
    const int vw = item_vwidth(v, it);
    int n = 0;
    if (rpc_fly) {
      // RPC sensor, direct mode: no column is formed here.  The block leaves the few numbers its focal and distortion columns are
      // made of (MVBA_XR doubles); phase b2 has one thread per (track, monomial) sum them over the track's blocks and store the
      // rows of G it alone owns -- no staging of 38 columns per block and no shared-memory atomics (a CAS loop each).
      const double* dist = a.x + sd.x_dist;
      const int deg = sd.rpc_deg;
      const RpcLayout L = rpc_layout(deg);
      const double fz = a.x[sd.x_focal] / it.Xc[2];
      const double ux = fz * it.Xc[0], uy = fz * it.Xc[1];
      double v0 = 0, v1 = 1, v2 = 0, v3 = 1;
      {
        double yi = 1.0;
        for (int i = 0; i <= deg; i++) {
          double m = yi;
          for (int aa = 0; aa + i <= deg; aa++) {
            const int o = rpc_tri(aa + i) + i;
            v0 = fma(dist[o], m, v0); v2 = fma(dist[L.o2 + o], m, v2);
            if (aa + i > 0) { v1 = fma(dist[L.o1 + o - 1], m, v1); v3 = fma(dist[L.o3 + o - 1], m, v3); }
            m *= ux;
          }
          yi *= uy;
        }
      }
      const double idx2 = 1.0 / v1, idy2 = 1.0 / v3;
      xr[0] = ux; xr[1] = uy; xr[2] = idx2; xr[3] = idy2;
      xr[4] = -(v0 * idx2) * idx2; xr[5] = -(v2 * idy2) * idy2;   // -ox / den_x, -oy / den_y
#pragma unroll
      for (int b = 0; b < 6; b++) xr[6 + b] = sr * Es[b];
      xr[12] = it.Jf[0]; xr[13] = it.Jf[1];
      n = -1;   // marks the record for phase b2
    } else if (Gk) {
      const int nv = vw - MVBA_VPOSE;
      int j = rot % nv;
      for (int i = 0; i < nv; i++) {
        const int vc = MVBA_VPOSE + j;
        if (++j == nv) j = 0;
        const int gc = item_gcol(v, it, vc);
        if (gc < 0) continue;
        double col[3] = {0, 0, 0};
        item_jcol(v, it, a.tab, a.dtab, vc, jd, jd + MVBA_MAXDZ, col);
        const int l2 = 7 * Wt + (gc - S.npc);
        const double sc = sr * cs[l2];
        double* g = Gk + (size_t)l2 * KK;
#pragma unroll
        for (int b = 0; b < 3; b++) {
          double x = col[0] * Es[b] + col[1] * Es[3 + b];
          if (it.nr == 3) x += col[2] * Es[6 + b];
          atomicAdd(g + b, sc * x);
        }
      }
    } else {
    for (int vc = MVBA_VPOSE; vc < vw; vc++) {
      const int gc = item_gcol(v, it, vc);
      if (gc < 0) continue;
      double col[3] = {0, 0, 0};
      item_jcol(v, it, a.tab, a.dtab, vc, jd, jd + MVBA_MAXDZ, col);
#pragma unroll
      for (int b = 0; b < 3; b++) {
        double x = col[0] * Es[b] + col[1] * Es[3 + b];
        if (it.nr == 3) x += col[2] * Es[6 + b];
        X[4 * n + b] = sr * x;
      }
      X[4 * n + 3] = (double)(7 * Wt + (gc - S.npc));
      n++;
    }
    }
    rec.nx = n;
  }
  return hr;
}

// One contributor's term of a gathered row: d' M with d = coef (translations), coef x Z (pose rotations),
// coef x (Xc - t_rc) (rig rotation) or dalpha (wa x Z + dt) (time offset; e = second coefficient row).
__device__ __forceinline__ void gather_term(const TRec& r, int role, int row, const double* c, const double* e, double* o) {
  double d0, d1, d2;
  if (role < 3 && row < 3) { d0 = c[0]; d1 = c[1]; d2 = c[2]; }
  else {
    const double* w = role == 2 ? r.y : r.Z;
    d0 = c[1] * w[2] - c[2] * w[1]; d1 = c[2] * w[0] - c[0] * w[2]; d2 = c[0] * w[1] - c[1] * w[0];
    if (role == 3) { d0 += e[0]; d1 += e[1]; d2 += e[2]; }
  }
  o[0] = fma(d0, r.M[0], fma(d1, r.M[3], fma(d2, r.M[6], o[0])));
  o[1] = fma(d0, r.M[1], fma(d1, r.M[4], fma(d2, r.M[7], o[1])));
  o[2] = fma(d0, r.M[2], fma(d1, r.M[5], fma(d2, r.M[8], o[2])));
}

// Phase b2: row `row` of group `gr` of G = F'E for one staged track, gathered over the group's contributors and
// Jacobi-scaled.  G points at the track's three columns of Gt (row stride KK).  The coefficient loads (L2) of the four inline
// contributors are issued together.
// (The group record is decoded from its raw 16 bytes with shifts: a TGrp object indexed by `u` would live in local memory.)
__device__ __forceinline__ unsigned grp_con(const int4& raw, int u) {
  const unsigned w = u < 2 ? (unsigned)raw.z : (unsigned)raw.w;
  return (w >> (16 * (u & 1))) & 0xffffu;
}
__device__ __forceinline__ void tile_gather(const TileArgs& ta, const int4 raw, int row, const TRec* __restrict__ rec,
                                            const double* __restrict__ cs, double* __restrict__ Gt, int KK) {
  const unsigned over0 = (unsigned)raw.x;
  const int lcol = raw.y & 0xffff, ktrk = (raw.y >> 16) & 0xff, nc = ((unsigned)raw.y >> 24);
  double* G = Gt + 3 * ktrk;
  const bool single = (nc & 0x80) != 0;
  const int ncon = nc & 0x7f;
  if (row >= (single ? 1 : 7)) return;
  double o[3] = {0, 0, 0};
  int c = 0;
  if (!single) {
    double cf[4][3];
    int role[4];
    const TRec* rp[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const unsigned cb = grp_con(raw, u);
      rp[u] = rec + (cb >> 2); role[u] = cb & 3;
      if (u < ncon) {
        const double* p = ta.coef[rp[u]->cid].c[7 * role[u] + row];
        cf[u][0] = __ldg(p); cf[u][1] = __ldg(p + 1); cf[u][2] = __ldg(p + 2);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (u < ncon) gather_term(*rp[u], role[u], row, cf[u], cf[u], o);
    c = 4;
  }
  for (; c < ncon; c++) {
    const unsigned cb = c < 4 ? grp_con(raw, c) : __ldg(ta.con + over0 + (c - 4));
    const TRec& r = rec[cb >> 2];
    const int rl = cb & 3;
    const double* p = ta.coef[r.cid].c[rl == 3 ? 21 : 7 * rl + row];
    const double c3[3] = {__ldg(p), __ldg(p + 1), __ldg(p + 2)};
    double e3[3] = {0, 0, 0};
    if (rl == 3) { e3[0] = __ldg(p + 3); e3[1] = __ldg(p + 4); e3[2] = __ldg(p + 5); }
    gather_term(r, rl, row, c3, e3, o);
  }
  const int l = lcol + row;
  const double sc = cs[l];
  double* g = G + (size_t)l * KK;
  g[0] = o[0] * sc; g[1] = o[1] * sc; g[2] = o[2] * sc;
}

#ifdef MVBA_TILE_PROFILE
__device__ unsigned long long g_tile_prof[8];
#define TPROF(slot) do { if (tid == 0) { long long now_ = clock64(); prof_[slot] += now_ - tp_; tp_ = now_; } } while (0)
#else
#define TPROF(slot) do { } while (0)
#endif

// NT threads, one CTA per SM.  Phase c keeps the tile's upper-triangular 8x8 blocks of C in DMMA accumulator
// fragments: every warp owns up to MAXS consecutive blocks (row-major over the upper block triangle).
// (MAXS = 4, 8 or 16: instantiated per window width so that narrow windows do not pay for 64 accumulator registers.)
template <int NT, int MAXS>
__global__ void __launch_bounds__(NT, 1) tile_kernel(View v, TileArgs a, Reduced S) {
  extern __shared__ double smem_d[];
  constexpr int NW = NT / 32;
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  const int K = a.K, KK = tile_kk(K);
  double* Gt = smem_d;
  double* Yt = Gt + (size_t)a.L8max * KK;
  double* ves = Yt + (size_t)a.L8max * KK;
  double* His = ves + KK;
  double* cs = His + (size_t)K * 6;
  double* Xs = cs + a.L8max;
  TRec* rec = reinterpret_cast<TRec*>(a.direct ? Xs + (size_t)a.icap * MVBA_XR : Xs + (size_t)2 * a.icap * 4 * a.nx);
  const int B = S.B;
  const int gid = lane >> 2, tig = lane & 3;
  double cost_acc = 0;
  double jd[2 * MVBA_MAXDZ];
#ifdef MVBA_TILE_PROFILE
  long long prof_[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tp_ = clock64();
#endif
  // Yt columns beyond a round's tracks are multiplied by zeros of Gt: they must be finite
  for (int i = tid; i < a.L8max * KK + KK; i += NT) Yt[i] = 0.0;

  for (int t = blockIdx.x; t < a.ntiles; t += gridDim.x) {
    const int w0 = a.tile_w0[t], Wt = a.tile_wn[t];
    const int Lt = 7 * Wt + B, L8 = (Lt + 7) & ~7;
    const int nb8 = L8 >> 3, nblk = nb8 * (nb8 + 1) / 2;
    const int per = (nblk + NW - 1) / NW;                       // <= MAXS by the host's window cap
    const int s0 = min(nblk, wib * per), ns = min(nblk, s0 + per) - s0;
    int bi0 = 0, bj0 = 0;                                       // first block of this warp
    { int rem = s0; while (bi0 < nb8 && rem >= nb8 - bi0) { rem -= nb8 - bi0; bi0++; } bj0 = bi0 + rem; }
    double acc[MAXS][2];
#pragma unroll
    for (int i = 0; i < MAXS; i++) { acc[i][0] = 0; acc[i][1] = 0; }
    double rhs_acc = 0;
    unsigned rowend = 0;               // wide windows: the slots after which the block row changes
    if constexpr (MAXS >= 24) {
      int bi = bi0, bj = bj0;
#pragma unroll
      for (int sl = 0; sl < MAXS; sl++)
        if (sl < ns && ++bj == nb8) { bi++; bj = bi; rowend |= 1u << sl; }
    }
    // Jacobi scale of every local column (the previous tile's readers are behind its last barrier)
    for (int l = tid; l < a.L8max; l += NT)
      cs[l] = (a.b.jscale_f && l < Lt) ? a.b.jscale_f[l < 7 * Wt ? 7 * w0 + l : S.npc + (l - 7 * Wt)] : 1.0;

    const int rd0 = a.tile_round0[t], rd1 = a.tile_round0[t + 1];
    const int rstep = a.direct ? 1 : 2;
    for (int rd = rd0; rd < rd1; rd += rstep) {
      // ---- phase b1 for TWO rounds at once (one thread per residual block, so that most of the CTA evaluates);
      //      the second round's records sit icap slots further.  Everyone clears Gt.
      //      Direct mode: ONE round; Gt is cleared first, then every block adds its free non-pose columns to it. ----
      const int nsub = min(rstep, rd1 - rd);
      if (a.direct) {
        double2* gz = reinterpret_cast<double2*>(Gt);
        const int nz = L8 * KK / 2;
        for (int i = tid; i < nz; i += NT) gz[i] = make_double2(0.0, 0.0);
        __syncthreads();   // also orders cs (written at the start of the tile) before its first reader
        const int blk0 = a.round_blk0[rd], nit = a.round_blk0[rd + 1] - blk0;
        for (int q = tid; q < nit; q += NT) {
          const TBlk bk = a.blk[blk0 + q];
          cost_acc += tile_eval<true>(v, a.b, S, bk, Wt, rec[q], nullptr, jd, Gt + 3 * ((bk.kind >> 8) & 0xff), cs, KK, 5 * q,
                                      Xs + (size_t)q * MVBA_XR);
        }
      } else {
        const int blk0 = a.round_blk0[rd], nit0 = a.round_blk0[rd + 1] - blk0, nit = a.round_blk0[rd + nsub] - blk0;
        if (a.nx > 0) {
          for (int q = tid; q < nit; q += NT) {
            const int slot = q < nit0 ? q : a.icap + (q - nit0);
            cost_acc += tile_eval<true>(v, a.b, S, a.blk[blk0 + q], Wt, rec[slot], Xs + (size_t)slot * 4 * a.nx, jd);
          }
        } else {
          for (int q = tid; q < nit; q += NT) {
            const int slot = q < nit0 ? q : a.icap + (q - nit0);
            cost_acc += tile_eval<false>(v, a.b, S, a.blk[blk0 + q], Wt, rec[slot], nullptr, nullptr);
          }
        }
      }
      if (!a.direct) {
        double2* gz = reinterpret_cast<double2*>(Gt);
        const int nz = L8 * KK / 2;
        for (int i = tid; i < nz; i += NT) gz[i] = make_double2(0.0, 0.0);
      }
      TPROF(0);
      __syncthreads();
      TPROF(1);
      for (int sub = 0; sub < nsub; sub++) {
      const int trk0 = a.round_trk0[rd + sub], kc = a.round_trk0[rd + sub + 1] - trk0;
      const int grp0 = a.round_grp0[rd + sub], ngrp = a.round_grp0[rd + sub + 1] - grp0;
      const TRec* rec_s = rec + (size_t)sub * a.icap;
      const double* Xs_s = Xs + (size_t)sub * a.icap * 4 * a.nx;
      if (sub > 0) {  // Gt of the first round was read by its phase c (behind that phase's closing barrier)
        double2* gz = reinterpret_cast<double2*>(Gt);
        const int nz = L8 * KK / 2;
        for (int i = tid; i < nz; i += NT) gz[i] = make_double2(0.0, 0.0);
        __syncthreads();
      }
      // ---- phase b2: one thread per (group, row) gathers a row of G; threads k < kc reduce and invert E'E ----
      {
        const int4* gtab = reinterpret_cast<const int4*>(a.grp + grp0);
        int i = tid;
        int4 raw = i < 8 * ngrp ? __ldg(gtab + (i >> 3)) : make_int4(0, 0, 0, 0);
        while (i < 8 * ngrp) {
          const int inext = i + NT;
          const int4 rawn = inext < 8 * ngrp ? __ldg(gtab + (inext >> 3)) : make_int4(0, 0, 0, 0);
          tile_gather(a, raw, i & 7, rec_s, cs, Gt, KK);
          raw = rawn; i = inext;
        }
      }
      if (a.direct && a.nmono > 0) {
        // focal and RPC distortion rows of G: one thread per (track, monomial | focal) sums over the track's RPC blocks from their
        // MVBA_XR records and adds to rows of Gt that no other thread touches (runs of blocks of one sensor are summed in registers)
        const int per = a.nmono + 1, ntask = kc * per;
        for (int t = tid; t < ntask; t += NT) {
          const int k = t / per, mi = t - k * per;
          const TTrk tk = a.ttrk[trk0 + k];
          double* G = Gt + 3 * k;
          int d = 0;
          while (rpc_tri(d + 1) <= mi) d++;
          const int iy = mi - rpc_tri(d), ax = d - iy;
          const bool focal = mi == a.nmono;
          int cur = -1;
          double s0[3] = {0, 0, 0}, s1[3] = {0, 0, 0}, s2[3] = {0, 0, 0}, s3[3] = {0, 0, 0};
          auto flush = [&](int sensor) {
            const SensorDev& sd = v.sens[sensor];
            auto put = [&](int gc, const double* sv) {
              const int l2 = 7 * Wt + (gc - S.npc);
              const double sc = cs[l2];
              double* g = G + (size_t)l2 * KK;
              g[0] += sv[0] * sc; g[1] += sv[1] * sc; g[2] += sv[2] * sc;
            };
            if (focal) { if (sd.c_focal >= 0) put(sd.c_focal, s0); }
            else if (sd.c_dist >= 0 && d <= sd.rpc_deg) {
              const RpcLayout L = rpc_layout(sd.rpc_deg);
              put(sd.c_dist + mi, s0);
              put(sd.c_dist + L.o2 + mi, s2);
              if (d > 0) { put(sd.c_dist + L.o1 + mi - 1, s1); put(sd.c_dist + L.o3 + mi - 1, s3); }
            }
#pragma unroll
            for (int b = 0; b < 3; b++) { s0[b] = 0; s1[b] = 0; s2[b] = 0; s3[b] = 0; }
          };
          for (int i = 0; i < tk.n; i++) {
            const TRec& r = rec_s[tk.q0 + i];
            if (r.nx != -1) continue;
            const int sensor = v.cam_sensor[r.cid];
            if (sensor != cur) { if (cur >= 0) flush(cur); cur = sensor; }
            const double* x = Xs + (size_t)(tk.q0 + i) * MVBA_XR;
            if (focal) {
#pragma unroll
              for (int b = 0; b < 3; b++) s0[b] += x[12] * x[6 + b] + x[13] * x[9 + b];
            } else {
              double m = 1.0;
              for (int e = 0; e < ax; e++) m *= x[0];
              for (int e = 0; e < iy; e++) m *= x[1];
              const double a0 = m * x[2], a1 = m * x[4], a2 = m * x[3], a3 = m * x[5];
#pragma unroll
              for (int b = 0; b < 3; b++) { s0[b] += a0 * x[6 + b]; s1[b] += a1 * x[6 + b]; s2[b] += a2 * x[9 + b]; s3[b] += a3 * x[9 + b]; }
            }
          }
          if (cur >= 0) flush(cur);
        }
      }
      if (tid < kc) {
        const TTrk tk = a.ttrk[trk0 + tid];
        double H[6] = {0, 0, 0, 0, 0, 0}, gE[3] = {0, 0, 0};
        for (int i = 0; i < tk.n; i++) {
          const TRec& r = rec_s[tk.q0 + i];
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
        for (int k = 0; k < 6; k++) { a.b.Hinv[6 * (size_t)tk.p + k] = Hi[k]; His[k * K + tid] = Hi[k]; }
#pragma unroll
        for (int k = 0; k < 3; k++) a.b.ge[3 * (size_t)tk.p + k] = gE[k];
        ves[3 * tid] = Hi[0] * gE[0] + Hi[1] * gE[1] + Hi[2] * gE[2];
        ves[3 * tid + 1] = Hi[1] * gE[0] + Hi[3] * gE[1] + Hi[4] * gE[2];
        ves[3 * tid + 2] = Hi[2] * gE[0] + Hi[4] * gE[1] + Hi[5] * gE[2];
      }
      if (a.nx > 0 && !a.direct) {  // free intrinsics / depth columns staged by phase b1: a warp per track, disjoint from the gathered columns
        for (int k = wib; k < kc; k += NW) {
          const TTrk tk = a.ttrk[trk0 + k];
          double* G = Gt + 3 * k;
          for (int i = 0; i < tk.n; i++) {
            const int nxi = rec_s[tk.q0 + i].nx;
            const double* X = Xs_s + (size_t)(tk.q0 + i) * 4 * a.nx;
            for (int kk = lane; kk < nxi; kk += 32) {
              const int l2 = (int)X[4 * kk + 3];
              const double sc = cs[l2];
              double* g = G + (size_t)l2 * KK;
              g[0] += X[4 * kk] * sc; g[1] += X[4 * kk + 1] * sc; g[2] += X[4 * kk + 2] * sc;
            }
            __syncwarp();
          }
        }
      }
      TPROF(2);
      __syncthreads();
      TPROF(3);
      // ---- phase b3: Y = G Hinv and rhs += G Hinv E'r (a warp per row of Gt, lanes over the staged tracks);
      //      lane j of warp w keeps the rhs entry of row w + NW j ----
      {
        int j = 0;
        for (int l = wib; l < L8; l += NW, j++) {
          double part = 0;
          for (int k = lane; k < kc; k += 32) {
            const double* g = Gt + (size_t)l * KK + 3 * k;
            double* y = Yt + (size_t)l * KK + 3 * k;
            const double g0 = g[0], g1 = g[1], g2 = g[2];
            const double h0 = His[k], h1 = His[K + k], h2 = His[2 * K + k], h3 = His[3 * K + k], h4 = His[4 * K + k], h5 = His[5 * K + k];
            y[0] = h0 * g0 + h1 * g1 + h2 * g2;
            y[1] = h1 * g0 + h3 * g1 + h4 * g2;
            y[2] = h2 * g0 + h4 * g1 + h5 * g2;
            part = fma(g0, ves[3 * k], fma(g1, ves[3 * k + 1], fma(g2, ves[3 * k + 2], part)));
          }
          part = warp_sum(part);
          if (lane == j) rhs_acc += part;
        }
      }
      __syncthreads();
      TPROF(4);
      // ---- phase c: C += Yt Gt' on the FP64 tensor cores, K dimension = 3 kc (padded with zero columns of Gt) ----
      if (ns > 0) {
        const int kend = (3 * kc + 3) & ~3;
        const double* ybase = Yt + (size_t)gid * KK + tig;
        const double* gbase = Gt + (size_t)gid * KK + tig;
        if constexpr (MAXS >= 24) {
          // wide windows, branch-free inside groups of eight slots: the slots that end a block row were marked once per tile; the
          // B fragment walks down Gt by a constant stride and restarts on the diagonal at a row end; a slot beyond the warp's
          // last block multiplies by zero (its loads read row 0)
          const int rstride = 8 * KK;
          for (int k0 = 0; k0 < kend; k0 += 4) {
            const double* yk = ybase + k0;
            const double* gk = gbase + k0;
            int arow = bi0 * rstride, grow = bj0 * rstride;
#pragma unroll
            for (int g8 = 0; g8 < MAXS; g8 += 8) {
              if (g8 < ns) {
#pragma unroll
                for (int sl = g8; sl < g8 + 8; sl++) {
                  const bool on = sl < ns;
                  const double av = on ? yk[arow] : 0.0;
                  const double bv = gk[on ? grow : 0];
                  dmma884(acc[sl][0], acc[sl][1], av, bv);
                  const bool end = (rowend >> sl) & 1u;
                  arow += end ? rstride : 0;
                  grow = end ? arow : grow + rstride;
                }
              }
            }
          }
        } else {
#pragma unroll(MAXS <= 4 ? 4 : MAXS <= 8 ? 2 : 1)
        for (int k0 = 0; k0 < kend; k0 += 4) {
          int bi = bi0, bj = bj0;
          double av = ybase[(size_t)(8 * bi) * KK + k0];
#pragma unroll
          for (int sl = 0; sl < MAXS; sl++) {
            if (sl < ns) {
              const double bv = gbase[(size_t)(8 * bj) * KK + k0];
              dmma884(acc[sl][0], acc[sl][1], av, bv);
              if (++bj == nb8) { bi++; bj = bi; if (sl + 1 < ns) av = ybase[(size_t)(8 * bi) * KK + k0]; }
            }
          }
        }
        }
      }
      TPROF(5);
      __syncthreads();
      TPROF(6);
      }  // sub
    }
    // ---- flush: local column -> reduced column; window poses are consecutive pose blocks ----
    {
      int bi = bi0, bj = bj0;
#pragma unroll
      for (int sl = 0; sl < MAXS; sl++) {
        if (sl < ns) {
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int li = 8 * bi + gid, lj = 8 * bj + 2 * tig + e;
            const double val = acc[sl][e];
            if (li <= lj && lj < Lt && val != 0.0) {
              const int gi = li < 7 * Wt ? 7 * w0 + li : S.npc + (li - 7 * Wt);
              const int gj = lj < 7 * Wt ? 7 * w0 + lj : S.npc + (lj - 7 * Wt);
              addS(S, gi, gj, -val);
            }
          }
          if (++bj == nb8) { bi++; bj = bi; }
        }
      }
    }
    {
      const int l = wib + NW * lane;  // the row whose rhs entry this lane accumulated in phase b3
      if (l < Lt && rhs_acc != 0.0) {
        const int gi = l < 7 * Wt ? 7 * w0 + l : S.npc + (l - 7 * Wt);
        atomicAdd(S.rhs + gi, -rhs_acc);
      }
    }
  }
  cost_acc = warp_sum(cost_acc);
  if (lane == 0 && cost_acc != 0.0) atomicAdd(a.b.cost, cost_acc);
#ifdef MVBA_TILE_PROFILE
  TPROF(7);
  if (tid == 0) for (int i = 0; i < 8; i++) atomicAdd(&g_tile_prof[i], (unsigned long long)prof_[i]);
#endif
}

}  // namespace mvba

// kernels_build.cuh -- pass B of the normal-equation build, generic version: one warp per track.
// Residuals + analytic Jacobians are evaluated in registers, G = F'E and the 3x3 landmark block E'E
// are reduced in shared memory, the landmark block is inverted and the track is Schur-eliminated
// straight into the reduced camera system with atomics.  Jacobians never touch HBM.  This kernel
// takes any track (any number of poses / columns); kernels_tile.cuh is the fast path for tracks
// whose poses fall in a narrow window.  F'F and F'r come from pass A (kernels_camera.cuh).
// (north-star kernel (b); Ceres 1.14 equivalent: SchurEliminator::Eliminate, made explicit.)
#pragma once
#include "items.cuh"

namespace mvba {

#define MVBA_CHUNK 8  // items evaluated per warp pass (one lane each)

// Reduced camera system  S = F'F + D_f^2 - sum_tracks (F'E)(E'E + D_e^2)^-1 (E'F), stored as
//   Spp  : 7x7 blocks for pose-block pairs u1 <= u2 (slot = pairtab[u1*nb+u2]); diagonal blocks upper
//   Spb  : npose_cols x B  (pose rows x border columns)
//   Sbb  : B x B, upper triangle
struct Reduced {
  double *Spp, *Spb, *Sbb;
  double *Fpp, *Fbb;     // block diagonal of F'F (Ceres JACOBI preconditioner); may be null
  double *g, *diagF, *rhs;  // nf each: F'r, diag(F'F), reduced right-hand side
  const int32_t* pairtab;
  int nb, npc, B, nf;
};

struct BuildArgs {
  const double* x;
  const CamTab* tab;
  const DepthTab* dtab;
  const double *jscale_f, *jscale_e;  // Jacobi column scaling (null: 1)
  double radius, min_diag, max_diag;
  double* Hinv;   // 6 per landmark
  double* ge;     // 3 per landmark: E' r (scaled)
  double* diagE;  // 3 per landmark: diag(E'E) (COLNORM mode)
  double* cost;   // sum of 0.5 rho
  int mode;       // 0 full build, 1 column norms only
  int want_cost;  // accumulate sum 0.5 rho (only the first build of a solve needs it: later costs come from the candidate pass)
  int maxposes;   // per-track capacity of the local pose list
  int maxdz;      // max nd_nz
  int vwmax;      // max virtual width
  const int32_t* trk_list;  // tracks to process (null: all of [0, ntrk))
  int ntrk;
};

__device__ __forceinline__ void addS(const Reduced& S, int g1, int g2, double val) {
  if (g1 > g2) { int t = g1; g1 = g2; g2 = t; }
  if (g2 < S.npc) {
    int u1 = g1 / 7, u2 = g2 / 7;
    int slot = S.pairtab[(size_t)u1 * S.nb + u2];
    atomicAdd(S.Spp + (size_t)slot * 49 + (g1 - 7 * u1) * 7 + (g2 - 7 * u2), val);
  } else if (g1 < S.npc) {
    atomicAdd(S.Spb + (size_t)g1 * S.B + (g2 - S.npc), val);
  } else {
    atomicAdd(S.Sbb + (size_t)(g1 - S.npc) * S.B + (g2 - S.npc), val);
  }
}
__device__ __forceinline__ void addFdiag(const Reduced& S, int g1, int g2, double val) {
  if (g1 > g2) { int t = g1; g1 = g2; g2 = t; }
  if (g2 < S.npc) {
    int u1 = g1 / 7, u2 = g2 / 7;
    if (u1 == u2) atomicAdd(S.Fpp + (size_t)u1 * 49 + (g1 - 7 * u1) * 7 + (g2 - 7 * u2), val);
  } else if (g1 >= S.npc) {
    atomicAdd(S.Fbb + (size_t)(g1 - S.npc) * S.B + (g2 - S.npc), val);
  }
}

// What phase 1 leaves in shared memory for the column phase.
struct Stash {
  int kind, nr, cid, sensor;
  double r[3], A[9], E[9], Z[3], Xc[3], Jf[3], m[3], Jc, sr;
};

// triangular index -> (i <= j) with idx = j (j+1)/2 + i
__device__ __forceinline__ void tri_decode(int idx, int& i, int& j) {
  j = (int)((sqrtf(8.0f * (float)idx + 1.0f) - 1.0f) * 0.5f);
  while ((j + 1) * (j + 2) / 2 <= idx) j++;
  while (j * (j + 1) / 2 > idx) j--;
  i = idx - j * (j + 1) / 2;
}

// Dynamic shared memory per warp (doubles first, then ints):
//   stash[CHUNK] | Jd[CHUNK][2][maxdz] | G[nloc][3] | Y[nloc][3] | lcol[nloc] (int) | lposes[maxposes] (int)
__host__ __device__ inline size_t build_smem_per_warp(int maxposes, int B, int maxdz, int vwmax) {
  size_t nloc = (size_t)7 * maxposes + B;
  size_t d = sizeof(Stash) * MVBA_CHUNK / 8 + (size_t)MVBA_CHUNK * 2 * maxdz + nloc * 6;
  size_t i = nloc + maxposes;
  return d * 8 + ((i * 4 + 7) / 8) * 8;
}

__global__ void __launch_bounds__(256) build_kernel(View v, BuildArgs a, Reduced S) {
  extern __shared__ double smem_d[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int nlocmax = 7 * a.maxposes + S.B;
  const size_t per_warp = build_smem_per_warp(a.maxposes, S.B, a.maxdz, a.vwmax) / 8;
  double* base = smem_d + per_warp * wib;
  Stash* stash = reinterpret_cast<Stash*>(base);
  double* Jd = base + sizeof(Stash) * MVBA_CHUNK / 8;
  double* G = Jd + (size_t)MVBA_CHUNK * 2 * a.maxdz;
  double* Y = G + 3 * (size_t)nlocmax;
  int* lcol = reinterpret_cast<int*>(Y + 3 * (size_t)nlocmax);
  int* lposes = lcol + nlocmax;
  const int B = S.B;
  double cost_acc = 0;

  for (int ti = blockIdx.x * wpb + wib; ti < a.ntrk; ti += gridDim.x * wpb) {
    const int p = a.trk_list ? a.trk_list[ti] : ti;
    const int64_t o0 = v.trk_obs0[p], o1 = v.trk_obs0[p + 1];
    const int64_t d0 = v.trk_dtri0[p], d1 = v.trk_dtri0[p + 1];
    const int64_t m0 = v.trk_dmesh0[p], m1 = v.trk_dmesh0[p + 1];
    const int64_t pm0 = v.trk_pm0[p], pm1 = v.trk_pm0[p + 1];
    const int64_t pt0 = v.trk_pt0[p], pt1 = v.trk_pt0[p + 1];
    const int n_obs = (int)(o1 - o0), n_dt = (int)(d1 - d0), n_dm = (int)(m1 - m0), n_pm = (int)(pm1 - pm0), n_pt = (int)(pt1 - pt0);
    const int nitems = n_obs + n_dt + n_dm + n_pm + n_pt;
    if (nitems == 0) continue;
    // local system starts with the border columns
    int npl = 0;
    for (int li = lane; li < nlocmax; li += 32) { G[3 * li] = G[3 * li + 1] = G[3 * li + 2] = 0; lcol[li] = li < B ? S.npc + li : -1; }
    double H[6] = {0, 0, 0, 0, 0, 0}, gE[3] = {0, 0, 0};
    double se[3] = {1, 1, 1};
    if (a.jscale_e) { se[0] = a.jscale_e[3 * (size_t)p]; se[1] = a.jscale_e[3 * (size_t)p + 1]; se[2] = a.jscale_e[3 * (size_t)p + 2]; }
    __syncwarp();

    for (int c0 = 0; c0 < nitems; c0 += MVBA_CHUNK) {
      const int cnt = min(MVBA_CHUNK, nitems - c0);
      // ---- phase 1: one lane per item ----
      if (lane < cnt) {
        int q = c0 + lane;
        Item it;
        double* jd0 = Jd + (size_t)lane * 2 * a.maxdz;
        if (q < n_obs) eval_cam<true>(v, a.x, a.tab, o0 + q, it, jd0, jd0 + a.maxdz);
        else if (q < n_obs + n_dt) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_TRI, d0 + (q - n_obs), it);
        else if (q < n_obs + n_dt + n_dm) eval_depth(v, a.x, a.tab, a.dtab, MVBA_BLOCK_DEPTH_MESH, m0 + (q - n_obs - n_dt), it);
        else if (q < n_obs + n_dt + n_dm + n_pm) eval_prior(v, a.x, pm0 + (q - n_obs - n_dt - n_dm), it);
        else eval_prior(v, a.x, pt0 + (q - n_obs - n_dt - n_dm - n_pm), it);
        double s = 0;
        for (int k = 0; k < it.nr; k++) s += it.r[k] * it.r[k];
        double hr, sr;
        cauchy(it.loss_a, s, &hr, &sr);
        cost_acc += hr;
        Stash& st = stash[lane];
        st.kind = it.kind; st.nr = it.nr; st.cid = it.cid; st.sensor = it.sensor; st.sr = sr; st.Jc = it.Jc;
#pragma unroll
        for (int k = 0; k < 3; k++) { st.r[k] = k < it.nr ? sr * it.r[k] : 0.0; st.Z[k] = it.Z[k]; st.Xc[k] = it.Xc[k]; st.Jf[k] = it.Jf[k]; st.m[k] = it.m[k]; }
#pragma unroll
        for (int k = 0; k < 9; k++) { st.A[k] = it.A[k]; st.E[k] = (k / 3 < it.nr) ? sr * it.E[k] * se[k % 3] : 0.0; }
        for (int r = 0; r < it.nr; r++) {
          const double* e = st.E + 3 * r;
          H[0] += e[0] * e[0]; H[1] += e[0] * e[1]; H[2] += e[0] * e[2]; H[3] += e[1] * e[1]; H[4] += e[1] * e[2]; H[5] += e[2] * e[2];
          gE[0] += e[0] * st.r[r]; gE[1] += e[1] * st.r[r]; gE[2] += e[2] * st.r[r];
        }
      }
      __syncwarp();
      if (a.mode == 1) continue;  // column norms: only diag(E'E) is needed here
      // ---- phase 2: the warp walks the chunk's items, lanes over Jacobian columns: G += J_f' E ----
      for (int j = 0; j < cnt; j++) {
        const Stash& st = stash[j];
        if (st.kind == MVBA_BLOCK_MESH_TRI || st.kind == MVBA_BLOCK_TRI) continue;  // no f columns
        Item it;  // rebuilt from the stash (registers)
        it.kind = st.kind; it.nr = st.nr; it.cid = st.cid; it.sensor = st.sensor; it.Jc = st.Jc;
#pragma unroll
        for (int k = 0; k < 3; k++) { it.Z[k] = st.Z[k]; it.Xc[k] = st.Xc[k]; it.Jf[k] = st.Jf[k]; it.m[k] = st.m[k]; }
#pragma unroll
        for (int k = 0; k < 9; k++) it.A[k] = st.A[k];
        // local pose indices of this item's bracketing poses (warp-uniform)
        int lp[2] = {-1, -1};
        for (int w = 0; w < 2; w++) {
          int pose = w == 0 ? v.cam_p0[st.cid] : v.cam_p1[st.cid];
          if (pose < 0) continue;
          int pc = v.pose_col[pose];
          if (pc < 0) continue;
          int found = -1;
          for (int b0 = 0; b0 < npl; b0 += 32) {
            int idx = b0 + lane;
            unsigned mk = __ballot_sync(0xffffffffu, idx < npl && lposes[idx] == pc);
            if (mk) { found = b0 + __ffs(mk) - 1; break; }
          }
          if (found < 0) {
            found = npl;
            if (lane == 0) lposes[npl] = pc;
            if (lane < 7) lcol[B + 7 * npl + lane] = pc + lane;
            npl++;
            __syncwarp();
          }
          lp[w] = found;
        }
        const int vw = item_vwidth(v, it);
        const double* jd0 = Jd + (size_t)j * 2 * a.maxdz;
        for (int v0 = 0; v0 < vw; v0 += 32) {
          int vc = v0 + lane;
          int gc = -1;
          if (vc < vw) gc = item_gcol(v, it, vc);
          if (gc >= 0) {
            double col[3] = {0, 0, 0};
            item_jcol(v, it, a.tab, a.dtab, vc, jd0, jd0 + a.maxdz, col);
            double sc = st.sr * (a.jscale_f ? a.jscale_f[gc] : 1.0);
            int li = gc >= S.npc ? gc - S.npc : B + 7 * (vc < 7 ? lp[0] : lp[1]) + (vc < 7 ? vc : vc - 7);
            double g0 = 0, g1 = 0, g2 = 0;
            for (int r = 0; r < st.nr; r++) {
              double jv = col[r] * sc;
              g0 += jv * st.E[3 * r]; g1 += jv * st.E[3 * r + 1]; g2 += jv * st.E[3 * r + 2];
            }
            G[3 * li] += g0; G[3 * li + 1] += g1; G[3 * li + 2] += g2;
          }
        }
        __syncwarp();
      }
    }
    // ---- landmark block ----
#pragma unroll
    for (int k = 0; k < 6; k++) H[k] = warp_sum(H[k]);
#pragma unroll
    for (int k = 0; k < 3; k++) gE[k] = warp_sum(gE[k]);
    const int nloc = B + 7 * npl;
    if (a.mode == 1) {
      if (lane == 0) { a.diagE[3 * (size_t)p] = H[0]; a.diagE[3 * (size_t)p + 1] = H[3]; a.diagE[3 * (size_t)p + 2] = H[5]; }
      __syncwarp();
      continue;
    }
    H[0] += fmin(fmax(H[0], a.min_diag), a.max_diag) / a.radius;
    H[3] += fmin(fmax(H[3], a.min_diag), a.max_diag) / a.radius;
    H[5] += fmin(fmax(H[5], a.min_diag), a.max_diag) / a.radius;
    double Hi[6];
    inv_sym3(H, Hi);
    if (lane < 6) a.Hinv[6 * (size_t)p + lane] = Hi[lane];
    if (lane < 3) a.ge[3 * (size_t)p + lane] = gE[lane];
    const double ve[3] = {Hi[0] * gE[0] + Hi[1] * gE[1] + Hi[2] * gE[2], Hi[1] * gE[0] + Hi[3] * gE[1] + Hi[4] * gE[2],
                          Hi[2] * gE[0] + Hi[4] * gE[1] + Hi[5] * gE[2]};
    for (int li = lane; li < nloc; li += 32) {
      double g0 = G[3 * li], g1 = G[3 * li + 1], g2 = G[3 * li + 2];
      Y[3 * li] = Hi[0] * g0 + Hi[1] * g1 + Hi[2] * g2;
      Y[3 * li + 1] = Hi[1] * g0 + Hi[3] * g1 + Hi[4] * g2;
      Y[3 * li + 2] = Hi[2] * g0 + Hi[4] * g1 + Hi[5] * g2;
      if (g0 != 0.0 || g1 != 0.0 || g2 != 0.0) atomicAdd(S.rhs + lcol[li], -(g0 * ve[0] + g1 * ve[1] + g2 * ve[2]));
    }
    __syncwarp();
    // ---- Schur elimination: S -= G Hinv G' over the track's local columns ----
    const int npairs = nloc * (nloc + 1) / 2;
    for (int idx = lane; idx < npairs; idx += 32) {
      int l1, l2;
      tri_decode(idx, l1, l2);
      double val = Y[3 * l1] * G[3 * l2] + Y[3 * l1 + 1] * G[3 * l2 + 1] + Y[3 * l1 + 2] * G[3 * l2 + 2];
      if (val != 0.0) addS(S, lcol[l1], lcol[l2], -val);
    }
    __syncwarp();
  }
  cost_acc = warp_sum(cost_acc);
  if (lane == 0 && a.mode == 0) atomicAdd(a.cost, cost_acc);
}

}  // namespace mvba

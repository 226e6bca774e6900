// solver.cuh -- host orchestration of the device Levenberg-Marquardt loop:
//   fused build (Jacobian + J'J + Schur) -> [NCCL allreduce] -> damping + block-Jacobi -> PCG
//   -> landmark back-substitution -> candidate cost -> accept/reject + radius on device.
// Follows Ceres 1.14 TrustRegionMinimizer with ITERATIVE_SCHUR as configured by the reference
// (rig_calibrator.cc:1910-1918); the Schur complement is formed explicitly instead of implicitly.
#pragma once
#include <dlfcn.h>

#include <chrono>

#include "handle.cuh"
#include "kernels_lm.cuh"
#include "kernels_camera.cuh"
#include "kernels_camq.cuh"
#include "kernels_tile.cuh"
#include "kernels_tile2.cuh"
#include "kernels_chol.cuh"

namespace mvba {

// ---- NCCL, loaded at run time so that libmvba.so has no link-time dependency on it ----
struct NcclId { char b[128]; };
struct Nccl {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId, int) = nullptr;  // ncclUniqueId is passed BY VALUE (a 128-byte struct)
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
inline Nccl& nccl() {
  static Nccl n;
  if (!n.lib) {
    n.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!n.lib) n.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (n.lib) {
      n.GetUniqueId = (int (*)(void*))dlsym(n.lib, "ncclGetUniqueId");
      n.CommInitRank = (int (*)(void**, int, NcclId, int))dlsym(n.lib, "ncclCommInitRank");
      n.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(n.lib, "ncclAllReduce");
      n.CommDestroy = (int (*)(void*))dlsym(n.lib, "ncclCommDestroy");
      n.GetErrorString = (const char* (*)(int))dlsym(n.lib, "ncclGetErrorString");
    }
  }
  return n;
}
inline int comm_unique_id(uint8_t* id128) {
  Nccl& n = nccl();
  if (!n.lib || !n.GetUniqueId) return fail(MVBA_ENCCL, "mvba: libnccl.so.2 not loadable");
  int rc = n.GetUniqueId(id128);
  return rc ? fail(MVBA_ENCCL, std::string("ncclGetUniqueId: ") + n.GetErrorString(rc)) : MVBA_OK;
}
inline int comm_init(mvba_problem* h, int rank, int nranks, const uint8_t* id128) {
  if (nranks <= 1) { h->rank = 0; h->nranks = 1; return MVBA_OK; }
  Nccl& n = nccl();
  if (!n.lib || !n.CommInitRank) return fail(MVBA_ENCCL, "mvba: libnccl.so.2 not loadable");
  NcclId id; std::memcpy(id.b, id128, 128);
  int rc = n.CommInitRank(&h->nccl_comm, nranks, id, rank);
  if (rc) return fail(MVBA_ENCCL, std::string("ncclCommInitRank: ") + n.GetErrorString(rc));
  h->rank = rank; h->nranks = nranks;
  h->built = false;  // the column layout becomes the union over ranks (setup.cuh)
  return MVBA_OK;
}
// op: 0 sum, 2 max (ncclRedOp_t); dtype 8 = ncclFloat64, 5 = ncclUint64
inline int allreduce(mvba_problem* h, void* buf, size_t n, int dtype, int op) {
  if (h->nranks <= 1 || n == 0) return MVBA_OK;
  Nccl& nc = nccl();
  int rc = nc.AllReduce(buf, buf, n, dtype, op, h->nccl_comm, h->stream);
  return rc ? fail(MVBA_ENCCL, std::string("ncclAllReduce: ") + nc.GetErrorString(rc)) : MVBA_OK;
}

// Collective agreement on a return code: every rank passes its own; when any rank failed, the healthy ranks get MVBA_EPEER
// and return too instead of blocking in the next collective (the failing rank keeps its own code and message).
inline int rank_agree(mvba_problem* h, int rc) {
  if (h->nranks <= 1) return rc;
  int32_t flag = rc != 0, any = 0;
  DBuf<int32_t> d;
  if (d.alloc(1) != cudaSuccess) return rc ? rc : fail(MVBA_ECUDA, "mvba: rank_agree allocation failed");
  cudaMemcpyAsync(d.p, &flag, sizeof(flag), cudaMemcpyHostToDevice, h->stream);
  const int arc = allreduce(h, d.p, 1, 2, 2);  // ncclInt32, ncclMax
  cudaMemcpyAsync(&any, d.p, sizeof(any), cudaMemcpyDeviceToHost, h->stream);
  cudaStreamSynchronize(h->stream);
  if (rc) return rc;
  if (arc) return arc;
  return any ? fail(MVBA_EPEER, "mvba: another rank failed before a collective; see that rank's error") : MVBA_OK;
}
// Element-wise OR of a byte table across ranks (ncclUint8 = 1, ncclMax = 2): layout decisions every rank must share.
inline int allreduce_flags(mvba_problem* h, std::vector<uint8_t>& flags) {
  if (h->nranks <= 1 || flags.empty()) return MVBA_OK;
  DBuf<uint8_t> d;
  MVBA_CUDA(d.upload(flags));
  int rc = allreduce(h, d.p, flags.size(), 1, 2); if (rc) return rc;
  MVBA_CUDA(cudaStreamSynchronize(h->stream));
  MVBA_CUDA(cudaMemcpy(flags.data(), d.p, flags.size(), cudaMemcpyDeviceToHost));
  return MVBA_OK;
}

struct Timers;
struct Solver {
  Timers* tm = nullptr;
  int build_launches = 0;
  int nb = 0, npc = 0, B = 0, nf = 0;
  int64_t nslots = 0, nnzb = 0;
  int maxposes = 1, maxitems = 1, maxdz = 1, vwmax = MVBA_VPOSE + 4;
  DBuf<int32_t> pairtab, row_ptr, col, slot, xcol, border_blk;
  DBuf<double> red;  // [Spp | Spb | Sbb | Fpp | Fbb | g | diagF | rhs | cost]  (one allreduce)
  DBuf<double> redA; // same layout: what pass A (the camera part F'F, F'r) adds, kept so that the build after a REJECTED step -- same
                     // point, new damping -- starts from it instead of evaluating pass A again (Ceres does not re-evaluate either)
  bool redA_valid = false;
  size_t o_Spp = 0, o_Spb = 0, o_Sbb = 0, o_Fpp = 0, o_Fbb = 0, o_g = 0, o_diagF = 0, o_rhs = 0, o_cost = 0, o_gmaxe = 0, red_n = 0;
  DBuf<double> Dsq, Mpp, Mbb, work, jscale_f, jscale_e, Hinv, ge, diagE, lo, hi;
  DBuf<double> y, r, z, p, p2, q, ye, lm, pcg_scal;
  DBuf<int> pcg_flags;
  DBuf<CamDir> camdir;
  DBuf<SensDir> sensdir;
  mvba_iteration* rec = nullptr;  // pinned
  int* flags_host = nullptr;      // pinned
  size_t smem_build = 0, smem_backsub = 0;
  int wpb_build = 8, wpb_backsub = 8, grid_build = 1, grid_backsub = 1;
  int pcg_coop_grid = 0;  // > 0: CTAs of the single-launch PCG (cooperative); 0: step kernels
  int pcg_rows_per_cta = 0;  // > 0: every CTA keeps the blocks of that many block rows in shared memory for the whole solve
  size_t smem_pcg = 0;
  int pcg_threads = 512;  // measured on config 5 (profiles/r02f_pcg_geometry_sweep.txt): 512 x 148..296 CTAs 0.231 ms per LM iteration, 256 x 296 0.240
  bool with_F = false;
  // pass A (camera-major F'F): block codes grouped by image, split into chunks
  DBuf<int32_t> citems, chunk_cam, chunk_beg;
  int nchunks = 0, cam_batch = 0, cam_ldj = 0, grid_cam = 1;
  size_t smem_cam = 0;
  // pass A without free intrinsics / depth columns: per-image 6 x 6 accumulators (kernels_camq.cuh)
  bool cam_fast = false;
  DBuf<double> camQ;
  DBuf<TSlot> crec;  // camera-major packed records (pass A without staged rows)
  int grid_camq = 1, grid_camcost = 1;
  // pass B fast path: tracks sorted by first pose block and cut into window tiles; the rest fall back
  // (static round tables: staged blocks, per-track slots, gather groups and their contributors, kernels_tile.cuh)
  DBuf<int32_t> tile_w0, tile_wn, tile_round0, round_trk0, round_blk0, round_grp0, fallback;
  DBuf<TBlk> t_blk; DBuf<TTrk> t_trk; DBuf<TGrp> t_grp; DBuf<uint16_t> t_con;
  int ntiles = 0, nrounds = 0, nfallback = 0, tile_nt = 512, tile_pr = 4, tile_K = 1, tile_icap = 64, tile_nx = 0, tile_L8max = 8, grid_tile = 1, tile_direct = 0, tile_nmono = 0;
  size_t smem_tile = 0;
  // pass B, staged variant (kernels_tile2.cuh): per-round blobs fetched by bulk copy, two CTAs per SM
  bool any_x = false, tile2 = false;
  DBuf<int32_t> t2_c0, t2_nc, t2_rounds, tile_ctr;
  DBuf<int64_t> t2_blob0; DBuf<uint32_t> t2_bytes0; DBuf<uint8_t> t2_blob;
  int t2_K = 1, t2_icap = 32, t2_ccap = 0, t2_L8max = 8, t2_blob_cap = 64, t2_maxs = 8, grid_tile2 = 1;
  size_t smem_tile2 = 0;
  // direct solver (linear_solver = 1): dense lower triangle + right-hand-side row, inverses of the diagonal blocks
  DBuf<double> cholA, cholLinv;
  Reduced R{}, RA{};
  Bsr A{};
};

inline void solver_destroy(mvba_problem* h) {
  if (!h->solver) return;
  if (h->solver->rec) cudaFreeHost(h->solver->rec);
  if (h->solver->flags_host) cudaFreeHost(h->solver->flags_host);
  delete h->solver; h->solver = nullptr;
}

// Structure analysis: which pose-block pairs are co-observed by some track, shared-memory sizing.
inline int solver_prepare(mvba_problem* h, bool with_F) {
  if (h->solver && h->solver->with_F == with_F) return MVBA_OK;
  struct PrepTimer {
    mvba_problem* h; std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    ~PrepTimer() { h->seconds_prepare = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
  } prep_timer{h};
  solver_destroy(h);
  AgreePoint agree_pattern(h);  // rank-local failures below must not leave the peers alone in the collective
  Solver* s = new Solver();
  h->solver = s;
  const View& v = h->view;
  s->with_F = with_F;
  s->npc = v.npose_cols; s->nb = v.npose_cols / 7; s->B = v.B; s->nf = v.nf;
  const int nb = s->nb;
  if ((size_t)nb * nb > ((size_t)1 << 31)) return fail(MVBA_EINVAL, "mvba: too many free pose blocks for the dense pair table");
  std::vector<uint8_t> flag((size_t)nb * nb, 0);
  std::vector<int> uniq;
  std::vector<int32_t> trk_umin(v.P, 0), trk_span(v.P, 0);
  std::vector<uint8_t> trk_any(v.P, 0);
  std::vector<int32_t> trk_items(v.P, 0);
  int maxposes = 1, maxitems = 1;
  for (int p = 0; p < v.P; p++) {
    uniq.clear();
    for (int64_t i = h->h_trk_obs0[p]; i < h->h_trk_obs0[p + 1]; i++) {
      int c = h->ocid[i];
      for (int w = 0; w < 2; w++) {
        int pose = w == 0 ? h->h_cam_p0[c] : h->h_cam_p1[c];
        if (pose < 0) continue;
        int pc = h->h_pose_col[pose];
        if (pc < 0) continue;
        int u = pc / 7;
        if (std::find(uniq.begin(), uniq.end(), u) == uniq.end()) uniq.push_back(u);
      }
    }
    for (size_t a = 0; a < uniq.size(); a++) for (size_t b = 0; b < uniq.size(); b++) { int u1 = uniq[a], u2 = uniq[b]; if (u1 <= u2) flag[(size_t)u1 * nb + u2] = 1; }
    maxposes = std::max(maxposes, (int)uniq.size());
    if (!uniq.empty()) {
      int lo = *std::min_element(uniq.begin(), uniq.end()), hi = *std::max_element(uniq.begin(), uniq.end());
      trk_umin[p] = lo; trk_span[p] = hi - lo + 1;
    }
    int items = (int)(h->h_trk_obs0[p + 1] - h->h_trk_obs0[p] + h->h_trk_dtri0[p + 1] - h->h_trk_dtri0[p] + h->h_trk_dmesh0[p + 1] - h->h_trk_dmesh0[p] +
                      h->h_trk_pm0[p + 1] - h->h_trk_pm0[p] + h->h_trk_pt0[p + 1] - h->h_trk_pt0[p]);
    maxitems = std::max(maxitems, items);
    trk_any[p] = items > 0; trk_items[p] = items;
  }
  for (int u = 0; u < nb; u++) flag[(size_t)u * nb + u] = 1;  // diagonal always present (damping)
  { int rc = agree_pattern.pass(); if (rc) return rc; }
  AgreePoint agree_end(h);
  { int rc = allreduce_flags(h, flag); if (rc) return rc; }      // block pattern of S = union over the ranks' landmark shards
  std::vector<int32_t> pairtab((size_t)nb * nb, -1), row_ptr(nb + 1, 0), col, slot;
  int64_t ns = 0;
  std::vector<std::vector<std::pair<int, int>>> rows(nb);
  for (int u1 = 0; u1 < nb; u1++)
    for (int u2 = u1; u2 < nb; u2++)
      if (flag[(size_t)u1 * nb + u2]) {
        pairtab[(size_t)u1 * nb + u2] = (int32_t)ns;
        rows[u1].push_back({u2, (int)ns});
        if (u1 != u2) rows[u2].push_back({u1, -(int)ns - 1});
        ns++;
      }
  for (int u = 0; u < nb; u++) {
    std::sort(rows[u].begin(), rows[u].end());
    row_ptr[u + 1] = row_ptr[u] + (int)rows[u].size();
    for (auto& e : rows[u]) { col.push_back(e.first); slot.push_back(e.second); }
  }
  s->nslots = ns; s->nnzb = (int64_t)col.size();
  s->maxposes = maxposes; s->maxitems = maxitems;
  int maxdz = 1, vw = MVBA_VPOSE + 3;
  for (int t = 0; t < v.T; t++) { maxdz = std::max(maxdz, v.sens[t].nd_nz); vw = std::max(vw, MVBA_VPOSE + 3 + v.sens[t].nd_nz); }
  vw = std::max(vw, MVBA_VPOSE + v.nd2i + 1);
  s->maxdz = maxdz; s->vwmax = vw;
  MVBA_CUDA(s->pairtab.upload(pairtab)); MVBA_CUDA(s->row_ptr.upload(row_ptr)); MVBA_CUDA(s->col.upload(col)); MVBA_CUDA(s->slot.upload(slot));
  MVBA_CUDA(s->xcol.upload(h->h_xcol));
  {  // parameter-block id of every border column (Ceres' block-Jacobi partition)
    std::vector<int32_t> blk(std::max(s->B, 1), 0);
    int id = 0;
    for (int t = 0; t < v.T; t++) {
      const SensorDev& sd = v.sens[t];
      const int c0[7] = {sd.c_rig, sd.c_off, sd.c_focal, sd.c_center, sd.c_dist, sd.c_d2i, sd.c_scale};
      const int w[7] = {7, 1, 1, 2, sd.nd, v.nd2i, 1};
      for (int k = 0; k < 7; k++) if (c0[k] >= 0) { for (int j = 0; j < w[k]; j++) blk[c0[k] - s->npc + j] = id; id++; }
    }
    MVBA_CUDA(s->border_blk.upload(blk));
  }
  if (h->bounded) { MVBA_CUDA(s->lo.upload(h->h_lo)); MVBA_CUDA(s->hi.upload(h->h_hi)); }
  // one contiguous buffer for everything that is summed across ranks
  size_t o = 0;
  auto take = [&](size_t n) { size_t r = o; o += (n + 1) & ~(size_t)1; return r; };
  s->o_Spp = take((size_t)ns * 49); s->o_Spb = take((size_t)s->npc * s->B); s->o_Sbb = take((size_t)s->B * s->B);
  s->o_Fpp = take(with_F ? (size_t)nb * 49 : 0); s->o_Fbb = take(with_F ? (size_t)s->B * s->B : 0);
  s->o_g = take(s->nf); s->o_diagF = take(s->nf); s->o_rhs = take(s->nf); s->o_cost = take(2);
  s->o_gmaxe = take(std::max(h->nranks, 1));  // one slot per rank: the landmark-part gradient maximum rides in the summed buffer
  s->red_n = o;
  MVBA_CUDA(s->red.alloc(o));
  const int nf = std::max(s->nf, 1);
  MVBA_CUDA(s->Dsq.alloc(nf)); MVBA_CUDA(s->Mpp.alloc((size_t)std::max(nb, 1) * 49)); MVBA_CUDA(s->Mbb.alloc((size_t)std::max(s->B * s->B, 1)));
  MVBA_CUDA(s->work.alloc((size_t)std::max(s->B * s->B, 1)));
  MVBA_CUDA(s->jscale_f.alloc(nf)); MVBA_CUDA(s->jscale_e.alloc((size_t)3 * std::max(v.P, 1)));
  MVBA_CUDA(s->Hinv.alloc((size_t)6 * std::max(v.P, 1))); MVBA_CUDA(s->ge.alloc((size_t)3 * std::max(v.P, 1)));
  MVBA_CUDA(s->diagE.alloc((size_t)3 * std::max(v.P, 1))); MVBA_CUDA(s->ye.alloc((size_t)3 * std::max(v.P, 1)));
  MVBA_CUDA(s->y.alloc(nf)); MVBA_CUDA(s->r.alloc(nf)); MVBA_CUDA(s->z.alloc(nf)); MVBA_CUDA(s->p.alloc(nf)); MVBA_CUDA(s->p2.alloc(nf)); MVBA_CUDA(s->q.alloc(nf));
  MVBA_CUDA(s->lm.alloc(LM_NSLOTS)); MVBA_CUDA(s->pcg_flags.alloc(4));
  MVBA_CUDA(cudaMallocHost((void**)&s->rec, sizeof(mvba_iteration)));
  MVBA_CUDA(cudaMallocHost((void**)&s->flags_host, 4 * sizeof(int)));
  Reduced& R = s->R;
  double* rb = s->red.p;
  R.Spp = rb + s->o_Spp; R.Spb = rb + s->o_Spb; R.Sbb = rb + s->o_Sbb;
  R.Fpp = with_F ? rb + s->o_Fpp : nullptr; R.Fbb = with_F ? rb + s->o_Fbb : nullptr;
  R.g = rb + s->o_g; R.diagF = rb + s->o_diagF; R.rhs = rb + s->o_rhs;
  R.pairtab = s->pairtab.p; R.nb = nb; R.npc = s->npc; R.B = s->B; R.nf = s->nf;
  MVBA_CUDA(s->redA.alloc(o)); s->redA_valid = false;
  s->RA = R;
  {
    double* ra = s->redA.p;
    Reduced& A2 = s->RA;
    A2.Spp = ra + s->o_Spp; A2.Spb = ra + s->o_Spb; A2.Sbb = ra + s->o_Sbb;
    A2.Fpp = with_F ? ra + s->o_Fpp : nullptr; A2.Fbb = with_F ? ra + s->o_Fbb : nullptr;
    A2.g = ra + s->o_g; A2.diagF = ra + s->o_diagF; A2.rhs = ra + s->o_rhs;
  }
  s->A.row_ptr = s->row_ptr.p; s->A.col = s->col.p; s->A.slot = s->slot.p; s->A.nb = nb;
  std::vector<int32_t> cam_items_host;
  {  // ---- pass A: residual blocks with f columns grouped by camera image (counting sort), chunked ----
    const int64_t nit = v.N + v.n_dtri + v.n_dmesh;
    if (nit >= ((int64_t)1 << 31)) return fail(MVBA_EINVAL, "mvba: more than 2^31 residual blocks on one device");
    std::vector<int32_t> cnt(v.C + 1, 0);
    auto cam_of = [&](int64_t code) -> int {
      if (code < v.N) return h->ocid[code];
      if (code < v.N + v.n_dtri) return h->ocid[h->h_dtri_obs[code - v.N]];
      return h->ocid[h->h_dmesh_obs[code - v.N - v.n_dtri]];
    };
    for (int64_t q = 0; q < nit; q++) cnt[cam_of(q) + 1]++;
    for (int c = 0; c < v.C; c++) cnt[c + 1] += cnt[c];
    std::vector<int32_t> items((size_t)nit), fill(cnt.begin(), cnt.end() - 1);
    for (int64_t q = 0; q < nit; q++) items[fill[cam_of(q)]++] = (int32_t)q;  // codes ascend inside an image
    cam_items_host = items;
    // widest image: active columns + r, padded to blocks of 4
    int maxact = 0;
    for (int t = 0; t < v.T; t++) {
      const SensorDev& sd = v.sens[t];
      int n = (v.npose_cols ? 14 : 0) + (sd.c_rig >= 0 ? 7 : 0) + (sd.c_off >= 0) + (sd.c_focal >= 0) + (sd.c_center >= 0 ? 2 : 0) +
              (sd.c_dist >= 0 ? sd.nd_nz : 0) + (sd.c_d2i >= 0 ? v.nd2i : 0) + (sd.c_scale >= 0);
      maxact = std::max(maxact, n);
    }
    if (maxact + 1 > MVBA_CAM_MAXACT) return fail(MVBA_EINVAL, "mvba: one image touches more than 87 free non-landmark columns");
    // row stride: = 12 (mod 16) doubles when every image fits the 24-column tensor fold (conflict-free DMMA fragment loads)
    s->cam_ldj = ((maxact + 1 + 3) / 4) * 4;
    if (maxact + 1 <= 24) { s->cam_ldj = ((maxact + 1 + 7) / 8) * 8; while ((s->cam_ldj & 15) != 12) s->cam_ldj += 4; }
    const int64_t target = std::max<int64_t>(256, std::min<int64_t>(4096, nit / std::max(1, 4 * h->sm_count)));
    std::vector<int32_t> ccam, cbeg;
    for (int c = 0; c < v.C; c++) {
      int n = cnt[c + 1] - cnt[c];
      if (n == 0) continue;
      int parts = (int)((n + target - 1) / target);
      for (int k = 0; k < parts; k++) { ccam.push_back(c); cbeg.push_back(cnt[c] + (int)((int64_t)n * k / parts)); }
    }
    cbeg.push_back((int32_t)nit);
    s->nchunks = (int)ccam.size();
    MVBA_CUDA(s->citems.upload(items)); MVBA_CUDA(s->chunk_cam.upload(ccam)); MVBA_CUDA(s->chunk_beg.upload(cbeg));
  }
  // launch geometry: as many warps per CTA as shared memory allows, persistent grid
  size_t pw = build_smem_per_warp(s->maxposes, s->B, s->maxdz, s->vwmax);
  int dev_smem = 0;
  MVBA_CUDA(cudaDeviceGetAttribute(&dev_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
  if (pw > (size_t)dev_smem) return fail(MVBA_EINVAL, "mvba: a track touches too many parameter columns for one warp's shared memory");
  s->wpb_build = (int)std::max<size_t>(1, std::min<size_t>(8, (size_t)dev_smem / 2 / pw));
  s->smem_build = pw * s->wpb_build;
  MVBA_CUDA(cudaFuncSetAttribute(build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_build));
  if (s->maxdz > MVBA_MAXDZ) return fail(MVBA_EINVAL, "mvba: distortion block too long");
  MVBA_CUDA(s->camdir.alloc(std::max(v.C, 1))); MVBA_CUDA(s->sensdir.alloc(std::max(v.T, 1)));
  {  // pass A: stage as many blocks per round as ~100 KB allows (two CTAs per SM)
    const size_t budget = std::min<size_t>((size_t)dev_smem, 112 * 1024) - cam_pass_smem(0, s->cam_ldj);
    int rows = (int)(budget / ((size_t)s->cam_ldj * 8));
    rows = std::min(3 * MVBA_CAM_THREADS, (rows / 12) * 12);
    if (rows < 96) return fail(MVBA_EINVAL, "mvba: pass A staging does not fit shared memory");
    s->cam_batch = rows;
    s->smem_cam = cam_pass_smem(rows, s->cam_ldj);
    MVBA_CUDA(cudaFuncSetAttribute(camera_pass_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_cam));
    int occ_cam = 1;
    MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_cam, camera_pass_kernel, MVBA_CAM_THREADS, s->smem_cam));
    s->grid_cam = std::max(1, std::min(s->nchunks, h->sm_count * std::max(occ_cam, 1)));
  }
  {  // pass A, images whose f columns are all pose / rig / offset columns: 27 accumulators per image instead of staged rows
    bool any_x = false;
    for (int t = 0; t < v.T; t++) {
      const SensorDev& sd = v.sens[t];
      any_x = any_x || sd.c_focal >= 0 || sd.c_center >= 0 || sd.c_dist >= 0 || sd.c_d2i >= 0 || sd.c_scale >= 0;
    }
    s->any_x = any_x;
    const char* ev = std::getenv("MVBA_CAM_STAGED");  // developer switch: force the staged-row kernel
    s->cam_fast = !any_x && !(ev && ev[0] == '1');
    if (s->cam_fast) {
      MVBA_CUDA(s->camQ.alloc((size_t)std::max(v.C, 1) * MVBA_CAMQ_N));
      {  // the camera-major block list once more, as 16-byte records the kernel reads coalesced
        std::vector<TSlot> rec(cam_items_host.size());
        for (size_t i = 0; i < cam_items_host.size(); i++) {
          const int64_t code = cam_items_host[i];
          TSlot sl{};
          if (code < v.N) { sl.a = h->opid[code]; sl.b = 0; sl.px = h->opix[2 * (size_t)code]; sl.py = h->opix[2 * (size_t)code + 1]; }
          else if (code < v.N + v.n_dtri) { sl.a = (int32_t)(code - v.N); sl.b = 1u << 24; }
          else { sl.a = (int32_t)(code - v.N - v.n_dtri); sl.b = 2u << 24; }
          rec[i] = sl;
        }
        if (rec.empty()) rec.push_back(TSlot{});
        MVBA_CUDA(s->crec.upload(rec));
      }
      int occ_q = 1;
      MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_q, camq_kernel, 256, 0));
      s->grid_camq = std::max(1, std::min(s->nchunks, h->sm_count * std::max(occ_q, 1)));
      int occ_c = 1;
      MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_c, camcost_kernel, 256, 0));
      s->grid_camcost = std::max(1, std::min(s->nchunks, h->sm_count * std::max(occ_c, 1)));
    }
  }
  {  // ---- pass B fast path: window tiles ----
    // 16 warps x <= 16 accumulator blocks of 8x8 cover the upper block triangle up to L8 = 176 local columns
    int maxspan = 0;
    for (int p = 0; p < v.P; p++) if (trk_any[p]) maxspan = std::max(maxspan, (int)trk_span[p]);
    // staged kernel (tile2): 8 warps x <= 16 accumulator blocks cover L8 = 120 local columns including the v row
    bool use_t2 = !s->any_x && v.C < (1 << 24);
    { const char* ev = std::getenv("MVBA_TILE_UNSTAGED"); if (ev && ev[0] == '1') use_t2 = false; }  // developer switch
    const int capw2 = (119 - s->B) / 7;
    if (use_t2) {  // nearly every track must fit its window, otherwise the wider windows of the unstaged kernel win
      int64_t wide = 0, any = 0;
      for (int p = 0; p < v.P; p++) if (trk_any[p]) { any++; wide += (trk_span[p] > capw2 || trk_items[p] > 32); }
      if (s->B > 119 || capw2 < 1 || wide * 100 > any) use_t2 = false;
    }
    int capw = use_t2 ? capw2 : (176 - s->B) / 7;
    if (const char* ev = std::getenv("MVBA_TILE_WCAP")) { const int w = std::atoi(ev); if (!use_t2 && w >= 1) capw = std::min(capw, w); }  // developer switch
    int W = std::min(maxspan + 2, capw);
    if (s->B > 176 || W < 0) W = -1;  // nothing fits: every track takes the generic kernel
    std::vector<int32_t> order, fb;
    for (int p = 0; p < v.P; p++) if (trk_any[p]) { if (W >= 0 && trk_span[p] <= W && trk_items[p] <= 32) order.push_back(p); else fb.push_back(p); }
    // by first pose block, then by span: tracks of one tile then need the same window, and narrow tracks get narrow windows
    // (the dense window update costs (7 Wt + B)^2 per track whatever the track touches)
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return trk_umin[a] != trk_umin[b] ? trk_umin[a] < trk_umin[b] : trk_span[a] < trk_span[b]; });
    std::vector<int32_t> t0, tw0, twn;
    // at least ~6 tiles per CTA so that the static tile -> CTA assignment balances (small shards of a multi-GPU run)
    int tile_div = 6;  // tiles per SM (tuning knob: MVBA_TILE_DIV)
    if (const char* ev = std::getenv("MVBA_TILE_DIV")) { const int d = std::atoi(ev); if (d >= 1 && d <= 64) tile_div = d; }
    const int tile_max = (int)std::max<size_t>(48, std::min<size_t>(512, order.size() / (size_t)(tile_div * (use_t2 ? 2 : 1) * std::max(h->sm_count, 1))));
    int tile_min = std::min(tile_max, 128);  // tracks a tile must hold before a window growth may close it (MVBA_TILE_MIN)
    if (const char* ev = std::getenv("MVBA_TILE_MIN")) { const int d = std::atoi(ev); if (d >= 1) tile_min = std::min(tile_max, d); }
    size_t i = 0;
    int wn_max = 0;
    while (i < order.size()) {
      const int w0 = trk_umin[order[i]];
      int hi = w0;  // one past the last pose block used
      size_t j = i;
      while (j < order.size() && (int)(j - i) < tile_max) {
        int p = order[j], e = trk_umin[p] + trk_span[p];
        if (trk_span[p] > 0 && e - w0 > W) break;
        if (trk_span[p] > 0 && e > hi && hi > w0 && (int)(j - i) >= tile_min) break;  // a wider track starts the next tile
        if (trk_span[p] > 0) hi = std::max(hi, e);
        j++;
      }
      t0.push_back((int32_t)i); tw0.push_back(w0); twn.push_back(hi - w0);
      wn_max = std::max(wn_max, hi - w0);
      i = j;
    }
    t0.push_back((int32_t)order.size());
    s->ntiles = (int)tw0.size(); s->nfallback = (int)fb.size();
    s->tile_L8max = std::max(8, (7 * wn_max + s->B + 7) & ~7);
    MVBA_CUDA(s->tile_w0.upload(tw0)); MVBA_CUDA(s->tile_wn.upload(twn)); MVBA_CUDA(s->fallback.upload(fb));
    // shared-memory plan: <= icap staged residual blocks and <= K staged tracks per round
    int nx = 0;
    for (int t = 0; t < v.T; t++) {
      const SensorDev& sd = v.sens[t];
      nx = std::max(nx, (sd.c_focal >= 0) + (sd.c_center >= 0 ? 2 : 0) + (sd.c_dist >= 0 ? sd.nd_nz : 0));
      nx = std::max(nx, (sd.c_d2i >= 0 ? v.nd2i : 0) + (sd.c_scale >= 0));
    }
    s->tile_nx = nx;
    double mean_items = 1.0;
    { double tot = 0; for (int p : order) tot += trk_items[p]; if (!order.empty()) mean_items = std::max(1.0, tot / order.size()); }
    // image range of every tile (the staged kernel copies the window's CamPack rows to shared memory in one piece)
    std::vector<int32_t> tc0(s->ntiles, 0), tnc(s->ntiles, 0);
    if (use_t2) {
      int maxnc = 0;
      for (int t = 0; t < s->ntiles; t++) {
        int lo = INT32_MAX, hi = -1;
        for (int ti = t0[t]; ti < t0[t + 1]; ti++) {
          const int p = order[ti];
          auto see = [&](int64_t obs) { const int c = h->ocid[obs]; lo = std::min(lo, c); hi = std::max(hi, c); };
          for (int64_t i = h->h_trk_obs0[p]; i < h->h_trk_obs0[p + 1]; i++) see(i);
          for (int64_t j = h->h_trk_dtri0[p]; j < h->h_trk_dtri0[p + 1]; j++) see(h->h_dtri_obs[j]);
          for (int64_t j = h->h_trk_dmesh0[p]; j < h->h_trk_dmesh0[p + 1]; j++) see(h->h_dmesh_obs[j]);
        }
        if (hi >= 0) { tc0[t] = lo; tnc[t] = hi - lo + 1; maxnc = std::max(maxnc, tnc[t]); }
      }
      // two CTAs per SM: half of the SM's shared memory each
      int sm_smem = 0;
      MVBA_CUDA(cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, h->device));
      const size_t budget = std::min<size_t>((size_t)dev_smem, ((size_t)sm_smem - 2 * 1024) / 2);
      s->t2_L8max = std::max(8, (7 * wn_max + s->B + 1 + 7) & ~7);
      s->t2_ccap = std::min(maxnc, 40);
      int K = 32, icap = 0, bcap = 0;
      for (; K >= 1; K--) {
        icap = (int)std::min<double>((double)MVBA_T2_THREADS, std::max(32.0, std::ceil(1.15 * mean_items * K)));
        icap = std::min(MVBA_T2_THREADS, (icap + 1) & ~1);
        bcap = (int)t2_blob_bytes(K, K, icap, MVBA_T2_GCAP);
        if (tile2_smem(K, icap, s->t2_ccap, s->t2_L8max, bcap) <= budget) break;
      }
      if (K < 4) use_t2 = false;  // (the window cap above was the staged kernel's: still valid for the unstaged one)
      else {
        while (icap + 2 <= MVBA_T2_THREADS && icap < 2 * mean_items * K &&
               tile2_smem(K, icap + 2, s->t2_ccap, s->t2_L8max, (int)t2_blob_bytes(K, K, icap + 2, MVBA_T2_GCAP)) <= budget) icap += 2;
        s->t2_K = K; s->t2_icap = icap; s->t2_blob_cap = (int)t2_blob_bytes(K, K, icap, MVBA_T2_GCAP);
        for (int t = 0; t < s->ntiles; t++) if (tnc[t] > s->t2_ccap) tnc[t] = 0;  // range too wide: that tile reads the global table
      }
    }
    s->tile2 = use_t2;
    s->tile_direct = 0;
    if (use_t2) { s->tile_K = s->t2_K; s->tile_icap = s->t2_icap; }
    else {
      // the largest K (tracks per round) whose staging fits, with room for ~1.15 x the mean number of blocks per track
      // wide free blocks (an RPC distortion block: 38 columns, 1.2 KB of staging per residual block) go from the evaluating thread
      // straight into the round's G (shared-memory atomics) instead of through a staging buffer that would hold the round to a few tracks
      bool direct = nx >= 16;
      if (const char* ev = std::getenv("MVBA_TILE_DIRECT")) direct = nx > 0 && ev[0] == '1';   // developer switch
      s->tile_direct = direct;
      s->tile_nmono = 0;
      if (direct)
        for (int t = 0; t < v.T; t++) {
          const SensorDev& sd = v.sens[t];
          if (sd.rpc_deg > 0 && (sd.c_dist >= 0 || sd.c_focal >= 0)) s->tile_nmono = std::max(s->tile_nmono, (sd.rpc_deg + 1) * (sd.rpc_deg + 2) / 2);
        }
      int K = 32, icap = 0;  // phase b3 maps the staged tracks to the lanes of a warp
      for (; K >= 1; K--) {
        icap = (int)std::min<double>(511.0, std::max(32.0, std::ceil(1.15 * mean_items * K)));
        if (tile_smem(K, icap, nx, s->tile_L8max, direct) + 1024 <= (size_t)dev_smem) break;
      }
      if (K < 1) return fail(MVBA_EINVAL, "mvba: tile staging does not fit shared memory");
      while (icap < 511 && tile_smem(K, icap + 1, nx, s->tile_L8max, direct) + 1024 <= (size_t)dev_smem && icap < 2 * mean_items * K) icap++;
      s->tile_K = K; s->tile_icap = icap;
      s->smem_tile = tile_smem(s->tile_K, s->tile_icap, nx, s->tile_L8max, direct);
      // accumulator blocks per warp: upper block triangle of the widest window over 16 warps.  Beyond 8 per warp the 512-thread
      // kernel (128 registers per thread) keeps part of its accumulators in local memory and every DMMA of those slots is a
      // load / store round trip: such windows run on 256 threads (255 registers) with 24 or 32 blocks per warp
      const int nb8 = s->tile_L8max / 8, nblk = nb8 * (nb8 + 1) / 2, per = (nblk + 15) / 16, per8 = (nblk + 7) / 8;
      s->tile_nt = 512;
      s->tile_pr = per <= 4 ? 4 : per <= 8 ? 8 : per <= 12 ? 12 : 16;
      bool wide256 = per > 8 && per8 <= 32;
      if (const char* ev = std::getenv("MVBA_TILE_256")) wide256 = wide256 && ev[0] != '0';   // developer switch
      if (wide256) { s->tile_nt = 256; s->tile_pr = per8 <= 24 ? 24 : 32; }
      if (s->tile_nt == 256 && s->tile_pr == 24) MVBA_CUDA(cudaFuncSetAttribute(tile_kernel<256, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile));
      else if (s->tile_nt == 256) MVBA_CUDA(cudaFuncSetAttribute(tile_kernel<256, 32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile));
      else if (s->tile_pr == 4) MVBA_CUDA(cudaFuncSetAttribute(tile_kernel<512, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile));
      else if (s->tile_pr == 8) MVBA_CUDA(cudaFuncSetAttribute(tile_kernel<512, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile));
      else if (s->tile_pr == 12) MVBA_CUDA(cudaFuncSetAttribute(tile_kernel<512, 12>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile));
      else MVBA_CUDA(cudaFuncSetAttribute(tile_kernel<512, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile));
    }
    const int gcap = use_t2 ? MVBA_T2_GCAP : INT32_MAX;  // gather groups of one round
    // ---- static round tables ----
    std::vector<int32_t> tile_round0{0}, round_trk0{0}, round_blk0{0}, round_grp0{0};
    std::vector<TBlk> blk; std::vector<TTrk> ttrk; std::vector<TGrp> grp; std::vector<uint16_t> con;
    blk.reserve((size_t)(v.N + v.n_dtri + v.n_dmesh + v.n_prior)); ttrk.reserve(order.size());
    struct Key { int lcol; int single; std::vector<uint16_t> c; };
    std::vector<Key> keys;
    std::vector<int> gkeys;
    auto count_groups = [&](int p) -> int {  // distinct (pose block | rig transform | offset) blocks touched by the track's f blocks
      gkeys.clear();
      auto add = [&](int key) { if (std::find(gkeys.begin(), gkeys.end(), key) == gkeys.end()) gkeys.push_back(key); };
      auto f_block = [&](int64_t obs) {
        const int c = h->ocid[obs];
        const SensorDev& sd = v.sens[h->cam_sensor[c]];
        const int p0 = h->h_cam_p0[c], p1 = h->h_cam_p1[c];
        const int cb = h->h_pose_col[p0], ce = p1 >= 0 ? h->h_pose_col[p1] : -1;
        if (cb >= 0) add(cb);
        if (ce >= 0) add(ce);
        if (sd.c_rig >= 0) add(sd.c_rig);
        if (sd.c_off >= 0) add(sd.c_off);
      };
      for (int64_t i = h->h_trk_obs0[p]; i < h->h_trk_obs0[p + 1]; i++) f_block(i);
      for (int64_t j = h->h_trk_dtri0[p]; j < h->h_trk_dtri0[p + 1]; j++) f_block(h->h_dtri_obs[j]);
      for (int64_t j = h->h_trk_dmesh0[p]; j < h->h_trk_dmesh0[p + 1]; j++) f_block(h->h_dmesh_obs[j]);
      return (int)gkeys.size();
    };
    for (int t = 0; t < s->ntiles; t++) {
      const int w0 = tw0[t], Wt = twn[t];
      int ti = t0[t];
      while (ti < t0[t + 1]) {
        int kc = 0, nit = 0, ng = 0;
        // staged kernel: the slots of a track must be lanes of one warp (its landmark block is summed with shuffles), so the slot
        // list is padded to the next multiple of 32 where a track would straddle
        auto slot_of = [&](int at, int n) { return use_t2 && (at & 31) + n > 32 ? (at + 31) & ~31 : at; };
        while (ti + kc < t0[t + 1] && kc < s->tile_K && slot_of(nit, trk_items[order[ti + kc]]) + trk_items[order[ti + kc]] <= s->tile_icap) {
          if (use_t2) {  // one gather thread per group: count the track's distinct parameter blocks
            const int n = count_groups(order[ti + kc]);
            if (ng + n > gcap) break;
            ng += n;
          }
          nit = slot_of(nit, trk_items[order[ti + kc]]) + trk_items[order[ti + kc]]; kc++;
        }
        if (kc == 0) return fail(MVBA_EINVAL, "mvba: a tile track does not fit the staging buffer");
        int q = 0;
        for (int k = 0; k < kc; k++) {
          const int p = order[ti + k];
          for (const int q1 = slot_of(q, trk_items[p]); q < q1; q++) blk.push_back(TBlk{0, 7});  // padding slots (kind 7: nothing to evaluate)
          ttrk.push_back(TTrk{p, (int16_t)q, (int16_t)trk_items[p]});
          keys.clear();
          auto contribute = [&](int lcol, int single, int slot, int role) {
            size_t i = 0;
            while (i < keys.size() && keys[i].lcol != lcol) i++;
            if (i == keys.size()) keys.push_back(Key{lcol, single, {}});
            keys[i].c.push_back((uint16_t)((slot << 2) | role));
          };
          auto f_block = [&](int64_t obs, int slot) {
            const int c = h->ocid[obs];
            const SensorDev& sd = v.sens[h->cam_sensor[c]];
            const int p0 = h->h_cam_p0[c], p1 = h->h_cam_p1[c];
            const int cb = h->h_pose_col[p0], ce = p1 >= 0 ? h->h_pose_col[p1] : -1;
            if (cb >= 0) contribute(cb - 7 * w0, 0, slot, 0);
            if (ce >= 0) contribute(ce - 7 * w0, 0, slot, 1);
            if (sd.c_rig >= 0) contribute(7 * Wt + (sd.c_rig - s->npc), 0, slot, 2);
            if (sd.c_off >= 0) contribute(7 * Wt + (sd.c_off - s->npc), 1, slot, 3);
          };
          const int kslot = s->tile_direct ? (k << 8) : 0;   // direct mode: the evaluating thread needs the track's position in the round
          for (int64_t i = h->h_trk_obs0[p]; i < h->h_trk_obs0[p + 1]; i++) { blk.push_back(TBlk{(int32_t)i, 0 | kslot}); f_block(i, q++); }
          for (int64_t j = h->h_trk_dtri0[p]; j < h->h_trk_dtri0[p + 1]; j++) { blk.push_back(TBlk{(int32_t)j, 1 | kslot}); f_block(h->h_dtri_obs[j], q++); }
          for (int64_t j = h->h_trk_dmesh0[p]; j < h->h_trk_dmesh0[p + 1]; j++) { blk.push_back(TBlk{(int32_t)j, 2 | kslot}); f_block(h->h_dmesh_obs[j], q++); }
          for (int64_t j = h->h_trk_pm0[p]; j < h->h_trk_pm0[p + 1]; j++) { blk.push_back(TBlk{(int32_t)j, 3}); q++; }
          for (int64_t j = h->h_trk_pt0[p]; j < h->h_trk_pt0[p + 1]; j++) { blk.push_back(TBlk{(int32_t)j, 4}); q++; }
          for (const Key& key : keys) {
            if (con.size() >= ((size_t)1 << 32) - 64) return fail(MVBA_EINVAL, "mvba: gather table too large");
            TGrp g{};
            g.over0 = (uint32_t)con.size(); g.lcol = (uint16_t)key.lcol; g.k = (uint8_t)k;
            g.ncon = (uint8_t)(key.c.size() | (key.single ? 0x80 : 0));
            for (size_t u = 0; u < key.c.size(); u++) { if (u < 4) g.c[u] = key.c[u]; else con.push_back(key.c[u]); }
            grp.push_back(g);
          }
        }
        ti += kc;
        round_trk0.push_back((int32_t)ttrk.size()); round_blk0.push_back((int32_t)blk.size()); round_grp0.push_back((int32_t)grp.size());
      }
      tile_round0.push_back((int32_t)round_trk0.size() - 1);
    }
    s->nrounds = (int)round_trk0.size() - 1;
    if (con.empty()) con.push_back(0);
    if (grp.empty()) grp.push_back(TGrp{});
    if (blk.empty()) blk.push_back(TBlk{0, 0});
    if (ttrk.empty()) ttrk.push_back(TTrk{0, 0, 0});
    MVBA_CUDA(s->t_con.upload(con));
    if (!use_t2) {
      MVBA_CUDA(s->tile_round0.upload(tile_round0)); MVBA_CUDA(s->round_trk0.upload(round_trk0)); MVBA_CUDA(s->round_blk0.upload(round_blk0));
      MVBA_CUDA(s->round_grp0.upload(round_grp0)); MVBA_CUDA(s->t_blk.upload(blk)); MVBA_CUDA(s->t_trk.upload(ttrk)); MVBA_CUDA(s->t_grp.upload(grp));
      s->grid_tile = std::max(1, std::min(s->ntiles, h->sm_count));
    } else {
      // ---- one contiguous blob per round: TRoundHdr | TTrk[kc] | next round's landmarks | TSlot[nit] | TGrp[ngrp] ----
      auto rkc = [&](int r) { return round_trk0[r + 1] - round_trk0[r]; };
      auto rbytes = [&](int r, bool last) {
        return t2_blob_bytes(rkc(r), last ? 0 : rkc(r + 1), round_blk0[r + 1] - round_blk0[r], round_grp0[r + 1] - round_grp0[r]);
      };
      std::vector<int32_t> trounds(s->ntiles, 0);
      std::vector<int64_t> tblob0(s->ntiles, 0);
      std::vector<uint32_t> tbytes0(s->ntiles, 0);
      size_t total = 0;
      uint32_t maxbytes = 64;
      for (int t = 0; t < s->ntiles; t++)
        for (int r = tile_round0[t]; r < tile_round0[t + 1]; r++) { const uint32_t nb = rbytes(r, r + 1 == tile_round0[t + 1]); total += nb; maxbytes = std::max(maxbytes, nb); }
      if ((size_t)maxbytes > (size_t)s->t2_blob_cap) return fail(MVBA_EINVAL, "mvba: round table larger than planned");
      std::vector<uint8_t> blob(std::max<size_t>(total, 16), 0);
      std::vector<int> rank_of, perm;
      size_t off = 0;
      for (int t = 0; t < s->ntiles; t++) {
        trounds[t] = tile_round0[t + 1] - tile_round0[t];
        tblob0[t] = (int64_t)off;
        for (int r = tile_round0[t]; r < tile_round0[t + 1]; r++) {
          const bool last = r + 1 == tile_round0[t + 1];
          const int kc = rkc(r), kn = last ? 0 : rkc(r + 1), nit = round_blk0[r + 1] - round_blk0[r], ng = round_grp0[r + 1] - round_grp0[r];
          const uint32_t nb = rbytes(r, last);
          if (r == tile_round0[t]) tbytes0[t] = nb;
          uint8_t* b = blob.data() + off;
          TRoundHdr hd{};
          hd.kc = kc; hd.nit = nit; hd.ngrp = ng; hd.kc_next = kn; hd.next_bytes = last ? 0 : rbytes(r + 1, r + 2 == tile_round0[t + 1]);
          std::memcpy(b, &hd, sizeof(hd));
          uint8_t* q = b + 32;
          std::memcpy(q, ttrk.data() + round_trk0[r], (size_t)kc * sizeof(TTrk)); q += t2_align16(8u * kc);
          for (int k = 0; k < kn; k++) { const int32_t pn = ttrk[round_trk0[r + 1] + k].p; std::memcpy(q + 4 * k, &pn, 4); }
          q += t2_align16(4u * kn);
          for (int i = 0; i < nit; i++) { TSlot pad{}; pad.b = 7u << 24; std::memcpy(q + 16 * (size_t)i, &pad, 16); }
          for (int k = 0; k < kc; k++) {
            const TTrk& tk = ttrk[round_trk0[r] + k];
            for (int i = 0; i < tk.n; i++) {
              const TBlk& bk = blk[round_blk0[r] + tk.q0 + i];
              TSlot sl{};
              int cid = 0;
              if (bk.kind == 0) { cid = h->ocid[bk.idx]; sl.a = h->opid[bk.idx]; sl.px = h->opix[2 * (size_t)bk.idx]; sl.py = h->opix[2 * (size_t)bk.idx + 1]; }
              else { sl.a = bk.idx; if (bk.kind == 1) cid = h->ocid[h->h_dtri_obs[bk.idx]]; else if (bk.kind == 2) cid = h->ocid[h->h_dmesh_obs[bk.idx]]; }
              sl.b = (uint32_t)cid | ((uint32_t)bk.kind << 24) | ((uint32_t)k << 27);
              std::memcpy(q + 16 * (size_t)(tk.q0 + i), &sl, 16);
            }
          }
          q += 16 * (size_t)nit;
          {  // groups ordered by their rank inside their track, then by track: neighbouring gather threads then write different
             // COLUMNS of Gs (tracks), which spreads over the banks; neighbouring groups of one track would write rows 7 apart
             // of the same column, 8-way conflicts for a row stride = 4 (mod 16)
            rank_of.assign(ng, 0); perm.resize(ng);
            int cnt[32] = {0};
            for (int g = 0; g < ng; g++) { rank_of[g] = cnt[grp[round_grp0[r] + g].k & 31]++; perm[g] = g; }
            std::stable_sort(perm.begin(), perm.end(), [&](int x, int y) { return rank_of[x] < rank_of[y]; });
            for (int g = 0; g < ng; g++) std::memcpy(q + 16 * (size_t)g, &grp[round_grp0[r] + perm[g]], sizeof(TGrp));
          }
          off += nb;
        }
      }
      MVBA_CUDA(s->t2_blob.upload(blob)); MVBA_CUDA(s->t2_blob0.upload(tblob0)); MVBA_CUDA(s->t2_bytes0.upload(tbytes0));
      MVBA_CUDA(s->t2_rounds.upload(trounds)); MVBA_CUDA(s->t2_c0.upload(tc0)); MVBA_CUDA(s->t2_nc.upload(tnc));
      MVBA_CUDA(s->tile_ctr.alloc(1));
      MVBA_CUDA(h->d_pack.alloc(std::max(v.C, 1)));
      s->smem_tile2 = tile2_smem(s->t2_K, s->t2_icap, s->t2_ccap, s->t2_L8max, s->t2_blob_cap);
      const int nb8 = s->t2_L8max / 8, per = (nb8 * (nb8 + 1) / 2 + 7) / 8;
      s->t2_maxs = 16;
      for (int m : {4, 5, 6, 7, 8, 10, 12, 16}) if (m >= per) { s->t2_maxs = m; break; }
      int occ2 = 1;
#define MVBA_T2_PLAN(M)                                                                                                                   \
  case M:                                                                                                                                 \
    MVBA_CUDA(cudaFuncSetAttribute(tile2_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s->smem_tile2));                      \
    MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ2, tile2_kernel<M>, MVBA_T2_THREADS, s->smem_tile2));                       \
    break;
      switch (s->t2_maxs) { MVBA_T2_PLAN(4) MVBA_T2_PLAN(5) MVBA_T2_PLAN(6) MVBA_T2_PLAN(7) MVBA_T2_PLAN(8) MVBA_T2_PLAN(10) MVBA_T2_PLAN(12) MVBA_T2_PLAN(16) }
#undef MVBA_T2_PLAN
      s->grid_tile2 = std::max(1, std::min(s->ntiles, h->sm_count * std::max(occ2, 1)));
    }
  }
  int occ = 1;
  MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, build_kernel, 32 * s->wpb_build, s->smem_build));
  s->grid_build = std::max(1, std::min((v.P + s->wpb_build - 1) / s->wpb_build, h->sm_count * std::max(occ, 1)));
  s->grid_backsub = std::max(1, std::min((v.P + 127) / 128, h->sm_count * 16));
  {  // single-launch PCG when the device can co-schedule the grid (MVBA_PCG_STEPS=1 forces the step kernels; developer switch)
    int coop = 0, occ_pcg = 0;
    MVBA_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device));
    if (const char* evt = std::getenv("MVBA_PCG_THREADS")) { const int t = std::atoi(evt); if (t == 128 || t == 256 || t == 512) s->pcg_threads = t; }
    MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_pcg, pcg_fused_kernel, s->pcg_threads, 0));
    const char* ev = std::getenv("MVBA_PCG_STEPS");
    if (coop && occ_pcg > 0 && s->nf > 0 && !(ev && ev[0] == '1')) {
      const int want = std::max((s->nb + s->B + 7) / 8, (s->nf + 255) / 256);
      s->pcg_coop_grid = std::max(1, std::min(want, std::min(occ_pcg, 2) * h->sm_count));
      if (const char* evg = std::getenv("MVBA_PCG_GRID")) { const int g = std::atoi(evg); if (g >= 1) s->pcg_coop_grid = std::min(s->pcg_coop_grid, g); }
      // the reduced system resident in shared memory across the grid (a few MB over up to 296 CTAs): S is then read from L2
      // once per solve instead of once per CG iteration.  MVBA_PCG_RESIDENT=0 keeps it in global memory (developer switch)
      const char* evr = std::getenv("MVBA_PCG_RESIDENT");
      if (nb > 0 && !(evr && evr[0] == '0')) {
        for (int grid : {s->pcg_coop_grid, h->sm_count}) {
          grid = std::max(1, std::min(grid, s->pcg_coop_grid));
          const int rpc = (nb + grid - 1) / grid;
          size_t need = 0;
          for (int r0 = 0; r0 < nb; r0 += rpc) {
            const int r1 = std::min(nb, r0 + rpc);
            need = std::max(need, ((size_t)(row_ptr[r1] - row_ptr[r0]) * 49 + (size_t)7 * (r1 - r0) * s->B) * 8);
          }
          if (need + 1024 > (size_t)dev_smem) continue;
          if (cudaFuncSetAttribute(pcg_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need) != cudaSuccess) { cudaGetLastError(); continue; }
          int occ_r = 0;
          MVBA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_r, pcg_fused_kernel, s->pcg_threads, need));
          if (occ_r * h->sm_count < grid) continue;
          s->pcg_coop_grid = grid; s->pcg_rows_per_cta = rpc; s->smem_pcg = need;
          break;
        }
      }
    }
  }
  return agree_end.pass();
}

}  // namespace mvba

#include "solver_loop.cuh"

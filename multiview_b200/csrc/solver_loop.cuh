// solver_loop.cuh -- the LM iteration itself (see solver.cuh for the overview).
#pragma once

namespace mvba {

struct Timers {
  struct Iv { cudaEvent_t a, b; int phase; };
  std::vector<Iv> iv;
  std::vector<size_t> open;
  cudaStream_t s;
  explicit Timers(cudaStream_t st) : s(st) {}
  ~Timers() { for (auto& i : iv) { cudaEventDestroy(i.a); if (i.b) cudaEventDestroy(i.b); } }
  void begin(int phase) { Iv i{nullptr, nullptr, phase}; cudaEventCreate(&i.a); cudaEventRecord(i.a, s); iv.push_back(i); open.push_back(iv.size() - 1); }
  void end() { Iv& i = iv[open.back()]; open.pop_back(); cudaEventCreate(&i.b); cudaEventRecord(i.b, s); }
  double total(int phase) {
    double t = 0;
    for (auto& i : iv) if (i.phase == phase && i.b) { float ms = 0; cudaEventElapsedTime(&ms, i.a, i.b); t += ms; }
    return t * 1e-3;
  }
};

// mode 0: full fused build at x for the given radius; mode 1: squared column norms only.
// reuse_A: the point has not moved since the previous mode-0 build (a rejected step): pass A's sums are taken from the copy kept then.
// tables_current: the per-image tables on the device already belong to this point (the candidate of the previous iteration was
// accepted: its cost evaluation prepared them).
inline int run_build(mvba_problem* h, const mvba_solve_options& o, double radius, int mode, bool scaled, bool want_cost = true, bool reuse_A = false,
                     bool tables_current = false) {
  Solver* s = h->solver;
  const View& v = h->view;
  reuse_A = reuse_A && s->redA_valid && mode == 0;
  if (!reuse_A) MVBA_CUDA(cudaMemsetAsync(s->redA.p, 0, s->red_n * sizeof(double), h->stream));
  if (!tables_current) run_prep(h, h->d_x.p);
  BuildArgs a{};
  a.x = h->d_x.p; a.tab = h->d_tab.p; a.dtab = h->d_dtab.p;
  a.jscale_f = scaled ? s->jscale_f.p : nullptr; a.jscale_e = scaled ? s->jscale_e.p : nullptr;
  a.radius = radius; a.min_diag = o.min_lm_diagonal; a.max_diag = o.max_lm_diagonal;
  a.Hinv = s->Hinv.p; a.ge = s->ge.p; a.diagE = s->diagE.p; a.cost = s->red.p + s->o_cost;
  a.mode = mode; a.want_cost = want_cost ? 1 : 0; a.maxposes = s->maxposes; a.maxdz = s->maxdz; a.vwmax = s->vwmax;
  a.trk_list = nullptr; a.ntrk = v.P;
  if (s->tm) s->tm->begin(9);
  if (reuse_A) {
  } else if (s->nchunks > 0 && s->nf > 0 && s->cam_fast) {  // pass A, pose-type columns only: per-image 6 x 6 accumulators, then T'QT
    MVBA_CUDA(s->camQ.zero(h->stream));
    CamQArgs qa{};
    qa.x = h->d_x.p; qa.tab = h->d_tab.p; qa.dtab = h->d_dtab.p;
    qa.chunk_cam = s->chunk_cam.p; qa.chunk_beg = s->chunk_beg.p; qa.citems = s->citems.p; qa.crec = s->crec.p; qa.nchunks = s->nchunks; qa.Q = s->camQ.p;
    camq_kernel<<<s->grid_camq, 256, 0, h->stream>>>(v, qa);
    CamFoldArgs fa{};
    fa.tab = h->d_tab.p; fa.coef = h->d_coef.p; fa.Q = s->camQ.p; fa.jscale_f = a.jscale_f; fa.mode = mode;
    camfold_kernel<<<(v.C + 3) / 4, 128, 0, h->stream>>>(v, fa, s->RA);
    h->launches += 2;
  } else if (s->nchunks > 0 && s->nf > 0) {  // pass A: F'F, F'r, diag(F'F) by camera image
    CamPassArgs ca{};
    ca.x = h->d_x.p; ca.tab = h->d_tab.p; ca.dtab = h->d_dtab.p; ca.jscale_f = a.jscale_f;
    ca.chunk_cam = s->chunk_cam.p; ca.chunk_beg = s->chunk_beg.p; ca.citems = s->citems.p; ca.nchunks = s->nchunks;
    ca.batch_rows = s->cam_batch; ca.ldj = s->cam_ldj; ca.mode = mode; ca.coef = h->d_coef.p;
    camera_pass_kernel<<<s->grid_cam, MVBA_CAM_THREADS, s->smem_cam, h->stream>>>(v, ca, s->RA);
    h->launches++;
  }
  s->redA_valid = mode == 0;
  MVBA_CUDA(cudaMemcpyAsync(s->red.p, s->redA.p, s->red_n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));  // pass B adds to a copy
  if (v.P > 0 && mode == 1) {  // column norms: diag(E'E) per landmark
    colnorm_e_kernel<<<s->grid_backsub, 128, 0, h->stream>>>(v, h->d_x.p, h->d_tab.p, h->d_dtab.p, s->diagE.p);
    h->launches++;
  } else if (v.P > 0) {  // pass B: landmark blocks, Schur elimination
    if (s->ntiles > 0 && s->tile2) {  // staged variant: per-round blobs by bulk copy, two CTAs per SM, dynamic tile queue
      MVBA_CUDA(cudaMemsetAsync(s->tile_ctr.p, 0, sizeof(int), h->stream));
      Tile2Args ta{};
      ta.b = a; ta.pack = h->d_pack.p; ta.tile_w0 = s->tile_w0.p; ta.tile_wn = s->tile_wn.p; ta.tile_c0 = s->t2_c0.p; ta.tile_nc = s->t2_nc.p;
      ta.tile_rounds = s->t2_rounds.p; ta.tile_blob0 = s->t2_blob0.p; ta.tile_bytes0 = s->t2_bytes0.p; ta.blob = s->t2_blob.p; ta.con = s->t_con.p;
      ta.tile_ctr = s->tile_ctr.p; ta.ntiles = s->ntiles; ta.K = s->t2_K; ta.icap = s->t2_icap; ta.ccap = s->t2_ccap; ta.L8max = s->t2_L8max;
      ta.blob_cap = s->t2_blob_cap;
#define MVBA_T2_LAUNCH(M) case M: tile2_kernel<M><<<s->grid_tile2, MVBA_T2_THREADS, s->smem_tile2, h->stream>>>(v, ta, s->R); break;
      switch (s->t2_maxs) { MVBA_T2_LAUNCH(4) MVBA_T2_LAUNCH(5) MVBA_T2_LAUNCH(6) MVBA_T2_LAUNCH(7) MVBA_T2_LAUNCH(8) MVBA_T2_LAUNCH(10) MVBA_T2_LAUNCH(12) MVBA_T2_LAUNCH(16) }
#undef MVBA_T2_LAUNCH
      h->launches++;
    } else if (s->ntiles > 0) {
      TileArgs ta{};
      ta.b = a; ta.tile_w0 = s->tile_w0.p; ta.tile_wn = s->tile_wn.p; ta.tile_round0 = s->tile_round0.p;
      ta.round_trk0 = s->round_trk0.p; ta.round_blk0 = s->round_blk0.p; ta.round_grp0 = s->round_grp0.p;
      ta.blk = s->t_blk.p; ta.ttrk = s->t_trk.p; ta.grp = s->t_grp.p; ta.con = s->t_con.p;
      ta.ntiles = s->ntiles; ta.K = s->tile_K; ta.icap = s->tile_icap; ta.nx = s->tile_nx; ta.L8max = s->tile_L8max; ta.direct = s->tile_direct; ta.nmono = s->tile_nmono; ta.coef = h->d_coef.p;
      if (s->tile_nt == 256 && s->tile_pr == 24) tile_kernel<256, 24><<<s->grid_tile, 256, s->smem_tile, h->stream>>>(v, ta, s->R);
      else if (s->tile_nt == 256) tile_kernel<256, 32><<<s->grid_tile, 256, s->smem_tile, h->stream>>>(v, ta, s->R);
      else if (s->tile_pr == 4) tile_kernel<512, 4><<<s->grid_tile, 512, s->smem_tile, h->stream>>>(v, ta, s->R);
      else if (s->tile_pr == 8) tile_kernel<512, 8><<<s->grid_tile, 512, s->smem_tile, h->stream>>>(v, ta, s->R);
      else if (s->tile_pr == 12) tile_kernel<512, 12><<<s->grid_tile, 512, s->smem_tile, h->stream>>>(v, ta, s->R);
      else tile_kernel<512, 16><<<s->grid_tile, 512, s->smem_tile, h->stream>>>(v, ta, s->R);
      h->launches++;
    }
    if (s->nfallback > 0) {
      a.trk_list = s->fallback.p; a.ntrk = s->nfallback;
      int grid = std::max(1, std::min((s->nfallback + s->wpb_build - 1) / s->wpb_build, s->grid_build));
      build_kernel<<<grid, 32 * s->wpb_build, s->smem_build, h->stream>>>(v, a, s->R);
      h->launches++;
    }
  }
  if (s->tm) s->tm->end();
#ifdef MVBA_TILE_PROFILE
  if (mode == 0 && s->ntiles > 0 && s->tile2) {
    unsigned long long pr[8], z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaStreamSynchronize(h->stream);
    cudaMemcpyFromSymbol(pr, g_tile_prof, sizeof(pr));
    cudaMemcpyToSymbol(g_tile_prof, z, sizeof(z));
    double tot = 0; for (int i = 0; i < 8; i++) tot += (double)pr[i];
    std::printf("tile2 prof (thread 0 of each CTA, %% of cycles): prologue+flush %.1f | P1 + landmark blocks %.1f | bar %.1f | P2 gather %.1f | (unused) %.1f | bar %.1f | P3 %.1f | bar %.1f  (Mcycles/CTA %.2f)\n",
                100 * pr[0] / tot, 100 * pr[1] / tot, 100 * pr[2] / tot, 100 * pr[3] / tot, 100 * pr[4] / tot, 100 * pr[5] / tot, 100 * pr[6] / tot, 100 * pr[7] / tot,
                tot / s->grid_tile2 * 1e-6);
    static bool once2 = false;
    if (!once2) { once2 = true; std::printf("tile2 plan: K %d icap %d ccap %d L8max %d blob_cap %d maxs %d ntiles %d nrounds %d nfallback %d grid %d smem %zu B %d\n", s->t2_K, s->t2_icap, s->t2_ccap, s->t2_L8max, s->t2_blob_cap, s->t2_maxs, s->ntiles, s->nrounds, s->nfallback, s->grid_tile2, s->smem_tile2, s->B); }
  } else if (mode == 0 && s->ntiles > 0) {
    unsigned long long pr[8], z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    cudaStreamSynchronize(h->stream);
    cudaMemcpyFromSymbol(pr, g_tile_prof, sizeof(pr));
    cudaMemcpyToSymbol(g_tile_prof, z, sizeof(z));
    double tot = 0; for (int i = 0; i < 8; i++) tot += (double)pr[i];
    std::printf("tile prof (thread 0 of each CTA, %% of cycles): b1 %.1f | bar %.1f | b2 %.1f | bar %.1f | b3+bar %.1f | c %.1f | bar %.1f | flush+tile setup %.1f  (Mcycles/CTA %.2f)\n",
                100 * pr[0] / tot, 100 * pr[1] / tot, 100 * pr[2] / tot, 100 * pr[3] / tot, 100 * pr[4] / tot, 100 * pr[5] / tot, 100 * pr[6] / tot, 100 * pr[7] / tot,
                tot / s->grid_tile * 1e-6);
    static bool once = false;
    if (!once) { once = true; std::printf("tile plan: nt %d pr %d K %d icap %d L8max %d ntiles %d nrounds %d nfallback %d grid %d smem %zu B %d\n", s->tile_nt, s->tile_pr, s->tile_K, s->tile_icap, s->tile_L8max, s->ntiles, s->nrounds, s->nfallback, s->grid_tile, s->smem_tile, s->B); }
  }
#endif
  s->build_launches++;
  if (mode == 0 && h->nranks > 1 && v.P > 0) {  // this rank's landmark-part gradient maximum, into its own slot of the summed buffer
    const int64_t ne = 3 * (int64_t)v.P;
    gradmax_kernel<<<(unsigned)std::max<int64_t>(1, std::min<int64_t>((ne + 255) / 256, (int64_t)8 * h->sm_count)), 256, 0, h->stream>>>(0, ne, nullptr, s->ge.p, nullptr, a.jscale_e, nullptr, 0,
                                                                         s->red.p + s->o_gmaxe + h->rank);
    h->launches++;
  }
  MVBA_CUDA(cudaGetLastError());
  return allreduce(h, s->red.p, s->red_n, 8, 0);
}

// Preconditioned CG on the explicit reduced system; returns the iteration count.
inline int run_pcg(mvba_problem* h, const mvba_solve_options& o, int* iters_out) {
  Solver* s = h->solver;
  const int nf = s->nf;
  *iters_out = 0;
  if (nf == 0) return MVBA_OK;
  const int max_it = std::max(1, o.max_linear_solver_iterations);
  MVBA_CUDA(s->pcg_scal.alloc((size_t)4 * (max_it + 2) + 2));
  MVBA_CUDA(s->pcg_scal.zero(h->stream));
  MVBA_CUDA(cudaMemsetAsync(s->pcg_flags.p, 0, 4 * sizeof(int), h->stream));
  PcgScal sc{s->pcg_scal.p, s->pcg_flags.p, s->pcg_flags.p + 1, s->pcg_scal.p + (size_t)4 * (max_it + 2)};
  const int tb = 256, gb = (nf + tb - 1) / tb;
  const int spmv_warps = s->nb + s->B, spmv_grid = (spmv_warps * 32 + 255) / 256;
  if (s->pcg_coop_grid > 0) {  // the whole solve in one cooperative launch, no host round trip
    const double* Mpp = s->Mpp.p; const double* Mbb = s->Mbb.p; const double* bb = s->R.rhs;
    double *xx = s->y.p, *rr = s->r.p, *zz = s->z.p, *pp = s->p.p, *qq = s->q.p;
    double eta = o.eta; int min_it = o.min_linear_solver_iterations, mx = max_it;
    int rpc = s->pcg_rows_per_cta;
    double* pp2 = s->p2.p;
    void* args[] = {&s->R, &s->A, &Mpp, &Mbb, &bb, &xx, &rr, &zz, &pp, &qq, &sc, &eta, &min_it, &mx, &rpc, &pp2};
    MVBA_CUDA(cudaLaunchCooperativeKernel((const void*)pcg_fused_kernel, dim3(s->pcg_coop_grid), dim3(s->pcg_threads), args, s->smem_pcg, h->stream));
    h->launches++;
    // the iteration count lands in pinned memory; the caller reads it after the LM iteration's own synchronisation (-1: deferred)
    MVBA_CUDA(cudaMemcpyAsync(s->flags_host, s->pcg_flags.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    *iters_out = -1;
    return MVBA_OK;
  }
  pcg_init_kernel<<<gb, tb, 0, h->stream>>>(nf, s->npc, s->R.rhs, s->y.p, s->r.p, s->q.p, sc);
  h->launches++;
  const int batch = 8;
  int it = 1;
  while (it <= max_it) {
    int last = std::min(max_it, it + batch - 1);
    for (; it <= last; it++) {
      pcg_precond_kernel<<<gb, tb, 0, h->stream>>>(s->R, s->Mpp.p, s->Mbb.p, s->r.p, s->z.p, sc, it);
      pcg_dir_kernel<<<gb, tb, 0, h->stream>>>(nf, s->npc, s->z.p, s->p.p, s->q.p, sc, it);
      spmv_kernel<<<spmv_grid, 256, 0, h->stream>>>(s->R, s->A, s->p.p, s->q.p, sc.done);
      pcg_dot_kernel<<<gb, tb, 0, h->stream>>>(nf, s->p.p, s->q.p, sc, it);
      pcg_update_kernel<<<gb, tb, 0, h->stream>>>(nf, s->R.rhs, s->p.p, s->q.p, s->y.p, s->r.p, sc, it);
      h->launches += 5;
      int refreshed = 0;
      if (it % 10 == 0) {  // residual_reset_period
        MVBA_CUDA(cudaMemsetAsync(s->q.p + s->npc, 0, (size_t)s->B * sizeof(double), h->stream));
        spmv_kernel<<<spmv_grid, 256, 0, h->stream>>>(s->R, s->A, s->y.p, s->q.p, sc.done);
        pcg_refresh_kernel<<<gb, tb, 0, h->stream>>>(nf, s->R.rhs, s->q.p, s->y.p, s->r.p, sc, it);
        h->launches += 2;
        refreshed = 1;
      }
      pcg_check_kernel<<<1, 1, 0, h->stream>>>(sc, it, refreshed, o.eta, o.min_linear_solver_iterations, max_it);
      h->launches++;
    }
    MVBA_CUDA(cudaMemcpyAsync(s->flags_host, s->pcg_flags.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    MVBA_CUDA(cudaStreamSynchronize(h->stream));
    if (s->flags_host[0]) break;
  }
  MVBA_CUDA(cudaGetLastError());
  *iters_out = s->flags_host[1];
  return MVBA_OK;
}

// Direct solve of the reduced camera system: dense blocked Cholesky on the FP64 tensor cores (kernels_chol.cuh).
inline int run_dense(mvba_problem* h, int* iters_out) {
  Solver* s = h->solver;
  const int nf = s->nf;
  *iters_out = 0;
  if (nf == 0) return MVBA_OK;
  if (nf > MVBA_CHOL_MAX_N) return fail(MVBA_EINVAL, "mvba_solve: linear_solver = 1 (dense Cholesky) takes reduced systems of at most 8192 columns");
  const int NB = MVBA_CHOL_NB, N = (nf + NB - 1) / NB * NB;
  MVBA_CUDA(s->cholA.alloc((size_t)(N + NB) * N));
  MVBA_CUDA(s->cholLinv.alloc((size_t)(N / NB) * NB * NB));
  static bool attr = false;
  if (!attr) {
    MVBA_CUDA(cudaFuncSetAttribute(chol_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MVBA_CHOL_SMEM_DIAG));
    MVBA_CUDA(cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MVBA_CHOL_SMEM_GEMM));
    MVBA_CUDA(cudaFuncSetAttribute(chol_syrk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MVBA_CHOL_SMEM_GEMM));
    MVBA_CUDA(cudaFuncSetAttribute(chol_backsolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (MVBA_CHOL_MAX_N + 17 * NB) * 8));
    attr = true;
  }
  MVBA_CUDA(s->cholA.zero(h->stream));
  chol_expand_kernel<<<std::max(1, std::min(4 * h->sm_count, (int)(((int64_t)s->nnzb * 49 + (int64_t)s->npc * s->B + 255) / 256 + 1))), 256, 0, h->stream>>>(
      s->R, s->A, s->R.rhs, s->cholA.p, N);
  h->launches++;
  for (int k = 0; k < N; k += NB) {
    chol_diag_kernel<<<1, 256, MVBA_CHOL_SMEM_DIAG, h->stream>>>(s->cholA.p, N, k, s->cholLinv.p + (size_t)(k / NB) * NB * NB);
    const int m = (N - k) / NB;  // tile rows below the panel, the right-hand-side tile included
    chol_panel_kernel<<<m, 256, MVBA_CHOL_SMEM_GEMM, h->stream>>>(s->cholA.p, N, k, s->cholLinv.p + (size_t)(k / NB) * NB * NB);
    h->launches += 2;
    if (m > 1) { chol_syrk_kernel<<<m * (m + 1) / 2, 256, MVBA_CHOL_SMEM_GEMM, h->stream>>>(s->cholA.p, N, k, m); h->launches++; }
  }
  chol_backsolve_kernel<<<1, 1024, (size_t)(N + 17 * NB) * 8, h->stream>>>(s->cholA.p, N, nf, s->cholLinv.p, s->y.p);
  h->launches++;
  MVBA_CUDA(cudaGetLastError());
  *iters_out = 1;
  return MVBA_OK;
}

inline int run_cost(mvba_problem* h, const double* x, double* out_slot) {
  const View& v = h->view;
  run_prep(h, x);
  Solver* s = h->solver;
  static const bool by_items = [] { const char* e = std::getenv("MVBA_COST_BY_ITEMS"); return e && e[0] == '1'; }();
  if (s && s->cam_fast && s->nchunks > 0 && !by_items) {  // camera-major records: image data broadcast per warp, priors by the item kernel
    CamQArgs qa{};
    qa.x = x; qa.tab = h->d_tab.p; qa.dtab = h->d_dtab.p;
    qa.chunk_cam = s->chunk_cam.p; qa.chunk_beg = s->chunk_beg.p; qa.citems = s->citems.p; qa.crec = s->crec.p; qa.nchunks = s->nchunks; qa.Q = nullptr;
    camcost_kernel<<<s->grid_camcost, 256, 0, h->stream>>>(v, qa, out_slot);
    h->launches++;
    if (v.n_prior > 0) {
      const int gp = (int)std::max<int64_t>(1, std::min<int64_t>((v.n_prior + 255) / 256, (int64_t)h->sm_count * 4));
      residual_kernel<<<gp, 256, 0, h->stream>>>(v, x, h->d_tab.p, h->d_dtab.p, 1, out_slot, nullptr, v.N + v.n_dtri + v.n_dmesh);
      h->launches++;
    }
    return MVBA_OK;
  }
  int64_t total = v.N + v.n_dtri + v.n_dmesh + v.n_prior;
  static const int occ = [] { int o = 0; return cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, residual_kernel, 256, 0) == cudaSuccess && o > 0 ? o : 4; }();
  int grid = (int)std::max<int64_t>(1, std::min<int64_t>((total + 255) / 256, (int64_t)h->sm_count * occ));  // one resident wave: a second, partial wave cost 20 %
  residual_kernel<<<grid, 256, 0, h->stream>>>(v, x, h->d_tab.p, h->d_dtab.p, 1, out_slot, nullptr);
  h->launches++;
  return MVBA_OK;
}

inline int solver_solve(mvba_problem* h, const mvba_solve_options& o, mvba_summary* S, mvba_iteration* its, int max_its) {
  auto t0 = std::chrono::steady_clock::now();
  std::memset(S, 0, sizeof(*S));
  const int64_t launches0 = h->launches;
  int rc = solver_prepare(h, o.preconditioner == 0); if (rc) { solver_destroy(h); return rc; }
  Solver* s = h->solver;
  const View& v = h->view;
  const bool timing = std::getenv("MVBA_TIMING") != nullptr;
  auto lap = [&](const char* what) { if (timing) { cudaStreamSynchronize(h->stream); std::fprintf(stderr, "[mvba timing] %s at %.3f ms\n", what, std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() * 1e3); } };
  lap("prepared");
  rc = upload_params(h); if (rc) return rc;
  lap("uploaded");
  MVBA_CUDA(cudaMemsetAsync(h->d_err.p, 0, sizeof(int), h->stream));
  MVBA_CUDA(cudaMemsetAsync(s->lm.p, 0, LM_NSLOTS * sizeof(double), h->stream));
  const double init[2] = {o.initial_trust_region_radius, 2.0};
  MVBA_CUDA(cudaMemcpyAsync(s->lm.p + LM_RADIUS, init, 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  Timers tm(h->stream);
  s->tm = &tm; s->build_launches = 0;
  enum { PH_ALL = 0, PH_BUILD = 1, PH_LINEAR = 2, PH_UPDATE = 3 };
  cudaEvent_t e_all0, e_all1;
  MVBA_CUDA(cudaEventCreate(&e_all0)); MVBA_CUDA(cudaEventCreate(&e_all1));
  MVBA_CUDA(cudaEventRecord(e_all0, h->stream));
  const int tb = 256;
  const int64_t ne = 3 * (int64_t)v.P;
  int njac = 0, nres = 0, nlin = 0, nsucc = 0, iter = 0, term = MVBA_TERM_NO_CONVERGENCE;
  double radius = o.initial_trust_region_radius, gmax = 0, cost = 0;
  const bool scaled = o.jacobi_scaling != 0;
  if (scaled) {
    tm.begin(PH_BUILD);
    rc = run_build(h, o, radius, 1, false); if (rc) return rc;
    if (s->nf) { jscale_kernel<<<(s->nf + tb - 1) / tb, tb, 0, h->stream>>>(s->nf, s->R.diagF, s->jscale_f.p); h->launches++; }
    if (ne) { jscale_kernel<<<(unsigned)((ne + tb - 1) / tb), tb, 0, h->stream>>>(ne, s->diagE.p, s->jscale_e.p); h->launches++; }
    tm.end();
  }
  LmOpts lo{o.min_relative_decrease, o.parameter_tolerance, o.function_tolerance, o.max_trust_region_radius, o.min_trust_region_radius, o.gradient_tolerance};
  bool have_cost = false, last_rejected = false, last_accepted = false;
  while (iter < o.max_num_iterations) {
    iter++;
    // ---- fused Jacobian + normal equations + Schur elimination at x ----
    tm.begin(PH_BUILD);
    rc = run_build(h, o, radius, 0, scaled, !have_cost, have_cost && last_rejected, last_accepted); if (rc) return rc;
    njac++;
    if (!have_cost) { MVBA_CUDA(cudaMemcpyAsync(s->lm.p + LM_COST, s->red.p + s->o_cost, sizeof(double), cudaMemcpyDeviceToDevice, h->stream)); }
    MVBA_CUDA(cudaMemsetAsync(s->lm.p + LM_GMAX, 0, sizeof(double), h->stream));
    if (h->nranks > 1)  // the landmark parts of all ranks came with the allreduce of the build (run_build)
      gradmax_kernel<<<(unsigned)((s->nf + h->nranks + tb - 1) / tb), tb, 0, h->stream>>>(s->nf, 0, s->R.g, nullptr, scaled ? s->jscale_f.p : nullptr, nullptr,
                                                                                         s->red.p + s->o_gmaxe, h->nranks, s->lm.p + LM_GMAX);
    else
      gradmax_kernel<<<(unsigned)std::max<int64_t>(1, std::min<int64_t>((s->nf + ne + tb - 1) / tb, (int64_t)8 * h->sm_count)), tb, 0, h->stream>>>(s->nf, ne, s->R.g, s->ge.p, scaled ? s->jscale_f.p : nullptr,
                                                                                    scaled ? s->jscale_e.p : nullptr, nullptr, 0, s->lm.p + LM_GMAX);
    h->launches++;
    if (s->nf) {
      damp_kernel<<<(s->nf + tb - 1) / tb, tb, 0, h->stream>>>(s->R, radius, o.min_lm_diagonal, o.max_lm_diagonal, s->Dsq.p);
      if (s->nb) precond_pose_kernel<<<(s->nb + 63) / 64, 64, 0, h->stream>>>(s->R, s->with_F ? 1 : 0, s->Mpp.p);
      if (s->B) precond_border_kernel<<<1, 256, 0, h->stream>>>(s->R, s->with_F ? 1 : 0, o.preconditioner == 2 ? nullptr : s->border_blk.p, s->Mbb.p, s->work.p);
      h->launches += 3;
    }
    tm.end();
    if (!have_cost) {
      double init_rec[LM_NSLOTS];
      MVBA_CUDA(cudaMemcpyAsync(init_rec, s->lm.p, LM_NSLOTS * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
      rc = check_device_error(h); if (rc) return rc;
      cost = init_rec[LM_COST]; gmax = init_rec[LM_GMAX];
      if (!std::isfinite(cost)) return fail(MVBA_ENUMERIC, "mvba: non finite cost at the initial point");
      S->initial_cost = cost;
      if (its && max_its > 0) its[0] = mvba_iteration{cost, 0, gmax, 0, 0, radius, 0, 1};
      if (o.verbose) std::printf("iter      cost      cost_change  |gradient|   |step|    tr_ratio  tr_radius  ls_iter\n   0  %.6e    0.00e+00    %.2e   0.00e+00   0.00e+00  %.2e        0\n", cost, gmax, radius);
      have_cost = true;
      if (gmax <= o.gradient_tolerance) { term = MVBA_TERM_CONVERGENCE; iter--; break; }
    }
    // ---- reduced camera system ----
    tm.begin(PH_LINEAR);
    int cgit = 0;
    rc = o.linear_solver == 1 ? run_dense(h, &cgit) : run_pcg(h, o, &cgit); if (rc) return rc;
    tm.end();
    // ---- back-substitution, candidate point and its cost ----
    tm.begin(PH_UPDATE);
    MVBA_CUDA(cudaMemsetAsync(s->lm.p + LM_CAND, 0, 4 * sizeof(double), h->stream));
    MVBA_CUDA(cudaMemsetAsync(s->lm.p + LM_STEP2, 0, 2 * sizeof(double), h->stream));
    if (v.P > 0) {
      const int ndir = v.C + v.T;
      dir_cameras_kernel<<<(ndir + 127) / 128, 128, 0, h->stream>>>(v, h->d_tab.p, h->d_dtab.p, s->y.p, scaled ? s->jscale_f.p : nullptr,
                                                                      s->camdir.p, s->sensdir.p);
      BacksubArgs b{};
      b.x = h->d_x.p; b.tab = h->d_tab.p; b.dtab = h->d_dtab.p; b.cd = s->camdir.p; b.sdir = s->sensdir.p;
      b.jscale_f = scaled ? s->jscale_f.p : nullptr; b.jscale_e = scaled ? s->jscale_e.p : nullptr;
      b.Hinv = s->Hinv.p; b.ge = s->ge.p; b.yf = s->y.p; b.ye = s->ye.p; b.lm = s->lm.p; b.lm_mcc = LM_MCC;
      bool dz = false;
      for (int t = 0; t < v.T; t++) dz = dz || v.sens[t].c_dist >= 0;
      if (dz) backsub_kernel<true><<<s->grid_backsub, 128, 0, h->stream>>>(v, b);
      else backsub_kernel<false, true><<<s->grid_backsub, 128, 0, h->stream>>>(v, b);  // R kept in registers: the kernel is bound by L1 requests (measured: 0.02 ms per iteration)
      h->launches += 2;
    }
    update_kernel<<<(unsigned)std::max<int64_t>(1, std::min<int64_t>((v.nx + tb - 1) / tb, (int64_t)8 * h->sm_count)), tb, 0, h->stream>>>(v.nx, s->nf, s->xcol.p, h->d_x.p, s->y.p, s->ye.p, scaled ? s->jscale_f.p : nullptr,
                                                              scaled ? s->jscale_e.p : nullptr, h->bounded ? s->lo.p : nullptr,
                                                              h->bounded ? s->hi.p : nullptr, h->d_xnew.p, s->lm.p);
    h->launches++;
    run_cost(h, h->d_xnew.p, s->lm.p + LM_CAND);
    nres++;
    rc = allreduce(h, s->lm.p + LM_CAND, 4, 8, 0); if (rc) return rc;
    lm_decide_kernel<<<1, 1, 0, h->stream>>>(s->lm.p, lo, s->rec, 0);
    commit_kernel<<<(v.nx + tb - 1) / tb, tb, 0, h->stream>>>(v.nx, s->lm.p, h->d_xnew.p, h->d_x.p);
    h->launches += 2;
    tm.end();
    MVBA_CUDA(cudaStreamSynchronize(h->stream));
    MVBA_CUDA(cudaGetLastError());
    if (cgit < 0) cgit = s->flags_host[1];
    nlin += cgit;
    mvba_iteration rec = *s->rec;
    int dev_term = rec.linear_iterations;
    rec.linear_iterations = cgit;
    radius = rec.radius; cost = rec.cost; gmax = rec.gradient_max_norm;
    if (dev_term == 3) {  // gradient tolerance met at the point the previous step reached: this pass over the loop was no iteration
      iter--; nlin -= cgit; term = MVBA_TERM_CONVERGENCE; break;
    }
    if (rec.step_accepted) nsucc++;
    last_rejected = !rec.step_accepted; last_accepted = rec.step_accepted != 0;
    if (its && iter < max_its) its[iter] = rec;
    if (o.verbose) std::printf("%4d  %.6e   %9.2e    %.2e   %.2e  %9.2e  %.2e   %6d %s\n", iter, rec.cost, rec.cost_change, rec.gradient_max_norm,
                               rec.step_norm, rec.relative_decrease, rec.radius, cgit, rec.step_accepted ? "" : "(rejected)");
    if (dev_term) { term = dev_term == 5 ? MVBA_TERM_FAILURE : MVBA_TERM_CONVERGENCE; break; }
  }
  // gradient tolerance at the final point is checked by the next build in Ceres; the reference's
  // 1e-16 tolerance is never reached in practice, so the last recorded norm is reported.
  MVBA_CUDA(cudaEventRecord(e_all1, h->stream));
  // an interpolation that left its bracket at ANY evaluated point (the reference aborts there, rig_calibrator.cc:346) is
  // reported before the write-back: the caller's arrays keep the values they came in with
  rc = check_device_error(h); if (rc) return rc;
  lap("loop done");
  rc = download_params(h); if (rc) return rc;
  lap("downloaded");
  float ms_all = 0;
  MVBA_CUDA(cudaEventElapsedTime(&ms_all, e_all0, e_all1));
  cudaEventDestroy(e_all0); cudaEventDestroy(e_all1);
  S->final_cost = cost; S->num_iterations = iter; S->num_successful_steps = nsucc; S->num_linear_iterations = nlin;
  S->termination = term; S->num_jacobian_evals = njac + (scaled ? 1 : 0); S->num_residual_evals = nres;
  S->num_kernel_launches = h->launches - launches0;
  S->seconds_device = ms_all * 1e-3; S->seconds_build = tm.total(PH_BUILD); S->seconds_linear = tm.total(PH_LINEAR); S->seconds_update = tm.total(PH_UPDATE);
  S->final_gradient_max_norm = gmax; S->final_radius = radius;
  S->seconds_build_kernel = tm.total(9); S->num_build_launches = s->build_launches; s->tm = nullptr;
  S->seconds_total = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return MVBA_OK;
}

// Timed fused build on device-resident parameters (bench.py roofline leg).
inline int solver_bench_build(mvba_problem* h, int repeats, float* ms_total) {
  mvba_solve_options o; mvba_solve_options_init(&o);
  int rc = solver_prepare(h, false); if (rc) return rc;
  Solver* s = h->solver;
  s->tm = nullptr;
  rc = run_build(h, o, o.initial_trust_region_radius, 1, false); if (rc) return rc;
  const int tb = 256; const int64_t ne = 3 * (int64_t)h->view.P;
  if (s->nf) jscale_kernel<<<(s->nf + tb - 1) / tb, tb, 0, h->stream>>>(s->nf, s->R.diagF, s->jscale_f.p);
  if (ne) jscale_kernel<<<(unsigned)((ne + tb - 1) / tb), tb, 0, h->stream>>>(ne, s->diagE.p, s->jscale_e.p);
  rc = run_build(h, o, o.initial_trust_region_radius, 0, true); if (rc) return rc;  // warm-up
  cudaEvent_t e0, e1;
  MVBA_CUDA(cudaEventCreate(&e0)); MVBA_CUDA(cudaEventCreate(&e1));
  MVBA_CUDA(cudaEventRecord(e0, h->stream));
  for (int i = 0; i < repeats; i++) { rc = run_build(h, o, o.initial_trust_region_radius, 0, true); if (rc) return rc; }
  MVBA_CUDA(cudaEventRecord(e1, h->stream));
  MVBA_CUDA(cudaEventSynchronize(e1));
  MVBA_CUDA(cudaEventElapsedTime(ms_total, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return MVBA_OK;
}

}  // namespace mvba

// kernels_tile2.cuh -- pass B of the normal-equation build, window tiles, for passes without free intrinsics / depth
// columns (every f column is a pose, rig-transform or time-offset column).  Same tiling and static round tables as
// kernels_tile.cuh (tracks sorted by first pose block, windows of consecutive pose blocks, rounds of <= K tracks), but
//   * two CTAs of 256 threads per SM (<= 113 KB of shared memory each): while one CTA runs its tensor-core phase the
//     other evaluates residuals, so the FP64 pipe, the DMMA pipe and the memory system overlap without explicit
//     warp roles;
//   * everything a round reads is STAGED ASYNCHRONOUSLY: the round's packed table (TRoundHdr | tracks | next round's
//     landmarks | one 16-byte record {landmark slot, image, pixel} per residual block | gather groups) arrives by one
//     1-D bulk copy (cp.async.bulk -> mbarrier) issued a phase earlier, the window's per-image data (CamPack: R, t,
//     t - t_rc, the 23 coefficient rows, sensor) by one bulk copy per tile, and the next round's landmark coordinates
//     by cp.async while the current round is gathered.  No dependent global load is left in the evaluation;
//   * the landmark block is factored, E'E + D = R'R, and U = R^-1 (so Hinv = U U') is folded into the gathered rows,
//     Gs = G U: ONE staged matrix feeds both DMMA operands (C += Gs Gs'), the Y = G Hinv phase and its buffer are gone,
//     and the reduced right-hand side comes out of the same GEMM through one extra row  v = U' E'r  (its column of C
//     is  Gs v = G Hinv E'r);
//   * one thread gathers a whole (track, parameter block) group (7 rows x 3): the block's M = (sr A)'E and
//     N = [Z]x M are loaded once per contributor instead of once per row;
//   * tiles are handed out by an atomic counter (no static tile -> CTA map, so small multi-GPU shards balance).
// Per round: P1 evaluate (thread per residual block) | P2 landmark inverse + Cholesky (warp 0) next to the gather
// (warps 1-7) | scale by L, write Gs | P3 DMMA.  Four CTA barriers per round.
#pragma once
#include "kernels_tile.cuh"

namespace mvba {

#define MVBA_T2_THREADS 256
#define MVBA_T2_GCAP MVBA_T2_THREADS  // gather groups of one round (one thread each)

struct alignas(16) TRoundHdr { int32_t kc, nit, ngrp, kc_next; uint32_t next_bytes; int32_t pad[3]; };
static_assert(sizeof(TRoundHdr) == 32, "round blob layout");
__host__ __device__ inline uint32_t t2_align16(uint32_t n) { return (n + 15u) & ~15u; }
// blob: TRoundHdr | TTrk[kc] | int32 next_p[kc_next] | TSlot[nit] | TGrp[ngrp], every part 16-byte aligned
__host__ __device__ inline uint32_t t2_blob_bytes(int kc, int kc_next, int nit, int ngrp) {
  return 32u + t2_align16(8u * kc) + t2_align16(4u * kc_next) + 16u * nit + 16u * ngrp;
}

// 13 doubles: an ODD stride in 8-byte units, so that the same field of consecutive slots falls in different banks.  What the
// gather needs of a block: M = (sr A)'E and Z (the rig rotation acts on Z + dl, dl per image); the block's share of E'E and E'r
// never leaves the registers (it is summed over the track's lanes with shuffles).
struct TRec2 { double M[9], Z[3], pad; };
static_assert(sizeof(TRec2) == 104, "TRec2 layout");
// Where slot q's record lives: one record is skipped after every eight.  Neighbouring gather threads read the records of
// neighbouring TRACKS, usually eight slots apart: 8 x 13 doubles is 8 (mod 16) -- two bank positions, 16-way conflicts --
// while 9 x 13 is 5 (mod 16).
__host__ __device__ inline int t2_rslot(int q) { return q + (q >> 3); }
__host__ __device__ inline int t2_rcap(int icap) { return icap + (icap >> 3) + 1; }

struct Tile2Args {
  BuildArgs b;
  const CamPack* pack;
  const int32_t* tile_w0;      // first pose block of the window
  const int32_t* tile_wn;      // number of pose blocks in the window
  const int32_t* tile_c0;      // first image of the tile's image range
  const int32_t* tile_nc;      // images staged (<= ccap)
  const int32_t* tile_rounds;  // rounds of the tile
  const int64_t* tile_blob0;   // byte offset of the tile's first round blob
  const uint32_t* tile_bytes0; // size of that blob
  const uint8_t* blob;
  const uint16_t* con;         // overflow contributors (groups with more than four)
  int* tile_ctr;               // work counter, zero at launch
  int ntiles;
  int K, icap, ccap, L8max, blob_cap;
};

// Shared memory: Gs[L8max][KK] | pack[ccap] | blob | Xs[2][K][6] | Ls[K][6] | cs[L8max] | rec[icap + icap / 8 + 1] | rcid[same] | mbarriers
// (the bulk-copy destinations first: they need 16-byte alignment)
// Row length of Gs: 3 K rounded up to 8 (mod 16) doubles.  The tensor-core phase reads every fragment as ONE 128-bit load that
// serves two DMMAs (columns k0 + 2 tig and k0 + 2 tig + 1: the product Gs Gs' does not care in which order the columns are
// contracted, as long as both operands use the same one); a quarter warp (2 rows x 4 x 16 bytes) then covers all banks once.
__host__ __device__ inline int tile2_kk(int K) { int kk = 3 * K; while ((kk & 15) != 8) kk++; return kk; }

__host__ __device__ inline size_t tile2_smem(int K, int icap, int ccap, int L8max, int blob_cap) {
  const size_t KK = tile2_kk(K);
  size_t n = (size_t)L8max * KK * 8 + (size_t)ccap * sizeof(CamPack) + (size_t)blob_cap;
  n += ((size_t)2 * K * 6 + (size_t)K * 6 + (size_t)L8max) * 8;
  n += (size_t)t2_rcap(icap) * sizeof(TRec2);
  n += (((size_t)t2_rcap(icap) * 4 + 7) & ~(size_t)7) + 16;
  return n + 16;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}" : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Landmark coordinates and Jacobi scale of one track -> Xs row (6 doubles), asynchronously.
__device__ __forceinline__ void t2_fetch_landmark(const View& v, const BuildArgs& a, int p, double* dst) {
  const double* X = a.x + v.x_xyz + 3 * (int64_t)p;
  cp_async8(dst, X); cp_async8(dst + 1, X + 1); cp_async8(dst + 2, X + 2);
  if (a.jscale_e) {
    const double* s = a.jscale_e + 3 * (size_t)p;
    cp_async8(dst + 3, s); cp_async8(dst + 4, s + 1); cp_async8(dst + 5, s + 2);
  } else { dst[3] = 1.0; dst[4] = 1.0; dst[5] = 1.0; }
}


// The image data of one residual block: from the tile's staged copy (STAGED, plain shared-memory loads) or, in tiles whose
// image range does not fit the staging area, straight from the global table.
template <bool STAGED>
__device__ __forceinline__ const CamPack* t2_pack(const CamPack* pack_s, const CamPack* __restrict__ pack_g, int cid, int c0) {
  return STAGED ? pack_s + (cid - c0) : pack_g + cid;
}

// One gather group: rows of G = F'E of one (track, parameter block) summed over its contributors.
template <bool STAGED>
__device__ __forceinline__ void t2_gather(const int4 raw, const uint16_t* __restrict__ con, const TRec2* rec, const int* rcid,
                                          const CamPack* pack_s, const CamPack* __restrict__ pack_g, int c0, double (*o)[3]) {
  const unsigned over0 = (unsigned)raw.x;
  const int ncon = (int)(((unsigned)raw.y >> 24) & 0x7f);
  for (int c = 0; c < ncon; c++) {
    const unsigned cb = c < 4 ? grp_con(raw, c) : __ldg(con + over0 + (c - 4));
    const int slot = cb >> 2, role = cb & 3;
    const TRec2& r = rec[t2_rslot(slot)];
    const CamPack* cp = t2_pack<STAGED>(pack_s, pack_g, rcid[t2_rslot(slot)], c0);
    double M[9];
#pragma unroll
    for (int k = 0; k < 9; k++) M[k] = r.M[k];
    double W0 = r.Z[0], W1 = r.Z[1], W2 = r.Z[2];
    if (role == 2) { W0 += cp->dl[0]; W1 += cp->dl[1]; W2 += cp->dl[2]; }  // the rig rotation acts on Xc - t_rc = Z + (t - t_rc)
    double N[9];  // N[:,b] = W x M[:,b]
#pragma unroll
    for (int b = 0; b < 3; b++) {
      N[b] = W1 * M[6 + b] - W2 * M[3 + b];
      N[3 + b] = W2 * M[b] - W0 * M[6 + b];
      N[6 + b] = W0 * M[3 + b] - W1 * M[b];
    }
    if (role == 3) {  // time offset: d = dalpha (wa x Z + dt)
      const double* cw = cp->c[21];
      const double* ce = cp->c[22];
#pragma unroll
      for (int b = 0; b < 3; b++)
        o[0][b] += cw[0] * N[b] + cw[1] * N[3 + b] + cw[2] * N[6 + b] + ce[0] * M[b] + ce[1] * M[3 + b] + ce[2] * M[6 + b];
    } else {
      const double(*cr)[3] = cp->c + 7 * role;
#pragma unroll
      for (int row = 0; row < 3; row++) {
        const double f0 = cr[row][0], f1 = cr[row][1], f2 = cr[row][2];
#pragma unroll
        for (int b = 0; b < 3; b++) o[row][b] = fma(f0, M[b], fma(f1, M[3 + b], fma(f2, M[6 + b], o[row][b])));
      }
#pragma unroll
      for (int row = 3; row < 7; row++) {
        const double f0 = cr[row][0], f1 = cr[row][1], f2 = cr[row][2];
#pragma unroll
        for (int b = 0; b < 3; b++) o[row][b] = fma(f0, N[b], fma(f1, N[3 + b], fma(f2, N[6 + b], o[row][b])));
      }
    }
  }
}

#ifdef MVBA_TILE_PROFILE
#define T2PROF(slot) do { if (tid == 0) { long long now_ = clock64(); prof_[slot] += now_ - tp_; tp_ = now_; } } while (0)
#else
#define T2PROF(slot) do { } while (0)
#endif

// MAXS: accumulator blocks (8 x 8) per warp; the eight warps cover the upper block triangle of the window.
template <int MAXS>
__global__ void __launch_bounds__(MVBA_T2_THREADS, 2) tile2_kernel(View v, Tile2Args a, Reduced S) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int NT = MVBA_T2_THREADS, NW = NT / 32;
  const int tid = threadIdx.x, lane = tid & 31, wib = tid >> 5;
  const int K = a.K, KK = tile2_kk(K);
  double* Gs = reinterpret_cast<double*>(smem_raw);
  CamPack* pack_s = reinterpret_cast<CamPack*>(Gs + (size_t)a.L8max * KK);
  unsigned char* blob_s = reinterpret_cast<unsigned char*>(pack_s + a.ccap);
  double* Xs = reinterpret_cast<double*>(blob_s + a.blob_cap);
  double* Ls = Xs + (size_t)2 * K * 6;
  double* cs = Ls + (size_t)K * 6;
  TRec2* rec = reinterpret_cast<TRec2*>(cs + a.L8max);
  int* rcid = reinterpret_cast<int*>(rec + t2_rcap(a.icap));
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(rcid) + (((size_t)t2_rcap(a.icap) * 4 + 7) & ~(size_t)7));
  uint64_t* bar_blob = bars;
  uint64_t* bar_pack = bars + 1;
  __shared__ int s_tile;
  const int B = S.B;
  const int gid = lane >> 2, tig = lane & 3;
  double cost_acc = 0;
  uint32_t ph_blob = 0, ph_pack = 0;
#ifdef MVBA_TILE_PROFILE
  long long prof_[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tp_ = clock64();
#endif
  if (tid == 0) {
    mbar_init(bar_blob, 1); mbar_init(bar_pack, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  for (;;) {
    if (tid == 0) s_tile = atomicAdd(a.tile_ctr, 1);
    __syncthreads();
    const int t = s_tile;
    if (t >= a.ntiles) break;
    const int w0 = a.tile_w0[t], Wt = a.tile_wn[t];
    const int c0 = a.tile_c0[t], nc = a.tile_nc[t];
    const bool staged = nc > 0;  // the tile's image range fits the staging area (otherwise every access goes to the global table)
    const int nrounds = a.tile_rounds[t];
    const int Lt = 7 * Wt + B;               // local f columns; row Lt carries v = L' E'r
    const int L8 = (Lt + 1 + 7) & ~7;
    const int nb8 = L8 >> 3, nblk = nb8 * (nb8 + 1) / 2;
    const int per = (nblk + NW - 1) / NW;    // <= MAXS by the host's window cap
    const int s0 = min(nblk, wib * per), ns = min(nblk, s0 + per) - s0;
    int bi0 = 0, bj0 = 0;                    // first block of this warp (row-major over the upper block triangle)
    { int rem = s0; while (bi0 < nb8 && rem >= nb8 - bi0) { rem -= nb8 - bi0; bi0++; } bj0 = bi0 + rem; }
    double acc[MAXS][2];
#pragma unroll
    for (int i = 0; i < MAXS; i++) { acc[i][0] = 0; acc[i][1] = 0; }
    int offa[MAXS], offb[MAXS];  // row offsets (doubles) of the two fragments of every block this warp accumulates
    unsigned newrow = 0;         // bit sl: block sl starts a new block row (its A fragment has to be loaded)
    {
      int bi = bi0, bj = bj0, prev = -1;
#pragma unroll
      for (int sl = 0; sl < MAXS; sl++) {
        if (sl < ns) {
          offa[sl] = 8 * bi * KK; offb[sl] = 8 * bj * KK;
          if (bi != prev) newrow |= 1u << sl;
          prev = bi;
          if (++bj == nb8) { bi++; bj = bi; }
        } else { offa[sl] = sl ? offa[0] : 0; offb[sl] = sl ? offb[0] : 0; }
      }
    }
    // ---- tile prologue: the window's images and the first round's table by bulk copy; Jacobi scale of the local columns
    int64_t boff = a.tile_blob0[t];
    uint32_t bbytes = a.tile_bytes0[t];
    if (tid == 0) {
      if (nc > 0) { mbar_expect_tx(bar_pack, (uint32_t)nc * (uint32_t)sizeof(CamPack)); bulk_g2s(pack_s, a.pack + c0, (uint32_t)nc * (uint32_t)sizeof(CamPack), bar_pack); }
      mbar_expect_tx(bar_blob, bbytes); bulk_g2s(blob_s, a.blob + boff, bbytes, bar_blob);
    }
    for (int l = tid; l < a.L8max; l += NT)
      cs[l] = (a.b.jscale_f && l < Lt) ? a.b.jscale_f[l < 7 * Wt ? 7 * w0 + l : S.npc + (l - 7 * Wt)] : 1.0;
    mbar_wait(bar_blob, ph_blob);
    {  // landmarks of the first round (later rounds are fetched one round ahead)
      const TRoundHdr* hd = reinterpret_cast<const TRoundHdr*>(blob_s);
      const TTrk* trk = reinterpret_cast<const TTrk*>(blob_s + 32);
      if (tid < hd->kc) t2_fetch_landmark(v, a.b, trk[tid].p, Xs + 6 * tid);
      cp_async_wait_all();
    }
    if (nc > 0) { mbar_wait(bar_pack, ph_pack); ph_pack ^= 1; }
    __syncthreads();
    T2PROF(0);

    for (int rd = 0; rd < nrounds; rd++) {
      if (rd > 0) mbar_wait(bar_blob, ph_blob);
      ph_blob ^= 1;
      const TRoundHdr hd = *reinterpret_cast<const TRoundHdr*>(blob_s);
      const int kc = hd.kc, nit = hd.nit, ngrp = hd.ngrp, kc_next = hd.kc_next;
      const TTrk* trk = reinterpret_cast<const TTrk*>(blob_s + 32);
      const int32_t* next_p = reinterpret_cast<const int32_t*>(blob_s + 32 + t2_align16(8u * kc));
      const TSlot* slots = reinterpret_cast<const TSlot*>(reinterpret_cast<const unsigned char*>(next_p) + t2_align16(4u * kc_next));
      const int4* gtab = reinterpret_cast<const int4*>(slots + nit);
      const double* Xcur = Xs + (size_t)(rd & 1) * K * 6;
      double* Xnext = Xs + (size_t)((rd + 1) & 1) * K * 6;
      // ---- P1: clear Gs, then one thread per residual block ----
      {
        double2* gz = reinterpret_cast<double2*>(Gs);
        const int nz = L8 * KK / 2;
        const int v0 = Lt * KK / 2, v1 = v0 + (3 * kc + 1) / 2;  // row Lt, columns < 3 kc: written by the track heads below, not cleared
        for (int i = tid; i < nz; i += NT) if (i < v0 || i >= v1) gz[i] = make_double2(0.0, 0.0);
      }
      double H[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};  // this block's share of E'E and E'r (zero for padding slots)
      int my_k = -1;                                          // track slot of this thread's block
      if (tid < nit && ((slots[tid].b >> 24) & 7) != 7) {
        const TSlot sl = slots[tid];
        const int kind = (sl.b >> 24) & 7, ks = sl.b >> 27, cid = sl.b & 0xffffffu;
        my_k = ks;
        const double* Xk = Xcur + 6 * ks;
        Item it;
        if (kind == 0) {
          it.nr = 2; it.loss_a = v.robust_th;
          const double X0 = Xk[0], X1 = Xk[1], X2 = Xk[2];
          double Rm[9], tv[3];
          int sensor;
          if (staged) {
            const CamPack* cp = t2_pack<true>(pack_s, a.pack, cid, c0);
#pragma unroll
            for (int i = 0; i < 9; i++) Rm[i] = cp->R[i];
#pragma unroll
            for (int i = 0; i < 3; i++) tv[i] = cp->t[i];
            sensor = cp->sensor;
          } else {
            const CamPack* cp = t2_pack<false>(pack_s, a.pack, cid, c0);
#pragma unroll
            for (int i = 0; i < 9; i++) Rm[i] = __ldg(cp->R + i);
#pragma unroll
            for (int i = 0; i < 3; i++) tv[i] = __ldg(cp->t + i);
            sensor = __ldg(&cp->sensor);
          }
#pragma unroll
          for (int i = 0; i < 3; i++) {
            it.Z[i] = Rm[3 * i] * X0 + Rm[3 * i + 1] * X1 + Rm[3 * i + 2] * X2;
            it.Xc[i] = it.Z[i] + tv[i];
          }
          project_distort<true>(v.sens[sensor], a.b.x, it.Xc, make_float2(sl.px, sl.py), it.r, it.A, it.Jf, &it.Jc, nullptr, nullptr, v.fov_tab + 2 * sensor);
          it.r[2] = 0;
#pragma unroll
          for (int i = 0; i < 2; i++)
#pragma unroll
            for (int b = 0; b < 3; b++) it.E[3 * i + b] = it.A[3 * i] * Rm[b] + it.A[3 * i + 1] * Rm[3 + b] + it.A[3 * i + 2] * Rm[6 + b];
        } else if (kind <= 2) {
          eval_depth(v, a.b.x, a.b.tab, a.b.dtab, kind == 1 ? MVBA_BLOCK_DEPTH_TRI : MVBA_BLOCK_DEPTH_MESH, sl.a, it);
        } else {
          eval_prior(v, a.b.x, sl.a, it);
        }
        double s = it.r[0] * it.r[0] + it.r[1] * it.r[1];
        if (it.nr == 3) s += it.r[2] * it.r[2];
        double hr = 0, sr;
        if (a.b.want_cost) cauchy(it.loss_a, s, &hr, &sr);
        else sr = cauchy_sr(it.loss_a, s);
        cost_acc += hr;
        const double se[3] = {Xk[3], Xk[4], Xk[5]};
        double Es[9];
#pragma unroll
        for (int k = 0; k < 9; k++) Es[k] = (k / 3 < it.nr) ? sr * it.E[k] * se[k % 3] : 0.0;
#pragma unroll
        for (int r = 0; r < 3; r++) {
          const double* e = Es + 3 * r;
          const double rs = r < it.nr ? sr * it.r[r] : 0.0;
          H[0] += e[0] * e[0]; H[1] += e[0] * e[1]; H[2] += e[0] * e[2]; H[3] += e[1] * e[1]; H[4] += e[1] * e[2]; H[5] += e[2] * e[2];
          g[0] += e[0] * rs; g[1] += e[1] * rs; g[2] += e[2] * rs;
        }
        TRec2& rc = rec[t2_rslot(tid)];
        rcid[t2_rslot(tid)] = kind >= 3 ? -1 : cid;
        if (kind < 3) {
#pragma unroll
          for (int k = 0; k < 3; k++) rc.Z[k] = it.Z[k];
#pragma unroll
          for (int b = 0; b < 3; b++)
#pragma unroll
            for (int aa = 0; aa < 3; aa++) {
              double tt = it.A[aa] * Es[b] + it.A[3 + aa] * Es[3 + b];
              if (it.nr == 3) tt += it.A[6 + aa] * Es[6 + b];
              rc.M[3 * aa + b] = sr * tt;
            }
        }
      }
      // ---- landmark block of every track: its slots are consecutive lanes of ONE warp (the host pads the slot list so that no
      //      track straddles a warp), so E'E and E'r are summed with shuffles and the track's first lane factors the block ----
      {
        int seg_end = lane + 1;
        bool head = false;
        TTrk tk = TTrk{0, 0, 0};
        if (my_k >= 0) { tk = trk[my_k]; seg_end = tk.q0 + tk.n - (tid - lane); head = tk.q0 == tid; }
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
          const bool in = lane + d < seg_end;
          if (!__any_sync(0xffffffffu, in)) break;  // no track of this warp is longer than d slots: the wider steps add nothing
#pragma unroll
          for (int k = 0; k < 6; k++) { const double t = __shfl_down_sync(0xffffffffu, H[k], d); if (in) H[k] += t; }
#pragma unroll
          for (int k = 0; k < 3; k++) { const double t = __shfl_down_sync(0xffffffffu, g[k], d); if (in) g[k] += t; }
        }
        if (head) {
          const double ir = 1.0 / a.b.radius;
          H[0] += fmin(fmax(H[0], a.b.min_diag), a.b.max_diag) * ir;
          H[3] += fmin(fmax(H[3], a.b.min_diag), a.b.max_diag) * ir;
          H[5] += fmin(fmax(H[5], a.b.min_diag), a.b.max_diag) * ir;
          // H = R'R (R upper triangular) with reciprocal square roots, U = R^-1:  Hinv = U U'  and  G Hinv G' = (G U)(G U)'
          const double i00 = rsqrt(H[0]);
          const double r01 = H[1] * i00, r02 = H[2] * i00;
          const double i11 = rsqrt(H[3] - r01 * r01);
          const double r12 = (H[4] - r01 * r02) * i11;
          const double i22 = rsqrt(H[5] - r02 * r02 - r12 * r12);
          const double u01 = -r01 * i00 * i11, u12 = -r12 * i11 * i22;
          const double u02 = -(r02 * i00 + r12 * u01) * i22;
          double Hi[6];  // xx xy xz yy yz zz
          Hi[0] = i00 * i00 + u01 * u01 + u02 * u02; Hi[1] = u01 * i11 + u02 * u12; Hi[2] = u02 * i22;
          Hi[3] = i11 * i11 + u12 * u12; Hi[4] = u12 * i22; Hi[5] = i22 * i22;
#pragma unroll
          for (int k = 0; k < 6; k++) a.b.Hinv[6 * (size_t)tk.p + k] = Hi[k];
#pragma unroll
          for (int k = 0; k < 3; k++) a.b.ge[3 * (size_t)tk.p + k] = g[k];
          double* L = Ls + 6 * my_k;
          L[0] = i00; L[1] = u01; L[2] = i11; L[3] = u02; L[4] = u12; L[5] = i22;
          double* vr = Gs + (size_t)Lt * KK + 3 * my_k;  // v = U' E'r
          vr[0] = i00 * g[0];
          vr[1] = u01 * g[0] + i11 * g[1];
          vr[2] = u02 * g[0] + u12 * g[1] + i22 * g[2];
          if (my_k == kc - 1 && ((3 * kc) & 1)) vr[3] = 0.0;  // the uncleared pair of row Lt ends one column past the last track
        }
      }
      T2PROF(1);
      __syncthreads();
      T2PROF(2);
      // ---- P2: one thread per (track, parameter block) group gathers its rows of G = F'E and stores Gs = diag(cs) G U ----
      if (tid < kc_next) t2_fetch_landmark(v, a.b, next_p[tid], Xnext + 6 * tid);  // next round's landmarks, waited for behind P3
      if (tid < ngrp) {
        const int4 raw = gtab[tid];
        const bool single = (((unsigned)raw.y >> 24) & 0x80) != 0;
        const int g_lcol = raw.y & 0xffff, g_k = (raw.y >> 16) & 0xff, g_rows = single ? 1 : 7;
        double o[7][3];
#pragma unroll
        for (int r = 0; r < 7; r++) { o[r][0] = 0; o[r][1] = 0; o[r][2] = 0; }
        if (staged) t2_gather<true>(raw, a.con, rec, rcid, pack_s, a.pack, c0, o);
        else t2_gather<false>(raw, a.con, rec, rcid, pack_s, a.pack, c0, o);
        const double* L = Ls + 6 * g_k;
        const double u00 = L[0], u01 = L[1], u11 = L[2], u02 = L[3], u12 = L[4], u22 = L[5];
#pragma unroll
        for (int r = 0; r < 7; r++) {
          if (r < g_rows) {
            const int l = g_lcol + r;
            const double sc = cs[l];
            double* gp = Gs + (size_t)l * KK + 3 * g_k;
            gp[0] = sc * (o[r][0] * u00);
            gp[1] = sc * (o[r][0] * u01 + o[r][1] * u11);
            gp[2] = sc * (o[r][0] * u02 + o[r][1] * u12 + o[r][2] * u22);
          }
        }
      }
      T2PROF(3);
      __syncthreads();
      T2PROF(5);
      // the round's table is consumed: fetch the next one while the tensor cores run
      if (rd + 1 < nrounds) {
        boff += bbytes; bbytes = hd.next_bytes;
        if (tid == 0) { mbar_expect_tx(bar_blob, bbytes); bulk_g2s(blob_s, a.blob + boff, bbytes, bar_blob); }
      }
      // ---- P3: C += Gs Gs' on the FP64 tensor cores, K dimension = 3 kc.  The row offsets of every block's two fragments
      //      are fixed for the tile: the inner loop is two shared loads and one DMMA per block, nothing else (slots past
      //      the warp's last block repeat block 0 into accumulators that are never flushed) ----
      {
        const int kend = (3 * kc + 7) & ~7;  // columns past 3 kc are zero (cleared with the rest of Gs)
        const double* base = Gs + (size_t)gid * KK + 2 * tig;
#pragma unroll 2
        for (int k0 = 0; k0 < kend; k0 += 8) {
          const double* kb = base + k0;
          double2 av = *reinterpret_cast<const double2*>(kb + offa[0]);
          double by[MAXS], ay[MAXS];
#pragma unroll
          for (int sl = 0; sl < MAXS; sl++) {
            // (slots past the warp's last block repeat block 0 into accumulators that are never flushed: a warp-uniform guard
            //  around the DMMA was measured 10 % slower than the idle DMMAs it saves)
            if (sl > 0 && ((newrow >> sl) & 1)) av = *reinterpret_cast<const double2*>(kb + offa[sl]);  // the A fragment stays while the warp's blocks stay in one block row
            const double2 bv = *reinterpret_cast<const double2*>(kb + offb[sl]);
            dmma884(acc[sl][0], acc[sl][1], av.x, bv.x);
            ay[sl] = av.y; by[sl] = bv.y;
          }
#pragma unroll
          for (int sl = 0; sl < MAXS; sl++) dmma884(acc[sl][0], acc[sl][1], ay[sl], by[sl]);
        }
      }
      T2PROF(6);
      cp_async_wait_all();  // the next round's landmarks (fetched during P2) are in place
      __syncthreads();
      T2PROF(7);
    }
    // ---- flush: local column -> reduced column; column Lt of C is the tile's share of G Hinv E'r ----
    {
      int bi = bi0, bj = bj0;
#pragma unroll
      for (int sl = 0; sl < MAXS; sl++) {
        if (sl < ns) {
#pragma unroll
          for (int e = 0; e < 2; e++) {
            const int li = 8 * bi + gid, lj = 8 * bj + 2 * tig + e;
            const double val = acc[sl][e];
            if (li <= lj && lj <= Lt && li < Lt && val != 0.0) {
              const int gi = li < 7 * Wt ? 7 * w0 + li : S.npc + (li - 7 * Wt);
              if (lj == Lt) atomicAdd(S.rhs + gi, -val);
              else {
                const int gj = lj < 7 * Wt ? 7 * w0 + lj : S.npc + (lj - 7 * Wt);
                addS(S, gi, gj, -val);
              }
            }
          }
          if (++bj == nb8) { bi++; bj = bi; }
        }
      }
    }
  }
  cost_acc = warp_sum(cost_acc);
  if (lane == 0 && cost_acc != 0.0) atomicAdd(a.b.cost, cost_acc);
#ifdef MVBA_TILE_PROFILE
  if (tid == 0) for (int i = 0; i < 8; i++) atomicAdd(&g_tile_prof[i], (unsigned long long)prof_[i]);
#endif
}

}  // namespace mvba

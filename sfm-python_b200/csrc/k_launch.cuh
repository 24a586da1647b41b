// Host-side launch wrappers (included at the end of ba_kernels.cu, namespace b200ba).

namespace {

struct Slots { int s[8]; };

__global__ void k_reduce_partials2(int n, int nsum, int nval, const double* __restrict__ part,
                                   double* __restrict__ out_sum, double* __restrict__ out_max, Slots slots) {
  // single CTA, fixed order: thread t sums partials t, t+256, ...; then a fixed tree
  __shared__ double sm[32 * 8];
  double v[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) v[k] = 0.0;
  for (int b = threadIdx.x; b < n; b += blockDim.x)
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (k < nval) {
        const double t = part[(int64_t)b * 8 + k];
        v[k] = (k < nsum) ? v[k] + t : fmax(v[k], t);
      }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const double r = (k < nsum) ? warp_sum(v[k]) : warp_max(v[k]);
    if (lane == 0) sm[wid * 8 + k] = r;
  }
  __syncthreads();
  if (threadIdx.x < nval) {
    const int k = threadIdx.x;
    double r = 0.0;
    for (int w = 0; w < nw; ++w) r = (k < nsum) ? r + sm[w * 8 + k] : fmax(r, sm[w * 8 + k]);
    if (k < nsum) out_sum[slots.s[k]] = r; else out_max[slots.s[k]] = r;
  }
}

inline unsigned cdiv(int64_t n, int t) { return (unsigned)std::max<int64_t>((n + t - 1) / t, 1); }

// FP64 FMA peak probe: 8 independent DFMA chains per thread (MEASURED_PEAKS.json holds no FP64 figure; bench.py
// divides the fused kernel's algorithmic flops by this number measured in the same run).
__global__ void __launch_bounds__(256) k_fp64_peak(int iters, double seed, double* __restrict__ sink) {
  double a[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = seed + k + threadIdx.x * 1e-9;
  const double m = 1.0 - 1e-12, c = 1e-13;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 8; ++u)
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] = a[k] * m + c;
  }
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  if (s == 123.456) sink[0] = s;  // never true: keeps the chains alive
}

}  // namespace

#ifndef REDUCE_T
#define REDUCE_T 128
#endif
static inline int point_grid(const Handle* h) {
  return (int)std::min<int64_t>(cdiv(h->pb.n_wt, LIN_WARPS), (int64_t)h->sm_count * 16);
}
static inline int obs_grid(const Handle* h) {
  return (int)std::min<int64_t>(cdiv(h->pb.N, 256), (int64_t)h->sm_count * 16);
}
static inline int fused_grid(const Handle* h) {
  if (h->pb.flow) return (int)std::min<int64_t>(std::max(h->pb.n_st, 1), (int64_t)h->sm_count);
  const int per_sm = h->pb.tile_obs <= FUSED_TILE_2CTA ? std::max(2, std::min(8, (FUSED_TILE_1CTA * FUSED_REC_BYTES) /
                                                                              std::max(h->pb.tile_obs * FUSED_REC_BYTES, 1)))
                                                       : 1;
  return (int)std::min<int64_t>(std::max(h->pb.n_st, 1), (int64_t)h->sm_count * per_sm);
}
int partial_capacity(const Handle* h) {
  return std::max(std::max(point_grid(h), obs_grid(h)), fused_grid(h));
}

// profiling helper: time the launches between begin() and end() with CUDA events on the library stream
struct KernelTimer {
  Handle* h; bool on;
  explicit KernelTimer(Handle* h_) : h(h_), on(h_->profile) {}
  void begin() { if (on) cudaEventRecord(h->ev2, h->stream); }
  void end(double* acc) {
    if (!on) return;
    cudaEventRecord(h->ev3, h->stream);
    cudaEventSynchronize(h->ev3);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev2, h->ev3);
    *acc += ms;
  }
};

void launch_cam_prep(Handle* h, const double* x_cams, const double* dc, double* x_out, double* camprep,
                     const double* camprep_old, double* cstep) {
  const int C = (int)h->pb.C;
  if (C == 0) return;
  k_cam_prep<<<cdiv(C, 128), 128, 0, h->stream>>>(C, h->pb.Cs, x_cams, dc, x_out, camprep, camprep_old, cstep);
  h->st.launches += 1;
}

void launch_residual(Handle* h, const double* x, const double* camprep, double* r_out, int scal_slot) {
  const int g = obs_grid(h);
  KernelTimer t(h);
  t.begin();
  k_residual<<<g, 256, 0, h->stream>>>(h->pb, x, camprep, r_out, h->w.part);
  t.end(&h->st.ms_residual_kernel);
  Slots sl{};
  sl.s[0] = scal_slot;
  k_reduce_partials2<<<1, 256, 0, h->stream>>>(g, 1, 1, h->w.part, h->w.scal, h->w.scal, sl);
  h->st.launches += 2;
}

void launch_pack_cameras(Handle* h, int64_t n, const double* H_dev, double* params_dev) {
  if (n <= 0) return;
  k_pack_cameras<<<cdiv(n, 128), 128, 0, h->stream>>>(n, H_dev, params_dev);
  h->st.launches += 1;
}
void launch_unpack_cameras(Handle* h, int64_t n, const double* params_dev, double* H_dev) {
  if (n <= 0) return;
  k_unpack_cameras<<<cdiv(n, 128), 128, 0, h->stream>>>(n, params_dev, H_dev);
  h->st.launches += 1;
}

void launch_pnp(Handle* h, int n_problems, const void* probs_dev, const double* K_dev, const double* guess_dev, int max_iter,
                double* out_dev) {
  if (n_problems <= 0) return;
  k_pnp<<<n_problems, PNP_T, 0, h->stream>>>(reinterpret_cast<const PnpProblem*>(probs_dev), K_dev, guess_dev, max_iter, out_dev);
  h->st.launches += 1;
}

void launch_triangulate(Handle* h, int64_t n, const double* M1, const double* M2, const double* pts1, const double* pts2,
                        const double* vis1, const double* vis2, double* X_out, unsigned char* mask_out) {
  if (n <= 0) return;
  k_triangulate<<<cdiv(n, 128), 128, 0, h->stream>>>(n, M1, M2, pts1, pts2, vis1, vis2, X_out, mask_out);
  h->st.launches += 1;
}

void launch_triangulate_refine(Handle* h, int64_t n, const double* M1, const double* M2, const double* pts1, const double* pts2,
                               int max_iter, double* X) {
  if (n <= 0) return;
  k_triangulate_refine<<<cdiv(n, 128), 128, 0, h->stream>>>(n, M1, M2, pts1, pts2, max_iter, X);
  h->st.launches += 1;
}

void launch_reproj_err(Handle* h, const double* Rt, const double* X, double* err_out) {
  const int g = obs_grid(h);
  KernelTimer t(h);
  t.begin();
  k_reproj_err<<<g, 256, 0, h->stream>>>(h->pb, Rt, X, err_out, h->w.part);
  t.end(&h->st.ms_reproj);
  Slots sl{};
  sl.s[0] = S_REPROJ_SUM;
  k_reduce_partials2<<<1, 256, 0, h->stream>>>(g, 1, 1, h->w.part, h->w.scal, h->w.scal, sl);
  h->st.launches += 2;
}

void launch_jacobian(Handle* h, const double* x, const double* camprep, double* r_out, double* Jc, double* Jp) {
  if (h->pb.N == 0) return;
  KernelTimer t(h);
  t.begin();
  k_jacobian<<<cdiv(h->pb.N, JT), JT, 0, h->stream>>>(h->pb, x, camprep, r_out, Jc, Jp);
  t.end(&h->st.ms_jacobian);
  h->st.launches += 1;
}

// residual + Jacobian blocks + J^T J / J^T r blocks + point-block Schur elimination at x (camprep must match x)
void launch_linearize(Handle* h, const double* x, double lambda) {
  Problem& pb = h->pb;
  Work& w = h->w;
  cudaStream_t st = h->stream;
  KernelTimer t(h);
  if (pb.fused) {
    static bool attr_dev[MAX_DEVICES] = {};   // function attributes are per device: one flag per ordinal
    bool& attr_set = attr_dev[h->device % MAX_DEVICES];
    if (!attr_set) {
      cudaFuncSetAttribute(k_lin_fused<256, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, FUSED_TILE_2CTA * FUSED_REC_BYTES);
      cudaFuncSetAttribute(k_lin_fused<512, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, FUSED_TILE_1CTA * FUSED_REC_BYTES);
      attr_set = true;
    }
    const int smem = pb.tile_obs * FUSED_REC_BYTES;
    const int fg = fused_grid(h);
    t.begin();
    if (pb.flow) {
      static bool flow_dev[MAX_DEVICES] = {};
      bool& flow_attr = flow_dev[h->device % MAX_DEVICES];
      if (!flow_attr) {
        cudaFuncSetAttribute(k_lin_flow, cudaFuncAttributeMaxDynamicSharedMemorySize, FlowSmem::TOTAL);
        flow_attr = true;
      }
      k_lin_flow<<<fg, FLOW_THREADS, FlowSmem::TOTAL, st>>>(pb, x, w.camprep, lambda, w.pp, w.gp, w.spart, w.dpart, w.part,
                                                         reinterpret_cast<int*>(w.scal + S_FLOW_ERR),
                                                         FLOW_PROF ? reinterpret_cast<long long*>(w.Y) : nullptr);
    } else if (pb.tile_obs <= FUSED_TILE_2CTA)
      k_lin_fused<256, 2><<<fg, 256, smem, st>>>(pb, x, w.camprep, lambda, w.pp, w.gp, w.spart, w.dpart, w.part);
    else
      k_lin_fused<512, 1><<<fg, 512, smem, st>>>(pb, x, w.camprep, lambda, w.pp, w.gp, w.spart, w.dpart, w.part);
    t.end(&h->st.ms_lin_fused);
    Slots sl{};
    sl.s[0] = S_COST2;
    sl.s[1] = S_GPMAX;
    // sparse exchange: this rank's contribution goes to xsend = [local blocks | b | g_c | diag U | sums | gpmax]
    double* xt = pb.xchg ? w.xsend + (int64_t)pb.xl_max * 36 : nullptr;
    double* o_b = pb.xchg ? xt : w.b;
    double* o_gc = pb.xchg ? xt + 6 * pb.C : w.gc;
    double* o_ud = pb.xchg ? xt + 12 * pb.C : w.udiag;
    double* o_sums = pb.xchg ? xt + 18 * pb.C : w.sums;
    k_reduce_partials2<<<1, 256, 0, st>>>(fg, 1, 2, w.part, o_sums, w.scal, sl);
    if (pb.xchg)
      cudaMemcpyAsync(o_sums + S_NSUM, w.scal + S_GPMAX, sizeof(double), cudaMemcpyDeviceToDevice, st);
    t.begin();
    k_cam_sums<<<cdiv(pb.C * 40, 128), 128, 0, st>>>((int)pb.C, pb.cam_dslot_off, pb.cam_dslots, w.dpart, w.camsum, o_ud,
                                                    o_gc, o_b);
    if (pb.xchg) {
      const char* e_dp = getenv("B200BA_DIRECT_PUSH");   // 0: stage in xsend and copy afterwards (A/B)
      PeerDst dst = (e_dp && atoi(e_dp) == 0) ? PeerDst{} : p2p_next_dst(h);
      h->direct_push = dst.n > 0;
      k_slot_reduce_S<<<cdiv((int64_t)pb.xl_n * 36, REDUCE_T), REDUCE_T, 0, st>>>(pb.xl_n, pb.xl_blk, pb.blk_r, pb.blk_c,
                                                                        pb.blk_slot_off, pb.blk_slots, w.spart, w.camsum,
                                                                        lambda, w.xsend, dst);
    } else {
      k_slot_reduce_S<<<cdiv((int64_t)pb.n_blk * 36, REDUCE_T), REDUCE_T, 0, st>>>(pb.n_blk, nullptr, pb.blk_r, pb.blk_c,
                                                                         pb.blk_slot_off, pb.blk_slots, w.spart, w.camsum,
                                                                         lambda, w.Sup, PeerDst{});
    }
    t.end(&h->st.ms_slot_reduce);
    h->st.launches += 4;
    h->st.n_linearize += 1;
    return;
  }
  const int pg = point_grid(h);
  t.begin();
  k_lin_points<<<pg, LIN_WARPS * 32, 0, st>>>(pb, x, w.camprep, lambda, w.Y, w.rr, w.pp, w.gp, w.part);
  t.end(&h->st.ms_linearize_points);
  Slots sl{};
  sl.s[0] = S_COST2;
  sl.s[1] = S_GPMAX;
  k_reduce_partials2<<<1, 256, 0, st>>>(pg, 1, 2, w.part, w.sums, w.scal, sl);
  const int cg = (int)std::min<int64_t>(cdiv(pb.n_cch, 8), (int64_t)h->sm_count * 16);
  t.begin();
  k_cam_side<<<cg, 256, 0, st>>>(pb, x, w.camprep, w.rr, w.cpart);
  k_cam_reduce<<<cdiv(pb.C * 33, 256), 256, 0, st>>>((int)pb.C, pb.cam_choff, w.cpart, w.U, w.gc, w.b, w.udiag);
  t.end(&h->st.ms_camera_blocks);
  static int schur_dev[MAX_DEVICES];
  static bool schur_init = false;
  if (!schur_init) { for (int i = 0; i < MAX_DEVICES; ++i) schur_dev[i] = -1; schur_init = true; }
  int& schur_mode = schur_dev[h->device % MAX_DEVICES];
  if (schur_mode < 0) {
    cudaFuncSetAttribute(k_schur<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, SCH_SMEM);
    cudaFuncSetAttribute(k_schur<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, SCH_SMEM);
    cudaFuncSetAttribute(k_schur<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SCH_SMEM);
    cudaFuncSetAttribute(k_schur_tiles, cudaFuncAttributeMaxDynamicSharedMemorySize, SCHT_SMEM_MAX);
    const char* e = getenv("B200BA_SCHUR_DIAG");  // diagnostic only (wrong results): 1 = fetch only, 2 = multiply only
    schur_mode = e ? atoi(e) : 0;
  }
  if (pb.schur_tiles) {
    const int smem = pb.tile_obs * 144 + SCHT_ENT_CAP * 4;
    const int per_sm = std::max(1, std::min(4, (220 * 1024) / std::max(smem, 1)));
    const int tg = (int)std::min<int64_t>(std::max(pb.n_st, 1), (int64_t)h->sm_count * per_sm);
    t.begin();
    k_schur_tiles<<<tg, SCHT_THREADS, smem, st>>>(pb, w.Y, w.spart);
    t.end(&h->st.ms_schur);
    k_schur_reduce_tiles<<<cdiv((int64_t)pb.n_blk * 36, 256), 256, 0, st>>>(
        pb.n_blk, pb.blk_r, pb.blk_c, pb.blk_slot_off, pb.blk_slots, w.spart, w.U, lambda, w.Sup);
  } else {
    const int sg = (int)std::min<int64_t>(cdiv(pb.n_sch, SCH_WARPS), (int64_t)h->sm_count);
    t.begin();
    if (schur_mode == 1) k_schur<1><<<sg, SCH_WARPS * 32, SCH_SMEM, st>>>(pb, w.Y, w.spart);
    else if (schur_mode == 2) k_schur<2><<<sg, SCH_WARPS * 32, SCH_SMEM, st>>>(pb, w.Y, w.spart);
    else k_schur<0><<<sg, SCH_WARPS * 32, SCH_SMEM, st>>>(pb, w.Y, w.spart);
    t.end(&h->st.ms_schur);
    k_schur_reduce<<<cdiv((int64_t)pb.n_blk * 36, 256), 256, 0, st>>>(pb.n_blk, pb.blk_r, pb.blk_c, pb.blk_choff,
                                                                      w.spart, w.U, lambda, w.Sup);
  }
  h->st.launches += 6;
  h->st.n_linearize += 1;
}

void launch_xchg_sum(Handle* h) {
  Problem& pb = h->pb;
  Work& w = h->w;
  const int64_t n = (int64_t)pb.n_blk * 36 + 18 * pb.C + 1;
  k_xchg_sum<<<cdiv(n, 256), 256, 0, h->stream>>>(pb.n_blk, (int)pb.C, h->nranks, pb.x_stride, (long long)pb.xl_max * 36,
                                                 pb.xc_off, pb.xc_src, w.xrecv, w.comm, w.sums, w.scal);
  h->st.launches += 1;
}

void launch_fix_diag(Handle* h) {
  k_fix_diag<<<cdiv(6 * h->pb.C, 256), 256, 0, h->stream>>>((int)h->pb.C, h->pb.diag_blk, h->w.udiag, h->w.Sup);
  h->st.launches += 1;
}

void launch_fix_cameras(Handle* h) {
  Problem& pb = h->pb;
  if (!pb.cam_fixed) return;
  const int64_t n = std::max<int64_t>((int64_t)pb.n_blk * 36, 6 * pb.C);
  k_fix_cameras<<<cdiv(n, 256), 256, 0, h->stream>>>(pb.n_blk, (int)pb.C, pb.blk_r, pb.blk_c, pb.cam_fixed, h->w.Sup, h->w.b,
                                                     h->w.gc);
  h->st.launches += 1;
}

// delta_p, x_new (points), trial cost partial sums -> w.sums[S_NEWCOST2 .. S_XP2]; camprep_n / xn cameras must be set
void launch_backsub_trial(Handle* h, double lambda) {
  (void)lambda;
  Work& w = h->w;
  const int pg = point_grid(h);
  k_backsub_trial<<<pg, LIN_WARPS * 32, 0, h->stream>>>(h->pb, w.x, w.xn, w.camprep, w.camprep_n, w.cstep, w.pp, w.part);
  Slots sl{};
  sl.s[0] = S_NEWCOST2; sl.s[1] = S_GP_DP; sl.s[2] = S_DP_D_DP; sl.s[3] = S_DP2; sl.s[4] = S_XP2;
  k_reduce_partials2<<<1, 256, 0, h->stream>>>(pg, 5, 5, w.part, w.sums, w.scal, sl);
  h->st.launches += 2;
}

void launch_cam_dots(Handle* h) {
  Work& w = h->w;
  const int n = (int)(6 * h->pb.C);
  const int nparts = std::max(1, std::min(std::min(w.n_part, 64), (n + 1023) / 1024));
  const int chunk = (n + nparts - 1) / nparts;
  k_cam_dots<<<nparts, 1024, 0, h->stream>>>(n, chunk, w.gc, w.dc, w.udiag, w.x, w.part, w.scal);
  if (nparts > 1) k_cam_dots_final<<<1, 32, 0, h->stream>>>(nparts, w.part, w.scal);
  h->st.launches += nparts > 1 ? 2 : 1;
}

void launch_densify(Handle* h, double* dense) {
  const int64_t n = 6 * h->pb.C;
  cudaMemsetAsync(dense, 0, (size_t)(n * n) * sizeof(double), h->stream);
  k_densify<<<cdiv((int64_t)h->pb.n_blk * 36, 256), 256, 0, h->stream>>>(h->pb.n_blk, (int)h->pb.C, h->pb.blk_r,
                                                                        h->pb.blk_c, h->w.Sup, dense);
  h->st.launches += 1;
}

// measured FP64 FMA throughput of the device in TFLOP/s (best of 5 timed launches after a warm-up)
double measure_fp64_peak(Handle* h) {
  const char* e_bps = getenv("B200BA_PEAK_BLOCKS_PER_SM");  // occupancy experiments (default: 8 x 256 threads = full)
  const int bps = (e_bps && atoi(e_bps) > 0) ? atoi(e_bps) : 8;
  const int iters = 4096, blocks = h->sm_count * bps, threads = 256;
  float best = 1e30f;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(h->ev0, h->stream);
    k_fp64_peak<<<blocks, threads, 0, h->stream>>>(iters, 1.0, h->w.scal);
    cudaEventRecord(h->ev1, h->stream);
    cudaEventSynchronize(h->ev1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    if (rep > 0 && ms < best) best = ms;
  }
  h->st.launches += 6;
  const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
  return flops / ((double)best * 1e-3) / 1e12;
}

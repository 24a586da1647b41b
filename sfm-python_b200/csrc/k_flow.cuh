// Dataflow version of the fused linearisation + Schur kernel (included inside ba_kernels.cu's anonymous namespace).
//
// Same arithmetic and the same slot partials as k_lin_fused (k_fused.cuh), different schedule.  ONE persistent CTA of
// 512 threads per SM walks its tiles k = 0, 1, 2, ... with no CTA-wide barrier:
//
//   * two record buffers in shared memory: while the slots of tile k are multiplied out of buffer k % 2, the records of
//     tile k + 1 are produced into the other one;
//   * all work is cut into warp-sized ITEMS in one monotone queue (a shared-memory counter):
//         step k  =  [ A(k+1): one warp tile of tile k+1 -> records ]  [ C(k): 8 pair slots ]  [ B(k): 8 self slots ]
//     a warp claims the next item, checks the item's dependency on shared-memory completion counters (acquire /
//     release), executes it and signals.  An item only depends on items earlier in the queue, so the schedule cannot
//     deadlock and latency-bound A items overlap the FP64-bound B / C items of the previous tile;
//   * the inputs of a tile are staged by the bulk-copy engine (cp.async.bulk + mbarrier complete_tx, the non-tensor
//     TMA path): the cam[] / pt[] windows, the tile's camera window of the field-major prep table (21 rows) and the
//     tile's metadata blob (slot descriptors + entries).  The warp whose signal frees a staging slot issues the
//     copies for the tile two steps ahead.  u, v and the point coordinates are read with plain coalesced loads.
//
// Outputs are bit-identical to k_lin_fused: every slot partial is produced by one lane group in a fixed order.

#ifndef B200BA_FLOW_THREADS
#define B200BA_FLOW_THREADS 512
#endif
constexpr int FLOW_THREADS = B200BA_FLOW_THREADS;
constexpr int FLOW_WARPS = FLOW_THREADS / 32;
constexpr int FLOW_SPI = 32 / FG;  // slots per item
constexpr int FLOW_MAX_WT = 2 * FLOW_TILE / 32 + 4;  // warp tiles per tile: two consecutive ones hold > 32 observations
constexpr int FLOW_WTO_INTS = 48;                    // staged wt_off window: FLOW_MAX_WT + 1 offsets + alignment slack, 16-byte units

struct FlowSmem {
  // byte offsets inside the dynamic shared memory block
  static constexpr int REC = 0;                                              // [2][FLOW_TILE * 22] doubles
  static constexpr int CAMP = REC + 2 * FLOW_TILE * FREC * 8;                // [2][21 * 64] doubles
  static constexpr int IDXC = CAMP + 2 * CAMPREP_FIELDS * FLOW_CAM_SPAN * 8; // [2][FLOW_IDX_BYTES] cam window
  static constexpr int IDXP = IDXC + 2 * FLOW_IDX_BYTES;                     // [2][FLOW_IDX_BYTES] pt window
  static constexpr int META = IDXP + 2 * FLOW_IDX_BYTES;                     // [2][FLOW_META_CAP]
  static constexpr int BAR = META + 2 * FLOW_META_CAP;                       // 8 mbarriers: inA[2], meta[2], doneA[2], doneBC[2]
  static constexpr int INFO = BAR + 8 * 8;                                   // [2][4] int4: staged tile_info of the A tile
  static constexpr int CTR = INFO + 2 * 64;                                  // q_next, doneA[2], doneBC[2], error
  static constexpr int RED = CTR + 8 * 4;                                    // [2][FLOW_MAX_WT][2] doubles: (sum r^2, max |g_p|)
  static constexpr int TSUM = RED + 2 * FLOW_MAX_WT * 2 * 8;                 // [64][2] doubles: tile sums, ring by k % 64
  static constexpr int WST = TSUM + 64 * 2 * 8;                              // [FLOW_WARPS][6] int4: per-warp walk state ring
  static constexpr int WTO = WST + FLOW_WARPS * 96;                          // [2][FLOW_WTO_INTS] staged wt_off window
  static constexpr int TOTAL = WTO + 2 * FLOW_WTO_INTS * 4;
};
static_assert(FlowSmem::TOTAL <= 232448, "dataflow kernel shared memory exceeds the 227 KB opt-in limit");

__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ int ld_acquire_s(const int* p) {
  int v;
  asm volatile("ld.acquire.cta.shared.s32 %0, [%1];\n" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ int add_release_s(int* p, int v) {
  int old;
  asm volatile("atom.add.acq_rel.cta.shared.s32 %0, [%1], %2;\n" : "=r"(old) : "r"(smem_u32(p)), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, P1;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

#ifndef FLOW_PROF
#define FLOW_PROF 0
#endif
#ifndef FLOW_MBAR_DEPS
#define FLOW_MBAR_DEPS 1   // item dependencies wait on mbarriers (hardware-suspended try_wait) instead of polling counters
#endif
#ifndef FLOW_POLL_NS
#define FLOW_POLL_NS 32
#endif
#ifndef FLOW_COOP_POINT_STORE
#define FLOW_COOP_POINT_STORE 1
#endif
#ifndef FLOW_WALK_RING
#define FLOW_WALK_RING 0   // walk state prefetched with cp.async (0: blocking global load at every step boundary)
#endif
#ifndef FLOW_STAGE_WTO
#define FLOW_STAGE_WTO 0   // the tile's warp-tile offsets staged in shared memory with the other inputs
#endif
constexpr int FLOW_SPIN_LIMIT = 1 << 24;  // bail out instead of hanging the GPU if a dependency never resolves

__device__ __forceinline__ void l2_prefetch(const void* p, unsigned bytes) {  // p 16-byte aligned, bytes % 16 == 0
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;\n" ::"l"(p), "r"(bytes) : "memory");
}

struct FlowCtx {
  unsigned char* sm;
  const Problem* pb;
  const double* camprep;
  const double* xp;
  int* ctr;                    // [0] q_next, [1..2] doneA, [3..4] doneBC, [5] error
  unsigned long long* bars;    // [0..1] inA, [2..3] meta
};

// one elected lane: stage cam[] / pt[] windows and the camera window of tile t (sequence index k) into slot k % 2
__device__ __forceinline__ void flow_issue_inA(const FlowCtx& c, int t, int k) {
  const Problem& pb = *c.pb;
  const int4 i1 = pb.tile_info[4 * t + 1], i2 = pb.tile_info[4 * t + 2];
  const int j = k & 1;
  const int a0 = i2.x, n = i2.y, a0a = a0 & ~3;
  const unsigned ib = (unsigned)(((a0 + n - a0a) * 4 + 15) & ~15);
  const int cmin = i1.z, cspan = (pb.tile_info[4 * t].w & 2) ? i1.w : 0;
  unsigned long long* bar = c.bars + j;
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  const int wt0a = i2.z & ~3;
  const int nA = pb.tile_info[4 * t].x;
  const unsigned wb = (FLOW_STAGE_WTO && nA + 1 <= FLOW_MAX_WT + 1) ? (unsigned)(((i2.z + nA + 1 - wt0a) * 4 + 15) & ~15) : 0u;
  mbar_expect_tx(bar, 2 * ib + (unsigned)(CAMPREP_FIELDS * cspan * 8) + 64u + wb);
  bulk_g2s(c.sm + FlowSmem::INFO + j * 64, pb.tile_info + 4 * t, 64u, bar);
  if (wb) bulk_g2s(c.sm + FlowSmem::WTO + j * (FLOW_WTO_INTS * 4), pb.wt_off + wt0a, wb, bar);
  bulk_g2s(c.sm + FlowSmem::IDXC + j * FLOW_IDX_BYTES, pb.cam + a0a, ib, bar);
  bulk_g2s(c.sm + FlowSmem::IDXP + j * FLOW_IDX_BYTES, pb.pt + a0a, ib, bar);
  if (cspan > 0) {
    double* dst = reinterpret_cast<double*>(c.sm + FlowSmem::CAMP) + j * (CAMPREP_FIELDS * FLOW_CAM_SPAN);
#pragma unroll 1
    for (int f = 0; f < CAMPREP_FIELDS; ++f)
      bulk_g2s(dst + f * FLOW_CAM_SPAN, c.camprep + (int64_t)f * pb.Cs + cmin, (unsigned)(cspan * 8), bar);
  }
  // u, v and the tile's points are read with plain loads two steps from now: pull them into L2 meanwhile
  const int4 i3 = pb.tile_info[4 * t + 3];
  const int a0e = a0 & ~1;
  const unsigned ub = (unsigned)(((a0 + n - a0e) * 8 + 15) & ~15);
  l2_prefetch(pb.u + a0e, ub);
  l2_prefetch(pb.v + a0e, ub);
  const long long x0 = (3ll * i3.z) & ~1ll;
  l2_prefetch(c.xp + x0, (unsigned)(((3ll * (i3.z + i3.w) - x0) * 8 + 15) & ~15ll));
}

__device__ __forceinline__ void flow_issue_meta(const FlowCtx& c, int t, int k) {
  const Problem& pb = *c.pb;
  const int4 i0 = pb.tile_info[4 * t], i1 = pb.tile_info[4 * t + 1];
  unsigned long long* bar = c.bars + 2 + (k & 1);
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  if (i0.w & 1) {
    mbar_expect_tx(bar, (unsigned)i1.y * 16u);
    bulk_g2s(c.sm + FlowSmem::META + (k & 1) * FLOW_META_CAP, pb.blob + i1.x, (unsigned)i1.y * 16u, bar);
  } else {
    mbar_expect_tx(bar, 0u);  // blob too large for the staging slot: consumers read it from global memory
  }
}

// all lanes: wait until the mbarrier generation `gen` of `bar` has completed
__device__ __forceinline__ bool flow_wait_bar(unsigned long long* bar, int gen, int* err) {
  int spins = 0;
  while (!mbar_try_wait(bar, (unsigned)(gen & 1))) {
    if (spins > 64) __nanosleep(64);
    if (++spins > FLOW_SPIN_LIMIT) { *err = 1; return false; }
  }
  return true;
}
// all lanes: wait until *ctr >= target
__device__ __forceinline__ bool flow_wait_ctr(const int* ctr, int target, int* err) {
  int spins = 0;
  while (ld_acquire_s(ctr) < target) {
    __nanosleep(FLOW_POLL_NS);
    if (++spins > FLOW_SPIN_LIMIT) { *err = 1; return false; }
  }
  return true;
}

__global__ void __launch_bounds__(FLOW_THREADS, 1)
k_lin_flow(const __grid_constant__ Problem pb, const double* __restrict__ x, const double* __restrict__ camprep, double lambda,
           double* __restrict__ pp, double* __restrict__ gp, double* __restrict__ spart, double* __restrict__ dpart,
           double* __restrict__ part, int* __restrict__ gerr, long long* __restrict__ prof) {
  extern __shared__ __align__(128) unsigned char fsm[];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int g = lane % FG, sub = lane / FG;
  const double* xp = x + 6 * pb.C;
  int* ctr = reinterpret_cast<int*>(fsm + FlowSmem::CTR);
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(fsm + FlowSmem::BAR);
  double* sred = reinterpret_cast<double*>(fsm + FlowSmem::RED);
  FlowCtx cx{fsm, &pb, camprep, xp, ctr, bars};
  const int K = (pb.n_st - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // tiles of this CTA
  auto tile_of = [&](int k) { return (int)blockIdx.x + k * (int)gridDim.x; };

  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) ctr[i] = 0;
    for (int i = 0; i < 128; ++i) reinterpret_cast<double*>(fsm + FlowSmem::TSUM)[i] = 0.0;
    for (int i = 0; i < 8; ++i) mbar_init(bars + i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 0; k < 2 && k < K; ++k) { flow_issue_inA(cx, tile_of(k), k); flow_issue_meta(cx, tile_of(k), k); }
  }

  long long tA = 0, tC = 0, tB = 0, tW = 0, tQ = 0, nIt = 0;  // FLOW_PROF: cycles per item kind, waits, queue
  long long tA1 = 0, tA2 = 0, tA3 = 0, cA = 0;              // A item: loads + projection | segmented scan | 3x3 factor
  (void)tA1; (void)tA2; (void)tA3; (void)cA;
  (void)tA; (void)tC; (void)tB; (void)tW; (void)tQ; (void)nIt;
  // ---- item bodies ----------------------------------------------------------------------------------------------
  // A item: warp tile r of tile kk (sequence index) -> records in buffer kk & 1
  auto do_A = [&](int kk, int r, int nA, int tgtA, int tgtBC_prev) -> bool {
    const int j = kk & 1, t = tile_of(kk);
    // record buffer j was last read by the B / C items of tile kk - 2; the staged inputs of tile kk must have landed
    const long long c0 = FLOW_PROF ? clock64() : 0;
    bool ok = (FLOW_MBAR_DEPS ? (kk < 2 || flow_wait_bar(bars + 6 + j, (kk - 2) >> 1, &ctr[5]))
                              : flow_wait_ctr(&ctr[3 + j], tgtBC_prev, &ctr[5])) &&
              flow_wait_bar(bars + j, kk >> 1, &ctr[5]);
    ok = __all_sync(FULL, ok);
    if (!ok) return false;
    const long long c1 = FLOW_PROF ? clock64() : 0;
    const int4* sinfo = reinterpret_cast<const int4*>(fsm + FlowSmem::INFO + j * 64);
    const int4 i0 = sinfo[0], i1 = sinfo[1], i2 = sinfo[2];
    const int a0 = i2.x, wt = i2.z + r;
    const int* s_cam = reinterpret_cast<const int*>(fsm + FlowSmem::IDXC + j * FLOW_IDX_BYTES) + (a0 & 3);
    const int* s_pt = reinterpret_cast<const int*>(fsm + FlowSmem::IDXP + j * FLOW_IDX_BYTES) + (a0 & 3);
    const double* s_cp = reinterpret_cast<const double*>(fsm + FlowSmem::CAMP) + j * (CAMPREP_FIELDS * FLOW_CAM_SPAN);
    const bool cams_staged = (i0.w & 2) != 0;
    const int cmin = i1.z;
    double* rec = reinterpret_cast<double*>(fsm + FlowSmem::REC) + j * (FLOW_TILE * FREC);
    // the tile's warp-tile offsets come from the staged window (one dependent trip to global memory less per item)
    const int* s_wto = reinterpret_cast<const int*>(fsm + FlowSmem::WTO + j * (FLOW_WTO_INTS * 4)) + (i2.z & 3);
    const bool wto_staged = FLOW_STAGE_WTO && i0.x <= FLOW_MAX_WT;
    const int a = wto_staged ? s_wto[r] : pb.wt_off[wt], b = wto_staged ? s_wto[r + 1] : pb.wt_off[wt + 1], n = b - a;
    double cost2 = 0.0, gmax = 0.0;
    auto get_cam = [&](int c, double* cp) {
      if (cams_staged) {
#pragma unroll
        for (int f = 0; f < CAMPREP_FIELDS; ++f) cp[f] = s_cp[f * FLOW_CAM_SPAN + (c - cmin)];
      } else {
        load_cam(camprep, pb.Cs, c, cp);
      }
    };
    // cost / gradient-norm partial of this warp tile; the last finisher of the tile sums them in index order, so the
    // result does not depend on which warp ran which item
    auto flow_signal_A = [&](double cost2_, double gmax_) {
      cost2_ = warp_sum(cost2_);
      gmax_ = warp_max(gmax_);
      double* tred = sred + j * (2 * FLOW_MAX_WT);
      if (lane == 0) { tred[2 * r] = cost2_; tred[2 * r + 1] = gmax_; }
      __threadfence_block();
      __syncwarp();
      if (lane == 0) {
        const int old = add_release_s(&ctr[1 + j], 1);
        if (old + 1 == tgtA - 1) {
          double s2 = 0.0, m2 = 0.0;
          for (int w = 0; w < nA; ++w) { s2 += tred[2 * w]; m2 = fmax(m2, tred[2 * w + 1]); }
          // per-CTA ring: tiles k and k + 64 are strictly ordered by the dependency chain => fixed summation order
          double* ts = reinterpret_cast<double*>(fsm + FlowSmem::TSUM) + 2 * (kk & 63);
          ts[0] += s2;
          ts[1] = fmax(ts[1], m2);
          add_release_s(&ctr[1 + j], 1);  // now the consumers of the tile may start (and, later, reuse of tred)
          if (FLOW_MBAR_DEPS) mbar_arrive(bars + 4 + j);
          // last A item of tile kk: its staging slot is free -> stage the inputs of tile kk + 2
          if (kk + 2 < K) flow_issue_inA(cx, tile_of(kk + 2), kk + 2);
        }
      }
    };
    if (n <= 32) {
      const bool valid = lane < n;
      const int i = a + (valid ? lane : 0);
      const double uo = pb.u[i], vo = pb.v[i];
      const int c = s_cam[i - a0], p = s_pt[i - a0];
      const double X0 = xp[3 * p], X1 = xp[3 * p + 1], X2 = xp[3 * p + 2];
      double r0, r1, Jc[12], Jp[6], vals[9], cp[CAMPREP_FIELDS];
      get_cam(c, cp);
      double c2;
      project_jac_loss(cp, pb.K, pb.loss, X0, X1, X2, uo, vo, r0, r1, Jc, Jp, c2);
      point_products(Jp, r0, r1, vals);
      if (!valid) {
#pragma unroll
        for (int q2 = 0; q2 < 9; ++q2) vals[q2] = 0.0;
      }
      if (FLOW_PROF) { const long long cc = clock64(); tA1 += cc - c1; cA = cc; }
      const SegInfo sg = seg_info(p, valid, lane);
      seg_allsum<9>(vals, sg, lane);
      if (FLOW_PROF) { const long long cc = clock64(); tA2 += cc - cA; cA = cc; }
      double Li[6], z[3];
      point_factor(vals, lambda, Li);
      z[0] = Li[0] * vals[6];
      z[1] = Li[1] * vals[6] + Li[2] * vals[7];
      z[2] = Li[3] * vals[6] + Li[4] * vals[7] + Li[5] * vals[8];
      if (FLOW_PROF) { const long long cc = clock64(); tA3 += cc - cA; cA = cc; }
      if (valid) {
        cost2 += c2;
        store_record(rec + FREC * (i - a0), Jc, Jp, Li, z, r0, r1);
        if (lane == sg.first) gmax = fmax(gmax, fmax(fabs(vals[6]), fmax(fabs(vals[7]), fabs(vals[8]))));
      }
      // the records are complete: signal first, write the per-point outputs to global memory afterwards (the fence
      // below must not wait for them)
      flow_signal_A(cost2, gmax);
      if (FLOW_COOP_POINT_STORE) {
        if (valid) store_point_coop(pp, gp, p, Li, z, vals, lane - sg.first, sg.last - sg.first + 1);
      } else if (valid && lane == sg.first) store_point(pp, gp, p, Li, z, vals);
      if (FLOW_PROF) { const long long ce = clock64(); tW += c1 - c0; tA += ce - c1; }
      return true;
    } else {
      const int p = s_pt[a - a0];
      const double X0 = xp[3 * p], X1 = xp[3 * p + 1], X2 = xp[3 * p + 2];
      double vals[9];
#pragma unroll
      for (int q2 = 0; q2 < 9; ++q2) vals[q2] = 0.0;
      for (int i = a + lane; i < b; i += 32) {
        double r0, r1, Jc[12], Jp[6], v9[9], cp[CAMPREP_FIELDS];
        get_cam(s_cam[i - a0], cp);
        double c2;
        project_jac_loss(cp, pb.K, pb.loss, X0, X1, X2, pb.u[i], pb.v[i], r0, r1, Jc, Jp, c2);
        point_products(Jp, r0, r1, v9);
#pragma unroll
        for (int q2 = 0; q2 < 9; ++q2) vals[q2] += v9[q2];
        cost2 += c2;
      }
#pragma unroll
      for (int q2 = 0; q2 < 9; ++q2) vals[q2] = warp_sum(vals[q2]);
      double Li[6], z[3];
      point_factor(vals, lambda, Li);
      z[0] = Li[0] * vals[6];
      z[1] = Li[1] * vals[6] + Li[2] * vals[7];
      z[2] = Li[3] * vals[6] + Li[4] * vals[7] + Li[5] * vals[8];
      if (lane == 0) {
        store_point(pp, gp, p, Li, z, vals);
        gmax = fmax(gmax, fmax(fabs(vals[6]), fmax(fabs(vals[7]), fabs(vals[8]))));
      }
      for (int i = a + lane; i < b; i += 32) {
        double r0, r1, Jc[12], Jp[6], cp[CAMPREP_FIELDS];
        get_cam(s_cam[i - a0], cp);
        double c2;
        project_jac_loss(cp, pb.K, pb.loss, X0, X1, X2, pb.u[i], pb.v[i], r0, r1, Jc, Jp, c2);
        store_record(rec + FREC * (i - a0), Jc, Jp, Li, z, r0, r1);
      }
    }
    flow_signal_A(cost2, gmax);
    if (FLOW_PROF) { const long long ce = clock64(); tW += c1 - c0; tA += ce - c1; }
    return true;
  };

  // B / C item r of tile k: 8 slots, 4 lanes each, records read from buffer k & 1
  auto do_BC = [&](auto staged_tag, int k, int r, int4 i0, int tgtA, int tgtBC) -> bool {
    constexpr bool STAGED = decltype(staged_tag)::value;  // blob in shared memory (LDS) or too large: global memory
    const int j = k & 1, t = tile_of(k);
    const long long c0 = FLOW_PROF ? clock64() : 0;
    bool ok = (FLOW_MBAR_DEPS ? flow_wait_bar(bars + 4 + j, k >> 1, &ctr[5]) : flow_wait_ctr(&ctr[1 + j], tgtA, &ctr[5])) &&
              flow_wait_bar(bars + 2 + j, k >> 1, &ctr[5]);
    ok = __all_sync(FULL, ok);
    if (!ok) return false;
    __threadfence_block();
    const long long c1 = FLOW_PROF ? clock64() : 0;
    const int* blob;
    if (STAGED) blob = reinterpret_cast<const int*>(fsm + FlowSmem::META + j * FLOW_META_CAP);
    else blob = reinterpret_cast<const int*>(pb.blob + pb.tile_info[4 * t + 1].x);
    const int np = i0.y, nd = i0.z, nC = (np + FLOW_SPI - 1) / FLOW_SPI;
    const int pbase = blob[4], dbase = blob[5];
    const double* rec = reinterpret_cast<const double*>(fsm + FlowSmem::REC) + j * (FLOW_TILE * FREC);
    const int* pdesc = blob + 8;
    const int* ddesc = pdesc + 2 * np;
    const unsigned* pent = reinterpret_cast<const unsigned*>(ddesc + 2 * nd);
    const unsigned short* dent = reinterpret_cast<const unsigned short*>(pent + blob[2]);
    if (r < nC) {
      const int s = r * FLOW_SPI + sub;
      const bool active = s < np;
      const int ea = active ? pdesc[2 * s] : 0, en = active ? pdesc[2 * s + 1] : 0;
      double acc[36];
#pragma unroll
      for (int q2 = 0; q2 < 36; ++q2) acc[q2] = 0.0;
      pair_slot_body(rec, pent + ea, en, g, acc);
      slot_reduce_store<36>(acc, spart + 36 * (int64_t)(pbase + s), g, active);
    } else {
      const int s = (r - nC) * FLOW_SPI + sub;
      const bool active = s < nd;
      const int ea = active ? ddesc[2 * s] : 0, en = active ? ddesc[2 * s + 1] : 0;
      double acc[40];
#pragma unroll
      for (int q2 = 0; q2 < 40; ++q2) acc[q2] = 0.0;
      self_slot_body(rec, dent + ea, en, g, acc);
      slot_reduce_store<40>(acc, dpart + DPART * (int64_t)(dbase + s), g, active);
    }
    __syncwarp();
    if (lane == 0) {
      const int old = add_release_s(&ctr[3 + j], 1);
      // last consumer of tile k: its metadata slot is free -> stage the blob of tile k + 2
      if (old + 1 == tgtBC) {
        if (FLOW_MBAR_DEPS) mbar_arrive(bars + 6 + j);   // the record buffer of tile k is free for the A items of tile k + 2
        if (k + 2 < K) flow_issue_meta(cx, tile_of(k + 2), k + 2);
      }
    }
    if (FLOW_PROF) { const long long ce = clock64(); tW += c1 - c0; if (r < nC) tC += ce - c1; else tB += ce - c1; }
    return true;
  };

  // ---- walk (identical in every warp).  Tiles are handled in PAIRS so that an item never depends on the group of
  // items right before it in the queue:   step 2m   = [ A(2m) | A(2m+1) ]      step 2m+1 = [ C,B(2m) | C,B(2m+1) ]
  // A(2m+2) needs B/C(2m) (one whole group earlier), B/C(2m) needs A(2m) (ditto): a warp that claims an item normally
  // finds its dependency already satisfied.
  int s = -1, base = 0, n0 = 0, n1 = 0;                  // current step, first queue index, item counts of its two halves
  const int4 zero4 = make_int4(0, 0, 0, 0);
  // per-warp walk state in shared memory (keeps 16 registers free for the item bodies): [0] / [1] = tile_info[.][0] of
  // the even / odd tile of the current pair, [2] / [3] = the pair after it, loaded one pair ahead so that the walk
  // never waits on global memory
  // ring of three pairs: [0] / [1] current, [2] / [3] next (arrived), [4] / [5] in flight (cp.async, no registers held,
  // no stall on the global load when a step boundary is crossed)
  int4* wst = reinterpret_cast<int4*>(fsm + FlowSmem::WST) + 6 * wid;
  auto walk_fetch = [&](int slot, int k) {   // lane 0 only
    if (k < K && !FLOW_WALK_RING) {
      wst[slot] = pb.tile_info[4 * tile_of(k)];
    } else if (k < K) {
      asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(smem_u32(wst + slot)), "l"(pb.tile_info + 4 * tile_of(k))
                   : "memory");
    } else {
      wst[slot] = zero4;
    }
  };
  if (lane == 0) {
    wst[0] = zero4; wst[1] = zero4;
    wst[2] = K > 0 ? pb.tile_info[4 * tile_of(0)] : zero4;
    wst[3] = K > 1 ? pb.tile_info[4 * tile_of(1)] : zero4;
    walk_fetch(4, 2);
    walk_fetch(5, 3);
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  }
  __syncwarp();
  // cumulative completion targets per buffer parity (A targets count the items plus the finisher's extra signal)
  int cumA0 = 0, cumA1 = 0, cumBC0 = 0, cumBC1 = 0, prevBC0 = 0, prevBC1 = 0;
  const int n_steps = 2 * ((K + 1) / 2);
  bool ok = true;
  const long long cstart = FLOW_PROF ? clock64() : 0;
  while (ok) {
    const long long cq0 = FLOW_PROF ? clock64() : 0;
    int q = 0;
    if (lane == 0) q = atomicAdd(&ctr[0], 1);
    q = __shfl_sync(FULL, q, 0);
    bool done = false;
    while (q >= base + n0 + n1) {
      base += n0 + n1;
      ++s;
      if (s >= n_steps) { done = true; break; }
      const int m = s >> 1;
      if (!(s & 1)) {
        __syncwarp();
        if (lane == 0) {
          asm volatile("cp.async.wait_group 0;\n" ::: "memory");   // issued one pair of steps ago: long since landed
          wst[0] = wst[2];
          wst[1] = wst[3];
          wst[2] = wst[4];
          wst[3] = wst[5];
          walk_fetch(4, 2 * m + 4);
          walk_fetch(5, 2 * m + 5);
          asm volatile("cp.async.commit_group;\n" ::: "memory");
        }
        __syncwarp();
        n0 = wst[0].x;
        n1 = wst[1].x;
        prevBC0 = cumBC0; prevBC1 = cumBC1;
        cumA0 += n0 + 1;
        cumA1 += n1 > 0 ? n1 + 1 : 0;
      } else {
        const int4 ie = wst[0], io = wst[1];
        n0 = (ie.y + FLOW_SPI - 1) / FLOW_SPI + (ie.z + FLOW_SPI - 1) / FLOW_SPI;
        n1 = (io.y + FLOW_SPI - 1) / FLOW_SPI + (io.z + FLOW_SPI - 1) / FLOW_SPI;
        cumBC0 += n0;
        cumBC1 += n1;
      }
    }
    if (done) break;
    if (FLOW_PROF) { tQ += clock64() - cq0; ++nIt; }
    const int r = q - base, m = s >> 1;
    if (!(s & 1)) {
      ok = (r < n0) ? do_A(2 * m, r, n0, cumA0, prevBC0) : do_A(2 * m + 1, r - n0, n1, cumA1, prevBC1);
    } else {
      const bool even = r < n0;
      const int4 ii = wst[even ? 0 : 1];
      const int kk = even ? 2 * m : 2 * m + 1, rr = even ? r : r - n0;
      const int tA = even ? cumA0 : cumA1, tBC = even ? cumBC0 : cumBC1;
      if (ii.w & 1) ok = do_BC(std::true_type{}, kk, rr, ii, tA, tBC);
      else ok = do_BC(std::false_type{}, kk, rr, ii, tA, tBC);
    }
  }

  if (FLOW_PROF && prof && lane == 0) {
    long long* o = prof + 16 * ((int64_t)blockIdx.x * FLOW_WARPS + wid);
    o[0] = tA; o[1] = tC; o[2] = tB; o[3] = tW; o[4] = tQ; o[5] = nIt; o[6] = clock64() - cstart; o[7] = K;
    o[8] = tA1; o[9] = tA2; o[10] = tA3;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const double* ts = reinterpret_cast<const double*>(fsm + FlowSmem::TSUM);
    double s2 = 0.0, m2 = 0.0;
    for (int i = 0; i < 64; ++i) { s2 += ts[2 * i]; m2 = fmax(m2, ts[2 * i + 1]); }
    part[blockIdx.x * 8] = s2;
    part[blockIdx.x * 8 + 1] = m2;
    if (ctr[5]) atomicExch(gerr, 1);
  }
}

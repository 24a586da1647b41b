// Internal declarations of libb200ba (not part of the C ABI; see include/b200ba.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200ba.h"
#include "ba_math.cuh"

namespace b200ba {

constexpr int LIN_WARPS = 8;       // warps per CTA in the point-major kernels
constexpr int CCHUNK = 256;        // observations per camera-side chunk (one warp)
constexpr int SCHUNK = 512;
constexpr int SCHUR_TILE_MAX = 1344;  // observations per Schur tile (unfused path): 1344 * 144 B = 193.5 KB of shared memory
constexpr int FUSED_REC_BYTES = 176;  // per-observation record of k_lin_fused (22 doubles)
constexpr int FUSED_TILE_2CTA = 640;  // 640 * 176 B = 110 KB: two CTAs of 256 threads per SM
constexpr int FUSED_TILE_1CTA = 1312; // 1312 * 176 B = 225.5 KB: one CTA of 512 threads per SM
// dataflow kernel (k_lin_flow): one persistent CTA of 512 threads per SM, two record buffers, bulk-copy staging
#ifndef B200BA_FLOW_TILE
#define B200BA_FLOW_TILE 544
#endif
#ifndef B200BA_FLOW_CAM_SPAN
#define B200BA_FLOW_CAM_SPAN 32
#endif
constexpr int FLOW_TILE = B200BA_FLOW_TILE;        // observations per tile: 2 * 544 * 176 B = 187 KB of records (measured: 448 -> 512 -> 544
                                                   // observations = 2.67 -> 2.43 -> 2.35 ms at config D; 544 is what fits 227 KB)
constexpr int FLOW_CAM_SPAN = B200BA_FLOW_CAM_SPAN;     // staged camera window per tile (21 x 32 doubles; wider tiles read the table from global memory)
constexpr int FLOW_META_CAP = 8192;   // staged tile metadata (slot descriptors + entries), bytes
constexpr int FLOW_IDX_BYTES = (FLOW_TILE + 8) * 4;  // staged cam[] / pt[] window incl. alignment slack
constexpr int NSCAL = 32;          // device scalar slots
constexpr int DENSE_MAX_CAM = 1024; // dense Cholesky limit (n = 6144: 302 MB of dense S)
constexpr int DENSE_RESIDENT_CAM = 512;  // dense S export limit; the buffer is always allocated up to here
constexpr int MAX_DEVICES = 64;    // per-device caches of function attributes / occupancy

// scalar slots (device array `scal`); "sum" slots are all-reduced with ncclSum
enum Scal {
  S_COST2 = 0,     // sum r^2 at x (linearisation point)
  S_NEWCOST2 = 1,  // sum r^2 at x + delta
  S_GP_DP = 2,     // sum_p g_p . delta_p
  S_DP_D_DP = 3,   // sum_p delta_p^T diag(V_p) delta_p
  S_DP2 = 4,       // ||delta_p||^2
  S_XP2 = 5,       // ||x_p||^2
  S_NSUM = 8,      // number of sum slots sent through the all-reduce
  S_GPMAX = 8,     // max |g_p|   (ncclMax)
  S_GC_DC = 10,    // camera-side dots (computed from all-reduced data, identical on all ranks)
  S_DC_D_DC = 11,
  S_DC2 = 12,
  S_XC2 = 13,
  S_GCMAX = 14,
  S_DC_RES = 15,   // delta_c . (b - S delta_c)   (PCG residual term of the predicted reduction)
  S_PCG_ITERS = 16,
  S_SOLVE_FAIL = 17,
  S_RES_COST2 = 18, // sum r^2 of b200ba_residuals
  S_REPROJ_SUM = 20, // sum of per-observation reprojection error norms (b200ba_reproj_errors)
  S_TL_OK = 19,     // 1 = the coarse inverse of the two-level preconditioner is valid
  S_P2P_ERR = 30,   // (int) set by k_p2p_wait when a peer's contribution did not arrive in time
  S_FLOW_ERR = 31   // (int) set by k_lin_flow when a dependency wait timed out
};

// One shard of a bundle-adjustment problem, resident in HBM.
struct Problem {
  int64_t C = 0, P = 0, N = 0;
  int64_t Cs = 0;           // row stride of the field-major camera prep table (C rounded up to even: 16-byte rows)
  Intrinsics K;
  Loss loss{0, 1.0, 1.0};   // robust loss of the solve (b200ba_set_loss); linear by default
  // observations, point-major (sorted by point, stable)
  double* u = nullptr;      // [N]
  double* v = nullptr;      // [N]
  int* cam = nullptr;       // [N]
  int* pt = nullptr;        // [N] local point id
  int* perm = nullptr;      // [N] sorted position -> caller's observation index; null = identity
  int* pt_off = nullptr;    // [P+1]
  int* wt_off = nullptr;    // [n_wt+1] warp tiles: whole points, <= 32 observations (or one big point)
  int n_wt = 0;
  // camera-sorted view
  int* cperm = nullptr;     // [N] camera-sorted position -> sorted observation index
  int* cam_off = nullptr;   // [C+1]
  int* cch_cam = nullptr;   // [n_cch] camera of each chunk
  int* cch_a = nullptr;     // [n_cch+1]-style begin / end in cperm
  int* cch_b = nullptr;
  int* cam_choff = nullptr; // [C+1] chunk range of each camera
  int n_cch = 0;
  // Schur pair entries sorted by (global) upper block
  int64_t n_ent = 0;
  long long* ent = nullptr; // [n_ent] (i << 32) | j, sorted observation indices of one point
  int n_blk = 0;            // upper blocks incl. every diagonal block (global structure, identical on all ranks)
  int* blk_r = nullptr;     // [n_blk]
  int* blk_c = nullptr;
  int* sch_blk = nullptr;   // [n_sch]
  int* sch_a = nullptr;
  int* sch_b = nullptr;
  int* blk_choff = nullptr; // [n_blk+1]
  int n_sch = 0;
  int* diag_blk = nullptr;  // [C] index of block (c,c)
  // tile mode of the Schur kernel (default): observations cut into tiles of whole points that fit shared memory;
  // a "slot" = all pair entries of one (tile, block); slots of a tile are ordered by decreasing entry count
  bool schur_tiles = false;
  int tile_obs = 0;           // observations per tile (<= SCHUR_TILE_MAX)
  int n_st = 0;
  int* st_off = nullptr;      // [n_st+1] observation offsets of the tiles
  unsigned* tent = nullptr;   // [n_ent] (i_local << 16) | j_local, grouped by (tile, block)
  int n_slot = 0;
  int* slot_a = nullptr;      // [n_slot] first entry
  int* slot_n = nullptr;      // [n_slot] entry count
  int* tile_slot_off = nullptr;  // [n_st+1]
  int* tile_ent_off = nullptr;   // [n_st+1] entry range of each tile
  int* blk_slot_off = nullptr;   // [n_blk+1] reduce lists: slots of each block
  int* blk_slots = nullptr;      // [n_slot]
  // fused kernel (k_lin_fused): pair entries exclude i == j; the tile's observations grouped by camera form "self" slots
  bool fused = false;
  int* st_wt_off = nullptr;      // [n_st+1] warp-tile range of each CTA tile
  unsigned short* dent = nullptr;  // [N] tile-local observation index, grouped by (tile, camera)
  int n_dslot = 0;
  int* dslot_a = nullptr;        // [n_dslot] first entry in dent
  int* dslot_n = nullptr;        // [n_dslot]
  int* tile_dslot_off = nullptr; // [n_st+1]
  int* cam_dslot_off = nullptr;  // [C+1] reduce lists: self slots of each camera
  int* cam_dslots = nullptr;     // [n_dslot]
  // dataflow kernel: per-tile info (4 x int4) and one contiguous metadata blob per tile (16-byte units)
  bool flow = false;
  // sparse exchange of the reduced system (nranks > 1): every rank sends only the blocks of S it has contributions to
  // (all-gather), then every rank sums the contributions of each block in rank order (identical bits everywhere)
  bool xchg = false;
  int xl_n = 0, xl_max = 0;          // local blocks of this rank, max over ranks
  int* xl_blk = nullptr;             // [xl_n] union index of every local block (ascending)
  int* xc_off = nullptr;             // [n_blk+1] contributions of every union block ...
  long long* xc_src = nullptr;       // ... as offsets into the gathered buffer, ascending rank
  long long x_stride = 0;            // doubles per rank in the gathered buffer: [xl_max*36 | b | g_c | diag U | sums | gpmax]
  int4* tile_info = nullptr;     // [n_st*4] {nA, n_pslot, n_dslot, flags} {blob_off16, blob_units16, cmin, cspan}
                                 //           {obs0, n_obs, wt0, n_pent} {pslot_base, dslot_base, pt0, n_pts}
  uint4* blob = nullptr;         // header(32 B) | pair slots (a, n) | self slots (a, n) | pair entries u32 | self entries u16
  unsigned char* cam_fixed = nullptr;  // [C] 1 = held fixed (b200ba_set_fixed_cameras); null = none
  int n_fixed = 0;
  // two-level preconditioner (k_twolevel.cuh): aggregates = connected groups of cameras, 7 coarse unknowns each
  bool tl = false;
  int n_agg = 0, nc = 0, ncp = 0;  // nc = 7 n_agg, ncp = nc rounded up to a multiple of 64
  int* cam_agg = nullptr;          // [C]
  int* agg_off = nullptr;          // [n_agg+1] ranges into agg_cams
  int* agg_cams = nullptr;         // [C] cameras grouped by aggregate (ascending inside one)
  int* sp_of = nullptr;            // [nnzb] block of S P that block q of the BSR view of S contributes to
  int* sp_rowptr = nullptr;        // [C+1] block-CSR structure of S P (6 x 7 blocks)
  int* sp_col = nullptr;           // [nnz_sp] aggregate of each block
  int nnz_sp = 0;
  // full (both triangles) block-CSR view for the PCG mat-vec
  int* bsr_rowptr = nullptr;  // [C+1]
  int* bsr_col = nullptr;     // [nnzb]
  int* bsr_src = nullptr;     // [nnzb] upper block index; bit 31 set = use the transpose
  int nnzb = 0;
};

struct Work {
  double* x = nullptr;        // [6C+3P] current parameters
  double* xn = nullptr;       // [6C+3P] trial parameters
  double* camprep = nullptr;  // [C*24]
  double* camprep_n = nullptr;
  double* cstep = nullptr;    // [6][Cs] per-camera part of the back-substitution: J_r dw, dT
  double* Y = nullptr;        // [Y_count] 18 N (unfused pipeline / Jacobian export) or a 2 N staging area
  int64_t Y_count = 0;
  double* rr = nullptr;       // [N*4]  r0 r1 rt0 rt1
  double* pp = nullptr;       // [P*12] Linv(6) z(3) diagV(3)
  double* gp = nullptr;       // [P*3]
  double* cpart = nullptr;    // [n_cch*33]
  double* spart = nullptr;    // [n_sch*36]
  double* U = nullptr;        // [C*21]
  double* xsend = nullptr;    // [x_stride] sparse exchange: this rank's blocks + camera vectors + scalars
  double* xrecv = nullptr;    // [nranks * x_stride] gathered contributions of this exchange (NCCL buffer or peer-memory slots)
  double* xrecv_nccl = nullptr;
  double* dpart = nullptr;    // [n_dslot*40] self-slot partials of the fused kernel
  double* camsum = nullptr;   // [C*40] per-camera sums of the self-slot partials
  // ---- contiguous all-reduce buffer: Sup | b | gc | udiag | sums
  double* comm = nullptr;
  double* Sup = nullptr;      // [n_blk*36]
  double* b = nullptr;        // [6C]
  double* gc = nullptr;       // [6C]
  double* udiag = nullptr;    // [6C]
  double* sums = nullptr;     // [S_NSUM]
  int64_t comm_count = 0;
  // ----
  double* scal = nullptr;     // [NSCAL]
  double* part = nullptr;     // [n_part * 8] per-CTA partials of the point-major kernels
  int n_part = 0;
  double* dc = nullptr;       // [6C] camera step
  double* pcg = nullptr;      // [6C*5 + C*36] r u p w s Minv
  double* pcg_part = nullptr; // [grid*4]
  double* dense = nullptr;    // [(6C)^2] dense S (resident for n_cam <= DENSE_RESIDENT_CAM, else allocated by the first dense solve)
  double* chol = nullptr;     // multi-CTA Cholesky: padded matrix, rhs, solution, inverses of the diagonal factors
  int chol_np = 0;
  // two-level preconditioner
  double* tl_G = nullptr;     // [C*42]   G_c
  double* tl_SP = nullptr;    // [nnz_sp*42]
  double* tl_A = nullptr;     // [ncp*ncp] coarse matrix, inverted in place
  double* tl_cb = nullptr;    // [ncp*32] Gauss-Jordan column panel
  double* tl_rb = nullptr;    // [32*ncp] Gauss-Jordan row panel
  double* tl_dsc = nullptr;   // [ncp]
  double* tl_cen = nullptr;   // [n_agg*3]
  double* tl_rc = nullptr;    // [ncp]
  double* tl_zc = nullptr;    // [ncp]
  double* tl_uf = nullptr;    // [6C]
  double* tl_prof = nullptr;  // [8] phase nanoseconds of k_pcg2 (debug_fetch "tl_prof")
  double* h_scal = nullptr;   // pinned host mirror [NSCAL]
};

// Exchange of the reduced camera system through peer memory (NVLink / NVSwitch), one process per GPU: every rank owns a
// cudaMalloc'ed buffer [flags (64 x u64) | parity 0: nranks x x_stride doubles | parity 1: ...], opened on every other
// rank through CUDA IPC.  k_p2p_push stores this rank's contribution straight into slot [parity][rank] of every peer
// and then publishes the epoch in the peers' flag words; k_p2p_wait spins (bounded) on the local flags.
constexpr int P2P_MAX_RANKS = 8;
// destinations of a direct push from inside a producing kernel: base of this rank's receive slot in every rank's buffer
// (own rank included), ordered so that rank r starts with rank r + 1 (at any moment the ranks write to different peers)
struct PeerDst { double* p[P2P_MAX_RANKS]; int n = 0; };
struct P2P {
  bool on = false;
  void* base = nullptr;              // local buffer
  size_t bytes = 0;
  void* peer[P2P_MAX_RANKS] = {};    // base of every rank's buffer as mapped here (own entry = base)
  unsigned int* done = nullptr;      // CTA counter of k_p2p_push
  unsigned long long epoch = 0;      // exchanges done since the buffer was (re)allocated; identical on every rank
  bool direct = false;               // peers live in this process (b200ba_create_multi): raw pointers + peer access, no IPC
};

// Host-built block-CSR view of S and the camera aggregates of the two-level preconditioner (set_problem step 7).
struct HostBsr {
  std::vector<int> br, bc, diag, sb, sa, se, choff, rowptr, col, src;
  std::vector<int> cam_agg, agg_off, agg_cams, sp_rowptr, sp_col, sp_of;
  int n_agg = 0;
  bool tl = false;
};

struct Handle {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;   // set_problem: upload of the pixel array underneath the index pipeline
  cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr;
  cudaMemPool_t pool = nullptr;   // stream-ordered pool that keeps freed blocks: re-running set_problem costs no cudaMalloc
  std::string err;
  bool has_problem = false;
  Problem pb;
  Work w;
  std::vector<void*> allocs;      // problem-lifetime device allocations
  int64_t bytes_device = 0;
  // NCCL
  void* comm = nullptr;
  int rank = 0, nranks = 1;
  bool in_process = false;         // all ranks are handles of one process (b200ba_create_multi)
  // page-locked staging for the small host-built arrays of set_problem: copies from it are truly asynchronous, so the
  // per-array stream syncs of pageable uploads disappear (16+ per set_problem); reset at the end of setup_problem
  char* stage = nullptr;
  size_t stage_bytes = 0, stage_off = 0;
  HostBsr hb;                       // set_problem scratch kept between calls (capacity reuse)
  char* xfer = nullptr;             // page-locked slots of the threaded pageable <-> device copies (big_copy), lazily allocated
  size_t xfer_bytes = 0;
  bool direct_push = false;        // this linearisation's slot reduction stored its blocks straight into the peers' slots
  int64_t n_obs_global = 0;
  P2P p2p;
  // stats
  b200ba_stats st{};
  bool profile = false;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // PhaseTimer
  cudaEvent_t ev2 = nullptr, ev3 = nullptr;   // KernelTimer (may run inside a PhaseTimer bracket)
  // two-level preconditioner state of the running solve (identical on every rank: driven by PCG iteration counts)
  int tl_mode = 1;            // 0 block-Jacobi only, 1 automatic, 2 two-level from the first solve, rebuilt every solve
  bool tl_built = false;      // G / A_c^-1 hold a coarse operator of this solve
  int tl_age = 0;             // camera solves since the last rebuild
  int tl_ref_iters = 0;       // PCG iterations of the first solve after the last rebuild
  int tl_last_iters = 0;      // PCG iterations of the previous camera solve
};

// ---- helpers (ba_engine.cu)
int fail(Handle* h, int code, const std::string& msg);
int ensure_Y(Handle* h, int64_t count);  // grow w.Y (staging / Jacobian export)
int cuda_fail(Handle* h, cudaError_t e, const char* where);
// Large copy between host memory and the device, ordered after the work already queued on `st`.
// Page-locked (or managed) host memory: one cudaMemcpyAsync on `st`.  PAGEABLE host memory of 8 MB or more: the driver
// would stage it through one internal buffer on the calling thread (about 10 GB/s); here a few worker threads copy
// their chunks through page-locked slots of their own and issue the DMA on their own streams.  The workers run in the
// background: wait() (or the destructor) joins them and returns the first error; until then neither side may be touched.
struct BigCopy {
  std::vector<std::thread> th;
  std::vector<int> rcs;
  cudaEvent_t start = nullptr;
  int wait();
  ~BigCopy() { wait(); }
};
int big_copy_start(Handle* h, BigCopy& bc, void* dst, const void* src, size_t bytes, bool to_device, cudaStream_t st);
void* dalloc(Handle* h, size_t bytes);  // nullptr on failure (h->err set)
void* pool_alloc(Handle* h, size_t bytes);  // raw stream-ordered allocation from h->pool (nullptr on failure)
void pool_free(Handle* h, void* p);
void free_problem(Handle* h);
#define BA_CUDA(h, call)                                             \
  do {                                                               \
    cudaError_t e__ = (call);                                        \
    if (e__ != cudaSuccess) return cuda_fail((h), e__, #call);       \
  } while (0)

// ---- collectives (ba_engine.cu); no-ops when nranks == 1
int comm_allreduce_sum(Handle* h, double* buf, int64_t count);
int comm_allreduce_max(Handle* h, double* buf, int64_t count);
int comm_allgather_i64(Handle* h, const long long* send, long long* recv, int64_t count);
int comm_allgather_f64(Handle* h, const double* send, double* recv, int64_t count);
int p2p_setup(Handle* h);      // collective; after setup_problem (needs pb.x_stride); leaves h->p2p.on
void p2p_release(Handle* h);
int p2p_exchange(Handle* h);   // xsend -> every rank's receive slots; sets w.xrecv to the slots of this epoch
PeerDst p2p_next_dst(Handle* h);   // where the NEXT exchange's contribution of this rank goes (n = 0: no peer memory)

// ---- ba_setup.cu
int setup_problem(Handle* h, int64_t C, int64_t P, int64_t N, const int64_t* cam_idx, const int64_t* pt_idx,
                  const double* uv, const double* K);

// ---- ba_kernels.cu (all launches on h->stream)
void launch_cam_prep(Handle* h, const double* x_cams, const double* dc /*nullable*/, double* x_out /*nullable*/,
                     double* camprep, const double* camprep_old = nullptr, double* cstep = nullptr);
void launch_residual(Handle* h, const double* x, const double* camprep, double* r_out /*nullable, caller order*/,
                     int scal_slot);
void launch_pack_cameras(Handle* h, int64_t n, const double* H_dev, double* params_dev);
void launch_unpack_cameras(Handle* h, int64_t n, const double* params_dev, double* H_dev);
// probs_dev: n_problems x {const double* X, const double* uv, int n, pad} (device pointers), out_dev: n_problems x 9
void launch_pnp(Handle* h, int n_problems, const void* probs_dev, const double* K_dev, const double* guess_dev, int max_iter,
                double* out_dev);
void launch_triangulate(Handle* h, int64_t n, const double* M1, const double* M2, const double* pts1, const double* pts2,
                        const double* vis1, const double* vis2, double* X_out, unsigned char* mask_out);
void launch_triangulate_refine(Handle* h, int64_t n, const double* M1, const double* M2, const double* pts1, const double* pts2,
                               int max_iter, double* X);
void launch_reproj_err(Handle* h, const double* Rt, const double* X, double* err_out /*nullable, caller order*/);
void launch_jacobian(Handle* h, const double* x, const double* camprep, double* r_out, double* Jc, double* Jp);
void launch_xchg_sum(Handle* h);  // sparse exchange: gathered per-rank contributions -> Sup, b, gc, udiag, cost, gpmax
void launch_linearize(Handle* h, const double* x, double lambda);  // points + cameras + schur -> Sup, b, gc, ...
void launch_fix_diag(Handle* h);
void launch_fix_cameras(Handle* h);  // no-op unless b200ba_set_fixed_cameras marked some
void launch_backsub_trial(Handle* h, double lambda);
void launch_cam_dots(Handle* h);
void launch_densify(Handle* h, double* dense /*6C x 6C*/);
double measure_fp64_peak(Handle* h);
// k_noalloc.cuh: residual + cost with all scratch in a static device array (no allocation on the call path)
int noalloc_setup(Handle* h);   // once per device, device current: cudaGetSymbolAddress of the scratch array
int residuals_noalloc(Handle* h, int64_t C, int64_t P, int64_t N, const int* cam_dev, const int* pt_dev, const double* uv_dev,
                      const double* K, const double* x_dev, double* r_out_dev, double* cost_out_dev);

// ---- ba_solve.cu
int launch_pcg(Handle* h, double rtol, int max_iter);
int launch_tl_build(Handle* h, bool rebuild_coarse);   // SP = S P (always); G, A_c, A_c^-1 when rebuild_coarse
int launch_pcg2(Handle* h, double rtol, int max_iter);
int launch_cholesky(Handle* h);

}  // namespace b200ba

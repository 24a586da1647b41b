// Problem setup: replaces the host-side packing of sfm_lite/sfm.py:248-261 and the Jacobian
// sparsity mask of sfm_lite/bundle_adjustment.py:6-17 by device-resident index structures:
// point-major SoA observations, per-point CSR, warp tiles, camera-sorted permutation + chunks,
// and the Schur pair-entry list sorted by reduced-camera-system block (atomics-free accumulation).
// One-off per problem; CUB sort/scan/run-length are used here (not on the per-iteration path).
#include <algorithm>
#include <cub/cub.cuh>

#include "ba_internal.h"

namespace b200ba {

namespace {

__global__ void k_check_convert(int64_t N, int64_t C, int64_t P, const int64_t* __restrict__ cam64,
                                const int64_t* __restrict__ pt64, int* __restrict__ cam32, int* __restrict__ pt32,
                                int* __restrict__ iota, int* __restrict__ flags) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t c = cam64[i], p = pt64[i];
  int f = 0;
  if (c < 0 || c >= C || p < 0 || p >= P) { f |= 1; c = 0; p = 0; }
  if (i > 0 && pt64[i - 1] > pt64[i]) f |= 2;
  if (f) atomicOr(flags, f);
  cam32[i] = (int)c;
  pt32[i] = (int)p;
  iota[i] = (int)i;
}

__global__ void k_gather_obs(int64_t N, const int* __restrict__ perm, const int* __restrict__ cam_in,
                             const double* __restrict__ uv, int* __restrict__ cam, double* __restrict__ u,
                             double* __restrict__ v) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t s = perm ? perm[i] : i;
  cam[i] = cam_in[s];
  u[i] = uv[2 * s];
  v[i] = uv[2 * s + 1];
}

// off[q] = first index i with key[i] >= q, for q in [0, M]; keys sorted ascending.
__global__ void k_offsets(int64_t N, int M, const int* __restrict__ key, int* __restrict__ off) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int cur = key[i];
  int prev = i > 0 ? key[i - 1] : -1;
  for (int q = prev + 1; q <= cur; ++q) off[q] = (int)i;
  if (i == N - 1)
    for (int q = cur + 1; q <= M; ++q) off[q] = (int)N;
}

__global__ void k_iota(int64_t N, int* __restrict__ a) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < N) a[i] = (int)i;
}

// ordered pairs (i,j) of one point's observations that land in the upper triangle (cam_i <= cam_j)
__global__ void k_pair_count(int P, const int* __restrict__ pt_off, const int* __restrict__ cam,
                             long long* __restrict__ cnt) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int a = pt_off[p], b = pt_off[p + 1];
  long long n = 0;
  for (int i = a; i < b; ++i) {
    int ci = cam[i];
    for (int j = a; j < b; ++j) n += (ci <= cam[j]);
  }
  cnt[p] = n;
}

__global__ void k_pair_fill(int P, long long C, const int* __restrict__ pt_off, const int* __restrict__ cam,
                            const long long* __restrict__ ent_off, long long* __restrict__ keys,
                            long long* __restrict__ vals) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int a = pt_off[p], b = pt_off[p + 1];
  long long e = ent_off[p];
  for (int i = a; i < b; ++i) {
    int ci = cam[i];
    for (int j = a; j < b; ++j) {
      int cj = cam[j];
      if (ci <= cj) {
        keys[e] = (long long)ci * C + cj;
        vals[e] = ((long long)i << 32) | (unsigned int)j;
        ++e;
      }
    }
  }
}

__global__ void k_lower_bound(int n_blk, const long long* __restrict__ gkeys, long long n_ent,
                              const long long* __restrict__ keys_sorted, long long* __restrict__ blk_off) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g > n_blk) return;
  if (g == n_blk) { blk_off[g] = n_ent; return; }
  long long key = gkeys[g];
  long long lo = 0, hi = n_ent;
  while (lo < hi) {
    long long mid = (lo + hi) >> 1;
    if (keys_sorted[mid] < key) lo = mid + 1; else hi = mid;
  }
  blk_off[g] = lo;
}

inline unsigned nblk(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

template <class T>
int upload(Handle* h, T** dst, const std::vector<T>& src) {
  size_t n = std::max<size_t>(src.size(), 1);
  *dst = (T*)dalloc(h, n * sizeof(T));
  if (!*dst) return B200BA_OOM;
  if (!src.empty())
    BA_CUDA(h, cudaMemcpyAsync(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice, h->stream));
  BA_CUDA(h, cudaStreamSynchronize(h->stream));  // src may be a temporary
  return 0;
}

struct Tmp {  // scratch freed at the end of setup
  std::vector<void*> p;
  ~Tmp() { for (void* q : p) cudaFree(q); }
  void* get(size_t bytes) {
    void* q = nullptr;
    if (cudaMalloc(&q, std::max<size_t>(bytes, 16)) != cudaSuccess) return nullptr;
    p.push_back(q);
    return q;
  }
};

}  // namespace

#define TMP_ALLOC(var, type, count)                                          \
  type* var = (type*)tmp.get((size_t)(count) * sizeof(type));                \
  if (!var) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)")
#define DEV_ALLOC(var, type, count)                                          \
  var = (type*)dalloc(h, (size_t)std::max<int64_t>((count), 1) * sizeof(type)); \
  if (!var) return B200BA_OOM

int setup_problem(Handle* h, int64_t C, int64_t P, int64_t N, const int64_t* cam_idx, const int64_t* pt_idx,
                  const double* uv, const double* K) {
  Problem& pb = h->pb;
  cudaStream_t st = h->stream;
  pb = Problem();
  pb.C = C; pb.P = P; pb.N = N;
  for (int i = 0; i < 9; ++i) pb.K.k[i] = K[i];
  Tmp tmp;
  const int T = 256;

  // ---- 1. raw inputs to the device, range / order check, int32 indices
  TMP_ALLOC(d_cam64, int64_t, N);
  TMP_ALLOC(d_pt64, int64_t, N);
  TMP_ALLOC(d_uv, double, 2 * N);
  TMP_ALLOC(d_cam_in, int, N);
  TMP_ALLOC(d_pt_in, int, N);
  TMP_ALLOC(d_iota, int, N);
  TMP_ALLOC(d_flags, int, 1);
  BA_CUDA(h, cudaMemsetAsync(d_flags, 0, sizeof(int), st));
  if (N > 0) {
    BA_CUDA(h, cudaMemcpyAsync(d_cam64, cam_idx, N * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    BA_CUDA(h, cudaMemcpyAsync(d_pt64, pt_idx, N * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    BA_CUDA(h, cudaMemcpyAsync(d_uv, uv, 2 * N * sizeof(double), cudaMemcpyHostToDevice, st));
    k_check_convert<<<nblk(N, T), T, 0, st>>>(N, C, P, d_cam64, d_pt64, d_cam_in, d_pt_in, d_iota, d_flags);
  }
  int flags = 0;
  BA_CUDA(h, cudaMemcpyAsync(&flags, d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
  BA_CUDA(h, cudaStreamSynchronize(st));
  if (flags & 1) return fail(h, B200BA_INDEX_RANGE, "camera or point index out of range");

  // ---- 2. point-major order (stable); identity when the input is already sorted (graph.py:186-191)
  DEV_ALLOC(pb.u, double, N);
  DEV_ALLOC(pb.v, double, N);
  DEV_ALLOC(pb.cam, int, N);
  DEV_ALLOC(pb.pt, int, N);
  DEV_ALLOC(pb.pt_off, int, P + 1);
  BA_CUDA(h, cudaMemsetAsync(pb.pt_off, 0, (P + 1) * sizeof(int), st));
  if (flags & 2) {
    DEV_ALLOC(pb.perm, int, N);
    size_t bytes = 0;
    int end_bit = 1;
    while (((int64_t)1 << end_bit) < P && end_bit < 31) ++end_bit;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_pt_in, pb.pt, d_iota, pb.perm, (int)N, 0, end_bit, st);
    TMP_ALLOC(d_sort, char, bytes);
    BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort, bytes, d_pt_in, pb.pt, d_iota, pb.perm, (int)N, 0, end_bit, st));
  } else if (N > 0) {
    BA_CUDA(h, cudaMemcpyAsync(pb.pt, d_pt_in, N * sizeof(int), cudaMemcpyDeviceToDevice, st));
  }
  if (N > 0) {
    k_gather_obs<<<nblk(N, T), T, 0, st>>>(N, pb.perm, d_cam_in, d_uv, pb.cam, pb.u, pb.v);
    k_offsets<<<nblk(N, T), T, 0, st>>>(N, (int)P, pb.pt, pb.pt_off);
  }

  // ---- 3. warp tiles (host greedy packing of whole points into <= 32 observations)
  std::vector<int> h_pt_off(P + 1, 0);
  BA_CUDA(h, cudaMemcpyAsync(h_pt_off.data(), pb.pt_off, (P + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));
  BA_CUDA(h, cudaStreamSynchronize(st));
  {
    std::vector<int> wt;
    wt.reserve(N / 24 + 16);
    wt.push_back(0);
    int start = 0;
    for (int64_t p = 0; p < P; ++p) {
      int e = h_pt_off[p + 1];
      if (e - start > 32) {
        if (h_pt_off[p] > start) wt.push_back(h_pt_off[p]);  // close the tile before this point
        start = wt.back();
        if (e - start > 32) { wt.push_back(e); start = e; }  // a point with > 32 observations: its own tile
      }
    }
    if ((int)N > start) wt.push_back((int)N);
    pb.n_wt = (int)wt.size() - 1;
    if (int rc = upload(h, &pb.wt_off, wt)) return rc;
  }

  // ---- 4. camera-sorted permutation and chunks
  DEV_ALLOC(pb.cperm, int, N);
  DEV_ALLOC(pb.cam_off, int, C + 1);
  BA_CUDA(h, cudaMemsetAsync(pb.cam_off, 0, (C + 1) * sizeof(int), st));
  {
    TMP_ALLOC(d_cam_sorted, int, N);
    if (N > 0) {
      k_iota<<<nblk(N, T), T, 0, st>>>(N, d_iota);
      size_t bytes = 0;
      int end_bit = 1;
      while (((int64_t)1 << end_bit) < C && end_bit < 31) ++end_bit;
      cub::DeviceRadixSort::SortPairs(nullptr, bytes, pb.cam, d_cam_sorted, d_iota, pb.cperm, (int)N, 0, end_bit, st);
      TMP_ALLOC(d_sort, char, bytes);
      BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort, bytes, pb.cam, d_cam_sorted, d_iota, pb.cperm, (int)N, 0,
                                                 end_bit, st));
      k_offsets<<<nblk(N, T), T, 0, st>>>(N, (int)C, d_cam_sorted, pb.cam_off);
    }
    std::vector<int> h_cam_off(C + 1, 0);
    BA_CUDA(h, cudaMemcpyAsync(h_cam_off.data(), pb.cam_off, (C + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
    std::vector<int> cc, ca, cb, coff(C + 1, 0);
    for (int64_t c = 0; c < C; ++c) {
      coff[c] = (int)cc.size();
      for (int a = h_cam_off[c]; a < h_cam_off[c + 1]; a += CCHUNK) {
        cc.push_back((int)c); ca.push_back(a); cb.push_back(std::min(a + CCHUNK, h_cam_off[c + 1]));
      }
    }
    coff[C] = (int)cc.size();
    pb.n_cch = (int)cc.size();
    if (int rc = upload(h, &pb.cch_cam, cc)) return rc;
    if (int rc = upload(h, &pb.cch_a, ca)) return rc;
    if (int rc = upload(h, &pb.cch_b, cb)) return rc;
    if (int rc = upload(h, &pb.cam_choff, coff)) return rc;
  }

  // ---- 5. Schur pair entries, sorted by block key = cam_i * C + cam_j (cam_i <= cam_j)
  TMP_ALLOC(d_cnt, long long, P + 1);
  TMP_ALLOC(d_ent_off, long long, P + 1);
  BA_CUDA(h, cudaMemsetAsync(d_cnt, 0, (P + 1) * sizeof(long long), st));
  if (P > 0) k_pair_count<<<nblk(P, T), T, 0, st>>>((int)P, pb.pt_off, pb.cam, d_cnt);
  {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, d_cnt, d_ent_off, (int)(P + 1), st);
    TMP_ALLOC(d_scan, char, bytes);
    BA_CUDA(h, cub::DeviceScan::ExclusiveSum(d_scan, bytes, d_cnt, d_ent_off, (int)(P + 1), st));
  }
  long long n_ent = 0;
  BA_CUDA(h, cudaMemcpyAsync(&n_ent, d_ent_off + P, sizeof(long long), cudaMemcpyDeviceToHost, st));
  BA_CUDA(h, cudaStreamSynchronize(st));
  if (n_ent >= ((long long)1 << 31)) return fail(h, B200BA_BAD_ARG, "too many Schur pair entries for one shard (>= 2^31)");
  pb.n_ent = n_ent;
  DEV_ALLOC(pb.ent, long long, n_ent);
  TMP_ALLOC(d_keys_sorted, long long, n_ent);
  std::vector<long long> keys_host;  // unique local keys
  {
    TMP_ALLOC(d_keys, long long, n_ent);
    TMP_ALLOC(d_vals, long long, n_ent);
    if (P > 0) k_pair_fill<<<nblk(P, T), T, 0, st>>>((int)P, (long long)C, pb.pt_off, pb.cam, d_ent_off, d_keys, d_vals);
    if (n_ent > 0) {
      int end_bit = 1;
      while (end_bit < 63 && ((long long)1 << end_bit) < (long long)C * C) ++end_bit;
      size_t bytes = 0;
      cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_keys, d_keys_sorted, d_vals, pb.ent, (int)n_ent, 0, end_bit, st);
      TMP_ALLOC(d_sort, char, bytes);
      BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort, bytes, d_keys, d_keys_sorted, d_vals, pb.ent, (int)n_ent, 0,
                                                 end_bit, st));
      TMP_ALLOC(d_uniq, long long, n_ent);
      TMP_ALLOC(d_runs, int, n_ent);
      TMP_ALLOC(d_nruns, int, 1);
      bytes = 0;
      cub::DeviceRunLengthEncode::Encode(nullptr, bytes, d_keys_sorted, d_uniq, d_runs, d_nruns, (int)n_ent, st);
      TMP_ALLOC(d_rle, char, bytes);
      BA_CUDA(h, cub::DeviceRunLengthEncode::Encode(d_rle, bytes, d_keys_sorted, d_uniq, d_runs, d_nruns, (int)n_ent, st));
      int nruns = 0;
      BA_CUDA(h, cudaMemcpyAsync(&nruns, d_nruns, sizeof(int), cudaMemcpyDeviceToHost, st));
      BA_CUDA(h, cudaStreamSynchronize(st));
      keys_host.resize(nruns);
      BA_CUDA(h, cudaMemcpy(keys_host.data(), d_uniq, (size_t)nruns * sizeof(long long), cudaMemcpyDeviceToHost));
    }
  }

  // ---- 6. global block structure: union over ranks of the local keys + every diagonal block
  if (h->nranks > 1) {
    long long my_n = (long long)keys_host.size();
    TMP_ALLOC(d_n, long long, h->nranks + 1);
    BA_CUDA(h, cudaMemcpyAsync(d_n + h->nranks, &my_n, sizeof(long long), cudaMemcpyHostToDevice, st));
    if (int rc = comm_allgather_i64(h, d_n + h->nranks, d_n, 1)) return rc;
    std::vector<long long> counts(h->nranks);
    BA_CUDA(h, cudaMemcpyAsync(counts.data(), d_n, h->nranks * sizeof(long long), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
    long long mx = 1;
    for (long long c : counts) mx = std::max(mx, c);
    TMP_ALLOC(d_send, long long, mx);
    TMP_ALLOC(d_recv, long long, mx * h->nranks);
    BA_CUDA(h, cudaMemsetAsync(d_send, 0, mx * sizeof(long long), st));
    if (my_n)
      BA_CUDA(h, cudaMemcpyAsync(d_send, keys_host.data(), my_n * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (int rc = comm_allgather_i64(h, d_send, d_recv, mx)) return rc;
    std::vector<long long> all((size_t)mx * h->nranks);
    BA_CUDA(h, cudaMemcpyAsync(all.data(), d_recv, all.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
    keys_host.clear();
    for (int r = 0; r < h->nranks; ++r)
      keys_host.insert(keys_host.end(), all.begin() + (size_t)r * mx, all.begin() + (size_t)r * mx + counts[r]);
  }
  for (int64_t c = 0; c < C; ++c) keys_host.push_back((long long)c * C + c);
  std::sort(keys_host.begin(), keys_host.end());
  keys_host.erase(std::unique(keys_host.begin(), keys_host.end()), keys_host.end());
  const int n_blk = (int)keys_host.size();
  pb.n_blk = n_blk;

  // entry range of every global block in the local sorted entry list
  std::vector<long long> h_blk_off(n_blk + 1, 0);
  {
    TMP_ALLOC(d_gkeys, long long, n_blk);
    TMP_ALLOC(d_blk_off, long long, n_blk + 1);
    BA_CUDA(h, cudaMemcpyAsync(d_gkeys, keys_host.data(), (size_t)n_blk * sizeof(long long), cudaMemcpyHostToDevice, st));
    k_lower_bound<<<nblk(n_blk + 1, T), T, 0, st>>>(n_blk, d_gkeys, n_ent, d_keys_sorted, d_blk_off);
    BA_CUDA(h, cudaMemcpyAsync(h_blk_off.data(), d_blk_off, (size_t)(n_blk + 1) * sizeof(long long),
                               cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
  }

  // ---- 7. host: block coordinates, Schur chunks, full BSR view
  {
    std::vector<int> br(n_blk), bc(n_blk), diag(C, -1), sb, sa, se, choff(n_blk + 1, 0);
    std::vector<int> rowcnt(C + 1, 0);
    for (int g = 0; g < n_blk; ++g) {
      br[g] = (int)(keys_host[g] / C);
      bc[g] = (int)(keys_host[g] % C);
      if (br[g] == bc[g]) { diag[br[g]] = g; rowcnt[br[g]]++; }
      else { rowcnt[br[g]]++; rowcnt[bc[g]]++; }
      choff[g] = (int)sb.size();
      for (long long a = h_blk_off[g]; a < h_blk_off[g + 1]; a += SCHUNK) {
        sb.push_back(g); sa.push_back((int)a); se.push_back((int)std::min<long long>(a + SCHUNK, h_blk_off[g + 1]));
      }
    }
    choff[n_blk] = (int)sb.size();
    pb.n_sch = (int)sb.size();
    std::vector<int> rowptr(C + 1, 0);
    for (int64_t c = 0; c < C; ++c) rowptr[c + 1] = rowptr[c] + rowcnt[c];
    pb.nnzb = rowptr[C];
    std::vector<int> col(pb.nnzb), src(pb.nnzb), fill(rowptr.begin(), rowptr.end() - 1);
    // keys are sorted by (r,c): pushing the direct block into row r and the transposed one into row c keeps
    // every row's columns ascending (transposed entries of row c come from r < c, direct ones from c' >= c)
    for (int pass = 0; pass < 2; ++pass)
      for (int g = 0; g < n_blk; ++g) {
        if (pass == 0 && br[g] != bc[g]) { int q = fill[bc[g]]++; col[q] = br[g]; src[q] = g | (int)0x80000000; }
        if (pass == 1) { int q = fill[br[g]]++; col[q] = bc[g]; src[q] = g; }
      }
    if (int rc = upload(h, &pb.blk_r, br)) return rc;
    if (int rc = upload(h, &pb.blk_c, bc)) return rc;
    if (int rc = upload(h, &pb.diag_blk, diag)) return rc;
    if (int rc = upload(h, &pb.sch_blk, sb)) return rc;
    if (int rc = upload(h, &pb.sch_a, sa)) return rc;
    if (int rc = upload(h, &pb.sch_b, se)) return rc;
    if (int rc = upload(h, &pb.blk_choff, choff)) return rc;
    if (int rc = upload(h, &pb.bsr_rowptr, rowptr)) return rc;
    if (int rc = upload(h, &pb.bsr_col, col)) return rc;
    if (int rc = upload(h, &pb.bsr_src, src)) return rc;
  }
  BA_CUDA(h, cudaStreamSynchronize(st));
  BA_CUDA(h, cudaGetLastError());
  h->st.schur_entries = pb.n_ent;
  h->st.schur_blocks = pb.n_blk;
  return 0;
}

}  // namespace b200ba

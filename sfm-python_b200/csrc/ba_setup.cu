// Problem setup: replaces the host-side packing of sfm_lite/sfm.py:248-261 and the Jacobian
// sparsity mask of sfm_lite/bundle_adjustment.py:6-17 by device-resident index structures:
// point-major SoA observations, per-point CSR, warp tiles, camera-sorted permutation + chunks,
// and the Schur pair-entry list sorted by reduced-camera-system block (atomics-free accumulation).
// One-off per problem; CUB sort/scan/run-length are used here (not on the per-iteration path).
#include <chrono>
#include <thread>
#include <algorithm>
#include <cub/cub.cuh>

#include "ba_internal.h"

namespace b200ba {

namespace {

constexpr int SLOT_CAP = 32;  // pair entries per Schur slot (tile mode)

// point indices: range check, order check, int32 copy, identity permutation (flags: 1 out of range, 2 unsorted)
__global__ void k_check_convert_pt(int64_t N, int64_t P, const int64_t* __restrict__ pt64, int* __restrict__ pt32,
                                   int* __restrict__ iota, int* __restrict__ flags) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t p = pt64[i];
  int f = 0;
  if (p < 0 || p >= P) { f |= 1; p = 0; }
  if (i > 0 && pt64[i - 1] > pt64[i]) f |= 2;
  if (f) atomicOr(flags, f);
  pt32[i] = (int)p;
  iota[i] = (int)i;
}

// camera indices: range check and int32 copy; runs on the second stream behind their upload (flags[1]: 1 out of range)
__global__ void k_check_convert_cam(int64_t N, int64_t C, const int64_t* __restrict__ cam64, int* __restrict__ cam32,
                                    int* __restrict__ flags) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t c = cam64[i];
  if (c < 0 || c >= C) { atomicOr(flags + 1, 1); c = 0; }
  cam32[i] = (int)c;
}

__global__ void k_gather_obs(int64_t N, const int* __restrict__ perm, const int* __restrict__ cam_in,
                             int* __restrict__ cam) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t s = perm ? perm[i] : i;
  cam[i] = cam_in[s];
}

// pixels into point-major order; runs at the END of set_problem: their upload travels on a second stream underneath
// the index pipeline (tiles, camera permutation, pair entries need indices only)
__global__ void k_gather_uv(int64_t N, const int* __restrict__ perm, const double* __restrict__ uv, double* __restrict__ u,
                            double* __restrict__ v) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int64_t s = perm ? perm[i] : i;
  const double2 p = reinterpret_cast<const double2*>(uv)[s];
  u[i] = p.x;
  v[i] = p.y;
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

__global__ void k_cam_seen(int64_t N, const int* __restrict__ cam, int* __restrict__ seen) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < N) seen[cam[i]] = 1;   // every writer stores the same value
}

__global__ void k_iota(int64_t N, int* __restrict__ a) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < N) a[i] = (int)i;
}

// ordered pairs (i,j) of one point's observations that land in the upper triangle (cam_i <= cam_j)
__global__ void k_pair_count(int P, const int* __restrict__ pt_off, const int* __restrict__ cam, bool skip_self,
                             long long* __restrict__ cnt) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int a = pt_off[p], b = pt_off[p + 1];
  long long n = 0;
  for (int i = a; i < b; ++i) {
    int ci = cam[i];
    for (int j = a; j < b; ++j) n += (ci <= cam[j]) && !(skip_self && i == j);
  }
  cnt[p] = n;
}

__global__ void k_pair_fill(int P, long long C, const int* __restrict__ pt_off, const int* __restrict__ cam,
                            bool skip_self, const long long* __restrict__ ent_off, const int* __restrict__ pt_tile,
                            long long* __restrict__ keys, long long* __restrict__ vals, const int* __restrict__ st_off,
                            unsigned* __restrict__ vals32) {
  int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int a = pt_off[p], b = pt_off[p + 1];
  long long e = ent_off[p];
  const long long tbase = pt_tile ? (long long)pt_tile[p] * C * C : 0;
  // tile mode: the value is the pair of tile-local 16-bit observation indices the kernels read (4 instead of 8 bytes
  // through the sort, no repacking pass afterwards)
  const int obase = vals32 ? st_off[pt_tile[p]] : 0;
  for (int i = a; i < b; ++i) {
    int ci = cam[i];
    for (int j = a; j < b; ++j) {
      int cj = cam[j];
      if (ci <= cj && !(skip_self && i == j)) {
        keys[e] = tbase + (long long)ci * C + cj;
        if (vals32) vals32[e] = ((unsigned)(i - obase) << 16) | (unsigned)(j - obase);
        else vals[e] = ((long long)i << 32) | (unsigned int)j;
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

__global__ void k_mod_keys(int n, const long long* __restrict__ keys, long long CC, long long* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = keys[i] % CC;
}

// per (tile, block) cell: global block id and the ordering key (tile asc, entry count desc)
__global__ void k_cell_meta(int n, const long long* __restrict__ cellkey, const int* __restrict__ cnt, long long CC,
                            const long long* __restrict__ gkeys, int n_blk, int* __restrict__ cell_blk,
                            int* __restrict__ cell_tile, int* __restrict__ nsub) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n) return;
  const long long key = cellkey[c];
  const long long tile = key / CC, bk = key - tile * CC;
  int lo = 0, hi = n_blk;
  if (!gkeys) lo = (int)bk;  // identity: the id is the group index
  else
    while (lo < hi) {
      int mid = (lo + hi) >> 1;
      if (gkeys[mid] < bk) lo = mid + 1; else hi = mid;
    }
  cell_blk[c] = lo;
  cell_tile[c] = (int)tile;
  nsub[c] = (cnt[c] + SLOT_CAP - 1) / SLOT_CAP;
}

// split every cell into sub-slots of at most SLOT_CAP entries (bounded, balanced work per lane group)
__global__ void k_sub_fill(int n_cell, const int* __restrict__ sub_off, const int* __restrict__ cell_off,
                           const int* __restrict__ cnt, const int* __restrict__ cell_blk,
                           const int* __restrict__ cell_tile, int* __restrict__ sub_a, int* __restrict__ sub_n,
                           int* __restrict__ sub_blk, long long* __restrict__ key2) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cell) return;
  const int q0 = sub_off[c], nq = sub_off[c + 1] - q0, a = cell_off[c], n = cnt[c];
  for (int k = 0; k < nq; ++k) {
    const int m = min(SLOT_CAP, n - SLOT_CAP * k);
    sub_a[q0 + k] = a + SLOT_CAP * k;
    sub_n[q0 + k] = m;
    sub_blk[q0 + k] = cell_blk[c];
    key2[q0 + k] = ((long long)cell_tile[c] << 32) | (long long)(0x7fffffff - m);
  }
}

__global__ void k_slot_fill(int n, const int* __restrict__ order, const int* __restrict__ sub_a,
                            const int* __restrict__ sub_n, const int* __restrict__ sub_blk, int* __restrict__ slot_a,
                            int* __restrict__ slot_n, int* __restrict__ slot_blk) {
  int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n) return;
  const int q = order[s];
  slot_a[s] = sub_a[q];
  slot_n[s] = sub_n[q];
  slot_blk[s] = sub_blk[q];
}

__global__ void k_tile_slot_off(int n_st, int n_slot, const long long* __restrict__ key2s, int* __restrict__ off) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > n_st) return;
  const long long key = (long long)t << 32;
  int lo = 0, hi = n_slot;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (key2s[mid] < key) lo = mid + 1; else hi = mid;
  }
  off[t] = lo;
}

__global__ void k_tile_ent_off(int n_st, int n_ent, const long long* __restrict__ keys_sorted, long long CC,
                               int* __restrict__ off) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t > n_st) return;
  const long long key = (long long)t * CC;
  int lo = 0, hi = n_ent;
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (keys_sorted[mid] < key) lo = mid + 1; else hi = mid;
  }
  off[t] = lo;
}

// tile of every point: the last CTA tile whose first observation is <= the point's first observation
__global__ void k_pt_tile(int P, const int* __restrict__ pt_off, const int* __restrict__ st_off, int n_st,
                          int* __restrict__ pt_tile) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int i = pt_off[p];
  int lo = 0, hi = n_st;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (st_off[mid] <= i) lo = mid; else hi = mid;
  }
  pt_tile[p] = lo;
}

__global__ void k_pack_entries(long long n, const long long* __restrict__ ent64, const int* __restrict__ st_off,
                               int n_st, unsigned* __restrict__ tent) {
  long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (e >= n) return;
  const long long ij = ent64[e];
  const int i = (int)(ij >> 32), j = (int)(ij & 0xffffffffll);
  int lo = 0, hi = n_st;  // last tile with st_off[t] <= i
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (st_off[mid] <= i) lo = mid; else hi = mid;
  }
  const int base = st_off[lo];
  tent[e] = ((unsigned)(i - base) << 16) | (unsigned)(j - base);
}

// fused kernel: key (tile, camera) of every observation; tile-local 16-bit index of the camera-grouped list
__global__ void k_self_keys(int64_t N, long long C, const int* __restrict__ cam, const int* __restrict__ st_off, int n_st,
                            long long* __restrict__ key) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= N) return;
  int lo = 0, hi = n_st;  // last tile with st_off[t] <= i
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (st_off[mid] <= i) lo = mid; else hi = mid;
  }
  key[i] = (long long)lo * C + cam[i];
}

__global__ void k_pack_self(int64_t N, const int* __restrict__ val, const int* __restrict__ st_off, int n_st,
                            unsigned short* __restrict__ dent) {
  int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e >= N) return;
  const int i = val[e];
  int lo = 0, hi = n_st;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (st_off[mid] <= i) lo = mid; else hi = mid;
  }
  dent[e] = (unsigned short)(i - st_off[lo]);
}

// ---- dataflow kernel metadata -------------------------------------------------------------------------------
// per tile: counts, camera window, blob size (16-byte units)
__global__ void k_tile_info(int n_st, long long C, const int* __restrict__ st_off, const int* __restrict__ st_wt_off,
                            const int* __restrict__ tile_slot_off, const int* __restrict__ tile_dslot_off,
                            const int* __restrict__ tile_ent_off, const int* __restrict__ pt, int n_cell,
                            const long long* __restrict__ dcell, int4* __restrict__ info, int* __restrict__ units) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_st) return;
  const int nobs = st_off[t + 1] - st_off[t];
  const int np = tile_slot_off[t + 1] - tile_slot_off[t], nd = tile_dslot_off[t + 1] - tile_dslot_off[t];
  const int npent = tile_ent_off[t + 1] - tile_ent_off[t];
  // first / last (tile, camera) cell of the tile in the sorted cell list
  int lo = 0, hi = n_cell;
  const long long k0 = (long long)t * C, k1 = (long long)(t + 1) * C;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (dcell[mid] < k0) lo = mid + 1; else hi = mid; }
  const int first = lo;
  hi = n_cell;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (dcell[mid] < k1) lo = mid + 1; else hi = mid; }
  const int last = lo - 1;
  int cmin = 0, cspan = 0;
  if (last >= first) {
    cmin = (int)(dcell[first] - k0) & ~1;
    const int cmax = (int)(dcell[last] - k0);
    cspan = (cmax - cmin + 2) & ~1;
  }
  const int bytes = 32 + 8 * (np + nd) + 4 * npent + 2 * nobs;
  const int u = (bytes + 15) / 16;
  int flags = 0;
  if (u * 16 <= FLOW_META_CAP) flags |= 1;
  if (cspan <= FLOW_CAM_SPAN) flags |= 2;
  info[4 * t + 0] = make_int4(st_wt_off[t + 1] - st_wt_off[t], np, nd, flags);
  info[4 * t + 1] = make_int4(0, u, cmin, cspan);
  info[4 * t + 2] = make_int4(st_off[t], nobs, st_wt_off[t], npent);
  const int p0 = nobs > 0 ? pt[st_off[t]] : 0, p1 = nobs > 0 ? pt[st_off[t + 1] - 1] : -1;
  info[4 * t + 3] = make_int4(tile_slot_off[t], tile_dslot_off[t], p0, p1 - p0 + 1);
  units[t] = u;
}

// Bank-aware entry order.  The slot kernels read per-observation records (22 doubles, stride 11 x 16 bytes) with 16-byte
// loads; the eight lanes of a quarter-warp (4 consecutive slots x 2 lanes, lane g of a slot takes entries g, g + 2, ...)
// collide in shared memory exactly when two of their records have the same index modulo 8.  The order of the entries
// INSIDE a slot is free (it only fixes the order of a sum), so it is chosen greedily here: round by round, every lane
// takes the remaining entry of its slot whose record classes are the least used in this round.  One thread per group of
// four slots; random order gives about 2.4 shared-memory wavefronts per ideal one, this order about 1.5.
template <class E>
__device__ __forceinline__ int entry_cost(E v, unsigned mi, unsigned mj);
template <>
__device__ __forceinline__ int entry_cost<unsigned>(unsigned v, unsigned mi, unsigned mj) {
  return (int)((mi >> (4 * ((v >> 16) & 7))) & 15u) + (int)((mj >> (4 * (v & 7))) & 15u);
}
template <>
__device__ __forceinline__ int entry_cost<unsigned short>(unsigned short v, unsigned mi, unsigned) {
  return (int)((mi >> (4 * (v & 7))) & 15u);
}
template <class E>
__device__ __forceinline__ void entry_mark(E v, unsigned& mi, unsigned& mj);
template <>
__device__ __forceinline__ void entry_mark<unsigned>(unsigned v, unsigned& mi, unsigned& mj) {
  mi += 1u << (4 * ((v >> 16) & 7));
  mj += 1u << (4 * (v & 7));
}
template <>
__device__ __forceinline__ void entry_mark<unsigned short>(unsigned short v, unsigned& mi, unsigned&) {
  mi += 1u << (4 * (v & 7));
}

template <class E>
__global__ void k_bank_order(int n_st, const int* __restrict__ tile_slot_off, const int* __restrict__ slot_a,
                             const int* __restrict__ slot_n, long long a_shift_unused, E* __restrict__ ent) {
  (void)a_shift_unused;
  for (int t = blockIdx.x; t < n_st; t += gridDim.x) {
    const int ps = tile_slot_off[t], np = tile_slot_off[t + 1] - ps;
    for (int qg = threadIdx.x; 4 * qg < np; qg += blockDim.x) {
      const int s0 = ps + 4 * qg, ns = min(4, np - 4 * qg);
      int a[4], n[4], maxn = 0;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        a[k] = k < ns ? slot_a[s0 + k] : 0;
        n[k] = k < ns ? slot_n[s0 + k] : 0;
        maxn = max(maxn, n[k]);
      }
      for (int e0 = 0; e0 < maxn; e0 += 2) {
        unsigned mi = 0, mj = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          for (int g = 0; g < 2; ++g) {
            const int e = e0 + g;
            if (e >= n[k]) break;
            E* base = ent + a[k];
            int best = e, bestc = 1 << 20;
            E bestv = base[e];
            for (int c = e; c < n[k]; ++c) {
              const E v = base[c];
              const int cost = entry_cost<E>(v, mi, mj);
              if (cost < bestc) { bestc = cost; best = c; bestv = v; if (cost == 0) break; }
            }
            if (best != e) { base[best] = base[e]; base[e] = bestv; }
            entry_mark<E>(bestv, mi, mj);
          }
        }
      }
    }
  }
}

// one CTA per tile: pack the tile's slot descriptors and entries into its blob
__global__ void k_blob_fill(int n_st, const int* __restrict__ blob_off, int4* __restrict__ info,
                            const int* __restrict__ slot_a, const int* __restrict__ slot_n,
                            const int* __restrict__ dslot_a, const int* __restrict__ dslot_n,
                            const int* __restrict__ tile_ent_off, const unsigned* __restrict__ tent,
                            const unsigned short* __restrict__ dent, uint4* __restrict__ blob) {
  for (int t = blockIdx.x; t < n_st; t += gridDim.x) {
    const int4 i0 = info[4 * t], i2 = info[4 * t + 2], i3 = info[4 * t + 3];
    const int np = i0.y, nd = i0.z, obs0 = i2.x, nobs = i2.y, npent = i2.w, ps = i3.x, ds = i3.y;
    const int off = blob_off[t];
    if (threadIdx.x == 0) info[4 * t + 1].x = off;
    int* w = reinterpret_cast<int*>(blob + off);
    if (threadIdx.x < 8) {
      const int hdr[8] = {np, nd, npent, nobs, ps, ds, 0, 0};
      w[threadIdx.x] = hdr[threadIdx.x];
    }
    int* pd = w + 8;
    const int e0 = tile_ent_off[t];
    for (int q = threadIdx.x; q < np; q += blockDim.x) { pd[2 * q] = slot_a[ps + q] - e0; pd[2 * q + 1] = slot_n[ps + q]; }
    int* dd = pd + 2 * np;
    for (int q = threadIdx.x; q < nd; q += blockDim.x) { dd[2 * q] = dslot_a[ds + q] - obs0; dd[2 * q + 1] = dslot_n[ds + q]; }
    unsigned* pe = reinterpret_cast<unsigned*>(dd + 2 * nd);
    for (int q = threadIdx.x; q < npent; q += blockDim.x) pe[q] = tent[e0 + q];
    unsigned short* de = reinterpret_cast<unsigned short*>(pe + npent);
    for (int q = threadIdx.x; q < nobs; q += blockDim.x) de[q] = dent[obs0 + q];
  }
}

inline unsigned nblk(int64_t n, int t) { return (unsigned)((n + t - 1) / t); }

template <class T>
int upload(Handle* h, T** dst, const std::vector<T>& src) {
  size_t n = std::max<size_t>(src.size(), 1);
  *dst = (T*)dalloc(h, n * sizeof(T));
  if (!*dst) return B200BA_OOM;
  if (src.empty()) return 0;
  const size_t bytes = src.size() * sizeof(T), pad = (bytes + 255) & ~(size_t)255;
  if (h->stage && h->stage_off + pad <= h->stage_bytes) {
    // through the page-locked staging area: no stream sync, src may die at once
    memcpy(h->stage + h->stage_off, src.data(), bytes);
    BA_CUDA(h, cudaMemcpyAsync(*dst, h->stage + h->stage_off, bytes, cudaMemcpyHostToDevice, h->stream));
    h->stage_off += pad;
    return 0;
  }
  BA_CUDA(h, cudaMemcpyAsync(*dst, src.data(), bytes, cudaMemcpyHostToDevice, h->stream));
  BA_CUDA(h, cudaStreamSynchronize(h->stream));  // src may be a temporary
  return 0;
}

struct Tmp {  // scratch freed (stream-ordered, back into the handle's pool) at the end of setup
  Handle* h;
  std::vector<void*> p;
  explicit Tmp(Handle* hh) : h(hh) {}
  ~Tmp() {
    if (h->stream2) cudaStreamSynchronize(h->stream2);   // an early return must not free d_uv under a running upload
    for (void* q : p) pool_free(h, q);
  }
  void* get(size_t bytes) {
    void* q = pool_alloc(h, std::max<size_t>(bytes, 16));
    if (q) p.push_back(q);
    return q;
  }
};

#define TMP_ALLOC(var, type, count)                                          \
  type* var = (type*)tmp.get((size_t)(count) * sizeof(type));                \
  if (!var) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)")
#define DEV_ALLOC(var, type, count)                                          \
  var = (type*)dalloc(h, (size_t)std::max<int64_t>((count), 1) * sizeof(type)); \
  if (!var) return B200BA_OOM


// ------------------------------------------------------------------ greedy packing on the device
// Consecutive items with prefix offsets off[0..n] (non-decreasing, off[n] = total) are packed greedily into bins of at
// most `cap`: a bin that starts at offset s ends at the largest item boundary <= s + cap, or -- when the first item alone
// is larger -- after that item.  This is the host loop "close the bin when the next item would overflow it", restated
// without the loop-carried dependence: nxt[p] (the item that starts the next bin) is a binary search per item, the bin
// starts are the orbit of item 0 under nxt, and the orbit is enumerated by pointer doubling (J_k = nxt^(2^k)): bin i
// starts at the item reached by the jumps of the set bits of i.  O(n log n) work in log n parallel passes instead of
// n dependent steps on one host core (config D: 4 M points, 8-10 ms on the host).
__global__ void k_gb_next(int n, const int* __restrict__ off, int cap, int* __restrict__ nxt) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p > n) return;
  const int total = off[n], s = off[p];
  if (s >= total) { nxt[p] = n; return; }
  int lo = p, hi = n + 1;                 // last index j in [p, n] with off[j] <= s + cap
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (off[mid] <= s + cap) lo = mid; else hi = mid; }
  int j = lo;
  if (off[j] <= s) {                      // the first item alone overflows the bin: first index with off > s
    lo = p; hi = n;                       // off[n] = total > s
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (off[mid] <= s) lo = mid; else hi = mid; }
    j = hi;
  }
  nxt[p] = j;
}

__global__ void k_gb_double(int m, const int* __restrict__ prev, int* __restrict__ next) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < m) next[p] = prev[prev[p]];
}

// number of bins = 1 + number of jumps from item 0 that do not yet reach the end
__global__ void k_gb_count(int n, const int* __restrict__ off, const int* __restrict__ J, int K, int* __restrict__ out) {
  if (blockIdx.x || threadIdx.x) return;
  const int total = off[n];
  if (total <= 0) { *out = 0; return; }
  int p = 0, cnt = 0;
  for (int k = K - 1; k >= 0; --k) {
    const int q = J[(int64_t)k * (n + 1) + p];
    if (off[q] < total) { p = q; cnt += 1 << k; }
  }
  *out = cnt + 1;
}

// bin_item[i] = first item of bin i (i < n_bins), bin_item[n_bins] = n;  bin_off[i] = off[bin_item[i]]
__global__ void k_gb_starts(int n_bins, int n, const int* __restrict__ off, const int* __restrict__ J, int K,
                            int* __restrict__ bin_item, int* __restrict__ bin_off) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n_bins) return;
  int p = 0;
  if (i == n_bins) p = n;
  else
    for (int k = 0; k < K; ++k)
      if ((i >> k) & 1) p = J[(int64_t)k * (n + 1) + p];
  if (bin_item) bin_item[i] = p;
  bin_off[i] = off[p];
}

__global__ void k_max_degree(int P, const int* __restrict__ off, int* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  int d = (p < P) ? off[p + 1] - off[p] : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) d = max(d, __shfl_xor_sync(0xffffffffu, d, o));
  if ((threadIdx.x & 31) == 0 && d > 0) atomicMax(out, d);
}

// bins of at most `cap` over the items of off[0..n]; *bin_off_out (problem lifetime, n_bins + 1 offsets) and, if asked
// for, *bin_item_out (problem lifetime, n_bins + 1 item indices); returns the bin count through *n_bins_out
static int greedy_bins(Handle* h, Tmp& tmp, const int* d_off, int n, int cap, int* n_bins_out, int** bin_off_out,
                       int** bin_item_out) {
  cudaStream_t st = h->stream;
  const int T = 256;
  int K = 1;
  while (((int64_t)1 << K) <= (int64_t)n + 1 && K < 31) ++K;
  TMP_ALLOC(d_J, int, (int64_t)K * (n + 1));
  TMP_ALLOC(d_cnt, int, 1);
  k_gb_next<<<(n + 1 + T - 1) / T, T, 0, st>>>(n, d_off, cap, d_J);
  for (int k = 1; k < K; ++k)
    k_gb_double<<<(n + 1 + T - 1) / T, T, 0, st>>>(n + 1, d_J + (int64_t)(k - 1) * (n + 1), d_J + (int64_t)k * (n + 1));
  k_gb_count<<<1, 32, 0, st>>>(n, d_off, d_J, K, d_cnt);
  int n_bins = 0;
  BA_CUDA(h, cudaMemcpyAsync(&n_bins, d_cnt, sizeof(int), cudaMemcpyDeviceToHost, st));
  BA_CUDA(h, cudaStreamSynchronize(st));
  *n_bins_out = n_bins;
  *bin_off_out = (int*)dalloc(h, (size_t)(n_bins + 1 + 8) * sizeof(int));   // + slack: staged as 16-byte aligned windows
  if (!*bin_off_out) return B200BA_OOM;
  int* items = nullptr;
  if (bin_item_out) {
    items = (int*)dalloc(h, (size_t)(n_bins + 1) * sizeof(int));
    if (!items) return B200BA_OOM;
    *bin_item_out = items;
  }
  k_gb_starts<<<(n_bins + 1 + T - 1) / T, T, 0, st>>>(n_bins, n, d_off, d_J, K, items, *bin_off_out);
  BA_CUDA(h, cudaGetLastError());
  return 0;
}

// Slots of a tile-grouped entry list.  Input: the (tile, id) cells of the list in sorted order (cellkey = tile * M + id,
// cnt = entries per cell; both with room for n_cell + 1 elements).  Every cell is cut into slots of <= SLOT_CAP entries;
// the slots of a tile are ordered by decreasing entry count (lanes of a warp finish together); grp_* lists the slots
// of every group (block of S / camera) in slot order for the fixed-order reduction.  gkeys maps id -> group (binary
// search in the sorted global key list) or is null when the id is the group index.
struct SlotList {
  int n_slot = 0;
  int* slot_a = nullptr;
  int* slot_n = nullptr;
  int* tile_off = nullptr;   // [n_st+1]
  int* grp_off = nullptr;    // [n_grp+1]
  int* grp_slots = nullptr;  // [n_slot]
};

int build_slots(Handle* h, Tmp& tmp, int n_cell, const long long* d_cellkey, int* d_cnt, long long M,
                const long long* d_gkeys, int n_grp, int n_st, SlotList* out) {
  cudaStream_t st = h->stream;
  const int T = 256;
  DEV_ALLOC(out->tile_off, int, n_st + 1);
  DEV_ALLOC(out->grp_off, int, n_grp + 1);
  BA_CUDA(h, cudaMemsetAsync(out->tile_off, 0, (n_st + 1) * sizeof(int), st));
  BA_CUDA(h, cudaMemsetAsync(out->grp_off, 0, (n_grp + 1) * sizeof(int), st));
  int n_slot = 0;
  int *d_sub_off = nullptr, *d_cell_off = nullptr, *d_cell_blk = nullptr, *d_cell_tile = nullptr;
  if (n_cell > 0) {
    TMP_ALLOC(a_cell_off, int, n_cell + 1);
    TMP_ALLOC(a_cell_blk, int, n_cell);
    TMP_ALLOC(a_cell_tile, int, n_cell);
    TMP_ALLOC(d_nsub, int, n_cell + 1);
    TMP_ALLOC(a_sub_off, int, n_cell + 1);
    d_sub_off = a_sub_off; d_cell_off = a_cell_off; d_cell_blk = a_cell_blk; d_cell_tile = a_cell_tile;
    BA_CUDA(h, cudaMemsetAsync(d_cnt + n_cell, 0, sizeof(int), st));
    BA_CUDA(h, cudaMemsetAsync(d_nsub + n_cell, 0, sizeof(int), st));
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, d_cnt, d_cell_off, n_cell + 1, st);
    TMP_ALLOC(d_scan2, char, bytes);
    BA_CUDA(h, cub::DeviceScan::ExclusiveSum(d_scan2, bytes, d_cnt, d_cell_off, n_cell + 1, st));
    k_cell_meta<<<nblk(n_cell, T), T, 0, st>>>(n_cell, d_cellkey, d_cnt, M, d_gkeys, n_grp, d_cell_blk, d_cell_tile, d_nsub);
    BA_CUDA(h, cub::DeviceScan::ExclusiveSum(d_scan2, bytes, d_nsub, d_sub_off, n_cell + 1, st));
    BA_CUDA(h, cudaMemcpyAsync(&n_slot, d_sub_off + n_cell, sizeof(int), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
  }
  out->n_slot = n_slot;
  DEV_ALLOC(out->slot_a, int, n_slot);
  DEV_ALLOC(out->slot_n, int, n_slot);
  DEV_ALLOC(out->grp_slots, int, n_slot);
  if (n_slot > 0) {
    TMP_ALLOC(d_sub_a, int, n_slot);
    TMP_ALLOC(d_sub_n, int, n_slot);
    TMP_ALLOC(d_sub_blk, int, n_slot);
    TMP_ALLOC(d_key2, long long, n_slot);
    TMP_ALLOC(d_key2s, long long, n_slot);
    TMP_ALLOC(d_si, int, n_slot);
    TMP_ALLOC(d_order, int, n_slot);
    TMP_ALLOC(d_slot_blk, int, n_slot);
    TMP_ALLOC(d_blk_sorted, int, n_slot);
    k_sub_fill<<<nblk(n_cell, T), T, 0, st>>>(n_cell, d_sub_off, d_cell_off, d_cnt, d_cell_blk, d_cell_tile, d_sub_a, d_sub_n,
                                              d_sub_blk, d_key2);
    k_iota<<<nblk(n_slot, T), T, 0, st>>>(n_slot, d_si);
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_key2, d_key2s, d_si, d_order, n_slot, 0, 63, st);
    TMP_ALLOC(d_sort3, char, bytes);
    BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort3, bytes, d_key2, d_key2s, d_si, d_order, n_slot, 0, 63, st));
    k_slot_fill<<<nblk(n_slot, T), T, 0, st>>>(n_slot, d_order, d_sub_a, d_sub_n, d_sub_blk, out->slot_a, out->slot_n,
                                               d_slot_blk);
    k_tile_slot_off<<<nblk(n_st + 1, T), T, 0, st>>>(n_st, n_slot, d_key2s, out->tile_off);
    int eb = 1;
    while (eb < 31 && (1 << eb) < n_grp) ++eb;
    bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_slot_blk, d_blk_sorted, d_si, out->grp_slots, n_slot, 0, eb, st);
    TMP_ALLOC(d_sort4, char, bytes);
    BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort4, bytes, d_slot_blk, d_blk_sorted, d_si, out->grp_slots, n_slot, 0, eb,
                                               st));
    k_offsets<<<nblk(n_slot, T), T, 0, st>>>(n_slot, n_grp, d_blk_sorted, out->grp_off);
  }
  BA_CUDA(h, cudaStreamSynchronize(st));
  return 0;
}

}  // namespace

int setup_problem(Handle* h, int64_t C, int64_t P, int64_t N, const int64_t* cam_idx, const int64_t* pt_idx,
                  const double* uv, const double* K) {
  Problem& pb = h->pb;
  cudaStream_t st = h->stream;
  pb = Problem();
  pb.C = C; pb.P = P; pb.N = N;
  cudaStreamSynchronize(st);   // nothing of an earlier call still reads the staging area
  h->stage_off = 0;
  pb.Cs = (C + 1) & ~(int64_t)1;
  for (int i = 0; i < 9; ++i) pb.K.k[i] = K[i];
  Tmp tmp(h);
  const int T = 256;
  // B200BA_SETUP_PROF=1: wall time of every setup phase to stderr (each mark synchronises the stream)
  const bool prof = getenv("B200BA_SETUP_PROF") && atoi(getenv("B200BA_SETUP_PROF"));
  auto t_prev = std::chrono::steady_clock::now();
  auto mark = [&](const char* what) {
    if (!prof) return;
    cudaStreamSynchronize(st);
    auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "[b200ba setup] %-28s %9.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t_prev).count());
    t_prev = t;
  };

  // ---- 1. raw inputs to the device, range / order check, int32 indices
  BigCopy uv_copy;   // declared after `tmp`: its workers are joined before the scratch (d_uv) is freed on an early return
  TMP_ALLOC(d_cam64, int64_t, N);
  TMP_ALLOC(d_pt64, int64_t, N);
  TMP_ALLOC(d_uv, double, 2 * N);
  TMP_ALLOC(d_cam_in, int, N);
  TMP_ALLOC(d_pt_in, int, N);
  TMP_ALLOC(d_iota, int, N);
  TMP_ALLOC(d_flags, int, 2);
  BA_CUDA(h, cudaMemsetAsync(d_flags, 0, 2 * sizeof(int), st));
  if (N > 0) {
    // Page-locked sources go as one DMA each; pageable ones (what numpy hands over) through big_copy's worker threads.
    // Only the point indices are needed at once (order, offsets, tiles): the camera indices and then the pixels travel
    // on the second stream underneath steps 2 and 3, the camera check / conversion runs there too.
    BA_CUDA(h, cudaEventRecord(h->ev_a, st));                 // the scratch arrays are stream-ordered allocations of `st`
    BA_CUDA(h, cudaStreamWaitEvent(h->stream2, h->ev_a, 0));
    {
      BigCopy bc;
      if (int rc = big_copy_start(h, bc, d_pt64, pt_idx, N * sizeof(int64_t), true, st)) return rc;
      if (bc.wait()) return fail(h, B200BA_CUDA, "upload of the point indices failed");
      k_check_convert_pt<<<nblk(N, T), T, 0, st>>>(N, P, d_pt64, d_pt_in, d_iota, d_flags);
      if (int rc = big_copy_start(h, bc, d_cam64, cam_idx, N * sizeof(int64_t), true, h->stream2)) return rc;
      if (bc.wait()) return fail(h, B200BA_CUDA, "upload of the camera indices failed");
    }
    k_check_convert_cam<<<nblk(N, T), T, 0, h->stream2>>>(N, C, d_cam64, d_cam_in, d_flags);
    BA_CUDA(h, cudaEventRecord(h->ev_c, h->stream2));
    // the pixel array (half of the upload): nothing needs it before k_gather_uv at the end
    if (int rc = big_copy_start(h, uv_copy, d_uv, uv, 2 * N * sizeof(double), true, h->stream2)) return rc;
    if (uv_copy.th.empty()) BA_CUDA(h, cudaEventRecord(h->ev_b, h->stream2));
  }
  int flags = 0;
  BA_CUDA(h, cudaMemcpyAsync(&flags, d_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
  BA_CUDA(h, cudaStreamSynchronize(st));
  if (flags & 1) return fail(h, B200BA_INDEX_RANGE, "camera or point index out of range");

  mark("1 upload+check (points)");
  // ---- 2. point-major order (stable); identity when the input is already sorted (graph.py:186-191)
  DEV_ALLOC(pb.u, double, N + 2);  // + slack for 16-byte aligned prefetch windows
  DEV_ALLOC(pb.v, double, N + 2);
  DEV_ALLOC(pb.cam, int, N + 16);  // + slack: the dataflow kernel bulk-copies 16-byte aligned windows
  DEV_ALLOC(pb.pt, int, N + 16);
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
  if (N > 0) k_offsets<<<nblk(N, T), T, 0, st>>>(N, (int)P, pb.pt, pb.pt_off);

  mark("2 point-major order");
  // ---- 3. warp tiles: greedy packing of whole points into <= 32 observations (a point with more is its own tile),
  //         CTA tiles: greedy packing of warp tiles into <= tile_obs observations -- both on the device (greedy_bins)
  {
    int max_deg = 0;
    {
      TMP_ALLOC(d_md, int, 1);
      BA_CUDA(h, cudaMemsetAsync(d_md, 0, sizeof(int), st));
      if (P > 0) k_max_degree<<<nblk(P, T), T, 0, st>>>((int)P, pb.pt_off, d_md);
      BA_CUDA(h, cudaMemcpyAsync(&max_deg, d_md, sizeof(int), cudaMemcpyDeviceToHost, st));
    }
    if (int rc = greedy_bins(h, tmp, pb.pt_off, (int)P, 32, &pb.n_wt, &pb.wt_off, nullptr)) return rc;   // syncs: max_deg is in
    // CTA tiles: runs of warp tiles (whole points) with at most tile_obs observations (records fit shared memory)
    const char* e_fused = getenv("B200BA_FUSED");        // 0: unfused A/B path (k_lin_points + k_cam_side + k_schur_tiles)
    const char* e_tile = getenv("B200BA_TILE_OBS");      // tuning sweeps
    const bool want_fused = !(e_fused && atoi(e_fused) == 0);
    const int cap = want_fused ? FUSED_TILE_1CTA : SCHUR_TILE_MAX;
    const char* e_flow0 = getenv("B200BA_FLOW");
    const bool want_flow = want_fused && !(e_flow0 && atoi(e_flow0) == 0);
    const int tile_max = want_flow ? FLOW_TILE : want_fused ? FUSED_TILE_2CTA : SCHUR_TILE_MAX;
    int tile_obs = (int)std::min<int64_t>(tile_max, std::max<int64_t>(256, (N / (4 * (int64_t)h->sm_count) + 31) / 32 * 32));
    if (e_tile && atoi(e_tile) >= 32) tile_obs = std::min(cap, atoi(e_tile) / 32 * 32);  // exact override (tests, sweeps)
    if (max_deg > tile_obs) tile_obs = std::min(cap, (max_deg + 31) / 32 * 32);  // one big point must fit a tile
    const char* force_stream = getenv("B200BA_SCHUR_STREAM");
    pb.schur_tiles = max_deg <= tile_obs && !(force_stream && atoi(force_stream));
    if (pb.schur_tiles) {
      int n_st = 0;
      int* st_off = nullptr;
      int* st_wt = nullptr;
      if (int rc = greedy_bins(h, tmp, pb.wt_off, pb.n_wt, tile_obs, &n_st, &st_off, &st_wt)) return rc;
      pb.n_st = n_st;
      pb.tile_obs = tile_obs;
      if ((double)(pb.n_st + 1) * (double)C * (double)C >= 4.0e18) pb.schur_tiles = false;  // key would overflow
      else {
        pb.fused = want_fused;
        pb.st_off = st_off;
        pb.st_wt_off = st_wt;
      }
    }
  }

  mark("3 warp/CTA tiles");
  // ---- 3b. camera indices into point-major order (their upload and check ran underneath steps 2 and 3)
  if (N > 0) {
    BA_CUDA(h, cudaStreamWaitEvent(st, h->ev_c, 0));
    k_gather_obs<<<nblk(N, T), T, 0, st>>>(N, pb.perm, d_cam_in, pb.cam);
    int cflag = 0;
    BA_CUDA(h, cudaMemcpyAsync(&cflag, d_flags + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
    if (cflag & 1) return fail(h, B200BA_INDEX_RANGE, "camera or point index out of range");
  }
  mark("3b camera indices");
  // ---- 4. camera-sorted permutation and chunks: only the unfused pipeline (k_cam_side / k_cam_reduce) reads them; the
  //         fused kernels group a tile's observations by camera through the self slots of step 9 (4.6 ms saved at config D)
  if (!pb.fused) {
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
  }

  mark("4 camera permutation");
  // ---- 5. Schur pair entries sorted by key: [tile *] C*C + cam_i * C + cam_j  (cam_i <= cam_j)
  const long long CC = (long long)C * C;
  int* d_pt_tile = nullptr;
  if (pb.schur_tiles) {
    d_pt_tile = (int*)tmp.get(std::max<size_t>(P, 1) * sizeof(int));
    if (!d_pt_tile) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)");
    if (P > 0) k_pt_tile<<<nblk(P, T), T, 0, st>>>((int)P, pb.pt_off, pb.st_off, pb.n_st, d_pt_tile);
  }
  TMP_ALLOC(d_cnt, long long, P + 1);
  TMP_ALLOC(d_ent_off, long long, P + 1);
  BA_CUDA(h, cudaMemsetAsync(d_cnt, 0, (P + 1) * sizeof(long long), st));
  if (P > 0) k_pair_count<<<nblk(P, T), T, 0, st>>>((int)P, pb.pt_off, pb.cam, pb.fused, d_cnt);
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
  long long* d_ent64 = nullptr;  // sorted (i << 32) | j : kept in stream mode, scratch in tile mode
  if (pb.schur_tiles) {
    DEV_ALLOC(pb.tent, unsigned, n_ent);   // the sort writes the packed tile-local entries straight into place
  } else {
    DEV_ALLOC(pb.ent, long long, n_ent);
    d_ent64 = pb.ent;
  }
  TMP_ALLOC(d_keys_sorted, long long, n_ent);
  TMP_ALLOC(d_uniq, long long, n_ent + 1);   // run keys: blocks (stream mode) or (tile, block) cells (tile mode)
  TMP_ALLOC(d_runs, int, n_ent + 1);         // run lengths
  int nruns = 0;
  std::vector<long long> keys_host;  // unique local block keys
  {
    TMP_ALLOC(d_keys, long long, n_ent);
    long long* d_vals = nullptr;
    unsigned* d_vals32 = nullptr;
    if (pb.schur_tiles) {
      d_vals32 = (unsigned*)tmp.get(std::max<size_t>(n_ent, 1) * sizeof(unsigned));
      if (!d_vals32) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)");
    } else {
      d_vals = (long long*)tmp.get(std::max<size_t>(n_ent, 1) * sizeof(long long));
      if (!d_vals) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)");
    }
    if (P > 0)
      k_pair_fill<<<nblk(P, T), T, 0, st>>>((int)P, (long long)C, pb.pt_off, pb.cam, pb.fused, d_ent_off, d_pt_tile, d_keys,
                                            d_vals, pb.st_off, d_vals32);
    if (n_ent > 0) {
      const double kmax = (double)CC * (double)(pb.schur_tiles ? pb.n_st : 1);
      int end_bit = 1;
      while (end_bit < 63 && (double)((long long)1 << end_bit) < kmax) ++end_bit;
      size_t bytes = 0;
      char* d_sort = nullptr;
      if (pb.schur_tiles) {
        cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_keys, d_keys_sorted, d_vals32, pb.tent, (int)n_ent, 0, end_bit, st);
        d_sort = (char*)tmp.get(std::max<size_t>(bytes, 1));
        if (!d_sort) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)");
        BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort, bytes, d_keys, d_keys_sorted, d_vals32, pb.tent, (int)n_ent, 0,
                                                   end_bit, st));
      } else {
        cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_keys, d_keys_sorted, d_vals, d_ent64, (int)n_ent, 0, end_bit, st);
        d_sort = (char*)tmp.get(std::max<size_t>(bytes, 1));
        if (!d_sort) return fail(h, B200BA_OOM, "out of device memory in set_problem (scratch)");
        BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort, bytes, d_keys, d_keys_sorted, d_vals, d_ent64, (int)n_ent, 0,
                                                   end_bit, st));
      }
      TMP_ALLOC(d_nruns, int, 1);
      bytes = 0;
      cub::DeviceRunLengthEncode::Encode(nullptr, bytes, d_keys_sorted, d_uniq, d_runs, d_nruns, (int)n_ent, st);
      TMP_ALLOC(d_rle, char, bytes);
      BA_CUDA(h, cub::DeviceRunLengthEncode::Encode(d_rle, bytes, d_keys_sorted, d_uniq, d_runs, d_nruns, (int)n_ent, st));
      BA_CUDA(h, cudaMemcpyAsync(&nruns, d_nruns, sizeof(int), cudaMemcpyDeviceToHost, st));
      BA_CUDA(h, cudaStreamSynchronize(st));
      if (!pb.schur_tiles) {
        keys_host.resize(nruns);
        BA_CUDA(h, cudaMemcpy(keys_host.data(), d_uniq, (size_t)nruns * sizeof(long long), cudaMemcpyDeviceToHost));
      } else {
        // block keys of the cells -> sort -> unique
        TMP_ALLOC(d_bk, long long, nruns);
        TMP_ALLOC(d_bk_sorted, long long, nruns);
        TMP_ALLOC(d_bk_uniq, long long, nruns);
        TMP_ALLOC(d_bk_cnt, int, nruns);
        k_mod_keys<<<nblk(nruns, T), T, 0, st>>>(nruns, d_uniq, CC, d_bk);
        int eb = 1;
        while (eb < 63 && (double)((long long)1 << eb) < (double)CC) ++eb;
        bytes = 0;
        cub::DeviceRadixSort::SortKeys(nullptr, bytes, d_bk, d_bk_sorted, nruns, 0, eb, st);
        TMP_ALLOC(d_sort2, char, bytes);
        BA_CUDA(h, cub::DeviceRadixSort::SortKeys(d_sort2, bytes, d_bk, d_bk_sorted, nruns, 0, eb, st));
        bytes = 0;
        cub::DeviceRunLengthEncode::Encode(nullptr, bytes, d_bk_sorted, d_bk_uniq, d_bk_cnt, d_nruns, nruns, st);
        TMP_ALLOC(d_rle2, char, bytes);
        BA_CUDA(h, cub::DeviceRunLengthEncode::Encode(d_rle2, bytes, d_bk_sorted, d_bk_uniq, d_bk_cnt, d_nruns, nruns, st));
        int nb = 0;
        BA_CUDA(h, cudaMemcpyAsync(&nb, d_nruns, sizeof(int), cudaMemcpyDeviceToHost, st));
        BA_CUDA(h, cudaStreamSynchronize(st));
        keys_host.resize(nb);
        BA_CUDA(h, cudaMemcpy(keys_host.data(), d_bk_uniq, (size_t)nb * sizeof(long long), cudaMemcpyDeviceToHost));
      }
    }
  }

  mark("5 pair entries");
  // ---- 6. global block structure: union over ranks of the local keys + every diagonal block
  std::vector<long long> x_all;       // every rank's local block keys (sparse exchange), padded to x_mx per rank
  std::vector<long long> x_counts;
  long long x_mx = 0;
  bool x_ok = false;
  if (h->nranks > 1) {
    // sparse exchange needs the fused reducers on every rank; local blocks then include the diagonal blocks of the
    // cameras this rank observes (self-slot contributions)
    const char* e_x = getenv("B200BA_XCHG");  // 0: always all-reduce, 1: always sparse exchange (when possible)
    const bool want_x = pb.fused && pb.schur_tiles && !(e_x && atoi(e_x) == 0);
    if (want_x) {
      // cameras this rank observes (the camera-sorted offsets of step 4 only exist in the unfused pipeline)
      std::vector<int> h_seen(C + 1, 0);
      TMP_ALLOC(d_seen, int, C + 1);
      BA_CUDA(h, cudaMemsetAsync(d_seen, 0, (C + 1) * sizeof(int), st));
      if (N > 0) k_cam_seen<<<nblk(N, T), T, 0, st>>>(N, pb.cam, d_seen);
      BA_CUDA(h, cudaMemcpyAsync(h_seen.data(), d_seen, (C + 1) * sizeof(int), cudaMemcpyDeviceToHost, st));
      BA_CUDA(h, cudaStreamSynchronize(st));
      // both lists are ascending (radix-sorted unique keys; camera order): a linear merge, not a sort
      std::vector<long long> dg, mg;
      for (int64_t c = 0; c < C; ++c)
        if (h_seen[c]) dg.push_back((long long)c * C + c);
      mg.resize(keys_host.size() + dg.size());
      std::merge(keys_host.begin(), keys_host.end(), dg.begin(), dg.end(), mg.begin());
      mg.erase(std::unique(mg.begin(), mg.end()), mg.end());
      keys_host.swap(mg);
    }
    long long my_n[2] = {(long long)keys_host.size(), want_x ? 1 : 0};
    TMP_ALLOC(d_n, long long, 2 * (h->nranks + 1));
    BA_CUDA(h, cudaMemcpyAsync(d_n + 2 * h->nranks, my_n, 2 * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (int rc = comm_allgather_i64(h, d_n + 2 * h->nranks, d_n, 2)) return rc;
    std::vector<long long> cnt2(2 * h->nranks);
    BA_CUDA(h, cudaMemcpyAsync(cnt2.data(), d_n, 2 * h->nranks * sizeof(long long), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
    std::vector<long long> counts(h->nranks);
    x_ok = true;
    for (int r = 0; r < h->nranks; ++r) { counts[r] = cnt2[2 * r]; x_ok = x_ok && cnt2[2 * r + 1] != 0; }
    long long mx = 1;
    for (long long c : counts) mx = std::max(mx, c);
    TMP_ALLOC(d_send, long long, mx);
    TMP_ALLOC(d_recv, long long, mx * h->nranks);
    BA_CUDA(h, cudaMemsetAsync(d_send, 0, mx * sizeof(long long), st));
    if (my_n[0])
      BA_CUDA(h, cudaMemcpyAsync(d_send, keys_host.data(), my_n[0] * sizeof(long long), cudaMemcpyHostToDevice, st));
    if (int rc = comm_allgather_i64(h, d_send, d_recv, mx)) return rc;
    std::vector<long long> all((size_t)mx * h->nranks);
    BA_CUDA(h, cudaMemcpyAsync(all.data(), d_recv, all.size() * sizeof(long long), cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
    // every rank's list is ascending: merge them one after the other (duplicates are removed with the diagonal below)
    keys_host.clear();
    {
      std::vector<long long> tmp;
      for (int r = 0; r < h->nranks; ++r) {
        const long long* a0 = all.data() + (size_t)r * mx;
        tmp.resize(keys_host.size() + (size_t)counts[r]);
        std::merge(keys_host.begin(), keys_host.end(), a0, a0 + counts[r], tmp.begin());
        keys_host.swap(tmp);
      }
    }
    if (x_ok) { x_all.swap(all); x_counts = counts; x_mx = mx; }
    mark("6a gather block keys");
  }
  {
    // union with every diagonal block.  One rank: the local list is already sorted (device run-length encode) and so
    // are the diagonal keys -> a linear merge instead of a 170 k-element sort (2.5 ms at config D)
    std::vector<long long> diag((size_t)C);
    for (int64_t c = 0; c < C; ++c) diag[(size_t)c] = (long long)c * C + c;
    if (!std::is_sorted(keys_host.begin(), keys_host.end())) std::sort(keys_host.begin(), keys_host.end());
    std::vector<long long> merged(keys_host.size() + diag.size());
    std::merge(keys_host.begin(), keys_host.end(), diag.begin(), diag.end(), merged.begin());
    merged.erase(std::unique(merged.begin(), merged.end()), merged.end());
    keys_host.swap(merged);
  }
  const int n_blk = (int)keys_host.size();
  pb.n_blk = n_blk;
  mark("6b union of keys");
  if (x_ok) {
    // worth it when the gathered buffer (nranks x the largest local list) is smaller than what an all-reduce moves
    // (about twice the union); the decision depends only on gathered counts, so every rank takes the same one
    const char* e_x = getenv("B200BA_XCHG");
    const bool force = e_x && atoi(e_x) == 1;
    if (force || (double)x_mx * h->nranks <= 1.5 * (double)n_blk) {
      std::vector<int> xoff(n_blk + 1, 0), mine;
      std::vector<std::vector<int>> idx(h->nranks);
      for (int r = 0; r < h->nranks; ++r) {
        idx[r].resize(x_counts[r]);
        int g = 0;   // the rank's keys ascend and are all in the union: one forward walk instead of a search per key
        for (long long k = 0; k < x_counts[r]; ++k) {
          const long long key = x_all[(size_t)r * x_mx + k];
          while (g + 1 < n_blk && keys_host[g] < key) ++g;
          idx[r][k] = g;
          xoff[g + 1]++;
        }
      }
      for (int g = 0; g < n_blk; ++g) xoff[g + 1] += xoff[g];
      pb.xl_max = (int)x_mx;
      pb.xl_n = (int)x_counts[h->rank];
      pb.x_stride = ((long long)pb.xl_max * 36 + 18 * C + S_NSUM + 1 + 1) & ~1ll;
      std::vector<long long> xsrc(xoff[n_blk]);
      std::vector<int> fill(xoff.begin(), xoff.end() - 1);
      for (int r = 0; r < h->nranks; ++r)  // ascending rank inside every block: the summation order of k_xchg_sum
        for (long long k = 0; k < x_counts[r]; ++k) xsrc[fill[idx[r][k]]++] = r * pb.x_stride + k * 36;
      if (int rc = upload(h, &pb.xl_blk, idx[h->rank])) return rc;
      if (int rc = upload(h, &pb.xc_off, xoff)) return rc;
      if (int rc = upload(h, &pb.xc_src, xsrc)) return rc;
      pb.xchg = true;
    }
    mark("6c exchange lists");
  }

  // entry range of every global block in the local sorted entry list (stream mode only)
  std::vector<long long> h_blk_off(n_blk + 1, 0);
  TMP_ALLOC(d_gkeys, long long, n_blk);
  BA_CUDA(h, cudaMemcpyAsync(d_gkeys, keys_host.data(), (size_t)n_blk * sizeof(long long), cudaMemcpyHostToDevice, st));
  if (!pb.schur_tiles) {
    TMP_ALLOC(d_blk_off, long long, n_blk + 1);
    k_lower_bound<<<nblk(n_blk + 1, T), T, 0, st>>>(n_blk, d_gkeys, n_ent, d_keys_sorted, d_blk_off);
    BA_CUDA(h, cudaMemcpyAsync(h_blk_off.data(), d_blk_off, (size_t)(n_blk + 1) * sizeof(long long),
                               cudaMemcpyDeviceToHost, st));
    BA_CUDA(h, cudaStreamSynchronize(st));
  }

  mark("6 block structure");
  // ---- 7. host: block coordinates, Schur chunks, full BSR view, camera aggregates of the two-level preconditioner.
  // Pure host work on the gathered keys (about 5 ms at config D): it runs on a worker thread while this thread drives
  // the device through steps 8 - 10, which only need the keys already uploaded in step 6; the results are uploaded
  // after the join (step 10b).
  // The result vectors live in the handle: a fresh multi-megabyte vector per call is a fresh mmap, and its page faults
  // cost as much as the work (7.3 -> 4.9 ms for this step at config D sizes).
  HostBsr& hb = h->hb;
  hb.sb.clear(); hb.sa.clear(); hb.se.clear(); hb.sp_col.clear();
  hb.n_agg = 0;
  hb.tl = false;
  const char* e_tl = getenv("B200BA_TWOLEVEL");   // 0: block-Jacobi only
  const char* e_agg = getenv("B200BA_AGG");       // cameras per aggregate (default max(16, C / 320))
  const bool want_tl = C >= 128 && !(e_tl && atoi(e_tl) == 0);
  const int agg_env = (e_agg && atoi(e_agg) >= 2) ? atoi(e_agg) : 0;
  auto host_bsr = [&hb, &keys_host, &h_blk_off, n_blk, C, want_tl, agg_env]() {
    std::vector<int>&br = hb.br, &bc = hb.bc, &diag = hb.diag, &sb = hb.sb, &sa = hb.sa, &se = hb.se, &choff = hb.choff;
    std::vector<int>&rowptr = hb.rowptr, &col = hb.col, &src = hb.src;
    br.resize(n_blk); bc.resize(n_blk); diag.assign(C, -1); choff.assign(n_blk + 1, 0);
    std::vector<int> rowcnt(C + 1, 0);
    long long row = 0;   // keys ascend: follow the row instead of a 64-bit division per block
    for (int g = 0; g < n_blk; ++g) {
      while (keys_host[g] >= (row + 1) * (long long)C) ++row;
      br[g] = (int)row;
      bc[g] = (int)(keys_host[g] - row * (long long)C);
      if (br[g] == bc[g]) { diag[br[g]] = g; rowcnt[br[g]]++; }
      else { rowcnt[br[g]]++; rowcnt[bc[g]]++; }
      choff[g] = (int)sb.size();
      for (long long a = h_blk_off[g]; a < h_blk_off[g + 1]; a += SCHUNK) {
        sb.push_back(g); sa.push_back((int)a); se.push_back((int)std::min<long long>(a + SCHUNK, h_blk_off[g + 1]));
      }
    }
    choff[n_blk] = (int)sb.size();
    rowptr.assign(C + 1, 0);
    for (int64_t c = 0; c < C; ++c) rowptr[c + 1] = rowptr[c] + rowcnt[c];
    const int nnzb = rowptr[C];
    col.resize(nnzb); src.resize(nnzb);
    std::vector<int> fill(rowptr.begin(), rowptr.end() - 1);
    // keys are sorted by (r,c): pushing the direct block into row r and the transposed one into row c keeps
    // every row's columns ascending (transposed entries of row c come from r < c, direct ones from c' >= c)
    for (int pass = 0; pass < 2; ++pass)
      for (int g = 0; g < n_blk; ++g) {
        if (pass == 0 && br[g] != bc[g]) { int q = fill[bc[g]]++; col[q] = br[g]; src[q] = g | (int)0x80000000; }
        if (pass == 1) { int q = fill[br[g]]++; col[q] = bc[g]; src[q] = g; }
      }
    // two-level preconditioner (k_twolevel.cuh): camera aggregates and the block structure of S P
    if (!want_tl) return;
    int per = (int)std::max<int64_t>(16, (C + 319) / 320);
    if (agg_env) per = agg_env;
    int n_agg = (int)((C + per - 1) / per);
    if (n_agg > 1 && C - (int64_t)(n_agg - 1) * per < 2) --n_agg;   // a one-camera aggregate has only 6 free motions
    if ((int64_t)n_agg * 7 > 3584) return;
    // Aggregates = connected groups of about `per` cameras of the graph of S (block (i, j) present <=> the two cameras
    // share a point), grown greedily from the lowest unassigned camera by strongest connection to the group.  On
    // an ordered sequence this gives index runs; on an arbitrary camera numbering (incremental SfM registers cameras
    // in match order) it still groups cameras that see the same points, which is what makes a locally rigid motion
    // cheap.  Groups smaller than half the target are merged into the neighbouring aggregate they touch most.
    std::vector<int>& cam_agg = hb.cam_agg;
    cam_agg.assign(C, -1);
    std::vector<int> agg_size;
    {
      // greedy growth: repeatedly add the unassigned camera with the most blocks of S into the aggregate so far
      // (ties: lowest index) -- on a sequence this walks outwards from the seed, whatever the numbering
      std::vector<int> score(C, 0), cand;
      for (int64_t seed = 0; seed < C; ++seed) {
        if (cam_agg[seed] >= 0) continue;
        const int I = (int)agg_size.size();
        cand.clear();
        int size = 0, next = (int)seed;
        while (next >= 0) {
          cam_agg[next] = I;
          ++size;
          for (int q = rowptr[next]; q < rowptr[next + 1]; ++q) {
            const int j = col[q];
            if (cam_agg[j] >= 0) continue;
            if (score[j]++ == 0) cand.push_back(j);
          }
          next = -1;
          if (size >= per) break;
          int best = -1;
          for (size_t k = 0; k < cand.size(); ++k) {
            const int j = cand[k];
            if (cam_agg[j] >= 0) continue;
            if (best < 0 || score[j] > score[best] || (score[j] == score[best] && j < best)) best = j;
          }
          next = best;
        }
        for (int j : cand) score[j] = 0;
        agg_size.push_back(size);
      }
      // merge fragments (a one-camera aggregate has only 6 free motions for 7 coarse unknowns)
      const int n0 = (int)agg_size.size();
      std::vector<int> target(n0);
      for (int I = 0; I < n0; ++I) target[I] = I;
      if (n0 > 1) {
        // cameras of every first-pass aggregate (counting sort), so that a fragment only walks its own cameras and
        // the aggregates it touches: linear in the fragments' blocks, not fragments x cameras
        std::vector<int> aoff(n0 + 1, 0), acam(C), touched;
        for (int64_t c = 0; c < C; ++c) aoff[cam_agg[c] + 1]++;
        for (int I = 0; I < n0; ++I) aoff[I + 1] += aoff[I];
        {
          std::vector<int> afill(aoff.begin(), aoff.end() - 1);
          for (int64_t c = 0; c < C; ++c) acam[afill[cam_agg[c]]++] = (int)c;
        }
        std::vector<int> votes(n0, 0);
        for (int I = 0; I < n0; ++I) {
          if (agg_size[I] * 2 >= per) continue;
          touched.clear();
          for (int kc = aoff[I]; kc < aoff[I + 1]; ++kc) {
            const int c = acam[kc];
            for (int q = rowptr[c]; q < rowptr[c + 1]; ++q) {
              int J = cam_agg[col[q]];
              while (target[J] != J) J = target[J];
              if (J != I && votes[J]++ == 0) touched.push_back(J);
            }
          }
          int best = -1;   // most votes, ties to the lowest aggregate index (as a scan over all aggregates would pick)
          for (int J : touched) if (best < 0 || votes[J] > votes[best] || (votes[J] == votes[best] && J < best)) best = J;
          for (int J : touched) votes[J] = 0;
          if (best < 0 && agg_size[I] < 2) best = (I > 0) ? I - 1 : 1;   // isolated camera: any partner
          if (best >= 0) {
            while (target[best] != best) best = target[best];
            if (best != I) { target[I] = best; agg_size[best] += agg_size[I]; agg_size[I] = 0; }
          }
        }
      }
      std::vector<int> renum(n0, -1);
      int n_new = 0;
      for (int I = 0; I < n0; ++I) { int J = I; while (target[J] != J) J = target[J]; if (renum[J] < 0) renum[J] = n_new++; }
      for (int64_t c = 0; c < C; ++c) { int J = cam_agg[c]; while (target[J] != J) J = target[J]; cam_agg[c] = renum[J]; }
      n_agg = n_new;
    }
    if ((int64_t)n_agg * 7 > 3584) return;
    std::vector<int>&agg_off = hb.agg_off, &agg_cams = hb.agg_cams, &sp_rowptr = hb.sp_rowptr, &sp_col = hb.sp_col, &sp_of = hb.sp_of;
    agg_off.assign(n_agg + 1, 0); agg_cams.resize(C); sp_rowptr.assign(C + 1, 0); sp_of.resize(nnzb);
    for (int64_t c = 0; c < C; ++c) agg_off[cam_agg[c] + 1]++;
    for (int I = 0; I < n_agg; ++I) agg_off[I + 1] += agg_off[I];
    {
      std::vector<int> afill(agg_off.begin(), agg_off.end() - 1);
      for (int64_t c = 0; c < C; ++c) agg_cams[afill[cam_agg[c]]++] = (int)c;   // ascending camera index inside an aggregate
    }
    // row c of S P: the distinct aggregates of the row's cameras in ascending order.  A row touches a handful of
    // aggregates: they are collected through a marker array and insertion-sorted (no sort of the whole row, no search)
    std::vector<int> pos(n_agg, -1), rowJ;
    sp_col.reserve((size_t)C * 4);
    for (int64_t c = 0; c < C; ++c) {
      rowJ.clear();
      for (int q = rowptr[c]; q < rowptr[c + 1]; ++q) {
        const int J = cam_agg[col[q]];
        if (pos[J] < 0) {
          pos[J] = 0;
          size_t k = rowJ.size();
          rowJ.push_back(J);
          while (k > 0 && rowJ[k - 1] > J) { rowJ[k] = rowJ[k - 1]; --k; }
          rowJ[k] = J;
        }
      }
      const int base = (int)sp_col.size();
      for (size_t k = 0; k < rowJ.size(); ++k) pos[rowJ[k]] = base + (int)k;
      for (int q = rowptr[c]; q < rowptr[c + 1]; ++q) sp_of[q] = pos[cam_agg[col[q]]];
      for (int J : rowJ) pos[J] = -1;
      sp_col.insert(sp_col.end(), rowJ.begin(), rowJ.end());
      sp_rowptr[c + 1] = (int)sp_col.size();
    }
    hb.n_agg = n_agg;
    hb.tl = true;
  };
  struct Joiner {   // an early return must not destroy a running thread
    std::thread t;
    ~Joiner() { if (t.joinable()) t.join(); }
  } worker;
  if (n_blk >= 4096) worker.t = std::thread(host_bsr);   // small problems: starting a thread costs more than the work
  else host_bsr();
  // ---- 8. tile mode: pair slots = (tile, block) cells ordered inside each tile by decreasing entry count,
  //         per-block reduce lists, entries packed to tile-local 16-bit indices
  if (pb.schur_tiles) {
    DEV_ALLOC(pb.tile_ent_off, int, pb.n_st + 1);
    BA_CUDA(h, cudaMemsetAsync(pb.tile_ent_off, 0, (pb.n_st + 1) * sizeof(int), st));
    SlotList sl;
    if (int rc = build_slots(h, tmp, nruns, d_uniq, d_runs, CC, d_gkeys, n_blk, pb.n_st, &sl)) return rc;
    pb.n_slot = sl.n_slot; pb.slot_a = sl.slot_a; pb.slot_n = sl.slot_n;
    pb.tile_slot_off = sl.tile_off; pb.blk_slot_off = sl.grp_off; pb.blk_slots = sl.grp_slots;
    if (n_ent > 0) {
      k_tile_ent_off<<<nblk(pb.n_st + 1, T), T, 0, st>>>(pb.n_st, (int)n_ent, d_keys_sorted, CC, pb.tile_ent_off);
    }
  }
  mark("8 pair slots");
  // ---- 9. fused kernel: self slots = the observations of one camera inside one tile
  if (pb.fused) {
    DEV_ALLOC(pb.dent, unsigned short, N);
    int n_cell = 0;
    TMP_ALLOC(d_dkey, long long, N);
    TMP_ALLOC(d_dkey_sorted, long long, N);
    TMP_ALLOC(d_dval, int, N);
    TMP_ALLOC(d_dcell, long long, N + 1);
    TMP_ALLOC(d_dcnt, int, N + 1);
    if (N > 0) {
      k_self_keys<<<nblk(N, T), T, 0, st>>>(N, (long long)C, pb.cam, pb.st_off, pb.n_st, d_dkey);
      k_iota<<<nblk(N, T), T, 0, st>>>(N, d_iota);
      const double kmax = (double)C * (double)pb.n_st;
      int end_bit = 1;
      while (end_bit < 63 && (double)((long long)1 << end_bit) < kmax) ++end_bit;
      size_t bytes = 0;
      cub::DeviceRadixSort::SortPairs(nullptr, bytes, d_dkey, d_dkey_sorted, d_iota, d_dval, (int)N, 0, end_bit, st);
      TMP_ALLOC(d_sort5, char, bytes);
      BA_CUDA(h, cub::DeviceRadixSort::SortPairs(d_sort5, bytes, d_dkey, d_dkey_sorted, d_iota, d_dval, (int)N, 0, end_bit, st));
      TMP_ALLOC(d_nruns2, int, 1);
      bytes = 0;
      cub::DeviceRunLengthEncode::Encode(nullptr, bytes, d_dkey_sorted, d_dcell, d_dcnt, d_nruns2, (int)N, st);
      TMP_ALLOC(d_rle3, char, bytes);
      BA_CUDA(h, cub::DeviceRunLengthEncode::Encode(d_rle3, bytes, d_dkey_sorted, d_dcell, d_dcnt, d_nruns2, (int)N, st));
      BA_CUDA(h, cudaMemcpyAsync(&n_cell, d_nruns2, sizeof(int), cudaMemcpyDeviceToHost, st));
      BA_CUDA(h, cudaStreamSynchronize(st));
      k_pack_self<<<nblk(N, T), T, 0, st>>>(N, d_dval, pb.st_off, pb.n_st, pb.dent);
    }
    SlotList sl;
    if (int rc = build_slots(h, tmp, n_cell, d_dcell, d_dcnt, (long long)C, nullptr, (int)C, pb.n_st, &sl)) return rc;
    pb.n_dslot = sl.n_slot; pb.dslot_a = sl.slot_a; pb.dslot_n = sl.slot_n;
    pb.tile_dslot_off = sl.tile_off; pb.cam_dslot_off = sl.grp_off; pb.cam_dslots = sl.grp_slots;
    mark("9 self slots");
    {
      const char* e_bo = getenv("B200BA_BANK_ORDER");   // 0: keep the sorted entry order (A/B measurements)
      if (!(e_bo && atoi(e_bo) == 0) && pb.n_st > 0 && pb.schur_tiles) {
        const int g = (int)std::min<int64_t>(pb.n_st, 65535 * 16);
        if (n_ent > 0) k_bank_order<unsigned><<<g, 32, 0, st>>>(pb.n_st, pb.tile_slot_off, pb.slot_a, pb.slot_n, 0, pb.tent);
        if (N > 0) k_bank_order<unsigned short><<<g, 32, 0, st>>>(pb.n_st, pb.tile_dslot_off, pb.dslot_a, pb.dslot_n, 0, pb.dent);
        mark("9b bank-aware entry order");
      }
    }
    // ---- 10. dataflow kernel: per-tile info + contiguous metadata blobs (bulk-copied to shared memory per tile)
    const char* e_flow = getenv("B200BA_FLOW");
    if (pb.tile_obs <= FLOW_TILE && !(e_flow && atoi(e_flow) == 0) && pb.n_st > 0) {
      DEV_ALLOC(pb.tile_info, int4, 4 * (int64_t)pb.n_st);
      TMP_ALLOC(d_units, int, pb.n_st + 1);
      TMP_ALLOC(d_boff, int, pb.n_st + 1);
      BA_CUDA(h, cudaMemsetAsync(d_units, 0, (pb.n_st + 1) * sizeof(int), st));
      k_tile_info<<<nblk(pb.n_st, T), T, 0, st>>>(pb.n_st, (long long)C, pb.st_off, pb.st_wt_off, pb.tile_slot_off,
                                                  pb.tile_dslot_off, pb.tile_ent_off, pb.pt, n_cell, d_dcell, pb.tile_info, d_units);
      size_t bytes = 0;
      cub::DeviceScan::ExclusiveSum(nullptr, bytes, d_units, d_boff, pb.n_st + 1, st);
      TMP_ALLOC(d_scan6, char, bytes);
      BA_CUDA(h, cub::DeviceScan::ExclusiveSum(d_scan6, bytes, d_units, d_boff, pb.n_st + 1, st));
      int total_units = 0;
      BA_CUDA(h, cudaMemcpyAsync(&total_units, d_boff + pb.n_st, sizeof(int), cudaMemcpyDeviceToHost, st));
      BA_CUDA(h, cudaStreamSynchronize(st));
      DEV_ALLOC(pb.blob, uint4, (int64_t)total_units + 1);
      k_blob_fill<<<std::min(pb.n_st, 4096), 128, 0, st>>>(pb.n_st, d_boff, pb.tile_info, pb.slot_a, pb.slot_n, pb.dslot_a,
                                                         pb.dslot_n, pb.tile_ent_off, pb.tent, pb.dent, pb.blob);
      pb.flow = true;
    }
  }
  mark("10 dataflow blobs");
  // ---- 10b. results of the host worker (step 7)
  if (worker.t.joinable()) worker.t.join();
  mark("7 host BSR (wait for the worker)");
  pb.n_sch = (int)hb.sb.size();
  pb.nnzb = (int)hb.col.size();
  if (int rc = upload(h, &pb.blk_r, hb.br)) return rc;
  if (int rc = upload(h, &pb.blk_c, hb.bc)) return rc;
  if (int rc = upload(h, &pb.diag_blk, hb.diag)) return rc;
  if (int rc = upload(h, &pb.sch_blk, hb.sb)) return rc;
  if (int rc = upload(h, &pb.sch_a, hb.sa)) return rc;
  if (int rc = upload(h, &pb.sch_b, hb.se)) return rc;
  if (int rc = upload(h, &pb.blk_choff, hb.choff)) return rc;
  if (int rc = upload(h, &pb.bsr_rowptr, hb.rowptr)) return rc;
  if (int rc = upload(h, &pb.bsr_col, hb.col)) return rc;
  if (int rc = upload(h, &pb.bsr_src, hb.src)) return rc;
  if (hb.tl) {
    pb.n_agg = hb.n_agg; pb.nc = 7 * hb.n_agg; pb.ncp = (pb.nc + 63) / 64 * 64; pb.nnz_sp = (int)hb.sp_col.size();
    if (int rc = upload(h, &pb.cam_agg, hb.cam_agg)) return rc;
    if (int rc = upload(h, &pb.agg_off, hb.agg_off)) return rc;
    if (int rc = upload(h, &pb.agg_cams, hb.agg_cams)) return rc;
    if (int rc = upload(h, &pb.sp_rowptr, hb.sp_rowptr)) return rc;
    if (int rc = upload(h, &pb.sp_col, hb.sp_col)) return rc;
    if (int rc = upload(h, &pb.sp_of, hb.sp_of)) return rc;
    pb.tl = true;
  }
  mark("10b BSR uploads");
  if (N > 0) {
    if (!uv_copy.th.empty()) {   // pageable pixels: the worker threads finish the upload (host-synchronous)
      if (uv_copy.wait()) return fail(h, B200BA_CUDA, "upload of the pixel array failed");
    } else {
      BA_CUDA(h, cudaStreamWaitEvent(st, h->ev_b, 0));
    }
    k_gather_uv<<<nblk(N, T), T, 0, st>>>(N, pb.perm, d_uv, pb.u, pb.v);
  }
  mark("11 pixels");
  BA_CUDA(h, cudaStreamSynchronize(st));
  BA_CUDA(h, cudaGetLastError());
  h->st.schur_entries = pb.n_ent;
  h->st.n_tiles = pb.n_st;
  h->st.n_pair_slots = pb.n_slot;
  h->st.n_self_slots = pb.n_dslot;
  h->st.schur_blocks = pb.n_blk;
  return 0;
}

}  // namespace b200ba

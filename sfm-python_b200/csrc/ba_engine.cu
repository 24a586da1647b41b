// Handle lifetime, device memory, NCCL plumbing and the Levenberg-Marquardt loop of libb200ba.
//
// The LM loop is the replacement for scipy.optimize.least_squares(method='trf', x_scale='jac') as called at
// sfm_lite/sfm.py:266-268: Marquardt damping lambda*diag(J^T J) plays the role of x_scale='jac'
// (scipy _lsq/common.py compute_jac_scale), a step is accepted iff the cost decreases (trf.py inner loop),
// and termination uses scipy's tests and status codes (common.py check_termination).
#include <dlfcn.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include <nvtx3/nvToolsExt.h>

#include "ba_internal.h"

namespace b200ba {

// NVTX range of one LM phase (header-only NVTX 3: a no-op unless a profiler is attached)
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

int partial_capacity(const Handle* h);
int pcg_grid(Handle* h);

int fail(Handle* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}
int cuda_fail(Handle* h, cudaError_t e, const char* where) {
  std::string m = std::string("CUDA error: ") + cudaGetErrorString(e) + " at " + where;
  cudaGetLastError();
  return fail(h, e == cudaErrorMemoryAllocation ? B200BA_OOM : B200BA_CUDA, m);
}

// ------------------------------------------------------------------ threaded pageable <-> device copies
int BigCopy::wait() {
  for (std::thread& t : th)
    if (t.joinable()) t.join();
  th.clear();
  if (start) { cudaEventDestroy(start); start = nullptr; }
  int rc = 0;
  for (int r : rcs)
    if (r && !rc) rc = r;
  rcs.clear();
  return rc;
}

int big_copy_start(Handle* h, BigCopy& bc, void* dst, const void* src, size_t bytes, bool to_device, cudaStream_t st) {
  if (bytes == 0) return 0;
  const cudaMemcpyKind kind = to_device ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
  const void* host = to_device ? src : dst;
  constexpr size_t CH = (size_t)2 << 20, MIN_BYTES = (size_t)8 << 20;
  int want = 0;
  if (bytes >= MIN_BYTES) {
    cudaPointerAttributes at;
    const cudaError_t e = cudaPointerGetAttributes(&at, host);
    if (e != cudaSuccess) cudaGetLastError();
    if (e != cudaSuccess || at.type == cudaMemoryTypeUnregistered) {
      const char* env = getenv("B200BA_COPY_THREADS");   // 0: leave pageable copies to the driver
      const unsigned hw = std::thread::hardware_concurrency();
      want = env ? atoi(env) : (int)std::min<unsigned>(8, std::max<unsigned>(2, hw / 4));
      want = std::max(0, std::min(want, 16));
    }
  }
  if (want > 0) {
    const size_t need = (size_t)want * 2 * CH;
    if (h->xfer_bytes < need) {
      if (h->xfer) cudaFreeHost(h->xfer);
      h->xfer = nullptr;
      h->xfer_bytes = 0;
      if (cudaMallocHost((void**)&h->xfer, need) == cudaSuccess) h->xfer_bytes = need;
      else { cudaGetLastError(); want = 0; }
    }
  }
  if (want == 0) {
    BA_CUDA(h, cudaMemcpyAsync(dst, src, bytes, kind, st));
    return 0;
  }
  BA_CUDA(h, cudaEventCreateWithFlags(&bc.start, cudaEventDisableTiming));
  BA_CUDA(h, cudaEventRecord(bc.start, st));
  const size_t nch = (bytes + CH - 1) / CH;
  const int T = (int)std::min<size_t>((size_t)want, nch);
  bc.rcs.assign(T, 0);
  const int device = h->device;
  char* slots = h->xfer;
  cudaEvent_t start = bc.start;
  for (int t = 0; t < T; ++t) {
    int* rc = &bc.rcs[t];
    bc.th.emplace_back([=]() {
      auto bad = [&](cudaError_t e) { if (e != cudaSuccess) { cudaGetLastError(); *rc = B200BA_CUDA; return true; } return false; };
      if (bad(cudaSetDevice(device))) return;
      cudaStream_t s = nullptr;
      cudaEvent_t ev[2] = {nullptr, nullptr};
      if (bad(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking))) return;
      bool ok = !bad(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming)) &&
                !bad(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming)) && !bad(cudaStreamWaitEvent(s, start, 0));
      char* slot[2] = {slots + (size_t)(2 * t) * CH, slots + (size_t)(2 * t + 1) * CH};
      size_t prev_c = 0;
      bool have_prev = false;
      int k = 0;
      for (size_t c = (size_t)t; ok && c < nch; c += (size_t)T, ++k) {
        const size_t off = c * CH, len = std::min(CH, bytes - off);
        const int b = k & 1;
        if (to_device) {
          if (k >= 2 && bad(cudaEventSynchronize(ev[b]))) { ok = false; break; }   // the slot's previous DMA has read it
          memcpy(slot[b], (const char*)src + off, len);
          ok = !bad(cudaMemcpyAsync((char*)dst + off, slot[b], len, cudaMemcpyHostToDevice, s)) && !bad(cudaEventRecord(ev[b], s));
        } else {
          // slot b was drained two chunks ago (below): start this chunk's DMA, then drain the previous chunk
          ok = !bad(cudaMemcpyAsync(slot[b], (const char*)src + off, len, cudaMemcpyDeviceToHost, s)) && !bad(cudaEventRecord(ev[b], s));
          if (ok && have_prev) {
            const size_t poff = prev_c * CH, plen = std::min(CH, bytes - poff);
            if (bad(cudaEventSynchronize(ev[b ^ 1]))) { ok = false; break; }
            memcpy((char*)dst + poff, slot[b ^ 1], plen);
          }
          prev_c = c;
          have_prev = true;
        }
      }
      if (ok && !to_device && have_prev) {
        const int b = (k - 1) & 1;
        const size_t poff = prev_c * CH, plen = std::min(CH, bytes - poff);
        if (!bad(cudaEventSynchronize(ev[b]))) memcpy((char*)dst + poff, slot[b], plen);
      }
      bad(cudaStreamSynchronize(s));
      if (ev[0]) cudaEventDestroy(ev[0]);
      if (ev[1]) cudaEventDestroy(ev[1]);
      cudaStreamDestroy(s);
    });
  }
  return 0;
}

void* pool_alloc(Handle* h, size_t bytes) {
  void* p = nullptr;
  cudaError_t e = h->pool ? cudaMallocFromPoolAsync(&p, bytes, h->pool, h->stream) : cudaMalloc(&p, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

void pool_free(Handle* h, void* p) {
  if (!p) return;
  if (h->pool) cudaFreeAsync(p, h->stream);
  else cudaFree(p);
}

void* dalloc(Handle* h, size_t bytes) {
  bytes = std::max<size_t>(bytes, 256);
  void* p = pool_alloc(h, bytes);
  if (!p) {
    h->err = "out of device memory (allocation of " + std::to_string(bytes) + " bytes failed)";
    return nullptr;
  }
  h->allocs.push_back(p);
  h->bytes_device += (int64_t)bytes;
  return p;
}

void free_problem(Handle* h) {
  for (void* p : h->allocs) pool_free(h, p);
  h->allocs.clear();
  h->bytes_device = 0;
  h->pb = Problem();
  double* hs = h->w.h_scal;
  h->w = Work();
  h->w.h_scal = hs;
  h->has_problem = false;
}

// ------------------------------------------------------------------ NCCL through dlopen (torch ships libnccl.so.2)
struct Id128 { char b[128]; };
namespace {
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, Id128, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
constexpr int NCCL_INT64 = 4, NCCL_DOUBLE = 8, NCCL_SUM = 0, NCCL_MAX = 2;

bool nccl_load(std::string* err) {
  if (g_nccl.lib) return true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  void* lib = nullptr;
  for (const char* n : names) {
    lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (lib) break;
  }
  if (!lib) { *err = std::string("cannot dlopen libnccl.so.2: ") + dlerror(); return false; }
  g_nccl.GetUniqueId = (int (*)(void*))dlsym(lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(void**, int, Id128, int))dlsym(lib, "ncclCommInitRank");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(lib, "ncclAllReduce");
  g_nccl.AllGather = (int (*)(const void*, void*, size_t, int, void*, cudaStream_t))dlsym(lib, "ncclAllGather");
  g_nccl.CommDestroy = (int (*)(void*))dlsym(lib, "ncclCommDestroy");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce || !g_nccl.AllGather) {
    *err = "libnccl.so.2 lacks a required symbol";
    return false;
  }
  g_nccl.lib = lib;
  return true;
}
int nccl_fail(Handle* h, int rc, const char* where) {
  return fail(h, B200BA_NCCL, std::string("NCCL error: ") +
                                  (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?") + " at " + where);
}
}  // namespace

int comm_allreduce_sum(Handle* h, double* buf, int64_t count) {
  if (h->nranks <= 1 || count <= 0) return 0;
  int rc = g_nccl.AllReduce(buf, buf, (size_t)count, NCCL_DOUBLE, NCCL_SUM, h->comm, h->stream);
  return rc ? nccl_fail(h, rc, "ncclAllReduce(sum)") : 0;
}
int comm_allreduce_max(Handle* h, double* buf, int64_t count) {
  if (h->nranks <= 1 || count <= 0) return 0;
  int rc = g_nccl.AllReduce(buf, buf, (size_t)count, NCCL_DOUBLE, NCCL_MAX, h->comm, h->stream);
  return rc ? nccl_fail(h, rc, "ncclAllReduce(max)") : 0;
}
int comm_allgather_i64(Handle* h, const long long* send, long long* recv, int64_t count) {
  if (h->nranks <= 1) {
    cudaMemcpyAsync(recv, send, count * sizeof(long long), cudaMemcpyDeviceToDevice, h->stream);
    return 0;
  }
  int rc = g_nccl.AllGather(send, recv, (size_t)count, NCCL_INT64, h->comm, h->stream);
  return rc ? nccl_fail(h, rc, "ncclAllGather") : 0;
}

int comm_allgather_f64(Handle* h, const double* send, double* recv, int64_t count) {
  if (h->nranks <= 1) {
    cudaMemcpyAsync(recv, send, count * sizeof(double), cudaMemcpyDeviceToDevice, h->stream);
    return 0;
  }
  int rc = g_nccl.AllGather(send, recv, (size_t)count, NCCL_DOUBLE, h->comm, h->stream);
  return rc ? nccl_fail(h, rc, "ncclAllGather") : 0;
}

// ------------------------------------------------------------------ exchange through peer memory (see struct P2P)
namespace {
struct PeerPtrs { double* p[P2P_MAX_RANKS]; unsigned long long* f[P2P_MAX_RANKS]; };

__global__ void __launch_bounds__(256) k_p2p_push(const double2* __restrict__ src, long long n2, int rank, int nranks,
                                                  PeerPtrs peers, long long slot_off, unsigned long long epoch,
                                                  unsigned int* __restrict__ done) {
  // four 16-byte loads in flight per thread; the peer loop starts at the next rank so that at any moment the ranks
  // write to different destinations
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i0 < n2; i0 += 4 * stride) {
    double2 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = (i0 + k * stride < n2) ? src[i0 + k * stride] : make_double2(0.0, 0.0);
    for (int q = 1; q <= nranks; ++q) {
      const int p = (rank + q) % nranks;
      double2* dst = reinterpret_cast<double2*>(peers.p[p] + slot_off);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (i0 + k * stride < n2) dst[i0 + k * stride] = v[k];
    }
  }
  __threadfence_system();   // this thread's stores are visible system-wide before the CTA is counted as done
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(done, 1u);
    if (prev == gridDim.x - 1) {   // last CTA: every CTA's data is out -> publish the epoch on every rank
      *done = 0;
      __threadfence_system();
      for (int p = 0; p < nranks; ++p)
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peers.f[p] + rank), "l"(epoch) : "memory");
    }
  }
}

__global__ void k_p2p_wait(const unsigned long long* __restrict__ flags, int nranks, unsigned long long epoch,
                           int* __restrict__ err) {
  const int p = threadIdx.x;
  if (p >= nranks) return;
  const long long t0 = clock64();
  unsigned long long v = 0;
  while (true) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flags + p) : "memory");
    if (v >= epoch) break;
    if (clock64() - t0 > 20000000000ll) { *err = 1 + p; break; }   // ~10 s: report instead of hanging the device
    __nanosleep(200);
  }
}
}  // namespace

void p2p_release(Handle* h) {
  P2P& q = h->p2p;
  for (int r = 0; r < P2P_MAX_RANKS; ++r)
    if (q.peer[r] && q.peer[r] != q.base && !q.direct) cudaIpcCloseMemHandle(q.peer[r]);
  if (q.base) cudaFree(q.base);
  if (q.done) cudaFree(q.done);
  cudaGetLastError();
  q = P2P();
}

int p2p_setup(Handle* h) {
  P2P& q = h->p2p;
  Problem& pb = h->pb;
  const char* e = getenv("B200BA_P2P");   // 0: always NCCL all-gather
  const bool want = h->nranks > 1 && h->nranks <= P2P_MAX_RANKS && pb.xchg && !(e && atoi(e) == 0);
  if (!want) { q.on = false; return 0; }   // the same decision on every rank (xchg and nranks are global)
  const size_t need = 512 + (size_t)2 * h->nranks * (size_t)pb.x_stride * sizeof(double);
  if (q.base && q.bytes >= need) { q.on = true; return 0; }   // x_stride is global, so every rank takes this branch together
  // (re)allocate on every rank, exchange the IPC handles through the communicator
  cudaStreamSynchronize(h->stream);
  p2p_release(h);
  long long ok = 1;
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (cudaMalloc(&q.base, need) != cudaSuccess || cudaMalloc((void**)&q.done, 256) != cudaSuccess ||
      cudaMemset(q.base, 0, need) != cudaSuccess || cudaMemset(q.done, 0, 256) != cudaSuccess ||
      (!h->in_process && cudaIpcGetMemHandle(&mine, q.base) != cudaSuccess)) {
    cudaGetLastError();
    ok = 0;
  }
  q.bytes = need;
  q.direct = h->in_process;
  if (q.direct) {
    // every rank is a handle of this process: the 64-byte slot carries {buffer address, device ordinal} instead of an
    // IPC handle (an IPC handle cannot be opened by the process that exported it); peer access is enabled below
    memset(&mine, 0, sizeof(mine));
    long long v[2] = {(long long)(uintptr_t)q.base, (long long)h->device};
    memcpy(&mine, v, sizeof(v));
  }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  const int per = 9;   // 8 longs of handle + 1 status
  long long* d = (long long*)pool_alloc(h, (size_t)per * (h->nranks + 1) * sizeof(long long));
  if (!d) return fail(h, B200BA_OOM, "out of device memory");
  long long send[per];
  memcpy(send, &mine, 64);
  send[8] = ok;
  BA_CUDA(h, cudaMemcpyAsync(d + (size_t)per * h->nranks, send, sizeof(send), cudaMemcpyHostToDevice, h->stream));
  if (int rc = comm_allgather_i64(h, d + (size_t)per * h->nranks, d, per)) { pool_free(h, d); return rc; }
  std::vector<long long> all((size_t)per * h->nranks);
  BA_CUDA(h, cudaMemcpyAsync(all.data(), d, all.size() * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
  BA_CUDA(h, cudaStreamSynchronize(h->stream));
  for (int r = 0; r < h->nranks; ++r) ok = ok && all[(size_t)per * r + 8];
  if (ok) {
    for (int r = 0; r < h->nranks; ++r) {
      if (r == h->rank) { q.peer[r] = q.base; continue; }
      cudaIpcMemHandle_t hh;
      memcpy(&hh, &all[(size_t)per * r], 64);
      if (q.direct) {
        long long v[2];
        memcpy(v, &hh, sizeof(v));
        int can = 0;
        cudaError_t e = cudaDeviceCanAccessPeer(&can, h->device, (int)v[1]);
        if (e == cudaSuccess && can) {
          e = cudaDeviceEnablePeerAccess((int)v[1], 0);
          if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
        }
        if (e != cudaSuccess || !can) { cudaGetLastError(); q.peer[r] = nullptr; ok = 0; }
        else q.peer[r] = (void*)(uintptr_t)v[0];
        continue;
      }
      if (cudaIpcOpenMemHandle(&q.peer[r], hh, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError();
        q.peer[r] = nullptr;
        ok = 0;
      }
    }
  }
  // second round: everybody must have opened everybody, otherwise every rank falls back to NCCL together
  long long mine_ok = ok;
  BA_CUDA(h, cudaMemcpyAsync(d + h->nranks, &mine_ok, sizeof(long long), cudaMemcpyHostToDevice, h->stream));
  if (int rc = comm_allgather_i64(h, d + h->nranks, d, 1)) { pool_free(h, d); return rc; }
  BA_CUDA(h, cudaMemcpyAsync(all.data(), d, h->nranks * sizeof(long long), cudaMemcpyDeviceToHost, h->stream));
  BA_CUDA(h, cudaStreamSynchronize(h->stream));
  pool_free(h, d);
  for (int r = 0; r < h->nranks; ++r) ok = ok && all[r];
  if (getenv("B200BA_SETUP_PROF") && atoi(getenv("B200BA_SETUP_PROF")))
    fprintf(stderr, "[b200ba p2p] rank %d/%d: %s\n", h->rank, h->nranks,
            !ok ? "peer mapping failed -> NCCL all-gather" : (q.direct ? "direct peer pointers (one process)" : "CUDA IPC"));
  if (!ok) { p2p_release(h); return 0; }
  q.epoch = 0;
  q.on = true;
  return 0;
}

// Receive slot of this rank for the exchange that the next p2p_exchange call performs, in every rank's buffer.
PeerDst p2p_next_dst(Handle* h) {
  PeerDst d;
  P2P& q = h->p2p;
  if (!q.on) return d;
  const unsigned long long e = q.epoch + 1;
  const long long slots = 64 + (long long)(e & 1) * h->nranks * h->pb.x_stride;
  for (int k = 1; k <= h->nranks; ++k) {
    const int r = (h->rank + k) % h->nranks;
    d.p[k - 1] = (double*)q.peer[r] + slots + (long long)h->rank * h->pb.x_stride;
  }
  d.n = h->nranks;
  return d;
}

int p2p_exchange(Handle* h) {
  P2P& q = h->p2p;
  Problem& pb = h->pb;
  Work& w = h->w;
  ++q.epoch;
  const long long parity = (long long)(q.epoch & 1);
  const long long slots = 64 + parity * h->nranks * pb.x_stride;        // doubles from the base of a buffer
  PeerPtrs pp{};
  for (int r = 0; r < h->nranks; ++r) { pp.p[r] = (double*)q.peer[r]; pp.f[r] = (unsigned long long*)q.peer[r]; }
  // the blocks of S were stored into the peers' slots by the slot reduction itself (launch_linearize, PeerDst); what is
  // pushed here is the tail [b | g_c | diag U | scalars] -- and the epoch flag, which covers both (the reduction kernel
  // has completed before this one starts: stream order)
  const long long head = h->direct_push ? ((long long)pb.xl_max * 36) : 0;   // xl_max * 36 is even
  const long long n2 = (pb.x_stride - head) / 2;
  const int grid = (int)std::max<long long>(1, std::min<long long>((n2 + 1023) / 1024, (long long)h->sm_count * 8));
  k_p2p_push<<<grid, 256, 0, h->stream>>>(reinterpret_cast<const double2*>(w.xsend + head), n2, h->rank, h->nranks, pp,
                                          slots + (long long)h->rank * pb.x_stride + head, q.epoch, q.done);
  k_p2p_wait<<<1, 32, 0, h->stream>>>((const unsigned long long*)q.base, h->nranks, q.epoch,
                                      reinterpret_cast<int*>(w.scal + S_P2P_ERR));
  h->st.launches += 2;
  w.xrecv = (double*)q.base + slots;
  BA_CUDA(h, cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------ per-problem work buffers
static int alloc_work(Handle* h) {
  Problem& pb = h->pb;
  Work& w = h->w;
  const int64_t C = pb.C, P = pb.P, N = pb.N, nx = 6 * C + 3 * P;
#define WALLOC(ptr, count)                                                                    \
  ptr = (double*)dalloc(h, (size_t)std::max<int64_t>((count), 1) * sizeof(double));           \
  if (!ptr) return B200BA_OOM
  WALLOC(w.x, nx + 2);
  WALLOC(w.xn, nx + 2);
  WALLOC(w.camprep, (C + 2) * CAMPREP_STRIDE);   // field-major [21][Cs], Cs <= C + 1
  WALLOC(w.camprep_n, (C + 2) * CAMPREP_STRIDE);
  WALLOC(w.cstep, (C + 2) * 6);
  // Y: 18 N doubles for the unfused pipeline (per-observation W L^-T blocks); the fused kernels keep those records in
  // shared memory, so only a 2 N staging area (residuals / reprojection errors in the caller's order) is held and the
  // Jacobian-block export grows it on demand (ensure_Y).  rr / cpart belong to the unfused pipeline only.
  w.Y_count = pb.fused ? std::max<int64_t>(2 * N, 16 * 16 * (int64_t)h->sm_count) : 18 * N;
  WALLOC(w.Y, w.Y_count);
  if (!pb.fused) { WALLOC(w.rr, N * 4); }
  WALLOC(w.pp, P * 12);
  WALLOC(w.gp, P * 3);
  if (!pb.fused) { WALLOC(w.cpart, (int64_t)pb.n_cch * 33); }
  WALLOC(w.spart, (int64_t)std::max(pb.n_sch, pb.n_slot) * 36);
  WALLOC(w.U, C * 21);
  WALLOC(w.dpart, (int64_t)pb.n_dslot * 40);
  WALLOC(w.camsum, C * 40);
  w.comm_count = (int64_t)pb.n_blk * 36 + 18 * C + S_NSUM;
  WALLOC(w.comm, w.comm_count);
  w.Sup = w.comm;
  w.b = w.Sup + (int64_t)pb.n_blk * 36;
  w.gc = w.b + 6 * C;
  w.udiag = w.gc + 6 * C;
  w.sums = w.udiag + 6 * C;
  if (pb.xchg) {
    WALLOC(w.xsend, pb.x_stride);
    WALLOC(w.xrecv_nccl, pb.x_stride * h->nranks);
    w.xrecv = w.xrecv_nccl;
    BA_CUDA(h, cudaMemsetAsync(w.xsend, 0, (size_t)pb.x_stride * sizeof(double), h->stream));
  }
  WALLOC(w.scal, NSCAL);
  w.n_part = partial_capacity(h);
  WALLOC(w.part, (int64_t)w.n_part * 8);
  WALLOC(w.dc, 6 * C);
  WALLOC(w.pcg, 6 * C * 5 + C * 36);
  WALLOC(w.pcg_part, (int64_t)2 * 4 * std::max(pcg_grid(h), 1));
  if (C <= DENSE_RESIDENT_CAM) { WALLOC(w.dense, 36 * C * C); }
  if (pb.tl) {
    WALLOC(w.tl_G, C * 42);
    WALLOC(w.tl_SP, (int64_t)pb.nnz_sp * 42);
    WALLOC(w.tl_A, (int64_t)pb.ncp * pb.ncp);
    WALLOC(w.tl_cb, (int64_t)pb.ncp * 32);
    WALLOC(w.tl_rb, (int64_t)pb.ncp * 32);
    WALLOC(w.tl_dsc, pb.ncp);
    WALLOC(w.tl_cen, (int64_t)pb.n_agg * 3);
    WALLOC(w.tl_rc, pb.ncp);
    WALLOC(w.tl_zc, pb.ncp);
    WALLOC(w.tl_uf, 6 * C);
    WALLOC(w.tl_prof, 8);
  }
  h->tl_built = false; h->tl_age = 0; h->tl_ref_iters = 0; h->tl_last_iters = 0;
#undef WALLOC
  cudaStream_t st = h->stream;
  BA_CUDA(h, cudaMemsetAsync(w.pp, 0, (size_t)std::max<int64_t>(P * 12, 1) * sizeof(double), st));
  BA_CUDA(h, cudaMemsetAsync(w.gp, 0, (size_t)std::max<int64_t>(P * 3, 1) * sizeof(double), st));
  BA_CUDA(h, cudaMemsetAsync(w.comm, 0, (size_t)w.comm_count * sizeof(double), st));
  BA_CUDA(h, cudaMemsetAsync(w.scal, 0, NSCAL * sizeof(double), st));
  BA_CUDA(h, cudaMemsetAsync(w.dc, 0, (size_t)std::max<int64_t>(6 * C, 1) * sizeof(double), st));
  if (pb.tl) {   // the padding entries nc .. ncp-1 stay zero (the coarse mat-vec runs over whole padded rows)
    BA_CUDA(h, cudaMemsetAsync(w.tl_rc, 0, (size_t)pb.ncp * sizeof(double), st));
    BA_CUDA(h, cudaMemsetAsync(w.tl_zc, 0, (size_t)pb.ncp * sizeof(double), st));
    BA_CUDA(h, cudaMemsetAsync(w.tl_prof, 0, 8 * sizeof(double), st));
  }
  BA_CUDA(h, cudaStreamSynchronize(st));
  return 0;
}

int ensure_Y(Handle* h, int64_t count) {
  Work& w = h->w;
  if (w.Y_count >= count) return 0;
  double* p = (double*)dalloc(h, (size_t)std::max<int64_t>(count, 1) * sizeof(double));
  if (!p) return B200BA_OOM;
  w.Y = p;   // the smaller block stays on the handle's list until the problem is released
  w.Y_count = count;
  return 0;
}

// ------------------------------------------------------------------ phase timing (only when profiling is on)
struct PhaseTimer {
  Handle* h; double* acc; bool on;
  PhaseTimer(Handle* h_, double* acc_) : h(h_), acc(acc_), on(h_->profile) {
    if (on) cudaEventRecord(h->ev0, h->stream);
  }
  ~PhaseTimer() {
    if (!on) return;
    cudaEventRecord(h->ev1, h->stream);
    cudaEventSynchronize(h->ev1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    *acc += ms;
  }
};

// linearise at w.x (camprep must match) and reduce across ranks; leaves Sup, b, gc, udiag, sums ready
int linearize_all(Handle* h, double lambda) {
  NvtxRange nv("b200ba:linearize+schur+exchange");
  Work& w = h->w;
  launch_linearize(h, w.x, lambda);
  {
    PhaseTimer t(h, &h->st.ms_allreduce);
    if (h->pb.xchg) {
      if (h->p2p.on) {
        if (int rc = p2p_exchange(h)) return rc;
      } else {
        w.xrecv = w.xrecv_nccl;
        if (int rc = comm_allgather_f64(h, w.xsend, w.xrecv, h->pb.x_stride)) return rc;
      }
      launch_xchg_sum(h);
    } else {
      if (int rc = comm_allreduce_sum(h, w.comm, w.comm_count)) return rc;
      if (int rc = comm_allreduce_max(h, w.scal + S_GPMAX, 1)) return rc;
    }
  }
  launch_fix_diag(h);
  launch_fix_cameras(h);
  return 0;
}

int camera_solve(Handle* h, int solver, double rtol, int max_iter) {
  NvtxRange nv("b200ba:camera_solve");
  PhaseTimer t(h, &h->st.ms_camera_solve);
  if (solver == B200BA_SOLVER_AUTO) solver = (h->pb.C <= 64) ? B200BA_SOLVER_CHOLESKY : B200BA_SOLVER_PCG;
  if (solver == B200BA_SOLVER_CHOLESKY) return launch_cholesky(h);
  // Two-level preconditioner: starts once block-Jacobi needed more than TL_START iterations (far from the optimum the
  // right-hand side has little weight in the weak modes and block-Jacobi converges in a handful); the coarse
  // operator is rebuilt when it is TL_MAX_AGE solves old or the iteration count has tripled since it was built
  // (measured: a coarse operator that is several LM iterations and a factor 30 in lambda old preconditions as well
  // as a fresh one; the count grows with shrinking lambda and the tighter forcing term, not with staleness).
  // The decisions depend only on iteration counts, which are identical on every rank.
  constexpr int TL_START = 24, TL_MAX_AGE = 12;
  if (h->pb.tl && h->tl_mode != 0 && (h->tl_built || h->tl_last_iters > TL_START || h->tl_mode == 2)) {
    const bool rebuild = h->tl_mode == 2 || !h->tl_built || h->tl_age >= TL_MAX_AGE ||
                         (h->tl_age >= 2 && h->tl_last_iters > 3 * h->tl_ref_iters + 30);
    if (int rc = launch_tl_build(h, rebuild)) return rc;
    if (rebuild) { h->tl_built = true; h->tl_age = 0; h->st.n_tl_builds += 1; }
    h->tl_age += 1;
    return launch_pcg2(h, rtol, max_iter);
  }
  return launch_pcg(h, rtol, max_iter);
}

// bookkeeping of the two-level policy after a camera solve (iters = PCG iterations it took)
void tl_note_iters(Handle* h, int iters) {
  h->tl_last_iters = iters;
  if (h->tl_built && h->tl_age == 1) h->tl_ref_iters = iters;
}

// trial point: cameras x + dc -> xn, camprep_n; points by back-substitution; sums all-reduced
int trial_step(Handle* h, double lambda) {
  NvtxRange nv("b200ba:backsub+trial_cost");
  Work& w = h->w;
  PhaseTimer t(h, &h->st.ms_backsub_trial);
  launch_cam_prep(h, w.x, w.dc, w.xn, w.camprep_n, w.camprep, w.cstep);
  launch_backsub_trial(h, lambda);
  if (int rc = comm_allreduce_sum(h, w.sums + 1, S_NSUM - 1)) return rc;  // slot 0 (cost at x) is already global
  launch_cam_dots(h);
  return 0;
}

int fetch_scalars(Handle* h) {
  NvtxRange nv("b200ba:fetch_scalars(host sync)");
  Work& w = h->w;
  BA_CUDA(h, cudaMemcpyAsync(w.h_scal, w.scal, NSCAL * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  BA_CUDA(h, cudaMemcpyAsync(w.h_scal + NSCAL, w.sums, S_NSUM * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  BA_CUDA(h, cudaStreamSynchronize(h->stream));
  int flow_err = 0;
  memcpy(&flow_err, w.h_scal + S_FLOW_ERR, sizeof(int));
  int p2p_err = 0;
  memcpy(&p2p_err, w.h_scal + S_P2P_ERR, sizeof(int));
  // the error words are reported once and cleared, so one transient timeout does not fail every later call on the handle
  if (flow_err || p2p_err) cudaMemsetAsync(w.scal + S_P2P_ERR, 0, 2 * sizeof(double), h->stream);
  if (flow_err) return fail(h, B200BA_CUDA, "k_lin_flow: a dependency wait timed out (results invalid)");
  if (p2p_err) return fail(h, B200BA_NCCL, "peer-memory exchange: the contribution of rank " + std::to_string(p2p_err - 1) +
                                           " did not arrive (results invalid)");
  return 0;
}

// test hooks and stats: the error word of the dataflow kernel (set instead of hanging the device)
int check_flow_error(Handle* h) {
  if (!h->has_problem || !h->pb.flow) return 0;
  int flow_err = 0;
  BA_CUDA(h, cudaMemcpyAsync(&flow_err, h->w.scal + S_FLOW_ERR, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  BA_CUDA(h, cudaStreamSynchronize(h->stream));
  if (flow_err) {
    cudaMemsetAsync(h->w.scal + S_FLOW_ERR, 0, sizeof(double), h->stream);   // reported once
    return fail(h, B200BA_CUDA, "k_lin_flow: a dependency wait timed out (results invalid)");
  }
  return 0;
}

}  // namespace b200ba

#include "ba_lm.inc"
#include "ba_capi.inc"
#include "ba_multi.inc"

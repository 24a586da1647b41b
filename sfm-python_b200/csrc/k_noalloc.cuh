// Allocation-free residual entry (included at the end of ba_kernels.cu, namespace b200ba).
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the residual + cost of
// sfm_lite/bundle_adjustment.py:35-42 (the k_cam_prep + k_residual pair behind b200ba_residuals) with ALL scratch --
// the per-camera prep table and the per-CTA cost partials -- in one file-scope device array of fixed size, so the call
// makes no cudaMalloc / cudaFree / pool allocation and needs no b200ba_set_problem.  Every GPU that loads the library
// holds the array (NOALLOC_SCRATCH_DOUBLES * 8 bytes = 0.77 MB) until the process exits.
//
// The array is shared by every call on the device and keeps the last call's contents.  Nothing in it has to start at
// zero: each call rewrites the camera rows it reads and every one of the NOALLOC_GRID partial slots before the final
// reduction reads them.  Calls that overlap on the GPU (two streams, two handles on one device) race on it.

constexpr int NOALLOC_MAX_CAM = 4096;                                  // cameras per call
constexpr int NOALLOC_GRID = 592;                                      // CTAs of the residual kernel (4 x 148 SMs)
constexpr int NOALLOC_CAMPREP = CAMPREP_FIELDS * NOALLOC_MAX_CAM;      // field-major [21][4096]
constexpr int NOALLOC_SCRATCH_DOUBLES = NOALLOC_CAMPREP + 1024;        // + partial slots (592 used)

static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];

namespace {

// residual in the CALLER's observation order from raw index arrays (no Problem, no sort); out-of-range indices give NaN
__global__ void __launch_bounds__(256) k_residual_noalloc(int64_t N, int C, int64_t P, Intrinsics K,
                                                          const int* __restrict__ cam, const int* __restrict__ pt,
                                                          const double* __restrict__ uv, const double* __restrict__ x,
                                                          const double* __restrict__ camprep, double* __restrict__ r_out,
                                                          double* __restrict__ part) {
  __shared__ double sm[32];
  const double* xp = x + 6 * (int64_t)C;
  double acc = 0.0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += stride) {
    const int c = cam[i], p = pt[i];
    const double2 o = reinterpret_cast<const double2*>(uv)[i];
    double r0, r1;
    if (c < 0 || c >= C || p < 0 || p >= P) {
      r0 = r1 = nan("");
    } else {
      double cp[CAMPREP_FIELDS];
      load_rt(camprep, NOALLOC_MAX_CAM, c, cp);
      project_residual(cp, K, xp[3 * (int64_t)p], xp[3 * (int64_t)p + 1], xp[3 * (int64_t)p + 2], o.x, o.y, r0, r1);
    }
    acc += r0 * r0 + r1 * r1;
    if (r_out) reinterpret_cast<double2*>(r_out)[i] = make_double2(r0, r1);
  }
  block_sum<1>(&acc, sm);
  if (threadIdx.x == 0) part[blockIdx.x] = acc;   // every slot is written on every call
}

// cost = 0.5 * sum of the partials, fixed order (one CTA)
__global__ void __launch_bounds__(256) k_noalloc_cost(int n, const double* __restrict__ part, double* __restrict__ cost_out) {
  __shared__ double sm[32];
  double acc = 0.0;
  for (int b = threadIdx.x; b < n; b += blockDim.x) acc += part[b];
  block_sum<1>(&acc, sm);
  if (threadIdx.x == 0) cost_out[0] = 0.5 * acc;
}

double* g_noalloc_ptr[MAX_DEVICES] = {};   // device address of g_noalloc_scratch per ordinal (host code cannot use the name)

}  // namespace

// Set-up, once per device, with that device current: resolve the device address of the scratch array.
int noalloc_setup(Handle* h) {
  if (h->device < 0 || h->device >= MAX_DEVICES) return fail(h, B200BA_BAD_ARG, "device ordinal beyond the per-device table");
  if (g_noalloc_ptr[h->device]) return 0;
  double* p = nullptr;
  cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
  if (e != cudaSuccess || !p) return cuda_fail(h, e != cudaSuccess ? e : cudaErrorInvalidSymbol, "cudaGetSymbolAddress(g_noalloc_scratch)");
  g_noalloc_ptr[h->device] = p;
  return 0;
}

int residuals_noalloc(Handle* h, int64_t C, int64_t P, int64_t N, const int* cam_dev, const int* pt_dev, const double* uv_dev,
                      const double* K, const double* x_dev, double* r_out_dev, double* cost_out_dev) {
  if (C < 0 || P < 0 || N < 0 || !K || !x_dev || (N > 0 && (!cam_dev || !pt_dev || !uv_dev)))
    return fail(h, B200BA_BAD_ARG, "null pointer or negative size");
  if (C > NOALLOC_MAX_CAM)
    return fail(h, B200BA_BAD_ARG, "b200ba_residuals_noalloc: at most " + std::to_string(NOALLOC_MAX_CAM) +
                                       " cameras fit the static scratch (got " + std::to_string(C) + "); use b200ba_residuals");
  if (N >= ((int64_t)1 << 40) || P >= ((int64_t)1 << 31)) return fail(h, B200BA_BAD_ARG, "problem too large");
  if (int rc = noalloc_setup(h)) return rc;
  double* scratch = g_noalloc_ptr[h->device];
  double* camprep = scratch;
  double* part = scratch + NOALLOC_CAMPREP;
  Intrinsics Kd;
  for (int i = 0; i < 9; ++i) Kd.k[i] = K[i];
  cudaStream_t st = h->stream;
  if (C > 0)
    k_cam_prep<<<cdiv(C, 128), 128, 0, st>>>((int)C, (int64_t)NOALLOC_MAX_CAM, x_dev, nullptr, nullptr, camprep, nullptr, nullptr);
  k_residual_noalloc<<<NOALLOC_GRID, 256, 0, st>>>(N, (int)C, P, Kd, cam_dev, pt_dev, uv_dev, x_dev, camprep, r_out_dev, part);
  if (cost_out_dev) k_noalloc_cost<<<1, 256, 0, st>>>(NOALLOC_GRID, part, cost_out_dev);
  h->st.launches += 2 + (cost_out_dev ? 1 : 0);
  BA_CUDA(h, cudaGetLastError());
  BA_CUDA(h, cudaStreamSynchronize(st));   // outputs are valid on return, like every other entry point
  return 0;
}

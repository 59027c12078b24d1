// mt_gemm_static_scratch: mt_gemm's decode-orientation split-K GEMM for callers that cannot allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// mt_gemm takes its split-K scratch (fp32 partial slabs [splits][M][N]) from the caller (mt_gemm_args.workspace).  Here the
// scratch is a file-scope device array of MT_GEMM_STATIC_SCRATCH_BYTES that every GPU loading the library holds until the
// process exits.  This entry point calls no cudaMalloc / cudaFree / allocating library routine: it resolves the array's
// device address once per device (cudaGetSymbolAddress, with that device current) and runs the same kernels as mt_gemm
// (mt::linear -> gemm_dec_kernel + splitk_reduce_kernel) with that address as the workspace.
//
// In HOST code the name `g_splitk_scratch` is a host-side stand-in for the symbol, not a device address: it is only ever
// handed to cudaGetSymbolAddress, never to a kernel.
//
// Sharing: every call on a device uses the same array, and the array keeps the previous call's contents.  Nothing has to
// start at zero: every (split, m < M, n < N) element the reduce kernel reads is written by this call's GEMM kernel first
// (stream order), so no clearing is needed.  Calls that OVERLAP on the GPU (different streams, or graphs replayed
// concurrently) race on the array; callers must serialise them on one stream or with events.
#include "gemm.cuh"

#include <mutex>

namespace {

static __device__ __align__(256) unsigned char g_splitk_scratch[MT_GEMM_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
float* g_scratch_ptr[kMaxDevices] = {};      // device address of g_splitk_scratch on each device (nullptr = not resolved yet)
std::mutex g_scratch_mutex;

// Device address of the scratch array on the CURRENT device, resolved once per device.
int scratch_address(float** out) {
  int dev = -1;
  MT_CUDA(cudaGetDevice(&dev));
  MT_CHECK(dev >= 0 && dev < kMaxDevices, "mt_gemm_static_scratch: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    MT_CUDA(cudaGetSymbolAddress(&p, g_splitk_scratch));      // the symbol, not `&g_splitk_scratch`: see the header comment
    MT_CHECK(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "mt_gemm_static_scratch: bad scratch address");
    g_scratch_ptr[dev] = reinterpret_cast<float*>(p);
  }
  *out = g_scratch_ptr[dev];
  return MT_OK;
}

}  // namespace

using namespace mt;

extern "C" int mt_gemm_static_scratch(const mt_gemm_args* a, mt_stream s) {
  MT_CHECK(a, "mt_gemm_static_scratch: null args");
  MT_CHECK(a->dtype_in == MT_DTYPE_BF16 || a->dtype_in == MT_DTYPE_F32 || a->dtype_in == MT_DTYPE_F16, "mt_gemm_static_scratch: bad dtype_in");
  MT_CHECK(a->dtype_out == MT_DTYPE_BF16 || a->dtype_out == MT_DTYPE_F32 || a->dtype_out == MT_DTYPE_F16, "mt_gemm_static_scratch: bad dtype_out");
  MT_CHECK(a->x && a->w && a->out, "mt_gemm_static_scratch: null operand");
  MT_CHECK(a->M > 0 && a->N > 0 && a->K > 0, "mt_gemm_static_scratch: empty problem M=%d N=%d K=%d", a->M, a->N, a->K);
  MT_CHECK(a->workspace == nullptr && a->out_splits == nullptr,
           "mt_gemm_static_scratch: workspace and out_splits must be NULL (the scratch is the library's own and the result is always the finished tensor)");
  MT_CHECK(a->M <= 256, "mt_gemm_static_scratch: M=%d exceeds 256 (only the decode-orientation GEMM uses split-K scratch; use mt_gemm)", a->M);
  const long long slab = static_cast<long long>(a->M) * a->N * static_cast<long long>(sizeof(float));
  MT_CHECK(2 * slab <= static_cast<long long>(MT_GEMM_STATIC_SCRATCH_BYTES),
           "mt_gemm_static_scratch: M*N = %d*%d needs %lld bytes for a two-way split, more than the static scratch of %d bytes "
           "(MT_GEMM_STATIC_SCRATCH_BYTES); use mt_gemm with a caller workspace",
           a->M, a->N, 2 * slab, MT_GEMM_STATIC_SCRATCH_BYTES);
  float* scratch = nullptr;
  MT_TRY(scratch_address(&scratch));
  LinearPlan p;
  p.dtype_in = a->dtype_in; p.dtype_out = a->dtype_out; p.M = a->M; p.N = a->N; p.K = a->K;
  p.x = a->x; p.ldx = a->ldx; p.w = a->w; p.ldw = a->ldw; p.out = a->out; p.ldo = a->ldo;
  p.bias = a->bias; p.act = a->act; p.residual = a->residual; p.ldr = a->ldr; p.scale = a->scale;
  p.round_bf16 = a->round_bf16;
  p.workspace = scratch; p.workspace_bytes = MT_GEMM_STATIC_SCRATCH_BYTES;
  p.impl = a->impl & MT_GEMM_SIMT;
  p.w_static = (a->impl & MT_GEMM_W_STATIC) != 0;
  p.w_tiled = (a->impl & MT_GEMM_W_TILED) != 0;
  return linear(p, reinterpret_cast<cudaStream_t>(s));
}

// extern "C" surface that is not tied to an engine handle (see include/mosstts.h).
#include "gemm.cuh"
#include "sampling.cuh"

namespace mt { const char* last_error(); void set_pdl(int on); void set_trace(void* buf, long long words); long long trace_used(); void set_attn_cluster(int cs); }
using namespace mt;

extern "C" int mt_version(void) { return 100; }
extern "C" const char* mt_last_error(void) { return mt::last_error(); }
extern "C" long long mt_launch_count(void) { return g_launches; }
extern "C" int mt_debug_force_simt(int on) { g_force_simt = on != 0; return MT_OK; }

extern "C" int mt_gemm_debug_set_pair(int mode) { mt::g_gemm_pair_mode = mode; return MT_OK; }

extern "C" int mt_debug_set_pdl(int on) { mt::set_pdl(on); return MT_OK; }
extern "C" int mt_debug_set_attn_cluster(int cs) {
  if (cs != 0 && cs != 1 && cs != 2 && cs != 4 && cs != 8) return MT_ERR_ARG;
  mt::set_attn_cluster(cs);
  return MT_OK;
}
extern "C" long long mt_debug_set_trace(void* buf, long long words) { const long long used = mt::trace_used(); mt::set_trace(buf, words); return used; }

extern "C" int mt_gemm(const mt_gemm_args* a, mt_stream s) {
  MT_CHECK(a, "mt_gemm: null args");
  MT_CHECK(a->dtype_in == MT_DTYPE_BF16 || a->dtype_in == MT_DTYPE_F32 || a->dtype_in == MT_DTYPE_F16, "mt_gemm: bad dtype_in");
  MT_CHECK(a->dtype_out == MT_DTYPE_BF16 || a->dtype_out == MT_DTYPE_F32 || a->dtype_out == MT_DTYPE_F16, "mt_gemm: bad dtype_out");
  LinearPlan p;
  p.dtype_in = a->dtype_in; p.dtype_out = a->dtype_out; p.M = a->M; p.N = a->N; p.K = a->K;
  p.x = a->x; p.ldx = a->ldx; p.w = a->w; p.ldw = a->ldw; p.out = a->out; p.ldo = a->ldo;
  p.bias = a->bias; p.act = a->act; p.residual = a->residual; p.ldr = a->ldr; p.scale = a->scale;
  p.round_bf16 = a->round_bf16; p.workspace = a->workspace; p.workspace_bytes = a->workspace_bytes;
  p.impl = a->impl & MT_GEMM_SIMT;
  p.out_splits = a->out_splits;
  MT_CHECK(!a->out_splits || (a->M <= 256 && a->workspace), "mt_gemm: out_splits needs M <= 256 and a workspace");
  p.w_static = (a->impl & MT_GEMM_W_STATIC) != 0;
  p.w_tiled = (a->impl & MT_GEMM_W_TILED) != 0;
  return linear(p, reinterpret_cast<cudaStream_t>(s));
}

extern "C" int mt_sample_tokens(const void* logits, long long ld, int N, int V, float temperature,
                                const uint32_t* prev_bitmap, float repetition_penalty, float top_p, int top_k,
                                int do_sample, const float* uniforms, int64_t* out, mt_stream s) {
  MT_CHECK(logits && out, "mt_sample_tokens: null argument");
  return sample_tokens(reinterpret_cast<const bf16*>(logits), ld, N, V, temperature, prev_bitmap, repetition_penalty,
                       top_p, top_k, do_sample, uniforms, out, reinterpret_cast<cudaStream_t>(s));
}
extern "C" long long mt_sample_full_scratch_bytes(int N) { return static_cast<long long>(N) * static_cast<long long>(sample_full_scratch_bytes()); }
extern "C" int mt_sample_tokens_full(const void* logits, long long ld, int N, int V, float temperature,
                                     const uint32_t* prev_bitmap, float repetition_penalty, float top_p, int top_k,
                                     int do_sample, const float* uniforms, int64_t* out, void* scratch, long long scratch_bytes,
                                     mt_stream s) {
  MT_CHECK(logits && out, "mt_sample_tokens_full: null argument");
  return sample_tokens(reinterpret_cast<const bf16*>(logits), ld, N, V, temperature, prev_bitmap, repetition_penalty,
                       top_p, top_k, do_sample, uniforms, out, reinterpret_cast<cudaStream_t>(s), scratch, scratch_bytes);
}
extern "C" int mt_find_last_equal(const int64_t* t, int B, int S, long long ld, long long value, int64_t* out,
                                  mt_stream s) {
  MT_CHECK(t && out, "mt_find_last_equal: null argument");
  return find_last_equal(t, B, S, ld, value, out, reinterpret_cast<cudaStream_t>(s));
}
extern "C" int mt_apply_delay_pattern(const int64_t* codes, int T, int n, long long pad_code, int64_t* out, mt_stream s) {
  return apply_delay_pattern(codes, T, n, pad_code, out, reinterpret_cast<cudaStream_t>(s));
}
extern "C" int mt_apply_de_delay_pattern(const int64_t* delayed, int Td, int n, int64_t* out, mt_stream s) {
  return apply_de_delay_pattern(delayed, Td, n, out, reinterpret_cast<cudaStream_t>(s));
}

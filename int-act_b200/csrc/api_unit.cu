// C-ABI entry points for single kernels (unit/parity tests call these through ctypes).
// See include/pi0b200.h for the contract of every function.
#include "../../include/pi0b200.h"
#include "elementwise.cuh"

using namespace pi0;

extern "C" {

int pi0_abi_version(void) { return PI0_ABI_VERSION; }
const char* pi0_last_error(void) { return get_error(); }

long long pi0_launch_count(int reset) {
  long long v = g_launch_count;
  if (reset) g_launch_count = 0;
  return v;
}

static int problem_from_desc(const pi0_gemm_desc* d, GemmProblem& g) {
  PI0_REQUIRE(d != nullptr, "pi0_gemm: null descriptor");
  g = gemm_problem(d->A, d->lda, d->W, d->ldw, d->M, d->N, d->K);
  g.a_z0 = d->a_z0; g.a_z1 = d->a_z1; g.w_z0 = d->w_z0; g.w_z1 = d->w_z1;
  g.nz0 = d->nz0 > 0 ? d->nz0 : 1;
  g.nz1 = d->nz1 > 0 ? d->nz1 : 1;
  g.epi_mode = d->epi_mode;
  g.bn = d->bn;
  g.splits = d->splits;
  EpiParams& e = g.epi;
  e.out = d->out; e.ldc = d->ldc; e.out_z0 = d->out_z0; e.out_z1 = d->out_z1;
  e.bias = d->bias; e.res = d->res; e.ldr = d->ldr; e.res_row_mod = d->res_row_mod;
  e.alpha = d->alpha; e.post_div = d->post_div; e.post_mul = d->post_mul; e.act = d->act;
  e.rpb_in = d->rpb_in; e.rpb_out = d->rpb_out; e.row_off = d->row_off;
  if (d->vt != nullptr) {
    e.vt = d->vt; e.vt_col0 = d->vt_col0; e.vt_hd = d->vt_hd; e.vt_heads = d->vt_heads;
    e.vt_pitch = d->vt_pitch;
  }
  e.vt_rpb = d->vt_rpb > 0 ? d->vt_rpb : 1;
  e.vt_toff = d->vt_toff;
  e.rope_cos = d->rope_cos; e.rope_sin = d->rope_sin; e.pos_ids = d->pos_ids;
  e.kcache = d->kcache; e.n_q_heads = d->n_q_heads; e.kv_pitch = d->kv_pitch;
  e.dt = d->dt;
  e.rowsq_out = d->rowsq_out; e.rowsq_ld = d->rowsq_ld;
  g.norm_gain = d->norm_gain; g.norm_rowsq = d->norm_rowsq; g.norm_slots = d->norm_slots;
  g.norm_eps = d->norm_eps != 0.0f ? d->norm_eps : 1e-6f;
  return 0;
}

int pi0_gemm(const pi0_gemm_desc* d, int dtype, int impl, int num_sms, void* stream) {
  GemmProblem g;
  PI0_TRY(problem_from_desc(d, g));
  return gemm_launch(g, dtype, impl, num_sms, (cudaStream_t)stream);
}

int pi0_gemm_f32x3(const pi0_gemm_desc* d, void* scratch_a, size_t bytes_a, void* scratch_w, size_t bytes_w, void* scratch_p,
                   size_t bytes_p, int num_sms, void* stream) {
  GemmProblem g;
  PI0_TRY(problem_from_desc(d, g));
  const int rc = gemm_x3_launch(g, scratch_a, bytes_a, scratch_w, bytes_w, scratch_p, bytes_p, num_sms > 0 ? num_sms : 148,
                                (cudaStream_t)stream);
  PI0_REQUIRE(rc != 1, "pi0_gemm_f32x3: the problem does not fit the scratch buffers (A %zu, W %zu bytes needed) or N=%d is "
              "not a multiple of 8", gemm_x3_operand_bytes((int64_t)g.nz0 * g.nz1, g.M, g.K),
              gemm_x3_operand_bytes((int64_t)g.nz0 * g.nz1, g.N, g.K), g.N);
  return rc;
}

size_t pi0_gemm_f32x3_scratch_bytes(int64_t nz, int64_t rows, int K) { return gemm_x3_operand_bytes(nz, rows, K); }

int pi0_rmsnorm(const void* x, void* y, const float* g, int M, int D, int dtype, void* stream) {
  if (dtype == 1) return rmsnorm<float>((const float*)x, (float*)y, g, M, D, 1e-6f, 0, 0, 0, (cudaStream_t)stream);
  return rmsnorm<bf16>((const bf16*)x, (bf16*)y, g, M, D, 1e-6f, 0, 0, 0, (cudaStream_t)stream);
}

int pi0_layernorm(const void* x, void* y, const float* w, const float* b, int M, int D, int dtype, void* stream) {
  if (dtype == 1) return layernorm<float>((const float*)x, (float*)y, w, b, M, D, 1e-6f, (cudaStream_t)stream);
  return layernorm<bf16>((const bf16*)x, (bf16*)y, w, b, M, D, 1e-6f, (cudaStream_t)stream);
}

int pi0_softmax(void* scores, int64_t rows_total, int rows_per_b, int64_t pitch, int n_keys, int n_prefix,
                const uint8_t* pad, int pad_pitch, const int32_t* cum, int heads, int dtype, void* stream) {
  if (dtype == 1)
    return softmax_rows<float>((float*)scores, rows_total, rows_per_b, pitch, n_keys, n_prefix, pad, pad_pitch, cum,
                               heads, (cudaStream_t)stream);
  return softmax_rows<bf16>((bf16*)scores, rows_total, rows_per_b, pitch, n_keys, n_prefix, pad, pad_pitch, cum, heads,
                            (cudaStream_t)stream);
}

int pi0_decode_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                         const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch, int num_sms, void* stream) {
  return decode_attention(q, kcache, vtcache, att, pad, cum, B, S, P, kv_pitch, vt_pitch, num_sms, (cudaStream_t)stream);
}

int pi0_prefill_attention(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, int B,
                          int T, int P, int kv_pitch, int vt_pitch, int num_sms, void* stream) {
  return prefill_attention(q, kcache, vtcache, att, pad, B, T, P, kv_pitch, vt_pitch, num_sms > 0 ? num_sms : 148,
                           (cudaStream_t)stream);
}

int pi0_attention_long(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad, const int32_t* cum,
                       int B, int T, int P, int n_suffix, int kv_pitch, int vt_pitch, int num_sms, void* stream) {
  return attention_long(q, kcache, vtcache, att, pad, cum, B, T, P, n_suffix, kv_pitch, vt_pitch, num_sms > 0 ? num_sms : 148,
                        (cudaStream_t)stream);
}

int pi0_siglip_attention(const void* qkv, void* att, int B, void* stream) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return siglip_attention(qkv, att, B, sms, (cudaStream_t)stream);
}

int pi0_decode_attention_batch(const void* q, const void* kcache, const void* vtcache, void* att, const uint8_t* pad,
                               const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch, void* stream) {
  return decode_attention_batch(q, kcache, vtcache, att, pad, cum, B, S, P, kv_pitch, vt_pitch, (cudaStream_t)stream);
}

int pi0_decode_attention_split(const void* q, const void* kcache, const void* vtcache, void* att, void* scratch,
                               const uint8_t* pad, const int32_t* cum, int B, int S, int P, int kv_pitch, int vt_pitch,
                               void* stream) {
  return decode_attention_split(q, kcache, vtcache, att, scratch, pad, cum, B, S, P, kv_pitch, vt_pitch,
                                (cudaStream_t)stream);
}


int pi0_position_ids(const uint8_t* pad, int B, int P, int S, int32_t* prefix_pos, int32_t* suffix_pos, void* stream) {
  return position_ids(pad, B, P, S, prefix_pos, suffix_pos, (cudaStream_t)stream);
}

}  // extern "C"

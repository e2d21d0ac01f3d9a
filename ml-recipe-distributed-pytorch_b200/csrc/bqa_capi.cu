// bqa_capi.cu — extern "C" surface of libbqa (declared in include/bqa.h).
#include "../../include/bqa.h"

#include "bqa_host.cuh"
#include "gemm_sm100.cuh"

using namespace bqa;

extern "C" {

const char* bqa_last_error(void) { return bqa::last_error(); }

int bqa_abi_version(void) { return 1; }

int bqa_gemm_bf16(const bqa_gemm_desc* d, void* stream) {
  if (d == nullptr) {
    set_error("bqa_gemm_bf16: null descriptor");
    return 1;
  }
  int num_sms = 0;
  if (device_sm_count(&num_sms)) return 1;
  GemmProblem p{};
  p.M = d->M;
  p.N = d->N;
  p.K = d->K;
  p.A = d->A;
  p.lda = d->lda;
  p.a_mn_major = d->a_mn_major;
  p.B = d->B;
  p.ldb = d->ldb;
  p.b_mn_major = d->b_mn_major;
  p.C = d->C;
  p.ldc = d->ldc;
  p.C2 = d->C2;
  p.epilogue = d->epilogue;
  p.bias = d->bias;
  p.residual = d->residual;
  p.ld_res = d->ld_res;
  p.aux = d->aux;
  p.ld_aux = d->ld_aux;
  p.k_splits = d->k_splits;
  p.drop = make_drop_key(d->drop_seed, d->drop_site, d->dropout_p);
  GemmPlan plan;
  if (gemm_plan(&plan, p, num_sms)) return 1;
  return gemm_launch(plan, reinterpret_cast<cudaStream_t>(stream));
}

}  // extern "C"

// extern "C" surface of libhydrogen_b200.so (declared in include/hydrogen_b200.h).
#include "../../include/hydrogen_b200.h"
#include "hyd_internal.h"

namespace hyd {
const char* last_error();
int probe_peak(int kind, int iters, int ctas_per_sm, double* tflops);
}  // namespace hyd

extern "C" {

const char* hyd_last_error(void) { return hyd::last_error(); }
int hyd_version(void) { return 100; }

int hyd_lgemm_f64(int tA, int tB, int64_t m, int64_t n, int64_t k, double alpha, const double* A,
                  int64_t lda, const double* B, int64_t ldb, double beta, double* C, int64_t ldc,
                  hyd_stream_t stream) {
    return hyd::lgemm_f64(tA, tB, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc, (cudaStream_t)stream);
}

int hyd_probe_peak(int kind, int iters, int ctas_per_sm, double* tflops) {
    return hyd::probe_peak(kind, iters, ctas_per_sm, tflops);
}

}  // extern "C"

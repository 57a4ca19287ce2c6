/* TEST INFRASTRUCTURE ONLY -- extended-precision arbiter for the theta-step loop.
 *
 * NOT IN THE REFERENCE.  Same recurrence as the reference's time loop
 *     rhs = (I + k(1-theta) alpha M) w + beta ;  w = spsolve(A, rhs)      [fvm_classes.py:408-411]
 * on the reference's own fp64 operators A and B (oracle/fvm_oracle.py: theta_operators,
 * fvm_classes.py:376-377), but with a no-pivot Thomas solve carried in __float128 (or long
 * double), so that its result is the exact solution of those operators to ~1e-30.  It decides who
 * is right when SuperLU and the GPU differ by O(eps * cond), and it is the C twin of
 * fvm_oracle.theta_run_thomas (numpy longdouble), which is too slow for 10^4-step runs.
 *
 * Built by oracle/Makefile into oracle/_build/libthomas_arbiter.so; loaded with ctypes by
 * fvm_oracle.theta_run_arbiter.  Only tests/ and tests/golden/make_golden.py use it. */
#include <stdint.h>
#include <stdlib.h>


#define DEFINE_RUN(NAME, T)                                                                        \
  int NAME(int64_t J, int64_t nsteps, const double* alo, const double* adi, const double* aup,     \
           const double* blo, const double* bdi, const double* bup, const double* beta,            \
           const double* w0, int64_t nsnap, const int64_t* snap_steps, double* out) {              \
    T* m = malloc(sizeof(T) * J); T* inv = malloc(sizeof(T) * J);                                  \
    T* w = malloc(sizeof(T) * J); T* r = malloc(sizeof(T) * J);                                    \
    if (!m || !inv || !w || !r) return 1;                                                          \
    inv[0] = 1 / (T)adi[0]; m[0] = 0;                                                              \
    for (int64_t j = 1; j < J; ++j) {                                                              \
      m[j] = alo[j] * inv[j - 1];                                                                  \
      inv[j] = 1 / ((T)adi[j] - m[j] * aup[j - 1]);                                                \
    }                                                                                              \
    for (int64_t j = 0; j < J; ++j) w[j] = w0[j];                                                  \
    int64_t k = 0;                                                                                 \
    for (int64_t n = 1; n <= nsteps; ++n) {                                                        \
      for (int64_t j = 0; j < J; ++j) {                                                            \
        T s = (T)bdi[j] * w[j] + beta[j];                                                          \
        if (j > 0) s += (T)blo[j] * w[j - 1];                                                      \
        if (j < J - 1) s += (T)bup[j] * w[j + 1];                                                  \
        r[j] = s;                                                                                  \
      }                                                                                            \
      for (int64_t j = 1; j < J; ++j) r[j] -= m[j] * r[j - 1];                                     \
      w[J - 1] = r[J - 1] * inv[J - 1];                                                            \
      for (int64_t j = J - 2; j >= 0; --j) w[j] = (r[j] - aup[j] * w[j + 1]) * inv[j];             \
      if (k < nsnap && snap_steps[k] == n) {                                                       \
        for (int64_t j = 0; j < J; ++j) out[k * J + j] = (double)w[j];                             \
        ++k;                                                                                       \
      }                                                                                            \
    }                                                                                              \
    free(m); free(inv); free(w); free(r);                                                          \
    return 0;                                                                                      \
  }

DEFINE_RUN(arbiter_run_quad, __float128)
DEFINE_RUN(arbiter_run_longdouble, long double)

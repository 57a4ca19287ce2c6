// On-device diagnostics of a batch: extrema of the cell Peclet number and of the CFL number (the
// quantities behind the reference's warnings, fvm_classes.py:103-114,135-139, printed by the demo
// at :379-380) and the conserved "area" sum_j w_j h_j the demo prints every step (:420-421).
// One warp per problem, strided over the cells, shuffle reduction.
#include "fvm_device.cuh"

namespace fvm {

__global__ void __launch_bounds__(256) diagnostics_kernel(fvm_problem pr, const double* __restrict__ dt,
                                                          const double* __restrict__ w, int64_t w_ps, int64_t w_cs,
                                                          double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= pr.P) return;
  double pe_lo = INFINITY, pe_hi = -INFINITY, cf_lo = INFINITY, cf_hi = -INFINITY, mass = 0.0;
  const double k = dt ? dt[p] : 0.0;
  for (int64_t j = lane; j < pr.J; j += 32) {
    const double h = sub(pr.faces[j + 1], pr.faces[j]);
    const double a = pr.a[p * pr.a_ps + j * pr.a_cs];
    const double d = pr.d[p * pr.d_ps + j * pr.d_cs];
    const double mu = dvd(mul(a, h), d);           // peclet_number() on the unmodified d, :103,135-136
    const double c = dvd(mul(a, k), h);            // CFL_condition(), :138-139
    pe_lo = fmin(pe_lo, mu); pe_hi = fmax(pe_hi, mu);
    cf_lo = fmin(cf_lo, c); cf_hi = fmax(cf_hi, c);
    if (w) mass = fma(w[p * w_ps + j * w_cs], h, mass);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    pe_lo = fmin(pe_lo, __shfl_xor_sync(0xffffffffu, pe_lo, s));
    pe_hi = fmax(pe_hi, __shfl_xor_sync(0xffffffffu, pe_hi, s));
    cf_lo = fmin(cf_lo, __shfl_xor_sync(0xffffffffu, cf_lo, s));
    cf_hi = fmax(cf_hi, __shfl_xor_sync(0xffffffffu, cf_hi, s));
    mass += __shfl_xor_sync(0xffffffffu, mass, s);
  }
  if (lane == 0) {
    double* o = out + p * 5;
    o[0] = pe_lo; o[1] = pe_hi; o[2] = cf_lo; o[3] = cf_hi; o[4] = mass;
  }
}

}  // namespace fvm

extern "C" int fvm_diagnostics(const fvm_problem* prob, const double* dt, const double* w, int64_t w_ps,
                               int64_t w_cs, double* out, void* stream) {
  if (!prob || !prob->faces || !prob->a || !prob->d || !out) return FVM_ERR_INVALID_ARG;
  if (prob->J < 1 || prob->P < 1) return FVM_ERR_INVALID_ARG;
  const int64_t blocks = (prob->P + 7) / 8;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  fvm::diagnostics_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(*prob, dt, w, w_ps, w_cs, out);
  return (int)cudaGetLastError();
}

// Mesh validation on the device, for meshes too large for the host-side check of Mesh.__init__
// (len(faces) != len(set(faces)), fvm_classes.py:36-37, hashes every face: infeasible at 2^30).
// counts[0] = cells of zero width (adjacent duplicate faces: what the reference rejects for a
// sorted mesh), counts[1] = cells of negative width (unsorted faces: the reference accepts them
// silently); counts are accumulated, so clear them first.
namespace fvm {
__global__ void __launch_bounds__(256) mesh_check_kernel(const double* __restrict__ faces, int64_t n_cells,
                                                         unsigned long long* __restrict__ counts) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  unsigned long long zero = 0, neg = 0;
  for (int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n_cells; j += stride) {
    const double h = sub(faces[j + 1], faces[j]);
    zero += (h == 0.0);
    neg += (h < 0.0);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    zero += __shfl_xor_sync(0xffffffffu, zero, s);
    neg += __shfl_xor_sync(0xffffffffu, neg, s);
  }
  if ((threadIdx.x & 31) == 0) {
    if (zero) atomicAdd(&counts[0], zero);
    if (neg) atomicAdd(&counts[1], neg);
  }
}
}  // namespace fvm

extern "C" int fvm_mesh_check(const double* faces, int64_t n_cells, uint64_t* counts, void* stream) {
  if (!faces || !counts || n_cells < 1) return FVM_ERR_INVALID_ARG;
  int64_t blocks = (n_cells + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  fvm::mesh_check_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(faces, n_cells,
                                                                             (unsigned long long*)counts);
  return (int)cudaGetLastError();
}

// Distance between two states of a batch on the same mesh: the measure behind steady-state
// detection (w_n against w_{n-m}) and behind the comparison with an analytic solution that the
// example driver plots (fvm_examples.py:71).  out[p*4 + 0..3] = max_j |a_j - b_j|,
// sum_j |a_j - b_j| h_j, max_j |b_j|, sum_j b_j h_j.  One warp per problem; the result is
// deterministic (fixed strided order, shuffle tree).
namespace fvm {
__global__ void __launch_bounds__(256) state_compare_kernel(int64_t J, int64_t P, const double* __restrict__ faces,
                                                            const double* __restrict__ a, int64_t a_ps, int64_t a_cs,
                                                            const double* __restrict__ b, int64_t b_ps, int64_t b_cs,
                                                            double* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int64_t p = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= P) return;
  double dmax = 0.0, dl1 = 0.0, bmax = 0.0, bmass = 0.0;
  bool bad = false;
  for (int64_t j = lane; j < J; j += 32) {
    const double h = sub(faces[j + 1], faces[j]);
    const double av = a[p * a_ps + j * a_cs], bv = b[p * b_ps + j * b_cs];
    const double e = fabs(sub(av, bv));
    bad |= !(e == e);                              // fmax would drop a NaN: carry it explicitly
    dmax = fmax(dmax, e);
    dl1 = fma(e, h, dl1);
    bmax = fmax(bmax, fabs(bv));
    bmass = fma(bv, h, bmass);
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, s));
    dl1 += __shfl_xor_sync(0xffffffffu, dl1, s);
    bmax = fmax(bmax, __shfl_xor_sync(0xffffffffu, bmax, s));
    bmass += __shfl_xor_sync(0xffffffffu, bmass, s);
  }
  bad = __any_sync(0xffffffffu, bad);
  if (lane == 0) {
    double* o = out + p * 4;
    o[0] = bad ? NAN : dmax; o[1] = dl1; o[2] = bmax; o[3] = bmass;
  }
}
}  // namespace fvm

extern "C" int fvm_state_compare(int64_t J, int64_t P, const double* faces, const double* a, int64_t a_ps,
                                 int64_t a_cs, const double* b, int64_t b_ps, int64_t b_cs, double* out,
                                 void* stream) {
  if (!faces || !a || !b || !out || J < 1 || P < 1) return FVM_ERR_INVALID_ARG;
  const int64_t blocks = (P + 7) / 8;
  if (blocks > 0x7fffffffLL) return FVM_ERR_UNSUPPORTED;
  fvm::state_compare_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(J, P, faces, a, a_ps, a_cs, b, b_ps,
                                                                                b_cs, out);
  return (int)cudaGetLastError();
}

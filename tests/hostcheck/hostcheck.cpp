// TEST HARNESS ONLY: compiles the product's __host__ __device__ element math
// (vcnf_b200/csrc/*_math.cuh) for the CPU so that the formulas can be checked against the oracle
// in the GPU-less build container.  Never linked into libvcnf_b200.so.
#include <cstdint>
#include "../../vcnf_b200/csrc/rqs_math.cuh"
#include "../../vcnf_b200/csrc/rqs_fast.cuh"

using namespace vcnf;

template <typename T, int KMAX>
static void run_fwd(const T* prm, const T* v, const T* bound, int64_t n, int P, T scale,
                    RqsElemCfg c, T* out, T* lad, int32_t* bin, int32_t* flag) {
  for (int64_t i = 0; i < n; ++i) {
    int b = -1, f = 0;
    rqs_element<T, KMAX>(prm + i * P, v[i], bound[i], scale, c, out + i, lad + i, &b, &f);
    bin[i] = b;
    *flag |= f;
  }
}

template <typename T, int KMAX>
static void run_bwd(const T* prm, const T* v, const T* bound, int64_t n, int P, T scale,
                    RqsElemCfg c, const T* gout, const T* glad, T* gv, T* gprm) {
  for (int64_t i = 0; i < n; ++i)
    rqs_element_bwd<T, KMAX>(prm + i * P, v[i], bound[i], scale, c, gout[i], glad[i], gv + i, gprm + i * P);
}


// batch-shared parameters: one feature's table, n inputs
template <typename T, int KMAX>
static void run_table_fwd(const T* prm, const T* v, T bound, int64_t n, RqsElemCfg c, T* out, T* lad,
                          int32_t* bin, int32_t* flag) {
  T tab[3 * (KMAX + 1)];
  rqs_build_table<T, KMAX>(prm, c.K, c.mode, bound, tab);
  for (int64_t i = 0; i < n; ++i) {
    int b = -1, f = 0;
    rqs_table_element<T, KMAX>(tab, v[i], bound, c, out + i, lad + i, &b, &f);
    bin[i] = b;
    *flag |= f;
  }
}

template <typename T, int KMAX>
static void run_table_bwd(const T* prm, const T* v, T bound, int64_t n, RqsElemCfg c, const T* gout,
                          const T* glad, T* gv, T* gprm) {
  T tab[3 * (KMAX + 1)], gtab[3 * (KMAX + 1)];
  rqs_build_table<T, KMAX>(prm, c.K, c.mode, bound, tab);
  for (int i = 0; i < 3 * (KMAX + 1); ++i) gtab[i] = T(0);
  for (int64_t i = 0; i < n; ++i)
    rqs_table_element_bwd<T, KMAX>(tab, v[i], bound, c, gout[i], glad[i], gv + i, gtab);
  rqs_table_to_raw_grad<T, KMAX>(prm, c.K, c.mode, bound, gtab, gprm);
}

#define DISPATCH(T, call)                                        \
  if (K <= 8) { call(T, 8); } else if (K <= 16) { call(T, 16); } else { call(T, 32); }

extern "C" {

int hostcheck_rqs_fwd(const void* prm, const void* v, const void* bound, int64_t n, int K, int mode,
                      int inverse, int outside_zero, double scale, int is_double,
                      void* out, void* lad, int32_t* bin) {
  RqsElemCfg c{K, mode, inverse, outside_zero};
  const int P = 2 * K + rqs_num_derivs(K, mode);
  int32_t flag = 0;
#define CALL(T, KM) run_fwd<T, KM>((const T*)prm, (const T*)v, (const T*)bound, n, P, (T)scale, c, (T*)out, (T*)lad, bin, &flag)
  if (is_double) { DISPATCH(double, CALL) } else { DISPATCH(float, CALL) }
#undef CALL
  return flag;
}

int hostcheck_rqs_bwd(const void* prm, const void* v, const void* bound, int64_t n, int K, int mode,
                      int inverse, int outside_zero, double scale, int is_double,
                      const void* gout, const void* glad, void* gv, void* gprm) {
  RqsElemCfg c{K, mode, inverse, outside_zero};
  const int P = 2 * K + rqs_num_derivs(K, mode);
#define CALL(T, KM) run_bwd<T, KM>((const T*)prm, (const T*)v, (const T*)bound, n, P, (T)scale, c, (const T*)gout, (const T*)glad, (T*)gv, (T*)gprm)
  if (is_double) { DISPATCH(double, CALL) } else { DISPATCH(float, CALL) }
#undef CALL
  return 0;
}
int hostcheck_rqs_table_fwd(const void* prm, const void* v, double bound, int64_t n, int K, int mode,
                            int inverse, int outside_zero, int is_double, void* out, void* lad, int32_t* bin) {
  RqsElemCfg c{K, mode, inverse, outside_zero};
  int32_t flag = 0;
#define CALL(T, KM) run_table_fwd<T, KM>((const T*)prm, (const T*)v, (T)bound, n, c, (T*)out, (T*)lad, bin, &flag)
  if (is_double) { DISPATCH(double, CALL) } else { DISPATCH(float, CALL) }
#undef CALL
  return flag;
}

int hostcheck_rqs_table_bwd(const void* prm, const void* v, double bound, int64_t n, int K, int mode,
                            int inverse, int outside_zero, int is_double, const void* gout, const void* glad,
                            void* gv, void* gprm) {
  RqsElemCfg c{K, mode, inverse, outside_zero};
#define CALL(T, KM) run_table_bwd<T, KM>((const T*)prm, (const T*)v, (T)bound, n, c, (const T*)gout, (const T*)glad, (T*)gv, (T*)gprm)
  if (is_double) { DISPATCH(double, CALL) } else { DISPATCH(float, CALL) }
#undef CALL
  return 0;
}

// ---- fp32 / K = 8 fast path (rqs_fast.cuh; the MUFU approximations are libm calls on the host) ----
int hostcheck_rqs_fast_fwd(const float* prm, const float* v, const float* bound, int64_t n, int mode, int inverse,
                           int outside_zero, double scale, float* out, float* lad, int32_t* bin) {
  int flag = 0;
  const int P = 16 + rqs_num_derivs(8, mode);
  for (int64_t i = 0; i < n; ++i) {
    int b = -1;
    if (inverse) fast::element<8, true>(prm + i * P, v[i], bound[i], (float)scale, mode, outside_zero, out + i, lad + i, &b, &flag);
    else fast::element<8, false>(prm + i * P, v[i], bound[i], (float)scale, mode, outside_zero, out + i, lad + i, &b, &flag);
    bin[i] = b;
  }
  return flag;
}
int hostcheck_rqs_fast_bwd(const float* prm, const float* v, const float* bound, int64_t n, int mode, int inverse,
                           int outside_zero, double scale, const float* gout, const float* glad, float* gv,
                           float* gprm) {
  const int P = 16 + rqs_num_derivs(8, mode);
  for (int64_t i = 0; i < n; ++i) {
    if (inverse) fast::element_bwd<8, true>(prm + i * P, v[i], bound[i], (float)scale, mode, outside_zero, gout[i], glad[i], gv + i, gprm + i * P);
    else fast::element_bwd<8, false>(prm + i * P, v[i], bound[i], (float)scale, mode, outside_zero, gout[i], glad[i], gv + i, gprm + i * P);
  }
  return 0;
}
int hostcheck_rqs_fast_table_fwd(const float* prm, const float* v, double bound, int64_t n, int mode, int inverse,
                                 int outside_zero, float* out, float* lad, int32_t* bin) {
  fast::Knot4 rec[9];
  fast::build_records<8>(prm, mode, (float)bound, rec);
  int flag = 0;
  for (int64_t i = 0; i < n; ++i) {
    int b = -1;
    if (inverse) fast::table_element<8, true>(rec, v[i], (float)bound, outside_zero, out + i, lad + i, &b, &flag);
    else fast::table_element<8, false>(rec, v[i], (float)bound, outside_zero, out + i, lad + i, &b, &flag);
    bin[i] = b;
  }
  return flag;
}
int hostcheck_rqs_fast_table_bwd(const float* prm, const float* v, double bound, int64_t n, int mode, int inverse,
                                 int outside_zero, const float* gout, const float* glad, float* gv, float* gprm) {
  float gtab[27];
  fast::Knot4 rec[9];
  fast::build_records<8>(prm, mode, (float)bound, rec);
  for (int i = 0; i < 27; ++i) gtab[i] = 0.f;
  for (int64_t i = 0; i < n; ++i) {
    if (inverse) fast::table_element_bwd<8, true>(rec, v[i], (float)bound, outside_zero, gout[i], glad[i], gv + i, gtab);
    else fast::table_element_bwd<8, false>(rec, v[i], (float)bound, outside_zero, gout[i], glad[i], gv + i, gtab);
  }
  rqs_table_to_raw_grad<float, 8>(prm, 8, mode, (float)bound, gtab, gprm);
  return 0;
}
}

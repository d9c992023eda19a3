/* TEST INFRASTRUCTURE ONLY -- CPU restatement (plain C + OpenMP) of the spreading / interpolation
 * stages that FINUFFT performs inside pytorch_finufft.functional.finufft_type1 / finufft_type2,
 * the two external calls under the reference's NUFFT class (efgpnd.py:1496-1499, 1533-1549, 1679).
 *
 * FINUFFT (finufft>=1.2.0, pytorch-finufft, unpinned: setup.py:17-18) is an un-vendored dependency
 * that is absent from /root/reference and from this image; this file restates its PUBLISHED
 * algorithm (Barnett, Magland, af Klinteberg 2019): "exponential of semicircle" kernel
 * phi(z) = exp(beta (sqrt(1 - z^2) - 1)) of width w on a sigma = 2 fine grid, evaluated with a
 * piecewise-polynomial (Horner) table, complex strengths, OpenMP over points with thread-private
 * grids.  The FFT and deconvolution stages are done by numpy in oracle/fast_nufft.py.
 *
 * It is used (a) as the timed CPU baseline of bench.py (cpu_baseline / --impl reference) and (b) as a
 * second, scalable checker for sizes where the exact O(N M) NUDFT of oracle/pytorch_finufft is too slow.
 * It is validated against that exact NUDFT in tests/test_oracle.py.  The product never links it.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define CN_MAXW 16
#define CN_MAXDEG 24

/* weights k[i] = sum_j coef[i][j] s^j, s = 2 t - 1, t in [0,1) the offset of the footprint start */
static inline void cn_eval(const double* coef, int w, int ncoef, double s, double* k) {
  for (int i = 0; i < w; ++i) {
    const double* c = coef + (size_t)i * ncoef;
    double v = c[ncoef - 1];
    for (int j = ncoef - 2; j >= 0; --j) v = v * s + c[j];
    k[i] = v;
  }
}

static inline int64_t cn_wrap(int64_t i, int64_t n) {
  i %= n;
  return i < 0 ? i + n : i;
}

/* footprint of a point at fine-grid coordinate X (any real): first index i0 = ceil(X - w/2),
 * t = i0 - (X - w/2) in [0,1).  Grid point l = i0 + i sits at distance (t + i - w/2) from X. */
static inline int64_t cn_foot(double X, int w, double* s) {
  double a = X - 0.5 * w;
  double i0 = ceil(a);
  *s = 2.0 * (i0 - a) - 1.0;
  return (int64_t)i0;
}

int cn_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* grid (nf0 x nf1 x nf2, C order, interleaved complex) += sum_n c_n psi(l - X_n).
 * X: (dim, N) fine-grid coordinates; c: (N) interleaved complex, or NULL for all ones. */
void cn_spread(int dim, int64_t N, const double* X, const double* c, int w, int ncoef, const double* coef,
               const int64_t* nf, double* grid, int nthreads) {
  const int64_t n0 = nf[0], n1 = dim > 1 ? nf[1] : 1, n2 = dim > 2 ? nf[2] : 1;
  const int64_t cells = n0 * n1 * n2;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
  nthreads = 1;
#endif
  /* thread-private grids when they are small next to N, otherwise atomics on the shared grid */
  int priv = nthreads > 1 && (double)cells * nthreads <= 4.0 * (double)N + 1e6;
  double* scratch = NULL;
  if (priv) scratch = (double*)calloc((size_t)cells * 2 * (nthreads - 1), sizeof(double));
  if (priv && !scratch) priv = 0;
#pragma omp parallel num_threads(nthreads)
  {
#ifdef _OPENMP
    int tid = omp_get_thread_num();
#else
    int tid = 0;
#endif
    double* g = (priv && tid > 0) ? scratch + (size_t)(tid - 1) * cells * 2 : grid;
    const int use_atomic = !priv && nthreads > 1;
    double k0[CN_MAXW], k1[CN_MAXW], k2[CN_MAXW];
    int64_t j1[CN_MAXW], j2[CN_MAXW];
#pragma omp for schedule(static)
    for (int64_t n = 0; n < N; ++n) {
      double s;
      double cr = c ? c[2 * n] : 1.0, ci = c ? c[2 * n + 1] : 0.0;
      int64_t a0 = cn_foot(X[n], w, &s);
      cn_eval(coef, w, ncoef, s, k0);
      int64_t a1 = 0, a2 = 0;
      k1[0] = 1.0;
      k2[0] = 1.0;
      int w1 = 1, w2 = 1;
      if (dim > 1) {
        a1 = cn_foot(X[N + n], w, &s);
        cn_eval(coef, w, ncoef, s, k1);
        w1 = w;
      }
      if (dim > 2) {
        a2 = cn_foot(X[2 * N + n], w, &s);
        cn_eval(coef, w, ncoef, s, k2);
        w2 = w;
      }
      for (int i = 0; i < w1; ++i) j1[i] = cn_wrap(a1 + i, n1);
      for (int i = 0; i < w2; ++i) j2[i] = cn_wrap(a2 + i, n2);
      for (int i0 = 0; i0 < w; ++i0) {
        int64_t r0 = cn_wrap(a0 + i0, n0) * n1;
        double vr0 = cr * k0[i0], vi0 = ci * k0[i0];
        for (int i1 = 0; i1 < w1; ++i1) {
          double* row = g + 2 * ((r0 + j1[i1]) * n2);
          double vr = vr0 * k1[i1], vi = vi0 * k1[i1];
          if (!use_atomic) {
            for (int i2 = 0; i2 < w2; ++i2) {
              row[2 * j2[i2]] += vr * k2[i2];
              row[2 * j2[i2] + 1] += vi * k2[i2];
            }
          } else {
            for (int i2 = 0; i2 < w2; ++i2) {
#pragma omp atomic
              row[2 * j2[i2]] += vr * k2[i2];
#pragma omp atomic
              row[2 * j2[i2] + 1] += vi * k2[i2];
            }
          }
        }
      }
    }
    if (priv) {
#pragma omp for schedule(static)
      for (int64_t q = 0; q < 2 * cells; ++q) {
        double acc = grid[q];
        for (int t = 0; t < nthreads - 1; ++t) acc += scratch[(size_t)t * cells * 2 + q];
        grid[q] = acc;
      }
    }
  }
  free(scratch);
}

/* out_n = sum_l grid[l] psi(l - X_n);  out: (N) interleaved complex */
void cn_interp(int dim, int64_t N, const double* X, int w, int ncoef, const double* coef, const int64_t* nf,
               const double* grid, double* out, int nthreads) {
  const int64_t n0 = nf[0], n1 = dim > 1 ? nf[1] : 1, n2 = dim > 2 ? nf[2] : 1;
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
  nthreads = 1;
#endif
#pragma omp parallel num_threads(nthreads)
  {
    double k0[CN_MAXW], k1[CN_MAXW], k2[CN_MAXW];
    int64_t j1[CN_MAXW], j2[CN_MAXW];
#pragma omp for schedule(static)
    for (int64_t n = 0; n < N; ++n) {
      double s;
      int64_t a0 = cn_foot(X[n], w, &s);
      cn_eval(coef, w, ncoef, s, k0);
      int64_t a1 = 0, a2 = 0;
      k1[0] = 1.0;
      k2[0] = 1.0;
      int w1 = 1, w2 = 1;
      if (dim > 1) {
        a1 = cn_foot(X[N + n], w, &s);
        cn_eval(coef, w, ncoef, s, k1);
        w1 = w;
      }
      if (dim > 2) {
        a2 = cn_foot(X[2 * N + n], w, &s);
        cn_eval(coef, w, ncoef, s, k2);
        w2 = w;
      }
      for (int i = 0; i < w1; ++i) j1[i] = cn_wrap(a1 + i, n1);
      for (int i = 0; i < w2; ++i) j2[i] = cn_wrap(a2 + i, n2);
      double ar = 0, ai = 0;
      for (int i0 = 0; i0 < w; ++i0) {
        int64_t r0 = cn_wrap(a0 + i0, n0) * n1;
        double br = 0, bi = 0;
        for (int i1 = 0; i1 < w1; ++i1) {
          const double* row = grid + 2 * ((r0 + j1[i1]) * n2);
          double cr = 0, ci = 0;
          for (int i2 = 0; i2 < w2; ++i2) {
            cr += row[2 * j2[i2]] * k2[i2];
            ci += row[2 * j2[i2] + 1] * k2[i2];
          }
          br += cr * k1[i1];
          bi += ci * k1[i1];
        }
        ar += br * k0[i0];
        ai += bi * k0[i0];
      }
      out[2 * n] = ar;
      out[2 * n + 1] = ai;
    }
  }
}

/* ORACLE / CPU BASELINE PORT (test infrastructure, never on the product path).
 *
 * A plain C + OpenMP restatement of the improved w-stacking gridder that the
 * reference reaches through ducc0.wgridder.ms2dirty
 * (src/ska_sdp_cip/invert.py:170-183; ducc0 == 0.34.0 is a third-party wheel
 * that is absent here, so this follows the published algorithm -- Arras et al.
 * 2021 "Efficient wide-field radio interferometric response", stages G1-G4 of
 * SURVEY.md section 8a -- and is labelled "port", not "reference").
 * PARITY UNPINNED against ducc0 itself; validated against oracle/csrc/dft.c.
 *
 * The file is compiled twice through wgrid_impl.h: once with float grids
 * (what ms2dirty does for complex64 input) and once with double grids.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define C_LIGHT 299792458.0

typedef struct {
  int nx, ny, nu, nv, W, nplanes;
  double px, py, sigma, beta; /* phi(x)=exp(beta*W*(sqrt(1-x^2)-1)) */
  double dw, w0, nshift;      /* plane p sits at w0 + p*dw */
} oracle_plan;

/* G1: scan. active = weight != 0 and visibility != 0; |w| range over active samples. */
void oracle_wg_scan(const double* uvw, int64_t nrow, const double* freq, int64_t nchan,
                    const float* vis /* (nrow,nchan,2) */, const float* wgt /* may be NULL */,
                    double* wmin, double* wmax, int64_t* nactive, double* sumw)
{
  double lo = 1e300, hi = -1e300, sw = 0.0;
  int64_t na = 0;
#pragma omp parallel for reduction(min : lo) reduction(max : hi) reduction(+ : sw, na) schedule(static)
  for (int64_t r = 0; r < nrow; ++r) {
    double aw = fabs(uvw[3 * r + 2]);
    for (int64_t c = 0; c < nchan; ++c) {
      int64_t k = r * nchan + c;
      float g = wgt ? wgt[k] : 1.0f;
      sw += g;
      if (g == 0.0f || (vis[2 * k] == 0.0f && vis[2 * k + 1] == 0.0f)) continue;
      double w = aw * (freq[c] / C_LIGHT);
      if (w < lo) lo = w;
      if (w > hi) hi = w;
      ++na;
    }
  }
  *wmin = lo; *wmax = hi; *nactive = na; *sumw = sw;
}

typedef struct {
  int32_t i0, j0, p0;   /* first grid cell (shifted coordinates, may be negative) and first plane */
  double fu, fv, fw;    /* offsets of the first cell/plane from the left edge of the support, in [0,1) */
  double vr, vi;        /* weighted, conjugated-if-flipped, n-shift-rotated visibility */
} oracle_rec;

static inline double es_kernel(double x, double betaW)
{
  double t = 1.0 - x * x;
  return t > 0.0 ? exp(betaW * (sqrt(t) - 1.0)) : exp(-betaW) * (t == 0.0);
}

/* Per-sample geometry.  Returns 0 for inactive samples. */
static inline int make_rec(const oracle_plan* P, const double* uvw3, double finv,
                           float vre, float vim, float g, oracle_rec* R)
{
  if (g == 0.0f || (vre == 0.0f && vim == 0.0f)) return 0;
  double u = uvw3[0] * finv, v = uvw3[1] * finv, w = uvw3[2] * finv;
  double re = (double)vre * g, im = (double)vim * g;
  if (w < 0) { u = -u; v = -v; w = -w; im = -im; }
  /* n-shift: V' = V exp(+2 pi i w s) */
  if (P->nshift != 0.0) {
    double ph = w * P->nshift; ph -= rint(ph);
    double c = cos(2 * M_PI * ph), s = sin(2 * M_PI * ph);
    double t = re * c - im * s; im = re * s + im * c; re = t;
  }
  double xu = u * P->px; xu -= rint(xu);
  double xv = v * P->py; xv -= rint(xv);
  double us = xu * P->nu + 0.5 * P->nu - 0.5 * P->W;   /* left edge of the support, shifted coords */
  double vs = xv * P->nv + 0.5 * P->nv - 0.5 * P->W;
  double ps = (w - P->w0) / P->dw - 0.5 * P->W;
  double ci = ceil(us), cj = ceil(vs), cp = ceil(ps);
  R->i0 = (int32_t)ci; R->j0 = (int32_t)cj; R->p0 = (int32_t)cp;
  R->fu = ci - us; R->fv = cj - vs; R->fw = cp - ps;
  R->vr = re; R->vi = im;
  return 1;
}

#define TS 32 /* colour-tile edge; must be >= W-1 */

/* Build tile-sorted records.  Returns number of active records; offsets has ntu*ntv+1 entries. */
int64_t oracle_wg_bin(const oracle_plan* P, const double* uvw, int64_t nrow, const double* freq,
                      int64_t nchan, const float* vis, const float* wgt,
                      oracle_rec* recs /* capacity nrow*nchan */, int64_t* offsets, int ntu, int ntv)
{
  int64_t ntile = (int64_t)ntu * ntv;
  int64_t* cnt = (int64_t*)calloc(ntile + 1, sizeof(int64_t));
  int32_t* key = (int32_t*)malloc(sizeof(int32_t) * nrow * nchan);
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < nrow; ++r)
    for (int64_t c = 0; c < nchan; ++c) {
      int64_t k = r * nchan + c;
      oracle_rec R;
      float g = wgt ? wgt[k] : 1.0f;
      if (!make_rec(P, uvw + 3 * r, freq[c] / C_LIGHT, vis[2 * k], vis[2 * k + 1], g, &R)) { key[k] = -1; continue; }
      int tu = (R.i0 + 16) / TS, tv = (R.j0 + 16) / TS; /* +16 keeps the dividend non-negative (W<=16) */
      key[k] = tu * ntv + tv;
#pragma omp atomic
      cnt[key[k] + 1]++;
    }
  for (int64_t t = 0; t < ntile; ++t) cnt[t + 1] += cnt[t];
  memcpy(offsets, cnt, sizeof(int64_t) * (ntile + 1));
  /* stable serial scatter keeps (row, chan) order inside a tile => deterministic sums */
  for (int64_t r = 0; r < nrow; ++r)
    for (int64_t c = 0; c < nchan; ++c) {
      int64_t k = r * nchan + c;
      if (key[k] < 0) continue;
      oracle_rec R;
      float g = wgt ? wgt[k] : 1.0f;
      make_rec(P, uvw + 3 * r, freq[c] / C_LIGHT, vis[2 * k], vis[2 * k + 1], g, &R);
      recs[cnt[key[k]]++] = R;
    }
  int64_t n = offsets[ntile];
  free(cnt); free(key);
  return n;
}

#define REAL float
#define SUFFIX(n) n##_f32
#include "wgrid_impl.h"
#undef REAL
#undef SUFFIX
#define REAL double
#define SUFFIX(n) n##_f64
#include "wgrid_impl.h"
#undef REAL
#undef SUFFIX

/* kernel transform c(xi) = int_{-W/2}^{W/2} phi(2t/W) cos(2 pi t xi) dt by Gauss-Legendre
   on theta (t = W/2 sin theta); nodes/weights for [-1,1] supplied by the caller. */
void oracle_wg_chat(const oracle_plan* P, const double* gl_x, const double* gl_w, int nq,
                    const double* xi, int64_t n, double* out)
{
  double bW = P->beta * P->W, h = 0.5 * P->W;
#pragma omp parallel for schedule(static)
  for (int64_t k = 0; k < n; ++k) {
    double acc = 0;
    for (int q = 0; q < nq; ++q) {
      double th = gl_x[q] * (M_PI / 2);
      double ct = cos(th);
      acc += gl_w[q] * exp(bW * (ct - 1.0)) * ct * cos(2 * M_PI * h * sin(th) * xi[k]);
    }
    out[k] = acc * h * (M_PI / 2);
  }
}

/* G4 final: divide by c_u c_v c_w n; cu/cv tables given; cw via quadrature per pixel. */
void oracle_wg_correct(const oracle_plan* P, const double* gl_x, const double* gl_w, int nq,
                       const double* cu, const double* cv, double* img /* (nx,ny) */)
{
  double bW = P->beta * P->W, h = 0.5 * P->W;
#pragma omp parallel for schedule(dynamic, 8)
  for (int i = 0; i < P->nx; ++i)
    for (int j = 0; j < P->ny; ++j) {
      double l = (i - P->nx / 2) * P->px, m = (j - P->ny / 2) * P->py, e = l * l + m * m;
      if (e >= 1.0) { img[(int64_t)i * P->ny + j] = 0; continue; }
      double nm1 = -e / (sqrt(1.0 - e) + 1.0);
      double xi = (nm1 + P->nshift) * P->dw, acc = 0;
      for (int q = 0; q < nq; ++q) {
        double th = gl_x[q] * (M_PI / 2), ct = cos(th);
        acc += gl_w[q] * exp(bW * (ct - 1.0)) * ct * cos(2 * M_PI * h * sin(th) * xi);
      }
      double cw = acc * h * (M_PI / 2);
      img[(int64_t)i * P->ny + j] /= cu[i] * cv[j] * cw * (nm1 + 1.0);
    }
}

int oracle_num_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* The timed CPU baseline uses every host core even when the launcher exported OMP_NUM_THREADS=1
 * (torchrun does); n <= 0 leaves the setting alone. */
void oracle_set_num_threads(int n)
{
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

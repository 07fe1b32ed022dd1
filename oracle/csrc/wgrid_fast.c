/* CPU BASELINE (test infrastructure, never on the product path): the timed CPU arm of bench.py.
 *
 * The same improved w-stacking algorithm as wgrid.c (the published scheme behind ducc0.wgridder.ms2dirty, reference
 * call site src/ska_sdp_cip/invert.py:170-183; ducc0 itself is not installable here), organised for speed on many
 * cores rather than for checking:
 *   - records are binned by uv tile with a two-pass parallel counting sort (per-thread counts, no serial scatter);
 *   - a record's 3 W kernel taps are evaluated ONCE, in float32 and eight at a time (a vectorised polynomial
 *     exponential instead of libm calls), and applied to all its W planes -- the planes of a block are resident
 *     together -- instead of re-evaluating them for every plane it touches; a row of the support is updated as 16
 *     consecutive floats (two 8-wide fused multiply-adds);
 *   - tiles of 32 x 32 cells are processed in four colour phases so that concurrent threads never share a cell;
 *   - the w correction comes from a table in (n - 1) with linear interpolation instead of a quadrature per pixel.
 * wgrid.c stays the float64 checker; tests/test_oracle.py compares both with the DFT.  PARITY UNPINNED against ducc0.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define C_LIGHT 299792458.0
#define FTS 32 /* tile edge in cells, >= 8 (a support of W <= 8 cells plus the halo of the colour scheme); 16 measured
                  no faster */
#define FTH (FTS / 2)
int oracle_fast_tile_edge(void) { return FTS; }

typedef struct {
  int nx, ny, nu, nv, W, nplanes;
  double px, py, sigma, beta;
  double dw, w0, nshift;
} oracle_plan;

typedef struct {
  int32_t i0, j0, p0;
  float fu, fv, fw;
  float vr, vi;
} fast_rec;

static inline int fast_make_rec(const oracle_plan* P, const double* uvw3, double finv, float vre, float vim, float g, fast_rec* R)
{
  if (g == 0.0f || (vre == 0.0f && vim == 0.0f)) return 0;
  double u = uvw3[0] * finv, v = uvw3[1] * finv, w = uvw3[2] * finv;
  double re = (double)vre * g, im = (double)vim * g;
  if (w < 0) { u = -u; v = -v; w = -w; im = -im; }
  if (P->nshift != 0.0) {
    double ph = w * P->nshift; ph -= rint(ph);
    double c = cos(2 * M_PI * ph), s = sin(2 * M_PI * ph);
    double t = re * c - im * s; im = re * s + im * c; re = t;
  }
  double xu = u * P->px; xu -= rint(xu);
  double xv = v * P->py; xv -= rint(xv);
  double us = xu * P->nu + 0.5 * P->nu - 0.5 * P->W;
  double vs = xv * P->nv + 0.5 * P->nv - 0.5 * P->W;
  double ps = (w - P->w0) / P->dw - 0.5 * P->W;
  double ci = ceil(us), cj = ceil(vs), cp = ceil(ps);
  R->i0 = (int32_t)ci; R->j0 = (int32_t)cj; R->p0 = (int32_t)cp;
  R->fu = (float)(ci - us); R->fv = (float)(cj - vs); R->fw = (float)(cp - ps);
  R->vr = (float)re; R->vi = (float)im;
  return 1;
}

/* Two-pass parallel counting sort of the active samples by tile.  offsets: ntu*ntv+1.  Returns the record count. */
int64_t oracle_fast_bin(const oracle_plan* P, const double* uvw, int64_t nrow, const double* freq, int64_t nchan,
                        const float* vis, const float* wgt, fast_rec* recs, int64_t* offsets, int ntu, int ntv)
{
  const int64_t ntile = (int64_t)ntu * ntv;
  int nth = 1;
#ifdef _OPENMP
  nth = omp_get_max_threads();
#endif
  int64_t* cnt = (int64_t*)calloc((size_t)nth * ntile, sizeof(int64_t));
  double* finv = (double*)malloc(sizeof(double) * nchan);
  for (int64_t c = 0; c < nchan; ++c) finv[c] = freq[c] / C_LIGHT;
#pragma omp parallel
  {
    int t = 0, n = 1;
#ifdef _OPENMP
    t = omp_get_thread_num(); n = omp_get_num_threads();
#endif
    int64_t* my = cnt + (int64_t)t * ntile;
    const int64_t r0 = nrow * t / n, r1 = nrow * (t + 1) / n;
    for (int64_t r = r0; r < r1; ++r)
      for (int64_t c = 0; c < nchan; ++c) {
        int64_t k = r * nchan + c;
        fast_rec R;
        if (!fast_make_rec(P, uvw + 3 * r, finv[c], vis[2 * k], vis[2 * k + 1], wgt ? wgt[k] : 1.0f, &R)) continue;
        my[(int64_t)((R.i0 + FTH) / FTS) * ntv + (R.j0 + FTH) / FTS]++;
      }
#pragma omp barrier
#pragma omp single
    {
      int64_t run = 0;
      for (int64_t tile = 0; tile < ntile; ++tile) {
        offsets[tile] = run;
        for (int q = 0; q < n; ++q) {
          int64_t v = cnt[(int64_t)q * ntile + tile];
          cnt[(int64_t)q * ntile + tile] = run;
          run += v;
        }
      }
      offsets[ntile] = run;
    }
    for (int64_t r = r0; r < r1; ++r)
      for (int64_t c = 0; c < nchan; ++c) {
        int64_t k = r * nchan + c;
        fast_rec R;
        if (!fast_make_rec(P, uvw + 3 * r, finv[c], vis[2 * k], vis[2 * k + 1], wgt ? wgt[k] : 1.0f, &R)) continue;
        recs[my[(int64_t)((R.i0 + FTH) / FTS) * ntv + (R.j0 + FTH) / FTS]++] = R;
      }
  }
  int64_t total = offsets[ntile];
  free(cnt); free(finv);
  return total;
}

static inline float es_f(float x, float bW)
{
  float t = 1.0f - x * x;
  return t > 0.0f ? expf(bW * (sqrtf(t) - 1.0f)) : 0.0f;
}

/* The W taps of one axis, eight at a time so that the compiler vectorises the loop: exp(y), y in [-beta W, 0], as
   2^round(z) * e^{(z - round(z)) ln 2} with a degree-6 Taylor polynomial on |.| <= ln(2)/2 (relative error 1.2e-7)
   instead of a libm call per tap -- the taps were 2/3 of the gridding time. */
static inline void es_taps8(float f, float tw, float bW, int W, float* restrict out)
{
#pragma omp simd
  for (int a = 0; a < 8; ++a) {
    const float x = (f + (float)a) * tw - 1.0f;
    const float t = 1.0f - x * x;
    const float tt = t > 0.0f ? t : 0.0f;
    const float z = bW * (sqrtf(tt) - 1.0f) * 1.4426950408889634f;
    const float zi = rintf(z), r = (z - zi) * 0.6931471805599453f;
    float p = 1.3888889e-3f;
    p = p * r + 8.3333333e-3f;
    p = p * r + 4.1666667e-2f;
    p = p * r + 1.6666667e-1f;
    p = p * r + 0.5f;
    p = p * r + 1.0f;
    p = p * r + 1.0f;
    union { float f; int32_t i; } u;
    u.f = p;
    u.i += (int32_t)zi << 23;
    out[a] = (a < W && t > 0.0f) ? u.f : 0.0f;
  }
}

#define FLB (FTS + 8) /* edge of a tile's private buffer: origins span FTS cells, supports reach 8 further */

/* One record into `dst`, a stack of planes with `rowlen` cells per row and `plane` floats per plane, whose cell
   (row i, column j) for this record's first support cell is (ib, jb); no wrap-around inside the support.  W <= 8:
   one row of the support = 16 consecutive floats (re, im of 8 cells; the taps beyond W are zero), two 8-wide fused
   multiply-adds per row. */
static inline void fast_add8(const fast_rec* R, const float* ku, const float* kv2, const float* kw, int W, int c_lo, int c_hi,
                             float* dst, int64_t plane, int64_t rowlen, int first_plane, int ib, int jb, int nu_wrap)
{
  for (int c = c_lo; c < c_hi; ++c) {
    float* pl = dst + (int64_t)(R->p0 + c - first_plane) * plane;
    const float vr = R->vr * kw[c], vi = R->vi * kw[c];
    for (int a = 0; a < W; ++a) {
      int i = ib + a; if (nu_wrap && i >= nu_wrap) i -= nu_wrap;
      const float tr = vr * ku[a], ti = vi * ku[a];
      float* restrict d = pl + 2 * ((int64_t)i * rowlen + jb);
#pragma omp simd
      for (int m = 0; m < 16; ++m) d[m] += kv2[m] * ((m & 1) ? ti : tr);
    }
  }
}

/* Planes [p_lo, p_hi) of the stack are resident in grid (p_hi - p_lo, nu, nv, 2) float32: every record adds its
   W x W x (planes inside the block) contributions with taps evaluated once.  Tiles are processed in colour phases,
   one thread per tile; the few tiles that hold a large share of the records (the core of an SKA layout lands in the
   four tiles around the uv origin: a quarter of all records) are first handled by ALL threads, each with a private
   copy of the tile that is summed into the grid afterwards -- without that the densest tile of each colour is the
   critical path. */
void oracle_fast_grid(const oracle_plan* P, const fast_rec* recs, const int64_t* offsets, int ntu, int ntv,
                      int p_lo, int p_hi, float* grid)
{
  const int W = P->W, nu = P->nu, nv = P->nv;
  const float bW = (float)(P->beta * W), tw = 2.0f / (float)W;
  const int64_t plane = (int64_t)nu * nv * 2;
  const int narrow = W <= 8;  /* the vectorised forms below hold 8 taps per axis */
  const int npl = p_hi - p_lo;
  const int64_t ntile = (int64_t)ntu * ntv;
  int nth = 1;
#ifdef _OPENMP
  nth = omp_get_max_threads();
#endif
  /* a tile left to one thread must not be a noticeable share of a colour phase: 1/64 of a thread's fair share at most */
  int64_t dense = offsets[ntile] / (64 * (int64_t)nth);
  if (dense < 4096) dense = 4096;
  unsigned char* is_dense = (unsigned char*)calloc((size_t)ntile, 1);
  if (narrow && nth > 1) {
    const int64_t pplane = (int64_t)FLB * FLB * 2, pstride = pplane * npl;
    float* priv = NULL;
    for (int64_t t = 0; t < ntile; ++t) {
      const int64_t n = offsets[t + 1] - offsets[t];
      const int tu = (int)(t / ntv), tv = (int)(t % ntv);
      if (n <= dense || tu >= ntu - 2 || tv >= ntv - 2) continue;
      is_dense[t] = 1;
      if (!priv) priv = (float*)malloc(sizeof(float) * (size_t)pstride * nth);
      const int oi = tu * FTS - FTH, oj = tv * FTS - FTH;  /* first origin of the tile */
#pragma omp parallel
      {
        int q = 0, nq = 1;
#ifdef _OPENMP
        q = omp_get_thread_num(); nq = omp_get_num_threads();
#endif
        float* mine = priv + (int64_t)q * pstride;
        memset(mine, 0, sizeof(float) * (size_t)pstride);
        for (int64_t k = offsets[t] + n * q / nq; k < offsets[t] + n * (q + 1) / nq; ++k) {
          const fast_rec* R = &recs[k];
          int c_lo = p_lo - R->p0, c_hi = p_hi - R->p0;
          if (c_lo < 0) c_lo = 0;
          if (c_hi > W) c_hi = W;
          if (c_lo >= c_hi) continue;
          float ku[8], kv[8], kw[8], kv2[16];
          es_taps8(R->fu, tw, bW, W, ku);
          es_taps8(R->fv, tw, bW, W, kv);
          es_taps8(R->fw, tw, bW, W, kw);
          for (int b = 0; b < 8; ++b) kv2[2 * b] = kv2[2 * b + 1] = kv[b];
          fast_add8(R, ku, kv2, kw, W, c_lo, c_hi, mine, pplane, FLB, p_lo, R->i0 - oi, R->j0 - oj, 0);
        }
#pragma omp barrier
#pragma omp for collapse(2) schedule(static)
        for (int pc = 0; pc < npl; ++pc)
          for (int li = 0; li < FLB; ++li) {
            const int gi = ((oi + li + nu / 2) % nu + nu) % nu;
            for (int lj = 0; lj < FLB; ++lj) {
              float sr = 0.0f, si = 0.0f;
              const int64_t o = (int64_t)pc * pplane + 2 * ((int64_t)li * FLB + lj);
              for (int w = 0; w < nq; ++w) { sr += priv[(int64_t)w * pstride + o]; si += priv[(int64_t)w * pstride + o + 1]; }
              if (sr != 0.0f || si != 0.0f) {
                const int gj = ((oj + lj + nv / 2) % nv + nv) % nv;
                float* d = grid + (int64_t)pc * plane + 2 * ((int64_t)gi * nv + gj);
                d[0] += sr; d[1] += si;
              }
            }
          }
      }
    }
    free(priv);
  }
  for (int colour = 0; colour < 5; ++colour) {
#pragma omp parallel for schedule(dynamic, 1) collapse(2) if (colour < 4)
    for (int tu = (colour < 4 ? (colour >> 1) : 0); tu < ntu; tu += (colour < 4 ? 2 : 1))
      for (int tv = (colour < 4 ? (colour & 1) : 0); tv < ntv; tv += (colour < 4 ? 2 : 1)) {
        int seam = (tu >= ntu - 2) || (tv >= ntv - 2);
        if ((colour < 4) == seam) continue;
        int64_t t = (int64_t)tu * ntv + tv;
        if (is_dense[t]) continue;
        for (int64_t k = offsets[t]; k < offsets[t + 1]; ++k) {
          const fast_rec* R = &recs[k];
          int c_lo = p_lo - R->p0, c_hi = p_hi - R->p0;
          if (c_lo < 0) c_lo = 0;
          if (c_hi > W) c_hi = W;
          if (c_lo >= c_hi) continue;
          float ku[16], kv[16], kw[16];
          if (narrow) {
            es_taps8(R->fu, tw, bW, W, ku);
            es_taps8(R->fv, tw, bW, W, kv);
            es_taps8(R->fw, tw, bW, W, kw);
          } else {
            for (int a = 0; a < W; ++a) {
              ku[a] = es_f((R->fu + a) * tw - 1.0f, bW);
              kv[a] = es_f((R->fv + a) * tw - 1.0f, bW);
              kw[a] = es_f((R->fw + a) * tw - 1.0f, bW);
            }
          }
          int ib = ((R->i0 + nu / 2) % nu + nu) % nu, jb = ((R->j0 + nv / 2) % nv + nv) % nv;
          const int wrap_j = jb + W > nv;
          if (narrow && jb + 8 <= nv) {
            /* (the cells the zero taps beyond W "touch" lie inside this tile's halo, which no other thread works on in
               this colour phase) */
            float kv2[16];
            for (int b = 0; b < 8; ++b) kv2[2 * b] = kv2[2 * b + 1] = kv[b];
            fast_add8(R, ku, kv2, kw, W, c_lo, c_hi, grid, plane, nv, p_lo, ib, jb, nu);
            continue;
          }
          for (int c = c_lo; c < c_hi; ++c) {
            float* pl = grid + (int64_t)(R->p0 + c - p_lo) * plane;
            const float vr = R->vr * kw[c], vi = R->vi * kw[c];
            for (int a = 0; a < W; ++a) {
              int i = ib + a; if (i >= nu) i -= nu;
              const float tr = vr * ku[a], ti = vi * ku[a];
              float* row = pl + 2 * (int64_t)i * nv;
              if (!wrap_j) {
                float* d = row + 2 * jb;
                for (int b = 0; b < W; ++b) { d[2 * b] += tr * kv[b]; d[2 * b + 1] += ti * kv[b]; }
              } else {
                for (int b = 0; b < W; ++b) {
                  int j = jb + b; if (j >= nv) j -= nv;
                  row[2 * j] += tr * kv[b]; row[2 * j + 1] += ti * kv[b];
                }
              }
            }
          }
        }
      }
  }
  free(is_dense);
}

/* img += Re( F[(i-nx/2) mod nu, (j-ny/2) mod nv] exp(-2 pi i w_p (n-1+s)) ), float32 transform, float64 image */
void oracle_fast_wscreen(const oracle_plan* P, const float* F, int p, double* img)
{
  const int nx = P->nx, ny = P->ny, nu = P->nu, nv = P->nv;
  const double wp = P->w0 + p * P->dw;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < nx; ++i) {
    int gi = (i - nx / 2 + nu) % nu;
    double l = (i - nx / 2) * P->px;
    for (int j = 0; j < ny; ++j) {
      int gj = (j - ny / 2 + nv) % nv;
      double m = (j - ny / 2) * P->py, e = l * l + m * m;
      if (e >= 1.0) continue;
      double nm1 = -e / (sqrt(1.0 - e) + 1.0);
      double ph = wp * (nm1 + P->nshift); ph -= rint(ph);
      float c = cosf((float)(2 * M_PI * ph)), s = sinf((float)(2 * M_PI * ph));
      const float* f = F + 2 * ((int64_t)gi * nv + gj);
      img[(int64_t)i * ny + j] += (double)(f[0] * c + f[1] * s);
    }
  }
}

/* divide by c_u c_v c_w n with c_w from a table over nm1 in [nm1_lo, 0] (ntab points, linear interpolation) */
void oracle_fast_correct(const oracle_plan* P, const double* cu, const double* cv, const double* cw_tab, int ntab,
                         double nm1_lo, double* img)
{
  const double scale = nm1_lo < 0 ? (ntab - 1) / -nm1_lo : 0.0;
#pragma omp parallel for schedule(static)
  for (int i = 0; i < P->nx; ++i)
    for (int j = 0; j < P->ny; ++j) {
      double l = (i - P->nx / 2) * P->px, m = (j - P->ny / 2) * P->py, e = l * l + m * m;
      if (e >= 1.0) { img[(int64_t)i * P->ny + j] = 0; continue; }
      double nm1 = -e / (sqrt(1.0 - e) + 1.0);
      double x = (nm1 - nm1_lo) * scale;
      int k = (int)x; if (k > ntab - 2) k = ntab - 2; if (k < 0) k = 0;
      double f = x - k, cw = cw_tab[k] * (1.0 - f) + cw_tab[k + 1] * f;
      img[(int64_t)i * P->ny + j] /= cu[i] * cv[j] * cw * (nm1 + 1.0);
    }
}

/* ORACLE (test infrastructure, never on the product path).
 *
 * Direct evaluation of the measurement equation that ducc0.wgridder.ms2dirty
 * (reference call site src/ska_sdp_cip/invert.py:170-183) approximates:
 *
 *   D[i,j] = 1/n * sum_k Re( V_k w_k exp(+2 pi i (u_k l + v_k m - w_k (n-1))) )
 *   l=(i-nx/2) px, m=(j-ny/2) py, n=sqrt(1-l^2-m^2), (u,v,w)_k = uvw_row f_chan/c
 *
 * float64 throughout.  Evaluated either on the full image or on a list of
 * sampled pixels.  PARITY UNPINNED against ducc0 itself (not installable here);
 * this is the definition ducc0's own test-suite checks ms2dirty against.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#define C_LIGHT 299792458.0

/* pix: npix flat indices (i*ny+j) or NULL for the whole image */
void oracle_dft_dirty(const double* uvw, int64_t nrow, const double* freq, int64_t nchan,
                      const double* vis_re, const double* vis_im, /* (nrow,nchan) already times weight */
                      int nx, int ny, double px, double py,
                      const int64_t* pix, int64_t npix, double* out)
{
  int64_t ntot = pix ? npix : (int64_t)nx * ny;
#pragma omp parallel for schedule(dynamic, 16)
  for (int64_t t = 0; t < ntot; ++t) {
    int64_t flat = pix ? pix[t] : t;
    int i = (int)(flat / ny), j = (int)(flat % ny);
    double l = (i - nx / 2) * px, m = (j - ny / 2) * py;
    double e = l * l + m * m;
    if (e >= 1.0) { out[t] = 0.0; continue; }
    double nm1 = -e / (sqrt(1.0 - e) + 1.0);
    double n = nm1 + 1.0;
    double acc = 0.0;
    for (int64_t r = 0; r < nrow; ++r) {
      double ph0 = uvw[3 * r] * l + uvw[3 * r + 1] * m - uvw[3 * r + 2] * nm1; /* metres */
      const double* vr = vis_re + r * nchan;
      const double* vi = vis_im + r * nchan;
      for (int64_t c = 0; c < nchan; ++c) {
        if (vr[c] == 0.0 && vi[c] == 0.0) continue;
        double ph = ph0 * (freq[c] / C_LIGHT);
        ph -= rint(ph);
        double a = 2.0 * M_PI * ph;
        acc += vr[c] * cos(a) - vi[c] * sin(a);
      }
    }
    out[t] = acc / n;
  }
}

/* ORACLE / CPU BASELINE PORT -- precision-generic part of wgrid.c (included twice). */

/* G2: add every record that touches plane p into grid (nu,nv,2) of REAL.
   Four colour phases over TS x TS tiles so concurrent tiles never overlap. */
void SUFFIX(oracle_wg_grid_plane)(const oracle_plan* P, const oracle_rec* recs, const int64_t* offsets,
                                  int ntu, int ntv, int p, REAL* grid)
{
  const int W = P->W, nu = P->nu, nv = P->nv;
  const double bW = P->beta * W;
  /* colour 4 = the wrap-around seam (last two tile rows/columns), done by one thread */
  for (int colour = 0; colour < 5; ++colour) {
#pragma omp parallel for schedule(dynamic, 1) collapse(2) if (colour < 4)
    for (int tu = (colour < 4 ? (colour >> 1) : 0); tu < ntu; tu += (colour < 4 ? 2 : 1))
      for (int tv = (colour < 4 ? (colour & 1) : 0); tv < ntv; tv += (colour < 4 ? 2 : 1)) {
        int seam = (tu >= ntu - 2) || (tv >= ntv - 2);
        if ((colour < 4) == seam) continue;
        int64_t t = (int64_t)tu * ntv + tv;
        for (int64_t k = offsets[t]; k < offsets[t + 1]; ++k) {
          const oracle_rec* R = &recs[k];
          int c = p - R->p0;
          if (c < 0 || c >= W) continue;
          double kw = es_kernel((2.0 * (R->fw + c) - W) / W, bW);
          REAL ku[16], kv[16];
          for (int a = 0; a < W; ++a) {
            ku[a] = (REAL)es_kernel((2.0 * (R->fu + a) - W) / W, bW);
            kv[a] = (REAL)es_kernel((2.0 * (R->fv + a) - W) / W, bW);
          }
          REAL vr = (REAL)(R->vr * kw), vi = (REAL)(R->vi * kw);
          /* shifted -> FFT-order index: (i - nu/2) mod nu */
          int ib = ((R->i0 + nu / 2) % nu + nu) % nu, jb = ((R->j0 + nv / 2) % nv + nv) % nv;
          for (int a = 0; a < W; ++a) {
            int i = ib + a; if (i >= nu) i -= nu;
            REAL tr = vr * ku[a], ti = vi * ku[a];
            REAL* row = grid + 2 * (int64_t)i * nv;
            for (int b = 0; b < W; ++b) {
              int j = jb + b; if (j >= nv) j -= nv;
              row[2 * j] += tr * kv[b];
              row[2 * j + 1] += ti * kv[b];
            }
          }
        }
      }
  }
}

/* G4 per plane: img += Re( F[(i-nx/2) mod nu, (j-ny/2) mod nv] * exp(-2 pi i w_p (n-1+s)) ) */
void SUFFIX(oracle_wg_wscreen)(const oracle_plan* P, const REAL* F /* (nu,nv,2) */, int p, double* img)
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
      double c = cos(2 * M_PI * ph), s = sin(2 * M_PI * ph);
      const REAL* f = F + 2 * ((int64_t)gi * nv + gj);
      img[(int64_t)i * ny + j] += (double)f[0] * c + (double)f[1] * s;
    }
  }
}

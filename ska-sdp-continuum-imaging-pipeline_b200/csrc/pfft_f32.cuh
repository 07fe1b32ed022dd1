// float32 passes of pfft.cu on PACKED PAIRS: every thread transforms two sequences at once, their real parts in one
// 64-bit register pair and their imaginary parts in another, so that each add / multiply / fma of the butterflies is one
// f32x2 instruction for both (the passes are bound by instruction issue and by the shared-memory pipe, not by HBM,
// unless the arithmetic is packed).  The two sequences are
//   * column passes: two neighbouring columns of the 16-column block (same twiddles: scalars, broadcast);
//   * row pass: the butterflies k and k + 1 of the same block (their operands are neighbours in shared memory; the
//     twiddles come packed from per-stage tables); the last stage (span = radix) is scalar, one butterfly per thread,
//     and stores straight to global memory.
// Shared memory and the work area G hold granules of two sequences / two neighbouring points, float4 (re0, re1, im0, im1):
// one 128-bit access per operand pair, in the register layout of the arithmetic.  Twiddles w_L^{qk} are read from
// per-stage tables [q - 1][k] built on the host (exact to half an ulp, coalesced over k) instead of being multiplied up.
#pragma once

namespace cip {
namespace {

using p2 = unsigned long long;
__device__ __forceinline__ p2 pk(float a, float b) { p2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void upk(p2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ p2 padd(p2 a, p2 b) { p2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ p2 psub(p2 a, p2 b) { p2 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ p2 pmul(p2 a, p2 b) { p2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ p2 pfma(p2 a, p2 b, p2 c) { p2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ p2 pdup(float x) { return pk(x, x); }

struct pc { p2 re, im; };  // two complex numbers
__device__ __forceinline__ pc pc_add(pc a, pc b) { pc r; r.re = padd(a.re, b.re); r.im = padd(a.im, b.im); return r; }
__device__ __forceinline__ pc pc_sub(pc a, pc b) { pc r; r.re = psub(a.re, b.re); r.im = psub(a.im, b.im); return r; }
// a (c + i s), the same scalar factor for both
__device__ __forceinline__ pc pc_mulc(pc a, float c, float s) {
  pc r;
  const p2 C = pdup(c), S = pdup(s), NS = pdup(-s);
  r.re = pfma(a.im, NS, pmul(a.re, C));
  r.im = pfma(a.im, C, pmul(a.re, S));
  return r;
}
// a (c + i s), a factor per sequence
__device__ __forceinline__ pc pc_mulp(pc a, p2 c, p2 s) {
  pc r;
  r.re = psub(pmul(a.re, c), pmul(a.im, s));
  r.im = pfma(a.im, c, pmul(a.re, s));
  return r;
}

__device__ __forceinline__ void dft4p(pc& x0, pc& x1, pc& x2, pc& x3) {
  const pc t0 = pc_add(x0, x2), t1 = pc_sub(x0, x2), t2 = pc_add(x1, x3);
  pc t3;  // i (x1 - x3)
  t3.re = psub(x3.im, x1.im);
  t3.im = psub(x1.re, x3.re);
  x0 = pc_add(t0, t2); x1 = pc_add(t1, t3); x2 = pc_sub(t0, t2); x3 = pc_sub(t1, t3);
}

template <int R> struct DftP;
template <> struct DftP<2> {
  static __device__ __forceinline__ void run(pc (&x)[2]) { const pc t = pc_sub(x[0], x[1]); x[0] = pc_add(x[0], x[1]); x[1] = t; }
};
template <> struct DftP<4> { static __device__ __forceinline__ void run(pc (&x)[4]) { dft4p(x[0], x[1], x[2], x[3]); } };
template <> struct DftP<3> {
  static __device__ __forceinline__ void run(pc (&x)[3]) {
    const float s3 = 0.86602540378443864676f;
    const pc t = pc_add(x[1], x[2]), d = pc_sub(x[1], x[2]);
    pc m;
    m.re = pfma(t.re, pdup(-0.5f), x[0].re);
    m.im = pfma(t.im, pdup(-0.5f), x[0].im);
    x[0] = pc_add(x[0], t);
    x[1].re = pfma(d.im, pdup(-s3), m.re); x[1].im = pfma(d.re, pdup(s3), m.im);   // m + i s3 d
    x[2].re = pfma(d.im, pdup(s3), m.re);  x[2].im = pfma(d.re, pdup(-s3), m.im);  // m - i s3 d
  }
};
template <> struct DftP<8> {
  static __device__ __forceinline__ void run(pc (&x)[8]) {
    const float h = 0.70710678118654752440f;
    dft4p(x[0], x[2], x[4], x[6]);
    dft4p(x[1], x[3], x[5], x[7]);
    const pc e0 = x[0], e1 = x[2], e2 = x[4], e3 = x[6], o0 = x[1], o1 = x[3], o2 = x[5], o3 = x[7];
    x[0] = pc_add(e0, o0); x[4] = pc_sub(e0, o0);
    const p2 d1 = psub(o1.re, o1.im), s1 = padd(o1.re, o1.im);  // w8 o1 = h (d1 + i s1)
    x[1].re = pfma(d1, pdup(h), e1.re);  x[1].im = pfma(s1, pdup(h), e1.im);
    x[5].re = pfma(d1, pdup(-h), e1.re); x[5].im = pfma(s1, pdup(-h), e1.im);
    x[2].re = psub(e2.re, o2.im); x[2].im = padd(e2.im, o2.re);  // i o2
    x[6].re = padd(e2.re, o2.im); x[6].im = psub(e2.im, o2.re);
    const p2 d3 = psub(o3.re, o3.im), s3 = padd(o3.re, o3.im);  // w8^3 o3 = h (-s3 + i d3)
    x[3].re = pfma(s3, pdup(-h), e3.re); x[3].im = pfma(d3, pdup(h), e3.im);
    x[7].re = pfma(s3, pdup(h), e3.re);  x[7].im = pfma(d3, pdup(-h), e3.im);
  }
};
template <> struct DftP<16> {
  static __device__ __forceinline__ void run(pc (&x)[16]) {
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f, h = 0.70710678118654752440f;
#pragma unroll
    for (int r1 = 0; r1 < 4; ++r1) dft4p(x[r1], x[r1 + 4], x[r1 + 8], x[r1 + 12]);  // x[r1 + 4 q2] = A_r1[q2]
    x[5] = pc_mulc(x[5], c1, s1);  x[9] = pc_mulc(x[9], h, h);    x[13] = pc_mulc(x[13], s1, c1);
    x[6] = pc_mulc(x[6], h, h);                                    x[14] = pc_mulc(x[14], -h, h);
    { const p2 t = x[10].re; x[10].re = psub(0ull, x[10].im); x[10].im = t; }  // w16^4 = i
    x[7] = pc_mulc(x[7], s1, c1);  x[11] = pc_mulc(x[11], -h, h); x[15] = pc_mulc(x[15], -c1, -s1);
#pragma unroll
    for (int q2 = 0; q2 < 4; ++q2) dft4p(x[4 * q2], x[4 * q2 + 1], x[4 * q2 + 2], x[4 * q2 + 3]);  // x[4 q2 + q1] = y[q2 + 4 q1]
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = a + 1; b < 4; ++b) {
        const pc t = x[4 * a + b];
        x[4 * a + b] = x[4 * b + a];
        x[4 * b + a] = t;
      }
  }
};
template <int R> struct DftPrimeP {
  static __device__ __forceinline__ void run(pc (&x)[R], const float (&c)[R], const float (&s)[R]) {
    pc y[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      pc acc = x[0];
#pragma unroll
      for (int r = 1; r < R; ++r) acc = pc_add(acc, pc_mulc(x[r], c[(q * r) % R], s[(q * r) % R]));
      y[q] = acc;
    }
#pragma unroll
    for (int q = 0; q < R; ++q) x[q] = y[q];
  }
};
template <> struct DftP<5> {
  static __device__ __forceinline__ void run(pc (&x)[5]) {
    const float c[5] = {1.f, 0.30901699437494742410f, -0.80901699437494742410f, -0.80901699437494742410f, 0.30901699437494742410f};
    const float s[5] = {0.f, 0.95105651629515357212f, 0.58778525229247312917f, -0.58778525229247312917f, -0.95105651629515357212f};
    DftPrimeP<5>::run(x, c, s);
  }
};
template <> struct DftP<7> {
  static __device__ __forceinline__ void run(pc (&x)[7]) {
    const float c[7] = {1.f, 0.62348980185873353053f, -0.22252093395631440429f, -0.90096886790241912624f,
                        -0.90096886790241912624f, -0.22252093395631440429f, 0.62348980185873353053f};
    const float s[7] = {0.f, 0.78183148246802980871f, 0.97492791218182360702f, 0.43388373911755812048f,
                        -0.43388373911755812048f, -0.97492791218182360702f, -0.78183148246802980871f};
    DftPrimeP<7>::run(x, c, s);
  }
};

// ---- asynchronous copies ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- column passes ---------------------------------------------------------------------------------------------------
// A tile is n points of P column pairs (P = 8, 16 or 32: 128 to 512 bytes per row).  The work area G holds the cropped
// rows as granules of two columns, float4 (re0, re1, im0, im1), `gpitch` granules per row: the layout of the packed
// arithmetic, so that tiles go to shared memory with plain 16-byte asynchronous copies.  AOS: the tile came straight from
// a plane, granules (re0, im0, re1, im1) -- the first stage (or, without stages, the output loop) reads it that way.
// One stage; tab: float2 (cos, sin) of w_L^{qk} at [(q - 1) sub + k].
template <int R, int P, bool AOS>
__device__ __forceinline__ void cols_stage(float4* sm, int n, int L, const float2* __restrict__ tab) {
  const int sub = L / R, total = (n / R) * P;
  for (int t = threadIdx.x; t < total; t += blockDim.x) {
    const int p = t % P, bf = t / P;
    const int block = bf / sub, k = bf - block * sub;
    float4* at = sm + (size_t)(block * L + k) * P + p;
    pc x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float4 v = at[r * sub * P];
      x[r].re = AOS ? pk(v.x, v.z) : pk(v.x, v.y);
      x[r].im = AOS ? pk(v.y, v.w) : pk(v.z, v.w);
    }
    DftP<R>::run(x);
    if (sub > 1 && k != 0) {
#pragma unroll
      for (int q = 1; q < R; ++q) {
        const float2 w = __ldg(tab + (q - 1) * sub + k);
        x[q] = pc_mulc(x[q], w.x, w.y);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) {
      float4 v;
      upk(x[q].re, v.x, v.y);
      upk(x[q].im, v.z, v.w);
      at[q * sub * P] = v;
    }
  }
}

template <int P, bool AOS>
__device__ __forceinline__ void cols_transform(float4* sm, const Stages& st, const float2* __restrict__ tabs) {
  int L = st.n;
  for (int i = 0; i < st.count; ++i) {
    const int R = st.radix[i];
    const float2* tab = tabs + st.toff[i];
    if (AOS && i == 0) {
      switch (R) {
        case 16: cols_stage<16, P, true>(sm, st.n, L, tab); break;
        case 8: cols_stage<8, P, true>(sm, st.n, L, tab); break;
        case 4: cols_stage<4, P, true>(sm, st.n, L, tab); break;
        case 2: cols_stage<2, P, true>(sm, st.n, L, tab); break;
        case 3: cols_stage<3, P, true>(sm, st.n, L, tab); break;
        case 5: cols_stage<5, P, true>(sm, st.n, L, tab); break;
        default: cols_stage<7, P, true>(sm, st.n, L, tab); break;
      }
    } else {
      switch (R) {
        case 16: cols_stage<16, P, false>(sm, st.n, L, tab); break;
        case 8: cols_stage<8, P, false>(sm, st.n, L, tab); break;
        case 4: cols_stage<4, P, false>(sm, st.n, L, tab); break;
        case 2: cols_stage<2, P, false>(sm, st.n, L, tab); break;
        case 3: cols_stage<3, P, false>(sm, st.n, L, tab); break;
        case 5: cols_stage<5, P, false>(sm, st.n, L, tab); break;
        default: cols_stage<7, P, false>(sm, st.n, L, tab); break;
      }
    }
    L /= R;
    __syncthreads();
  }
}

// nrows rows (row stride rstep granules) of one P-granule column block -> shared memory, asynchronously
template <int P>
__device__ __forceinline__ void cols_prefetch(float4* sm, const float4* __restrict__ g, int nrows, size_t rstep) {
  const int p = threadIdx.x % P;
  for (int k = threadIdx.x / P; k < nrows; k += blockDim.x / P) cp_async16(sm + k * P + p, g + (size_t)k * rstep + p);
  cp_async_commit();
}

// Where pass B1 reads: the work area G (granules; in place) or, AOS, a plane whose rows have been transformed along v
// (the retained columns jp < ny are the plane columns (jp - ny/2) mod nv; ny is a multiple of 4 so that no pair wraps).
struct Cols1Src {
  const float2* plane;  // AOS only
  int nv, ny;
  int keep_lo, keep_hi;  // AOS only: rows [keep_lo, keep_hi) of the plane are known to be zero and are not read
};

// Pass B1, persistent: tile = (column block cb, residue r); see pfft.cu.
template <int P, bool AOS>
__global__ void __launch_bounds__(128, 4) cols1_kernel_p(float4* __restrict__ g, Cols1Src src, Stages st, const float2* __restrict__ tabs,
                                                         const int* __restrict__ perm, const float2* __restrict__ twu, int n2,
                                                         int ncb, int gpitch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n1 = st.n;
  float4* buf[2] = {reinterpret_cast<float4*>(smem_raw), reinterpret_cast<float4*>(smem_raw) + (size_t)n1 * P};
  const size_t rstep = (size_t)n2 * gpitch;
  const int ntiles = ncb * n2;
  auto tile_base = [&](int tile) { return g + (size_t)(tile / ncb) * gpitch + (size_t)(tile % ncb) * P; };
  auto prefetch = [&](float4* sm, int tile) {
    if (AOS) {
      const int p = threadIdx.x % P, jp = (tile % ncb) * 2 * P + 2 * p, r = tile / ncb;
      int c = jp - src.ny / 2;
      c += c < 0 ? src.nv : 0;
      const float2* from = src.plane + (size_t)r * src.nv + c;
      const size_t step = (size_t)n2 * src.nv;
      if (jp < src.ny) {
        for (int k = threadIdx.x / P; k < n1; k += blockDim.x / P) {
          const int row = r + k * n2;
          if (row < src.keep_lo || row >= src.keep_hi) cp_async16(sm + k * P + p, from + (size_t)k * step);
          else sm[k * P + p] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      } else {
        for (int k = threadIdx.x / P; k < n1; k += blockDim.x / P) sm[k * P + p] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
      cp_async_commit();
    } else {
      cols_prefetch<P>(sm, tile_base(tile), n1, rstep);
    }
  };
  int tile = blockIdx.x, it = 0;
  if (tile < ntiles) prefetch(buf[0], tile);
  for (; tile < ntiles; tile += gridDim.x, ++it) {
    float4* sm = buf[it & 1];
    const int next = tile + gridDim.x;
    if (next < ntiles) {
      prefetch(buf[(it + 1) & 1], next);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    cols_transform<P, AOS>(sm, st, tabs);
    const bool aos_out = AOS && st.count == 0;
    const int r = tile / ncb, p = threadIdx.x % P;
    float4* base = tile_base(tile) + p;
    for (int q = threadIdx.x / P; q < n1; q += blockDim.x / P) {
      float4 v = sm[__ldg(perm + q) * P + p];  // re0 re1 im0 im1
      if (aos_out) v = make_float4(v.x, v.z, v.y, v.w);
      float2 w = make_float2(1.f, 0.f);
      if (r != 0 && q != 0) w = __ldg(twu + r * q);
      float4 o;
      o.x = v.x * w.x - v.z * w.y; o.y = v.y * w.x - v.w * w.y;
      o.z = v.x * w.y + v.z * w.x; o.w = v.y * w.y + v.w * w.x;
      base[(size_t)q * rstep] = o;
    }
    __syncthreads();  // the buffer is refilled two iterations later, but its prefetch is issued in the next one
  }
}

// Pass B2, persistent: tile = (column block cb, q).
template <int P, bool SCREEN>
__global__ void __launch_bounds__(128, 4) cols2_kernel_p(const float4* __restrict__ g, Stages st, const float2* __restrict__ tabs,
                                                         const int* __restrict__ perm, int n1, int ncb, int gpitch, ScreenArgs a,
                                                         void* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (SCREEN && *a.skip) return;
  const int n2 = st.n;
  float4* buf[2] = {reinterpret_cast<float4*>(smem_raw), reinterpret_cast<float4*>(smem_raw) + (size_t)n2 * P};
  const int ntiles = ncb * n1;
  auto tile_base = [&](int tile) { return g + (size_t)(tile / ncb) * n2 * gpitch + (size_t)(tile % ncb) * P; };
  int tile = blockIdx.x, it = 0;
  if (tile < ntiles) cols_prefetch<P>(buf[0], tile_base(tile), n2, (size_t)gpitch);
  for (; tile < ntiles; tile += gridDim.x, ++it) {
    float4* sm = buf[it & 1];
    const int next = tile + gridDim.x;
    if (next < ntiles) {
      cols_prefetch<P>(buf[(it + 1) & 1], tile_base(next), n2, (size_t)gpitch);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    cols_transform<P, false>(sm, st, tabs);
    const int q = tile / ncb, p = threadIdx.x % P, col = (tile % ncb) * 2 * P + 2 * p;
    if (col < a.ny) {
      const double m0 = (col - a.ny / 2) * a.py, m1 = (col + 1 - a.ny / 2) * a.py;
      for (int sp = threadIdx.x / P; sp < n2; sp += blockDim.x / P) {
        const int f = q + n1 * sp;
        int i;
        if (f < a.nx / 2) i = f + a.nx / 2;
        else if (f >= a.nu - a.nx / 2) i = f - (a.nu - a.nx / 2);
        else continue;
        const float4 v = sm[__ldg(perm + sp) * P + p];  // re0 re1 im0 im1
        if constexpr (SCREEN) {
          float* px = static_cast<float*>(out) + (size_t)i * a.ny + col;
          const double l = (i - a.nx / 2) * a.px, ll = l * l;
          const double e0 = ll + m0 * m0, e1 = ll + m1 * m1;
          float2 acc = make_float2(0.f, 0.f);
          // phase w (n - 1 + s) in turns, n - 1 = sqrt(1 - e) - 1: the cancellation costs absolute, not relative
          // accuracy (1e-16 turns per wavelength of w), which is all a phase needs
          if (e0 < 1.0) {
            double ph = a.wp * (a.do_wstacking ? sqrt(1.0 - e0) - 1.0 + a.nshift : 0.0);
            ph -= rint(ph);
            float sn, cs;
            sincospif(2.0f * (float)ph, &sn, &cs);
            acc.x = v.x * cs + v.z * sn;
          }
          if (e1 < 1.0) {
            double ph = a.wp * (a.do_wstacking ? sqrt(1.0 - e1) - 1.0 + a.nshift : 0.0);
            ph -= rint(ph);
            float sn, cs;
            sincospif(2.0f * (float)ph, &sn, &cs);
            acc.y = v.y * cs + v.w * sn;
          }
          // one pixel pair per thread and plane, planes in stream order: the reduction only spares the thread the
          // round trip of reading the image (the sums stay reproducible)
          asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(px), "f"(acc.x), "f"(acc.y) : "memory");
        } else {
          *reinterpret_cast<float4*>(static_cast<float2*>(out) + (size_t)i * a.ny + col) = make_float4(v.x, v.z, v.y, v.w);
        }
      }
    }
    __syncthreads();
  }
}

// ---- row pass ------------------------------------------------------------------------------------------------------
// Shared memory: the row as granules of two neighbouring points, float4 (re0, re1, im0, im1).  Stage with sub = L / R
// even: the butterflies k, k + 1 of a block are the two halves of the granules (base + r sub) / 2.  tab: cos at
// [(q - 1) sub + k], sin (R - 1) sub further.
// SWZ (the stage before the last): the outputs are the operands of the last stage, whose butterfly bf = block R + q
// reads its rlast / 2 granules with one thread; they are stored at bf rlast / 2 + (j ^ H), H = the low bits of the
// frequency F(bf) = fblock[block] + wf q the butterfly produces -- so that the last stage, whose lanes take
// consecutive F (coalesced stores to global memory), reads shared memory without bank conflicts.
template <int R, bool SWZ>
__device__ __forceinline__ void rows_stage(float4* sm, int n, int L, const float* __restrict__ tab, const int* __restrict__ fblock,
                                           int wf) {
  const int sub = L / R, hs = sub >> 1, npairs = n / R / 2;
  const float* tsin = tab + (R - 1) * sub;
  for (int t0 = 0; t0 < npairs; t0 += blockDim.x) {
    const bool active = t0 + (int)threadIdx.x < npairs;
    const int t = active ? t0 + threadIdx.x : npairs - 1;  // idle threads redo the last pair and store nothing
    const int bf = 2 * t;
    const int block = bf / sub, k = bf - block * sub;
    float4* at = sm + ((block * L + k) >> 1);
    pc x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float4 v = at[r * hs];
      x[r].re = pk(v.x, v.y);
      x[r].im = pk(v.z, v.w);
    }
    DftP<R>::run(x);
#pragma unroll
    for (int q = 1; q < R; ++q) {
      const p2 c = __ldg(reinterpret_cast<const p2*>(tab + (q - 1) * sub + k));
      const p2 s = __ldg(reinterpret_cast<const p2*>(tsin + (q - 1) * sub + k));
      x[q] = pc_mulp(x[q], c, s);
    }
    if (SWZ) {
      // The swizzled positions are the operands of OTHER threads of the same block: everybody reads first.  (The
      // threads of one pass of this loop cover whole blocks: blockDim is a multiple of the hs pairs of a block.)
      __syncthreads();
      if (active) {
        const int fb = __ldg(fblock + block), j = k >> 1, mask = hs - 1;
        float4* ob = sm + (size_t)block * R * hs;
#pragma unroll
        for (int q = 0; q < R; ++q) {
          float4 v;
          upk(x[q].re, v.x, v.y);
          upk(x[q].im, v.z, v.w);
          ob[q * hs + (j ^ ((fb + wf * q) & mask))] = v;
        }
      }
    } else if (active) {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        float4 v;
        upk(x[q].re, v.x, v.y);
        upk(x[q].im, v.z, v.w);
        at[q * hs] = v;
      }
    }
  }
}

// Last stage (L = R, no twiddles), straight to global memory: lane <-> frequency F < n / R, butterfly bf_of_f[F], outputs
// F + q n / R; only the retained columns are stored, in image order, into the granule layout of G.
template <int R>
__device__ __forceinline__ void rows_final(const float4* sm, int n, const int* __restrict__ bf_of_f, bool swz, float* __restrict__ dst,
                                           int ny) {
  const int nbf = n / R;
  constexpr int HG = R / 2;
  for (int F = threadIdx.x; F < nbf; F += blockDim.x) {
    const int bf = __ldg(bf_of_f + F);
    const int H = swz ? (F & (HG - 1)) : 0;
    float2 x[R];
#pragma unroll
    for (int j = 0; j < HG; ++j) {
      const float4 v = sm[bf * HG + (j ^ H)];
      x[2 * j] = make_float2(v.x, v.z);
      x[2 * j + 1] = make_float2(v.y, v.w);
    }
    Dft<float, R>::run(x);
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const int f = F + q * nbf;
      int jp;
      if (f < ny / 2) jp = f + ny / 2;
      else if (f >= n - ny / 2) jp = f - (n - ny / 2);
      else continue;
      float* o = dst + (jp >> 1) * 4 + (jp & 1);
      o[0] = x[q].x;
      o[2] = x[q].y;
    }
  }
}

struct RowsAux {
  const float* tabs;    // stage tables
  const int* fblock;    // stage before the last: frequency contribution of a block
  const int* bf_of_f;   // last stage: butterfly that produces the frequencies F + q n / R
  int wf;               // weight of the digit of the stage before the last
};

template <int R>
__device__ __forceinline__ void rows_any_stage(float4* sm, int n, int L, const float* __restrict__ tab, bool swz, const RowsAux& aux) {
  if (swz) rows_stage<R, true>(sm, n, L, tab, aux.fblock, aux.wf);
  else rows_stage<R, false>(sm, n, L, tab, aux.fblock, aux.wf);
}

template <int NT>
__global__ void __launch_bounds__(NT, NT == 256 ? 2 : 1) rows_kernel_p(const float2* __restrict__ planes, float* __restrict__ g,
                                                                        Stages st, RowsAux aux, int nv, int ny, int gpitch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float4* sm = reinterpret_cast<float4*>(smem_raw);
  const float4* src4 = reinterpret_cast<const float4*>(planes + (size_t)blockIdx.x * nv);  // nv is even
  for (int e2 = threadIdx.x; e2 < nv / 2; e2 += NT) {
    const float4 v = __ldcs(src4 + e2);
    sm[e2] = make_float4(v.x, v.z, v.y, v.w);
  }
  __syncthreads();
  int L = nv;
  for (int i = 0; i + 1 < st.count; ++i) {
    const int R = st.radix[i];
    const float* tab = aux.tabs + st.toff[i];
    const bool swz = i + 2 == st.count;
    switch (R) {
      case 16: rows_any_stage<16>(sm, nv, L, tab, swz, aux); break;
      case 8: rows_any_stage<8>(sm, nv, L, tab, swz, aux); break;
      case 4: rows_any_stage<4>(sm, nv, L, tab, swz, aux); break;
      case 2: rows_any_stage<2>(sm, nv, L, tab, swz, aux); break;
      case 3: rows_any_stage<3>(sm, nv, L, tab, swz, aux); break;
      case 5: rows_any_stage<5>(sm, nv, L, tab, swz, aux); break;
      default: rows_any_stage<7>(sm, nv, L, tab, swz, aux); break;
    }
    L /= R;
    __syncthreads();
  }
  float* dst = g + (size_t)blockIdx.x * gpitch * 4;
  const bool swz = st.count >= 2;
  switch (L) {  // = the last radix, a power of two
    case 16: rows_final<16>(sm, nv, aux.bf_of_f, swz, dst, ny); break;
    case 8: rows_final<8>(sm, nv, aux.bf_of_f, swz, dst, ny); break;
    case 4: rows_final<4>(sm, nv, aux.bf_of_f, swz, dst, ny); break;
    default: rows_final<2>(sm, nv, aux.bf_of_f, swz, dst, ny); break;
  }
}

}  // namespace
}  // namespace cip

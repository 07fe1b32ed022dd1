// Kernel 3, own transform: the per-plane 2-D FFT of stage G3 fused with the crop and the w-screen of stage G4
// (inside ducc0.wgridder.ms2dirty in the reference; call site src/ska_sdp_cip/invert.py:170-183).
//
// The image keeps nx x ny of the nu x nv cells of a transformed plane, so the transform is pruned and never
// materialised:
//
//   pass A   the transform along v of every grid row.  Default (float32, image width a multiple of 4): cuFFT's
//            contiguous 1-D transform in place -- its single-pass register-resident kernel runs at 4.4 TB/s on 6144-point
//            rows, the own row pass below reaches half of that.  CIP_FFT=own / float64: rows_kernel(_p), one CTA per row,
//            the whole row in shared memory, only the ny retained columns written, in image order, to the work area G.
//   pass B1  (cols1_kernel) the transform along u as a four-step transform nu = n1 n2 on blocks of 16 to 64 neighbouring
//            retained columns: for every residue r the n1-point transform over the rows r + n2 k, times the step
//            twiddle; reads the plane (default) or G, writes G.
//   pass B2  (cols2_kernel) for every q the n2-point transform over the n2 contiguous rows n2 q ..., whose outputs are
//            the frequencies q + n1 s; the retained ones are multiplied with the w-screen of the plane and their real
//            part is added to the image -- the transformed plane is never written.
//
// Per 6144^2 plane and a 4096^2 image: 604 MB (cuFFT rows) + 201 + 201 (B1) + 201 + 2 x 67 (B2) = 1.34 GB in three
// launches against 1.68 GB in four for cuFFT rows / transpose-crop / cuFFT rows / w-screen; measured 0.38 against
// 0.43 ms per plane (configs[1]), 7.0 against 8.9 ms at 24576^2 (configs[3]).
//
// The own passes use an in-place mixed-radix decimation-in-frequency transform in shared memory (radices 16, 8, 4, 2, 3,
// 5, 7: every size good_size() produces): stage of span L and radix R:
//     y_q[k] = (sum_r x[k + r L/R] w_R^{qr}) w_L^{qk}   stored at q L/R + k,   w_m = exp(+2 pi i / m)
// after which position sum_s d_s n / (R_1 .. R_s) holds frequency d_1 + R_1 d_2 + R_1 R_2 d_3 + ... (a table).  The
// float32 passes (pfft_f32.cuh) work on packed pairs with per-stage twiddle tables; the float64 passes below are the
// plain first version (one table of exp(2 pi i t / n) per dimension, powers multiplied up).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#include "common.h"

namespace cip {

namespace {

template <typename T> struct Cx;
template <> struct Cx<float> { using type = float2; };
template <> struct Cx<double> { using type = double2; };
template <typename T> using cx = typename Cx<T>::type;

template <typename C> __device__ __forceinline__ C cadd(C a, C b) { C r; r.x = a.x + b.x; r.y = a.y + b.y; return r; }
template <typename C> __device__ __forceinline__ C csub(C a, C b) { C r; r.x = a.x - b.x; r.y = a.y - b.y; return r; }
template <typename C> __device__ __forceinline__ C cmul(C a, C b) { C r; r.x = a.x * b.x - a.y * b.y; r.y = a.x * b.y + a.y * b.x; return r; }
template <typename C> __device__ __forceinline__ C mul_i(C a) { C r; r.x = -a.y; r.y = a.x; return r; }
template <typename C, typename T> __device__ __forceinline__ C cmulc(C a, T c, T s) { C r; r.x = a.x * c - a.y * s; r.y = a.x * s + a.y * c; return r; }

// ---- small transforms, in place, e^{+2 pi i q r / R} ------------------------------------------------------
template <typename C> __device__ __forceinline__ void dft2(C& a, C& b) { const C t = csub(a, b); a = cadd(a, b); b = t; }
template <typename C> __device__ __forceinline__ void dft4(C& x0, C& x1, C& x2, C& x3) {
  const C t0 = cadd(x0, x2), t1 = csub(x0, x2), t2 = cadd(x1, x3), t3 = mul_i(csub(x1, x3));
  x0 = cadd(t0, t2); x1 = cadd(t1, t3); x2 = csub(t0, t2); x3 = csub(t1, t3);
}

template <typename T, int R> struct Dft;
template <typename T> struct Dft<T, 2> { static __device__ __forceinline__ void run(cx<T> (&x)[2]) { dft2(x[0], x[1]); } };
template <typename T> struct Dft<T, 4> { static __device__ __forceinline__ void run(cx<T> (&x)[4]) { dft4(x[0], x[1], x[2], x[3]); } };
template <typename T> struct Dft<T, 3> {
  static __device__ __forceinline__ void run(cx<T> (&x)[3]) {
    const T s3 = (T)0.86602540378443864676;
    const cx<T> t = cadd(x[1], x[2]), d = csub(x[1], x[2]);
    cx<T> m; m.x = x[0].x - (T)0.5 * t.x; m.y = x[0].y - (T)0.5 * t.y;
    cx<T> e; e.x = -s3 * d.y; e.y = s3 * d.x;  // i s3 d
    x[0] = cadd(x[0], t); x[1] = cadd(m, e); x[2] = csub(m, e);
  }
};
template <typename T> struct Dft<T, 8> {
  static __device__ __forceinline__ void run(cx<T> (&x)[8]) {
    const T h = (T)0.70710678118654752440;
    dft4(x[0], x[2], x[4], x[6]);
    dft4(x[1], x[3], x[5], x[7]);
    cx<T> o1, o3;
    o1.x = h * (x[3].x - x[3].y); o1.y = h * (x[3].x + x[3].y);    // w8   * O1
    o3.x = -h * (x[7].x + x[7].y); o3.y = h * (x[7].x - x[7].y);   // w8^3 * O3
    const cx<T> o0 = x[1], o2 = mul_i(x[5]);
    const cx<T> e0 = x[0], e1 = x[2], e2 = x[4], e3 = x[6];
    x[0] = cadd(e0, o0); x[4] = csub(e0, o0);
    x[1] = cadd(e1, o1); x[5] = csub(e1, o1);
    x[2] = cadd(e2, o2); x[6] = csub(e2, o2);
    x[3] = cadd(e3, o3); x[7] = csub(e3, o3);
  }
};
template <typename T> struct Dft<T, 16> {
  static __device__ __forceinline__ void run(cx<T> (&x)[16]) {
    const T c1 = (T)0.92387953251128675613, s1 = (T)0.38268343236508977173, h = (T)0.70710678118654752440;
#pragma unroll
    for (int r1 = 0; r1 < 4; ++r1) dft4(x[r1], x[r1 + 4], x[r1 + 8], x[r1 + 12]);  // x[r1 + 4 q2] = A_r1[q2]
    // A_r1[q2] *= w16^{r1 q2}
    x[5] = cmulc(x[5], c1, s1);   x[9] = cmulc(x[9], h, h);     x[13] = cmulc(x[13], s1, c1);   // r1 = 1: m = 1, 2, 3
    x[6] = cmulc(x[6], h, h);     x[10] = mul_i(x[10]);         x[14] = cmulc(x[14], -h, h);    // r1 = 2: m = 2, 4, 6
    x[7] = cmulc(x[7], s1, c1);   x[11] = cmulc(x[11], -h, h);  x[15] = cmulc(x[15], -c1, -s1); // r1 = 3: m = 3, 6, 9
#pragma unroll
    for (int q2 = 0; q2 < 4; ++q2) dft4(x[4 * q2], x[4 * q2 + 1], x[4 * q2 + 2], x[4 * q2 + 3]);  // x[4 q2 + q1] = y[q2 + 4 q1]
    // transpose the 4 x 4 register array
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = a + 1; b < 4; ++b) {
        const cx<T> t = x[4 * a + b];
        x[4 * a + b] = x[4 * b + a];
        x[4 * b + a] = t;
      }
  }
};
// odd primes beyond 3: direct evaluation (rare sizes)
template <typename T, int R> struct DftPrime {
  static __device__ __forceinline__ void run(cx<T> (&x)[R], const T (&c)[R], const T (&s)[R]) {
    cx<T> y[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      cx<T> acc = x[0];
#pragma unroll
      for (int r = 1; r < R; ++r) acc = cadd(acc, cmulc(x[r], c[(q * r) % R], s[(q * r) % R]));
      y[q] = acc;
    }
#pragma unroll
    for (int q = 0; q < R; ++q) x[q] = y[q];
  }
};
template <typename T> struct Dft<T, 5> {
  static __device__ __forceinline__ void run(cx<T> (&x)[5]) {
    const T c[5] = {(T)1, (T)0.30901699437494742410, (T)-0.80901699437494742410, (T)-0.80901699437494742410, (T)0.30901699437494742410};
    const T s[5] = {(T)0, (T)0.95105651629515357212, (T)0.58778525229247312917, (T)-0.58778525229247312917, (T)-0.95105651629515357212};
    DftPrime<T, 5>::run(x, c, s);
  }
};
template <typename T> struct Dft<T, 7> {
  static __device__ __forceinline__ void run(cx<T> (&x)[7]) {
    const T c[7] = {(T)1, (T)0.62348980185873353053, (T)-0.22252093395631440429, (T)-0.90096886790241912624,
                    (T)-0.90096886790241912624, (T)-0.22252093395631440429, (T)0.62348980185873353053};
    const T s[7] = {(T)0, (T)0.78183148246802980871, (T)0.97492791218182360702, (T)0.43388373911755812048,
                    (T)-0.43388373911755812048, (T)-0.97492791218182360702, (T)-0.78183148246802980871};
    DftPrime<T, 7>::run(x, c, s);
  }
};

// ---- shared-memory layouts ----------------------------------------------------------------------------------
// C = 1: one sequence, one pad element per 16 (a stride of 16 elements becomes 17: the last stages are conflict-free);
// C = 16: 16 interleaved sequences, element e of sequence s at 16 e + s (a half-warp reads one 128-byte line).
template <int C> __device__ __forceinline__ int sidx(int e, int s) { return C == 1 ? e + (e >> 4) : e * C + s; }
__host__ __device__ inline size_t padded_len(size_t n) { return n + (n >> 4) + 1; }

constexpr int kMaxStages = 16;
struct Stages {
  int n = 0, count = 0;
  int radix[kMaxStages] = {};
  int toff[kMaxStages] = {};  // float32 passes: where the twiddle table of the stage starts (pfft_f32.cuh)
};

// One stage: n points per sequence, span L, radix R; table tw of exp(2 pi i t / N) with w_n = tw[tstep].
template <typename T, int R, int C>
__device__ __forceinline__ void stage(cx<T>* sm, int n, int L, const cx<T>* __restrict__ tw, int tstep) {
  const int sub = L / R, total = (n / R) * C;
  const int tl = tstep * (n / L);  // w_L = tw[tl]
  for (int t = threadIdx.x; t < total; t += blockDim.x) {
    const int s = C == 1 ? 0 : t % C, bf = C == 1 ? t : t / C;
    const int block = bf / sub, k = bf - block * sub;
    const int base = block * L + k;
    cx<T> x[R];
#pragma unroll
    for (int r = 0; r < R; ++r) x[r] = sm[sidx<C>(base + r * sub, s)];
    Dft<T, R>::run(x);
    if (sub > 1 && k != 0) {
      // y_q *= w_L^{q k}: w^1 from the table, w^2, w^4, w^8 by squaring, the others as products of those (few live
      // registers, at most six roundings)
      cx<T> wp[4];
      wp[0] = __ldg(tw + k * tl);
#pragma unroll
      for (int b = 1; b < 4; ++b)
        if ((1 << b) < R) wp[b] = cmul(wp[b - 1], wp[b - 1]);
#pragma unroll
      for (int q = 1; q < R; ++q) {
        cx<T> w;
        bool have = false;
#pragma unroll
        for (int b = 3; b >= 0; --b)
          if (q & (1 << b)) {
            w = have ? cmul(w, wp[b]) : wp[b];
            have = true;
          }
        x[q] = cmul(x[q], w);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) sm[sidx<C>(base + q * sub, s)] = x[q];
  }
}

template <typename T, int C>
__device__ __forceinline__ void transform(cx<T>* sm, const Stages& st, const cx<T>* __restrict__ tw, int tstep) {
  int L = st.n;
  for (int i = 0; i < st.count; ++i) {
    const int R = st.radix[i];
    switch (R) {
      case 16: stage<T, 16, C>(sm, st.n, L, tw, tstep); break;
      case 8: stage<T, 8, C>(sm, st.n, L, tw, tstep); break;
      case 4: stage<T, 4, C>(sm, st.n, L, tw, tstep); break;
      case 2: stage<T, 2, C>(sm, st.n, L, tw, tstep); break;
      case 3: stage<T, 3, C>(sm, st.n, L, tw, tstep); break;
      case 5: stage<T, 5, C>(sm, st.n, L, tw, tstep); break;
      default: stage<T, 7, C>(sm, st.n, L, tw, tstep); break;
    }
    L /= R;
    __syncthreads();
  }
}

// ---- pass A ---------------------------------------------------------------------------------------------------
// G[i][jp] = sum_j F[i][j] exp(+2 pi i j f / nv), f = (jp - ny/2) mod nv, for the ny retained columns.
template <typename T, int NT>
__global__ void __launch_bounds__(NT, NT == 256 ? (sizeof(T) == 4 ? 4 : 2) : 1) rows_kernel(const cx<T>* __restrict__ planes, cx<T>* __restrict__ g, Stages st,
                                                   const cx<T>* __restrict__ tw, const int* __restrict__ perm, int nv, int ny,
                                                   int pitch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cx<T>* sm = reinterpret_cast<cx<T>*>(smem_raw);
  const cx<T>* src = planes + (size_t)blockIdx.x * nv;
  if constexpr (sizeof(T) == 4) {
    const float4* src4 = reinterpret_cast<const float4*>(src);  // nv is even, rows are 16-byte aligned
    for (int e2 = threadIdx.x; e2 < nv / 2; e2 += blockDim.x) {
      const float4 v = __ldcs(src4 + e2);
      sm[sidx<1>(2 * e2, 0)] = make_float2(v.x, v.y);
      sm[sidx<1>(2 * e2 + 1, 0)] = make_float2(v.z, v.w);
    }
  } else {
    for (int e = threadIdx.x; e < nv; e += blockDim.x) sm[sidx<1>(e, 0)] = __ldcs(src + e);
  }
  __syncthreads();
  transform<T, 1>(sm, st, tw, 1);
  cx<T>* dst = g + (size_t)blockIdx.x * pitch;
  for (int jp = threadIdx.x; jp < ny; jp += blockDim.x) {
    int f = jp - ny / 2;
    f += f < 0 ? nv : 0;
    dst[jp] = sm[sidx<1>(__ldg(perm + f), 0)];
  }
}

// ---- pass B1 --------------------------------------------------------------------------------------------------
// For column block blockIdx.x and residue r = blockIdx.y:  Z[r][q] = w_nu^{q r} sum_k G[r + n2 k] w_n1^{q k} -> row r + n2 q.
template <typename T>
__global__ void __launch_bounds__(128, sizeof(T) == 4 ? 8 : 4) cols1_kernel(cx<T>* __restrict__ g, Stages st, const cx<T>* __restrict__ tw,
                                                    const int* __restrict__ perm, int n2, int ny, int pitch) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cx<T>* sm = reinterpret_cast<cx<T>*>(smem_raw);
  const int n1 = st.n, r = blockIdx.y, col0 = blockIdx.x * 16;
  const int s = threadIdx.x & 15;
  const bool in = col0 + s < ny;
  cx<T>* base = g + (size_t)r * pitch + col0 + s;
  const size_t rstep = (size_t)n2 * pitch;
  cx<T> zero;
  zero.x = zero.y = 0;
  for (int k = threadIdx.x >> 4; k < n1; k += blockDim.x >> 4) sm[k * 16 + s] = in ? base[k * rstep] : zero;
  __syncthreads();
  transform<T, 16>(sm, st, tw, n2);  // w_n1 = w_nu^{n2}
  if (!in) return;
  for (int q = threadIdx.x >> 4; q < n1; q += blockDim.x >> 4) {
    cx<T> v = sm[__ldg(perm + q) * 16 + s];
    if (r != 0 && q != 0) v = cmul(v, __ldg(tw + r * q));
    base[q * rstep] = v;
  }
}

// ---- pass B2 --------------------------------------------------------------------------------------------------
struct ScreenArgs {
  int nx, ny, nu, nv;
  double px, py, wp, nshift;  // wp: w of the plane
  int do_wstacking;
  const uint32_t* skip;  // != 0 on the device: a binning pass of this call overflowed, the image must stay untouched
};

}  // namespace
}  // namespace cip
#include "pfft_f32.cuh"  // the float32 passes, on packed pairs
namespace cip {
namespace {

// For column block blockIdx.x and q = blockIdx.y: X[q + n1 s'] = sum_r Z[r][q] w_n2^{s' r}; the retained frequencies
// (image rows) go through the w-screen into the image (SCREEN) or are written as they are (tests).
template <typename T, bool SCREEN>
__global__ void __launch_bounds__(128, sizeof(T) == 4 ? 8 : 4) cols2_kernel(const cx<T>* __restrict__ g, Stages st, const cx<T>* __restrict__ tw,
                                                    const int* __restrict__ perm, int n1, int pitch, ScreenArgs p,
                                                    void* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  cx<T>* sm = reinterpret_cast<cx<T>*>(smem_raw);
  if (SCREEN && *p.skip) return;
  const int n2 = st.n, q = blockIdx.y, col0 = blockIdx.x * 16;
  const int s = threadIdx.x & 15, col = col0 + s;
  const bool in = col < p.ny;
  const cx<T>* base = g + (size_t)q * n2 * pitch + col;
  cx<T> zero;
  zero.x = zero.y = 0;
  for (int r = threadIdx.x >> 4; r < n2; r += blockDim.x >> 4) sm[r * 16 + s] = in ? __ldcs(base + (size_t)r * pitch) : zero;
  __syncthreads();
  transform<T, 16>(sm, st, tw, n1);  // w_n2 = w_nu^{n1}
  if (!in) return;
  const double m = (col - p.ny / 2) * p.py;
  for (int sp = threadIdx.x >> 4; sp < n2; sp += blockDim.x >> 4) {
    const int f = q + n1 * sp;
    int i;
    if (f < p.nx / 2) i = f + p.nx / 2;
    else if (f >= p.nu - p.nx / 2) i = f - (p.nu - p.nx / 2);
    else continue;
    const cx<T> v = sm[__ldg(perm + sp) * 16 + s];
    if constexpr (SCREEN) {
      T* image = static_cast<T*>(out);
      const double l = (i - p.nx / 2) * p.px, e = l * l + m * m;
      if (e >= 1.0) continue;
      const double z = p.do_wstacking ? -e / (sqrt(1.0 - e) + 1.0) + p.nshift : 0.0;
      double ph = p.wp * z;
      ph -= rint(ph);
      T val;
      if constexpr (sizeof(T) == 4) {
        float sn, cs;
        sincospif(2.0f * (float)ph, &sn, &cs);
        val = v.x * cs + v.y * sn;
      } else {
        double sn, cs;
        sincospi(2.0 * ph, &sn, &cs);
        val = v.x * cs + v.y * sn;
      }
      image[(size_t)i * p.ny + col] += val;
    } else {
      static_cast<cx<T>*>(out)[(size_t)i * p.ny + col] = v;
    }
  }
}

// ---- plans ----------------------------------------------------------------------------------------------------
Stages make_stages(int n) {
  Stages st;
  st.n = n;
  // odd radices first (largest spans), the widest power-of-two radices last: the final stage needs no twiddles
  std::vector<int> odd, pow2;
  for (int p : {3, 5, 7})
    while (n % p == 0) { odd.push_back(p); n /= p; }
  while (n % 16 == 0) { pow2.push_back(16); n /= 16; }
  for (int p : {8, 4, 2})
    if (n % p == 0) { pow2.insert(pow2.begin(), p); n /= p; }
  if (n != 1 || odd.size() + pow2.size() > (size_t)kMaxStages) {
    st.count = -1;
    return st;
  }
  for (int r : odd) st.radix[st.count++] = r;
  for (int r : pow2) st.radix[st.count++] = r;
  return st;
}

std::vector<int> make_perm(const Stages& st) {  // position of frequency f after the stages
  std::vector<int> perm(st.n);
  for (int f = 0; f < st.n; ++f) {
    int ff = f, p = 0, wt = st.n;
    for (int i = 0; i < st.count; ++i) {
      wt /= st.radix[i];
      p += (ff % st.radix[i]) * wt;
      ff /= st.radix[i];
    }
    perm[f] = p;
  }
  return perm;
}

// float32 stage tables: for every stage with sub = L / R > 1 the factors w_L^{qk}, q = 1 .. R - 1, k < sub.
// soa: cos block then sin block of floats (row pass, read as packed pairs over k); else float2 (cos, sin).
// Returns the table; sets st->toff (in floats resp. float2).
std::vector<float> make_stage_tables(Stages* st, bool soa) {
  std::vector<float> tab;
  int L = st->n;
  for (int i = 0; i < st->count; ++i) {
    const int R = st->radix[i], sub = L / R;
    st->toff[i] = (int)(soa ? tab.size() : tab.size() / 2);
    if (sub > 1) {
      const size_t cnt = (size_t)(R - 1) * sub, at = tab.size();
      tab.resize(at + 2 * cnt);
      for (int q = 1; q < R; ++q)
        for (int k = 0; k < sub; ++k) {
          const double ang = 2.0 * M_PI * (double)(((long long)q * k) % L) / (double)L;
          const size_t e = (size_t)(q - 1) * sub + k;
          if (soa) {
            tab[at + e] = (float)cos(ang);
            tab[at + cnt + e] = (float)sin(ang);
          } else {
            tab[at + 2 * e] = (float)cos(ang);
            tab[at + 2 * e + 1] = (float)sin(ang);
          }
        }
    }
    L /= R;
  }
  return tab;
}

// the packed row pass needs an even number of butterflies per block in every stage but the last
bool rows_pairable(const Stages& st) {
  int L = st.n;
  for (int i = 0; i < st.count; ++i) {
    const int sub = L / st.radix[i];
    if (sub > 1 && sub % 2) return false;
    L /= st.radix[i];
  }
  return st.count > 0;
}

template <typename T>
void fill_twiddles(std::vector<unsigned char>& buf, size_t off, int n) {
  cx<T>* t = reinterpret_cast<cx<T>*>(buf.data() + off);
  const double step = 2.0 * M_PI / (double)n;
  for (int k = 0; k < n; ++k) {
    // octant symmetry keeps the arguments small
    int kk = k % n;
    double c, s;
    const int oct = (int)((8LL * kk) / n);
    switch (oct) {
      case 0: c = cos(step * kk); s = sin(step * kk); break;
      case 1: case 2: { const double a = step * kk - M_PI / 2; c = -sin(a); s = cos(a); } break;
      case 3: case 4: { const double a = step * kk - M_PI; c = -cos(a); s = -sin(a); } break;
      case 5: case 6: { const double a = step * kk - 3 * M_PI / 2; c = sin(a); s = -cos(a); } break;
      default: { const double a = step * kk - 2 * M_PI; c = cos(a); s = sin(a); } break;
    }
    t[k].x = (T)c;
    t[k].y = (T)s;
  }
}

}  // namespace

// Tables of one (nu, nv, precision): twiddles of both dimensions and the output permutations of the three transforms.
struct PfftPlan {
  DevBuf buf;
  Stages sa, s1, s2;
  int n1 = 0, n2 = 0;
  size_t off_twv = 0, off_twu = 0, off_pa = 0, off_p1 = 0, off_p2 = 0, off_ta = 0, off_t1 = 0, off_t2 = 0, off_fb = 0, off_bf = 0;
  int wf = 1;
  size_t smem_a = 0;
  bool ok = false;       // the column passes are available (float32: cuFFT can do the row pass)
  bool ok_rows = false;  // the own row pass too
};

void pfft_release(cip_ctx* ctx) {
  for (auto& kv : ctx->pfft_plans) {
    kv.second->buf.release();
    delete kv.second;
  }
  ctx->pfft_plans.clear();
}

static int split_columns(int nu) {  // n2: the divisor (<= 128) of nu that needs the fewest stages for n1 = nu / n2 and n2 together
  int best = 1, best_cost = 1 << 30;
  for (int d = 2; d <= 128 && d <= nu; ++d) {
    if (nu % d) continue;
    const Stages a = make_stages(d), b = make_stages(nu / d);
    if (a.count < 0 || b.count < 0) continue;
    const int cost = (a.count + b.count) * 4 + (d < 32 ? 3 : 0) + (nu / d > 512 ? 3 : 0);
    if (cost <= best_cost) {
      best_cost = cost;
      best = d;
    }
  }
  return best;
}

static cip_status get_pfft(cip_ctx* ctx, int nu, int nv, int fp64, PfftPlan** out) {
  const auto key = std::make_tuple(nu, nv, fp64);
  auto it = ctx->pfft_plans.find(key);
  if (it != ctx->pfft_plans.end()) {
    *out = it->second;
    return CIP_OK;
  }
  std::unique_ptr<PfftPlan> p(new PfftPlan);
  const size_t csize = fp64 ? 16 : 8;
  p->n2 = split_columns(nu);
  p->n1 = nu / p->n2;
  p->sa = make_stages(nv);
  p->s1 = make_stages(p->n1);
  p->s2 = make_stages(p->n2);
  p->smem_a = fp64 ? padded_len((size_t)nv) * csize : (size_t)nv * csize;
  std::vector<float> ta, t1, t2;
  std::vector<int> fblock, bf_of_f;
  if (!fp64 && p->sa.count > 0 && p->s1.count >= 0 && p->s2.count >= 0) {
    ta = make_stage_tables(&p->sa, true);
    t1 = make_stage_tables(&p->s1, false);
    t2 = make_stage_tables(&p->s2, false);
    // row pass: the last stage's butterfly per frequency, and the frequency contribution of the blocks of the stage
    // before it (pfft_f32.cuh, rows_stage / rows_final)
    const int m = p->sa.count;
    Stages head = p->sa;  // stages 1 .. m - 1 on n / R_m points
    head.count = m - 1;
    head.n = nv / p->sa.radix[m - 1];
    bf_of_f = make_perm(head);
    Stages head2 = head;  // stages 1 .. m - 2 on R_1 .. R_{m-2} points
    if (m >= 2) {
      head2.count = m - 2;
      head2.n = head.n / p->sa.radix[m - 2];
    } else {
      head2.count = 0;
      head2.n = 1;
    }
    const std::vector<int> pos = make_perm(head2);
    fblock.assign(head2.n, 0);
    for (int f = 0; f < head2.n; ++f) fblock[pos[f]] = f;
    p->wf = head2.n;
  }
  const size_t smem_b = (size_t)std::max(p->n1, p->n2) * 16 * csize * 2;  // double-buffered tiles (float32)
  p->ok = p->sa.count > 0 && p->s1.count >= 0 && p->s2.count >= 0 && smem_b <= ctx->smem_optin && p->n1 >= 1 && p->n2 >= 2 &&
          (int64_t)nu * nv < ((int64_t)1 << 40);
  p->ok_rows = p->ok && p->smem_a <= ctx->smem_optin && (fp64 || rows_pairable(p->sa));
  if (p->ok) {
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    p->off_twv = 0;
    p->off_twu = al(p->off_twv + csize * nv);
    p->off_pa = al(p->off_twu + csize * nu);
    p->off_p1 = al(p->off_pa + sizeof(int) * nv);
    p->off_p2 = al(p->off_p1 + sizeof(int) * p->n1);
    p->off_ta = al(p->off_p2 + sizeof(int) * p->n2);
    p->off_t1 = al(p->off_ta + sizeof(float) * ta.size());
    p->off_t2 = al(p->off_t1 + sizeof(float) * t1.size());
    p->off_fb = al(p->off_t2 + sizeof(float) * t2.size());
    p->off_bf = al(p->off_fb + sizeof(int) * fblock.size());
    const size_t total = al(p->off_bf + sizeof(int) * bf_of_f.size());
    std::vector<unsigned char> host(total, 0);
    if (fp64) {
      fill_twiddles<double>(host, p->off_twv, nv);
      fill_twiddles<double>(host, p->off_twu, nu);
    } else {
      fill_twiddles<float>(host, p->off_twv, nv);
      fill_twiddles<float>(host, p->off_twu, nu);
    }
    const std::vector<int> pa = make_perm(p->sa), p1 = make_perm(p->s1), p2 = make_perm(p->s2);
    memcpy(host.data() + p->off_pa, pa.data(), sizeof(int) * pa.size());
    memcpy(host.data() + p->off_p1, p1.data(), sizeof(int) * p1.size());
    memcpy(host.data() + p->off_p2, p2.data(), sizeof(int) * p2.size());
    if (!ta.empty()) memcpy(host.data() + p->off_ta, ta.data(), sizeof(float) * ta.size());
    if (!t1.empty()) memcpy(host.data() + p->off_t1, t1.data(), sizeof(float) * t1.size());
    if (!t2.empty()) memcpy(host.data() + p->off_t2, t2.data(), sizeof(float) * t2.size());
    if (!fblock.empty()) memcpy(host.data() + p->off_fb, fblock.data(), sizeof(int) * fblock.size());
    if (!bf_of_f.empty()) memcpy(host.data() + p->off_bf, bf_of_f.data(), sizeof(int) * bf_of_f.size());
    CIP_TRY(p->buf.ensure(total));
    CIP_CHECK_CUDA(cudaMemcpyAsync(p->buf.p, host.data(), total, cudaMemcpyHostToDevice, ctx->stream));
    CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));  // `host` goes out of scope
  }
  *out = p.get();
  ctx->pfft_plans.emplace(key, p.release());
  return CIP_OK;
}

static int pfft_pitch(int ny) { return (ny + 63) & ~63; }  // columns of the work area: whole tiles of up to 32 pairs
// pairs per tile row of a column pass over n points: the widest tile whose two buffers stay within 48 KB
static int cols_pairs(int n) { return 2 * (size_t)n * 32 * 16 <= 48 * 1024 ? 32 : (2 * (size_t)n * 16 * 16 <= 48 * 1024 ? 16 : 8); }
// CIP_FFT=own: the own row pass too; default: cuFFT for the contiguous row pass, own column passes
static bool pfft_hybrid() {
  const char* e = getenv("CIP_FFT");
  return !(e && strcmp(e, "own") == 0);
}

bool pfft_enabled() {
  const char* e = getenv("CIP_FFT");
  return !(e && strcmp(e, "cufft") == 0);
}

// 0: not available; 1: own row pass; 2: cuFFT row pass + own column passes (float32 only)
static int pfft_variant(const PfftPlan* p, int fp64, int ny) {
  if (!p->ok) return 0;
  if (fp64) return p->ok_rows ? 1 : 0;
  if (pfft_hybrid() && ny % 4 == 0) return 2;
  return p->ok_rows ? 1 : (ny % 4 == 0 ? 2 : 0);
}

int pfft_row_blocks(cip_ctx* ctx, const cip_plan& pl) {
  if (!pfft_enabled() || pl.accumulate_fp64 || getenv("CIP_FFT_2D") || pl.nu % kRowBlocks) return 0;
  const char* e = getenv("CIP_ROW_BLOCKS");  // 0: never, 1: always (tests on small planes)
  if (e ? atoi(e) == 0 : (size_t)pl.nu * pl.nv * 8 < ((size_t)128 << 20)) return 0;  // below 128 MB the extent read-back costs what it saves
  PfftPlan* p;
  if (get_pfft(ctx, pl.nu, pl.nv, 0, &p) != CIP_OK || pfft_variant(p, 0, pl.npix_y) != 2) return 0;
  return pl.nu / kRowBlocks;
}

// Whether the own transform handles this plan (else the cuFFT path of image.cu / fft.cu runs); creates the tables.
cip_status pfft_prepare(cip_ctx* ctx, const cip_plan& pl, bool* ok) {
  *ok = false;
  if (!pfft_enabled()) return CIP_OK;
  PfftPlan* p;
  CIP_TRY(get_pfft(ctx, pl.nu, pl.nv, pl.accumulate_fp64, &p));
  const int variant = pfft_variant(p, pl.accumulate_fp64, pl.npix_y);
  if (!variant) return CIP_OK;
  // float64: only on request (CIP_FFT=own) -- its passes are the unpacked first version, not tuned
  if (pl.accumulate_fp64 && pfft_hybrid()) return CIP_OK;
  CIP_TRY(ctx->fft_t.ensure((size_t)(pl.accumulate_fp64 ? 16 : 8) * pl.nu * pfft_pitch(pl.npix_y)));
  if (variant == 2) CIP_TRY(fft_rows(ctx, nullptr, pl.nv, pl.nu, 0));  // the plan and its work area
  *ok = true;
  return CIP_OK;
}

// row_block > 0 and kept > 0: only the first and the last `kept` blocks of row_block rows hold data (the rows between
// them were never written or cleared: pass A skips them, pass B1 takes them as zeros).
template <typename T, bool SCREEN>
static cip_status pfft_plane(cip_ctx* ctx, PfftPlan* p, const void* plane, int nu, int nv, int ny, const ScreenArgs& sa, void* out,
                             int row_block = 0, int kept = 0) {
  using C = cx<T>;
  const int pitch = pfft_pitch(ny);
  C* g = ctx->fft_t.as<C>();
  char* tb = static_cast<char*>(p->buf.p);
  const C* twv = reinterpret_cast<const C*>(tb + p->off_twv);
  const C* twu = reinterpret_cast<const C*>(tb + p->off_twu);
  const int* pa = reinterpret_cast<const int*>(tb + p->off_pa);
  const int* p1 = reinterpret_cast<const int*>(tb + p->off_p1);
  const int* p2 = reinterpret_cast<const int*>(tb + p->off_p2);
  cudaStream_t st = ctx->stream;
  const int ncb = pitch / 16;
  const size_t smem1 = (size_t)p->n1 * 16 * sizeof(C), smem2 = (size_t)p->n2 * 16 * sizeof(C);
  // rows of up to 56 KB: several CTAs of 256 threads per SM; longer ones: one or two CTAs of 512
  const bool small = p->smem_a <= 56 * 1024;
  if constexpr (sizeof(T) == 4) {
    const float* ta = reinterpret_cast<const float*>(tb + p->off_ta);
    const float2* t1 = reinterpret_cast<const float2*>(tb + p->off_t1);
    const float2* t2 = reinterpret_cast<const float2*>(tb + p->off_t2);
    // G as granules of two columns (pfft_f32.cuh); the column passes are persistent CTAs with double-buffered tiles
    const int gpitch = pitch / 2;
    float4* g4 = reinterpret_cast<float4*>(g);
    const bool hybrid = pfft_variant(p, 0, ny) == 2;
    if (hybrid) {
      // pass A by cuFFT: contiguous 1-D transforms along v, in place (its single-pass kernel runs near the HBM rate;
      // the own row pass of rows_kernel_p does not); pass B1 then reads the retained columns of the plane
      if (row_block > 0 && kept > 0) {
        // two transforms of kept * row_block rows each, one plan per distinct count (block-wise execution of one small
        // plan costs more than it saves: 4.34 ms for 32 blocks of a 24576^2 plane against 3.58 ms in one piece,
        // tools/experiments/fft_blocks.cu)
        const int rows = kept * row_block;
        float2* f = static_cast<float2*>(const_cast<void*>(plane));
        CIP_TRY(fft_rows(ctx, f, nv, rows, 0));
        CIP_TRY(fft_rows(ctx, f + (size_t)(nu - rows) * nv, nv, rows, 0));
      } else {
        CIP_TRY(fft_rows(ctx, const_cast<void*>(plane), nv, nu, 0));
      }
    } else {
      const RowsAux aux{ta, reinterpret_cast<const int*>(tb + p->off_fb), reinterpret_cast<const int*>(tb + p->off_bf), p->wf};
      auto ka = small ? rows_kernel_p<256> : rows_kernel_p<512>;
      CIP_CHECK_CUDA(cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem_a));
      ka<<<nu, small ? 256 : 512, p->smem_a, st>>>(static_cast<const float2*>(plane), reinterpret_cast<float*>(g), p->sa, aux, nv, ny, gpitch);
    }
    const bool pruned = hybrid && row_block > 0 && kept > 0;
    const Cols1Src src{static_cast<const float2*>(plane), nv, ny, pruned ? kept * row_block : nu, pruned ? nu - kept * row_block : nu};
    auto launch1 = [&](auto kern, int P) -> cip_status {
      const size_t smem = 2 * (size_t)p->n1 * P * sizeof(float4);
      const int ncb = gpitch / P;
      CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int per_sm = 1;
      CIP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
      kern<<<std::min(ncb * p->n2, ctx->sm_count * std::max(per_sm, 1)), 128, smem, st>>>(g4, src, p->s1, t1, p1, twu, p->n2, ncb, gpitch);
      return CIP_OK;
    };
    auto launch2 = [&](auto kern, int P) -> cip_status {
      const size_t smem = 2 * (size_t)p->n2 * P * sizeof(float4);
      const int ncb = gpitch / P;
      CIP_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int per_sm = 1;
      CIP_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 128, smem));
      kern<<<std::min(ncb * p->n1, ctx->sm_count * std::max(per_sm, 1)), 128, smem, st>>>(g4, p->s2, t2, p2, p->n1, ncb, gpitch, sa, out);
      return CIP_OK;
    };
    const int P1 = cols_pairs(p->n1), P2 = cols_pairs(p->n2);
    if (hybrid) {
      if (P1 == 32) CIP_TRY(launch1(cols1_kernel_p<32, true>, 32));
      else if (P1 == 16) CIP_TRY(launch1(cols1_kernel_p<16, true>, 16));
      else CIP_TRY(launch1(cols1_kernel_p<8, true>, 8));
    } else {
      if (P1 == 32) CIP_TRY(launch1(cols1_kernel_p<32, false>, 32));
      else if (P1 == 16) CIP_TRY(launch1(cols1_kernel_p<16, false>, 16));
      else CIP_TRY(launch1(cols1_kernel_p<8, false>, 8));
    }
    if (P2 == 32) CIP_TRY(launch2(cols2_kernel_p<32, SCREEN>, 32));
    else if (P2 == 16) CIP_TRY(launch2(cols2_kernel_p<16, SCREEN>, 16));
    else CIP_TRY(launch2(cols2_kernel_p<8, SCREEN>, 8));
    (void)pa;
    (void)ncb;
  } else {
    auto ka = small ? rows_kernel<T, 256> : rows_kernel<T, 512>;
    CIP_CHECK_CUDA(cudaFuncSetAttribute(ka, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p->smem_a));
    ka<<<nu, small ? 256 : 512, p->smem_a, st>>>(static_cast<const C*>(plane), g, p->sa, twv, pa, nv, ny, pitch);
    auto k1 = cols1_kernel<T>;
    CIP_CHECK_CUDA(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    k1<<<dim3(ncb, p->n2), 128, smem1, st>>>(g, p->s1, twu, p1, p->n2, ny, pitch);
    auto k2 = cols2_kernel<T, SCREEN>;
    CIP_CHECK_CUDA(cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
    k2<<<dim3(ncb, p->n1), 128, smem2, st>>>(g, p->s2, twu, p2, p->n1, pitch, sa, out);
  }
  ctx->launches += 3;
  CIP_CHECK_CUDA(cudaGetLastError());
  return CIP_OK;
}

// Transform, crop and w-screen of the resident planes [plane_base, plane_base + nslots) into the image.
cip_status run_pfft_screen(cip_ctx* ctx, const cip_plan& pl, void* grid, int plane_base, int nslots, void* image) {
  PfftPlan* p;
  CIP_TRY(get_pfft(ctx, pl.nu, pl.nv, pl.accumulate_fp64, &p));
  CIP_REQUIRE(pfft_variant(p, pl.accumulate_fp64, pl.npix_y), "own transform not available for %d x %d", pl.nu, pl.nv);
  const size_t csize = pl.accumulate_fp64 ? 16 : 8;
  CIP_TRY(ctx->fft_t.ensure(csize * pl.nu * pfft_pitch(pl.npix_y)));
  const int row_block = pfft_row_blocks(ctx, pl);
  CIP_REQUIRE(row_block > 0 || ctx->rows_kept_blocks == 0, "row blocks were skipped for a transform that reads every row");
  for (int s = 0; s < nslots; ++s) {
    const int plane = plane_base + s;
    if (plane >= pl.nplanes) break;
    ScreenArgs sa{pl.npix_x, pl.npix_y, pl.nu, pl.nv, pl.pixsize_x, pl.pixsize_y, pl.w0 + plane * pl.dw, pl.nshift,
                  pl.do_wstacking, ctx->call_flags.as<uint32_t>()};
    const char* f = static_cast<const char*>(grid) + csize * (size_t)s * pl.nu * pl.nv;
    if (pl.accumulate_fp64) CIP_TRY((pfft_plane<double, true>(ctx, p, f, pl.nu, pl.nv, pl.npix_y, sa, image)));
    else CIP_TRY((pfft_plane<float, true>(ctx, p, f, pl.nu, pl.nv, pl.npix_y, sa, image, row_block, ctx->rows_kept_blocks)));
  }
  return CIP_OK;
}

}  // namespace cip

// Stage G3 on its own: out[i][j] = sum F[u][v] exp(+2 pi i (u fu / nu + v fv / nv)), fu = (i - nx/2) mod nu,
// fv = (j - ny/2) mod nv -- the nx x ny cells of the transformed plane that the image keeps.  Device pointers.
extern "C" cip_status cip_plane_fft_crop(cip_ctx* ctx, const void* plane, int nu, int nv, int nx, int ny, int fp64, void* out) {
  using namespace cip;
  CIP_REQUIRE(ctx && plane && out, "NULL argument");
  CIP_REQUIRE(nu >= 2 && nv >= 2 && nu % 2 == 0 && nv % 2 == 0 && nx <= nu && ny <= nv && nx >= 2 && ny >= 2 && nx % 2 == 0 && ny % 2 == 0,
              "bad plane / image sizes");
  CIP_ENTER_DEVICE(ctx);
  PfftPlan* p;
  CIP_TRY(get_pfft(ctx, nu, nv, fp64, &p));
  CIP_REQUIRE(pfft_variant(p, fp64, ny), "own transform not available for %d x %d (fp64 = %d)", nu, nv, fp64);
  CIP_TRY(ctx->fft_t.ensure((size_t)(fp64 ? 16 : 8) * nu * pfft_pitch(ny)));
  ScreenArgs sa{nx, ny, nu, nv, 0.0, 0.0, 0.0, 0.0, 0, nullptr};
  if (fp64) CIP_TRY((pfft_plane<double, false>(ctx, p, plane, nu, nv, ny, sa, out)));
  else CIP_TRY((pfft_plane<float, false>(ctx, p, plane, nu, nv, ny, sa, out)));
  CIP_CHECK_CUDA(cudaStreamSynchronize(ctx->stream));
  return CIP_OK;
}

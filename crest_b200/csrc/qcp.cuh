// Per-pair epilogue math, all in registers (FP64).
//
// The reference obtains lambda_max of the 4x4 key matrix S(R) by Givens tridiagonalisation + SVD
// (reference src/ls_rmsd.f90:223-284).  Here lambda_max is the largest root of the characteristic
// quartic of S written in the invariants of the 3x3 covariance R (sigma_k = singular values):
//     P(l) = l^4 - 2*I1*l^2 - 8*det(R)*l + (I1^2 - 4*I2),
//     I1 = ||R||_F^2 = sum sigma^2,   I2 = ||cof R||_F^2 = sum_{a<b} sigma_a^2 sigma_b^2,
// whose roots are (+-sigma1 +-sigma2 +-sigma3) with an even number of minus signs times
// sign(det R).  Negating x of the second structure (the reference's mirror test,
// src/cregen.f90:1279-1284) flips the sign of det R only, so min(rmsd, rmsd_mirror) belongs to
// the quartic with -8*|det R|: one Newton solve per pair instead of two eigen-decompositions.
#pragma once
#include <cuda_runtime.h>

namespace crest {

__device__ __forceinline__ double rcp_approx(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));   // MUFU.RCP64H, ~20 bits
  return r;
}

struct Quartic {
  double c2, c1abs, c0;   // P(l) = l^4 + c2 l^2 -+ c1abs l + c0 ; c1abs = 8*|det R|
  double det;
};

// R row-major r[3*a+b]
__device__ __forceinline__ Quartic quartic_of(const double* r) {
  const double i1 = fma(r[8], r[8], fma(r[7], r[7], fma(r[6], r[6], fma(r[5], r[5],
                    fma(r[4], r[4], fma(r[3], r[3], fma(r[2], r[2], fma(r[1], r[1], r[0] * r[0]))))))));
  const double k00 = fma(r[4], r[8], -r[5] * r[7]);
  const double k01 = fma(r[5], r[6], -r[3] * r[8]);
  const double k02 = fma(r[3], r[7], -r[4] * r[6]);
  const double k10 = fma(r[2], r[7], -r[1] * r[8]);
  const double k11 = fma(r[0], r[8], -r[2] * r[6]);
  const double k12 = fma(r[1], r[6], -r[0] * r[7]);
  const double k20 = fma(r[1], r[5], -r[2] * r[4]);
  const double k21 = fma(r[2], r[3], -r[0] * r[5]);
  const double k22 = fma(r[0], r[4], -r[1] * r[3]);
  const double det = fma(r[0], k00, fma(r[1], k01, r[2] * k02));
  const double i2 = fma(k22, k22, fma(k21, k21, fma(k20, k20, fma(k12, k12, fma(k11, k11,
                    fma(k10, k10, fma(k02, k02, fma(k01, k01, k00 * k00))))))));
  Quartic q;
  q.c2 = -2.0 * i1;
  q.det = det;
  q.c1abs = 8.0 * fabs(det);
  q.c0 = fma(i1, i1, -4.0 * i2);
  return q;
}

// Cyclic Jacobi on the symmetric traceless 4x4 key matrix (ls_rmsd.f90:129-147); the rare,
// exact path for (near-)degenerate lambda_max where Newton on the quartic stalls.
static __device__ __noinline__ double lambda_max_jacobi(const double* r, double sign_x) {
  // sign_x = -1 builds S for the mirrored second structure (column... row 1 of y negated)
  double R[3][3];
  for (int a = 0; a < 3; ++a) {
    R[a][0] = r[3 * a + 0] * sign_x;   // y_x -> -y_x flips column 0 of R(a,b)=sum x_a y_b
    R[a][1] = r[3 * a + 1];
    R[a][2] = r[3 * a + 2];
  }
  double S[4][4];
  S[0][0] = R[0][0] + R[1][1] + R[2][2];
  S[1][0] = R[1][2] - R[2][1];
  S[2][0] = R[2][0] - R[0][2];
  S[3][0] = R[0][1] - R[1][0];
  S[1][1] = R[0][0] - R[1][1] - R[2][2];
  S[2][1] = R[0][1] + R[1][0];
  S[3][1] = R[0][2] + R[2][0];
  S[2][2] = -R[0][0] + R[1][1] - R[2][2];
  S[3][2] = R[1][2] + R[2][1];
  S[3][3] = -R[0][0] - R[1][1] + R[2][2];
  for (int i = 0; i < 4; ++i)
    for (int j = i + 1; j < 4; ++j) S[i][j] = S[j][i];
  for (int sweep = 0; sweep < 50; ++sweep) {
    double off = 0.0, diag = 0.0;
    for (int i = 0; i < 4; ++i) {
      diag += S[i][i] * S[i][i];
      for (int j = i + 1; j < 4; ++j) off += S[i][j] * S[i][j];
    }
    if (off <= 1e-34 * diag || off == 0.0) break;
    for (int p = 0; p < 3; ++p)
      for (int q = p + 1; q < 4; ++q) {
        const double apq = S[p][q];
        if (apq == 0.0) continue;
        const double theta = (S[q][q] - S[p][p]) / (2.0 * apq);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
        for (int k = 0; k < 4; ++k) {
          const double akp = S[k][p], akq = S[k][q];
          S[k][p] = c * akp - s * akq;
          S[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 4; ++k) {
          const double apk = S[p][k], aqk = S[q][k];
          S[p][k] = c * apk - s * aqk;
          S[q][k] = s * apk + c * aqk;
        }
      }
  }
  return fmax(fmax(S[0][0], S[1][1]), fmax(S[2][2], S[3][3]));
}

// Largest root of l^4 + c2 l^2 + c1 l + c0 by Newton from the upper bound l0 >= lambda_max.
// Monotone from the right for a real-rooted quartic.  Returns false if not converged (caller
// falls back to Jacobi).
__device__ __forceinline__ bool newton_lambda(double c2, double c1, double c0, double l0,
                                              double& lam_out) {
  double l = l0;
  const double tol = 1e-11 * l0;
  bool ok = false;
#pragma unroll 1
  for (int it = 0; it < 40; ++it) {
    const double x2 = l * l;
    const double a = x2 + c2;
    const double b = a * l;
    const double t = b + c1;
    const double f = fma(t, l, c0);
    const double fp = fma(x2 + x2, l, b) + t;
    const double d = f * rcp_approx(fp);
    l -= d;
    if (fabs(d) <= tol) { ok = true; break; }
  }
  // a (near-)double root shows as a derivative that vanished relative to its natural scale
  {
    const double x2 = l * l;
    const double fp = fma(4.0 * x2, l, fma(2.0 * c2, l, c1));
    if (!(fabs(fp) > 1e-6 * x2 * l)) ok = false;
  }
  lam_out = l;
  return ok;
}

// e = xn + yn - 2*lambda for the minimum of proper / mirror RMSD (mirror only if enantio).
// Also usable for a single chirality: pass enantio=0 (proper) or det-sign flipped via mirror=1.
__device__ __forceinline__ double pair_e(const double* r, double s, int enantio, int mirror) {
  const Quartic q = quartic_of(r);
  double c1;
  if (enantio) c1 = -q.c1abs;
  else c1 = mirror ? 8.0 * q.det : -8.0 * q.det;
  double lam;
  if (!newton_lambda(q.c2, c1, q.c0, 0.5 * s, lam)) {
    if (enantio) lam = fmax(lambda_max_jacobi(r, 1.0), lambda_max_jacobi(r, -1.0));
    else lam = lambda_max_jacobi(r, mirror ? -1.0 : 1.0);
  }
  return fma(-2.0, lam, s);
}

// Out-of-line exact solve for the rare candidate pairs of the tile epilogue (keeps the hot loop's
// instruction footprint small).
static __device__ __noinline__ double pair_e_exact(double r0, double r1, double r2, double r3,
                                                   double r4, double r5, double r6, double r7,
                                                   double r8, double s, int enantio) {
  const double r[9] = {r0, r1, r2, r3, r4, r5, r6, r7, r8};
  return pair_e(r, s, enantio, 0);
}

// Sink of the rmsd() callers beside cregen_CRE (common.cuh PairSink): r = FP64 covariance of pair (i, j),
// s = |x_i|^2 + |y_j|^2.  Out of line: it runs for every pair of a compare / intraconf sweep but must not
// bloat the CREGEN kernels' hot loops.
struct PairSink;
static __device__ __noinline__ void sink_pair(const PairSink& k, int i, int j, double r0, double r1, double r2,
                                              double r3, double r4, double r5, double r6, double r7, double r8,
                                              double s);

// Branch-free screen: true  =>  PROVEN  e = s - 2*lambda_max > Ecut  (the pair cannot pass the
// 2*rthr gate), false => undecided, run the exact solve.
// With lam = (s - Ecut)/2 the claim is lambda_max < lam.  For a real-rooted quartic P (the
// characteristic polynomial of a symmetric matrix) Budan-Fourier gives: P and its first, second
// and third derivative all > 0 at lam  <=>  no root above lam.  Each sign is required with a
// floor far above the rounding error of its evaluation (~1e-14 * scale), so a `true` is rigorous;
// anything closer goes to the exact path, and the caller passes Ecut with a 1e-8*s margin over
// the decision threshold so the screen never decides a pair whose exact-path result could round
// to the other side.
__device__ __forceinline__ bool screen_far(const Quartic& q, double c1, double s, double Ecut) {
  const double lam = 0.5 * (s - Ecut);
  const double l2 = lam * lam;
  const double i1 = -0.5 * q.c2;
  const double sc = fmax(l2, i1);              // scale^2
  const double a = l2 + q.c2;                  // l^2 + c2
  const double t = fma(a, lam, c1);            // l^3 + c2 l + c1
  const double P = fma(t, lam, q.c0);
  const double dP = fma(fma(4.0, l2, 2.0 * q.c2), lam, c1);
  const double ddP = fma(6.0, l2, q.c2);       // second derivative / 2
  const double fl = 1e-10 * sc;
  return (lam > 0.0) & (P > fl * sc) & (dP > fl * lam) & (ddP > fl);
}

// FP64 -> FP32 narrowing by four integer ops (truncation toward zero, relative error < 2^-23).
// F2F.F32.F64 is a slow-path instruction on B200 that also stalls the DMMA pipe
// (tools/microbench/cvt_bench.cu, profiles/r01_cvt_microbench.txt); ALU instructions run beside
// the tensor pipe.  |x| below the float normal range (incl. 0) gives a float denormal < 2^-126,
// i.e. an absolute error far below every floor of the screen.  The caller guarantees
// |x| < 2^127 (pair_tile checks s = |x|^2+|y|^2 against 2^100; |R_ab| <= s/2 by Cauchy-Schwarz).
__device__ __forceinline__ float f64_to_f32_trunc(double x) {
  const int hi = __double2hiint(x), lo = __double2loint(x);
  const int t = max((hi & 0x7fffffff) - 0x38000000, 0);
  const unsigned f = __funnelshift_l((unsigned)lo, (unsigned)t, 3);
  return __uint_as_float(f | ((unsigned)hi & 0x80000000u));
}

// FP32 version of the screen, run on the FFMA pipe (which co-issues with the DMMA/FP64 pipe,
// profiles/r01_fp64_pipes_microbench.txt).  Inputs are the covariance rounded to float and a
// test point lam <= (s - Ecut)/2 (the caller rounds it DOWN).  Every quantity is a short sum of
// products bounded by M^2 (P), M^1.5 (P') or M (P''/2) with M = max(lam^2, I1); the accumulated
// rounding error of the float evaluation, input truncation (2^-23) included, is below ~100 * 2^-24 of those
// scales, and each sign is required with a floor of 2e-4 of the scale (> 30x that bound), so a
// `true` still PROVES lambda_max < lam.  Overflow/NaN compare false => undecided.
__device__ __forceinline__ bool screen_far_f32(const float* r, float lam, int enantio) {
  const float i1 = fmaf(r[8], r[8], fmaf(r[7], r[7], fmaf(r[6], r[6], fmaf(r[5], r[5],
                   fmaf(r[4], r[4], fmaf(r[3], r[3], fmaf(r[2], r[2], fmaf(r[1], r[1], r[0] * r[0]))))))));
  const float k00 = fmaf(r[4], r[8], -r[5] * r[7]);
  const float k01 = fmaf(r[5], r[6], -r[3] * r[8]);
  const float k02 = fmaf(r[3], r[7], -r[4] * r[6]);
  const float k10 = fmaf(r[2], r[7], -r[1] * r[8]);
  const float k11 = fmaf(r[0], r[8], -r[2] * r[6]);
  const float k12 = fmaf(r[1], r[6], -r[0] * r[7]);
  const float k20 = fmaf(r[1], r[5], -r[2] * r[4]);
  const float k21 = fmaf(r[2], r[3], -r[0] * r[5]);
  const float k22 = fmaf(r[0], r[4], -r[1] * r[3]);
  const float det = fmaf(r[0], k00, fmaf(r[1], k01, r[2] * k02));
  const float i2 = fmaf(k22, k22, fmaf(k21, k21, fmaf(k20, k20, fmaf(k12, k12, fmaf(k11, k11,
                   fmaf(k10, k10, fmaf(k02, k02, fmaf(k01, k01, k00 * k00))))))));
  const float c2 = -2.0f * i1;
  const float c1 = enantio ? -8.0f * fabsf(det) : -8.0f * det;
  const float c0 = fmaf(i1, i1, -4.0f * i2);
  const float l2 = lam * lam;
  const float M = fmaxf(l2, i1);
  const float t = fmaf(l2 + c2, lam, c1);
  const float P = fmaf(t, lam, c0);
  const float dP = fmaf(fmaf(4.0f, l2, 2.0f * c2), lam, c1);
  const float ddP = fmaf(6.0f, l2, c2);
  const float fl = 2e-4f * M;
  return (lam > 0.0f) & (P > fl * M) & (dP > fl * lam) & (ddP > fl);
}

// ---- packed (2 pairs per instruction) version of the FP32 screen --------------------------------
// Blackwell issues FP32 arithmetic on 64-bit register pairs (FFMA2/FMUL2/FADD2, PTX .f32x2): same
// FMA-pipe time per element, HALF the issue slots, which is what the tensor-core screen kernel
// (pair_tf32.cu) is short of.  Written without operand negation (the packed forms have no negate
// modifier in PTX): I2 comes from the Gram matrix G = R^T R,
//     4*I2 = 2*(I1^2 - ||G||_F^2),  I1 = tr G,
// so the only differences of products left are the three cofactors of det R, which use
// sign-flipped copies of row 2 (integer XOR, ALU pipe).  Same test point and floors as screen_far_f32.
typedef unsigned long long f2_t;   // two floats: lo = first pair, hi = second pair
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) {
  f2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ f2_t f2_pack_u(unsigned lo, unsigned hi) {
  f2_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void f2_unpack(f2_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) {
  f2_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) {
  f2_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) {
  f2_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2_t f2_sub(f2_t a, f2_t b) {
  f2_t r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2_t f2_neg(f2_t a) {   // written per half so that ptxas can fold it into an operand modifier
  float lo, hi;
  f2_unpack(a, lo, hi);
  return f2_pack(-lo, -hi);
}
__device__ __forceinline__ f2_t f2_abs(f2_t a) { return a & 0x7FFFFFFF7FFFFFFFull; }

__device__ __forceinline__ f2_t f2_add_rd(f2_t a, f2_t b) {   // round toward -inf (test points)
  f2_t r;
  asm("add.rm.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// r[3a+b] packed covariance of two pairs, lam packed test points, which the CALLER guarantees to be
// positive.  Returns a packed word per pair whose SIGN BIT IS CLEAR iff the pair is PROVEN far, i.e.
// lambda_max < lam (the OR of the two margins below; +0 counts as far, the floors are 30x the
// evaluation error).  The caller also guarantees finite arithmetic (lam < 1e9, |R_ab| <= ~lam): no
// Inf/NaN can arise, so the sign test is a valid "both margins >= 0".
//
// Two signs instead of the four of Budan-Fourier: the key matrix is symmetric and traceless, so its
// eigenvalues l1 >= l2 >= l3 >= l4 satisfy sum = 0 and sum of squares = 4*I1 (the lambda^2
// coefficient of P).  With l1 >= l2 = a >= 0: l3 + l4 <= -2a, hence l3^2 + l4^2 >= 2a^2 and
// 4*I1 >= 4a^2, i.e.  l2 <= sqrt(I1)  always.  Therefore
//     lam > 0,  lam^2 > I1,  P(lam) > 0     =>     lam > l2 and P > 0     =>     lam > l1
// (between l2 and l1 the quartic is negative).  With both det signs wanted (enantio) P is evaluated
// with -8*|det|*lam, the smaller of the two.  Evaluation: u = lam^2 - I1,
//     P = u^2 + 2*(||G||^2 - I1^2) - 8*det*lam = u^2 + 2*x,  x = ||G||^2 - I1^2 - 4*det*lam,
// every term bounded by lam^4 once u > 0; FP32 evaluation error < 100 * 2^-24 of lam^2 (u) and
// lam^4 (P); both signs are required with a floor of 2e-4 of those scales.
// 46 packed FMA-pipe instructions per two pairs (52 for the four-sign form).
__device__ __forceinline__ f2_t screen_margin_f32x2(const f2_t* r, f2_t lam, int enantio) {
  // Gram matrix G = R^T R (columns of R): g_bc = sum_a r[3a+b] r[3a+c]
  const f2_t g00 = f2_fma(r[6], r[6], f2_fma(r[3], r[3], f2_mul(r[0], r[0])));
  const f2_t g11 = f2_fma(r[7], r[7], f2_fma(r[4], r[4], f2_mul(r[1], r[1])));
  const f2_t g22 = f2_fma(r[8], r[8], f2_fma(r[5], r[5], f2_mul(r[2], r[2])));
  const f2_t g01 = f2_fma(r[6], r[7], f2_fma(r[3], r[4], f2_mul(r[0], r[1])));
  const f2_t g02 = f2_fma(r[6], r[8], f2_fma(r[3], r[5], f2_mul(r[0], r[2])));
  const f2_t g12 = f2_fma(r[7], r[8], f2_fma(r[4], r[5], f2_mul(r[1], r[2])));
  const f2_t i1 = f2_add(f2_add(g00, g11), g22);
  const f2_t hoff = f2_fma(g12, g12, f2_fma(g02, g02, f2_mul(g01, g01)));
  const f2_t dsq = f2_fma(g22, g22, f2_fma(g11, g11, f2_mul(g00, g00)));
  const f2_t ng = f2_fma(f2_pack(2.0f, 2.0f), hoff, dsq);      // ||G||_F^2
  // det R by row 0; cofactors k0b = minors of rows 1,2 with n = -row 2
  const f2_t n6 = f2_neg(r[6]), n7 = f2_neg(r[7]), n8 = f2_neg(r[8]);
  const f2_t k00 = f2_fma(r[5], n7, f2_mul(r[4], r[8]));       // r4 r8 - r5 r7
  const f2_t k01 = f2_fma(r[3], n8, f2_mul(r[5], r[6]));       // r5 r6 - r3 r8
  const f2_t k02 = f2_fma(r[4], n6, f2_mul(r[3], r[7]));       // r3 r7 - r4 r6
  const f2_t det = f2_fma(r[0], k00, f2_fma(r[1], k01, f2_mul(r[2], k02)));
  const f2_t l2 = f2_mul(lam, lam);
  const f2_t u = f2_sub(l2, i1);
  const f2_t x0 = f2_fma(f2_neg(i1), i1, ng);                  // ||G||^2 - I1^2
  const f2_t lm4 = f2_mul(lam, f2_pack(-4.0f, -4.0f));
  const f2_t x = f2_fma(enantio ? f2_abs(det) : det, lm4, x0);
  const f2_t P = f2_fma(x, f2_pack(2.0f, 2.0f), f2_mul(u, u));
  // margins over the floors: d1 = u - 2e-4*lam^2, d2 = P - 2e-4*lam^4
  const f2_t d1 = f2_fma(l2, f2_pack(-2e-4f, -2e-4f), u);
  const f2_t d2 = f2_fma(f2_mul(l2, f2_pack(-2e-4f, -2e-4f)), l2, P);
  return d1 | d2;
}

// reference src/rotcompare.f90:109-122 (same operation order)
__device__ __forceinline__ double rotaniso_dev(double a, double b, double c) {
  const double av = __ddiv_rn(__dadd_rn(__dadd_rn(a, b), c), 3.0);
  const double da = __dsub_rn(a, av), db = __dsub_rn(b, av), dc = __dsub_rn(c, av);
  double r = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(da, da), __dmul_rn(db, db)), __dmul_rn(dc, dc)));
  r = __ddiv_rn(r, av);
  r = __ddiv_rn(r, __dmul_rn(3.0, sqrt(__ddiv_rn(2.0, 3.0))));
  return r;
}

// reference src/rotcompare.f90:127-142
__device__ __forceinline__ double bthrerf_dev(double bthr, double aniso, double bthrmax,
                                              double bthrshift) {
  const double c = __ddiv_rn(__dsub_rn(__dmul_rn(bthrmax, 100.0), __dmul_rn(bthr, 100.0)), 2.0);
  const double a = __dadd_rn(__dmul_rn(-erf(-2.5), c), __dmul_rn(bthr, 100.0));
  const double b = __ddiv_rn(4.0, 0.8);
  const double d = __ddiv_rn(bthrshift, 0.15);
  double thr = __dadd_rn(__dmul_rn(erf(__dsub_rn(__dmul_rn(aniso, b), d)), c), a);
  thr = __ddiv_rn(thr, 100.0);
  return thr;
}

// reference src/rotcompare.f90:146-169 ; division by a zero rot. constant gives Inf/NaN => false
__device__ __forceinline__ bool equalrotaniso_dev(const double* roti, const double* rotj,
                                                  double anisoi, double anisoj, double bthr,
                                                  double bthrmax, double bthrshift) {
  const double av = __ddiv_rn(__dadd_rn(anisoi, anisoj), 2.0);
  const double thr = bthrerf_dev(bthr, av, bthrmax, bthrshift);
  const bool r1 = fabs(__dsub_rn(__ddiv_rn(roti[0], rotj[0]), 1.0)) <= thr;
  const bool r2 = fabs(__dsub_rn(__ddiv_rn(roti[1], rotj[1]), 1.0)) <= thr;
  const bool r3 = fabs(__dsub_rn(__ddiv_rn(roti[2], rotj[2]), 1.0)) <= thr;
  return r1 && r2 && r3;
}

}  // namespace crest

namespace crest {
static __device__ __noinline__ void sink_pair(const PairSink& k, int i, int j, double r0, double r1, double r2,
                                              double r3, double r4, double r5, double r6, double r7, double r8,
                                              double s) {
  const double r[9] = {r0, r1, r2, r3, r4, r5, r6, r7, r8};
  const double invA = 1.0 / (double)k.A;
  if (k.mode == 1) {
    // compare_ensembles (ensemblecomp.f90:213-248): minimum all-atom RMSD per (conformer, conformer) block
    const int ci = k.blk_i[i], cj = k.blk_j[j];
    if (ci < 0 || cj < 0) return;
    const double val = sqrt(fmax(0.0, pair_e(r, s, 0, 0)) * invA);         // ls_rmsd.f90:163
    // non-negative doubles order like their bit patterns; the plain read only filters (the slot decreases
    // monotonically), the atomic decides
    unsigned long long* slot = k.rmat_bits + ci + (size_t)k.nconf1 * cj;
    const unsigned long long bits = (unsigned long long)__double_as_longlong(val);
    if (bits < *reinterpret_cast<volatile unsigned long long*>(slot)) atomicMin(slot, bits);
  } else {
    // intraconfRMSD (entropic.f90:1000-1047): rotamers of ONE conformer, j < i
    const int ci = k.blk_i[i];
    if (ci < 0 || ci != k.blk_j[j]) return;
    const double val = sqrt(fmax(0.0, pair_e(r, s, 0, 0)) * invA);
    if (val < k.rthr) k.has_earlier[i] = 1;                                // entropic.f90:1088 via uniqueCore
    if (val > k.rthr) {                                                    // entropic.f90:1024-1031
      const double vm = sqrt(fmax(0.0, pair_e(r, s, 0, 1)) * invA);
      if (vm < k.rthr_enantio) k.enantio[ci] = 1;
    }
  }
}
}  // namespace crest

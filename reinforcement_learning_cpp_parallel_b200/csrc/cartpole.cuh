// cartpole.cuh -- the cart-pole ODE and its 6-stage low-storage RK46-NL integrator, templated on the
// arithmetic type.  State lives in registers for the whole rollout; nothing here touches memory.
// Reference: include/cart_pole.h  (Diff :149-170 non-SWINGUP branch, rk46_nl :23-45, advance :107-118).
//
// One env step is 50 substeps x 6 stages = 300 evaluations of Diff, each needing sin/cos of the pole
// angle and one true division.  Calling libdevice sincos()/"/" 300 times costs ~150 instructions per
// stage, two thirds of them integer fix-up code (round-1 ncu, profiles/r1a_*: FP64 pipe 42 % busy, ALU
// 18 %, instruction fetch the top stall).  The double path therefore restates the expensive pieces
// so that a stage is 27 straight-line DFMA/DMUL/DADD (32 in round 1) and almost nothing else -- the kernel is bound by
// FP64 instruction ISSUE (a DFMA holds the port for two cycles), so every instruction removed is time:
//   * sin/cos are carried along with the state and ROTATED by the stage's angle increment d
//     (|d| ~ 1e-5..1e-3; |d| < 1.8e-3 is enforced): sin d and cos d - 1 from two-term Taylor series
//     (truncation below double rounding in that range), then the angle-addition formulas as two FMAs
//     per component.  The pair is re-anchored with a full-precision
//     sincos() once per env step, so rounding cannot accumulate beyond ~300 ulp.  While |thetadot| > 16
//     rad/s (|d| could leave the polynomial range) substeps run through an exact out-of-line version.
//   * 1/fac2 is carried as well and refreshed with one Newton step per stage (fac2 moves by < 1e-3
//     relative between stages, typically 1e-6), followed by one residual correction of the quotient:
//     error ~ e^4, i.e. < 1e-12 in the worst case and ~1e-24 typically.
//   * the RK accumulator is kept as W = w / dt and b_i dt is folded into one constant.
// Round 2 (same algebra, fewer instructions; every change re-checked against the 1e-9 replay gates):
//   * (sin, cos) ride along SCALED by g: the rotation is linear, so g s and g c rotate like s and c, and g folds
//     into the constants of fac2, F1 / M and the numerator (one DMUL per stage gone);
//   * the rotation by d is three shears, (s, c) -> s += c tau, c -= s sigma, s += c tau with tau = tan(d / 2),
//     sigma = sin d from their two-term series: an EXACT rotation in three FMAs instead of four;
//   * the carried 1 / fac2 takes one third-order step, r (1 + e + e^2) with e = 1 - fac2 r: |e| < 1.3e-4 while
//     |d| < 1.8e-3 (fac2 moves by 0.07 d relative per stage), so the quotient is good to e^3 < 2.2e-12 in the worst
//     case and ~1e-18 typically, and the residual correction (two instructions) is gone;
//   * x and xdot never feed back into Diff (include/cart_pole.h:161-168 reads theta and thetadot only), so they are not
//     carried through the low-storage recurrence: unrolling it gives u_final = u_0 + dt sum_j beta_j f_j, and for the
//     pair (x, xdot) that is one FMA each per stage with precomputed weights (rk_derived()) instead of two each.
// Measured against the unmodified header: <= 1e-12 relative per env step (tests/test_gpu_parity.py).
#pragma once
#include "../../include/b200ppo.h"
#include "common.cuh"

namespace b2 {

// constants include/cart_pole.h:65-70
struct CPC {
  static constexpr double mp = 0.1, mc = 1.0, l = 0.5, g = 9.81, dt = 4e-4;
  static constexpr int nsteps = 50;
  static constexpr double inv_tot = 1.0 / (mp + mc);
  static constexpr double mpl = mp * l;
  static constexpr double A0 = l * (4.0 / 3);          // fac2 = A0 - A1 c^2   (:162)
  static constexpr double A1 = l * mp / (mp + mc);
  static constexpr double PI_15 = 3.14159265358979323846 / 15;
};

// RK46-NL coefficients, include/cart_pole.h:25-30
#define B2_RK_A {0.000000000000, -0.737101392796, -1.634740794341, -0.744739003780, -1.469897351522, -2.813971388035}
#define B2_RK_B {0.032918605146, 0.823256998200, 0.381530948900, 0.200092213184, 1.718581042715, 0.270000000000}

template <typename R>
struct CartPoleDyn;

// ------------------------------------------------------------------------------------------ double
// Every double constant of the hot loop lives in one __constant__ table so that it reaches the DFMA as
// a c[bank][offset] operand (ptxas otherwise rebuilds each literal with two UMOVs per use: 312 of the
// 965 instructions of a substep in the first version).
// The low-storage recurrence w_i = a_i w_{i-1} + dt f_i, u_i = u_{i-1} + b_i w_i unrolled (include/cart_pole.h:39-43):
//   u_k = u_0 + dt sum_{j<=k} B[k][j] f_j,   B[k][j] = sum_{i=j..k} b_i prod_{m=j+1..i} a_m,   beta_j = B[5][j].
// For xdot, f_j = xddot_j.  For x, f_j is xdot BEFORE stage j, i.e. xdot_0 + dt sum_{m<j} B[j-1][m] xddot_m, hence
//   x_6 = x_0 + dt (sum beta) xdot_0 + dt^2 sum_m gamma_m xddot_m,   gamma_m = sum_{j>m} beta_j B[j-1][m].
struct RKDerived { double beta[6], gamma[6], cbeta; };
constexpr RKDerived rk_derived() {
  constexpr double a[6] = B2_RK_A;
  constexpr double b[6] = B2_RK_B;
  double B[6][6] = {};
  for (int j = 0; j < 6; j++)
    for (int k = j; k < 6; k++) {
      double s = 0;
      for (int i = j; i <= k; i++) {
        double p = 1;
        for (int m = j + 1; m <= i; m++) p *= a[m];
        s += b[i] * p;
      }
      B[k][j] = s;
    }
  RKDerived r{};
  for (int j = 0; j < 6; j++) { r.beta[j] = B[5][j]; r.cbeta += B[5][j]; }
  for (int m = 0; m < 6; m++) {
    double s = 0;
    for (int j = m + 1; j < 6; j++) s += B[5][j] * B[j - 1][m];
    r.gamma[m] = s;
  }
  return r;
}
constexpr RKDerived kRK = rk_derived();

struct CPTable {
  double a[6], bdt[6];          // RK46-NL a_i and b_i * dt  (the accumulator is kept as W = w / dt)
  double cv[6], cx[6];          // xdot += cv_j (xddot_j / g), x += cx_j (xddot_j / g):  g dt beta_j,  g dt^2 gamma_j
  double dtcb;                  // dt * sum beta:  x += dtcb * xdot_0 at the start of a substep
  double A0, A1g, k2;           // fac2 = A0 - A1g (g c)^2, A1g = A1 / g^2 ;  k2 = mp l / (M g^2)
  double t1, t3;                // tan(d / 2) = d (t1 + t3 d^2)        |error| <= d^5 / 240
  double s3;                    // sin d = d (1 + s3 d^2)              |error| <= d^5 / 120
};
#define B2_G 9.81
__device__ __constant__ CPTable kCP = {
    B2_RK_A,
    {0.032918605146 * 4e-4, 0.823256998200 * 4e-4, 0.381530948900 * 4e-4, 0.200092213184 * 4e-4,
     1.718581042715 * 4e-4, 0.270000000000 * 4e-4},
    {B2_G * 4e-4 * kRK.beta[0], B2_G * 4e-4 * kRK.beta[1], B2_G * 4e-4 * kRK.beta[2], B2_G * 4e-4 * kRK.beta[3],
     B2_G * 4e-4 * kRK.beta[4], B2_G * 4e-4 * kRK.beta[5]},
    {B2_G * 4e-4 * 4e-4 * kRK.gamma[0], B2_G * 4e-4 * 4e-4 * kRK.gamma[1], B2_G * 4e-4 * 4e-4 * kRK.gamma[2],
     B2_G * 4e-4 * 4e-4 * kRK.gamma[3], B2_G * 4e-4 * 4e-4 * kRK.gamma[4], B2_G * 4e-4 * 4e-4 * kRK.gamma[5]},
    4e-4 * kRK.cbeta,
    CPC::A0, CPC::A1 / (B2_G * B2_G), CPC::mpl* CPC::inv_tot / (B2_G * B2_G),
    0.5, 1.0 / 24,
    -1.0 / 6};
static_assert(CPC::g == B2_G && CPC::dt == 4e-4, "kCP is built from the same constants");

// Stage increments of the angle are |d| = |W2 b_i dt| <= 0.27 |thetadot| dt for this scheme (W2 is a
// convex-like blend of the stage velocities).  Below 16 rad/s that is |d| < 1.8e-3, where the two
// truncated series are exact to 1.5e-16 (sin) and 5e-20 (cos - 1), i.e. below double rounding.  An
// untrained policy peaks at ~11 rad/s (oracle, 2048 envs x 128 steps), so the exact path never runs
// in practice; it is there so that correctness does not depend on that observation.
constexpr double kOmegaFast = 16.0;

template <>
struct CartPoleDyn<double> {
  // Reference-shaped substep (full sincos + true division per stage), used only while the pole spins so
  // fast that the rotation polynomials below are out of range.  Kept out of line: it is never hot.
  // (by value in and out: passing u by reference would pin the state to the local-memory stack frame)
  static __device__ __noinline__ double4 substep_exact(double4 u4, double F) {
    constexpr double a[6] = B2_RK_A;
    constexpr double b[6] = B2_RK_B;
    double u[4] = {u4.x, u4.y, u4.z, u4.w};
    double w[4] = {0, 0, 0, 0};
#pragma unroll 1
    for (int i = 0; i < 6; i++) {
      double s, c;
      sincos(u[2], &s, &c);
      const double fac2 = CPC::l * (4.0 / 3 - (CPC::mp * c * c) * CPC::inv_tot);
      const double F1 = F + CPC::mpl * u[3] * u[3] * s;
      const double dd4 = (CPC::g * s - F1 * c * CPC::inv_tot) / fac2;
      const double dd2 = (F1 - CPC::mpl * dd4 * c) * CPC::inv_tot;
      const double d[4] = {u[1], dd2, u[3], dd4};
#pragma unroll
      for (int j = 0; j < 4; j++) {
        w[j] = w[j] * a[i] + d[j] * CPC::dt;
        u[j] = u[j] + w[j] * b[i];
      }
    }
    return make_double4(u[0], u[1], u[2], u[3]);
  }

  // state: u = (x, xdot, theta, thetadot).  (gs, gc, rinv) = (g sin theta, g cos theta, 1 / fac2) ride along.
  static __device__ __forceinline__ void advance(double (&u)[4], double F) {
    double gs, gc;
    sincos(u[2], &gs, &gc);                                // anchor (full precision), once per env step
    gs *= CPC::g; gc *= CPC::g;
    double rinv = 1.0 / fma(-kCP.A1g, gc * gc, kCP.A0);
    const double Fmg = F * (CPC::inv_tot / CPC::g);        // F / (M g)
#pragma unroll 1
    for (int sub = 0; sub < CPC::nsteps; sub++) {
      if (__builtin_expect(fabs(u[3]) > kOmegaFast, 0)) {  // |d| could leave the polynomial range: exact substep
        const double4 r = substep_exact(make_double4(u[0], u[1], u[2], u[3]), F);
        u[0] = r.x; u[1] = r.y; u[2] = r.z; u[3] = r.w;
        sincos(u[2], &gs, &gc);
        gs *= CPC::g; gc *= CPC::g;
        rinv = 1.0 / fma(-kCP.A1g, gc * gc, kCP.A0);
        continue;
      }
      u[0] = fma(kCP.dtcb, u[1], u[0]);                    // x += dt (sum beta) xdot_0; the xddot terms follow per stage
      double W2, W3;                                       // w / dt of the reference (:41), theta and thetadot only
#pragma unroll
      for (int i = 0; i < 6; i++) {
        // ---- Diff (:161-168) in units of g: F1mg = F1 / (M g), num = g s - (F1 / M) c ----
        const double fac2 = fma(-kCP.A1g, gc * gc, kCP.A0);                // :162
        const double om = u[3];
        const double F1mg = fma(kCP.k2 * (om * om), gs, Fmg);              // :163 / (M g)
        const double num = fma(-F1mg, gc, gs);                             // :164 numerator
        const double e = fma(-fac2, rinv, 1.0);                            // third-order step on the carried 1 / fac2
        rinv = fma(rinv, fma(e, e, e), rinv);
        const double dd4 = num * rinv;                                     // thetaddot (:164)
        const double dd2g = fma(-kCP.k2 * dd4, gc, F1mg);                  // xddot / g (:165)
        // ---- x, xdot: outside the recurrence (nothing reads them back), weights of the unrolled scheme ----
        u[1] = fma(kCP.cv[i], dd2g, u[1]);
        if (i < 5) u[0] = fma(kCP.cx[i], dd2g, u[0]);                      // (gamma_5 = 0)
        // ---- theta, thetadot: low-storage RK update (:41-42) on W = w / dt; w starts at 0 and a[0] = 0 ----
        if (i == 0) {
          W2 = om; W3 = dd4;
        } else {
          W2 = fma(W2, kCP.a[i], om); W3 = fma(W3, kCP.a[i], dd4);
        }
        const double d = W2 * kCP.bdt[i];                                  // the angle increment of this stage
        u[2] = u[2] + d;
        u[3] = fma(W3, kCP.bdt[i], u[3]);
        // ---- rotate (g sin, g cos) by d: three shears, tau = tan(d / 2), sigma = sin d ----
        const double d2 = d * d;
        const double tau = d * fma(d2, kCP.t3, kCP.t1);
        const double sig = d * fma(d2, kCP.s3, 1.0);
        gs = fma(gc, tau, gs);
        gc = fma(-gs, sig, gc);
        gs = fma(gc, tau, gs);
      }
    }
  }

  static __device__ __forceinline__ float reward(const double (&u)[4]) {  // getReward :142
    return (fabs(u[2]) > CPC::PI_15 || fabs(u[0]) > 2.4) ? 0.0f : 1.0f;
  }
  static __device__ __forceinline__ void observe(const double (&u)[4], float (&o)[6]) {  // getobs :123-130
    double s, c;
    sincos(u[2], &s, &c);
    o[0] = (float)u[0]; o[1] = (float)u[1]; o[2] = (float)u[3]; o[3] = (float)u[2];
    o[4] = (float)c;    o[5] = (float)s;
  }
};

// ------------------------------------------------------------------------------------------- float
// Throughput path: everything in f32, sin/cos/reciprocal on the MUFU pipe.  NOT parity-gated: f32 state
// drifts ~1e-5 (median) over a 100-step episode against the double reference.
template <>
struct CartPoleDyn<float> {
  static __device__ __forceinline__ void advance(float (&u)[4], float F) {
    constexpr double a[6] = B2_RK_A;
    constexpr double b[6] = B2_RK_B;
    constexpr float dt = (float)CPC::dt, A0 = (float)CPC::A0, A1 = (float)CPC::A1, mpl = (float)CPC::mpl,
                    inv_tot = (float)CPC::inv_tot, g = (float)CPC::g;
    // (the RK update on packed FP32 -- FFMA2 / FMUL2 on the (x, xdot), (theta, thetadot) pairs, 12 -> 6 instructions per
    // stage -- was measured SLOWER, 4.25 vs 3.93 ms: the stage is one dependency chain and the packed ops lengthen it)
#pragma unroll 1
    for (int sub = 0; sub < CPC::nsteps; sub++) {
      float w0, w1, w2, w3;
#pragma unroll
      for (int i = 0; i < 6; i++) {
        float s, c;
        __sincosf(u[2], &s, &c);
        const float fac2 = fmaf(-A1, c * c, A0);
        const float om = u[3];
        const float F1 = fmaf(mpl * (om * om), s, F);
        const float num = fmaf(g, s, -(F1 * c) * inv_tot);
        const float dd4 = __fdividef(num, fac2);
        const float dd2 = fmaf(-mpl * dd4, c, F1) * inv_tot;
        if (i == 0) {
          w0 = u[1] * dt; w1 = dd2 * dt; w2 = om * dt; w3 = dd4 * dt;
        } else {
          w0 = fmaf(w0, (float)a[i], u[1] * dt); w1 = fmaf(w1, (float)a[i], dd2 * dt);
          w2 = fmaf(w2, (float)a[i], om * dt);   w3 = fmaf(w3, (float)a[i], dd4 * dt);
        }
        u[0] = fmaf(w0, (float)b[i], u[0]); u[1] = fmaf(w1, (float)b[i], u[1]);
        u[2] = fmaf(w2, (float)b[i], u[2]); u[3] = fmaf(w3, (float)b[i], u[3]);
      }
    }
  }
  static __device__ __forceinline__ float reward(const float (&u)[4]) {
    return (fabsf(u[2]) > (float)CPC::PI_15 || fabsf(u[0]) > 2.4f) ? 0.0f : 1.0f;
  }
  static __device__ __forceinline__ void observe(const float (&u)[4], float (&o)[6]) {
    float s, c;
    sincosf(u[2], &s, &c);
    o[0] = u[0]; o[1] = u[1]; o[2] = u[3]; o[3] = u[2]; o[4] = c; o[5] = s;
  }
};

// ------------------------------------------------------------------------------------------- mixed
// B2_PREC_MIXED: the STATE and the RK accumulators stay double, Diff -- sin, cos, the
// divide and the dozen multiply-adds around them, include/cart_pole.h:161-168 -- runs in f32 on the FP32 / MUFU pipes.
// A stage is then 6 DFMA (the low-storage recurrence for theta / thetadot, the unrolled weights for x / xdot, as in the
// double path) + 4 conversions on the FP64 pipe instead of 27 instructions: the kernel is bound by FP64 ISSUE, so this
// is the one variant faster than the double path whose error is BOUNDED rather than accumulated in the state: every
// stage's rounding enters through f (relative 2^-24 .. 2^-21 of the accelerations), never through u itself.
// Gate (tests/test_gpu_parity.py): teacher-forced single step <= 1e-5 of ||u||_inf against the reference's double
// states (survey probe: 8e-6 worst); the free-running drift over an episode is reported, like the f32 path's.
struct MixedPrec {};
template <typename P> struct StateOf { using type = P; };
template <> struct StateOf<MixedPrec> { using type = double; };

struct CPTableMixed { double cvm[6], cxm[6]; };   // xdot += cvm_j xddot_j, x += cxm_j xddot_j:  dt beta_j,  dt^2 gamma_j
__device__ __constant__ CPTableMixed kCPM = {
    {4e-4 * kRK.beta[0], 4e-4 * kRK.beta[1], 4e-4 * kRK.beta[2], 4e-4 * kRK.beta[3], 4e-4 * kRK.beta[4], 4e-4 * kRK.beta[5]},
    {4e-4 * 4e-4 * kRK.gamma[0], 4e-4 * 4e-4 * kRK.gamma[1], 4e-4 * 4e-4 * kRK.gamma[2], 4e-4 * 4e-4 * kRK.gamma[3],
     4e-4 * 4e-4 * kRK.gamma[4], 4e-4 * 4e-4 * kRK.gamma[5]}};

template <>
struct CartPoleDyn<MixedPrec> {
  static __device__ __forceinline__ void advance(double (&u)[4], double F) {
    constexpr float A0 = (float)CPC::A0, A1 = (float)CPC::A1, mpl = (float)CPC::mpl, inv_tot = (float)CPC::inv_tot,
                    g = (float)CPC::g;
    const float Ff = (float)F;                             // exact: the force is an f32 widened (scale_action)
#pragma unroll 1
    for (int sub = 0; sub < CPC::nsteps; sub++) {
      u[0] = fma(kCP.dtcb, u[1], u[0]);                    // x += dt (sum beta) xdot_0; the xddot terms follow per stage
      double W2, W3;
#pragma unroll
      for (int i = 0; i < 6; i++) {
        const double omd = u[3];
        const float th = (float)u[2], om = (float)omd;
        float s, c;
        sincosf(th, &s, &c);                               // (the MUFU pair __sincosf costs 1.3e-5 teacher-forced: outside the gate)
        const float fac2 = fmaf(-A1, c * c, A0);           // :162
        const float F1 = fmaf(mpl * (om * om), s, Ff);     // :163
        const float num = fmaf(g, s, -(F1 * c) * inv_tot); // :164
        const float dd4 = __fdividef(num, fac2);
        const float dd2 = fmaf(-mpl * dd4, c, F1) * inv_tot;   // :165
        const double dd2d = (double)dd2, dd4d = (double)dd4;
        u[1] = fma(kCPM.cvm[i], dd2d, u[1]);
        if (i < 5) u[0] = fma(kCPM.cxm[i], dd2d, u[0]);    // (gamma_5 = 0)
        if (i == 0) {
          W2 = omd; W3 = dd4d;
        } else {
          W2 = fma(W2, kCP.a[i], omd); W3 = fma(W3, kCP.a[i], dd4d);
        }
        u[2] = fma(W2, kCP.bdt[i], u[2]);
        u[3] = fma(W3, kCP.bdt[i], u[3]);
      }
    }
  }
  static __device__ __forceinline__ float reward(const double (&u)[4]) { return CartPoleDyn<double>::reward(u); }
  static __device__ __forceinline__ void observe(const double (&u)[4], float (&o)[6]) { CartPoleDyn<double>::observe(u, o); }
};

// ---------------------------------------------------------------------------------------- variants
// The branches the reference header documents but ships disabled (b200ppo.h B2_ENV_*): SWINGUP dynamics / reward /
// is_over, early termination inside advance(), random initial state.  Opt-in and never on the benchmark path, so the
// physics is written reference-shaped (sincos + true divisions per stage, include/cart_pole.h:149-170) rather than
// with the carried-rotation stage above; double only.  Out of line: one copy per kernel.
struct CartPoleVar {
  static __device__ __forceinline__ bool is_over(int fl, const double (&u)[4], int step) {   // :99-105
    if (fl & B2_ENV_SWINGUP) return step >= 500 || fabs(u[0]) > 2.4;
    return fabs(u[0]) > 2.4 || fabs(u[2]) > CPC::PI_15;
  }
  // advance (:107-118) with line :115 live when B2_ENV_EARLY_TERM is set; step = the value after step++ (:109)
  struct Adv { double4 u; int term; };   // by value in and out (see substep_exact)
  static __device__ __noinline__ Adv advance(double4 u4, double F, int fl, int step) {
    constexpr double a[6] = B2_RK_A;
    constexpr double b[6] = B2_RK_B;
    double u[4] = {u4.x, u4.y, u4.z, u4.w};
    int term = 0;
#pragma unroll 1
    for (int sub = 0; sub < CPC::nsteps && !term; sub++) {
      double w[4] = {0, 0, 0, 0};
#pragma unroll 1
      for (int i = 0; i < 6; i++) {
        double s, c, dd2, dd4;
        sincos(u[2], &s, &c);
        const double om = u[3];
        if (fl & B2_ENV_SWINGUP) {                                                              // :155-159
          const double fac1 = 1.0 / (CPC::mc + CPC::mp * s * s);
          const double fac2 = fac1 / CPC::l;
          dd2 = fac1 * (F + CPC::mp * s * (CPC::l * om * om + CPC::g * c));
          dd4 = fac2 * (-F * c - CPC::mp * CPC::l * om * om * c * s - (CPC::mc + CPC::mp) * CPC::g * s);
        } else {                                                                                // :161-165
          const double tot = CPC::mp + CPC::mc;
          const double fac2 = CPC::l * (4.0 / 3 - (CPC::mp * c * c) / tot);
          const double F1 = F + CPC::mp * CPC::l * om * om * s;
          dd4 = (CPC::g * s - F1 * c / tot) / fac2;
          dd2 = (F1 - CPC::mp * CPC::l * dd4 * c) / tot;
        }
        const double d[4] = {u[1], dd2, om, dd4};
#pragma unroll
        for (int j = 0; j < 4; j++) {
          w[j] = w[j] * a[i] + d[j] * CPC::dt;
          u[j] = u[j] + w[j] * b[i];
        }
      }
      if ((fl & B2_ENV_EARLY_TERM) && is_over(fl, u, step)) term = 1;                           // :115
    }
    Adv r;
    r.u = make_double4(u[0], u[1], u[2], u[3]);
    r.term = term;
    return r;
  }
  static __device__ __forceinline__ float reward(int fl, const double (&u)[4]) {                // :135-147
    if (!(fl & B2_ENV_SWINGUP)) return CartPoleDyn<double>::reward(u);
    double angle = fmod(u[2], 2 * 3.14159265358979323846);
    angle = angle < 0 ? angle + 2 * 3.14159265358979323846 : angle;
    return fabs(angle - 3.14159265358979323846) < 3.14159265358979323846 / 6 ? 1.0f : 0.0f;
  }
  // reset (:77-90): zero state, or (B2_ENV_RANDOM_INIT) U(-r, r)^4 from a Philox stream of its own, counter
  // (global env id, tick) (the CPU checker restates the same draw)
  static __device__ __forceinline__ void init_state(int fl, uint64_t seed, uint64_t env, uint64_t tick, double (&u)[4]) {
    if (!(fl & B2_ENV_RANDOM_INIT)) { u[0] = u[1] = u[2] = u[3] = 0.0; return; }
    const double r = (fl & B2_ENV_SWINGUP) ? 0.1 : 0.05;
    uint32_t xs[4];
    philox4x32_10((uint32_t)env, (uint32_t)(env >> 32), (uint32_t)tick, (uint32_t)(tick >> 32),
                  (uint32_t)seed ^ 0x494E4954u, (uint32_t)(seed >> 32), xs);
#pragma unroll
    for (int j = 0; j < 4; j++)
      u[j] = __dmul_rn(__dadd_rn(__dmul_rn(__dmul_rn(__dadd_rn((double)xs[j], 0.5), 1.0 / 4294967296.0), 2.0), -1.0), r);
  }
};

// action -> force, src/main.cpp:57-60,69: (a + 1) in f32, the rest in double, stored to f32, widened
__device__ __forceinline__ double scale_action(float a) {
  float ap1 = __fadd_rn(a, 1.0f);
  float scaled = (float)__dadd_rn(__dmul_rn((double)ap1, 10.0), -10.0);  // *(hi-lo)/2 + lo, exact rewrites
  return (double)scaled;
}

}  // namespace b2

// cartpole.cuh -- the cart-pole ODE and its 6-stage low-storage RK46-NL integrator, templated on the
// arithmetic type.  State lives in registers for the whole rollout; nothing here touches memory.
// Reference: include/cart_pole.h  (Diff :149-170 non-SWINGUP branch, rk46_nl :23-45, advance :107-118).
//
// One env step is 50 substeps x 6 stages = 300 evaluations of Diff, each needing sin/cos of the pole
// angle and one true division.  Calling libdevice sincos()/"/" 300 times costs ~150 instructions per
// stage, two thirds of them integer fix-up code (round-1 ncu, profiles/r1a_*: FP64 pipe 42 % busy, ALU
// 18 %, instruction fetch the top stall).  The double path therefore restates the two expensive pieces
// so that a stage is 32 straight-line DFMA/DMUL/DADD and almost nothing else:
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
struct CPTable {
  double a[6], bdt[6];          // RK46-NL a_i and b_i * dt  (the accumulator is kept as W = w / dt)
  double A0, A1, k1, g;         // fac2 = A0 - A1 c^2 ; k1 = mp l / M ; gravity
  double s3;                    // sin d = d + s3 d^3                  |error| <= d^5 / 120
  double c2, c4;                // cos d - 1 = d^2 (c2 + c4 d^2)       |error| <= d^6 / 720
};
__device__ __constant__ CPTable kCP = {
    B2_RK_A,
    {0.032918605146 * 4e-4, 0.823256998200 * 4e-4, 0.381530948900 * 4e-4, 0.200092213184 * 4e-4,
     1.718581042715 * 4e-4, 0.270000000000 * 4e-4},
    CPC::A0, CPC::A1, CPC::mpl* CPC::inv_tot, CPC::g,
    -1.0 / 6,
    -0.5, 1.0 / 24};

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

  // state: u = (x, xdot, theta, thetadot).  (s, c, rinv) = (sin theta, cos theta, 1/fac2) ride along.
  static __device__ __forceinline__ void advance(double (&u)[4], double F) {
    double s, c;
    sincos(u[2], &s, &c);                                  // anchor (full precision), once per env step
    double rinv = 1.0 / fma(-kCP.A1, c * c, kCP.A0);
    const double Fm = F * CPC::inv_tot;                    // F / M
#pragma unroll 1
    for (int sub = 0; sub < CPC::nsteps; sub++) {
      if (__builtin_expect(fabs(u[3]) > kOmegaFast, 0)) {  // |d| could leave the polynomial range: exact substep
        const double4 r = substep_exact(make_double4(u[0], u[1], u[2], u[3]), F);
        u[0] = r.x; u[1] = r.y; u[2] = r.z; u[3] = r.w;
        sincos(u[2], &s, &c);
        rinv = 1.0 / fma(-kCP.A1, c * c, kCP.A0);
        continue;
      }
      double W0, W1, W2, W3;                               // w / dt of the reference (:41)
#pragma unroll
      for (int i = 0; i < 6; i++) {
        // ---- Diff (:161-168), with F1m = F1 / M ----
        const double fac2 = fma(-kCP.A1, c * c, kCP.A0);                   // :162
        const double om = u[3];
        const double F1m = fma(kCP.k1 * (om * om), s, Fm);                 // :163 / M
        const double num = fma(-F1m, c, kCP.g * s);                        // :164 numerator
        const double e = fma(-fac2, rinv, 1.0);                            // one Newton step on the carried 1/fac2
        rinv = fma(rinv, e, rinv);
        double dd4 = num * rinv;
        dd4 = fma(fma(-dd4, fac2, num), rinv, dd4);                        // residual correction: num / fac2
        const double dd2 = fma(-kCP.k1 * dd4, c, F1m);                     // :165
        // ---- low-storage RK update (:41-42) on W = w / dt; w starts at 0 and a[0] = 0 ----
        if (i == 0) {
          W0 = u[1]; W1 = dd2; W2 = om; W3 = dd4;
        } else {
          W0 = fma(W0, kCP.a[i], u[1]); W1 = fma(W1, kCP.a[i], dd2);
          W2 = fma(W2, kCP.a[i], om);   W3 = fma(W3, kCP.a[i], dd4);
        }
        const double d = W2 * kCP.bdt[i];                                  // the angle increment of this stage
        u[0] = fma(W0, kCP.bdt[i], u[0]); u[1] = fma(W1, kCP.bdt[i], u[1]);
        u[2] = u[2] + d;                  u[3] = fma(W3, kCP.bdt[i], u[3]);
        // ---- rotate (sin, cos) by d ----
        const double d2 = d * d;
        const double sd = fma(d * d2, kCP.s3, d);                          // sin d
        const double cm1 = d2 * fma(d2, kCP.c4, kCP.c2);                   // cos d - 1
        const double s_new = fma(s, cm1, fma(c, sd, s));                   // s cos d + c sin d
        const double c_new = fma(c, cm1, fma(-s, sd, c));                  // c cos d - s sin d
        s = s_new; c = c_new;
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

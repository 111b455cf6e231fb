// cartpole.cuh -- the cart-pole ODE and its 6-stage low-storage RK46-NL integrator, templated on the
// arithmetic type.  State lives in registers for the whole rollout; nothing here touches memory.
// Reference: include/cart_pole.h  (Diff :149-170 non-SWINGUP branch, rk46_nl :23-45, advance :107-118).
#pragma once
#include "common.cuh"

namespace b2 {

template <typename R>
struct Phys;

template <>
struct Phys<double> {
  static __device__ __forceinline__ void sincos_(double x, double* s, double* c) { sincos(x, s, c); }
  static __device__ __forceinline__ double div_(double a, double b) { return a / b; }
  static __device__ __forceinline__ double abs_(double x) { return fabs(x); }
};
template <>
struct Phys<float> {
  static __device__ __forceinline__ void sincos_(float x, float* s, float* c) { sincosf(x, s, c); }
  static __device__ __forceinline__ float div_(float a, float b) { return a / b; }
  static __device__ __forceinline__ float abs_(float x) { return fabsf(x); }
};

// constants include/cart_pole.h:65-70
template <typename R>
struct CartPoleDyn {
  static constexpr double mp = 0.1, mc = 1.0, l = 0.5, g = 9.81, dt = 4e-4;
  static constexpr int nsteps = 50;

  // Diff (:161-168): u = (x, xdot, theta, thetadot), F = force
  static __device__ __forceinline__ void diff(const R (&u)[4], R F, R (&d)[4]) {
    R s, c;
    Phys<R>::sincos_(u[2], &s, &c);
    const R inv_tot = (R)(1.0 / (mp + mc));   // the reference divides by totMass; 1 ulp apart
    const R w = u[3];
    const R fac2 = (R)l * ((R)(4.0 / 3) - ((R)mp * c * c) * inv_tot);  // :162
    const R F1 = F + (R)(mp * l) * w * w * s;                          // :163
    d[3] = Phys<R>::div_((R)g * s - F1 * c * inv_tot, fac2);           // :164
    d[1] = (F1 - (R)(mp * l) * d[3] * c) * inv_tot;                    // :165
    d[0] = u[1];                                                       // :167
    d[2] = u[3];                                                       // :168
  }

  // one env step = 50 x rk46_nl (:110-116).  Coefficients :25-30.
  static __device__ __forceinline__ void advance(R (&u)[4], R F) {
    constexpr double a[6] = {0.000000000000,  -0.737101392796, -1.634740794341,
                             -0.744739003780, -1.469897351522, -2.813971388035};
    constexpr double b[6] = {0.032918605146, 0.823256998200, 0.381530948900,
                             0.200092213184, 1.718581042715, 0.270000000000};
#pragma unroll 1
    for (int sub = 0; sub < nsteps; sub++) {
      R w[4];
#pragma unroll
      for (int i = 0; i < 6; i++) {
        R d[4];
        diff(u, F, d);
#pragma unroll
        for (int j = 0; j < 4; j++) {
          w[j] = (i == 0) ? d[j] * (R)dt : w[j] * (R)a[i] + d[j] * (R)dt;  // :41 (w starts at 0, a[0] = 0)
          u[j] = u[j] + w[j] * (R)b[i];                                    // :42
        }
      }
    }
  }

  // getReward (:142): 1 - (|theta| > pi/15 || |x| > 2.4)
  static __device__ __forceinline__ float reward(const R (&u)[4]) {
    return (Phys<R>::abs_(u[2]) > (R)(3.14159265358979323846 / 15) || Phys<R>::abs_(u[0]) > (R)2.4) ? 0.0f : 1.0f;
  }

  // getobs (:123-130) narrowed to f32 once (src/main.cpp:79); cos/sin from the full-precision angle
  static __device__ __forceinline__ void observe(const R (&u)[4], float (&o)[6]) {
    R s, c;
    Phys<R>::sincos_(u[2], &s, &c);
    o[0] = (float)u[0]; o[1] = (float)u[1]; o[2] = (float)u[3]; o[3] = (float)u[2];
    o[4] = (float)c;    o[5] = (float)s;
  }
};

// action -> force, src/main.cpp:57-60,69: (a + 1) in f32, the rest in double, stored to f32, widened
__device__ __forceinline__ double scale_action(float a) {
  float ap1 = __fadd_rn(a, 1.0f);
  float scaled = (float)__dadd_rn(__dmul_rn((double)ap1, 10.0), -10.0);  // *(hi-lo)/2 + lo, exact rewrites
  return (double)scaled;
}

}  // namespace b2

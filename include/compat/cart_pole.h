// include/compat/cart_pole.h -- the reference's environment interface (include/cart_pole.h) on top of
// libb200ppo.so.  Two classes:
//
//   CartPoleBatch  n simulators resident on one GPU, advanced together by one kernel launch per step
//                  (b2_env_step / b2_env_observe).  This is the shape the hot path wants.
//   CartPole       the reference's own one-environment class, same method names and public fields
//                  (reset, advance, getobs, getReward, is_failed, is_over; u, F, t, step), implemented as a
//                  batch of one.  It exists so that code written against the reference header keeps compiling
//                  and so that the parity tests read like the reference's own usage (src/main.cpp:20-79);
//                  one launch per env step is NOT the fast way to use a GPU -- b2_rollout is.
//
// All arithmetic (RK46-NL over the cart-pole ODE, observations, reward) runs on the device; nothing is
// re-implemented on the host.  Header-only, C++17.  Errors throw std::runtime_error(b2_last_error()).
#ifndef B200PPO_COMPAT_CART_POLE_H
#define B200PPO_COMPAT_CART_POLE_H
#include <cassert>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../b200ppo.h"

// The reference selects its environment variant at compile time (`#define SWINGUP 0`, include/cart_pole.h:17) and keeps
// the random initial state (:85) and the early exit of advance() (:115) as commented-out lines.  Same switches here,
// defined before including this header: SWINGUP 1, B2_CARTPOLE_RANDOM_INIT 1, B2_CARTPOLE_EARLY_TERM 1.
#ifndef SWINGUP
#define SWINGUP 0
#endif
#ifndef B2_CARTPOLE_RANDOM_INIT
#define B2_CARTPOLE_RANDOM_INIT 0
#endif
#ifndef B2_CARTPOLE_EARLY_TERM
#define B2_CARTPOLE_EARLY_TERM 0
#endif
#define B2_CARTPOLE_ENV_FLAGS \
  ((SWINGUP ? B2_ENV_SWINGUP : 0) | (B2_CARTPOLE_RANDOM_INIT ? B2_ENV_RANDOM_INIT : 0) | (B2_CARTPOLE_EARLY_TERM ? B2_ENV_EARLY_TERM : 0))

// include/cart_pole.h:47-61 -- plain 4-vector: y1 = x, y2 = xdot, y3 = theta, y4 = thetadot
struct Vec4 {
  double y1, y2, y3, y4;
  Vec4(double a = 0, double b = 0, double c = 0, double d = 0) : y1(a), y2(b), y3(c), y4(d) {}
  Vec4 operator*(double v) const { return Vec4(y1 * v, y2 * v, y3 * v, y4 * v); }
  Vec4 operator+(const Vec4& o) const { return Vec4(y1 + o.y1, y2 + o.y2, y3 + o.y3, y4 + o.y4); }
};

class CartPoleBatch {
 public:
  explicit CartPoleBatch(int64_t n, int device = 0, int precision = B2_PREC_F64, int env_flags = B2_CARTPOLE_ENV_FLAGS,
                         uint64_t seed = 0x5EED)
      : n_(n) {
    b2_config c;
    ck(b2_config_default(&c), "b2_config_default");
    c.n_envs = n;
    c.horizon = 1;
    c.device = device;
    c.precision = precision;
    c.env_flags = env_flags;
    c.seed = seed;
    ck(b2_create(&c, &h_), "CartPoleBatch");
  }
  ~CartPoleBatch() { if (h_) b2_destroy(h_); }
  CartPoleBatch(const CartPoleBatch&) = delete;
  CartPoleBatch& operator=(const CartPoleBatch&) = delete;

  int64_t size() const { return n_; }
  void reset() { ck(b2_env_reset(h_), "reset"); }                                          // :77-90 on every env
  void advance(const std::vector<double>& force) {                                         // :107-118
    if ((int64_t)force.size() != n_) throw std::runtime_error("advance: one force per environment expected");
    ck(b2_env_step(h_, force.data(), n_), "advance");
  }
  std::vector<double> getobs() {                                                           // :120-133, [n][6]
    std::vector<double> o((size_t)6 * n_);
    ck(b2_env_observe(h_, o.data(), nullptr, n_), "getobs");
    return o;
  }
  std::vector<double> getReward() {                                                        // :135-147, [n]
    std::vector<double> r((size_t)n_);
    ck(b2_env_observe(h_, nullptr, r.data(), n_), "getReward");
    return r;
  }
  std::vector<uint8_t> terminated() {                                                      // advance()'s return value per env
    std::vector<uint8_t> t((size_t)n_);
    ck(b2_env_terminated(h_, t.data(), n_), "terminated");
    return t;
  }
  std::vector<Vec4> state() {
    std::vector<double> u((size_t)4 * n_);
    ck(b2_env_get_state(h_, u.data(), nullptr, n_), "state");
    std::vector<Vec4> out((size_t)n_);
    for (int64_t e = 0; e < n_; e++) out[(size_t)e] = Vec4(u[4 * e], u[4 * e + 1], u[4 * e + 2], u[4 * e + 3]);
    return out;
  }
  void set_state(const std::vector<Vec4>& s, const std::vector<int32_t>* steps = nullptr) {
    if ((int64_t)s.size() != n_) throw std::runtime_error("set_state: one state per environment expected");
    if (steps && (int64_t)steps->size() != n_) throw std::runtime_error("set_state: one step counter per environment expected");
    std::vector<double> u((size_t)4 * n_);
    for (int64_t e = 0; e < n_; e++) {
      u[4 * e] = s[(size_t)e].y1; u[4 * e + 1] = s[(size_t)e].y2; u[4 * e + 2] = s[(size_t)e].y3; u[4 * e + 3] = s[(size_t)e].y4;
    }
    ck(b2_env_set_state(h_, u.data(), steps ? steps->data() : nullptr, n_), "set_state");
  }
  b2_handle handle() const { return h_; }

 private:
  static void ck(int rc, const char* what) {
    if (rc != B2_OK) throw std::runtime_error(std::string(what) + ": " + b2_last_error());
  }
  b2_handle h_ = nullptr;
  int64_t n_;
};

class CartPole {
 public:
  const double mp = 0.1, mc = 1, l = 0.5, g = 9.81, dt = 4e-4;  // include/cart_pole.h:65-69
  const int nsteps = 50;                                         // :70
  int step = 0;
  Vec4 u;
  double F = 0, t = 0;

  explicit CartPole(int device = 0) : sim_(1, device) {}

  void reset() {  // :77-90 (upstream the random initial state is commented out: exact zero state unless B2_CARTPOLE_RANDOM_INIT)
    sim_.reset();
    u = sim_.state()[0];
    step = 0; F = 0; t = 0;
  }
#if SWINGUP
  bool is_failed() const { return std::fabs(u.y1) > 2.4; }                                 // :93-94
  bool is_over() const { return step >= 500 || std::fabs(u.y1) > 2.4; }                    // :100-101
#else
  bool is_failed() const { return std::fabs(u.y1) > 2.4 || std::fabs(u.y3) > M_PI / 15; }  // :96
  bool is_over() const { return is_failed(); }                                             // :103
#endif
  int advance(std::vector<double> action) {  // :107-118; returns 0 unless B2_CARTPOLE_EARLY_TERM (the early exit, :115)
    F = action.at(0);
    const std::vector<int32_t> before{step};
    step++;
    sim_.set_state({u}, &before);  // `u` and `step` are public fields upstream: honour a caller that wrote to them
    sim_.advance({F});
    u = sim_.state()[0];
    t += dt * nsteps;
    return sim_.terminated()[0];
  }
  std::vector<double> getobs(const int size = 6) {  // :120-133
    assert(size == 4 || size == 6);
    std::vector<double> o = sim_.getobs();
    o.resize((size_t)size);
    return o;
  }
  double getReward() { return sim_.getReward()[0]; }  // :135-147

 private:
  CartPoleBatch sim_;
};

#endif  // B200PPO_COMPAT_CART_POLE_H

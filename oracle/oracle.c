/* oracle/oracle.c -- plain-C restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY
 * (see oracle.h).  Written from the behaviour of the reference, with the reference line that fixes
 * each arithmetic choice cited beside it (paths relative to /root/reference).  Build with
 * -ffp-contract=off: the reference build sets no FMA flags (CMakeLists.txt:11).
 */
#include "oracle.h"

#include <alloca.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#ifndef M_E
#define M_E 2.7182818284590452354
#endif

/* ------------------------------------------------------------------ shapes */
int orc_net_params(const orc_dims* d, int out) {
  return d->H1 * d->S + d->H1 + d->H2 * d->H1 + d->H2 + out * d->H2 + out;
}
int orc_n_params(const orc_dims* d) { return orc_net_params(d, d->A) + orc_net_params(d, 1); }

/* ------------------------------------------------------------- environment */
/* include/cart_pole.h:65-70 */
static const double MP = 0.1, MC = 1, LEN = 0.5, GRAV = 9.81, DT = 4e-4;
static const int NSUB = 50;

void orc_cartpole_diff(const double u[4], double F, double du[4]) {
  const double cosy = cos(u[2]), siny = sin(u[2]); /* :152 */
  const double w = u[3];                           /* :153 */
  const double totMass = MP + MC;                  /* :161 */
  const double fac2 = LEN * (4.0 / 3 - (MP * cosy * cosy) / totMass); /* :162 */
  const double F1 = F + MP * LEN * w * w * siny;                      /* :163 */
  du[3] = (GRAV * siny - F1 * cosy / totMass) / fac2;                 /* :164 */
  du[1] = (F1 - MP * LEN * du[3] * cosy) / totMass;                   /* :165 */
  du[0] = u[1];                                                       /* :167 */
  du[2] = u[3];                                                       /* :168 */
}

void orc_rk46_nl(double u[4], double dt, double F) {
  /* coefficients include/cart_pole.h:25-30; c[] (:31-33) only feeds the unused time argument */
  static const double a[6] = {0.000000000000,  -0.737101392796, -1.634740794341,
                              -0.744739003780, -1.469897351522, -2.813971388035};
  static const double b[6] = {0.032918605146, 0.823256998200, 0.381530948900,
                              0.200092213184, 1.718581042715, 0.270000000000};
  double w[4] = {0, 0, 0, 0}; /* Vec w; default-constructed zeros (:35, :51) */
  for (int i = 0; i < 6; i++) {
    double d[4];
    orc_cartpole_diff(u, F, d);
    for (int j = 0; j < 4; j++) {
      w[j] = w[j] * a[i] + d[j] * dt; /* :41 */
      u[j] = u[j] + w[j] * b[i];      /* :42 */
    }
  }
}

void orc_cartpole_advance(double u[4], double F) {
  for (int i = 0; i < NSUB; i++) orc_rk46_nl(u, DT, F); /* :110-116; never terminates early (:115,:117) */
}

void orc_cartpole_obs(const double u[4], float obs[6]) {
  obs[0] = (float)u[0];      /* :123 */
  obs[1] = (float)u[1];      /* :124 */
  obs[2] = (float)u[3];      /* :125  angular velocity BEFORE angle */
  obs[3] = (float)u[2];      /* :126 */
  obs[4] = (float)cos(u[2]); /* :129  from the fp64 angle, narrowed once (src/main.cpp:79) */
  obs[5] = (float)sin(u[2]); /* :130 */
}

double orc_cartpole_reward(const double u[4]) {
  return 1 - (fabs(u[2]) > M_PI / 15 || fabs(u[0]) > 2.4); /* :142 */
}

/* ---- variants (documented in the header, disabled in the shipped build) ---- */
void orc_cartpole_diff_v(int fl, const double u[4], double F, double du[4]) {
  if (!(fl & ORC_ENV_SWINGUP)) { orc_cartpole_diff(u, F, du); return; }
  const double cosy = cos(u[2]), siny = sin(u[2]); /* :152 */
  const double w = u[3];
  const double fac1 = 1 / (MC + MP * siny * siny);                                   /* :155 */
  const double fac2 = fac1 / LEN;                                                    /* :156 */
  du[1] = fac1 * (F + MP * siny * (LEN * w * w + GRAV * cosy));                      /* :157 */
  du[3] = fac2 * (-F * cosy - MP * LEN * w * w * cosy * siny - (MC + MP) * GRAV * siny); /* :158-159 */
  du[0] = u[1];
  du[2] = u[3];
}

int orc_cartpole_is_over(int fl, const double u[4], int step) {
  if (fl & ORC_ENV_SWINGUP) return step >= 500 || fabs(u[0]) > 2.4;                  /* :101 */
  return fabs(u[0]) > 2.4 || fabs(u[2]) > M_PI / 15;                                 /* :103 */
}

int orc_cartpole_advance_v(int fl, double u[4], int step, double F) {
  static const double a[6] = {0.000000000000,  -0.737101392796, -1.634740794341,
                              -0.744739003780, -1.469897351522, -2.813971388035};
  static const double b[6] = {0.032918605146, 0.823256998200, 0.381530948900,
                              0.200092213184, 1.718581042715, 0.270000000000};
  for (int s = 0; s < NSUB; s++) {
    double w[4] = {0, 0, 0, 0};
    for (int i = 0; i < 6; i++) {
      double d[4];
      orc_cartpole_diff_v(fl, u, F, d);
      for (int j = 0; j < 4; j++) {
        w[j] = w[j] * a[i] + d[j] * DT;
        u[j] = u[j] + w[j] * b[i];
      }
    }
    if ((fl & ORC_ENV_EARLY_TERM) && orc_cartpole_is_over(fl, u, step)) return 1;    /* :115 */
  }
  return 0;
}

double orc_cartpole_reward_v(int fl, const double u[4]) {
  if (!(fl & ORC_ENV_SWINGUP)) return orc_cartpole_reward(u);
  double angle = fmod(u[2], 2 * M_PI);                                               /* :137 */
  angle = angle < 0 ? angle + 2 * M_PI : angle;                                      /* :138 */
  return fabs(angle - M_PI) < M_PI / 6 ? 1 : 0;                                      /* :139 */
}

void orc_cartpole_init_state(int fl, uint64_t seed, uint64_t env, uint64_t tick, double u[4]) {
  if (!(fl & ORC_ENV_RANDOM_INIT)) { u[0] = u[1] = u[2] = u[3] = 0; return; }        /* :86 */
  const double r = (fl & ORC_ENV_SWINGUP) ? 0.1 : 0.05;                              /* :81, :83 */
  uint32_t ctr[4] = {(uint32_t)env, (uint32_t)(env >> 32), (uint32_t)tick, (uint32_t)(tick >> 32)};
  uint32_t key[2] = {(uint32_t)seed ^ 0x494E4954u, (uint32_t)(seed >> 32)};          /* a stream of its own: "INIT" */
  uint32_t x[4];
  orc_philox4x32_10(ctr, key, x);
  for (int j = 0; j < 4; j++) u[j] = (((double)x[j] + 0.5) * (1.0 / 4294967296.0) * 2 - 1) * r;
}

double orc_scale_action(float a) {
  /* src/main.cpp:57-60: float + int in float, then double arithmetic, stored to float; :69 widen */
  const double hi = 10, lo = -10; /* :18 */
  float ap1 = a + 1;
  float scaled = (float)((double)ap1 * (hi - lo) / 2 + lo);
  return (double)scaled;
}

/* ------------------------------------------------------------------ policy */
static void linear(const float* W, const float* b, int out, int in, const float* x, float* y) {
  for (int o = 0; o < out; o++) {
    float acc = b[o];
    for (int i = 0; i < in; i++) acc += W[o * in + i] * x[i]; /* torch Linear: y = x W^T + b, f32 */
    y[o] = acc;
  }
}

void orc_mlp_forward(const orc_dims* d, const float* net, int out, int final_tanh, const float* obs,
                     float* h1o, float* h2o, float* y) {
  const float* W1 = net;
  const float* b1 = W1 + d->H1 * d->S;
  const float* W2 = b1 + d->H1;
  const float* b2 = W2 + d->H2 * d->H1;
  const float* W3 = b2 + d->H2;
  const float* b3 = W3 + out * d->H2;
  float* h1 = h1o ? h1o : (float*)alloca(sizeof(float) * d->H1);
  float* h2 = h2o ? h2o : (float*)alloca(sizeof(float) * d->H2);
  linear(W1, b1, d->H1, d->S, obs, h1);
  for (int i = 0; i < d->H1; i++) h1[i] = tanhf(h1[i]); /* ppo_nn.h:85-90 */
  linear(W2, b2, d->H2, d->H1, h1, h2);
  for (int i = 0; i < d->H2; i++) h2[i] = tanhf(h2[i]); /* :91-96 */
  linear(W3, b3, out, d->H2, h2, y);
  if (final_tanh)
    for (int i = 0; i < out; i++) y[i] = tanhf(y[i]); /* :98-102 actor only */
}

float orc_logprob(float a, float mu, double sigma) {
  /* src/ppo_nn.cpp:20-49 with k = 1, Sigma = [[sigma^2]] held as f32 (ppo_nn.h:124-126) */
  float var = (float)(sigma * sigma);
  float diff = a - mu;                         /* :20 */
  float inv = 1.0f / var;                      /* :24 covar.inverse() */
  float numer = -0.5f * (diff * (inv * diff)); /* :30-31 */
  float denom = (logf(var) + (float)(log(2 * M_PI) * 1)) / 2; /* :42-44 */
  return numer - denom;                        /* :49 */
}

float orc_entropy_const(double sigma) {
  float var = (float)(sigma * sigma);
  float v1 = (float)pow(2 * M_PI * M_E, 1);    /* :56 -- adds (2 pi e)^k, not its log (Q8) */
  return 0.5f * (logf(var) + v1);              /* :64 */
}

void orc_act(const orc_dims* d, const float* params, const float* obs, int n, const float* eps,
             double sigma, float* mu, float* action, float* logp) {
  for (int i = 0; i < n; i++) {
    float m;
    orc_mlp_forward(d, params, d->A, 1, obs + (long)i * d->S, NULL, NULL, &m); /* :100 */
    float a = (float)((double)m + sigma * (double)(eps ? eps[i] : 0.f));       /* :107-108 */
    if (mu) mu[i] = m;
    if (action) action[i] = a;
    if (logp) logp[i] = orc_logprob(a, m, sigma);                              /* :115 */
  }
}

void orc_value(const orc_dims* d, const float* params, const float* obs, int n, float* v) {
  const float* critic = params + orc_net_params(d, d->A);
  for (int i = 0; i < n; i++)
    orc_mlp_forward(d, critic, 1, 0, obs + (long)i * d->S, NULL, NULL, v + i); /* :189 */
}

/* ----------------------------------------------------------------- returns */
void orc_returns(const orc_dims* d, const float* params, const float* obs, const float* reward,
                 const uint8_t* flags, long B, double gamma, float* R) {
  const float* critic = params + orc_net_params(d, d->A);
  float g = (float)gamma; /* gamma * f32 tensor: scalar is narrowed to the tensor dtype (Q6) */
  float disc = 0.0f;      /* :215 torch::tensor({0.0}) is Float */
  for (long k = B - 1; k >= 0; k--) { /* :211-214 reverse, :220 */
    if (flags[k] & 6) {               /* :226-228 timeup (or horizon cut): critic at the PRE-step state (Q5) */
      orc_mlp_forward(d, critic, 1, 0, obs + k * d->S, NULL, NULL, &disc);
    } else if (flags[k] & 1) {        /* :229-230 */
      disc = 0.0f;
    }
    float gd = g * disc;              /* :231 */
    disc = reward[k] + gd;
    R[k] = disc;                      /* :232 insert at front */
  }
}

/* ----------------------------------------------------------- loss and grad */
void orc_loss_grad(const orc_dims* d, const float* params, const float* obs, const float* action,
                   const float* old_logp, const float* R, const long* idx, long nb, double sigma,
                   double eps_clip, orc_loss* out, float* grads) {
  const int S = d->S, H1 = d->H1, H2 = d->H2;
  const int P = orc_n_params(d), PA = orc_net_params(d, d->A);
  double* G = (double*)calloc(P, sizeof(double));
  double s1 = 0, s2 = 0, nclip = 0;
  const float lo = (float)(1 - eps_clip), hi = (float)(1 + eps_clip);
  const float inv_var = 1.0f / (float)(sigma * sigma);
#pragma omp parallel
  {
    double* g = (double*)calloc(P, sizeof(double));
    float* h1 = (float*)malloc(sizeof(float) * H1);
    float* h2 = (float*)malloc(sizeof(float) * H2);
    float* d1 = (float*)malloc(sizeof(float) * H1);
    float* d2 = (float*)malloc(sizeof(float) * H2);
    double ls1 = 0, ls2 = 0, lclip = 0;
#pragma omp for schedule(static)
    for (long q = 0; q < nb; q++) {
      long i = idx ? idx[q] : q;
      const float* x = obs + i * S;
      for (int net = 0; net < 2; net++) { /* 0 actor, 1 critic: src/ppo_nn.cpp:130,189 */
        const float* base = params + (net ? PA : 0);
        double* gb = g + (net ? PA : 0);
        const float* W2 = base + H1 * S + H1;
        const float* W3 = W2 + H2 * H1 + H2;
        float y;
        orc_mlp_forward(d, base, 1, net == 0, x, h1, h2, &y);
        float dz3; /* d(mean loss)/d(pre-activation of the output unit) */
        if (net == 0) {
          float mu = y;
          float logp = orc_logprob(action[i], mu, sigma);
          float ratio = expf(logp - old_logp[i]);                  /* :263 */
          /* A uses the CURRENT critic, no grad (:274, Q7) */
          float V;
          orc_mlp_forward(d, params + PA, 1, 0, x, NULL, NULL, &V);
          float A = R[i] - V;
          float surr1 = ratio * A;                                 /* :284 */
          float rc = ratio < lo ? lo : (ratio > hi ? hi : ratio);  /* :285 */
          float surr2 = rc * A;
          float l1 = -(surr1 < surr2 ? surr1 : surr2);             /* :287 */
          ls1 += l1;
          /* autograd of min(): ties split 1/2+1/2, clamp passes grad only inside [lo,hi] */
          float gr;
          int inside = (ratio >= lo && ratio <= hi);
          if (surr1 < surr2) gr = A;
          else if (surr1 > surr2) gr = inside ? A : 0.0f;
          else gr = 0.5f * A + (inside ? 0.5f * A : 0.0f);
          if (!inside) lclip += 1;
          float dlogp = -gr * ratio / (float)nb;                   /* d(-min)/dlogp, mean over nb (:302) */
          float dmu = dlogp * (action[i] - mu) * inv_var;          /* :20-31 */
          dz3 = dmu * (1.0f - mu * mu);                            /* tanh3, ppo_nn.h:102 */
        } else {
          float V = y;
          float e = V - R[i];
          ls2 += (double)e * e;                                    /* :288 0.5*MSE */
          dz3 = e / (float)nb;                                     /* d(0.5*mean e^2)/dV */
        }
        /* backprop through Linear3, tanh2, Linear2, tanh1, Linear1 */
        double* gW1 = gb;
        double* gb1 = gW1 + H1 * S;
        double* gW2 = gb1 + H1;
        double* gb2 = gW2 + H2 * H1;
        double* gW3 = gb2 + H2;
        double* gb3 = gW3 + H2;
        for (int j = 0; j < H2; j++) {
          gW3[j] += (double)dz3 * h2[j];
          d2[j] = dz3 * W3[j] * (1.0f - h2[j] * h2[j]);
        }
        gb3[0] += dz3;
        for (int k = 0; k < H1; k++) d1[k] = 0.0f;
        for (int j = 0; j < H2; j++) {
          gb2[j] += d2[j];
          for (int k = 0; k < H1; k++) {
            gW2[j * H1 + k] += (double)d2[j] * h1[k];
            d1[k] += d2[j] * W2[j * H1 + k];
          }
        }
        for (int k = 0; k < H1; k++) {
          float dz1 = d1[k] * (1.0f - h1[k] * h1[k]);
          gb1[k] += dz1;
          for (int s = 0; s < S; s++) gW1[k * S + s] += (double)dz1 * x[s];
        }
      }
    }
#pragma omp critical
    {
      for (int p = 0; p < P; p++) G[p] += g[p];
      s1 += ls1; s2 += ls2; nclip += lclip;
    }
    free(g); free(h1); free(h2); free(d1); free(d2);
  }
  if (grads) for (int p = 0; p < P; p++) grads[p] = (float)G[p];
  if (out) {
    out->loss1 = s1 / nb;
    out->loss2 = 0.5 * s2 / nb;
    out->loss3 = -0.01 * orc_entropy_const(sigma); /* :289 */
    out->loss = out->loss1 + out->loss2 + out->loss3; /* :290,:302 */
    out->clip_frac = nclip / nb;
  }
  free(G);
}

/* -------------------------------------------------------------------- Adam */
void orc_adam(float* p, const float* g, float* m, float* v, long n, long step, double lr, double b1,
              double b2, double eps) {
  /* libtorch torch/csrc/api/src/optim/adam.cpp (non-amsgrad, weight_decay 0); scalars are doubles
   * narrowed to f32 inside the tensor kernels. */
  double bc1 = 1 - pow(b1, (double)step);
  double bc2 = 1 - pow(b2, (double)step);
  float fb1 = (float)b1, f1mb1 = (float)(1 - b1), fb2 = (float)b2, f1mb2 = (float)(1 - b2);
  float sq_bc2 = (float)sqrt(bc2), feps = (float)eps, step_size = (float)(-(lr / bc1));
  for (long i = 0; i < n; i++) {
    m[i] = m[i] * fb1 + f1mb1 * g[i];            /* exp_avg.mul_(beta1).add_(grad, 1-beta1) */
    v[i] = v[i] * fb2 + f1mb2 * (g[i] * g[i]);   /* exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2) */
    float denom = sqrtf(v[i]) / sq_bc2 + feps;   /* (exp_avg_sq.sqrt()/sqrt(bc2)).add_(eps) */
    p[i] = p[i] + step_size * (m[i] / denom);    /* p.addcdiv_(exp_avg, denom, -step_size) */
  }
}

/* ------------------------------------------------------------------ update */
void orc_update(const orc_dims* d, float* params, float* m, float* v, long* step, const float* obs,
                const float* action, const float* old_logp, const float* reward,
                const uint8_t* flags, long B, int K, double sigma, double eps_clip, double gamma,
                double lr, double b1, double b2, orc_loss* per_epoch, float* last_grads,
                float* returns_out) {
  int P = orc_n_params(d);
  float* R = (float*)malloc(sizeof(float) * B);
  float* g = (float*)malloc(sizeof(float) * P);
  orc_returns(d, params, obs, reward, flags, B, gamma, R); /* once, at entry (:207-236) */
  if (returns_out) memcpy(returns_out, R, sizeof(float) * B);
  for (int k = 0; k < K; k++) {                            /* :245 */
    orc_loss l;
    orc_loss_grad(d, params, obs, action, old_logp, R, NULL, B, sigma, eps_clip, &l, g);
    if (per_epoch) per_epoch[k] = l;
    *step += 1;                                            /* persists across calls (optimizer made once) */
    orc_adam(params, g, m, v, P, *step, lr, b1, b2, 1e-8);
  }
  if (last_grads) memcpy(last_grads, g, sizeof(float) * P);
  free(R); free(g);
}

/* ------------------------------------------------------------------- noise */
void orc_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3], k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; r++) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

float orc_gauss(uint64_t seed, uint64_t env, uint64_t step) {
  uint32_t ctr[4] = {(uint32_t)env, (uint32_t)(env >> 32), (uint32_t)step, (uint32_t)(step >> 32)};
  uint32_t key[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
  uint32_t x[4];
  orc_philox4x32_10(ctr, key, x);
  float u1 = ((float)(x[0] >> 9) + 0.5f) * 1.1920928955078125e-07f; /* (0,1), exact in f32 */
  float u2 = ((float)(x[1] >> 9) + 0.5f) * 1.1920928955078125e-07f;
  float r = sqrtf(-2.0f * logf(u1));
  return r * (float)cos(M_PI * (double)(2.0f * u2)); /* engine: r * cospif(2*u2) */
}

/* ----------------------------------------------------------------- rollout */
void orc_rollout_v(const orc_dims* d, const float* params, long N, int T, int n_timestep, double sigma,
                   uint64_t seed, uint64_t env0, uint64_t step0, const float* forced_action, int env_flags,
                   double* state, int* ep_step, double* ep_ret, float* obs, float* action,
                   float* logp, float* mu, float* reward, uint8_t* flags, double* state_trace) {
#pragma omp parallel for schedule(static)
  for (long e = 0; e < N; e++) {
    double u[4] = {state[4 * e], state[4 * e + 1], state[4 * e + 2], state[4 * e + 3]};
    int es = ep_step[e];
    double er = ep_ret ? ep_ret[e] : 0;
    for (int t = 0; t < T; t++) {
      long k = (long)t * N + e;
      float o[6], m, a;
      orc_cartpole_obs(u, o);
      orc_mlp_forward(d, params, d->A, 1, o, NULL, NULL, &m);
      if (forced_action) a = forced_action[k];
      else a = (float)((double)m + sigma * (double)orc_gauss(seed, env0 + e, step0 + t));
      float lp = orc_logprob(a, m, sigma);
      es += 1;                                                                   /* cart_pole.h:109 */
      int term = orc_cartpole_advance_v(env_flags, u, es, orc_scale_action(a));
      double r = orc_cartpole_reward_v(env_flags, u);
      er += r;
      int timeup = !term && (es == n_timestep);                                  /* src/main.cpp:81-88 */
      uint8_t f = (uint8_t)(term ? 1 : (timeup ? 2 : (t == T - 1 ? 4 : 0)));
      if (obs) memcpy(obs + k * 6, o, sizeof o);
      if (action) action[k] = a;
      if (logp) logp[k] = lp;
      if (mu) mu[k] = m;
      if (reward) reward[k] = (float)r;
      if (flags) flags[k] = f;
      if (state_trace) memcpy(state_trace + k * 4, u, sizeof u);
      if (term || timeup) {
        orc_cartpole_init_state(env_flags, seed, env0 + e, step0 + (uint64_t)t + 1, u);
        es = 0; er = 0;
      }
    }
    memcpy(state + 4 * e, u, sizeof u);
    ep_step[e] = es;
    if (ep_ret) ep_ret[e] = er;
  }
}

void orc_rollout(const orc_dims* d, const float* params, long N, int T, int n_timestep, double sigma,
                 uint64_t seed, uint64_t env0, uint64_t step0, const float* forced_action,
                 double* state, int* ep_step, double* ep_ret, float* obs, float* action,
                 float* logp, float* mu, float* reward, uint8_t* flags, double* state_trace) {
#pragma omp parallel for schedule(static)
  for (long e = 0; e < N; e++) {
    double u[4] = {state[4 * e], state[4 * e + 1], state[4 * e + 2], state[4 * e + 3]};
    int es = ep_step[e];
    double er = ep_ret ? ep_ret[e] : 0;
    for (int t = 0; t < T; t++) {
      long k = (long)t * N + e;
      float o[6], m, a;
      orc_cartpole_obs(u, o);
      orc_mlp_forward(d, params, d->A, 1, o, NULL, NULL, &m);
      if (forced_action) a = forced_action[k];
      else a = (float)((double)m + sigma * (double)orc_gauss(seed, env0 + e, step0 + t));
      float lp = orc_logprob(a, m, sigma);
      orc_cartpole_advance(u, orc_scale_action(a));
      double r = orc_cartpole_reward(u);
      es += 1;
      er += r;
      int timeup = (es == n_timestep); /* src/main.cpp:86-88 */
      uint8_t f = (uint8_t)((timeup ? 2 : 0) | ((!timeup && t == T - 1) ? 4 : 0));
      if (obs) memcpy(obs + k * 6, o, sizeof o);
      if (action) action[k] = a;
      if (logp) logp[k] = lp;
      if (mu) mu[k] = m;
      if (reward) reward[k] = (float)r;
      if (flags) flags[k] = f;
      if (state_trace) memcpy(state_trace + k * 4, u, sizeof u);
      if (timeup) { u[0] = u[1] = u[2] = u[3] = 0; es = 0; er = 0; } /* cart_pole.h:86; main.cpp:26-27 */
    }
    memcpy(state + 4 * e, u, sizeof u);
    ep_step[e] = es;
    if (ep_ret) ep_ret[e] = er;
  }
}

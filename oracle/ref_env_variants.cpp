// oracle/ref_env_variants.cpp -- mints known-answer vectors for the cart-pole variants the reference documents but
// ships disabled: the SWINGUP branches, the random initial state and the early
// termination inside advance().  TEST INFRASTRUCTURE: compiled by oracle/Makefile against the reference's own
// headers where they lie (CP_HEADER = the header under /root/reference to include):
//   swingup        include/cart_pole.h with its `#define SWINGUP 0` line switched to 1 by sed at build time
//                  (the generated file lives only in the git-ignored oracle/_ref/gen/ while the compile runs)
//   early          archived/single_environment/cart_pole.h as is: `if (is_over()) return 1;` live (:119)
//   early_swingup  the archived header with SWINGUP switched to 1 the same way
// Output: <dir>/envvar_<name>.f64 = [E][T][8] doubles: F, y1, y2, y3, y4, reward, terminate, step (post-step values);
// after a terminate the env is reset() and its state overwritten with the next scripted initial state, which the
// file <dir>/envvar_<name>_init.f64 = [E][R][4] lists in order of use (R slots; slot 0 = the first state).
#include CP_HEADER
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

static uint64_t lcg = 0x9E3779B97F4A7C15ull;
static double unif() {  // (0,1), deterministic
  lcg = lcg * 6364136223846793005ull + 1442695040888963407ull;
  return ((double)(lcg >> 11) + 0.5) / 9007199254740992.0;
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s <name> <outdir>\n", argv[0]); return 2; }
  const std::string name = argv[1], dir = argv[2];
  const int E = 8, T = 620, R = 64;
  const double range = SWINGUP ? 0.1 : 0.05;  // cart_pole.h:80-84
  std::vector<double> out((size_t)E * T * 8), init((size_t)E * R * 4);
  for (int e = 0; e < E; e++) {
    for (int r = 0; r < R; r++)
      for (int j = 0; j < 4; j++) init[((size_t)e * R + r) * 4 + j] = (r == 0 && e == 0) ? 0.0 : (2 * unif() - 1) * range;
    CartPole env;
    int slot = 0;
    env.reset();
    env.u = Vec4(init[((size_t)e * R) * 4], init[((size_t)e * R) * 4 + 1], init[((size_t)e * R) * 4 + 2], init[((size_t)e * R) * 4 + 3]);
    // force script: env 0..3 gentle (long episodes, reaches step 500), the rest random in [-10, 10] with a bias
    const double amp = e < 4 ? 0.5 : 10.0, bias = (e % 3 - 1) * (e < 4 ? 0.05 : 1.5);
    for (int t = 0; t < T; t++) {
      const double F = (double)(float)((2 * unif() - 1) * amp + bias);
      const int term = env.advance(std::vector<double>{F});
      double* o = &out[((size_t)e * T + t) * 8];
      o[0] = F; o[1] = env.u.y1; o[2] = env.u.y2; o[3] = env.u.y3; o[4] = env.u.y4;
      o[5] = env.getReward(); o[6] = term; o[7] = env.step;
      if (term) {
        env.reset();
        slot = (slot + 1) % R;
        const double* s = &init[((size_t)e * R + slot) * 4];
        env.u = Vec4(s[0], s[1], s[2], s[3]);
      }
    }
  }
  FILE* f = fopen((dir + "/envvar_" + name + ".f64").c_str(), "wb");
  if (!f) return 1;
  fwrite(out.data(), sizeof(double), out.size(), f); fclose(f);
  f = fopen((dir + "/envvar_" + name + "_init.f64").c_str(), "wb");
  if (!f) return 1;
  fwrite(init.data(), sizeof(double), init.size(), f); fclose(f);
  int terms = 0;
  for (size_t i = 6; i < out.size(); i += 8) terms += out[i] != 0;
  printf("envvar_%s: %d envs x %d steps, %d terminations\n", name.c_str(), E, T, terms);
  return 0;
}

// tools/compat_variants_check.cpp -- include/compat/cart_pole.h with the reference's compile-time env switches on
// (built with -DSWINGUP=1 -DB2_CARTPOLE_EARLY_TERM=1): one CartPole driven through the known-answer vector minted from
// the reference's own archived header with SWINGUP 1 (tests/golden/envvar_early_swingup*, oracle/ref_env_variants.cpp).
// usage: compat_variants_check <golden dir>     prints one JSON line, exit 0 iff every check holds.
#include <cmath>
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

#include "compat/cart_pole.h"

static std::vector<double> load(const std::string& path) {
  std::ifstream f(path, std::ios::binary | std::ios::ate);
  if (!f) throw std::runtime_error("cannot open " + path);
  std::vector<double> v((size_t)f.tellg() / sizeof(double));
  f.seekg(0);
  f.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(v.size() * sizeof(double)));
  return v;
}

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  try {
    const std::string dir = argv[1];
    const auto kat = load(dir + "/envvar_early_swingup.f64");        // [8][620][8]
    const auto init = load(dir + "/envvar_early_swingup_init.f64");  // [8][64][4]
    const int T = 620, R = 64;
    double worst = 0;
    int terms = 0, mism = 0;
    for (int e : {0, 5}) {   // one gentle env (runs to step 500) and one hard-driven env
      CartPole env;
      env.reset();
      int slot = 0;
      const double* s0 = &init[((size_t)e * R) * 4];
      env.u = Vec4(s0[0], s0[1], s0[2], s0[3]);
      for (int t = 0; t < T; t++) {
        const double* k = &kat[((size_t)e * T + t) * 8];
        const int term = env.advance({k[0]});
        const double ref[4] = {k[1], k[2], k[3], k[4]};
        const double got[4] = {env.u.y1, env.u.y2, env.u.y3, env.u.y4};
        double scale = 1e-3;
        for (double r : ref) scale = std::fmax(scale, std::fabs(r));
        for (int j = 0; j < 4; j++) worst = std::fmax(worst, std::fabs(got[j] - ref[j]) / scale);
        mism += term != (int)k[6];
        mism += env.getReward() != k[5];
        mism += env.step != (int)k[7];
        mism += (bool)term != env.is_over();                          // :115 and :100-101 agree by construction
        env.u = Vec4(ref[0], ref[1], ref[2], ref[3]);                 // stay on the reference's trajectory
        if (k[6] != 0) {
          terms++;
          env.reset();
          slot = (slot + 1) % R;
          const double* s = &init[((size_t)e * R + slot) * 4];
          env.u = Vec4(s[0], s[1], s[2], s[3]);
        }
      }
    }
    const bool ok = worst < 1e-9 && mism == 0 && terms > 0;
    std::printf("{\"ok\": %s, \"err_state\": %.3e, \"mismatches\": %d, \"terminations\": %d}\n", ok ? "true" : "false",
                worst, mism, terms);
    return ok ? 0 : 1;
  } catch (const std::exception& e) {
    std::printf("{\"ok\": false, \"error\": \"%s\"}\n", e.what());
    return 3;
  }
}

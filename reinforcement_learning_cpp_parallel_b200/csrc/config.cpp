// config.cpp -- reads the reference's 6-line `config` file verbatim.
//
// Format (config:1-6, parsed by read_parameters, src/main.cpp:356-382): one "<int> = <comment>" line
// per variable, POSITIONAL: control_vars, obs_vars, Nepisodes, N_timestep, Max_timestep, updateTimestep.
// The reference takes std::stoi of each whole line (leading integer, rest ignored) and aborts with an
// uncaught exception on a line that does not start with an integer; here that is B2_ERR_IO.
// Hyper-parameters are NOT in the file: the reference hard-codes them (src/main.cpp:196-203), so they
// live in b2_config with those defaults.
#include <errno.h>
#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <fstream>
#include <string>

#include "engine.h"

extern "C" int b2_config_default(b2_config* c) {
  if (!c) return b2::set_error(B2_ERR_INVALID, "null config");
  memset(c, 0, sizeof *c);
  // config:1-6
  c->control_vars = 1;
  c->obs_vars = 6;
  c->n_episodes = 10000;
  c->n_timestep = 100;
  c->max_timestep = 1000000;
  c->update_timestep = 2000;
  // src/main.cpp:196-203
  c->action_std = 0.5;
  c->k_epochs = 80;
  c->eps_clip = 0.2;
  c->gamma = 0.99;
  c->lr = 0.0003;
  c->beta1 = 0.9;
  c->beta2 = 0.999;
  // engine knobs
  c->hidden1 = 64;  // include/ppo_nn.h:85
  c->hidden2 = 32;  // include/ppo_nn.h:91
  c->n_envs = 1;
  c->horizon = 2000;
  c->n_minibatches = 1;
  c->precision = B2_PREC_F64;
  c->seed = 0x5EED;
  c->env_offset = 0;
  c->device = 0;
  c->gae_lambda = 1.0;  // plain discounted returns (src/ppo_nn.cpp:231)
  return B2_OK;
}

extern "C" int b2_config_read(const char* path, b2_config* c) {
  if (!path || !c) return b2::set_error(B2_ERR_INVALID, "null argument");
  std::ifstream f(path);
  if (!f) return b2::set_error(B2_ERR_IO, "config file not found: %s", path);  // src/main.cpp:379-381
  int numbers[6];
  int n = 0, lineno = 0;
  std::string line;
  while (n < 6 && std::getline(f, line)) {
    lineno++;
    const char* s = line.c_str();
    char* end = nullptr;
    errno = 0;
    long v = strtol(s, &end, 10);  // std::stoi: skip leading blanks, optional sign, digits
    if (end == s) return b2::set_error(B2_ERR_IO, "%s:%d: line does not start with an integer", path, lineno);
    if (errno == ERANGE || v > INT_MAX || v < INT_MIN) return b2::set_error(B2_ERR_IO, "%s:%d: integer out of range", path, lineno);
    numbers[n++] = (int)v;
  }
  if (n < 6) return b2::set_error(B2_ERR_IO, "%s: expected 6 lines, found %d", path, n);
  c->control_vars = numbers[0];     // src/main.cpp:372
  c->obs_vars = numbers[1];         // :373
  c->n_episodes = numbers[2];       // :374
  c->n_timestep = numbers[3];       // :375
  c->max_timestep = numbers[4];     // :376
  c->update_timestep = numbers[5];  // :377
  return B2_OK;
}

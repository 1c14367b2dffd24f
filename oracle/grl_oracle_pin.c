/* grl_oracle_pin.c -- pins of the oracle against reference-held vectors that need objects OFF the hot path.
 * TEST INFRASTRUCTURE ONLY (see grl_oracle.h).  Citations: /root/reference/base.
 *
 * orc_pin_cart_pole_pid: the reference's tests/cart_pole_balancing-pid.yaml (grld -s 1): environment/modeled with
 * dynamics/cart_pole + model/dynamical (RK4, control_step 0.05, 5 integration steps) under task/cart_pole/balancing,
 * driven by agent/fixed with a P-only policy/parameterized/pid.  No learning: every row of
 * tests/template/cart_pole_balancing-pid-0.txt depends only on the start draw, eom_cart_pole and the 5-state RK4,
 * which is exactly what the hot path shares with it (orc_eom / orc_rk4 below are the oracle's own). */
#define _DEFAULT_SOURCE
#define _XOPEN_SOURCE 600
#include "grl_oracle.h"
#include <math.h>
#include <stdlib.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* cart_pole.cpp:316-319 CartPoleBalancingTask::failed */
static int balancing_failed(const double *s)
{
  return fabs(s[0]) > 2.4 || fabs(s[1]) > 12 * M_PI / 180;
}

/* PIDPolicy::act(time, in, out) (policies/pid.cpp:134-175) with only P gains, setpoint 0, one output:
 * u = sum_ii p[ii] * (setpoint - in[ii]) in order; clamp to [action_min, action_max] (= -15, 15, cart_pole.cpp:255-256) */
static double pid_act(const double *p, const double *obs)
{
  double u = 0;
  for (int ii = 0; ii < 4; ++ii)
  {
    double err = 0 - obs[ii];
    u += p[ii] * err;
  }
  return fmin(15., fmax(u, -15.));
}

/* `trials` test episodes of online_learning.cpp:154-222 (test_interval 0: every trial is a test trial of the same
 * fixed agent); rewards[t] = the episode's total reward, steps[t] = its length.  RNG as grld -s 1: srand48(seed)
 * (deployer.cpp:70-74); the only generator on this path is the thread's RandGen, created on its first use by
 * CartPoleBalancingTask::start and seeded srand48_r(lrand48()) (utils.h:84-93,145-187). */
void orc_pin_cart_pole_pid(long seed, const double *p_gains, double timeout, int trials, double *rewards, int32_t *steps)
{
  struct drand48_data randgen;
  srand48(seed);
  srand48_r(lrand48(), &randgen);
  for (int t = 0; t < trials; ++t)
  {
    double state[ORC_MAX_STATE] = {0}, next[ORC_MAX_STATE], r, total = 0;
    int terminal = 0, n = 0;
    /* CartPoleBalancingTask::start (cart_pole.cpp:264-273) */
    drand48_r(&randgen, &r);
    state[1] = (r * 0.1) - 0.05;
    double action = pid_act(p_gains, state);            /* FixedAgent::start: observe copies state[0..3] */
    do
    {
      /* ModeledEnvironment::step (modeled.cpp:160-213): actuate is the identity (environment.h:94) */
      orc_rk4(ORC_DYN_CART_POLE, 0.05, 5, state, action, next);
      /* observe (cart_pole.cpp:275-296) */
      if (balancing_failed(next)) terminal = 2;
      else if (next[4] > timeout) terminal = 1;
      else terminal = 0;
      /* evaluate (cart_pole.cpp:298-306) */
      double reward = balancing_failed(next) ? 0 : 1 - (fabs(state[0]) + fabs(state[1])) / (2.4 + 12 * M_PI / 180);
      total += reward;
      ++n;
      for (int i = 0; i < 5; ++i) state[i] = next[i];
      if (terminal != 2) action = pid_act(p_gains, state);
    } while (!terminal);
    rewards[t] = total;
    if (steps) steps[t] = n;
  }
}

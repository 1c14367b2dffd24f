/* grl_oracle_env.c -- oracle primitives: tile hash, projector, dynamics, tasks.
 * TEST INFRASTRUCTURE ONLY (see grl_oracle.h).  Citations: /root/reference/base. */
#include "grl_oracle.h"
#include <math.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* ---- include/grl/projectors/tile_coding.h:78-114 createHashSum ---------- */
uint32_t orc_hash(const int32_t *ints, uint32_t n)
{
  const uint32_t m = 0x5bd1e995u;
  const int r = 24;
  uint32_t h = 449u ^ n;   /* seed 449 (tile_coding.h:118), xor COUNT of ints (:94) */
  for (uint32_t i = 0; i < n; ++i)
  {
    uint32_t k = (uint32_t)ints[i];
    k *= m;
    k ^= k >> r;
    k *= m;
    h *= m;
    h ^= k;
  }
  h ^= h >> 13;
  h *= m;
  h ^= h >> 15;
  return h;
}

/* Appleby's published MurmurHash2 over bytes (len multiple of 4 here, so no
 * tail block).  h0 = seed ^ len.  orc_hash(ints,n) == this with
 * seed' = 449 ^ n ^ (4n): an independent cross-check of the restatement. */
uint32_t orc_murmur2_bytes(const void *key, int len, uint32_t seed)
{
  const uint32_t m = 0x5bd1e995u;
  const unsigned char *data = (const unsigned char *)key;
  uint32_t h = seed ^ (uint32_t)len;
  while (len >= 4)
  {
    uint32_t k = (uint32_t)data[0] | ((uint32_t)data[1] << 8) |
                 ((uint32_t)data[2] << 16) | ((uint32_t)data[3] << 24);
    k *= m; k ^= k >> 24; k *= m;
    h *= m; h ^= k;
    data += 4; len -= 4;
  }
  h ^= h >> 13; h *= m; h ^= h >> 15;
  return h;
}

/* include/grl/utils.h:70-78 */
static int32_t safe_mod_i(int32_t X, int32_t Y)
{
  int32_t result = X % Y;
  return result < 0 ? result + Y : result;
}

/* src/projectors/tile_coding.cpp:66-79 (configure) */
void orc_scaling(const double *resolution, const double *wrapping, int dims, int tilings,
                 double *scaling, int32_t *iwrap)
{
  for (int ii = 0; ii < dims; ++ii)
  {
    scaling[ii] = tilings / resolution[ii];
    double w = wrapping[ii] * scaling[ii];
    iwrap[ii] = (int32_t)round(w);   /* reference throws if |w-round(w)| > 0.001 */
  }
}

/* src/projectors/tile_coding.cpp:103-149 (_project), safe = 0 */
void orc_project(const double *resolution, const double *wrapping, int dims, int tilings,
                 int32_t memory, const double *in, uint32_t *hash, uint32_t *index)
{
  double scaling[ORC_MAX_DIMS];
  int32_t iwrap[ORC_MAX_DIMS];
  int32_t qstate[ORC_MAX_DIMS], base[ORC_MAX_DIMS], coordinates[ORC_MAX_DIMS + 1];
  int i, j;

  orc_scaling(resolution, wrapping, dims, tilings, scaling, iwrap);

  for (i = 0; i < dims; ++i)
  {
    qstate[i] = (int32_t)floor(in[i] * scaling[i]);
    base[i] = 0;
  }
  for (j = 0; j < tilings; ++j)
  {
    for (i = 0; i < dims; ++i)
    {
      coordinates[i] = qstate[i] - safe_mod_i(qstate[i] - base[i], tilings);
      if (iwrap[i] != 0)
        coordinates[i] = safe_mod_i(coordinates[i], iwrap[i]);
      base[i] += 1 + (2 * i);
    }
    coordinates[i] = j;
    uint32_t h = orc_hash(coordinates, (uint32_t)(dims + 1));
    if (hash) hash[j] = h;
    index[j] = h % (uint32_t)memory;   /* tile_coding.h:118 */
  }
}

void orc_project_hashes(const double *resolution, const double *wrapping, int dims, int tilings,
                        const double *in, uint32_t *hash)
{
  uint32_t index[ORC_MAX_TILINGS];
  orc_project(resolution, wrapping, dims, tilings, 1, in, hash, index);
}

int orc_state_dims(int dynamics)
{
  return dynamics == ORC_DYN_PENDULUM ? 3 : dynamics == ORC_DYN_CART_POLE ? 5 : 7;
}

/* ---- equations of motion ------------------------------------------------ */

/* src/environments/pendulum.cpp:40-49,55-68 */
static void eom_pendulum(const double *state, double u, double *xd)
{
  const double J_ = 0.000191, m_ = 0.055, g_ = 9.81, l_ = 0.042, b_ = 0.000003,
               K_ = 0.0536, R_ = 9.5;
  double a = state[0];
  double ad = state[1];
  double add = (1 / J_) * (m_ * g_ * l_ * sin(a) - b_ * ad - (K_ * K_ / R_) * ad + (K_ / R_) * u);
  xd[0] = ad;
  xd[1] = add;
  xd[2] = 1;
}

/* src/environments/cart_pole.cpp:42-52,58-108 (end_stop_ = 1) */
static void eom_cart_pole(const double *state, double u, double *xd)
{
  const double g_ = 9.8, mass_cart_ = 1.0, mass_pole_ = 0.1, length_ = 0.5;
  const double total_mass_ = mass_cart_ + mass_pole_;
  const double pole_mass_length_ = mass_pole_ * length_;
  double acc, thetaacc, costheta, sintheta, temp;
  /* cart_pole.cpp:66: `theta = state[end_stop_], dtheta = state[3-2*end_stop_]`.  With end_stop_ = 1 (the default and
   * the only form the 5-state tasks accept) BOTH are state[1]: the reference feeds the pole ANGLE where the angular
   * velocity belongs.  Reproduced as is (found by compiling the reference's cart_pole.cpp, tests/test_oracle_vs_ref.py). */
  const int end_stop_ = 1;
  double theta = state[end_stop_], dtheta = state[3 - 2 * end_stop_];

  costheta = cos(theta);
  sintheta = sin(theta);
  temp = (u + pole_mass_length_ * dtheta * dtheta * sintheta) / total_mass_;
  thetaacc = (g_ * sintheta - costheta * temp) /
             (length_ * ((4. / 3.) - mass_pole_ * costheta * costheta / total_mass_));
  acc = temp - pole_mass_length_ * thetaacc * costheta / total_mass_;

  xd[0] = state[2];
  xd[1] = state[3];
  xd[2] = acc;
  xd[3] = thetaacc;
  xd[4] = 1;
  if (state[0] > 2.4 && state[2] > 0)
  {
    xd[0] = 0;
    if (acc > 0) xd[2] = 0;
  }
  else if (state[0] < -2.4 && state[2] < 0)
  {
    xd[0] = 0;
    if (acc < 0) xd[2] = 0;
  }
}

/* Eigen 3.3/3.4 Inverse (LU/InverseImpl.h, size-3 specialisation):
 * cofactor_3x3<i,j> = m(i1,j1)*m(i2,j2) - m(i1,j2)*m(i2,j1), i1=(i+1)%3 ...;
 * det = sum(cofactors_col0 .* m.col(0)); invdet = 1/det;
 * inv(r,c) = cofactor<c,r> * invdet.  Third-party (Eigen3, version unpinned
 * by the reference; CI installs libeigen3-dev 3.4.0): restated from its
 * published algorithm, sums of three taken as (a0+a1)+a2.  PARITY UNPINNED. */
static double cof3(const double m[3][3], int i, int j)
{
  int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
  return m[i1][j1] * m[i2][j2] - m[i1][j2] * m[i2][j1];
}

static void inverse3(const double m[3][3], double inv[3][3])
{
  double c0 = cof3(m, 0, 0), c1 = cof3(m, 1, 0), c2 = cof3(m, 2, 0);
  double det = (c0 * m[0][0] + c1 * m[1][0]) + c2 * m[2][0];
  double invdet = 1 / det;
  inv[0][0] = c0 * invdet; inv[0][1] = c1 * invdet; inv[0][2] = c2 * invdet;
  for (int r = 1; r < 3; ++r)
    for (int c = 0; c < 3; ++c)
      inv[r][c] = cof3(m, c, r) * invdet;
}

static void matvec3(const double m[3][3], const double v[3], double out[3])
{
  for (int r = 0; r < 3; ++r)
    out[r] = (m[r][0] * v[0] + m[r][1] * v[1]) + m[r][2] * v[2];
}

/* src/environments/cart_double_pole.cpp:46-71,78-127 */
static void eom_cart_double_pole(const double *state, double f, double *xd)
{
  const double m_ = 0.5, m1_ = 0.5, m2_ = 0.5, l1_ = 0.3, l2_ = 0.3, g_ = 9.82;
  const double L1_ = 2 * l1_, L2_ = 2 * l2_;
  const double J1 = m1_ * L1_ * L1_ / 3, J2 = m2_ * L2_ * L2_ / 3;
  const double h1_ = m_ + m1_ + m2_;
  const double h2_ = m1_ * l1_ + m2_ * L1_;
  const double h3_ = m2_ * l2_;
  const double h4_ = m1_ * l1_ * l1_ + m2_ * L1_ * L1_ + J1;
  const double h5_ = m2_ * l2_ * L1_;
  const double h6_ = m2_ * l2_ * l2_ + J2;
  const double h7_ = m1_ * l1_ * g_ + m2_ * L1_ * g_;
  const double h8_ = m2_ * l2_ * g_;

  double M[3][3], C[3][3], Mi[3][3], g[3], u[3], Cqd[3], rhs[3], qdd[3];
  double qd[3] = { state[3], state[4], state[5] };
  double c1 = cos(state[1]), c2 = cos(state[2]),
         s1 = sin(state[1]), s2 = sin(state[2]),
         c12 = cos(state[1] - state[2]),
         s12 = sin(state[1] - state[2]);

  M[0][0] = h1_;      M[0][1] = h2_ * c1;  M[0][2] = h3_ * c2;
  M[1][0] = h2_ * c1; M[1][1] = h4_;       M[1][2] = h5_ * c12;
  M[2][0] = h3_ * c2; M[2][1] = h5_ * c12; M[2][2] = h6_;
  C[0][0] = 0; C[0][1] = -h2_ * qd[1] * s1;  C[0][2] = -h3_ * qd[2] * s2;
  C[1][0] = 0; C[1][1] = 0;                  C[1][2] = h5_ * qd[2] * s12;
  C[2][0] = 0; C[2][1] = -h5_ * qd[1] * s12; C[2][2] = 0;
  g[0] = 0; g[1] = -h7_ * s1; g[2] = -h8_ * s2;
  u[0] = f; u[1] = 0; u[2] = 0;

  matvec3(C, qd, Cqd);
  for (int r = 0; r < 3; ++r)
    rhs[r] = (u[r] - Cqd[r]) - g[r];
  inverse3(M, Mi);
  matvec3(Mi, rhs, qdd);

  xd[0] = qd[0]; xd[1] = qd[1]; xd[2] = qd[2];
  xd[3] = qdd[0]; xd[4] = qdd[1]; xd[5] = qdd[2];
  xd[6] = 1;
  if (state[0] > 2.4 && state[3] > 0)
  {
    xd[0] = 0;
    if (xd[3] > 0) xd[3] = 0;
  }
  else if (state[0] < -2.4 && state[3] < 0)
  {
    xd[0] = 0;
    if (xd[3] < 0) xd[3] = 0;
  }
}

void orc_eom(int dynamics, const double *x, double u, double *xd)
{
  if (dynamics == ORC_DYN_PENDULUM) eom_pendulum(x, u, xd);
  else if (dynamics == ORC_DYN_CART_POLE) eom_cart_pole(x, u, xd);
  else eom_cart_double_pole(x, u, xd);
}

/* src/environments/modeled.cpp:254-276 DynamicalModel::step */
void orc_rk4(int dynamics, double tau, int steps, const double *x, double u, double *next)
{
  int n = orc_state_dims(dynamics);
  double xd[ORC_MAX_STATE], k1[ORC_MAX_STATE], k2[ORC_MAX_STATE], k3[ORC_MAX_STATE],
         k4[ORC_MAX_STATE], tmp[ORC_MAX_STATE];
  double h = tau / steps;
  int i;

  for (i = 0; i < n; ++i) next[i] = x[i];
  for (int ii = 0; ii < steps; ++ii)
  {
    orc_eom(dynamics, next, u, xd);
    for (i = 0; i < n; ++i) k1[i] = h * xd[i];
    for (i = 0; i < n; ++i) tmp[i] = next[i] + k1[i] / 2;
    orc_eom(dynamics, tmp, u, xd);
    for (i = 0; i < n; ++i) k2[i] = h * xd[i];
    for (i = 0; i < n; ++i) tmp[i] = next[i] + k2[i] / 2;
    orc_eom(dynamics, tmp, u, xd);
    for (i = 0; i < n; ++i) k3[i] = h * xd[i];
    for (i = 0; i < n; ++i) tmp[i] = next[i] + k3[i];
    orc_eom(dynamics, tmp, u, xd);
    for (i = 0; i < n; ++i) k4[i] = h * xd[i];
    /* Eigen array expression: ((k1 + 2*k2) + 2*k3) + k4, true division by 6 */
    for (i = 0; i < n; ++i)
      next[i] = next[i] + (((k1[i] + 2 * k2[i]) + 2 * k3[i]) + k4[i]) / 6;
  }
}

/* ---- tasks -------------------------------------------------------------- */

static int cp_failed(const double *s) { return fabs(s[0]) > 2.4; }
static int cdp_failed(const double *s)
{
  return fabs(s[0]) > 2.4 || fabs(s[1]) > 0.7 || fabs(s[2]) > 0.7;
}

/* pendulum.cpp:111-129, cart_pole.cpp:162-186, cart_double_pole.cpp:264-287 */
void orc_observe(const orc_config *c, const double *x, double *obs, int32_t *terminal)
{
  if (c->dynamics == ORC_DYN_PENDULUM)
  {
    double a = fmod(x[0] + M_PI, 2 * M_PI);
    if (a < 0) a += 2 * M_PI;
    obs[0] = a;
    obs[1] = x[1];
    *terminal = x[2] > c->timeout ? 1 : 0;
  }
  else if (c->dynamics == ORC_DYN_CART_POLE)
  {
    double a = fmod(x[1] + M_PI, 2 * M_PI);
    if (a < 0) a += 2 * M_PI;
    obs[0] = x[0]; obs[1] = a; obs[2] = x[2]; obs[3] = x[3];
    if (c->end_stop_penalty && cp_failed(x)) *terminal = 2;   /* cart_pole.cpp:177; default 1 (cart_pole.h:75) */
    else if (x[4] > c->timeout) *terminal = 1;
    else *terminal = 0;
  }
  else
  {
    for (int i = 0; i < 6; ++i) obs[i] = x[i];
    if (cdp_failed(x)) *terminal = 2;
    else if (x[6] > c->timeout) *terminal = 1;
    else *terminal = 0;
  }
}

/* Task::start with the uniform draw r (always consumed): pendulum.cpp:97-103, cart_pole.cpp:152-160,
 * cart_double_pole.cpp:250-262 */
void orc_task_start(const orc_config *c, int test, double r, double *state)
{
  const int S = orc_state_dims(c->dynamics);
  for (int i = 0; i < S; ++i) state[i] = 0;
  if (c->dynamics == ORC_DYN_PENDULUM)
    state[0] = M_PI + c->randomization * (test == 0) * r * 2 * M_PI;
  else if (c->dynamics == ORC_DYN_CART_POLE)
    state[1] = M_PI + (int)c->randomization * ((r * 0.1) - 0.05);   /* int randomization_ (cart_pole.h:73) */
  else
    state[1] = (r * 0.01) - 0.005;
}

/* cart_pole.cpp:229-235 */
static double cp_potential(const double *s)
{
  double a = fmod(fabs(s[1]), 2 * M_PI);
  if (a > M_PI) a -= 2 * M_PI;
  return -2 * pow(s[0], 2) - 0.1 * pow(s[2], 2) - pow(a, 2) - 0.1 * pow(s[3], 2);
}

/* pendulum.cpp:131-145, cart_pole.cpp:188-197, cart_double_pole.cpp:289-298 */
double orc_evaluate(const orc_config *c, const double *state, double action, const double *next)
{
  double reward;
  if (c->dynamics == ORC_DYN_PENDULUM)
  {
    double a = fmod(fabs(next[0]), 2 * M_PI);
    if (a > M_PI) a -= 2 * M_PI;
    reward = -5 * pow(a, 2) - 0.1 * pow(next[1], 2) - 1 * pow(action, 2);
    if ((next[2] - state[2]) != 1)
      reward *= (next[2] - state[2]) / 0.03;
  }
  else if (c->dynamics == ORC_DYN_CART_POLE)
  {
    /* shaping_ = 0, action_penalty_ = 0:
     * potential(next) - 0*pow(action/15,2)*2 - end_stop_penalty_*failed(next)*10000 (int product, cart_pole.cpp:196) */
    reward = cp_potential(next) - 0 * pow(action / 15, 2) * 2 - c->end_stop_penalty * cp_failed(next) * 10000;
  }
  else
  {
    if (cdp_failed(next)) reward = 0;
    else reward = 6.2 - fabs(state[0]) - fabs(state[1]) - fabs(state[2]);
  }
  return reward;
}

/* modeled.cpp:160-213 (window_ = 1, no delta_, discrete_time_ = 1) */
void orc_env_step(const orc_config *c, const double *x, double action,
                  double *next, double *obs, double *reward, int32_t *terminal)
{
  double actuation = action;
  if (c->dynamics == ORC_DYN_PENDULUM)
    actuation = fmin(fmax(action, -3), 3);       /* pendulum.cpp:105-109 */
  orc_rk4(c->dynamics, c->control_step, c->integration_steps, x, actuation, next);
  orc_observe(c, next, obs, terminal);
  *reward = orc_evaluate(c, x, action, next);
}

/* ---- Philox4x32-10 ------------------------------------------------------ */
void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                uint32_t k0, uint32_t k1, uint32_t out[4])
{
  for (int r = 0; r < 10; ++r)
  {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

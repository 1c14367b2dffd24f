// grlcu_device.cuh -- device functions of the batched GRL learning loop (sm_100a).
//
// All floating point is FP64 and this translation unit is compiled with -fmad=false:
// the reference binary has no FMA contraction (CMakeLists.txt:19 has no -march/-mfma),
// so every product and sum below rounds separately, in the reference's evaluation order.
// Citations: /root/reference/base.
#pragma once
#include <stdint.h>
#include <math.h>

namespace grlcu {

constexpr uint32_t kInvalidIndex = 0xFFFFFFFFu;   // (size_t)-1 of projection.h:73-76
constexpr int kMaxDims = 8;

// RNG stream ids (Philox counter word 2)
enum { STREAM_ACT = 0, STREAM_START = 1, STREAM_TABLE = 2, STREAM_TEST_ACT = 3,
       STREAM_ACTOR_TABLE = 4, STREAM_NOISE = 5 };

// Gather of one hashed table word.  A plain ld.global of a random 8-byte word makes L2 fetch the whole
// 128-byte line from DRAM (127 B per access measured); the .L2::64B prefetch-size qualifier halves that
// (63.8 B per access, profiles/r01_randbench3_ncu.txt) -- and DRAM bytes are what this path runs out of.
// volatile + "memory": the gathers keep their program order with respect to the table stores of the update.
__device__ __forceinline__ double ld_tbl(const double *p)
{
  double v;
  asm volatile("ld.global.L2::64B.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
// two adjacent table words with one 16-byte request (the actor-critic's interleaved {critic, actor} pairs)
__device__ __forceinline__ double2 ld_tbl2(const double *p)
{
  double2 v;
  asm volatile("ld.global.L2::64B.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p) : "memory");
  return v;
}

// ---- Philox4x32-10 (Salmon et al., SC'11) -------------------------------------------------
struct Philox4 { uint32_t w[4]; };

__device__ __forceinline__ Philox4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                 uint32_t k0, uint32_t k1)
{
#pragma unroll
  for (int r = 0; r < 10; ++r)
  {
    uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  Philox4 o; o.w[0] = c0; o.w[1] = c1; o.w[2] = c2; o.w[3] = c3;
  return o;
}

__device__ __forceinline__ Philox4 philox_for(uint64_t seed, uint64_t agent_id, int stream, uint64_t ctr)
{
  return philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), (uint32_t)stream, (uint32_t)(seed >> 32),
                       (uint32_t)agent_id, (uint32_t)seed ^ (uint32_t)(agent_id >> 32));
}

__device__ __forceinline__ double u53(uint32_t hi, uint32_t lo)
{
  uint64_t x = ((uint64_t)hi << 32) | lo;
  return (double)(x >> 11) * (1.0 / 9007199254740992.0);
}

// ---- tile coding: tile_coding.h:78-118 (MurmurHash2 over ints, h0 = 449 ^ count) ----------
template <int N>
__device__ __forceinline__ uint32_t tile_hash(const int32_t (&ints)[N])
{
  const uint32_t m = 0x5bd1e995u;
  uint32_t h = 449u ^ (uint32_t)N;
#pragma unroll
  for (int i = 0; i < N; ++i)
  {
    uint32_t k = (uint32_t)ints[i];
    k *= m; k ^= k >> 24; k *= m;
    h *= m; h ^= k;
  }
  h ^= h >> 13; h *= m; h ^= h >> 15;
  return h;
}

// utils.h:70-78
__device__ __forceinline__ int32_t safe_mod(int32_t X, int32_t Y)
{
  int32_t r = X % Y;
  return r < 0 ? r + Y : r;
}

// Projector parameters precomputed on the host exactly like TileCodingProjector::configure
// (tile_coding.cpp:66-79): scaling = tilings/resolution, wrap = round(wrapping*scaling).
struct Projector
{
  double  scaling[kMaxDims];
  int32_t wrap[kMaxDims];
  int32_t memory;
};

// tile_coding.cpp:103-149 (_project, safe = 0).  DIMS = obs dims + 1 (action last).
// T (tilings) is a power of two in every shipped config; safe_mod(x, T) handles the general case.
template <int DIMS, int T>
__device__ __forceinline__ void quantize(const Projector &P, const double (&in)[DIMS], int32_t (&q)[DIMS])
{
#pragma unroll
  for (int i = 0; i < DIMS; ++i)
    q[i] = (int32_t)floor(in[i] * P.scaling[i]);
}

template <int DIMS, int T>
__device__ __forceinline__ uint32_t tile_hash_of(const Projector &P, const int32_t (&q)[DIMS], int j)
{
  int32_t coords[DIMS + 1];
#pragma unroll
  for (int i = 0; i < DIMS; ++i)
  {
    int32_t base = j * (1 + 2 * i);             // base[i] after j increments of 1+2i
    int32_t c = q[i] - safe_mod(q[i] - base, T);
    if (P.wrap[i] != 0) c = safe_mod(c, P.wrap[i]);
    coords[i] = c;
  }
  coords[DIMS] = j;
  return tile_hash<DIMS + 1>(coords);
}

// ---- dynamics ------------------------------------------------------------------------------
template <int DYN> struct Dyn;

// pendulum.cpp:40-49,55-68
template <> struct Dyn<0>
{
  static constexpr int S = 3, OBS = 2;
  __device__ __forceinline__ static void eom(const double (&x)[3], double u, double (&xd)[3])
  {
    const double J = 0.000191, m = 0.055, g = 9.81, l = 0.042, b = 0.000003, K = 0.0536, R = 9.5;
    double a = x[0], ad = x[1];
    double add = (1 / J) * (m * g * l * sin(a) - b * ad - (K * K / R) * ad + (K / R) * u);
    xd[0] = ad; xd[1] = add; xd[2] = 1;
  }
  // pendulum.cpp:105-109
  __device__ __forceinline__ static double actuate(double action) { return fmin(fmax(action, -3.), 3.); }
};

// cart_pole.cpp:42-52,58-108 (end_stop_ = 1)
template <> struct Dyn<1>
{
  static constexpr int S = 5, OBS = 4;
  __device__ __forceinline__ static void eom(const double (&x)[5], double u, double (&xd)[5])
  {
    const double g = 9.8, mass_cart = 1.0, mass_pole = 0.1, length = 0.5;
    const double total_mass = mass_cart + mass_pole;
    const double pole_mass_length = mass_pole * length;
    // cart_pole.cpp:66: theta = state[end_stop_], dtheta = state[3-2*end_stop_]; with end_stop_ = 1 both are state[1]
    // (the reference uses the pole angle where the angular velocity belongs; reproduced as is)
    double theta = x[1], dtheta = x[3 - 2 * 1];
    double costheta, sintheta;
    sincos(theta, &sintheta, &costheta);        // one range reduction for both (the dynamics' bar is 1e-9 against glibc)
    double temp = (u + pole_mass_length * dtheta * dtheta * sintheta) / total_mass;
    double thetaacc = (g * sintheta - costheta * temp) /
                      (length * ((4. / 3.) - mass_pole * costheta * costheta / total_mass));
    double acc = temp - pole_mass_length * thetaacc * costheta / total_mass;
    xd[0] = x[2]; xd[1] = x[3]; xd[2] = acc; xd[3] = thetaacc; xd[4] = 1;
    if (x[0] > 2.4 && x[2] > 0)
    {
      xd[0] = 0;
      if (acc > 0) xd[2] = 0;
    }
    else if (x[0] < -2.4 && x[2] < 0)
    {
      xd[0] = 0;
      if (acc < 0) xd[2] = 0;
    }
  }
  __device__ __forceinline__ static double actuate(double action) { return action; }  // environment.h:94
};

// cart_double_pole.cpp:46-71,78-127; Eigen Matrix3d::inverse() = cofactors * (1/det)
template <> struct Dyn<2>
{
  static constexpr int S = 7, OBS = 6;
  __device__ __forceinline__ static double cof(const double (&m)[3][3], int i, int j)
  {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return m[i1][j1] * m[i2][j2] - m[i1][j2] * m[i2][j1];
  }
  __device__ __forceinline__ static void eom(const double (&x)[7], double f, double (&xd)[7])
  {
    const double m_ = 0.5, m1 = 0.5, m2 = 0.5, l1 = 0.3, l2 = 0.3, g_ = 9.82;
    const double L1 = 2 * l1, L2 = 2 * l2;
    const double J1 = m1 * L1 * L1 / 3, J2 = m2 * L2 * L2 / 3;
    const double h1 = m_ + m1 + m2, h2 = m1 * l1 + m2 * L1, h3 = m2 * l2,
                 h4 = m1 * l1 * l1 + m2 * L1 * L1 + J1, h5 = m2 * l2 * L1, h6 = m2 * l2 * l2 + J2,
                 h7 = m1 * l1 * g_ + m2 * L1 * g_, h8 = m2 * l2 * g_;
    double qd[3] = { x[3], x[4], x[5] };
    double c1, c2, s1, s2, c12, s12;
    sincos(x[1], &s1, &c1); sincos(x[2], &s2, &c2); sincos(x[1] - x[2], &s12, &c12);
    double M[3][3] = { { h1, h2 * c1, h3 * c2 }, { h2 * c1, h4, h5 * c12 }, { h3 * c2, h5 * c12, h6 } };
    double C[3][3] = { { 0, -h2 * qd[1] * s1, -h3 * qd[2] * s2 }, { 0, 0, h5 * qd[2] * s12 },
                       { 0, -h5 * qd[1] * s12, 0 } };
    double g[3] = { 0, -h7 * s1, -h8 * s2 };
    double u[3] = { f, 0, 0 };
    double rhs[3], Mi[3][3], qdd[3];
#pragma unroll
    for (int r = 0; r < 3; ++r)
    {
      double cq = (C[r][0] * qd[0] + C[r][1] * qd[1]) + C[r][2] * qd[2];
      rhs[r] = (u[r] - cq) - g[r];
    }
    double k0 = cof(M, 0, 0), k1 = cof(M, 1, 0), k2 = cof(M, 2, 0);
    double det = (k0 * M[0][0] + k1 * M[1][0]) + k2 * M[2][0];
    double invdet = 1 / det;
    Mi[0][0] = k0 * invdet; Mi[0][1] = k1 * invdet; Mi[0][2] = k2 * invdet;
#pragma unroll
    for (int r = 1; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c)
        Mi[r][c] = cof(M, c, r) * invdet;
#pragma unroll
    for (int r = 0; r < 3; ++r)
      qdd[r] = (Mi[r][0] * rhs[0] + Mi[r][1] * rhs[1]) + Mi[r][2] * rhs[2];
    xd[0] = qd[0]; xd[1] = qd[1]; xd[2] = qd[2];
    xd[3] = qdd[0]; xd[4] = qdd[1]; xd[5] = qdd[2]; xd[6] = 1;
    if (x[0] > 2.4 && x[3] > 0)
    {
      xd[0] = 0;
      if (xd[3] > 0) xd[3] = 0;
    }
    else if (x[0] < -2.4 && x[3] < 0)
    {
      xd[0] = 0;
      if (xd[3] < 0) xd[3] = 0;
    }
  }
  __device__ __forceinline__ static double actuate(double action) { return action; }
};

// modeled.cpp:254-276 DynamicalModel::step: classical RK4, `steps` sub-steps of h = tau/steps
template <int DYN>
__device__ __forceinline__ void rk4(double tau, int steps, double u, double (&x)[Dyn<DYN>::S])
{
  constexpr int S = Dyn<DYN>::S;
  const double h = tau / steps;
  for (int ii = 0; ii < steps; ++ii)
  {
    // The four stages go through ONE copy of the equations of motion (code size: the fused kernels are
    // instruction-fetch bound).  Same arithmetic as the reference's integrator (modeled.cpp:254-276):
    //   tmp = x + k/2 (stages 1,2: k * 0.5 == k / 2 exactly) or x + k (stage 3);
    //   x += (((k1 + 2 k2) + 2 k3) + k4) / 6, summed in that order.
    double tmp[S], acc[S];
#pragma unroll
    for (int i = 0; i < S; ++i) tmp[i] = x[i];
    // (the pendulum's equations are a handful of instructions: unrolled there)
#pragma unroll (DYN == 0 ? 4 : 1)
    for (int stage = 0; stage < 4; ++stage)
    {
      double xd[S];
      Dyn<DYN>::eom(tmp, u, xd);
      const double cn = stage < 2 ? 0.5 : 1.;     // weight of this k in the next stage's argument
      const double wk = (stage == 1 || stage == 2) ? 2. : 1.;
#pragma unroll
      for (int i = 0; i < S; ++i)
      {
        const double k = h * xd[i];
        acc[i] = stage == 0 ? k : acc[i] + wk * k;
        tmp[i] = x[i] + k * cn;
      }
    }
#pragma unroll
    for (int i = 0; i < S; ++i) x[i] = x[i] + acc[i] / 6;
  }
}

// ---- tasks: observe / evaluate / start --------------------------------------------------------
#define GRLCU_PI 3.14159265358979323846

template <int DYN>
__device__ __forceinline__ void observe(const double (&x)[Dyn<DYN>::S], double timeout,
                                        double (&obs)[Dyn<DYN>::OBS], int &terminal, int end_stop_penalty = 1)
{
  if (DYN == 0)
  {
    // pendulum.cpp:111-129
    double a = fmod(x[0] + GRLCU_PI, 2 * GRLCU_PI);
    if (a < 0) a += 2 * GRLCU_PI;
    obs[0] = a; obs[1] = x[1];
    terminal = x[2] > timeout ? 1 : 0;
  }
  else if (DYN == 1)
  {
    // cart_pole.cpp:162-186
    double a = fmod(x[1] + GRLCU_PI, 2 * GRLCU_PI);
    if (a < 0) a += 2 * GRLCU_PI;
    obs[0] = x[0]; obs[1] = a; obs[2] = x[2]; obs[3] = x[3];
    if (end_stop_penalty && fabs(x[0]) > 2.4) terminal = 2;
    else if (x[4] > timeout) terminal = 1;
    else terminal = 0;
  }
  else
  {
    // cart_double_pole.cpp:264-287
#pragma unroll
    for (int i = 0; i < 6; ++i) obs[i] = x[i];
    if (fabs(x[0]) > 2.4 || fabs(x[1]) > 0.7 || fabs(x[2]) > 0.7) terminal = 2;
    else if (x[6] > timeout) terminal = 1;
    else terminal = 0;
  }
}

template <int DYN>
__device__ __forceinline__ double evaluate(const double (&state)[Dyn<DYN>::S], double action,
                                           const double (&next)[Dyn<DYN>::S], int end_stop_penalty = 1)
{
  if (DYN == 0)
  {
    // pendulum.cpp:131-145; pow(x, 2) == x*x
    double a = fmod(fabs(next[0]), 2 * GRLCU_PI);
    if (a > GRLCU_PI) a -= 2 * GRLCU_PI;
    double reward = -5 * (a * a) - 0.1 * (next[1] * next[1]) - 1 * (action * action);
    double dt = next[2] - state[2];
    if (dt != 1) reward *= dt / 0.03;
    return reward;
  }
  else if (DYN == 1)
  {
    // cart_pole.cpp:188-197,229-235 (shaping 0, action_penalty 0; end_stop_penalty 1 by default, 0 in
    // cfg/cart_pole/ac_tc.yaml: the cart then rides the end stop of the dynamics without ending the episode)
    double a = fmod(fabs(next[1]), 2 * GRLCU_PI);
    if (a > GRLCU_PI) a -= 2 * GRLCU_PI;
    double potential = -2 * (next[0] * next[0]) - 0.1 * (next[2] * next[2]) - (a * a) - 0.1 * (next[3] * next[3]);
    double failed = fabs(next[0]) > 2.4 ? 1. : 0.;
    return potential - 0. - end_stop_penalty * failed * 10000;
  }
  else
  {
    // cart_double_pole.cpp:289-298
    if (fabs(next[0]) > 2.4 || fabs(next[1]) > 0.7 || fabs(next[2]) > 0.7) return 0;
    return 6.2 - fabs(state[0]) - fabs(state[1]) - fabs(state[2]);
  }
}

// pendulum.cpp:97-103, cart_pole.cpp:152-160, cart_double_pole.cpp:250-262
template <int DYN>
__device__ __forceinline__ void task_start(double r, double randomization, int test, double (&x)[Dyn<DYN>::S])
{
#pragma unroll
  for (int i = 0; i < Dyn<DYN>::S; ++i) x[i] = 0;
  if (DYN == 0) x[0] = GRLCU_PI + randomization * (test == 0) * r * 2 * GRLCU_PI;
  else if (DYN == 1) x[1] = GRLCU_PI + (int)randomization * ((r * 0.1) - 0.05);
  else x[1] = (r * 0.01) - 0.005;
}

}  // namespace grlcu

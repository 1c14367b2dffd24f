/* oracle/ref_env.cpp -- C harness over the REFERENCE'S OWN environment classes (compiled from /root/reference where
 * they lie, against the stand-in headers of oracle/ref_shim/; see that header for what is and is not the reference's).
 * TEST INFRASTRUCTURE ONLY: built into oracle/_ref/libgrlref_env.so by `make -C oracle ref`; loaded by
 * tests/test_oracle_vs_ref.py only.  dyn: 0 pendulum, 1 cart-pole, 2 cart-double-pole (hot-path tasks: pendulum swing-up,
 * cart-pole swing-up, cart-double-pole balancing). */
#include <grl/environments/pendulum.h>
#include <grl/environments/cart_pole.h>
#include <grl/environments/cart_double_pole.h>
#include <string.h>

using namespace grl;

namespace {
struct Objects
{
  PendulumDynamics pd; CartPoleDynamics cpd; CartDoublePoleDynamics cdpd;
  PendulumSwingupTask pt; CartPoleSwingupTask cpt; CartDoublePoleBalancingTask cdpt;
  int configured_esp; double configured_rand; bool ok;
  Objects() : configured_esp(-1), configured_rand(-1), ok(false) { }
};
Objects &objs() { static Objects o; return o; }

void configure(double timeout, double randomization, int end_stop_penalty)
{
  Objects &o = objs();
  Configuration c;
  /* the YAML keys each configure() reads (pendulum.cpp:70-95, cart_pole.cpp:110-148, cart_double_pole.cpp:233-247) */
  c.put("timeout", timeout); c.put("randomization", randomization);
  c.put("end_stop", 1);                                          /* dynamics/cart_pole default (cart_pole.h:52) */
  c.put("shaping", 0); c.put("gamma", 1); c.put("action_penalty", 0); c.put("end_stop_penalty", end_stop_penalty);
  c.put("wrap_angle", 1);
  o.pd.configure(c); o.cpd.configure(c); o.cdpd.configure(c);
  o.pt.configure(c); o.cpt.configure(c); o.cdpt.configure(c);
  o.ok = true;
}
int sdim(int dyn) { return dyn == 0 ? 3 : dyn == 1 ? 5 : 7; }
Vector vec(const double *x, int n) { Vector v(n); for (int i = 0; i < n; ++i) v[i] = x[i]; return v; }
Dynamics &dynamics(int dyn) { Objects &o = objs(); return dyn == 0 ? (Dynamics&)o.pd : dyn == 1 ? (Dynamics&)o.cpd : (Dynamics&)o.cdpd; }
Task &task(int dyn) { Objects &o = objs(); return dyn == 0 ? (Task&)o.pt : dyn == 1 ? (Task&)o.cpt : (Task&)o.cdpt; }
}

extern "C" {

int ref_configure(double timeout, double randomization, int end_stop_penalty)
{
  try { configure(timeout, randomization, end_stop_penalty); } catch (std::exception &e) { fprintf(stderr, "ref_configure: %s\n", e.what()); return 1; }
  return 0;
}

/* Dynamics::eom */
int ref_eom(int dyn, const double *x, double u, double *xd)
{
  try
  {
    Vector out;
    dynamics(dyn).eom(vec(x, sdim(dyn)), VectorConstructor(u), &out);
    for (size_t i = 0; i < out.size(); ++i) xd[i] = out[i];
  }
  catch (std::exception &e) { fprintf(stderr, "ref_eom: %s\n", e.what()); return 1; }
  return 0;
}

/* Task::start with the uniform draw `u` fed in place of RandGen::get() */
int ref_start(int dyn, int test, double u, double *state)
{
  RandGen::fed() = u;
  Vector s;
  task(dyn).start(test, &s);
  for (size_t i = 0; i < s.size(); ++i) state[i] = s[i];
  return (int)s.size();
}

/* Task::actuate */
double ref_actuate(int dyn, const double *x, double action)
{
  Vector a;
  task(dyn).actuate(vec(x, sdim(dyn)), vec(x, sdim(dyn)), Action(VectorConstructor(action)), &a);
  return a[0];
}

/* Task::observe */
int ref_observe(int dyn, const double *x, double *obs, int *terminal, int *absorbing)
{
  Observation o;
  task(dyn).observe(vec(x, sdim(dyn)), &o, terminal);
  for (size_t i = 0; i < o.size(); ++i) obs[i] = o[i];
  if (absorbing) *absorbing = o.absorbing;
  return (int)o.size();
}

/* Task::evaluate */
double ref_evaluate(int dyn, const double *x, double action, const double *next)
{
  double r = 0;
  task(dyn).evaluate(vec(x, sdim(dyn)), Action(VectorConstructor(action)), vec(next, sdim(dyn)), &r);
  return r;
}

}

/* oracle/ref_shim/grl/environment.h -- STAND-IN for the reference's <grl/environment.h> include chain.
 *
 * TEST INFRASTRUCTURE ONLY.  Purpose: compile the reference's OWN first-party environment sources
 *   /root/reference/base/src/environments/{pendulum,cart_pole,cart_double_pole}.cpp
 * (and their own class headers include/grl/environments/*.h) where they lie, WITHOUT Eigen 3 and without the
 * rest of libgrl, into oracle/_ref/libgrlref_env.so (oracle/Makefile, target `ref`), so that the C restatement in
 * oracle/grl_oracle_env.c can be checked against the reference's own arithmetic (tests/test_oracle_vs_ref.py).
 *
 * What is stand-in (written here, NOT the reference's): grl::Vector (a small fixed-capacity array of double with
 * the element access / resize / arithmetic the three sources use), Observation / Action, the Configurable plumbing
 * (Configuration, ConfigurationRequest, CRP, REGISTER_CONFIGURABLE, TYPEINFO: inert), Exception, RandGen (returns
 * the draw the harness feeds), RegulatorTask (inert base: the regulator tasks are off the hot path), and
 * Eigen::Matrix3d / Vector3d with the size-3 inverse restated from Eigen 3.4's InverseImpl.h (cofactors of column 0,
 * det = their dot product with column 0, inverse = cofactor(c, r) * (1/det)); products and 3-term sums are taken as
 * (a0 + a1) + a2.  That sum order is an ASSUMPTION about a third-party library that is absent here: the one thing
 * this build cannot pin.  Every formula in eom / start / observe / evaluate / failed / potential is the reference's.
 */
#ifndef GRL_ENVIRONMENT_H_
#define GRL_ENVIRONMENT_H_

#include <math.h>
#include <float.h>
#include <stddef.h>
#include <string>
#include <vector>
#include <iostream>
#include <initializer_list>

namespace grl {

class Exception : public std::exception
{
  std::string what_;
  public:
    Exception(const std::string &what) : what_(what) { }
    virtual ~Exception() throw() { }
    virtual const char *what() const throw() { return what_.c_str(); }
};
class bad_param : public Exception { public: bad_param(const std::string &w) : Exception(w) { } };

class Vector
{
  double d_[16];
  size_t n_;
  public:
    Vector() : n_(0) { }
    explicit Vector(size_t n) : n_(n) { for (size_t i = 0; i < 16; ++i) d_[i] = 0; }
    size_t size() const { return n_; }
    void resize(size_t n) { n_ = n; }
    double &operator[](size_t i) { return d_[i]; }
    const double &operator[](size_t i) const { return d_[i]; }
    struct Comma { Vector *v; Comma &operator,(double x) { (*v)[v->n_++] = x; return *this; } };
    Comma operator<<(double x) { n_ = 0; d_[n_++] = x; Comma c = { this }; return c; }
};
#define GRLREF_BIN(op) \
  inline Vector operator op(const Vector &a, const Vector &b) { Vector c(a.size()); for (size_t i = 0; i < a.size(); ++i) c[i] = a[i] op b[i]; return c; } \
  inline Vector operator op(const Vector &a, double b) { Vector c(a.size()); for (size_t i = 0; i < a.size(); ++i) c[i] = a[i] op b; return c; } \
  inline Vector operator op(double a, const Vector &b) { Vector c(b.size()); for (size_t i = 0; i < b.size(); ++i) c[i] = a op b[i]; return c; }
GRLREF_BIN(+) GRLREF_BIN(-) GRLREF_BIN(*) GRLREF_BIN(/)
#undef GRLREF_BIN
inline Vector operator-(const Vector &a) { Vector c(a.size()); for (size_t i = 0; i < a.size(); ++i) c[i] = -a[i]; return c; }

template <typename... Args> inline Vector VectorConstructor(Args... args)
{
  const double v[] = { (double)args... };
  Vector r(sizeof...(args));
  for (size_t i = 0; i < sizeof...(args); ++i) r[i] = v[i];
  return r;
}
inline Vector ConstantVector(size_t n, double d) { Vector r(n); for (size_t i = 0; i < n; ++i) r[i] = d; return r; }
inline Vector extend(const Vector &a, const Vector &b)
{
  Vector c(a.size() + b.size());
  for (size_t i = 0; i < a.size(); ++i) c[i] = a[i];
  for (size_t i = 0; i < b.size(); ++i) c[a.size() + i] = b[i];
  return c;
}

struct Matrix { };                                   /* rewardHessian (off the path) */
inline Matrix diagonal(const Vector &) { return Matrix(); }

class Observation
{
  public:
    Vector v; bool absorbing; Vector u;
    Observation() : absorbing(false) { }
    Observation(const Vector &x) : v(x), absorbing(false) { }
    double &operator[](size_t i) { return v[i]; }
    const double &operator[](size_t i) const { return v[i]; }
    size_t size() const { return v.size(); }
    operator Vector&() { return v; }
    operator const Vector&() const { return v; }
};
class Action
{
  public:
    Vector v; int type; double logp;
    Action() : type(0), logp(0) { }
    Action(const Vector &x) : v(x), type(0), logp(0) { }
    double &operator[](size_t i) { return v[i]; }
    const double &operator[](size_t i) const { return v[i]; }
    size_t size() const { return v.size(); }
    operator Vector&() { return v; }
    operator const Vector&() const { return v; }
};

/* ---- inert Configurable plumbing: configure() runs on the defaults the class constructors set ---- */
struct ConfigurationParameter
{
  double num; bool set;
  ConfigurationParameter() : num(0), set(false) { }
  template <typename T> operator T() const { return (T)num; }
  Vector v() const { return Vector(); }
  std::string str() const { return ""; }
  void *ptr() const { return 0; }
};
class Configuration
{
  std::vector<std::pair<std::string, double> > nums_;
  public:
    void put(const std::string &k, double v) { nums_.push_back(std::make_pair(k, v)); }
    ConfigurationParameter operator[](const std::string &k) const
    {
      ConfigurationParameter p;
      for (size_t i = 0; i < nums_.size(); ++i) if (nums_[i].first == k) { p.num = nums_[i].second; p.set = true; }
      if (!p.set) throw Exception("ref_shim: parameter '" + k + "' was not fed by the harness");
      return p;
    }
    bool has(const std::string &) const { return false; }
    template <typename T> bool get(const std::string &k, T &out) const
    {
      for (size_t i = 0; i < nums_.size(); ++i) if (nums_[i].first == k) { out = (T)nums_[i].second; return true; }
      return false;
    }
    template <typename T> void set(const std::string &, const T &) { }
};
struct CRP
{
  enum Mutability { Provided, System, Configuration, Online };
  template <typename... Args> CRP(Args...) { }
  template <typename... Args> CRP(const char *, const char *, const std::string &, Mutability, std::initializer_list<const char*>) { }
};
struct ConfigurationRequest { void push_back(const CRP &) { } };
#define REGISTER_CONFIGURABLE(x)
#define TYPEINFO(t, d)

class Configurable
{
  public:
    virtual ~Configurable() { }
    virtual void request(ConfigurationRequest *) { }
    virtual void configure(Configuration &) { }
    virtual void reconfigure(const Configuration &) { }
};

/* RandGen::get() (utils.h:145-187): the harness feeds the uniform draw */
struct RandGen { static double &fed() { static double u = 0; return u; } static double get() { return fed(); } };
struct Rand { double getNormal(double mu, double) { return mu; } };

class Dynamics : public Configurable
{
  public:
    virtual void eom(const Vector &state, const Vector &actuation, Vector *xd) const = 0;
};

class Task : public Configurable
{
  public:
    virtual void request(ConfigurationRequest *) { }
    virtual void start(int test, Vector *state) = 0;
    virtual bool actuate(const Vector &, const Vector &, const Action &action, Vector *actuation) const { *actuation = action; return true; }
    virtual void observe(const Vector &state, Observation *obs, int *terminal) const = 0;
    virtual void evaluate(const Vector &state, const Action &action, const Vector &next, double *reward) const = 0;
    virtual bool invert(const Observation &, Vector *, double = 0.) const { return false; }
    virtual Matrix rewardHessian(const Vector &, const Action &) const { return Matrix(); }
};

/* off the hot path: kept only so that the regulator classes in the three sources compile */
class RegulatorTask : public Task
{
  protected:
    Vector start_, goal_, goal_action_, stddev_, q_, r_, min_, max_;
    double timeout_;
  public:
    RegulatorTask() : timeout_(10) { }
    void request(ConfigurationRequest *) { }
    void configure(Configuration &) { }
    void reconfigure(const Configuration &) { }
    void start(int, Vector *state) { *state = start_; }
    void observe(const Vector &, Observation *, int *terminal) const { *terminal = 0; }
    void evaluate(const Vector &, const Action &, const Vector &, double *reward) const { *reward = 0; }
};

}  // namespace grl

/* ---- Eigen stand-in: exactly what cart_double_pole.cpp:78-104 uses ---- */
namespace Eigen {
template <int N> struct CommaInit
{
  double *p; int k;
  CommaInit &operator,(double v) { p[k++] = v; return *this; }
};
struct Vector3d
{
  double d[3];
  Vector3d() { d[0] = d[1] = d[2] = 0; }
  Vector3d(double a, double b, double c) { d[0] = a; d[1] = b; d[2] = c; }
  double &operator[](int i) { return d[i]; }
  const double &operator[](int i) const { return d[i]; }
  CommaInit<3> operator<<(double v) { d[0] = v; CommaInit<3> c = { d, 1 }; return c; }
};
inline Vector3d operator-(const Vector3d &a, const Vector3d &b) { return Vector3d(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
struct Matrix3d
{
  double m[3][3];                                  /* m[row][col]; the comma initialiser fills row by row */
  double flat[9];
  CommaInit<9> operator<<(double v) { flat[0] = v; CommaInit<9> c = { flat, 1 }; return c; }
  void sync() { for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) m[r][c] = flat[3 * r + c]; }
  static double cof(const double (&a)[3][3], int i, int j)
  {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    return a[i1][j1] * a[i2][j2] - a[i1][j2] * a[i2][j1];
  }
  /* Eigen 3.4 LU/InverseImpl.h, compute_inverse<MatrixType, ResultType, 3> */
  Matrix3d inverse() const
  {
    Matrix3d s = *this; s.sync();
    Matrix3d r;
    const double c0 = cof(s.m, 0, 0), c1 = cof(s.m, 1, 0), c2 = cof(s.m, 2, 0);
    const double det = (c0 * s.m[0][0] + c1 * s.m[1][0]) + c2 * s.m[2][0];
    const double invdet = 1 / det;
    r.m[0][0] = c0 * invdet; r.m[0][1] = c1 * invdet; r.m[0][2] = c2 * invdet;
    for (int i = 1; i < 3; ++i) for (int j = 0; j < 3; ++j) r.m[i][j] = cof(s.m, j, i) * invdet;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) r.flat[3 * i + j] = r.m[i][j];
    return r;
  }
};
inline Vector3d operator*(const Matrix3d &a_, const Vector3d &v)
{
  Matrix3d a = a_; a.sync();
  Vector3d r;
  for (int i = 0; i < 3; ++i) r[i] = (a.m[i][0] * v[0] + a.m[i][1] * v[1]) + a.m[i][2] * v[2];
  return r;
}
}  // namespace Eigen

#endif

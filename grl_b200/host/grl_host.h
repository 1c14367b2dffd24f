// grl_host.h -- the slice of GRL's Configurable plugin contract that the batched CUDA path sits behind.
//
// GRL objects are subclasses of grl::Configurable that announce their parameters in request(), receive
// them in configure(), accept messages in reconfigure() and self-register under a type string with
// REGISTER_CONFIGURABLE (reference: base/include/grl/configurable.h:247-250,704-826, factory.h:81-91).
// A YAML file is a tree of such objects; a scalar value may be a path to an earlier object or to one of its
// (provided) parameters, and `type:` may be any unique suffix of a registered type
// (base/src/configurable.cpp:106-125,732-752).  This header restates that contract -- same names, same
// argument meaning, same error behaviour (grl::bad_param / grl::Exception, no error codes) -- so that the
// object sources under grl_b200/host/ read like a GRL addon and can be moved into
// addons/cuda/ of a GRL checkout with only the #include changed.
// Nothing here is copied from the reference: the registry, the YAML subset reader and the configurator
// are written for this repository and support exactly what the configs on the hot path use.
#pragma once
#include <cstdint>
#include <cstdio>
#include <functional>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace grl {

typedef std::vector<double> Vector;

class Exception : public std::runtime_error
{
  public:
    explicit Exception(const std::string &what) : std::runtime_error(what) { }
};

/// Thrown by configure() when a parameter value is unacceptable ("type:param").
class bad_param : public Exception
{
  public:
    explicit bad_param(const std::string &which) : Exception("Parameter " + which + " has an illegal value") { }
};

extern unsigned char grl_log_verbosity__;
#define GRL_LOG(l, tag, m) do { if ((l) <= grl::grl_log_verbosity__) { std::ostringstream o__; o__ << m; \
  ((l) < 2 ? std::cerr : std::cout) << tag << " " << o__.str() << std::endl; } } while (0)
#define ERROR(m)   GRL_LOG(0, "[ERR]", m)
#define WARNING(m) GRL_LOG(1, "[WRN]", m)
#define NOTICE(m)  GRL_LOG(2, "[NTC]", m)
#define INFO(m)    GRL_LOG(3, "[INF]", m)
#define DEBUG(m)   GRL_LOG(4, "[DBG]", m)

class Configurable;

/// One value in a Configuration: a scalar/vector literal or a borrowed object pointer.
class ConfigurationParameter
{
  public:
    ConfigurationParameter() : ptr_(nullptr) { }
    explicit ConfigurationParameter(const std::string &v) : value_(v), ptr_(nullptr) { }
    explicit ConfigurationParameter(Configurable *p) : ptr_(p) { }
    const std::string &str() const { return value_; }
    Configurable *ptr() const { return ptr_; }
    Vector v() const;                       ///< "[a, b]" / "a" / "[]" -> Vector
    operator int() const;
    operator double() const;
    operator size_t() const { return (size_t)(int)(*this); }
  private:
    std::string value_;
    Configurable *ptr_;
};

/// Name -> value map handed to configure() / reconfigure().
class Configuration
{
  public:
    bool has(const std::string &key) const { return values_.count(key) != 0; }
    const ConfigurationParameter &operator[](const std::string &key) const;
    void set(const std::string &key, const std::string &v) { values_[key] = ConfigurationParameter(v); }
    void set(const std::string &key, const char *v) { set(key, std::string(v)); }
    void set(const std::string &key, double v);
    void set(const std::string &key, int v) { set(key, std::to_string(v)); }
    void set(const std::string &key, const Vector &v);
    void set(const std::string &key, Configurable *p) { values_[key] = ConfigurationParameter(p); }
    template <class T> bool get(const std::string &key, T &out) const
    {
      if (!has(key)) return false;
      out = (T)(*this)[key];
      return true;
    }
    const std::map<std::string, ConfigurationParameter> &values() const { return values_; }
  private:
    std::map<std::string, ConfigurationParameter> values_;
};

/// Configuration request parameter: what an object wants (or provides), with its default.
struct CRP
{
  enum Mutability { Configuration, System, Online, Provided };
  std::string name, type, description, value;
  Mutability mutability;
  bool optional;
  // scalar / vector parameter with a default
  CRP(const std::string &n, const std::string &d, int v, Mutability m = Configuration, int = 0, int = 0)
    : name(n), type("int"), description(d), value(std::to_string(v)), mutability(m), optional(true) { }
  CRP(const std::string &n, const std::string &d, double v, Mutability m = Configuration, double = 0, double = 0);
  CRP(const std::string &n, const std::string &d, const Vector &v, Mutability m = Configuration);
  // string parameter (name, description, value) -- or, with m == Provided, a provided parameter
  // (name, type, description): same signature, told apart by the mutability like the reference does
  CRP(const std::string &n, const std::string &d, const std::string &v, Mutability m = Configuration)
    : name(n), type("string"), description(d), value(v), mutability(m), optional(true)
  {
    if (m == Provided) { type = d; description = v; value.clear(); }
  }
  CRP(const std::string &n, const std::string &d, const char *v, Mutability m = Configuration)
    : CRP(n, d, std::string(v), m) { }
  // typed scalar ("int.memory", "vector.action_min", ...)
  CRP(const std::string &n, const std::string &t, const std::string &d, int v, Mutability m = Configuration)
    : name(n), type(t), description(d), value(std::to_string(v)), mutability(m), optional(true) { }
  CRP(const std::string &n, const std::string &t, const std::string &d, double v, Mutability m = Configuration, double = 0, double = 0);
  CRP(const std::string &n, const std::string &t, const std::string &d, const Vector &v, Mutability m = Configuration);
  // object parameter: accepts any registered type that starts with t (before the role suffix ".role")
  CRP(const std::string &n, const std::string &t, const std::string &d, Configurable *, bool opt = false)
    : name(n), type(t), description(d), mutability(Configuration), optional(opt) { is_object = true; }
  bool is_object = false;
};
typedef std::vector<CRP> ConfigurationRequest;

#define TYPEINFO(t, d)\
    static std::string s_type() { return t; }\
    virtual std::string d_type() const { return t; }\
    virtual std::string description() const { return d; }

class Configurable
{
  public:
    TYPEINFO("", "Base object")
    virtual ~Configurable() { }
    virtual void request(ConfigurationRequest * /*config*/) { }
    virtual void request(const std::string & /*role*/, ConfigurationRequest *config) { request(config); }
    virtual void configure(Configuration & /*config*/) { }
    virtual void reconfigure(const Configuration & /*config*/) { }
    virtual Configurable &copy(const Configurable & /*obj*/) { return *this; }
    const std::string &path() const { return path_; }
    void setPath(const std::string &p) { path_ = p; }
  private:
    std::string path_ = "<unknown>";
};

/// Type registry (reference: DECLARE_FACTORY / REGISTER_FACTORY, factory.h:45-91).
class ConfigurableFactory
{
  public:
    typedef std::function<Configurable *()> Creator;
    static std::map<std::string, Creator> &types();
    static bool add(const std::string &type, Creator c) { types()[type] = c; return true; }
    /// Full type for a (possibly abbreviated) YAML type string; throws if unknown or ambiguous.
    static std::string expand(const std::string &type);
    static Configurable *create(const std::string &type);
};
/// dlopen (RTLD_NOW | RTLD_GLOBAL) every `libaddon*.so` in `dir` so that the static REGISTER_CONFIGURABLE objects of
/// the addons register their types (reference: loadPlugins(), base/src/grl.cpp:73-111).  Returns the number loaded.
int loadPlugins(const std::string &dir);

#define REGISTER_CONFIGURABLE(subx) \
  static bool registered_##subx##__ = grl::ConfigurableFactory::add(subx::s_type(), []() -> grl::Configurable * { return new subx(); });

// ---- YAML subset ----------------------------------------------------------------------------------
/// Block mappings (two-space or any consistent indentation), plain scalars, flow sequences "[a, b]".
struct YamlNode
{
  enum Kind { Scalar, Map } kind = Scalar;
  std::string scalar;
  std::vector<std::pair<std::string, std::shared_ptr<YamlNode>>> children;
  const YamlNode *find(const std::string &key) const
  {
    for (auto &c : children) if (c.first == key) return c.second.get();
    return nullptr;
  }
};
std::shared_ptr<YamlNode> parseYAML(const std::string &text);
std::shared_ptr<YamlNode> loadYAMLFile(const std::string &file);

// ---- Configurator: instantiates the object tree of a YAML document ------------------------------------
class Configurator
{
  public:
    ~Configurator();
    /// Instantiate every top-level object section that has a registered type; unknown top-level sections
    /// (visualizer, visualization: graphics, unavailable here) are skipped with a notice.
    void instantiate(const YamlNode &root, const std::vector<std::string> &only = {});
    Configurable *object(const std::string &path) const;
    bool value(const std::string &path, std::string *out) const;
    /// Walk action (reset / save / load) over every object, in construction order.
    void walk(const Configuration &config);
    std::string yaml() const;     ///< resolved configuration, in construction order
  private:
    Configurable *build(const YamlNode &node, const std::string &path, const std::string &role);
    std::vector<std::pair<std::string, Configurable *>> objects_;   // construction order, owned
    std::map<std::string, Configurable *> by_path_;
    std::map<std::string, std::string> values_;                     // path/param -> literal (incl. provided)
    std::map<std::string, Configurable *> value_objects_;           // path/param -> referenced object
    std::map<std::string, std::string> aliases_;                    // path/param given as a reference -> path of the object it names
};

}  // namespace grl

// grl_host.cpp -- registry, YAML-subset reader and configurator behind grl_host.h.
#include "grl_host.h"

#include <dlfcn.h>
#include <glob.h>

#include <algorithm>
#include <cmath>
#include <fstream>
#include <iomanip>

namespace grl {

unsigned char grl_log_verbosity__ = 3;

// ---- values -----------------------------------------------------------------------------------------
static std::string trim(const std::string &s)
{
  size_t b = s.find_first_not_of(" \t\r\n"), e = s.find_last_not_of(" \t\r\n");
  return b == std::string::npos ? std::string() : s.substr(b, e - b + 1);
}

static std::string fmt(double v)
{
  std::ostringstream o;
  o << std::setprecision(17) << v;
  return o.str();
}

static std::string fmt(const Vector &v)
{
  std::string s = "[";
  for (size_t i = 0; i < v.size(); ++i) s += (i ? ", " : "") + fmt(v[i]);
  return s + "]";
}

static bool parse_double(const std::string &s, double *out)
{
  const std::string t = trim(s);
  if (t.empty()) return false;
  char *end = nullptr;
  double v = strtod(t.c_str(), &end);
  if (*end) return false;
  *out = v;
  return true;
}

Vector ConfigurationParameter::v() const
{
  std::string t = trim(value_);
  Vector out;
  if (t.empty()) return out;
  if (t.front() == '[')
  {
    if (t.back() != ']') throw Exception("Malformed vector '" + value_ + "'");
    t = t.substr(1, t.size() - 2);
  }
  std::stringstream ss(t);
  std::string item;
  while (std::getline(ss, item, ','))
  {
    if (trim(item).empty()) continue;
    double d;
    if (!parse_double(item, &d)) throw Exception("Malformed vector element '" + item + "'");
    out.push_back(d);
  }
  return out;
}

ConfigurationParameter::operator double() const
{
  double d;
  if (!parse_double(value_, &d))
  {
    Vector vv = v();                       // "[3]" is accepted where a scalar is wanted, like the reference
    if (vv.size() != 1) throw Exception("Parameter value '" + value_ + "' is not a number");
    d = vv[0];
  }
  return d;
}

ConfigurationParameter::operator int() const
{
  double d = (double)(*this);
  if (d != std::floor(d)) throw Exception("Parameter value '" + value_ + "' is not an integer");
  return (int)d;
}

const ConfigurationParameter &Configuration::operator[](const std::string &key) const
{
  static const ConfigurationParameter empty;
  auto it = values_.find(key);
  return it == values_.end() ? empty : it->second;
}
void Configuration::set(const std::string &key, double v) { set(key, fmt(v)); }
void Configuration::set(const std::string &key, const Vector &v) { set(key, fmt(v)); }

CRP::CRP(const std::string &n, const std::string &d, double v, Mutability m, double, double)
  : name(n), type("double"), description(d), value(fmt(v)), mutability(m), optional(true) { }
CRP::CRP(const std::string &n, const std::string &d, const Vector &v, Mutability m)
  : name(n), type("vector"), description(d), value(fmt(v)), mutability(m), optional(true) { }
CRP::CRP(const std::string &n, const std::string &t, const std::string &d, double v, Mutability m, double, double)
  : name(n), type(t), description(d), value(fmt(v)), mutability(m), optional(true) { }
CRP::CRP(const std::string &n, const std::string &t, const std::string &d, const Vector &v, Mutability m)
  : name(n), type(t), description(d), value(fmt(v)), mutability(m), optional(true) { }

// ---- factory ----------------------------------------------------------------------------------------
std::map<std::string, ConfigurableFactory::Creator> &ConfigurableFactory::types()
{
  static std::map<std::string, Creator> t;
  return t;
}

std::string ConfigurableFactory::expand(const std::string &type)
{
  if (types().count(type)) return type;
  std::vector<std::string> hits;
  const std::string suffix = "/" + type;
  for (auto &kv : types())
    if (kv.first.size() > suffix.size() && kv.first.compare(kv.first.size() - suffix.size(), suffix.size(), suffix) == 0)
      hits.push_back(kv.first);
  if (hits.size() == 1) return hits[0];
  if (hits.empty()) throw Exception("Unknown object type '" + type + "'");
  throw Exception("Ambiguous object type '" + type + "' (" + hits[0] + ", " + hits[1] + ", ...)");
}

Configurable *ConfigurableFactory::create(const std::string &type)
{
  return types().at(expand(type))();
}

// ---- YAML subset ------------------------------------------------------------------------------------
namespace {
struct Line { int indent; std::string key, value; };
}

std::shared_ptr<YamlNode> parseYAML(const std::string &text)
{
  std::vector<Line> lines;
  std::stringstream ss(text);
  std::string raw;
  int lineno = 0;
  while (std::getline(ss, raw))
  {
    ++lineno;
    // strip comments that are not inside a flow sequence or quotes (none of the path configs quote '#')
    size_t hash = raw.find('#');
    if (hash != std::string::npos && (hash == 0 || raw[hash - 1] == ' ')) raw = raw.substr(0, hash);
    if (trim(raw).empty() || trim(raw) == "---") continue;
    size_t ind = raw.find_first_not_of(' ');
    if (raw[ind] == '\t') throw Exception("YAML line " + std::to_string(lineno) + ": tabs are not allowed for indentation");
    size_t colon = raw.find(':', ind);
    if (colon == std::string::npos)
      throw Exception("YAML line " + std::to_string(lineno) + ": expected 'key: value' (block sequences are not supported)");
    Line l;
    l.indent = (int)ind;
    l.key = trim(raw.substr(ind, colon - ind));
    l.value = trim(raw.substr(colon + 1));
    if (l.value.size() >= 2 && ((l.value.front() == '"' && l.value.back() == '"') || (l.value.front() == '\'' && l.value.back() == '\'')))
      l.value = l.value.substr(1, l.value.size() - 2);
    lines.push_back(l);
  }
  auto root = std::make_shared<YamlNode>();
  root->kind = YamlNode::Map;
  std::vector<std::pair<int, YamlNode *>> stack;       // (indent of the node's children, node)
  stack.push_back({-1, root.get()});
  for (size_t i = 0; i < lines.size(); ++i)
  {
    const Line &l = lines[i];
    while (stack.size() > 1 && l.indent <= stack.back().first) stack.pop_back();
    auto node = std::make_shared<YamlNode>();
    const bool has_children = l.value.empty() && i + 1 < lines.size() && lines[i + 1].indent > l.indent;
    if (has_children) node->kind = YamlNode::Map;
    else { node->kind = YamlNode::Scalar; node->scalar = l.value; }
    stack.back().second->children.push_back({l.key, node});
    if (has_children) stack.push_back({l.indent, node.get()});
  }
  return root;
}

std::shared_ptr<YamlNode> loadYAMLFile(const std::string &file)
{
  std::ifstream ifs(file);
  if (!ifs.good()) throw Exception("Could not open configuration file '" + file + "'");
  std::stringstream ss;
  ss << ifs.rdbuf();
  return parseYAML(ss.str());
}

// ---- configurator -----------------------------------------------------------------------------------
Configurator::~Configurator()
{
  for (auto it = objects_.rbegin(); it != objects_.rend(); ++it) delete it->second;
}

int loadPlugins(const std::string &dir)
{
  glob_t globbuf;
  int loaded = 0;
  const std::string pattern = dir + "/libaddon*.so";
  NOTICE("Looking for plugins in '" << pattern << "'");
  if (glob(pattern.c_str(), 0, nullptr, &globbuf) == 0)
  {
    for (size_t ii = 0; ii < globbuf.gl_pathc; ++ii)
    {
      NOTICE("Loading plugin '" << globbuf.gl_pathv[ii] << "'");
      if (!dlopen(globbuf.gl_pathv[ii], RTLD_NOW | RTLD_GLOBAL))
        ERROR("Error loading plugin '" << globbuf.gl_pathv[ii] << "': " << dlerror());
      else
        ++loaded;
    }
    globfree(&globbuf);
  }
  return loaded;
}

Configurable *Configurator::object(const std::string &path_in) const
{
  // a leading '/' anchors the path at the root (configurable.cpp: absolute references, cfg/pendulum/multi_sarsa_tc.yaml);
  // every path here is already root-relative
  const std::string path = !path_in.empty() && path_in[0] == '/' ? path_in.substr(1) : path_in;
  auto it = by_path_.find(path);
  if (it != by_path_.end()) return it->second;
  auto jt = value_objects_.find(path);
  return jt == value_objects_.end() ? nullptr : jt->second;
}

bool Configurator::value(const std::string &path_in, std::string *out) const
{
  const std::string path = !path_in.empty() && path_in[0] == '/' ? path_in.substr(1) : path_in;
  auto it = values_.find(path);
  if (it != values_.end()) { *out = it->second; return true; }
  // a path through a parameter that was itself given as a reference ("…/critic/projector/memory" where
  // critic/projector names the policy's projector, cfg/pendulum/ac_tc.yaml): follow it like the reference's
  // ReferenceConfigurator does (configurable.cpp:318-346)
  for (size_t cut = path.rfind('/'); cut != std::string::npos && cut > 0; cut = path.rfind('/', cut - 1))
  {
    auto at = aliases_.find(path.substr(0, cut));
    if (at != aliases_.end()) return value(at->second + path.substr(cut), out);
  }
  return false;
}

static std::string base_type(const std::string &t)
{
  size_t dot = t.find('.');                 // "projector.pair" -> type "projector", role "pair"
  return dot == std::string::npos ? t : t.substr(0, dot);
}
static std::string role_of(const std::string &t)
{
  size_t dot = t.find('.');
  return dot == std::string::npos ? std::string() : t.substr(dot + 1);
}

Configurable *Configurator::build(const YamlNode &node, const std::string &path, const std::string &role)
{
  const YamlNode *tn = node.find("type");
  if (!tn || tn->kind != YamlNode::Scalar) throw Exception(path + ": object definition has no 'type'");
  const std::string type = ConfigurableFactory::expand(tn->scalar);
  std::unique_ptr<Configurable> obj(ConfigurableFactory::create(type));
  obj->setPath(path);
  ConfigurationRequest request;
  obj->request(role, &request);
  Configuration config;
  for (const CRP &crp : request)
  {
    if (crp.mutability == CRP::Provided) continue;
    const YamlNode *pn = node.find(crp.name);
    const std::string ppath = path + "/" + crp.name;
    if (crp.is_object)
    {
      Configurable *child = nullptr;
      if (pn && pn->kind == YamlNode::Map) child = build(*pn, ppath, role_of(crp.type));
      else if (pn && !pn->scalar.empty() && pn->scalar != "0")
      {
        child = object(pn->scalar);         // reference to an earlier object
        if (!child) throw Exception(ppath + ": '" + pn->scalar + "' does not name an object defined earlier");
        aliases_[ppath] = child->path();
      }
      if (child)
      {
        const std::string want = base_type(crp.type);
        if (child->d_type().compare(0, want.size(), want) != 0)
          throw Exception(ppath + ": object of type '" + child->d_type() + "' given where '" + want + "' is required");
        value_objects_[ppath] = child;
      }
      else if (!crp.optional) throw Exception(ppath + ": required object parameter is not set");
      config.set(crp.name, child);
      continue;
    }
    std::string v = crp.value;
    if (pn)
    {
      if (pn->kind != YamlNode::Scalar) throw Exception(ppath + ": expected a value, found a sub-tree");
      v = pn->scalar;
      std::string referenced;
      if (value(v, &referenced)) v = referenced;         // path to an earlier (possibly provided) parameter
    }
    values_[ppath] = v;
    config.set(crp.name, v);
  }
  for (auto &c : node.children)
  {
    if (c.first == "type") continue;
    bool known = false;
    for (const CRP &crp : request) known |= crp.name == c.first;
    if (!known) WARNING(path << ": ignoring unknown parameter '" << c.first << "'");
  }
  try { obj->configure(config); }
  catch (bad_param &e) { ERROR(path << ": " << e.what()); throw; }
  // provided parameters become referencable by later objects
  for (const CRP &crp : request)
    if (crp.mutability == CRP::Provided && config.has(crp.name))
    {
      if (config[crp.name].ptr()) value_objects_[path + "/" + crp.name] = config[crp.name].ptr();
      else values_[path + "/" + crp.name] = config[crp.name].str();
    }
  Configurable *raw = obj.release();
  objects_.push_back({path, raw});
  by_path_[path] = raw;
  return raw;
}

void Configurator::instantiate(const YamlNode &root, const std::vector<std::string> &only)
{
  for (auto &c : root.children)
  {
    if (!only.empty() && std::find(only.begin(), only.end(), c.first) == only.end()) continue;
    const YamlNode *tn = c.second->find("type");
    if (c.second->kind != YamlNode::Map || !tn)
    {
      values_[c.first] = c.second->scalar;
      continue;
    }
    try { ConfigurableFactory::expand(tn->scalar); }
    catch (Exception &)
    {
      NOTICE("skipping top-level section '" << c.first << "' of unregistered type '" << tn->scalar << "'");
      continue;
    }
    build(*c.second, c.first, "");
  }
}

void Configurator::walk(const Configuration &config)
{
  for (auto &o : objects_) o.second->reconfigure(config);
}

std::string Configurator::yaml() const
{
  std::ostringstream o;
  for (auto &obj : objects_)
  {
    o << obj.first << ": " << obj.second->d_type() << "\n";
    const std::string prefix = obj.first + "/";
    for (auto &kv : values_)
      if (kv.first.compare(0, prefix.size(), prefix) == 0 && kv.first.find('/', prefix.size()) == std::string::npos)
        o << "  " << kv.first.substr(prefix.size()) << ": " << kv.second << "\n";
  }
  return o.str();
}

}  // namespace grl

/* oracle/ref_shim_eigen/yaml-cpp/yaml.h -- STAND-IN: the reference's configurable.h names YAML::Node in one declaration
 * (configurable.h:700, loadYAML); no YAML is parsed by the test harness.  TEST INFRASTRUCTURE ONLY. */
#ifndef GRLREF_YAML_STANDIN
#define GRLREF_YAML_STANDIN
namespace YAML { class Node { }; }
#endif

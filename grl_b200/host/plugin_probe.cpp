// plugin_probe -- registration the way GRL does it (base/src/grl.cpp:73-111, factory.h:81-91): this executable links the
// framework library only (libgrl_b200.so: registry, configurator, YAML reader) and has NO link dependency on the addon;
// loadPlugins() globs and dlopens libaddon*.so, whose static REGISTER_CONFIGURABLE objects then register themselves.
// Prints one line per requested type: "<type> <d_type of the created object>"; exit code 1 if any is missing.
#include "grl_host.h"

#include <libgen.h>
#include <unistd.h>

int main(int argc, char **argv)
{
  char self[4096];
  ssize_t len = readlink("/proc/self/exe", self, sizeof(self) - 1);
  if (len < 0) return 2;
  self[len] = 0;
  const std::string dir = dirname(self);
  const size_t before = grl::ConfigurableFactory::types().size();
  const int loaded = grl::loadPlugins(dir);
  printf("plugins %d types_before %zu types_after %zu\n", loaded, before, grl::ConfigurableFactory::types().size());
  int missing = 0;
  for (int i = 1; i < argc; ++i)
  {
    try
    {
      grl::Configurable *obj = grl::ConfigurableFactory::create(argv[i]);
      printf("%s %s\n", argv[i], obj->d_type().c_str());
      delete obj;
    }
    catch (grl::Exception &e)
    {
      printf("%s MISSING (%s)\n", argv[i], e.what());
      ++missing;
    }
  }
  return missing ? 1 : 0;
}

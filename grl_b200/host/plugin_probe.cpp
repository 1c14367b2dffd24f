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
  // --requests <type>...: the parameters each type requests (Configurable::request), one line per parameter in the format of
  // the reference-side dump (oracle/ref_agent.cpp, refag_requests): "<type>: name|type|default|mutability|optional"
  if (argc > 1 && std::string(argv[1]) == "--requests")
  {
    for (int i = 2; i < argc; ++i)
    {
      try
      {
        grl::Configurable *obj = grl::ConfigurableFactory::create(argv[i]);
        grl::ConfigurationRequest req;
        obj->request("", &req);
        for (const grl::CRP &p : req)
        {
          const char m = p.mutability == grl::CRP::Provided ? 'P' : p.mutability == grl::CRP::System ? 'S'
                       : p.mutability == grl::CRP::Online ? 'O' : 'C';
          printf("%s: %s|%s|%s|%c|%d\n", argv[i], p.name.c_str(), p.type.c_str(), p.is_object ? "0" : p.value.c_str(), m, p.optional ? 1 : 0);
        }
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

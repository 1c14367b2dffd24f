// Host-side check of the physical table layouts (grl_b200/csrc/grlcu_layout.h), compiled and run by
// tests/test_host_layout.py -- no GPU.  Test infrastructure only.
//
//   layout_check tile <cfg.bin> <inputs.bin> <n> <out.bin>
//       builds the tile-major layout for the configuration, walks the kernel's slot computation on the host for
//       every input (obs[OBS] doubles, then the action index as a double) and writes, per input, the T LOGICAL
//       indices it arrives at (iperm[slot]); the test compares them with the oracle's projection of the same input.
//       Also verifies that perm / iperm are inverse bijections, that slotmap[] is the identity for words that own
//       their slot (the fused kernel reads slotmap[] unconditionally) and agrees with the ownership flags.
//   layout_check visits <cfg.bin> <seed> <n_keys>
//       builds the visit-ordered layout from synthetic pilot keys and verifies the bijection and the sector property.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include "../grl_b200/csrc/grlcu_layout.h"

using namespace grlcu_host;

static int fail(const char *what, long a = 0, long b = 0)
{
  fprintf(stderr, "layout_check: %s (%ld, %ld)\n", what, a, b);
  return 1;
}

static bool read_cfg(const char *path, grlcu_config &cfg)
{
  FILE *f = fopen(path, "rb");
  if (!f) return false;
  const size_t got = fread(&cfg, 1, sizeof(cfg), f);
  fclose(f);
  return got == sizeof(cfg);
}

static int check_bijection(const Layout &L, uint32_t mem)
{
  if (L.perm.size() != mem) return fail("perm size", (long)L.perm.size(), mem);
  if (L.iperm.size() != L.tstride) return fail("iperm size", (long)L.iperm.size(), L.tstride);
  std::vector<char> seen(L.tstride, 0);
  for (uint32_t x = 0; x < mem; ++x)
  {
    const uint32_t s = L.perm[x];
    if (s >= L.tstride) return fail("perm out of the row", x, s);
    if (seen[s]) return fail("two indices on one slot", x, s);
    seen[s] = 1;
    if (L.iperm[s] != x) return fail("iperm is not the inverse", x, s);
  }
  uint32_t padding = 0;
  for (uint32_t s = 0; s < L.tstride; ++s) if (!seen[s]) { ++padding; if (L.iperm[s] != 0xFFFFFFFFu) return fail("padding slot with an index", s, L.iperm[s]); }
  if (padding != L.tstride - mem) return fail("padding count", padding, L.tstride - mem);
  return 0;
}

static int run_tile(int argc, char **argv)
{
  if (argc < 6) return fail("usage: tile cfg inputs n out");
  grlcu_config cfg;
  if (!read_cfg(argv[2], cfg)) return fail("cannot read the configuration");
  const int T = cfg.tilings, OBS = cfg.obs_dims, A = cfg.n_actions, ia = OBS;
  double scaling[GRLCU_MAX_DIMS + 1];
  int32_t wrap[GRLCU_MAX_DIMS + 1];
  for (int i = 0; i <= OBS; ++i)      // TileCodingProjector::configure (tile_coding.cpp:66-79)
  {
    scaling[i] = cfg.tilings / cfg.resolution[i];
    wrap[i] = (int32_t)round(cfg.wrapping[i] * scaling[i]);
  }
  const Layout L = build_layout(cfg, scaling, wrap);
  const long n = atol(argv[4]);
  FILE *out = fopen(argv[5], "wb");
  if (!out) return fail("cannot write");
  const uint32_t enabled = L.enabled ? 1u : 0u;
  fwrite(&enabled, 4, 1, out);
  if (!L.enabled) { fclose(out); return 0; }
  if (int r = check_bijection(L, (uint32_t)cfg.memory)) return r;
  // slotmap: identity for owners, an earlier group's slot otherwise; consistent with gflags
  uint32_t owners = 0, sharers = 0;
  for (uint32_t group = 0; group < L.ncells * (uint32_t)T; ++group)
  {
    const uint32_t own = (L.gflags[group >> 3] >> ((group & 7u) * 4u)) & 15u;
    for (int aa = 0; aa < A; ++aa)
    {
      const uint32_t home = group * 4 + aa, s = L.slotmap[home];
      if (s >= L.tstride) return fail("slotmap out of the row", home, s);
      if ((own >> aa) & 1u) { ++owners; if (s != home) return fail("owner whose slotmap entry is not its own slot", home, s); }
      else { ++sharers; if (s == home) return fail("sharer mapped onto its own home", home, s); if (s > home) return fail("sharer mapped onto a LATER slot", home, s); }
    }
  }
  if (owners + sharers != L.inbox_words || sharers != L.shared_words) return fail("word counts", owners, sharers);

  std::vector<double> in((size_t)n * (OBS + 1));
  FILE *f = fopen(argv[3], "rb");
  if (!f || fread(in.data(), sizeof(double), in.size(), f) != in.size()) return fail("cannot read the inputs");
  fclose(f);
  std::vector<uint32_t> logical((size_t)n * T);
  long geometric = 0;
  for (long r = 0; r < n; ++r)
  {
    const double *obs = &in[(size_t)r * (OBS + 1)];
    const int aa = (int)obs[OBS];
    int32_t qz[GRLCU_MAX_DIMS];
    for (int i = 0; i < OBS; ++i) qz[i] = (int32_t)floor(obs[i] * scaling[i]);
    const int32_t qa = (int32_t)floor(cfg.actions[aa] * scaling[ia]);
    for (int j = 0; j < T; ++j)
    {
      // the kernel's geometric path (q_values, GEO): coarse cell of tiling j in every dimension
      bool inbox = true;
      uint32_t cell = 0;
      for (int i = 0; i < OBS; ++i)
      {
        int32_t k = (qz[i] - ((j * (1 + 2 * i)) & (T - 1))) >> 4;
        k += L.koff[i];
        if (L.wrapped[i]) k = k < 0 ? k + L.kn[i] : k >= L.kn[i] ? k - L.kn[i] : k;
        inbox = inbox && (uint32_t)k < (uint32_t)L.kn[i];
        cell = cell * (uint32_t)L.kn[i] + (uint32_t)k;
      }
      uint32_t slot;
      if (inbox) { slot = L.slotmap[(cell * T + j) * 4u + aa]; ++geometric; }
      else
      {
        // the hash path (tile_coding.cpp:122-140, tile_coding.h:78-118) + perm[]
        int32_t coords[GRLCU_MAX_DIMS + 2];
        for (int i = 0; i < OBS; ++i)
        {
          coords[i] = qz[i] - safe_mod_host(qz[i] - j * (1 + 2 * i), T);
          if (wrap[i] != 0) coords[i] = safe_mod_host(coords[i], wrap[i]);
        }
        coords[OBS] = qa - safe_mod_host(qa - j * (1 + 2 * ia), T);
        coords[OBS + 1] = j;
        slot = L.perm[tile_hash_host(coords, OBS + 2) % (uint32_t)cfg.memory];
      }
      logical[(size_t)r * T + j] = L.iperm[slot];
    }
  }
  fwrite(logical.data(), 4, logical.size(), out);
  fclose(out);
  printf("tile-major: cells %u, stride %u, in-box words %u, sharing %u, %ld of %ld lookups geometric\n",
         L.ncells, L.tstride, L.inbox_words, L.shared_words, geometric, n * T);
  return 0;
}

static int run_visits(int argc, char **argv)
{
  if (argc < 5) return fail("usage: visits cfg seed n_keys");
  grlcu_config cfg;
  if (!read_cfg(argv[2], cfg)) return fail("cannot read the configuration");
  uint64_t s = (uint64_t)atoll(argv[3]) * 0x9E3779B97F4A7C15ull + 1;
  auto next = [&s]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
  uint32_t kact[8][16], kj[16];
  for (int a = 0; a < 8; ++a) for (int j = 0; j < 16; ++j) kact[a][j] = (uint32_t)next();
  for (int j = 0; j < 16; ++j) kj[j] = (uint32_t)next();
  // a pilot stream with revisits: a few thousand distinct groups drawn with a skewed distribution
  const long n_keys = atol(argv[4]);
  const uint32_t distinct = (uint32_t)cfg.memory;          // more groups than sectors: some stay unranked
  std::vector<uint64_t> pool(distinct);
  for (auto &k : pool) k = ((next() & 0xFFFFFFFFull) << 8) | (next() & 15u);
  std::vector<uint64_t> keys((size_t)n_keys);
  for (auto &k : keys) { const uint64_t u = next() % distinct; k = pool[(size_t)((u * u) / distinct)]; }
  const Layout L = build_layout_from_visits(cfg, keys, kact, kj);
  if (!L.enabled) return fail("layout not enabled");
  if (L.tstride != (uint32_t)cfg.memory) return fail("stride", L.tstride, cfg.memory);
  if (int r = check_bijection(L, (uint32_t)cfg.memory)) return r;
  // sector property: the words of a ranked group that own their slots sit in that group's sector, action a at offset a
  const uint32_t m = 0x5bd1e995u, mem = (uint32_t)cfg.memory;
  long together = 0, groups_seen = 0;
  std::vector<uint64_t> uniq(keys);
  std::sort(uniq.begin(), uniq.end());
  uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
  for (uint64_t key : uniq)
  {
    const uint32_t h = (uint32_t)(key >> 8);
    const int j = (int)(key & 255u);
    uint32_t slot[3];
    for (int aa = 0; aa < 3; ++aa)
    {
      uint32_t g = (h * m) ^ kact[aa][j];
      g = (g * m) ^ kj[j];
      g ^= g >> 13; g *= m; g ^= g >> 15;
      slot[aa] = L.perm[g % mem];
    }
    ++groups_seen;
    // a whole group in one sector with three distinct slots: it is the group's own sector, word a at offset a (words
    // elsewhere share the slot of an earlier group, or the group was never ranked)
    if ((slot[0] >> 2) == (slot[1] >> 2) && (slot[1] >> 2) == (slot[2] >> 2) &&
        slot[0] != slot[1] && slot[1] != slot[2] && slot[0] != slot[2])
    {
      ++together;
      for (int aa = 0; aa < 3; ++aa) if ((slot[aa] & 3u) != (uint32_t)aa) return fail("word not at its action's offset", slot[aa], aa);
    }
  }
  if (L.ncells == 0 || L.ncells > mem / 4) return fail("sectors handed out", L.ncells, mem / 4);
  if (together * 4 < (long)L.ncells) return fail("fewer than a quarter of the sectors hold a whole group", together, L.ncells);
  printf("visit-ordered: %ld keys, %ld groups, %u sectors, %ld whole groups in one sector, words sharing %u of %u\n",
         n_keys, groups_seen, L.ncells, together, L.shared_words, L.inbox_words);
  return 0;
}

int main(int argc, char **argv)
{
  if (argc >= 2 && !strcmp(argv[1], "tile")) return run_tile(argc, argv);
  if (argc >= 2 && !strcmp(argv[1], "visits")) return run_visits(argc, argv);
  return fail("usage: layout_check tile|visits ...");
}

// grlcu_layout.h -- host side of the tile-major physical table layout (struct Geo in grlcu_agent.cuh, DESIGN.md
// section 4): builds, for one context, the permutation logical index -> physical slot shared by all its agents.
// The logical index is the reference's h % memory (tile_coding.h:116-118, tile_coding.cpp:103-149); nothing here
// changes which tiles share a weight, only where the weight is stored.
#pragma once
#include <stdint.h>
#include <math.h>
#include <vector>
#include <algorithm>
#include <utility>
#include "../../include/grlcu.h"

namespace grlcu_host {

struct Layout
{
  bool enabled = false;
  int  kn[GRLCU_MAX_DIMS] = {0}, koff[GRLCU_MAX_DIMS] = {0}, wrapped[GRLCU_MAX_DIMS] = {0};
  uint32_t ncells = 0, tstride = 0, inbox_words = 0, shared_words = 0;
  std::vector<uint32_t> perm, iperm, slotmap, gflags;
};

// MurmurHash2 over the int coordinates, tile_coding.h:78-118 (the same restatement as tile_hash<> on the device)
inline uint32_t tile_hash_host(const int32_t *ints, int n)
{
  const uint32_t m = 0x5bd1e995u;
  uint32_t h = 449u ^ (uint32_t)n;
  for (int i = 0; i < n; ++i)
  {
    uint32_t k = (uint32_t)ints[i];
    k *= m; k ^= k >> 24; k *= m;
    h *= m; h ^= k;
  }
  h ^= h >> 13; h *= m; h ^= h >> 15;
  return h;
}

inline int32_t safe_mod_host(int32_t x, int32_t y) { int32_t r = x % y; return r < 0 ? r + y : r; }

// The observation box of the task (what its configure() publishes as observation_min / observation_max):
// pendulum.cpp:83-84.  Observations outside the box are served by the hash path, so the box only has to cover
// where agents normally are.
inline bool task_box(const grlcu_config &cfg, double *lo, double *hi)
{
  const double pi = 3.14159265358979323846;
  if (cfg.dynamics == GRLCU_DYN_PENDULUM) { lo[0] = 0; hi[0] = 2 * pi; lo[1] = -12 * pi; hi[1] = 12 * pi; return true; }
  return false;
}

// scaling / wrap exactly as TileCodingProjector::configure (tile_coding.cpp:66-79)
inline Layout build_layout(const grlcu_config &cfg, const double *scaling, const int32_t *wrap)
{
  Layout L;
  const int T = cfg.tilings, OBS = cfg.obs_dims, A = cfg.n_actions, ia = OBS;
  double lo[GRLCU_MAX_DIMS], hi[GRLCU_MAX_DIMS];
  if (cfg.algo == GRLCU_ALGO_AC || T != 16 || A != 3 || (cfg.memory & 3) || !task_box(cfg, lo, hi)) return L;
  if (wrap[ia] != 0) return L;
  uint64_t ncells = 1;
  for (int i = 0; i < OBS; ++i)
  {
    if (wrap[i] != 0)
    {
      if (wrap[i] % T) return L;
      L.wrapped[i] = 1; L.kn[i] = wrap[i] / T; L.koff[i] = 0;
    }
    else
    {
      // q in [qlo, qhi]; tiling offsets b in [0, T): k = floor((q - b) / T)
      const int32_t qlo = (int32_t)floor(lo[i] * scaling[i]), qhi = (int32_t)floor(hi[i] * scaling[i]);
      const int32_t klo = (int32_t)floor((double)(qlo - (T - 1)) / T), khi = (int32_t)floor((double)qhi / T);
      L.koff[i] = -klo; L.kn[i] = khi - klo + 1;
    }
    ncells *= (uint64_t)L.kn[i];
    if (ncells > (1u << 24)) return L;
  }
  // worth it only while most tiles own their slot: tiles of the box <= table entries
  if (ncells * T * A > (uint64_t)cfg.memory) return L;
  L.ncells = (uint32_t)ncells;
  const uint32_t G = L.ncells * T * 4;
  L.tstride = G > (uint32_t)cfg.memory ? G : (uint32_t)cfg.memory;
  L.tstride = (L.tstride + 3u) & ~3u;
  const uint32_t UNSET = 0xFFFFFFFFu;
  L.perm.assign(cfg.memory, UNSET);
  L.iperm.assign(L.tstride, UNSET);
  L.slotmap.assign(G, UNSET);
  L.gflags.assign((L.ncells * T + 7) / 8, 0);
  // action coordinate per (action, tiling): tile_coding.cpp:122-140 on the discretizer's values
  std::vector<int32_t> ca(A * T);
  for (int aa = 0; aa < A; ++aa)
  {
    const int32_t qa = (int32_t)floor(cfg.actions[aa] * scaling[ia]);
    for (int j = 0; j < T; ++j) ca[aa * T + j] = qa - safe_mod_host(qa - j * (1 + 2 * ia), T);
  }
  std::vector<int> k(OBS, 0);
  for (uint32_t cell = 0; cell < L.ncells; ++cell)
  {
    uint32_t rem = cell;
    for (int i = OBS - 1; i >= 0; --i) { k[i] = (int)(rem % (uint32_t)L.kn[i]); rem /= (uint32_t)L.kn[i]; }
    for (int j = 0; j < T; ++j)
    {
      int32_t coords[GRLCU_MAX_DIMS + 2];
      for (int i = 0; i < OBS; ++i)
      {
        const int32_t b = (j * (1 + 2 * i)) & (T - 1);
        coords[i] = b + T * (k[i] - L.koff[i]);          // wrapped: koff = 0 and b + T*k is already in [0, wrap)
      }
      const uint32_t group = cell * T + j;
      uint32_t own = 0;
      for (int aa = 0; aa < A; ++aa)
      {
        coords[OBS] = ca[aa * T + j];
        coords[OBS + 1] = j;
        const uint32_t x = tile_hash_host(coords, OBS + 2) % (uint32_t)cfg.memory;
        const uint32_t home = group * 4 + aa;
        ++L.inbox_words;
        if (L.perm[x] == UNSET) { L.perm[x] = home; L.iperm[home] = x; own |= 1u << aa; }
        else ++L.shared_words;
        L.slotmap[home] = L.perm[x];
      }
      L.gflags[group >> 3] |= own << ((group & 7u) * 4u);
    }
  }
  // indices no tile of the box maps to: into the holes (padding words, homes of colliding tiles), then the tail
  uint32_t next = 0;
  for (uint32_t x = 0; x < (uint32_t)cfg.memory; ++x)
    if (L.perm[x] == UNSET)
    {
      while (L.iperm[next] != UNSET) ++next;
      L.perm[x] = next; L.iperm[next] = x;
    }
  L.enabled = true;
  return L;
}

// Visit-ordered layout for tasks whose tiles do not fit the table (the hash really compresses: cart-pole, cart-double-pole).
// `keys`: the groups a pilot run projected ((hash state after the observation coordinates) << 8 | tiling), with
// repetitions.  Groups are ranked by visit count; the A action words of the most visited groups get consecutive slots
// padded to a sector, exactly as in build_layout, until a quarter of the table's entries have become group bases; a word
// whose index is already placed shares that slot; everything else fills the holes.  The kernel finds a slot through
// perm[] alone (Geo.enabled = 2), so an unranked group costs nothing but its co-location.
// kact / kj: the pre-mixed action and tiling words of Params (grlcu_create).
inline Layout build_layout_from_visits(const grlcu_config &cfg, const std::vector<uint64_t> &keys,
                                       const uint32_t (*kact)[16], const uint32_t *kj)
{
  Layout L;
  const int A = cfg.n_actions;
  if (keys.empty() || A != 3 || cfg.tilings != 16 || (cfg.memory & 3)) return L;
  // visits per group, and the rank below which a group gets a sector of its own
  std::vector<uint64_t> sorted(keys);
  std::sort(sorted.begin(), sorted.end());
  std::vector<std::pair<uint32_t, uint64_t>> ranked;          // (visits, key)
  for (size_t i = 0; i < sorted.size();)
  {
    size_t e = i;
    while (e < sorted.size() && sorted[e] == sorted[i]) ++e;
    ranked.push_back(std::make_pair((uint32_t)(e - i), sorted[i]));
    i = e;
  }
  std::sort(ranked.begin(), ranked.end(), [](const std::pair<uint32_t, uint64_t> &a, const std::pair<uint32_t, uint64_t> &b)
            { return a.first != b.first ? a.first > b.first : a.second < b.second; });
  const uint32_t mem = (uint32_t)cfg.memory, max_groups = mem / 4;
  if (ranked.size() > max_groups) ranked.resize(max_groups);
  std::vector<uint64_t> chosen(ranked.size());
  for (size_t i = 0; i < ranked.size(); ++i) chosen[i] = ranked[i].second;
  std::sort(chosen.begin(), chosen.end());
  std::vector<char> placed(chosen.size(), 0);
  const uint32_t UNSET = 0xFFFFFFFFu, m = 0x5bd1e995u;
  L.perm.assign(mem, UNSET);
  L.iperm.assign(mem, UNSET);
  uint32_t groups = 0;
  // Sectors are handed out in FIRST-VISIT order along the pilot agents' own trajectories (`keys` is agent after agent, each
  // in projection order): the 16 groups of one observation are discovered together and a trajectory moves continuously,
  // so groups that are read together end up in neighbouring sectors (64-byte fetches and L2 lines get shared).
  for (size_t r = 0; r < keys.size(); ++r)
  {
    const size_t at = std::lower_bound(chosen.begin(), chosen.end(), keys[r]) - chosen.begin();
    if (at >= chosen.size() || chosen[at] != keys[r] || placed[at]) continue;
    placed[at] = 1;
    const uint32_t h = (uint32_t)(keys[r] >> 8);
    const int j = (int)(keys[r] & 255u);
    uint32_t x[3];
    bool any_free = false;
    for (int aa = 0; aa < A; ++aa)
    {
      uint32_t g = (h * m) ^ kact[aa][j];
      g = (g * m) ^ kj[j];
      g ^= g >> 13; g *= m; g ^= g >> 15;
      x[aa] = g % mem;
      any_free |= L.perm[x[aa]] == UNSET;
    }
    if (!any_free) { L.inbox_words += A; L.shared_words += A; continue; }
    const uint32_t base = groups++ * 4;
    for (int aa = 0; aa < A; ++aa)
    {
      ++L.inbox_words;
      if (L.perm[x[aa]] == UNSET) { L.perm[x[aa]] = base + aa; L.iperm[base + aa] = x[aa]; }
      else ++L.shared_words;
    }
  }
  uint32_t next = 0;
  for (uint32_t xx = 0; xx < mem; ++xx)
    if (L.perm[xx] == UNSET)
    {
      while (L.iperm[next] != UNSET) ++next;
      L.perm[xx] = next; L.iperm[next] = xx;
    }
  L.tstride = mem;
  L.ncells = groups;
  L.enabled = true;
  return L;
}

}  // namespace grlcu_host

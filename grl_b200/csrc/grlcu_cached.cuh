// grlcu_cached.cuh -- SARSA(lambda) update with the eligibility trace's weights held in a dense
// per-agent value cache instead of being read-modify-written in the hashed table every step.
//
// Why: the reference's trace sweep (representation.h:79-83) does ~100 random 8-byte RMWs per step into a
// 512 KiB table.  On B200 a random 8-byte RMW over tens of GiB runs at ~15 G/s and a random read at
// ~36 G/s (tools/randbench.cu, profiles/r01_randbench_*.txt), so the uncached path sits exactly on that
// ceiling (~1.05e8 agent-steps/s).  A weight that is in the replacing trace is touched every step for up
// to L steps, so it is kept in cv[slot][agent] (coalesced, streamed) while it is in the trace:
//
//   * directory in the table itself: while index x is cached in trace slot s, table[x] holds a NaN-boxed
//     sentinel {hi = kSentinelHi, lo = s}.  Every table read checks for it and, if present, takes the value
//     from cv[s].  That resolves Q(s',.), Q(p) and -- because the sentinel names the slot -- also does the
//     replacing trace's ssub (projection.h:93-104) with no index comparisons at all: the old slot that
//     holds one of p's indices is exactly the slot its sentinel names.
//   * replacing trace => a valid index lives in exactly one slot (ssub removed it from all older entries),
//     so cv[s] is the single copy of table[x]; duplicate indices inside one projection are folded onto
//     their first occurrence (alias slots), which preserves the reference's sequential += order.
//   * values return to the table when their entry is popped (trace.h:229-233), when the trace is cleared
//     (sarsa.cpp:126-132) and when the kernel exits, so between launches the table and the trace ring are
//     in the plain (uncached) representation the replay kernel and grlcu_get_table use.
//
// Every floating-point operation is the one the reference performs on that weight, in the same order:
// results are bit-identical to the uncached path (tests/test_gpu_cached.py) and hence to the oracle.
#pragma once
#include "grlcu_agent.cuh"

namespace grlcu {

constexpr int kSentinelHi = 0x7FF4C0DE;       // signalling-NaN pattern no arithmetic produces
constexpr uint32_t kAliasBit = 0x80000000u;   // ring value: duplicate of slot (value & 15) of the same entry

__device__ __forceinline__ bool is_sentinel(double v) { return __double2hiint(v) == kSentinelHi; }
__device__ __forceinline__ int sentinel_slot(double v) { return __double2loint(v); }
__device__ __forceinline__ double make_sentinel(int slot) { return __hiloint2double(kSentinelHi, slot); }

// Cache geometry: slot = pos*T + j (pos = ring position); cv[(a/32)*cap*T*32 + slot*32 + a%32]; vmask[(a/32)*cap*32 + pos*32 + a%32] bit j = valid.
template <int T>
struct Cache
{
  double   *cv;
  uint32_t *vmask;
  uint32_t *ring;    // trace_idx
  // all three arrays are blocked by warp (warp_rows, grlcu_agent.cuh): the 16 slots of a ring entry are 256 bytes apart
  // in cv and 128 in ring -- compile-time strides: one base address per entry and immediate offsets instead of a
  // multiply-add per access (the address arithmetic of val() was a fifth of the instructions the pendulum kernel
  // executed, profiles/r02_ncu_pendulum.txt), and what a warp streams per entry is contiguous in DRAM
  __device__ __forceinline__ void bind(const Params &p, int a)
  {
    cv = warp_rows(p.cache_val, a, (size_t)p.trace_cap * T);
    vmask = warp_rows(p.cache_mask, a, (size_t)p.trace_cap);
    ring = warp_rows(p.trace_idx, a, (size_t)p.trace_cap * T);
  }
  __device__ __forceinline__ double &val(int slot) const { return cv[(uint32_t)slot * kLanes]; }
  __device__ __forceinline__ uint32_t &mask(int pos) const { return vmask[(uint32_t)pos * kLanes]; }
  __device__ __forceinline__ uint32_t &idx(int pos, int j) const { return ring[(uint32_t)(pos * T + j) * kLanes]; }
};

// 4 bits per slot jj of ring entry `pos`: how many alias slots of the entry duplicate the (valid) slot jj
template <int T>
__device__ __noinline__ uint64_t alias_counts(const Cache<T> &c, int pos, uint32_t m)
{
  // all loads first: rolled, every iteration waited for its own ring word (15 round trips in a row for the warp)
  uint32_t x[T];
#pragma unroll
  for (int j = 1; j < T; ++j) x[j] = c.idx(pos, j);
  uint64_t counts = 0;
#pragma unroll
  for (int j = 1; j < T; ++j)
    if (!((m >> j) & 1u) && x[j] != kInvalidIndex && (x[j] & kAliasBit))
    {
      const int jj = (int)(x[j] & 15u);
      if ((m >> jj) & 1u) counts += 1ull << (4 * jj);
    }
  return counts;
}

// table read through the directory
template <int T>
__device__ __forceinline__ double resolve(const Cache<T> &c, double raw)
{
  return is_sentinel(raw) ? c.val(sentinel_slot(raw)) : raw;
}

// Write the cached values of one entry back into the table (plain representation).
template <int T>
__device__ __noinline__ uint32_t flush_entry(const Cache<T> c, double *tbl, int pos)
{
  const uint32_t m = c.mask(pos);
  // all loads first, then all stores: the compiler cannot hoist a load over a store that may alias it, and one
  // load -> store round trip per tiling was the largest stall of the fused kernels (profiles/r01_ncu_*_v6_*)
  uint32_t ix[T];
  double v[T];
#pragma unroll
  for (int j = 0; j < T; ++j) { ix[j] = c.idx(pos, j); v[j] = c.val(pos * T + j); }
#pragma unroll
  for (int j = 0; j < T; ++j)
    if ((m >> j) & 1u) tbl[ix[j]] = v[j];
  return (uint32_t)__popc(m);          // table stores issued
}

// Kernel entry: the ring is in the plain representation (kInvalidIndex = invalidated).  Build the valid
// masks, fold duplicate indices inside an entry onto their first occurrence, pull the cached weights out
// of the table and leave sentinels behind.
template <int T>
__device__ __noinline__ uint32_t cache_open_core(const Cache<T> c, double *tbl, int cap, int head, int len, uint64_t dupmask)
{
  uint32_t stores = 0;
  for (int k = 0; k < len; ++k)
  {
    const int pos = ring_pos(head, k, cap);
    uint32_t ix[T];
    uint32_t m = 0;
#pragma unroll
    for (int j = 0; j < T; ++j) { ix[j] = c.idx(pos, j); if (ix[j] != kInvalidIndex) m |= 1u << j; }
    if ((dupmask >> k) & 1ull)
    {
#pragma unroll 1
      for (int j = 1; j < T; ++j)
#pragma unroll 1
        for (int jj = j - 1; jj >= 0; --jj)
          if (((m >> j) & 1u) && ((m >> jj) & 1u) && ix[j] == ix[jj]) { m &= ~(1u << j); c.idx(pos, j) = kAliasBit | (uint32_t)jj; }
    }
    c.mask(pos) = m;
    stores += (uint32_t)__popc(m);
#pragma unroll
    for (int j = 0; j < T; ++j)
      if ((m >> j) & 1u)
      {
        c.val(pos * T + j) = ld_tbl(tbl + ix[j]);
        tbl[ix[j]] = make_sentinel(pos * T + j);
      }
  }
  return stores;
}

template <int DYN, int T>
__device__ __forceinline__ void cache_open(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl)
{
  ag.dtblw += cache_open_core<T>(c, tbl, p.trace_cap, ag.head, ag.len, ag.dupmask);
}

// Kernel exit: back to the plain representation (values in the table, kInvalidIndex in the ring for
// invalidated slots, alias slots showing the index they duplicate).
template <int T>
__device__ __noinline__ uint32_t cache_close_core(const Cache<T> c, double *tbl, int cap, int head, int len)
{
  uint32_t stores = 0;
  for (int k = 0; k < len; ++k)
  {
    const int pos = ring_pos(head, k, cap);
    const uint32_t m = c.mask(pos);
    stores += (uint32_t)__popc(m);
#pragma unroll
    for (int j = 0; j < T; ++j)
    {
      const uint32_t x = c.idx(pos, j);
      if ((m >> j) & 1u) tbl[x] = c.val(pos * T + j);
      else if (x != kInvalidIndex && (x & kAliasBit))
      {
        const int jj = (int)(x & 15u);
        c.idx(pos, j) = ((m >> jj) & 1u) ? c.idx(pos, jj) : kInvalidIndex;
      }
      else c.idx(pos, j) = kInvalidIndex;
    }
  }
  return stores;
}

template <int DYN, int T>
__device__ __forceinline__ void cache_close(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl)
{
  ag.dtblw += cache_close_core<T>(c, tbl, p.trace_cap, ag.head, ag.len);
}

// trace.clear() (sarsa.cpp:126-132 via TDAgent::start): cached weights go back to the table.
template <int T>
__device__ __noinline__ uint32_t cache_clear_core(const Cache<T> c, double *tbl, int cap, int head, int len)
{
  uint32_t stores = 0;
  for (int k = 0; k < len; ++k) stores += flush_entry<T>(c, tbl, ring_pos(head, k, cap));
  return stores;
}

template <int DYN, int T>
__device__ __forceinline__ void cache_clear(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl)
{
  if (ag.len > 0) ag.dtblw += cache_clear_core<T>(c, tbl, p.trace_cap, ag.head, ag.len);
  ag.len = 0; ag.total_decay = 1.; ag.dupmask = 0;
}

// SARSAPredictor::criticize (sarsa.cpp:98-124), replacing trace, through the value cache.
//   rawp[j]: table[pidx[j]] as read before any store of this step (value or sentinel).
struct NoEarlyStores { __device__ __forceinline__ void operator()(double) const { } };

//   early(delta): called as soon as the TD error is known, before the sweep: lets a caller issue its own stores to
//   table words it fetched together with rawp while they are still in L2 (the actor update of the actor-critic).
template <int DYN, int T, typename Early = NoEarlyStores>
__device__ __forceinline__ void sarsa_update_cached(const Params &p, Agent<DYN, T> &ag, const Cache<T> &c, double *tbl,
                                                    double reward, bool bootstrap, double q_next,
                                                    const double (&rawp)[T],
                                                    double &o_qp, double &o_target, double &o_delta, int &o_swept,
                                                    Early early = Early())
{
  const int cap = p.trace_cap;
  double target = reward;
  if (bootstrap) target += p.gamma * q_next;    // pow(gamma, 1)
  // read(p): resolve through the directory; remember which old slot holds each index
  double np[T];
  int sp[T];
#pragma unroll
  for (int j = 0; j < T; ++j)
  {
    sp[j] = is_sentinel(rawp[j]) ? sentinel_slot(rawp[j]) : -1;
    np[j] = sp[j] >= 0 ? c.val(sp[j]) : rawp[j];
  }
  double qp = 0;
#pragma unroll
  for (int j = 0; j < T; ++j) qp += np[j];
  qp /= T;
  const double delta = target - qp;
  const double dw = ag.alpha * (target - qp);   // linear.cpp:194
  const double e = ag.e;
  const double ad = ag.alpha * delta;
  const double dc = 0.01;
  const bool pdup = (ag.flags & 8) != 0;

  // write(p): += dw per occurrence, in tiling order.  Duplicates are folded onto the first occurrence.
  uint32_t owner = 0xFFFFu;                     // bit j: slot j owns its index (first occurrence)
  uint64_t dups = 0;                            // 4 bits per owner: number of its duplicates
  if (pdup)
  {
    uint32_t lp[T];
#pragma unroll
    for (int j = 0; j < T; ++j) lp[j] = ag.pidx[j];
    owner = dup_owner_mask(lp, T);
    dups = dup_counts(lp, owner, T);
  }
#pragma unroll
  for (int j = 0; j < T; ++j) np[j] += dw;
  if (pdup)
  {
#pragma unroll
    for (int j = 0; j < T; ++j)
      for (int r = (int)((dups >> (4 * j)) & 15u); r > 0; --r) np[j] += dw;
  }
  // The directory sentinels of p's new slots (ring position head + 1, where trace.add puts p below) are stored NOW,
  // while the sectors that Q(p)'s gather fetched a moment ago are still in L2: after the sweep they are gone (an L2
  // line lives ~40 us, a step of all agents takes ~200 us) and each store would pay a 32-byte fill from DRAM.
  // Nothing between here and the push reads these table words: the update works on rawp / sp / the dense cache, and
  // the flushes below only write indices that ssub has not invalidated, i.e. never one of p's.
  early(delta);
  const int nhead = ag.head + 1 == cap ? 0 : ag.head + 1;
#pragma unroll
  for (int j = 0; j < T; ++j)
    if ((owner >> j) & 1u) tbl[ag.pidx[j]] = make_sentinel(nhead * T + j);
  // indices that an older entry holds: that slot receives p's write now and the sweep's increment below
#pragma unroll
  for (int j = 0; j < T; ++j)
    if (((owner >> j) & 1u) && sp[j] >= 0) c.val(sp[j]) = np[j];

  // update(trace, alpha*delta, gamma*lambda) (representation.h:79-83): newest -> oldest, on the cache
  int swept = 0;
  {
    double weight = 1.;
    for (int k = 0; k < ag.len && weight > 0.001; ++k)
    {
      const int pos = ring_pos(ag.head, k, cap);
      const double inc = weight * ad * e;
      weight *= e;
      const uint32_t m = c.mask(pos);
      double v[T];
      // unconditional loads: the values do not wait for the mask (invalid slots are loaded and never stored)
#pragma unroll
      for (int j = 0; j < T; ++j) v[j] = c.val(pos * T + j);
#pragma unroll
      for (int j = 0; j < T; ++j) v[j] += inc;
      swept += __popc(m);
      if ((ag.dupmask >> k) & 1ull)
      {
        // alias slots: one more sequential += inc on the slot they duplicate (if it is still valid)
        const uint64_t ac_ = alias_counts<T>(c, pos, m);
#pragma unroll
        for (int j = 0; j < T; ++j)
          for (int r = (int)((ac_ >> (4 * j)) & 15u); r > 0; --r) { v[j] += inc; ++swept; }
      }
#pragma unroll
      for (int j = 0; j < T; ++j)
        if ((m >> j) & 1u) c.val(pos * T + j) = v[j];
    }
  }

  // trace.add(p, e) (trace.h:215-234).  ssub: every old slot named by one of p's sentinels is invalidated
  // and its weight moves into p's new slot.
#pragma unroll
  for (int j = 0; j < T; ++j)
    if (((owner >> j) & 1u) && sp[j] >= 0) np[j] = c.val(sp[j]);      // loads first (see flush_entry)
#pragma unroll
  for (int j = 0; j < T; ++j)
    if (((owner >> j) & 1u) && sp[j] >= 0)
    {
      // a reduction, not a read-modify-write: nothing waits for it (the load of each of these used to be a round trip of
      // its own, one after the other, because the compiler cannot tell whether two of them hit the same word); the
      // thread's own later loads and stores of the word are ordered after it (same thread, same address)
      const int pos = sp[j] / T, jj = sp[j] - pos * T;
      atomicAnd(&c.mask(pos), ~(1u << jj));
    }
  if (e < dc) cache_clear<DYN, T>(p, ag, c, tbl);                      // trace.h:220-221
  // push; then pop while total_decay < 0.01 (flushing what leaves)
  ag.head = ag.head + 1 == cap ? 0 : ag.head + 1;
  {
    const int pos = ag.head;
#pragma unroll
    for (int j = 0; j < T; ++j)
    {
      const uint32_t x = ag.pidx[j];
      if ((owner >> j) & 1u) c.val(pos * T + j) = np[j];      // (its sentinel went out before the sweep)
      c.idx(pos, j) = x;
    }
    if (pdup)
    {
      uint32_t lp[T];
#pragma unroll
      for (int j = 0; j < T; ++j) lp[j] = ag.pidx[j];
#pragma unroll 1
      for (int j = 1; j < T; ++j)
        if (!((owner >> j) & 1u)) c.idx(pos, j) = kAliasBit | (uint32_t)dup_owner_of(lp, owner, j);
    }
    c.mask(pos) = owner;
    ag.dtblw += (uint32_t)__popc(owner);           // directory sentinels written
  }
  ag.len++;
  ag.dupmask = (ag.dupmask << 1) | (pdup ? 1ull : 0ull);
  ag.total_decay *= e;
  while (ag.total_decay < dc && ag.len > 1)
  {
    ag.total_decay /= e;
    ag.dtblw += flush_entry<T>(c, tbl, ring_pos(ag.head, ag.len - 1, cap));
    ag.len--;
  }
  ag.dswept += (uint32_t)swept;
  o_qp = qp; o_target = target; o_delta = delta; o_swept = swept;
}

}  // namespace grlcu

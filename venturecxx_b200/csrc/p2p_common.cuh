// Peer-memory plumbing shared by the sharded resample (p2p_resample.cuh) and the
// observation-sharded subsampled MH (blr_subsampled.cuh, E3): the ranks of one NVSwitch node (one process per
// GPU) open each other's buffers once through CUDA IPC; afterwards scalars travel through MAILBOXES in peer
// memory -- the sender stores its payload into its slot of every peer's mailbox as 8-byte words that each carry
// the exchange's number next to 32 bits of payload (no fence, no ordering between the words); the receiver spins
// on its own (local) mailbox until every word shows the expected number -- and bulk state moves
// with plain peer loads / stores over NVLink.  No collective and no host round trip on either path.
//
// An exchange is executed by ONE warp, and it runs in the TAIL of the kernel that produced the payload (the
// "last block" of a grid-wide reduction), not in a kernel of its own: a sharded resample is the same four
// launches as the single-GPU one.
//
// Failure policy: a peer that never arrives trips a wall-clock bound (%globaltimer; VK_P2P_TIMEOUT_S, default
// 20 s).  The waiting lane then sets the STICKY error word par->err and its mirror in host-mapped pinned
// memory: every later kernel of the path returns at once when it is set (the gather poisons the weights with
// NaN), and the next host-synchronising entry point raises (p2p_check_error).  A timed-out slot is never
// consumed.
#pragma once
#include "vk_common.cuh"

namespace vk {

static const int P2P_MAX_WORLD = 16;
static const int P2P_MAX_ARRAYS = 4;
static const int P2P_PHASES = 2;          // 0: (max, min, NaN flag) of the weights; 1: (sum exp, live buffer, mass)
static const int P2P_SS_MAX_ROUNDS = 16;   // minibatch rounds evaluated per subsampled-MH exchange

struct P2PMail {                       // 64 bytes, written by ONE sender into its slot of a peer's mailbox
  // six 8-byte words {payload word (low 32 bits), exchange number (high 32 bits)}: an aligned 8-byte store is
  // atomic, so a word whose high half equals the expected number carries valid data -- no fence between
  // "payload" and "flag", no ordering between the words (the LL protocol of collective libraries)
  unsigned long long w[6];
  unsigned long long pad[2];
};
struct P2PParams {                     // device-resident results of the exchanges
  unsigned long long offset, total;    // this rank's mass offset, global mass
  int cur_idx[P2P_MAX_WORLD];          // which of its two physical buffers is each rank's live one
  int err;                             // sticky: 1 = a wait timed out (peer never arrived); mirrored to the host word
};
struct P2PTables {                     // device-resident peer pointers
  P2PMail *mail[P2P_MAX_WORLD];        // [parity][phase][sender] inside
  int32_t *anc[P2P_MAX_WORLD];         // [parity of the resample number][P] global ancestor of every child slot; -1 = not stored yet
  unsigned char *buf[P2P_MAX_ARRAYS][P2P_MAX_WORLD][2];
  double *ss_mail[P2P_MAX_WORLD];      // subsampled MH: [parity][sender][ss_stride] doubles, see blr_subsampled.cuh
  unsigned long long *ss_seq[P2P_MAX_WORLD];   // [parity][sender] sequence words of the ss mailboxes
};
// everything an exchange needs, by value in the kernel argument list (tab == nullptr: no exchange)
struct P2PXchg {
  P2PTables *tab;
  P2PParams *par;
  volatile int *err_host;              // host-mapped mirror of par->err (read by the host without a copy)
  int world, rank, my_cur_idx;
  unsigned long long seq;              // number of THIS exchange of its phase (per-phase counters)
  long long n_total;                   // particles on all ranks
  long long timeout_ns;
};

// what the observation-sharded subsampled MH (blr_subsampled_persistent.cuh) needs from the peer tables
struct P2PSSView {
  P2PTables *tab; P2PParams *par; int *err_host_dev;
  int world, rank;
  long long ss_stride, ss_pcap, timeout_ns;
};

__device__ __forceinline__ unsigned long long p2p_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ int p2p_mail_index(int parity, int phase, int sender) {
  return (parity * P2P_PHASES + phase) * P2P_MAX_WORLD + sender;
}

// One warp; lane r talks to rank r.  Phase 0 sends (max, min, nan) from red[0..2]; phase 1 sends
// (sum exp, live-buffer index, mass = tile_sums[n_tiles]).  Results: phase 0 -> red[0..2]; phase 1 -> par
// (offset, total, cur_idx) and red[3] (global sum exp), red[4] (LSE, utils.py:77-91 edge cases).
// Returns false (in every lane) when a peer timed out: the results are then NOT written.
__device__ __forceinline__ bool
p2p_exchange_warp(const P2PXchg &x, double *red, const uint64_t *tile_sums, int n_tiles, int phase) {
  const int lane = threadIdx.x & 31;
  const int parity = (int)(x.seq & 1ull);
  double a = 0.0, b = 0.0;
  unsigned long long c = 0ull;
  if (phase == 0) { a = red[0]; b = red[1]; c = red[2] != 0.0 ? 1ull : 0ull; }
  else { a = red[3]; b = (double)x.my_cur_idx; c = tile_sums[n_tiles]; }
  int bad = ((volatile P2PParams *)x.par)->err != 0;
  const unsigned long long flag = (x.seq & 0xffffffffull) << 32;
  if (lane < x.world && !bad) {
    volatile P2PMail *dst = x.tab->mail[lane] + p2p_mail_index(parity, phase, x.rank);
    const unsigned long long ua = (unsigned long long)__double_as_longlong(a), ub = (unsigned long long)__double_as_longlong(b);
    dst->w[0] = (ua & 0xffffffffull) | flag; dst->w[1] = (ua >> 32) | flag;
    dst->w[2] = (ub & 0xffffffffull) | flag; dst->w[3] = (ub >> 32) | flag;
    dst->w[4] = (c & 0xffffffffull) | flag;  dst->w[5] = (c >> 32) | flag;
  }
  double ra = 0.0, rb = 0.0;
  unsigned long long rc = 0ull;
  if (lane < x.world && !bad) {
    volatile P2PMail *src = x.tab->mail[x.rank] + p2p_mail_index(parity, phase, lane);
    const unsigned long long t0 = p2p_now_ns();
    unsigned long long v[6];
    int have = 0, spins = 0;
    while (have != 0x3f) {
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (!((have >> k) & 1)) { v[k] = src->w[k]; if ((v[k] & 0xffffffff00000000ull) == flag) have |= 1 << k; }
      if (have != 0x3f && (++spins & 15) == 0 && (long long)(p2p_now_ns() - t0) > x.timeout_ns) { bad = 1; break; }
    }
    if (!bad) {
      ra = __longlong_as_double((long long)((v[0] & 0xffffffffull) | (v[1] << 32)));
      rb = __longlong_as_double((long long)((v[2] & 0xffffffffull) | (v[3] << 32)));
      rc = (v[4] & 0xffffffffull) | (v[5] << 32);
    }
  }
  bad = __any_sync(0xffffffffu, bad);
  if (bad) {
    if (lane == 0) { x.par->err = 1; *x.err_host = 1; __threadfence_system(); }
    return false;
  }
  if (phase == 0) {
    double mx = lane < x.world ? ra : -INFINITY, mn = lane < x.world ? rb : INFINITY;
    unsigned long long nf = lane < x.world ? rc : 0ull;
    for (int o = 16; o > 0; o >>= 1) {
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      nf |= __shfl_xor_sync(0xffffffffu, nf, o);
    }
    if (lane == 0) { red[0] = mx; red[1] = mn; red[2] = nf ? 1.0 : 0.0; }
  } else {
    // rank-ordered sums (every lane walks the ranks) so that every GPU computes identical values
    unsigned long long off = 0ull, tot = 0ull;
    double se = 0.0;
    for (int r = 0; r < x.world; ++r) {
      unsigned long long m_r = __shfl_sync(0xffffffffu, rc, r);
      double se_r = __shfl_sync(0xffffffffu, ra, r);
      if (r < x.rank) off += m_r;
      tot += m_r;
      se += se_r;
    }
    if (lane < x.world) x.par->cur_idx[lane] = (int)rb;
    if (lane == 0) {
      x.par->offset = off; x.par->total = tot;
      double m = red[0], mn = red[1], nanf = red[2], lse;
      if (x.n_total == 0) lse = -INFINITY;
      else if (nanf != 0.0) lse = NAN;
      else if (isinf(m) && mn != -m) lse = m;
      else if (isinf(m)) lse = NAN;
      else lse = m + log(se);
      red[3] = se; red[4] = lse;
    }
  }
  return true;
}

}  // namespace vk

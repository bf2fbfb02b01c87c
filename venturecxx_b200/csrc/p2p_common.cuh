// Peer-memory plumbing shared by the sharded resample (p2p_resample.cuh) and the
// observation-sharded subsampled MH (blr_subsampled.cuh, E3): the ranks of one NVSwitch node (one process per
// GPU) open each other's buffers once through CUDA IPC; afterwards scalars travel through MAILBOXES in peer
// memory -- the sender stores its payload into its slot of every peer's mailbox as 8-byte words that each carry
// the exchange's number next to 32 bits of payload (no fence, no ordering between the words); the receiver spins
// on its own (local) mailbox until every word shows the expected number -- and bulk state moves
// with plain peer loads / stores over NVLink.  No collective and no host round trip on either path.
//
// An exchange has two halves, and neither is a kernel of its own (a sharded resample is the same four launches as
// the single-GPU one).  SEND: one warp, in the TAIL of the kernel that produced the payload (the "last block" of a
// grid-wide reduction), stores into the peers' mailboxes and returns without waiting.  COLLECT: the first warp of
// EVERY CTA of the kernel that consumes the values spins on the local mailbox in its prologue, so the peers' stores
// travel while the producer drains and the consumer launches, and no single CTA's wait serialises a kernel tail.
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

// SEND half of an exchange: one warp; lane r stores (a, b, c) into this rank's slot of rank r's mailbox.
__device__ __forceinline__ void p2p_send_warp(const P2PXchg &x, int phase, double a, double b, unsigned long long c) {
  const int lane = threadIdx.x & 31;
  const int parity = (int)(x.seq & 1ull);
  const unsigned long long flag = (x.seq & 0xffffffffull) << 32;
  if (lane < x.world && !((volatile P2PParams *)x.par)->err) {
    volatile P2PMail *dst = x.tab->mail[lane] + p2p_mail_index(parity, phase, x.rank);
    const unsigned long long ua = (unsigned long long)__double_as_longlong(a), ub = (unsigned long long)__double_as_longlong(b);
    dst->w[0] = (ua & 0xffffffffull) | flag; dst->w[1] = (ua >> 32) | flag;
    dst->w[2] = (ub & 0xffffffffull) | flag; dst->w[3] = (ub >> 32) | flag;
    dst->w[4] = (c & 0xffffffffull) | flag;  dst->w[5] = (c >> 32) | flag;
  }
}
// COLLECT half: one warp; lane r (< world) spins on rank r's slot of the LOCAL mailbox until all six words carry this
// exchange's number.  Any number of warps (one per CTA of the consuming kernel) may collect the same exchange: a
// slot is rewritten two exchanges of its phase later, and the sender cannot get there before every kernel of this
// rank that collects the present one has finished -- to send number seq + 2 of a phase the peer must have collected
// this rank's seq + 1 of BOTH phases, and this rank sends those after (in stream order) the kernels that collect seq.
// Returns false in every lane after a timeout (par->err set).
__device__ __forceinline__ bool p2p_collect_warp(const P2PXchg &x, int phase, double &ra, double &rb, unsigned long long &rc) {
  const int lane = threadIdx.x & 31;
  const int parity = (int)(x.seq & 1ull);
  const unsigned long long flag = (x.seq & 0xffffffffull) << 32;
  int bad = ((volatile P2PParams *)x.par)->err != 0;
  ra = 0.0; rb = 0.0; rc = 0ull;
  if (lane < x.world && !bad) {
    volatile P2PMail *src = x.tab->mail[x.rank] + p2p_mail_index(parity, phase, lane);
    const unsigned long long t0 = p2p_now_ns();
    unsigned long long v[6];
    int have = 0, spins = 0;
    while (have != 0x3f) {
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (!((have >> k) & 1)) { v[k] = src->w[k]; if ((v[k] & 0xffffffff00000000ull) == flag) have |= 1 << k; }
      if (have != 0x3f && (++spins & 15) == 0) {
        if (((volatile P2PParams *)x.par)->err || (long long)(p2p_now_ns() - t0) > x.timeout_ns) { bad = 1; break; }
      }
    }
    if (!bad) {
      ra = __longlong_as_double((long long)((v[0] & 0xffffffffull) | (v[1] << 32)));
      rb = __longlong_as_double((long long)((v[2] & 0xffffffffull) | (v[3] << 32)));
      rc = (v[4] & 0xffffffffull) | (v[5] << 32);
    }
  }
  bad = __any_sync(0xffffffffu, bad);
  if (bad && lane == 0) { x.par->err = 1; *x.err_host = 1; __threadfence_system(); }
  return !bad;
}
// phase 0 collected: (max, min, NaN flag) over all ranks, in every lane
__device__ __forceinline__ void p2p_reduce_minmax(const P2PXchg &x, double ra, double rb, unsigned long long rc,
                                                  double &mx, double &mn, double &nanf) {
  const int lane = threadIdx.x & 31;
  mx = lane < x.world ? ra : -INFINITY; mn = lane < x.world ? rb : INFINITY;
  unsigned long long nf = lane < x.world ? rc : 0ull;
  for (int o = 16; o > 0; o >>= 1) {
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    nf |= __shfl_xor_sync(0xffffffffu, nf, o);
  }
  nanf = nf ? 1.0 : 0.0;
}
// phase 1 collected: this rank's mass offset, the global mass and the global sum of exponentials, rank-ordered sums
// (every lane walks the ranks) so that every GPU computes identical values
__device__ __forceinline__ void p2p_reduce_mass(const P2PXchg &x, double ra, unsigned long long rc,
                                                unsigned long long &off, unsigned long long &tot, double &se) {
  off = 0ull; tot = 0ull; se = 0.0;
  for (int r = 0; r < x.world; ++r) {
    const unsigned long long m_r = __shfl_sync(0xffffffffu, rc, r);
    const double se_r = __shfl_sync(0xffffffffu, ra, r);
    if (r < x.rank) off += m_r;
    tot += m_r;
    se += se_r;
  }
}
__device__ __forceinline__ double p2p_lse(long long n_total, double m, double mn, double nanf, double se) {   // utils.py:77-91
  if (n_total == 0) return -INFINITY;
  if (nanf != 0.0) return NAN;
  if (isinf(m) && mn != -m) return m;
  if (isinf(m)) return NAN;
  return m + log(se);
}

}  // namespace vk

// kernels_dfs.cuh -- lane-owned depth-first traversal (sm_100a).
//
// Every LANE owns one pose and walks the two OBB trees depth first, exactly like FCL's collisionRecurse
// (/root/reference/third-party/fcl/include/fcl/narrowphase/detail/traversal/traversal_recurse-inl.h:83-128)
// including its early exit at the first intersecting triangle pair -- which is what keeps the number of box
// tests per pose at FCL's ~45 instead of the ~140 of a breadth-first frontier.  What makes it a GPU kernel:
//   * the warp stays converged: one iteration = every lane box-tests the node pair on top of its own stack;
//     a lane that finishes its pose pulls the next one from a per-warp buffer, so lanes never wait for the
//     slowest pose of a tile;
//   * new poses are prepared 32 at a time by the whole warp (coalesced state loads, fp64 pose maths, the root
//     box test) -- poses separated at the root never occupy a lane;
//   * triangle pairs are not tested where they are found: they go to a per-warp queue that is drained 32 pairs
//     at a time with the fp32-filtered SAT-17; the few pairs it leaves open (about one in 500 states with the
//     per-axis bounds of sat17.h) get the fp64 SAT-17 on the spot;
//   * the trees are 32-byte QNodes (qnode.h); their first levels (for alpha-1.2: both whole trees, 129 KB)
//     live in shared memory, where a 32-lane gather costs bank conflicts instead of one L1 wavefront per lane.
#pragma once
#include "kernels.cuh"
#include "qnode.h"
#include "sat17.h"

namespace osr {

struct DfsParams {
  const uint4* env_q;  // QNodes, 2 x uint4 each
  const uint4* rob_q;
  const double* env_tris;
  const double* rob_tris;
  int env_nodes, rob_nodes;
  int env_smem_nodes, rob_smem_nodes;  // leading nodes of each tree cached in shared memory
  Rigid env_tf;                        // environment transform (perturbation)
  double env_quat[4];
  float r_rob;                         // largest |vertex| of the robot model
  DevCounters* counters;
  unsigned long long* work_counter;
  int* error_flag;
  // stream-fed launches (FEED): number of leading states whose upload has completed, advanced by small H2D
  // copies queued behind each chunk of the input on the copy stream (engine.cu: osr_is_valid_states)
  const long long* ready;
  long long ready_floor;  // states certainly resident when the kernel starts (the launch is ordered behind them)
  unsigned long long feed_timeout_ns;  // give up waiting for an upload after this long (%globaltimer; OSR_FEED_TIMEOUT_S)
  // Work distribution: the counter hands out GROUPS of states -- 32 states below tail_start (a multiple of 32),
  // tail_group states from there on.  The last (warps x tail_group) states are spread over every warp of the grid, a
  // few states each, and the free lanes of a warp share its poses at once: the end of a launch (and the end of a
  // stream-fed call, which waits for its last chunk with every warp idle) costs ~30 us instead of ~85.
  long long tail_start;
  int tail_group;
};

// Tunables (the -D overrides exist for the A/B builds of scripts/k1_variants.sh; defaults = measured best)
#ifdef OSR_FEED_BLOCKING
constexpr bool kFeedDefer = false;   // A/B: a warp waits for its next group's upload with its poses stalled
#else
constexpr bool kFeedDefer = true;
#endif
#ifndef OSR_LEAF_FLUSH
#define OSR_LEAF_FLUSH 16
#endif
#ifndef OSR_RE_FLUSH
#define OSR_RE_FLUSH 8
#endif
#ifndef OSR_WAIT_FLUSH
#define OSR_WAIT_FLUSH 12
#endif
#ifndef OSR_PAUSE_ON_LEAF
#define OSR_PAUSE_ON_LEAF 2
#endif
constexpr int DFS_WARPS = 12;   // 3 per scheduler: 4 would need <= 128 registers (16 K registers per SM sub-partition)
constexpr int DFS_THREADS = DFS_WARPS * 32;
constexpr int LEAFQ_CAP = 96;    // >= LEAF_FLUSH - 1 + 64 (at most two new pairs per lane and iteration)
constexpr int REQ_CAP = 64;
constexpr int LEAF_FLUSH = OSR_LEAF_FLUSH;   // drain the triangle-pair queue once it holds this many pairs
constexpr int RE_FLUSH = OSR_RE_FLUSH;       // same for the queue of pairs awaiting the exact fp64 predicate
constexpr int WAIT_FLUSH = OSR_WAIT_FLUSH;   // ... or once this many lanes have nothing left but queued pairs
constexpr int REFILL_MIN = 1;    // pull new poses once this many lanes are idle
// 0: a lane keeps traversing while its triangle pairs wait in a queue; 1: it pauses while pairs wait for the fp32
// filter (with depth-ordered expansion the first pair found is usually the hit); 2: also while pairs wait for fp64
constexpr int PAUSE_ON_LEAF = OSR_PAUSE_ON_LEAF;
constexpr unsigned PAUSE_MASK = PAUSE_ON_LEAF >= 1 ? 4u : 0u;   // F_PL (K1; the segment variant also pauses on F_PR)

// work-stack entry = one OVERLAPPING (environment node, robot node) pair still to be expanded, stored as
// (first child of the node that gets split, the other node, which tree is split); 0 means "none".
template <typename E> struct TaskCodec;
template <> struct TaskCodec<uint32_t> {  // valid:1 | other node:15 | split-env:1 | first child:15
  static __device__ __forceinline__ uint32_t make(unsigned child, unsigned other, bool split_env) {
    return 0x80000000u | (other << 16) | (split_env ? 0x8000u : 0u) | child;
  }
  static __device__ __forceinline__ unsigned child(uint32_t e) { return e & 0x7fffu; }
  static __device__ __forceinline__ unsigned other(uint32_t e) { return (e >> 16) & 0x7fffu; }
  static __device__ __forceinline__ bool split_env(uint32_t e) { return (e & 0x8000u) != 0; }
};
template <> struct TaskCodec<unsigned long long> {  // valid:1 | other node:31 | split-env:1 | first child:31
  static __device__ __forceinline__ unsigned long long make(unsigned child, unsigned other, bool split_env) {
    return 0x8000000000000000ull | ((unsigned long long)other << 32) | (split_env ? 0x80000000ull : 0ull) | child;
  }
  static __device__ __forceinline__ unsigned child(unsigned long long e) { return (unsigned)e & 0x7fffffffu; }
  static __device__ __forceinline__ unsigned other(unsigned long long e) { return (unsigned)(e >> 32) & 0x7fffffffu; }
  static __device__ __forceinline__ bool split_env(unsigned long long e) { return ((unsigned)e & 0x80000000u) != 0; }
};

template <typename E, int DEPTH>
struct __align__(16) DfsWarpMem {
  E stack[DEPTH][32];
  unsigned long long leafq[LEAFQ_CAP];
  float pend[17][32];  // BoxPose of prepared poses: q[4] R[9] t[3] eta
  int pend_idx[32];
  int parts[32];       // work sharing at the end of the input: lanes still working on the pose of home lane i
  int hitf[32];        // ... and whether one of them has found an intersecting pair
  int ready_seen;      // stream-fed launches: states known to be resident (kept out of the register file)
  int pad_[3];
};

// leaf-queue entry: owner lane:5 | epoch:11 | env triangle:24 | robot triangle:24
__device__ __forceinline__ unsigned long long leaf_entry(unsigned lane, unsigned epoch, unsigned ta, unsigned tb) {
  return (unsigned long long)lane | ((unsigned long long)epoch << 5) | ((unsigned long long)ta << 16) |
         ((unsigned long long)tb << 40);
}

template <bool ALL_SMEM>
__device__ __forceinline__ QNode load_qnode(const uint4* sm, const uint4* gl, int nsm, unsigned idx) {
  uint4 a, b;
  if (ALL_SMEM || (int)idx < nsm) { a = sm[2 * idx]; b = sm[2 * idx + 1]; }
  else { a = __ldg(gl + 2 * (size_t)idx); b = __ldg(gl + 2 * (size_t)idx + 1); }
  QNode n;
  n.c[0] = __uint_as_float(a.x); n.c[1] = __uint_as_float(a.y); n.c[2] = __uint_as_float(a.z);
  n.link = (int)a.w;
  n.q01 = b.x; n.q23 = b.y; n.e01 = b.z; n.e2x = b.w;
  return n;
}

// Robot triangle into the environment frame with the reference's operation order
// (intersect-inl.h:598-617: Q' = R*Q + T; then :738-743 everything minus P1); transform in registers.
__device__ __forceinline__ void load_pair_exact_r(const double* __restrict__ Pt, const double* __restrict__ Qt,
                                                  const Rigid& rel, TriPair<double>& t) {
  double p1[3] = {__ldg(Pt), __ldg(Pt + 1), __ldg(Pt + 2)};
#pragma unroll
  for (int k = 0; k < 3; k++) {
    t.p2[k] = __ldg(Pt + 3 + k) - p1[k];
    t.p3[k] = __ldg(Pt + 6 + k) - p1[k];
  }
  double* qo[3] = {t.q1, t.q2, t.q3};
#pragma unroll
  for (int v = 0; v < 3; v++) {
    double x = __ldg(Qt + 3 * v), y = __ldg(Qt + 3 * v + 1), z = __ldg(Qt + 3 * v + 2);
#pragma unroll
    for (int k = 0; k < 3; k++) {
      double w = (rel.R[3 * k] * x + rel.R[3 * k + 1] * y + rel.R[3 * k + 2] * z) + rel.t[k];
      qo[v][k] = w - p1[k];
    }
  }
}

// Descent rule and entry of an overlapping pair (A = environment node ia, B = robot node ib): a leaf is never
// split; otherwise the box of larger volume is.
template <typename E>
__device__ __forceinline__ E expand_entry(const QNode& A, const QNode& B, unsigned ia, unsigned ib, float volA, float volB) {
  typedef TaskCodec<E> TC;
  if (B.link < 0 || (A.link >= 0 && volA > volB)) return TC::make((unsigned)A.link, ib, true);
  return TC::make((unsigned)B.link, ia, false);
}

// One expansion step of a lane: the pair in `cur` is known to overlap; its two child pairs (the split node's
// children against the other node) are box-tested together -- two independent tests per lane and iteration hide
// each other's latency -- triangle pairs are reported to the caller, and of two overlapping child pairs the one
// that overlaps DEEPER is expanded next while the other goes on the lane's stack.
template <typename E, bool ALL_SMEM, typename STACK>
__device__ __forceinline__ void dfs_expand_step(E& cur, int& sp, const int sb, STACK& stack, const int lane,
                                                const uint4* sm_env, const uint4* sm_rob, const DfsParams& P,
                                                const BoxPose& bp, bool& lf0, unsigned& ta0, unsigned& tb0, bool& lf1,
                                                unsigned& ta1, unsigned& tb1) {
  typedef TaskCodec<E> TC;
  const unsigned c = TC::child(cur), o = TC::other(cur);
  const bool se = TC::split_env(cur);
  const unsigned ia0 = se ? c : o, ia1 = se ? c + 1u : o, ib0 = se ? o : c, ib1 = se ? o : c + 1u;
  const QNode A0 = load_qnode<ALL_SMEM>(sm_env, P.env_q, P.env_smem_nodes, ia0);
  const QNode A1 = load_qnode<ALL_SMEM>(sm_env, P.env_q, P.env_smem_nodes, ia1);
  const QNode B0 = load_qnode<ALL_SMEM>(sm_rob, P.rob_q, P.rob_smem_nodes, ib0);
  const QNode B1 = load_qnode<ALL_SMEM>(sm_rob, P.rob_q, P.rob_smem_nodes, ib1);
  float s0, s1, va0, vb0, va1, vb1;
  const bool ov0 = obb_overlap_q2(A0, B0, bp, s0, va0, vb0);
  const bool ov1 = obb_overlap_q2(A1, B1, bp, s1, va1, vb1);
#ifndef OSR_DFS_BRANCHY
  // everything below is selects and predicated shared-memory accesses: the lanes of a warp take all combinations of
  // (overlap, leaf pair, which child first), and branches here cost more in reconvergence than the few ALU
  // instructions they skip
  const bool pl0 = A0.link < 0 && B0.link < 0, pl1 = A1.link < 0 && B1.link < 0;
  lf0 = ov0 && pl0; lf1 = ov1 && pl1;
  ta0 = (unsigned)(-(A0.link + 1)); tb0 = (unsigned)(-(B0.link + 1));
  ta1 = (unsigned)(-(A1.link + 1)); tb1 = (unsigned)(-(B1.link + 1));
  const E x0 = expand_entry<E>(A0, B0, ia0, ib0, va0, vb0), x1 = expand_entry<E>(A1, B1, ia1, ib1, va1, vb1);
  const E e0 = (ov0 && !pl0) ? x0 : (E)0, e1 = (ov1 && !pl1) ? x1 : (E)0;
  const bool both = e0 != 0 && e1 != 0, second_first = s1 > s0;
  if (both) stack[sp][lane] = second_first ? e0 : e1;
  E nxt = both ? (second_first ? e1 : e0) : (e0 != 0 ? e0 : e1);
  const bool pop = nxt == 0 && sp > sb;
  sp += both ? 1 : (pop ? -1 : 0);
  if (pop) nxt = stack[sp][lane];
  cur = nxt;
#else
  E e0 = 0, e1 = 0;
  if (ov0) {
    if (A0.link < 0 && B0.link < 0) { lf0 = true; ta0 = (unsigned)(-(A0.link + 1)); tb0 = (unsigned)(-(B0.link + 1)); }
    else e0 = expand_entry<E>(A0, B0, ia0, ib0, va0, vb0);
  }
  if (ov1) {
    if (A1.link < 0 && B1.link < 0) { lf1 = true; ta1 = (unsigned)(-(A1.link + 1)); tb1 = (unsigned)(-(B1.link + 1)); }
    else e1 = expand_entry<E>(A1, B1, ia1, ib1, va1, vb1);
  }
  if (e0 != 0 && e1 != 0) {
    const bool second_first = s1 > s0;
    stack[sp][lane] = second_first ? e0 : e1;
    sp++;
    cur = second_first ? e1 : e0;
  } else {
    E nxt = e0 != 0 ? e0 : e1;
    if (nxt == 0 && sp > sb) { sp--; nxt = stack[sp][lane]; }
    cur = nxt;
  }
#endif
}

// The entry a pose starts from once its root boxes are known to overlap (0 when both roots are leaves: the only
// triangle pair then goes straight to the pair queue).
template <typename E>
__device__ __forceinline__ E root_entry(const uint4* env_q, const uint4* rob_q) {
  QNode A = load_qnode<false>(nullptr, env_q, 0, 0u), B = load_qnode<false>(nullptr, rob_q, 0, 0u);
  if (A.link < 0 && B.link < 0) return 0;
  float ea0, ea1, ea2, eb0, eb1, eb2, spare;
  unpack_half2(A.e01, ea0, ea1); unpack_half2(A.e2x, ea2, spare);
  unpack_half2(B.e01, eb0, eb1); unpack_half2(B.e2x, eb2, spare);
  return expand_entry<E>(A, B, 0u, 0u, ea0 * ea1 * ea2, eb0 * eb1 * eb2);
}

// ---- pose sources -----------------------------------------------------------------------------------
// A source hands out poses by an integer tag and can rebuild the exact fp64 transform of a tag.
// T = double, or float for callers whose samples are float32 (the reference's own sampler,
// src/GP/pipeline/uw_random.py:8-17, makes float32 arrays that pybind11 widens on the host: widening here is
// the same exact conversion and halves the PCIe traffic of a host call).
template <typename T>
struct ValiditySource {
  const T* q;
  long long n;
  uint8_t* out;
  __device__ __forceinline__ void exact_pose(const DfsParams& P, int tag, Rigid& rel, double* quat) const {
    double s[7];
#pragma unroll
    for (int k = 0; k < 7; k++) s[k] = (double)__ldg(q + 7 * (long long)tag + k);
    Rigid rob;
    state_to_rigid(s, rob);
    relative_rigid(P.env_tf, rob, rel);
    quat[0] = s[3]; quat[1] = s[4]; quat[2] = s[5]; quat[3] = s[6];
  }
  __device__ __forceinline__ void result(int tag, bool hit) const { out[tag] = hit ? 0 : 1; }
};

}  // namespace osr

namespace osr {

// lane flags
constexpr unsigned F_HAS = 1u;  // the lane works on a pose
constexpr unsigned F_HIT = 2u;  // an intersecting triangle pair of that pose is known
constexpr unsigned F_PL = 4u;   // triangle pairs of the pose are waiting in the pair queue
constexpr unsigned F_PR = 8u;   // ... in the exact-recheck queue

// ---------------------------------------------------------------------------------------------------
// K1 (second generation): batched validity, out[i] = 1 iff state i is collision free
// (UnitWorld::isValid, /root/reference/lib/osr/unit_world.cc:215-247).  One persistent block per SM.
// The loop body is kept small on purpose: its instruction working set (box test + stack step + the two
// drains) has to stay inside the 32 KB L1.5 instruction cache shared by the 12 warps.
// ---------------------------------------------------------------------------------------------------
template <typename E, int DEPTH, bool ALL_SMEM, bool FEED, typename T = double>
__global__ void __launch_bounds__(DFS_THREADS, 1)
    k_valid_dfs(DfsParams P, const T* __restrict__ q, long long n, uint8_t* __restrict__ out) {
  typedef TaskCodec<E> TC;
  typedef DfsWarpMem<E, DEPTH> WM;
  extern __shared__ uint4 dfs_smem[];
  uint4* sm_env = dfs_smem;
  uint4* sm_rob = dfs_smem + 2 * (size_t)P.env_smem_nodes;
  WM& M = reinterpret_cast<WM*>(dfs_smem + 2 * (size_t)(P.env_smem_nodes + P.rob_smem_nodes))[threadIdx.x >> 5];
  for (int i = threadIdx.x; i < 2 * P.env_smem_nodes; i += blockDim.x) sm_env[i] = __ldg(P.env_q + i);
  for (int i = threadIdx.x; i < 2 * P.rob_smem_nodes; i += blockDim.x) sm_rob[i] = __ldg(P.rob_q + i);
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const unsigned ltmask = (1u << lane) - 1u;
  ValiditySource<T> src = {q, n, out};
  const E root_task = root_entry<E>(P.env_q, P.rob_q);
  const bool root_leafpair = root_task == 0;  // two single-triangle meshes

  // lane state
  BoxPose bp;
  int tag = 0;
  unsigned epoch = 0, fl = 0;
  E cur = 0;
  int sp = 0, sb = 0;   // the lane's stack is stack[sb .. sp-1][lane]
  int home = lane;      // lane that pulled the pose this lane works on (differs after a steal)
  // warp state (uniform)
  int pend_n = 0, nleaf = 0, since_flush = 0;
  bool exhausted = false, draining_leaf = false, steal_mode = false;
  int held = -1, held_cnt = 0;   // stream-fed launches: a group taken from the counter whose states have not arrived
  unsigned poll_tick = 0;
  unsigned n_bv = 0, n_leaf = 0, n_fp64 = 0;
  unsigned long long nposes = 0;
  if (FEED) M.ready_seen = (int)min(P.ready_floor, 0x7fffffffLL);  // same value from every lane
#ifdef OSR_DFS_STATS
  unsigned st_iters = 0, st_active = 0, st_wait = 0, st_idle = 0;
#endif

  while (true) {
    // (1) completed poses.  Once lanes share poses (steal_mode) a pose is finished by the last of its parts.
    {
      bool done = (fl & F_HAS) && ((fl & F_HIT) || (cur == 0 && !(fl & F_PL)));
      if (!steal_mode) {
        if (done) src.result(tag, (fl & F_HIT) != 0);
      } else {
        if ((fl & F_HAS) && M.hitf[home] != 0) { fl |= F_HIT; done = true; }
        if (done && (fl & F_HIT)) M.hitf[home] = 1;
        __syncwarp();
        if (done && atomicSub(&M.parts[home], 1) == 1) src.result(tag, M.hitf[home] != 0);
        __syncwarp();
      }
      if (done) { fl = 0; cur = 0; sp = 0; sb = 0; }
    }
    // (2) idle lanes: new poses, or (input exhausted) a share of another lane's pose
    unsigned idle = __ballot_sync(FULL, !(fl & F_HAS));
    if (idle != 0) {
      if (pend_n == 0 && !exhausted) {  // prepare the next 32 states with the whole warp
        unsigned long long base = 0;
        int cnt = 32;
        bool defer = false;
        if (FEED && held >= 0) {   // the group taken earlier, whose states had not arrived yet
          base = (unsigned long long)held;
          cnt = held_cnt;
        } else {
          if (lane == 0) base = atomicAdd(P.work_counter, 1ull);   // a group id (DfsParams::tail_start)
          base = __shfl_sync(FULL, base, 0);
          const unsigned long long bulk = (unsigned long long)(P.tail_start >> 5);
          if (base < bulk) base <<= 5;
          else {
            cnt = P.tail_group;
            base = (unsigned long long)P.tail_start + (base - bulk) * (unsigned)cnt;
          }
        }
        if (FEED && kFeedDefer && (long long)base < n && idle != FULL) {
          // Some lanes still hold poses: do not stop them for an upload.  One look at the word every 8th visit; if
          // the group's states are not there yet it is kept and the traversal goes on.  (The upload is slower than
          // the kernel, so warps are here most of the time: waiting with poses half done is what the END of a call
          // -- last chunk in, every warp still owing its stalled poses -- was made of.)
          const long long need = min((long long)base + cnt, n);
          long long seen = __shfl_sync(FULL, (long long)M.ready_seen, 0);
          if (seen < need && (poll_tick++ & 7) == 0) {
            long long v;
            asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(P.ready) : "memory");
            seen = __shfl_sync(FULL, v, 0);
            M.ready_seen = (int)min(seen, 0x7fffffffLL);
          }
          if (seen < need) {
            held = (int)base;
            held_cnt = cnt;
            defer = true;
          }
          __syncwarp();
        }
        if (FEED && !defer && (long long)base < n) {
          // Wait until the states [base, base+32) have been uploaded.  Written so that control flow stays
          // warp-uniform (every lane polls, the value of lane 0 decides): a one-lane polling loop or an
          // out-of-line call here costs the whole traversal loop its convergence (+25 % instructions, +8 % time).
          const long long need = min((long long)base + cnt, n);
          long long seen = __shfl_sync(FULL, (long long)M.ready_seen, 0);
          unsigned long long t0 = 0;   // the wait is bounded in TIME (a poll count depends on the clock and on tools)
          for (int spins = 0; seen < need; spins++) {
            long long v;
            asm volatile("ld.acquire.sys.global.s64 %0, [%1];" : "=l"(v) : "l"(P.ready) : "memory");
            seen = __shfl_sync(FULL, v, 0);
            if (seen >= need) { M.ready_seen = (int)min(seen, 0x7fffffffLL); break; }
            if ((spins & 255) == 255) {
              unsigned long long now;
              asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
              now = __shfl_sync(FULL, now, 0);
              if (t0 == 0) t0 = now;
              else if (now - t0 > P.feed_timeout_ns) {   // gave up: the host re-runs the call without the feed
                atomicExch(P.error_flag, 2);
                base = (unsigned long long)n;
                break;
              }
            }
            __nanosleep(1000);
          }
          __syncwarp();
        }
        if (FEED && !defer) held = -1;
        if (defer) {
        } else if ((long long)base >= n) exhausted = true;
        else {
          long long i = (long long)base + lane;
          bool have = lane < cnt && i < n;
          unsigned hm = __ballot_sync(FULL, have);
          // the root box test runs here, 32 poses at a time: poses separated at the root are free and never
          // occupy a lane; the others are compacted into the buffer of prepared poses
          bool keep = false;
          BoxPose np;
          if (have) {
            Rigid rel; double qd[4];
            src.exact_pose(P, (int)i, rel, qd);
            make_box_pose(rel, P.env_quat, qd, P.r_rob, np);
            const QNode RA = load_qnode<ALL_SMEM>(sm_env, P.env_q, P.env_smem_nodes, 0u);
            const QNode RB = load_qnode<ALL_SMEM>(sm_rob, P.rob_q, P.rob_smem_nodes, 0u);
            float sl, va, vb;
            keep = obb_overlap_q2(RA, RB, np, sl, va, vb);
            n_bv++;
            if (!keep) src.result((int)i, false);
          }
          const unsigned km = __ballot_sync(FULL, keep);
          if (keep) {
            const int slot = __popc(km & ltmask);
#pragma unroll
            for (int k = 0; k < 4; k++) M.pend[k][slot] = np.q[k];
#pragma unroll
            for (int k = 0; k < 9; k++) M.pend[4 + k][slot] = np.R[k];
#pragma unroll
            for (int k = 0; k < 3; k++) M.pend[13 + k][slot] = np.t[k];
            M.pend[16][slot] = np.eta;
            M.pend_idx[slot] = (int)i;
          }
          pend_n = __popc(km);
          nposes += __popc(hm);
          __syncwarp();
        }
      }
      if (pend_n > 0) {
        int k = __popc(idle & ltmask);
        bool pulled = !(fl & F_HAS) && k < pend_n;
        if (pulled) {
          int slot = pend_n - 1 - k;
#pragma unroll
          for (int j = 0; j < 4; j++) bp.q[j] = M.pend[j][slot];
#pragma unroll
          for (int j = 0; j < 9; j++) bp.R[j] = M.pend[4 + j][slot];
#pragma unroll
          for (int j = 0; j < 3; j++) bp.t[j] = M.pend[13 + j][slot];
          bp.eta = M.pend[16][slot];
          tag = M.pend_idx[slot];
          fl = F_HAS; sp = 0; sb = 0;
          epoch = (epoch + 1) & 0x7ffu;
          cur = root_task;
        }
        pend_n -= min(pend_n, __popc(idle));
        if (root_leafpair) {  // degenerate meshes: the only pair goes straight to the pair queue
          unsigned m2 = __ballot_sync(FULL, pulled);
          if (pulled) { M.leafq[nleaf + __popc(m2 & ltmask)] = leaf_entry(lane, epoch, 0, 0); fl |= F_PL; }
          nleaf += __popc(m2);
        }
        __syncwarp();
      } else if (exhausted) {
        // idle lanes take the OLDEST stack entry (the largest pending subtree) of lanes that still have one.
        // Without this the kernel cannot end before its longest single pose does (up to ~2000 box tests in
        // sequence on alpha-1.2) and every warp idles through its own longest pose.
        if (!steal_mode) {
          M.parts[lane] = (fl & F_HAS) ? 1 : 0;
          M.hitf[lane] = 0;
          home = lane;
          steal_mode = true;
          __syncwarp();
        }
        unsigned victims = __ballot_sync(FULL, (fl & F_HAS) && sp > sb);
        int ns = min(__popc(idle), __popc(victims));
        if (ns > 0) {
          int r = __popc(idle & ltmask);
          bool thief = !(fl & F_HAS) && r < ns;
          int v = thief ? (int)__fns(victims, 0, r + 1) : lane;
          BoxPose vb;
#pragma unroll
          for (int k = 0; k < 4; k++) vb.q[k] = __shfl_sync(FULL, bp.q[k], v);
#pragma unroll
          for (int k = 0; k < 9; k++) vb.R[k] = __shfl_sync(FULL, bp.R[k], v);
#pragma unroll
          for (int k = 0; k < 3; k++) vb.t[k] = __shfl_sync(FULL, bp.t[k], v);
          vb.eta = __shfl_sync(FULL, bp.eta, v);
          int vtag = __shfl_sync(FULL, tag, v), vhome = __shfl_sync(FULL, home, v), vsb = __shfl_sync(FULL, sb, v);
          bool robbed = (fl & F_HAS) && sp > sb && __popc(victims & ltmask) < ns;
          E stolen = 0;
          if (thief) stolen = M.stack[vsb][v];
          __syncwarp();
          if (robbed) sb++;
          if (thief) {
            bp = vb; tag = vtag; home = vhome;
            cur = stolen; sp = 0; sb = 0;
            fl = F_HAS;
            epoch = (epoch + 1) & 0x7ffu;
            atomicAdd(&M.parts[home], 1);
          }
          __syncwarp();
        }
      }
    }
    // (3) one box test per lane: the pair on top of the lane's stack; an overlapping pair is replaced by its
    // two child pairs (not while a queue is being drained: the queues' capacity assumes at most one iteration
    // of new triangle pairs on top)
    const bool runnable = cur != 0 && !(fl & PAUSE_MASK);
    const bool any_cur = __any_sync(FULL, runnable);
    if (any_cur && !draining_leaf) {
      bool lf0 = false, lf1 = false;
      unsigned ta0 = 0, tb0 = 0, ta1 = 0, tb1 = 0;
#ifdef OSR_DFS_STATS
      st_iters++;
      st_active += __popc(__ballot_sync(FULL, runnable));
      st_wait += __popc(__ballot_sync(FULL, (fl & F_HAS) && !runnable));
      st_idle += __popc(__ballot_sync(FULL, !(fl & F_HAS)));
#endif
      if (runnable) {
        dfs_expand_step<E, ALL_SMEM>(cur, sp, sb, M.stack, lane, sm_env, sm_rob, P, bp, lf0, ta0, tb0, lf1, ta1, tb1);
        n_bv += 2;
      }
      const unsigned m0 = __ballot_sync(FULL, lf0), m1 = __ballot_sync(FULL, lf1);
      if (m0 | m1) {  // queue the triangle pairs found
        if (lf0) M.leafq[nleaf + __popc(m0 & ltmask)] = leaf_entry(lane, epoch, ta0, tb0);
        if (lf1) M.leafq[nleaf + __popc(m0) + __popc(m1 & ltmask)] = leaf_entry(lane, epoch, ta1, tb1);
        if (lf0 | lf1) fl |= F_PL;
        nleaf += __popc(m0) + __popc(m1);
        __syncwarp();
      }
      since_flush++;
    }
    // (4) drain policy: is one chunk of up to 32 queued triangle pairs processed now?
    bool do_leaf = false;
    if (draining_leaf) do_leaf = true;
    else if (!any_cur) {
      if (nleaf > 0) do_leaf = true;
      else if (!__any_sync(FULL, (fl & F_HAS) != 0) && exhausted && pend_n == 0) break;
    } else if (nleaf > 0) {
      unsigned waiting = __ballot_sync(FULL, (fl & (F_HAS | F_PL)) == (F_HAS | F_PL) && (cur == 0 || (fl & PAUSE_MASK)));
      do_leaf = nleaf >= LEAF_FLUSH || __popc(waiting) >= WAIT_FLUSH || since_flush > 1024;
    }
    if (do_leaf) {
      // the fp32-filtered SAT-17 for queued pairs; a pair the filter leaves open gets the exact fp64 SAT-17 at once
      int cnt = min(32, nleaf);
      bool act = lane < cnt;
      unsigned long long e = act ? M.leafq[nleaf - 1 - lane] : 0ull;
      nleaf -= cnt;
      unsigned owner = (unsigned)e & 31u, ep = (unsigned)(e >> 5) & 0x7ffu;
      unsigned oep = __shfl_sync(FULL, epoch, owner), ofl = __shfl_sync(FULL, fl, owner);
      int otag = __shfl_sync(FULL, tag, owner);
      act = act && (ofl & (F_HAS | F_HIT)) == F_HAS && oep == ep;
      int r = 0;
      if (act) {
        Rigid rel; double qd[4];
        src.exact_pose(P, otag, rel, qd);
        TriPair<double> td;
        load_pair_exact_r(P.env_tris + (size_t)9 * ((unsigned)(e >> 16) & 0xffffffu),
                          P.rob_tris + (size_t)9 * (unsigned)(e >> 40), rel, td);
        TriPair<float> tf;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
          tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
        }
        r = sat17_filter_c(tf);
        n_leaf++;
        if (r == 2) {
          r = sat17_exact_c(td) ? 1 : 0;
          n_fp64++;
        }
      }
      unsigned hm = __reduce_or_sync(FULL, r == 1 ? (1u << owner) : 0u);
      if (hm >> lane & 1u) fl |= F_HIT;
      draining_leaf = nleaf > 0;
      if (nleaf == 0) { fl &= ~F_PL; since_flush = 0; }
      __syncwarp();
    }
  }
  // counters
  unsigned bv = __reduce_add_sync(FULL, n_bv), lf = __reduce_add_sync(FULL, n_leaf), f64 = __reduce_add_sync(FULL, n_fp64);
  if (lane == 0) {
    atomicAdd(&P.counters->bv, (unsigned long long)bv);
    atomicAdd(&P.counters->leaf, (unsigned long long)lf);
    atomicAdd(&P.counters->fp64, (unsigned long long)f64);
    atomicAdd(&P.counters->poses, nposes);
#ifdef OSR_DFS_STATS
    atomicAdd(&P.counters->warp_iters, (unsigned long long)st_iters);
    atomicAdd(&P.counters->lanes_active, (unsigned long long)st_active);
    atomicAdd(&P.counters->lanes_waiting, (unsigned long long)st_wait);
    atomicAdd(&P.counters->lanes_idle, (unsigned long long)st_idle);
#endif
  }
}

}  // namespace osr

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
//     at a time with the fp32-filtered SAT-17, undecided pairs go to a second queue for the fp64 SAT-17, so
//     the expensive exact predicate also runs with full lanes;
//   * the trees are 32-byte QNodes (qnode.h); their first levels (for alpha-1.2: both whole trees, 129 KB)
//     live in shared memory, where a 32-lane gather costs bank conflicts instead of one L1 wavefront per lane.
#pragma once
#include "kernels.cuh"
#include "qnode.h"

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
};

constexpr int DFS_WARPS = 12;
constexpr int DFS_THREADS = DFS_WARPS * 32;
constexpr int LEAFQ_CAP = 128;   // >= LEAF_FLUSH - 1 + 96 (up to three pairs per lane and iteration)
constexpr int REQ_CAP = 64;
constexpr int LEAF_FLUSH = 16;   // drain the triangle-pair queue once it holds this many pairs
constexpr int RE_FLUSH = 8;      // same for the queue of pairs awaiting the exact fp64 predicate
constexpr int WAIT_FLUSH = 4;    // ... or once this many lanes have nothing left but queued pairs
constexpr int REFILL_MIN = 1;    // pull new poses once this many lanes are idle

// work-stack entry = one (environment node, robot node) pair still to be box-tested; bit 31 / 63 marks a
// valid entry so that 0 means "none"
template <typename E> struct TaskCodec;
template <> struct TaskCodec<uint32_t> {  // valid:1 | robot node:15 | environment node:16
  static __device__ __forceinline__ uint32_t make(unsigned n1, unsigned n2) { return 0x80000000u | (n2 << 16) | n1; }
  static __device__ __forceinline__ unsigned env(uint32_t e) { return e & 0xffffu; }
  static __device__ __forceinline__ unsigned rob(uint32_t e) { return (e >> 16) & 0x7fffu; }
};
template <> struct TaskCodec<unsigned long long> {  // valid:1 | robot node:31 | environment node:32
  static __device__ __forceinline__ unsigned long long make(unsigned n1, unsigned n2) {
    return 0x8000000000000000ull | ((unsigned long long)n2 << 32) | n1;
  }
  static __device__ __forceinline__ unsigned env(unsigned long long e) { return (unsigned)e; }
  static __device__ __forceinline__ unsigned rob(unsigned long long e) { return (unsigned)(e >> 32) & 0x7fffffffu; }
};

template <typename E, int DEPTH>
struct __align__(16) DfsWarpMem {
  E stack[DEPTH][32];
  unsigned long long leafq[LEAFQ_CAP];
  unsigned long long req[REQ_CAP];
  float pend[17][32];  // BoxPose of prepared poses: q[4] R[9] t[3] eta
  int pend_idx[32];
  int parts[32];       // work sharing at the end of the input: lanes still working on the pose of home lane i
  int hitf[32];        // ... and whether one of them has found an intersecting pair
};

// leaf-queue entry: owner lane:5 | epoch:11 | env triangle:24 | robot triangle:24
__device__ __forceinline__ unsigned long long leaf_entry(unsigned lane, unsigned epoch, unsigned ta, unsigned tb) {
  return (unsigned long long)lane | ((unsigned long long)epoch << 5) | ((unsigned long long)ta << 16) |
         ((unsigned long long)tb << 40);
}

__device__ __forceinline__ QNode load_qnode(const uint4* sm, const uint4* gl, int nsm, unsigned idx) {
  uint4 a, b;
  if ((int)idx < nsm) { a = sm[2 * idx]; b = sm[2 * idx + 1]; }
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
                                                  const Rigid& rel, Tri6<double>& t) {
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

// ---- pose sources -----------------------------------------------------------------------------------
// A source hands out poses by an integer tag and can rebuild the exact fp64 transform of a tag.
struct ValiditySource {
  const double* q;
  long long n;
  uint8_t* out;
  __device__ __forceinline__ void exact_pose(const DfsParams& P, int tag, Rigid& rel, double* quat) const {
    double s[7];
#pragma unroll
    for (int k = 0; k < 7; k++) s[k] = __ldg(q + 7 * (long long)tag + k);
    Rigid rob;
    state_to_rigid(s, rob);
    relative_rigid(P.env_tf, rob, rel);
    quat[0] = s[3]; quat[1] = s[4]; quat[2] = s[5]; quat[3] = s[6];
  }
  __device__ __forceinline__ void result(int tag, bool hit) const { out[tag] = hit ? 0 : 1; }
};

}  // namespace osr

namespace osr {

// ---------------------------------------------------------------------------------------------------
// K1 (second generation): batched validity, out[i] = 1 iff state i is collision free
// (UnitWorld::isValid, /root/reference/lib/osr/unit_world.cc:215-247).  One persistent block per SM.
// ---------------------------------------------------------------------------------------------------
template <typename E, int DEPTH>
__global__ void __launch_bounds__(DFS_THREADS, 1)
    k_valid_dfs(DfsParams P, const double* __restrict__ q, long long n, uint8_t* __restrict__ out) {
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
  ValiditySource src = {q, n, out};

  // every pose that passes the root box test starts with the root pair already expanded (same descent rule
  // as everywhere else: bvh_collision_traversal_node-inl.h:78-89)
  E root_t0 = 0, root_t1 = 0;
  bool root_leafpair = false;
  {
    QNode a = load_qnode(sm_env, P.env_q, P.env_smem_nodes, 0), b = load_qnode(sm_rob, P.rob_q, P.rob_smem_nodes, 0);
    float ea0, ea1, ea2, eb0, eb1, eb2, sp_;
    unpack_half2(a.e01, ea0, ea1); unpack_half2(a.e2x, ea2, sp_);
    unpack_half2(b.e01, eb0, eb1); unpack_half2(b.e2x, eb2, sp_);
    float sa = fmaf(ea0, ea0, fmaf(ea1, ea1, ea2 * ea2)), sb = fmaf(eb0, eb0, fmaf(eb1, eb1, eb2 * eb2));
    if (a.link < 0 && b.link < 0) root_leafpair = true;
    else if (b.link < 0 || (a.link >= 0 && sa > sb)) { root_t0 = TC::make((unsigned)a.link, 0); root_t1 = TC::make((unsigned)a.link + 1u, 0); }
    else { root_t0 = TC::make(0, (unsigned)b.link); root_t1 = TC::make(0, (unsigned)b.link + 1u); }
  }

  // lane state
  BoxPose bp;
  int tag = 0;
  unsigned epoch = 0;
  bool has_pose = false, hit = false, pl = false, pr = false, lf_root = false;
  E cur = 0;
  int sp = 0, sb = 0;   // the lane's stack is stack[sb .. sp-1][lane]
  int home = lane;      // lane that pulled the pose this lane works on (differs after a steal)
  // warp state (uniform)
  int pend_n = 0, nleaf = 0, nre = 0, since_flush = 0, re_age = 0;
  bool exhausted = false, draining_leaf = false, draining_re = false, steal_mode = false;
  unsigned n_bv = 0, n_leaf = 0, n_fp64 = 0;
  unsigned long long nposes = 0;
  unsigned st_iters = 0, st_active = 0, st_wait = 0, st_idle = 0;

  // One drain step = one chunk of up to 32 queued entries.  Both drains appear exactly once in the code
  // (the kernel is instruction-cache sensitive: 12 warps at different places of a long loop).
  while (true) {
    // (1) completed poses.  Once lanes share poses (steal_mode) a pose is finished by the last of its parts.
    {
      bool done = has_pose && (hit || (cur == 0 && !pl && !pr && !lf_root));
      if (!steal_mode) {
        if (done) src.result(tag, hit);
      } else {
        if (has_pose && M.hitf[home] != 0) { hit = true; done = true; }
        if (done && hit) M.hitf[home] = 1;
        __syncwarp();
        if (done && atomicSub(&M.parts[home], 1) == 1) src.result(tag, M.hitf[home] != 0);
        __syncwarp();
      }
      if (done) { has_pose = false; hit = false; pl = false; pr = false; cur = 0; sp = 0; sb = 0; }
    }
    // (2) hand new poses to idle lanes
    unsigned idle = __ballot_sync(FULL, !has_pose);
    bool any_cur = __any_sync(FULL, cur != 0 || lf_root);
    if (idle != 0 && (__popc(idle) >= REFILL_MIN || !any_cur)) {
      while (pend_n == 0 && !exhausted) {  // prepare the next 32 states with the whole warp
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(P.work_counter, 32ull);
        base = __shfl_sync(FULL, base, 0);
        if ((long long)base >= n) { exhausted = true; break; }
        long long i = (long long)base + lane;
        bool have = i < n, ov = false;
        BoxPose np;
        if (have) {
          Rigid rel; double qd[4];
          src.exact_pose(P, (int)i, rel, qd);
          make_box_pose(rel, P.env_quat, qd, P.r_rob, np);
          QNode a = load_qnode(sm_env, P.env_q, P.env_smem_nodes, 0), b = load_qnode(sm_rob, P.rob_q, P.rob_smem_nodes, 0);
          float s1, s2;
          ov = obb_overlap_q(a, b, np, s1, s2);
          n_bv++;
          if (!ov) src.result((int)i, false);
        }
        nposes += __popc(__ballot_sync(FULL, have));
        unsigned sm = __ballot_sync(FULL, ov);
        if (ov) {
          int slot = pend_n + __popc(sm & ltmask);
#pragma unroll
          for (int k = 0; k < 4; k++) M.pend[k][slot] = np.q[k];
#pragma unroll
          for (int k = 0; k < 9; k++) M.pend[4 + k][slot] = np.R[k];
#pragma unroll
          for (int k = 0; k < 3; k++) M.pend[13 + k][slot] = np.t[k];
          M.pend[16][slot] = np.eta;
          M.pend_idx[slot] = (int)i;
        }
        pend_n += __popc(sm);
        __syncwarp();
      }
      if (pend_n > 0) {
        int k = __popc(idle & ltmask);
        if (!has_pose && k < pend_n) {
          int slot = pend_n - 1 - k;
#pragma unroll
          for (int j = 0; j < 4; j++) bp.q[j] = M.pend[j][slot];
#pragma unroll
          for (int j = 0; j < 9; j++) bp.R[j] = M.pend[4 + j][slot];
#pragma unroll
          for (int j = 0; j < 3; j++) bp.t[j] = M.pend[13 + j][slot];
          bp.eta = M.pend[16][slot];
          tag = M.pend_idx[slot];
          has_pose = true; hit = false; pl = false; pr = false; sp = 0; sb = 0;
          epoch = (epoch + 1) & 0x7ffu;
          cur = root_t0;
          if (root_t1 != 0) { M.stack[0][lane] = root_t1; sp = 1; }
          lf_root = root_leafpair;
        }
        pend_n -= min(pend_n, __popc(idle));
        __syncwarp();
        any_cur = __any_sync(FULL, cur != 0 || lf_root);
      }
    }
    // (2b) input exhausted and lanes idle: idle lanes take the OLDEST stack entry (the largest pending subtree)
    // of lanes that still have one.  Without this the kernel cannot end before its longest single pose does
    // (up to ~2000 box tests in sequence on alpha-1.2), and every warp idles through its own longest pose.
    if (exhausted && pend_n == 0) {
      unsigned thieves = __ballot_sync(FULL, !has_pose);
      if (thieves != 0) {
        if (!steal_mode) {
          M.parts[lane] = has_pose ? 1 : 0;
          M.hitf[lane] = 0;
          home = lane;
          steal_mode = true;
          __syncwarp();
        }
        unsigned victims = __ballot_sync(FULL, has_pose && sp > sb);
        int ns = min(__popc(thieves), __popc(victims));
        if (ns > 0) {
          int r = __popc(thieves & ltmask);
          bool thief = !has_pose && r < ns;
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
          bool robbed = has_pose && sp > sb && __popc(victims & ltmask) < ns;
          E stolen = 0;
          if (thief) stolen = M.stack[vsb][v];
          __syncwarp();
          if (robbed) sb++;
          if (thief) {
            bp = vb; tag = vtag; home = vhome;
            cur = stolen; sp = 0; sb = 0;
            has_pose = true; hit = false; pl = false; pr = false; lf_root = false;
            epoch = (epoch + 1) & 0x7ffu;
            atomicAdd(&M.parts[home], 1);
          }
          __syncwarp();
          any_cur = true;
        }
      }
    }
    // (3) one box test per lane: the pair on top of the lane's stack; an overlapping pair is replaced by its
    // two child pairs (not while a queue is being drained: the queues' capacity assumes at most one iteration
    // of new triangle pairs on top)
    if (any_cur && !draining_leaf && !draining_re) {
      bool lf0 = false;
      unsigned ta0 = 0, tb0 = 0;
      st_iters++;
      st_active += __popc(__ballot_sync(FULL, cur != 0));
      st_wait += __popc(__ballot_sync(FULL, has_pose && cur == 0));
      st_idle += __popc(__ballot_sync(FULL, !has_pose));
      if (cur != 0) {
        unsigned n1 = TC::env(cur), n2 = TC::rob(cur);
        QNode A = load_qnode(sm_env, P.env_q, P.env_smem_nodes, n1);
        QNode B = load_qnode(sm_rob, P.rob_q, P.rob_smem_nodes, n2);
        float sa, sb;
        bool ov = obb_overlap_q(A, B, bp, sa, sb);
        n_bv++;
        E nxt = 0;
        if (ov) {
          if (A.link < 0 && B.link < 0) { lf0 = true; ta0 = (unsigned)(-(A.link + 1)); tb0 = (unsigned)(-(B.link + 1)); }
          else if (B.link < 0 || (A.link >= 0 && sa > sb)) {
            nxt = TC::make((unsigned)A.link, n2);
            M.stack[sp][lane] = TC::make((unsigned)A.link + 1u, n2);
            sp++;
          } else {
            nxt = TC::make(n1, (unsigned)B.link);
            M.stack[sp][lane] = TC::make(n1, (unsigned)B.link + 1u);
            sp++;
          }
        }
        if (nxt == 0 && sp > sb) { sp--; nxt = M.stack[sp][lane]; }
        cur = nxt;
      }
      // queue the triangle pairs found
      unsigned m0 = __ballot_sync(FULL, lf0), m2 = __ballot_sync(FULL, lf_root);
      if (m0 | m2) {
        if (lf0) M.leafq[nleaf + __popc(m0 & ltmask)] = leaf_entry(lane, epoch, ta0, tb0);
        if (lf_root) M.leafq[nleaf + __popc(m0) + __popc(m2 & ltmask)] = leaf_entry(lane, epoch, 0, 0);
        nleaf += __popc(m0) + __popc(m2);
        pl = pl || lf0 || lf_root;
        lf_root = false;
        __syncwarp();
      }
      since_flush++;
      re_age = nre > 0 ? re_age + 1 : 0;
    }
    // (4) drain policy: which queue (if any) gets one chunk processed now
    bool do_leaf = false, do_re = false;
    if (draining_leaf) do_leaf = true;
    else if (draining_re) do_re = true;
    else if (!any_cur) {
      if (nleaf > 0) do_leaf = true;
      else if (nre > 0) do_re = true;
      else if (!__any_sync(FULL, has_pose) && exhausted && pend_n == 0) break;
    } else {
      if (nleaf > 0) {
        unsigned waiting = __ballot_sync(FULL, has_pose && cur == 0 && pl);
        do_leaf = nleaf >= LEAF_FLUSH || __popc(waiting) >= WAIT_FLUSH || since_flush > 1024;
      }
      if (!do_leaf && nre > 0) {
        unsigned waiting = __ballot_sync(FULL, has_pose && cur == 0 && !pl && pr);
        do_re = nre >= RE_FLUSH || __popc(waiting) >= WAIT_FLUSH || re_age > 1024;
      }
    }
    if (do_leaf && nre > REQ_CAP - 32) { do_leaf = false; do_re = true; draining_leaf = true; }  // make room first
    if (do_leaf | do_re) {
      // one chunk of the chosen queue: the fp32-filtered SAT-17 for queued pairs, the exact fp64 SAT-17 for
      // the pairs the filter could not decide
      unsigned long long* Q = do_leaf ? M.leafq : M.req;
      int& cntq = do_leaf ? nleaf : nre;
      int cnt = min(32, cntq);
      bool act = lane < cnt;
      unsigned long long e = act ? Q[cntq - 1 - lane] : 0ull;
      cntq -= cnt;
      unsigned owner = (unsigned)e & 31u, ep = (unsigned)(e >> 5) & 0x7ffu;
      unsigned oep = __shfl_sync(FULL, epoch, owner);
      int ohit = __shfl_sync(FULL, (int)hit, owner), ohas = __shfl_sync(FULL, (int)has_pose, owner);
      int otag = __shfl_sync(FULL, tag, owner);
      act = act && ohas && oep == ep && !ohit;
      int r = 0;
      if (act) {
        Rigid rel; double qd[4];
        src.exact_pose(P, otag, rel, qd);
        Tri6<double> td;
        load_pair_exact_r(P.env_tris + (size_t)9 * ((unsigned)(e >> 16) & 0xffffffu),
                          P.rob_tris + (size_t)9 * (unsigned)(e >> 40), rel, td);
        if (do_leaf) {
          Tri6<float> tf;
#pragma unroll
          for (int k = 0; k < 3; k++) {
            tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
            tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
          }
          r = sat17_filter(tf);
          n_leaf++;
        } else {
          r = sat17_exact(td) ? 1 : 0;
          n_fp64++;
        }
      }
      unsigned hm = __reduce_or_sync(FULL, r == 1 ? (1u << owner) : 0u);
      if (hm >> lane & 1u) hit = true;
      if (do_leaf) {
        unsigned um = __ballot_sync(FULL, r == 2);
        unsigned uo = __reduce_or_sync(FULL, r == 2 ? (1u << owner) : 0u);
        if (r == 2) M.req[nre + __popc(um & ltmask)] = e;
        nre += __popc(um);
        if (uo >> lane & 1u) pr = true;
        draining_leaf = nleaf > 0;
        if (nleaf == 0) { pl = false; since_flush = 0; }
      } else {
        draining_re = nre > 0;
        if (nre == 0) pr = false;
      }
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
    atomicAdd(&P.counters->warp_iters, (unsigned long long)st_iters);
    atomicAdd(&P.counters->lanes_active, (unsigned long long)st_active);
    atomicAdd(&P.counters->lanes_waiting, (unsigned long long)st_wait);
    atomicAdd(&P.counters->lanes_idle, (unsigned long long)st_idle);
  }
}

}  // namespace osr

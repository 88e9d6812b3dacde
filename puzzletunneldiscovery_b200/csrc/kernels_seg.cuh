// kernels_seg.cuh -- K2, second generation: segment walk / first contact on the lane-owned depth-first
// traversal of kernels_dfs.cuh (sm_100a).
//
// Reference semantics (UnitWorld::transitStateToWithContact, /root/reference/lib/osr/unit_world.cc:405-474;
// isValidTransition :477-486): walk the segment at
//     delta_1 = vd;  delta_{k+1} = delta_k < dist ? min(dist, delta_k + vd) : delta_k + vd;  tau_k = delta_k / dist
// and return the FIRST step whose interpolated state collides.  Validity along a segment is not monotone, so
// every step before the first hit has to be evaluated; steps are independent of each other though, which is
// what this kernel uses:
//   * a WARP owns one segment at a time, a LANE owns one step of it and runs the same private depth-first
//     (environment node, robot node) traversal as K1 -- near-contact steps (hundreds of box tests) and free
//     steps (a handful) no longer wait for each other as they did in the 32-step tiles of the first
//     generation, and a colliding step stops at its first intersecting triangle pair;
//   * steps are prepared 32 at a time by the whole warp (the running fp64 delta sum reproduced add by add,
//     Eigen's slerp, the fp32 box pose) and handed to idle lanes IN ORDER; the first hit found cancels every
//     later step in flight, the segment ends when all earlier steps have finished;
//   * triangle pairs go through the same two queues (fp32-filtered SAT-17, then exact fp64 SAT-17) as in K1,
//     with the exact pose taken from the lane's interpolated state kept in shared memory.
#pragma once
#include "kernels_dfs.cuh"

namespace osr {

template <typename E, int DEPTH>
struct __align__(16) SegWarpMem {
  E stack[DEPTH][32];
  unsigned long long leafq[LEAFQ_CAP];
  unsigned long long req[REQ_CAP];
  float pend[17][32];        // BoxPose of the prepared steps: q[4] R[9] t[3] eta
  double pend_state[7][32];  // their interpolated states
  double cur_state[7][32];   // state of the step every lane works on (exact pose of the triangle tests)
  double pend_tau[33];       // [0] tau of the step before the prepared batch, [1 + j] tau of prepared step j
  double lane_tau[2][32];    // tau of each lane's step and of the step before it
  double ab[14];             // from, to
  double hit_tau, hit_ptau;  // taus recorded with the earliest hit so far
  double last_tau;           // tau of the last step of the walk that was prepared
};

struct SegArgs {
  const double* from;
  const double* to;
  long long n;
  const double* vd_arr;
  long long vd_count;
  double vd_scalar;
  int transition_mode;  // vd := min(dist/2, vd) (isValidTransition)
  SegOut O;
};

// Results of one segment (all lanes compute the same values; lane 0 / lanes 0..6 store them).
__device__ __forceinline__ void seg_store(const SegArgs& A, long long seg, const double* ab, bool free_flag,
                                       double free_tau, double touch_tau, long long steps, int mode, int lane) {
  // mode 0: lf = from, fh = to (vd >= dist);  1: hit at the first step (lf = from);  2: hit later;  3: all free;
  // 4: lf = fh = from (NaN distance / verify_delta: the reference's loop body never runs)
  double lf[7], fh[7];
  if (mode == 0 || mode == 4) {
#pragma unroll
    for (int k = 0; k < 7; k++) { lf[k] = ab[k]; fh[k] = ab[(mode == 0 ? 7 : 0) + k]; }
  } else if (mode == 3) {
    interpolate(ab, ab + 7, free_tau, lf);
#pragma unroll
    for (int k = 0; k < 7; k++) fh[k] = lf[k];
  } else {
    interpolate(ab, ab + 7, touch_tau, fh);
    if (mode == 2) interpolate(ab, ab + 7, free_tau, lf);
    else {
#pragma unroll
      for (int k = 0; k < 7; k++) lf[k] = ab[k];
    }
  }
  if (lane == 0) {
    if (A.O.is_free) A.O.is_free[seg] = free_flag ? 1 : 0;
    if (A.O.free_tau) A.O.free_tau[seg] = free_tau;
    if (A.O.touch_tau) A.O.touch_tau[seg] = touch_tau;
    if (A.O.steps) A.O.steps[seg] = steps;
  }
  if (lane < 7) {
    if (A.O.last_free) A.O.last_free[7 * seg + lane] = lf[lane];
    if (A.O.first_hit) A.O.first_hit[7 * seg + lane] = fh[lane];
  }
}


// Prepare the next (up to) 32 steps of the warp's segment: lane j takes the (j+1)-th delta after `carry`.
// Returns the number of steps prepared (a prefix of the lanes); fewer than 32 means the walk ends with them.
template <typename WM>
__device__ __forceinline__ int seg_prepare(const DfsParams& P, WM& M, double vd, double dist, double& carry, bool first,
                                        int lane) {
  double d = carry, mine = 0.0;
  bool mine_ok = false, ok = true;
#pragma unroll 1
  for (int j = 0; j < 32; j++) {
    if (first && j == 0) d = vd;
    else if (d < dist) d = fmin(dist, d + vd);
    else d = d + vd;
    ok = ok && (d <= dist);
    if (j == lane) { mine = d; mine_ok = ok; }
  }
  carry = d;  // delta of slot 31 (only used when all 32 slots were valid)
  const double tau = mine * (1.0 / dist);
  double prev_last = M.pend_tau[32];
  __syncwarp();
  if (lane == 0 && !first) M.pend_tau[0] = prev_last;
  if (mine_ok) {
    double s[7];
    interpolate(M.ab, M.ab + 7, tau, s);
    Rigid rob, rel;
    state_to_rigid(s, rob);
    relative_rigid(P.env_tf, rob, rel);
    BoxPose np;
    make_box_pose(rel, P.env_quat, s + 3, P.r_rob, np);
#pragma unroll
    for (int k = 0; k < 4; k++) M.pend[k][lane] = np.q[k];
#pragma unroll
    for (int k = 0; k < 9; k++) M.pend[4 + k][lane] = np.R[k];
#pragma unroll
    for (int k = 0; k < 3; k++) M.pend[13 + k][lane] = np.t[k];
    M.pend[16][lane] = np.eta;
#pragma unroll
    for (int k = 0; k < 7; k++) M.pend_state[k][lane] = s[k];
    M.pend_tau[1 + lane] = tau;
  }
  unsigned alive = __ballot_sync(FULL, mine_ok);
  int cnt = __popc(alive);
  if (mine_ok && lane == cnt - 1) M.last_tau = tau;
  __syncwarp();
  return cnt;
}


constexpr unsigned NO_HIT = 0xffffffffu;

// ---------------------------------------------------------------------------------------------------
// K2 (second generation).  One persistent block per SM, WARPS warps; see the file header.
// ---------------------------------------------------------------------------------------------------
template <typename E, int DEPTH, bool ALL_SMEM, int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) k_transit_dfs(DfsParams P, SegArgs A) {
  typedef TaskCodec<E> TC;
  typedef SegWarpMem<E, DEPTH> WM;
  extern __shared__ uint4 dfs_smem[];
  uint4* sm_env = dfs_smem;
  uint4* sm_rob = dfs_smem + 2 * (size_t)P.env_smem_nodes;
  WM& M = reinterpret_cast<WM*>(dfs_smem + 2 * (size_t)(P.env_smem_nodes + P.rob_smem_nodes))[threadIdx.x >> 5];
  for (int i = threadIdx.x; i < 2 * P.env_smem_nodes; i += blockDim.x) sm_env[i] = __ldg(P.env_q + i);
  for (int i = threadIdx.x; i < 2 * P.rob_smem_nodes; i += blockDim.x) sm_rob[i] = __ldg(P.rob_q + i);
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const unsigned ltmask = (1u << lane) - 1u;
  const E root_task = root_entry<E>(P.env_q, P.rob_q);
  const bool root_leafpair = root_task == 0;

  // lane state
  BoxPose bp;
  unsigned kstep = 0, epoch = 0, fl = 0;
  E cur = 0;
  int sp = 0;
  // warp state (uniform)
  bool have_seg = false, walk_done = false, first = true;
  long long seg = 0;
  double vd = 0.0, dist = 0.0, carry = 0.0;
  unsigned k_next = 0, pend_base = 0, hit_k = NO_HIT;
  int pend_n = 0, pend_taken = 0, nleaf = 0, nre = 0, since_flush = 0, re_age = 0;
  bool draining_leaf = false, draining_re = false;
  unsigned n_bv = 0, n_leaf = 0, n_fp64 = 0;
  unsigned long long nposes = 0;

  while (true) {
    // (0) next segment
    if (!have_seg) {
      unsigned long long sg = 0;
      if (lane == 0) sg = atomicAdd(P.work_counter, 1ull);
      sg = __shfl_sync(FULL, sg, 0);
      if ((long long)sg >= A.n) break;
      seg = (long long)sg;
      __syncwarp();
      if (lane < 7) M.ab[lane] = __ldg(A.from + 7 * seg + lane);
      else if (lane < 14) M.ab[lane] = __ldg(A.to + 7 * seg + (lane - 7));
      __syncwarp();
      vd = A.vd_arr ? __ldg(A.vd_arr + (A.vd_count > 1 ? seg : 0)) : A.vd_scalar;
      dist = distance(M.ab, M.ab + 7);
      if (A.transition_mode) vd = fmin(dist / 2.0, vd);
      if (vd >= dist) {  // unit_world.cc:445-447
        seg_store(A, seg, M.ab, false, 0.0, 1.0, 0, 0, lane);
        continue;
      }
      if (!(dist == dist) || !(vd == vd) || !(vd > 0.0)) {   // see k_transit: NaN walk / verify_delta <= 0
        if (dist == dist && vd == vd && lane == 0) atomicExch(P.error_flag, 3);
        seg_store(A, seg, M.ab, false, 0.0, 0.0, 0, 4, lane);
        continue;
      }
      have_seg = true; walk_done = false; first = true;
      carry = 0.0; k_next = 0; pend_base = 0; pend_n = 0; pend_taken = 0; hit_k = NO_HIT;
      nleaf = 0; nre = 0; since_flush = 0; re_age = 0; draining_leaf = false; draining_re = false;
      if (lane == 0) { M.pend_tau[0] = 0.0; M.last_tau = 0.0; }
      __syncwarp();
    }
    // (1) finished steps; the earliest hit so far cancels every later step
    {
      bool done = (fl & F_HAS) && ((fl & F_HIT) || (cur == 0 && !(fl & (F_PL | F_PR))));
      bool hit_now = done && (fl & F_HIT);
      unsigned kmin = __reduce_min_sync(FULL, hit_now ? kstep : NO_HIT);
      kmin = __shfl_sync(FULL, kmin, 0);  // tells the compiler the value is warp-uniform (it steers the loop)
      if (kmin < hit_k) {
        hit_k = kmin;
        if (hit_now && kstep == kmin) { M.hit_tau = M.lane_tau[0][lane]; M.hit_ptau = M.lane_tau[1][lane]; }
        pend_n = 0;
        walk_done = true;
        __syncwarp();
      }
      if (done || ((fl & F_HAS) && kstep > hit_k)) { fl = 0; cur = 0; sp = 0; }
    }
    // (2) idle lanes take the next steps, in order
    unsigned idle = __ballot_sync(FULL, !(fl & F_HAS));
    if (idle != 0) {
      if (pend_n == 0 && !walk_done) {
        int cnt = seg_prepare(P, M, vd, dist, carry, first, lane);
        first = false;
        pend_base = k_next; k_next += cnt; pend_n = cnt; pend_taken = 0;
        if (cnt < 32) walk_done = true;
      }
      if (pend_n > 0) {
        int k = __popc(idle & ltmask);
        bool pulled = !(fl & F_HAS) && k < pend_n;
        if (pulled) {
          int slot = pend_taken + k;
#pragma unroll
          for (int j = 0; j < 4; j++) bp.q[j] = M.pend[j][slot];
#pragma unroll
          for (int j = 0; j < 9; j++) bp.R[j] = M.pend[4 + j][slot];
#pragma unroll
          for (int j = 0; j < 3; j++) bp.t[j] = M.pend[13 + j][slot];
          bp.eta = M.pend[16][slot];
#pragma unroll
          for (int j = 0; j < 7; j++) M.cur_state[j][lane] = M.pend_state[j][slot];
          M.lane_tau[0][lane] = M.pend_tau[1 + slot];
          M.lane_tau[1][lane] = M.pend_tau[slot];
          kstep = pend_base + (unsigned)slot;
          fl = F_HAS; sp = 0;
          epoch = (epoch + 1) & 0x7ffu;
          // the root box test (the traversal only holds pairs known to overlap): a step whose root boxes are
          // apart is a finished free step
          cur = 0;
          if (!root_leafpair) {
            const QNode RA = load_qnode<ALL_SMEM>(sm_env, P.env_q, P.env_smem_nodes, 0u);
            const QNode RB = load_qnode<ALL_SMEM>(sm_rob, P.rob_q, P.rob_smem_nodes, 0u);
            float sl, va, vb;
            if (obb_overlap_q2(RA, RB, bp, sl, va, vb)) cur = root_task;
            n_bv++;
          }
        }
        int taken = min(pend_n, __popc(idle));
        pend_taken += taken; pend_n -= taken;
        if (root_leafpair) {
          unsigned m2 = __ballot_sync(FULL, pulled);
          if (pulled) { M.leafq[nleaf + __popc(m2 & ltmask)] = leaf_entry(lane, epoch, 0, 0); fl |= F_PL; }
          nleaf += __popc(m2);
        }
        __syncwarp();
      }
    }
    // (2b) segment finished: no step in flight, none left to hand out
    if (walk_done && pend_n == 0 && !__any_sync(FULL, (fl & F_HAS) != 0)) {
      if (hit_k != NO_HIT) {
        nposes += (unsigned long long)hit_k + 1ull;
        seg_store(A, seg, M.ab, false, M.hit_ptau, M.hit_tau, (long long)hit_k + 1, hit_k > 0 ? 2 : 1, lane);
      } else {
        nposes += k_next;
        seg_store(A, seg, M.ab, true, M.last_tau, 0.0, (long long)k_next, 3, lane);
      }
      have_seg = false;
      continue;
    }
    // (3) one box test per lane: the pair on top of the lane's stack (as in K1)
    const bool any_cur = __any_sync(FULL, cur != 0);
    if (any_cur && !draining_leaf && !draining_re) {
      bool lf0 = false, lf1 = false;
      unsigned ta0 = 0, tb0 = 0, ta1 = 0, tb1 = 0;
      if (cur != 0) {
        dfs_expand_step<E, ALL_SMEM>(cur, sp, 0, M.stack, lane, sm_env, sm_rob, P, bp, lf0, ta0, tb0, lf1, ta1, tb1);
        n_bv += 2;
      }
      const unsigned m0 = __ballot_sync(FULL, lf0), m1 = __ballot_sync(FULL, lf1);
      if (m0 | m1) {
        if (lf0) M.leafq[nleaf + __popc(m0 & ltmask)] = leaf_entry(lane, epoch, ta0, tb0);
        if (lf1) M.leafq[nleaf + __popc(m0) + __popc(m1 & ltmask)] = leaf_entry(lane, epoch, ta1, tb1);
        if (lf0 | lf1) fl |= F_PL;
        nleaf += __popc(m0) + __popc(m1);
        __syncwarp();
      }
      since_flush++;
      re_age = nre > 0 ? re_age + 1 : 0;
    }
    // (4) drain policy (as in K1)
    bool do_leaf = false, do_re = false;
    if (draining_leaf) do_leaf = true;
    else if (draining_re) do_re = true;
    else if (!any_cur) {
      if (nleaf > 0) do_leaf = true;
      else if (nre > 0) do_re = true;
    } else {
      if (nleaf > 0) {
        unsigned waiting = __ballot_sync(FULL, (fl & (F_HAS | F_PL)) == (F_HAS | F_PL) && cur == 0);
        do_leaf = nleaf >= LEAF_FLUSH || __popc(waiting) >= WAIT_FLUSH || since_flush > 1024;
      }
      if (!do_leaf && nre > 0) {
        unsigned waiting = __ballot_sync(FULL, (fl & (F_HAS | F_PL | F_PR)) == (F_HAS | F_PR) && cur == 0);
        do_re = nre >= RE_FLUSH || __popc(waiting) >= WAIT_FLUSH || re_age > 1024;
      }
    }
    if (do_leaf && nre > REQ_CAP - 32) { do_leaf = false; do_re = true; draining_leaf = true; }  // make room first
    if (do_leaf | do_re) {
      unsigned long long* Q = do_leaf ? M.leafq : M.req;
      int& cntq = do_leaf ? nleaf : nre;
      int cnt = min(32, cntq);
      bool act = lane < cnt;
      unsigned long long e = act ? Q[cntq - 1 - lane] : 0ull;
      cntq -= cnt;
      unsigned owner = (unsigned)e & 31u, ep = (unsigned)(e >> 5) & 0x7ffu;
      unsigned oep = __shfl_sync(FULL, epoch, owner), ofl = __shfl_sync(FULL, fl, owner);
      act = act && (ofl & (F_HAS | F_HIT)) == F_HAS && oep == ep;
      int r = 0;
      if (act) {
        double st[7];
#pragma unroll
        for (int k = 0; k < 7; k++) st[k] = M.cur_state[k][owner];
        Rigid rob, rel;
        state_to_rigid(st, rob);
        relative_rigid(P.env_tf, rob, rel);
        TriPair<double> td;
        load_pair_exact_r(P.env_tris + (size_t)9 * ((unsigned)(e >> 16) & 0xffffffu),
                          P.rob_tris + (size_t)9 * (unsigned)(e >> 40), rel, td);
        if (do_leaf) {
          TriPair<float> tf;
#pragma unroll
          for (int k = 0; k < 3; k++) {
            tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
            tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
          }
          r = sat17_filter_c(tf);
          n_leaf++;
        } else {
          r = sat17_exact_c(td) ? 1 : 0;
          n_fp64++;
        }
      }
      unsigned hm = __reduce_or_sync(FULL, r == 1 ? (1u << owner) : 0u);
      if (hm >> lane & 1u) fl |= F_HIT;
      if (do_leaf) {
        unsigned um = __ballot_sync(FULL, r == 2);
        unsigned uo = __reduce_or_sync(FULL, r == 2 ? (1u << owner) : 0u);
        if (r == 2) M.req[nre + __popc(um & ltmask)] = e;
        nre += __popc(um);
        if (uo >> lane & 1u) fl |= F_PR;
        draining_leaf = nleaf > 0;
        if (nleaf == 0) { fl &= ~F_PL; since_flush = 0; }
      } else {
        draining_re = nre > 0;
        if (nre == 0) fl &= ~F_PR;
      }
      __syncwarp();
    }
  }
  unsigned bv = __reduce_add_sync(FULL, n_bv), lf = __reduce_add_sync(FULL, n_leaf), f64 = __reduce_add_sync(FULL, n_fp64);
  if (lane == 0) {
    atomicAdd(&P.counters->bv, (unsigned long long)bv);
    atomicAdd(&P.counters->leaf, (unsigned long long)lf);
    atomicAdd(&P.counters->fp64, (unsigned long long)f64);
    atomicAdd(&P.counters->poses, nposes);
  }
}

}  // namespace osr

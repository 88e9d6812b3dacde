// kernels_contacts.cuh -- K3c: the contact records of k_contacts put into the reference's row order ON THE DEVICE.
//
// k_contacts appends (state, env triangle, robot triangle, FCL row multiplicity) records in whatever order its warps
// find them.  The rows of UnitWorld::intersectingSegments (/root/reference/lib/osr/unit_world.cc:959-1012) are per
// state; inside a state this engine emits them ordered by (env triangle, robot triangle) -- FCL's own order is the
// DFS order of ITS tree, which no caller can rely on -- each record repeated `mult` times (the 0..2 contacts FCL
// reports for the pair, cdmodel.cc:213-227).  Round 1 sorted the records on the host (std::sort of ~1e6 16-byte
// records: 100 ms, three quarters of the whole call).  Here: count per state, exclusive scan (host, n integers),
// scatter into per-state buckets, then one warp per state ranks its bucket and lays out the rows.
#pragma once
#include "kernels.cuh"

namespace osr {

__global__ void k_rec_count(const ContactRec* __restrict__ recs, long long nrec, int* __restrict__ cnt,
                            int* __restrict__ msum) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrec) return;
  const ContactRec rec = recs[r];
  atomicAdd(cnt + rec.state, 1);
  if (rec.mult) atomicAdd(msum + rec.state, rec.mult);
}

__global__ void k_rec_scatter(const ContactRec* __restrict__ recs, long long nrec, const long long* __restrict__ rec_off,
                              int* __restrict__ cursor, ContactRec* __restrict__ out) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= nrec) return;
  const ContactRec rec = recs[r];
  out[rec_off[rec.state] + atomicAdd(cursor + rec.state, 1)] = rec;
}

// One warp per state: rank the records of the state's bucket by (env triangle, robot triangle) -- the pairs of a
// state are distinct, so ranks are a permutation -- and write them to `out` in that order; then expand the row list:
// row_rec[row_off[state] + j] = index (in `out`) of the record the j-th row of the state belongs to, rows beyond the
// state's budget row_off[state + 1] - row_off[state] (output capacity, decided on the host) dropped.
constexpr int REC_ORDER_WARP_MAX = 2048;   // larger buckets (deep penetrations) are ranked by a whole block

// the row list of one state, laid out by one warp from the ordered records
__device__ __forceinline__ void rec_rows(const ContactRec* out, long long base, int c, long long r0, int budget,
                                         int* __restrict__ row_rec, int lane) {
  int running = 0;
  for (int i0 = 0; i0 < c && running < budget; i0 += 32) {
    const int i = i0 + lane;
    const int m = i < c ? out[base + i].mult : 0;
    int incl = m;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int v = __shfl_up_sync(FULL, incl, d);
      if (lane >= d) incl += v;
    }
    const int start = running + incl - m;
    for (int k = 0; k < m; k++)
      if (start + k < budget) row_rec[r0 + start + k] = (int)(base + i);
    running += __shfl_sync(FULL, incl, 31);
  }
}

// The same for the few states whose bucket is large: one block per listed state, keys staged through shared memory.
__global__ void __launch_bounds__(1024)
    k_rec_order_big(const ContactRec* __restrict__ in, ContactRec* out, const long long* __restrict__ rec_off,
                    const int* __restrict__ big_states, const long long* __restrict__ row_off, int* __restrict__ row_rec) {
  __shared__ unsigned long long keys[1024];
  const long long st = big_states[blockIdx.x];
  const long long base = rec_off[st];
  const int c = (int)(rec_off[st + 1] - base);
  for (int i0 = 0; i0 < c; i0 += 1024) {
    const int i = i0 + (int)threadIdx.x;
    ContactRec me = {0, 0, 0, 0};
    if (i < c) me = in[base + i];
    const unsigned long long key = ((unsigned long long)(unsigned)me.env_tri << 32) | (unsigned)me.rob_tri;
    int rank = 0;
    for (int j0 = 0; j0 < c; j0 += 1024) {
      __syncthreads();
      const int j = j0 + (int)threadIdx.x;
      if (j < c) {
        const ContactRec o = in[base + j];
        keys[threadIdx.x] = ((unsigned long long)(unsigned)o.env_tri << 32) | (unsigned)o.rob_tri;
      }
      __syncthreads();
      const int cnt = min(1024, c - j0);
      for (int t = 0; t < cnt; t++) rank += keys[t] < key ? 1 : 0;
    }
    if (i < c) out[base + rank] = me;
  }
  __threadfence_block();
  __syncthreads();
  if (threadIdx.x < 32)
    rec_rows(out, base, c, row_off[st], (int)(row_off[st + 1] - row_off[st]), row_rec, (int)threadIdx.x);
}

__global__ void __launch_bounds__(128)
    k_rec_order(const ContactRec* __restrict__ in, ContactRec* out, const long long* __restrict__ rec_off,
                long long nstates, const long long* __restrict__ row_off, int* __restrict__ row_rec, int warp_max) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long st = warp; st < nstates; st += nwarps) {
    const long long base = rec_off[st];
    const int c = (int)(rec_off[st + 1] - base);
    if (c == 0 || c > warp_max) continue;   // (larger buckets: k_rec_order_big)
    for (int i = lane; i < c; i += 32) {
      const ContactRec me = in[base + i];
      const unsigned long long key = ((unsigned long long)(unsigned)me.env_tri << 32) | (unsigned)me.rob_tri;
      int rank = 0;
      for (int j = 0; j < c; j++) {
        const ContactRec o = in[base + j];
        rank += ((((unsigned long long)(unsigned)o.env_tri << 32) | (unsigned)o.rob_tri) < key) ? 1 : 0;
      }
      out[base + rank] = me;
    }
    __syncwarp();
    rec_rows(out, base, c, row_off[st], (int)(row_off[st + 1] - row_off[st]), row_rec, lane);
    __syncwarp();
  }
}

}  // namespace osr

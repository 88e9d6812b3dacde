// engine.cu -- libosrgpu.so: world handle, device management and the C ABI of include/osrgpu.h.
// Host-side replacement of osr::UnitWorld's query methods (/root/reference/lib/osr/unit_world.cc);
// every query runs on the GPU, there is no CPU fallback.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <thread>
#include <cmath>
#include <condition_variable>
#include <deque>
#include <functional>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <type_traits>
#include <vector>
#if defined(__x86_64__)
#include <emmintrin.h>
#include <sched.h>
#include <sys/mman.h>
#include <map>
#endif

#include "../../include/osrgpu.h"
#include "geometry.h"
#include "kernels.cuh"
#include "kernels_dfs.cuh"
#include "kernels_seg.cuh"
#include "kernels_surface.cuh"
#include "kernels_knn.cuh"
#include "kernels_raster.cuh"
#include "kernels_clearance.cuh"
#include "kernels_keyconf.cuh"
#include "kernels_contacts.cuh"

// Scratch of the allocation-free entry point osr_is_valid_transitions_device_noalloc (end of this file): the work-queue
// word of K2 and its per-warp traversal-stack spill areas.  File scope, in every GPU's memory that loads the library
// until the process exits: 256 B + 2048 warps x 1024 entries x 8 B = 16 MiB + 256 B.  In HOST code this name is a
// host-side stand-in, not a device address: init_slot asks cudaGetSymbolAddress for the address on each device.
#define OSR_NOALLOC_HEADER_BYTES 256
#define OSR_NOALLOC_MAX_WARPS 2048
#define OSR_NOALLOC_SPILL_CAP 1024
#define OSR_NOALLOC_SCRATCH_BYTES \
  ((size_t)OSR_NOALLOC_HEADER_BYTES + (size_t)OSR_NOALLOC_MAX_WARPS * OSR_NOALLOC_SPILL_CAP * sizeof(unsigned long long))
static __device__ __align__(256) unsigned char g_noalloc_scratch[OSR_NOALLOC_SCRATCH_BYTES];

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) {
  g_err = msg;
  return code;
}
#define CU(call)                                                                                         \
  do {                                                                                                   \
    cudaError_t e_ = (call);                                                                             \
    if (e_ != cudaSuccess)                                                                               \
      return fail(OSR_ECUDA, std::string(#call) + ": " + cudaGetErrorString(e_) + " @" + __FILE__ + ":" + \
                                 std::to_string(__LINE__));                                              \
  } while (0)

struct DevBuf {  // grow-only device buffer
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t need(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 4 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct Slot {  // one GPU
  int dev = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t stream2 = nullptr;       // copy stream of host calls (second lane of the chunked pipeline)
  cudaStream_t stream3 = nullptr;       // stream-fed launches: advances `ready` behind every uploaded chunk
  cudaEvent_t feed_ev[64] = {};         // one per upload chunk (created on first use)
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, evp = nullptr, evq = nullptr;
  cudaEvent_t trace_ev[3] = {nullptr, nullptr, nullptr};   // OSR_FEED_TRACE (development)
  long long* ready = nullptr;           // device word: states uploaded so far (stream-fed validity launches)
  long long* h_ready = nullptr;         // pinned source values of the copies that advance it
  char* stage = nullptr;                // pinned staging ring of host calls on pageable buffers
  size_t stage_cap = 0;
  float4* env_nodes = nullptr; double* env_tris = nullptr;
  float4* rob_nodes = nullptr; double* rob_tris = nullptr;
  uint4* env_qnodes = nullptr; uint4* rob_qnodes = nullptr;  // 32-byte QNodes (qnode.h)
  int dfs_variant = -1;        // 0: <u32,32>  1: <u32,64>  2: <u64,48>  -1: trees too deep, tile kernel only
  int dfs_env_smem = 0, dfs_rob_smem = 0;
  bool dfs_all_smem = false;   // both trees fit into shared memory entirely
  size_t dfs_smem_bytes = 0;
  int transit_minb = 3;        // tile K2: blocks per SM its registers are capped for (OSR_TRANSIT_BLOCKS)
  int seg_warps = 0;           // K2 on the depth-first traversal: warps per block (0: tile kernel only)
  int seg_env_smem = 0, seg_rob_smem = 0;
  bool seg_all_smem = false;
  size_t seg_smem_bytes = 0;
  int num_sms = 0;
  osr::DevCounters* counters = nullptr;
  unsigned long long* work = nullptr;   // [0] tile queue, [1] contact record count, [2..] queues of pipelined chunks
  unsigned long long* spill = nullptr;
  int* err = nullptr;
  int* h_err = nullptr;                 // pinned landing word of the error flag (one small async copy per call)
  unsigned char* noalloc = nullptr;     // device address of g_noalloc_scratch on this GPU (cudaGetSymbolAddress)
  double* small_in = nullptr;           // mapped page-locked buffers of the small-call path (SMALL_STATES states / labels)
  uint8_t* small_out = nullptr;
  unsigned char* small_seg = nullptr;   // ... and of small segment calls (SMALL_SEGMENTS x 264 bytes: inputs + outputs)
  int grid_valid = 0, grid_transit = 0, grid_contacts = 0;
  int spill_cap = 0;
  DevBuf in0, in1, vd, out[6], recs, rowrec, seg, faces, cop, bary;
  DevBuf surf_v, surf_f, surf_flags, surf_cnt, surf_off, surf_face, surf_l;  // K5 (kernels_surface.cuh)
  DevBuf clr_q0, clr_stats, clr_dist;                                         // K8 (kernels_clearance.cuh)
  DevBuf c_cnt, c_msum, c_cur, c_recoff, c_rowoff, c_recs2;                   // K3c (kernels_contacts.cuh)
};

}  // namespace

struct osr_world {
  std::mutex mu;
  osr::TriMesh env_mesh, rob_mesh;
  bool has_env = false, has_rob = false, finalized = false;
  bool host_ready = false;  // unit meshes + BVHs built (true even when no CUDA device could be opened)
  bool centre_enforced = false;
  float centre[3] = {0, 0, 0};        // centre used for the robot (Scene::center_)
  float ompl_centre[3] = {0, 0, 0};   // mean of vertices (Scene::mean_of_vertices_)
  float scene_scale = 1.0f;
  double recommended_cres = 0.001;
  double perturb[7] = {0, 0, 0, 1, 0, 0, 0};
  osr::UnitModel env, rob;
  std::vector<Slot> slots;
  osr_counters host_counters = {};
  bool seg_dfs = false;            // OSR_SEG_KERNEL=dfs: K2 on the lane-owned depth-first traversal
  bool force_tile_kernel = false;  // OSR_VALID_KERNEL=tile: first-generation kernel, kept for A/B measurements
  int sweep_tiles = 64;            // K2: largest swept window, tiles of 32 steps (OSR_SWEEP_TILES, 0 = off)
  double rob_radius = 0.0;         // max |p| over the robot's unit-mesh vertices
  double sweep_rmax = 0.02;        // K2: largest swept radius worth a traversal (OSR_SWEEP_RMAX, unit-scene lengths)
  double sweep_rstart = 0.005;     // K2: swept radius of the first windows of a walk (OSR_SWEEP_RSTART)
  int sweep_short_tile = 8;        // K2: steps of the exact tile after a blocked 4-step window (OSR_SWEEP_SHORT_TILE)
  int sweep_nb0 = 4;               // K2: windows in the first swept traversal of a walk (OSR_SWEEP_NB0)
  int rec_warp_max = osr::REC_ORDER_WARP_MAX;   // K3c: largest bucket ranked by one warp (OSR_REC_WARP_MAX; tests)
  double sweep_keep_rmax = 3e-4;   // K2: windows of a larger swept radius keep no pairs (OSR_SWEEP_KEEP_RMAX, unit-scene lengths)
  int sweep_resolve = 512;         // K2: exact (pair, step) tests one window may cost before it counts as blocked (OSR_SWEEP_RESOLVE, 0 = off)

  osr::KParams params(const Slot& s) const {
    osr::KParams P;
    P.env.nodes = s.env_nodes; P.env.tris = s.env_tris;
    P.env.num_nodes = env.bvh.num_nodes; P.env.num_tris = env.nf();
    P.rob.nodes = s.rob_nodes; P.rob.tris = s.rob_tris;
    P.rob.num_nodes = rob.bvh.num_nodes; P.rob.num_tris = rob.nf();
    osr::state_to_rigid(perturb, P.env_tf);
    P.rob_radius = rob_radius;
    P.sweep_tiles = sweep_tiles;
    P.sweep_rmax = sweep_rmax;
    P.sweep_rstart = sweep_rstart;
    P.sweep_short_tile = sweep_short_tile;
    P.sweep_nb0 = sweep_nb0;
    P.sweep_resolve = sweep_resolve;
    P.sweep_keep_rmax = sweep_keep_rmax;
    P.counters = s.counters;
    P.work_counter = s.work;
    P.spill = s.spill;
    P.spill_cap = s.spill_cap;
    P.error_flag = s.err;
    return P;
  }
};

namespace {

constexpr int MAX_FEED_CHUNKS = 64;
constexpr int64_t SMALL_SEGMENTS = 1024;  // host segment calls up to this size: mapped buffers, one launch, one wait
constexpr int64_t SMALL_STATES = 16384;  // host validity calls up to this size take the small-call path

// cuStreamWriteValue64 (driver API stream memory operation) through the runtime's entry-point lookup, so the
// library does not link libcuda: a stream-ordered 8-byte write costs ~1 us where a tiny cudaMemcpyAsync costs
// a DMA round trip.  Null when the driver does not offer it (then the small copy is used).
typedef int (*StreamWriteValue64)(cudaStream_t, unsigned long long, unsigned long long, unsigned int);
StreamWriteValue64 stream_write_value64() {
  static StreamWriteValue64 fn = [] {
    void* p = nullptr;
    if (getenv("OSR_NO_STREAM_MEMOPS")) return (StreamWriteValue64) nullptr;
#if CUDART_VERSION >= 12050
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPointByVersion("cuStreamWriteValue64", &p, 12000, cudaEnableDefault, &st) != cudaSuccess ||
        st != cudaDriverEntryPointSuccess)
      p = nullptr;
#else
    cudaDriverEntryPointQueryResult st;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue64", &p, cudaEnableDefault, &st) != cudaSuccess ||
        st != cudaDriverEntryPointSuccess)
      p = nullptr;
#endif
    cudaGetLastError();
    return (StreamWriteValue64)p;
  }();
  return fn;
}
constexpr int WORK_SLOTS = 4;  // [0],[1] see Slot::work; [2],[3] the two pipeline lanes of host calls
constexpr size_t SMEM_PER_BLOCK = sizeof(osr::WarpShared) * osr::WARPS_PER_BLOCK;

void free_slot(Slot& s) {
  cudaSetDevice(s.dev);
  if (s.stream) cudaStreamSynchronize(s.stream);
  cudaFree(s.env_nodes); cudaFree(s.env_tris); cudaFree(s.rob_nodes); cudaFree(s.rob_tris);
  cudaFree(s.env_qnodes); cudaFree(s.rob_qnodes);
  cudaFree(s.counters); cudaFree(s.work); cudaFree(s.spill); cudaFree(s.err);
  s.in0.release(); s.in1.release(); s.vd.release();
  for (auto& b : s.out) b.release();
  s.surf_v.release(); s.surf_f.release(); s.surf_flags.release(); s.surf_cnt.release(); s.surf_off.release();
  s.surf_face.release(); s.surf_l.release();
  s.clr_q0.release(); s.clr_stats.release(); s.clr_dist.release();
  s.recs.release(); s.rowrec.release(); s.seg.release(); s.faces.release(); s.cop.release(); s.bary.release();
  if (s.ev0) cudaEventDestroy(s.ev0);
  for (auto& e : s.trace_ev) if (e) cudaEventDestroy(e);
  if (s.ev1) cudaEventDestroy(s.ev1);
  if (s.evp) cudaEventDestroy(s.evp);
  if (s.evq) cudaEventDestroy(s.evq);
  cudaFree(s.ready);
  if (s.stage) cudaFreeHost(s.stage);
  if (s.h_ready) cudaFreeHost(s.h_ready);
  if (s.h_err) cudaFreeHost(s.h_err);
  if (s.small_in) cudaFreeHost(s.small_in);
  if (s.small_out) cudaFreeHost(s.small_out);
  if (s.small_seg) cudaFreeHost(s.small_seg);
  for (auto& e : s.feed_ev) if (e) cudaEventDestroy(e);
  if (s.stream3) cudaStreamDestroy(s.stream3);
  if (s.stream2) cudaStreamDestroy(s.stream2);
  if (s.stream) cudaStreamDestroy(s.stream);
  s = Slot();
}

// NUMA node the device's PCIe root hangs off (sysfs), -1 when unknown.
static int device_numa_node(int dev) {
  if (const char* e = getenv("OSR_NUMA_NODE")) return atoi(e);   // override (guests that hide the topology; tests)
  char id[32] = {0};
  if (cudaDeviceGetPCIBusId(id, (int)sizeof id, dev) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  for (char* c = id; *c; c++) *c = (char)tolower(*c);
  FILE* f = fopen((std::string("/sys/bus/pci/devices/") + id + "/numa_node").c_str(), "r");
  if (!f) return -1;
  int node = -1;
  if (fscanf(f, "%d", &node) != 1) node = -1;
  fclose(f);
  return node;
}

// While alive the calling thread runs on the CPUs of the device's NUMA node, so that host memory allocated and first
// touched inside lands next to the GPU's PCIe root: a DMA from the other socket's memory crosses the inter-socket link
// and runs at about half the PCIe rate (what the slowest rank of a 4- or 8-process run was bound by).  Only the
// calling thread's mask changes, and it is restored.  OSR_NO_NUMA_LOCAL=1 turns it off.
struct NumaScope {
  cpu_set_t old;
  bool changed = false;
  int node = -1;
  explicit NumaScope(int dev) {
    static const bool off = getenv("OSR_NO_NUMA_LOCAL") != nullptr;
    if (off) return;
    node = device_numa_node(dev);
    if (node < 0 || sched_getaffinity(0, sizeof old, &old) != 0) return;
    char path[96];
    snprintf(path, sizeof path, "/sys/devices/system/node/node%d/cpulist", node);
    FILE* f = fopen(path, "r");
    if (!f) return;
    cpu_set_t want;
    CPU_ZERO(&want);
    int a = 0, b = 0;
    while (fscanf(f, "%d", &a) == 1) {   // "0-31,64-95"
      b = a;
      int c = fgetc(f);
      if (c == '-') {
        if (fscanf(f, "%d", &b) != 1) break;
        c = fgetc(f);
      }
      for (int k = a; k <= b && k < CPU_SETSIZE; k++)
        if (CPU_ISSET(k, &old)) CPU_SET(k, &want);
      if (c != ',') break;
    }
    fclose(f);
    if (CPU_COUNT(&want) == 0) return;
    changed = sched_setaffinity(0, sizeof want, &want) == 0;
  }
  ~NumaScope() {
    if (changed) sched_setaffinity(0, sizeof old, &old);
  }
};

int init_slot(osr_world* w, Slot& s, int dev) {
  s.dev = dev;
  CU(cudaSetDevice(dev));
  NumaScope numa(dev);   // the slot's page-locked buffers (staging ring, mapped small-call buffers) next to the GPU
  CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&s.stream2, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&s.stream3, cudaStreamNonBlocking));
  CU(cudaEventCreate(&s.ev0));
  CU(cudaEventCreate(&s.ev1));
  CU(cudaEventCreateWithFlags(&s.evp, cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&s.evq, cudaEventDisableTiming));
  CU(cudaMalloc((void**)&s.ready, sizeof(long long)));
  CU(cudaHostAlloc((void**)&s.h_ready, MAX_FEED_CHUNKS * sizeof(long long), cudaHostAllocDefault));
  auto up = [&](void** d, const void* h, size_t bytes) -> cudaError_t {
    cudaError_t e = cudaMalloc(d, bytes);
    if (e != cudaSuccess) return e;
    return cudaMemcpy(*d, h, bytes, cudaMemcpyHostToDevice);
  };
  CU(up((void**)&s.env_nodes, w->env.bvh.nodes.data(), w->env.bvh.nodes.size() * 4));
  CU(up((void**)&s.env_tris, w->env.tris.data(), w->env.tris.size() * 8));
  CU(up((void**)&s.rob_nodes, w->rob.bvh.nodes.data(), w->rob.bvh.nodes.size() * 4));
  CU(up((void**)&s.rob_tris, w->rob.tris.data(), w->rob.tris.size() * 8));
  CU(up((void**)&s.env_qnodes, w->env.bvh.qnodes.data(), w->env.bvh.qnodes.size() * 4));
  CU(up((void**)&s.rob_qnodes, w->rob.bvh.qnodes.data(), w->rob.bvh.qnodes.size() * 4));
  CU(cudaMalloc((void**)&s.counters, sizeof(osr::DevCounters)));
  CU(cudaMemset(s.counters, 0, sizeof(osr::DevCounters)));
  CU(cudaMalloc((void**)&s.work, WORK_SLOTS * sizeof(unsigned long long)));
  CU(cudaMalloc((void**)&s.err, sizeof(int)));
  CU(cudaMemset(s.err, 0, sizeof(int)));
  CU(cudaHostAlloc((void**)&s.h_err, sizeof(int), cudaHostAllocDefault));
  CU(cudaGetSymbolAddress((void**)&s.noalloc, g_noalloc_scratch));   // this device's copy of the static scratch
  CU(cudaHostAlloc((void**)&s.small_in, SMALL_STATES * 56, cudaHostAllocMapped));
  CU(cudaHostAlloc((void**)&s.small_out, SMALL_STATES, cudaHostAllocMapped));
  CU(cudaHostAlloc((void**)&s.small_seg, SMALL_SEGMENTS * 264, cudaHostAllocMapped));
  // persistent grid: as many blocks as fit, on every SM
  CU(cudaFuncSetAttribute(osr::k_is_valid<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  CU(cudaFuncSetAttribute(osr::k_is_valid<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  CU(cudaFuncSetAttribute(osr::k_ompl_motion, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  CU(cudaFuncSetAttribute(osr::k_transit<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  CU(cudaFuncSetAttribute(osr::k_transit<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  CU(cudaFuncSetAttribute(osr::k_transit<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  {
    const char* e = getenv("OSR_TRANSIT_BLOCKS");
    s.transit_minb = e ? std::min(4, std::max(2, atoi(e))) : 3;
  }
  CU(cudaFuncSetAttribute(osr::k_contacts, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_PER_BLOCK));
  int nsm = 0, bpsm = 0, b2 = 0, b3 = 0;
  CU(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bpsm, osr::k_is_valid<false>, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK));
  if (s.transit_minb == 4) CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, osr::k_transit<4>, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK));
  else if (s.transit_minb == 3) CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, osr::k_transit<3>, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK));
  else CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, osr::k_transit<2>, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK));
  CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b3, osr::k_contacts, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK));
  s.grid_valid = nsm * std::max(1, bpsm);
  s.grid_transit = nsm * std::max(1, b2);
  s.grid_contacts = nsm * std::max(1, b3);
  s.num_sms = nsm;
  {  // lane-owned DFS kernel: pick the task width / stack depth variant and the shared-memory node budget
    int need = w->env.bvh.depth + w->rob.bvh.depth + 2;
    bool small = w->env.bvh.num_nodes < 32768 && w->rob.bvh.num_nodes < 32768;
    size_t wm = 0;
    if (small && need <= 32) { s.dfs_variant = 0; wm = sizeof(osr::DfsWarpMem<uint32_t, 32>); }
    else if (small && need <= 64) { s.dfs_variant = 1; wm = sizeof(osr::DfsWarpMem<uint32_t, 64>); }
    else if (need <= 48) { s.dfs_variant = 2; wm = sizeof(osr::DfsWarpMem<unsigned long long, 48>); }
    else s.dfs_variant = -1;
    if (s.dfs_variant >= 0) {
      int max_smem = 0;
      CU(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
      size_t fixed = wm * osr::DFS_WARPS;
      long long budget = ((long long)max_smem - (long long)fixed) / 32;  // nodes
      if (const char* cap = getenv("OSR_DFS_SMEM_NODES")) {                // A/B: leave more of the SM's memory to L1
        long long c = atoll(cap);
        if (c >= 2 && (w->env.bvh.num_nodes + w->rob.bvh.num_nodes) > budget) budget = std::min(budget, c);
      }
      if (budget < 2) s.dfs_variant = -1;
      else {
        long long ne = w->env.bvh.num_nodes, nr = w->rob.bvh.num_nodes;
        long long se = std::min(ne, budget / 2), sr = std::min(nr, budget / 2);
        se = std::min(ne, budget - sr);
        sr = std::min(nr, budget - se);
        s.dfs_env_smem = (int)se; s.dfs_rob_smem = (int)sr;
        s.dfs_smem_bytes = (size_t)(se + sr) * 32 + fixed;
        s.dfs_all_smem = se == ne && sr == nr;
#define OSR_DFS_ATTR(...)                                                                                         \
  CU(cudaFuncSetAttribute(osr::k_valid_dfs<__VA_ARGS__, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem)); \
  CU(cudaFuncSetAttribute(osr::k_valid_dfs<__VA_ARGS__, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));  \
  CU(cudaFuncSetAttribute(osr::k_valid_dfs<__VA_ARGS__, true, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem))
        OSR_DFS_ATTR(uint32_t, 32, true);
        OSR_DFS_ATTR(uint32_t, 32, false);
        OSR_DFS_ATTR(uint32_t, 64, false);
        OSR_DFS_ATTR(unsigned long long, 48, false);
#undef OSR_DFS_ATTR
      }
    }
  }
  if (s.dfs_variant >= 0) {  // K2 (kernels_seg.cuh): same tree variants, its own per-warp shared memory
    int max_smem = 0;
    CU(cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    const char* e = getenv("OSR_SEG_WARPS");
    int sw = (s.dfs_variant == 0) ? ((e && atoi(e) == 8) ? 8 : 12) : 8;
    size_t wm = s.dfs_variant == 0 ? sizeof(osr::SegWarpMem<uint32_t, 32>)
              : s.dfs_variant == 1 ? sizeof(osr::SegWarpMem<uint32_t, 64>)
                                   : sizeof(osr::SegWarpMem<unsigned long long, 48>);
    size_t fixed = wm * sw;
    long long budget = ((long long)max_smem - (long long)fixed) / 32;
    if (budget >= 2) {
      long long ne = w->env.bvh.num_nodes, nr = w->rob.bvh.num_nodes;
      long long se = std::min(ne, budget / 2), sr = std::min(nr, budget / 2);
      se = std::min(ne, budget - sr);
      sr = std::min(nr, budget - se);
      s.seg_warps = sw;
      s.seg_env_smem = (int)se; s.seg_rob_smem = (int)sr;
      s.seg_all_smem = se == ne && sr == nr;
      s.seg_smem_bytes = (size_t)(se + sr) * 32 + fixed;
#define OSR_SEG_ATTR(...) CU(cudaFuncSetAttribute(osr::k_transit_dfs<__VA_ARGS__>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem))
      OSR_SEG_ATTR(uint32_t, 32, true, 8);
      OSR_SEG_ATTR(uint32_t, 32, false, 8);
      OSR_SEG_ATTR(uint32_t, 32, true, 12);
      OSR_SEG_ATTR(uint32_t, 32, false, 12);
      OSR_SEG_ATTR(uint32_t, 64, false, 8);
      OSR_SEG_ATTR(unsigned long long, 48, false, 8);
#undef OSR_SEG_ATTR
    }
  }
  s.spill_cap = 8192;
  size_t warps = (size_t)std::max(s.grid_valid, std::max(s.grid_transit, s.grid_contacts)) * osr::WARPS_PER_BLOCK;
  CU(cudaMalloc((void**)&s.spill, warps * s.spill_cap * sizeof(unsigned long long)));
  return OSR_OK;
}

// contiguous index-range shards, the reference's partt.get_task_chunk rule
// (/root/reference/src/GP/pipeline/partt.py:22-35)
inline void shard(int64_t n, int parts, int i, int64_t& lo, int64_t& hi) {
  lo = n * i / parts;
  hi = n * (i + 1) / parts;
}

int check_ready(osr_world* w) {
  if (!w) return fail(OSR_EINVAL, "null world");
  if (!w->finalized) return fail(OSR_EINVAL, "osr_finalize() has not been called");
  return OSR_OK;
}

thread_local bool g_feed_timed_out = false;   // set when a stream-fed K1 launch gave up waiting for its input
int check_device_error(osr_world* w, Slot& s) {
  CU(cudaMemcpyAsync(s.h_err, s.err, sizeof(int), cudaMemcpyDeviceToHost, s.stream));
  CU(cudaStreamSynchronize(s.stream));
  const int flag = *s.h_err;
  if (flag) {
    cudaMemsetAsync(s.err, 0, sizeof(int), s.stream);
    if (flag == 2) {
      g_feed_timed_out = true;
      return fail(OSR_ECUDA, "validity kernel timed out waiting for its input upload");
    }
    if (flag == 3)
      return fail(OSR_EINVAL, "verify_delta <= 0 on a segment of non-zero length (the reference's walk never ends)");
    return fail(OSR_EOVERFLOW, "device traversal stack spill area exhausted");
  }
  (void)w;
  return OSR_OK;
}

}  // namespace

namespace {
// the four tree-size variants of K1 x {resident input, stream-fed input}
template <bool FEED, typename T>
void launch_dfs_variant(const Slot& s, const osr::DfsParams& D0, const T* d_q, long long n, uint8_t* d_out,
                        cudaStream_t st, int grid = 0) {
  const int g = grid > 0 ? grid : s.num_sms;   // persistent: one block per SM (fewer for the small-call path)
  // the last (warps x tail_group) states go out in groups of tail_group (DfsParams::tail_start); OSR_K1_TAIL_GROUP=0
  // = groups of 32 to the end (A/B)
  static const int tail_group = [] { const char* e = getenv("OSR_K1_TAIL_GROUP"); return e ? std::max(0, std::min(32, atoi(e))) : 8; }();
  osr::DfsParams D = D0;
  D.tail_group = tail_group > 0 ? tail_group : 32;
  const long long tail = (long long)g * (osr::DFS_THREADS / 32) * D.tail_group;
  D.tail_start = tail_group > 0 ? std::max(0LL, n - tail) / 32 * 32 : (n + 31) / 32 * 32;
  if (s.dfs_variant == 0 && s.dfs_all_smem)
    osr::k_valid_dfs<uint32_t, 32, true, FEED, T><<<g, osr::DFS_THREADS, s.dfs_smem_bytes, st>>>(D, d_q, n, d_out);
  else if (s.dfs_variant == 0)
    osr::k_valid_dfs<uint32_t, 32, false, FEED, T><<<g, osr::DFS_THREADS, s.dfs_smem_bytes, st>>>(D, d_q, n, d_out);
  else if (s.dfs_variant == 1)
    osr::k_valid_dfs<uint32_t, 64, false, FEED, T><<<g, osr::DFS_THREADS, s.dfs_smem_bytes, st>>>(D, d_q, n, d_out);
  else
    osr::k_valid_dfs<unsigned long long, 48, false, FEED, T><<<g, osr::DFS_THREADS, s.dfs_smem_bytes, st>>>(D, d_q, n, d_out);
}

}  // namespace

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

const char* osr_last_error(void) { return g_err.c_str(); }

int osr_create(osr_world** out) {
  if (!out) return fail(OSR_EINVAL, "null out");
  *out = new osr_world();
  const char* k = getenv("OSR_VALID_KERNEL");
  (*out)->force_tile_kernel = k && strcmp(k, "tile") == 0;
  const char* k2 = getenv("OSR_SEG_KERNEL");
  (*out)->seg_dfs = k2 && strcmp(k2, "dfs") == 0;
  if (const char* k3 = getenv("OSR_SWEEP_TILES")) (*out)->sweep_tiles = std::max(0, std::min(1 << 16, atoi(k3)));
  if (const char* k4 = getenv("OSR_SWEEP_RMAX")) (*out)->sweep_rmax = std::max(0.0, atof(k4));
  if (const char* k5 = getenv("OSR_SWEEP_RSTART")) (*out)->sweep_rstart = std::max(0.0, atof(k5));
  if (const char* k10 = getenv("OSR_REC_WARP_MAX")) (*out)->rec_warp_max = std::max(1, atoi(k10));
  if (const char* k9 = getenv("OSR_SWEEP_NB0")) (*out)->sweep_nb0 = std::max(1, std::min(32, atoi(k9)));
  if (const char* k6 = getenv("OSR_SWEEP_SHORT_TILE")) (*out)->sweep_short_tile = std::max(1, std::min(32, atoi(k6)));
  if (const char* k8 = getenv("OSR_SWEEP_KEEP_RMAX")) (*out)->sweep_keep_rmax = std::max(0.0, atof(k8));
  if (const char* k7 = getenv("OSR_SWEEP_RESOLVE")) (*out)->sweep_resolve = std::max(0, std::min(1 << 20, atoi(k7)));
  return OSR_OK;
}

void osr_destroy(osr_world* w) {
  if (!w) return;
  for (auto& s : w->slots) free_slot(s);
  delete w;
}

static int set_mesh(osr_world* w, bool rob, const float* V, int nV, const int32_t* F, int nF) {
  if (!w || !V || !F || nV <= 0 || nF <= 0) return fail(OSR_EINVAL, "bad mesh arguments");
  for (int i = 0; i < 3 * nF; i++)
    if (F[i] < 0 || F[i] >= nV) return fail(OSR_EINVAL, "face index out of range");
  osr::TriMesh& m = rob ? w->rob_mesh : w->env_mesh;
  m.V.assign(V, V + 3 * (size_t)nV);
  m.F.assign(F, F + 3 * (size_t)nF);
  (rob ? w->has_rob : w->has_env) = true;
  w->finalized = false;
  return OSR_OK;
}
int osr_set_env_mesh(osr_world* w, const float* V, int nV, const int32_t* F, int nF) { return set_mesh(w, false, V, nV, F, nF); }
int osr_set_rob_mesh(osr_world* w, const float* V, int nV, const int32_t* F, int nF) { return set_mesh(w, true, V, nV, F, nF); }

static int load_mesh(osr_world* w, bool rob, const char* path) {
  if (!w || !path) return fail(OSR_EINVAL, "bad arguments");
  std::string err;
  osr::TriMesh m;
  if (!osr::load_obj(path, m, err)) return fail(OSR_EIO, err);
  (rob ? w->rob_mesh : w->env_mesh) = std::move(m);
  (rob ? w->has_rob : w->has_env) = true;
  w->finalized = false;
  return OSR_OK;
}
int osr_load_env_obj(osr_world* w, const char* path) { return load_mesh(w, false, path); }
int osr_load_rob_obj(osr_world* w, const char* path) { return load_mesh(w, true, path); }

int osr_enforce_robot_center(osr_world* w, const double c[3]) {
  if (!w || !c) return fail(OSR_EINVAL, "bad arguments");
  for (int k = 0; k < 3; k++) w->centre[k] = (float)c[k];
  w->centre_enforced = true;
  w->finalized = false;
  return OSR_OK;
}

int osr_finalize(osr_world* w, const int* devices, int ndev) {
  if (!w) return fail(OSR_EINVAL, "null world");
  std::lock_guard<std::mutex> lk(w->mu);
  if (!w->has_env || !w->has_rob) return fail(OSR_EINVAL, "both the environment and the robot mesh must be loaded");
  w->finalized = false;
  w->host_ready = false;
  // UnitWorld::scaleToUnit (unit_world.cc:123-131)
  {
    float es = osr::bbox_span(w->env_mesh), rs = osr::bbox_span(w->rob_mesh);
    w->scene_scale = (float)(1.0 / std::max(es, rs));
    osr::mean_of_vertices(w->rob_mesh, w->ompl_centre);
    if (!w->centre_enforced)
      for (int k = 0; k < 3; k++) w->centre[k] = w->ompl_centre[k];
    // UnitWorld::angleModel(0,0) (unit_world.cc:137-161) -> CDModel (cdmodel.cc:85-160)
    osr::make_unit_model(w->env_mesh, w->scene_scale, nullptr, w->env);
    osr::make_unit_model(w->rob_mesh, w->scene_scale, w->centre, w->rob);
    double r2 = 0.0;
    for (size_t v = 0; v + 2 < w->rob.V.size(); v += 3)
      r2 = std::max(r2, w->rob.V[v] * w->rob.V[v] + w->rob.V[v + 1] * w->rob.V[v + 1] + w->rob.V[v + 2] * w->rob.V[v + 2]);
    w->rob_radius = std::sqrt(r2) * (1.0 + 1e-12);
    osr::build_gpu_bvh(w->env);
    osr::build_gpu_bvh(w->rob);
    if (w->env.bvh.num_nodes >= (1 << 27) || w->rob.bvh.num_nodes >= (1 << 27))
      return fail(OSR_EINVAL, "mesh too large for the 27-bit node index of the work stack");
    w->host_ready = true;
  }
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0)
    return fail(OSR_ENODEV, "no CUDA device available (the engine has no CPU fallback)");
  std::vector<int> devs;
  if (ndev <= 0 || !devices) {
    int cur = 0;
    cudaGetDevice(&cur);
    devs.push_back(cur);
  } else {
    for (int i = 0; i < ndev; i++) {
      if (devices[i] < 0 || devices[i] >= count) return fail(OSR_ENODEV, "device index out of range");
      devs.push_back(devices[i]);
    }
  }
  for (auto& s : w->slots) free_slot(s);
  w->slots.assign(devs.size(), Slot());
  for (size_t i = 0; i < devs.size(); i++) {
    int rc = init_slot(w, w->slots[i], devs[i]);
    if (rc != OSR_OK) return rc;
  }
  w->finalized = true;
  return OSR_OK;
}

int osr_set_perturbation(osr_world* w, const double q[7]) {
  if (!w || !q) return fail(OSR_EINVAL, "bad arguments");
  std::lock_guard<std::mutex> lk(w->mu);
  memcpy(w->perturb, q, sizeof(w->perturb));
  return OSR_OK;
}
double osr_scene_scale(const osr_world* w) { return w ? (double)w->scene_scale : 0.0; }
double osr_get_recommended_cres(const osr_world* w) { return w ? w->recommended_cres : 0.0; }
int osr_set_recommended_cres(osr_world* w, double v) {
  if (!w) return fail(OSR_EINVAL, "null world");
  w->recommended_cres = v;
  return OSR_OK;
}
int osr_get_robot_center(const osr_world* w, double c[3]) {
  if (!w || !c) return fail(OSR_EINVAL, "bad arguments");
  for (int k = 0; k < 3; k++) c[k] = w->centre[k];
  return OSR_OK;
}
int osr_get_ompl_center(const osr_world* w, int which, double c[3]) {  // Scene::getOMPLCenter = mean of vertices
  if (!w || !c || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  float m[3];
  if (which) for (int k = 0; k < 3; k++) m[k] = w->ompl_centre[k];
  else osr::mean_of_vertices(w->env_mesh, m);
  for (int k = 0; k < 3; k++) c[k] = m[k];
  return OSR_OK;
}
int osr_num_devices(const osr_world* w) { return w ? (int)w->slots.size() : 0; }
int osr_mesh_sizes(const osr_world* w, int which, int* nV, int* nF, int* nNodes, int* depth) {
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  const osr::UnitModel& m = which ? w->rob : w->env;
  if (nV) *nV = (int)(m.V.size() / 3);
  if (nF) *nF = m.nf();
  if (nNodes) *nNodes = m.bvh.num_nodes;
  if (depth) *depth = m.bvh.depth;
  return OSR_OK;
}
int osr_get_unit_mesh(const osr_world* w, int which, double* V, int32_t* F) {
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  const osr::UnitModel& m = which ? w->rob : w->env;
  if (V) memcpy(V, m.V.data(), m.V.size() * 8);
  if (F) memcpy(F, m.F.data(), m.F.size() * 4);
  return OSR_OK;
}

int osr_get_counters(osr_world* w, osr_counters* out, int reset) {
  int rc = check_ready(w);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(w->mu);
  osr_counters tot = w->host_counters;
  for (auto& s : w->slots) {
    CU(cudaSetDevice(s.dev));
    osr::DevCounters c;
    CU(cudaMemcpyAsync(&c, s.counters, sizeof(c), cudaMemcpyDeviceToHost, s.stream));
    CU(cudaStreamSynchronize(s.stream));
    tot.bv_tests += c.bv; tot.leaf_tests += c.leaf; tot.fp64_rechecks += c.fp64; tot.poses += c.poses;
    tot.spills += c.spills;
    tot.warp_iters += c.warp_iters; tot.lanes_active += c.lanes_active; tot.lanes_waiting += c.lanes_waiting;
    tot.lanes_idle += c.lanes_idle;
    if (reset) CU(cudaMemsetAsync(s.counters, 0, sizeof(c), s.stream));
  }
  if (out) *out = tot;
  if (reset) w->host_counters = osr_counters{};
  return OSR_OK;
}

// ---- launch helpers ---------------------------------------------------------------------------
// How long a stream-fed K1 launch waits for an upload before it gives up (the host then re-runs the call with
// every copy queued in front of the launch).  Generous: the wait only ends early when something is wrong.
static unsigned long long feed_timeout_ns() {
  static unsigned long long v = [] {
    const char* e = getenv("OSR_FEED_TIMEOUT_S");
    double sec = e ? atof(e) : 20.0;
    if (!(sec > 0.0)) sec = 20.0;
    return (unsigned long long)(sec * 1e9);
  }();
  return v;
}
// Tools that serialise launches (CUDA_LAUNCH_BLOCKING, or an injected profiler / sanitizer that replays kernels)
// would make a kernel launched BEFORE its uploads are queued wait for input that cannot arrive.
static bool launches_may_block() {
  static bool v = [] {
    const char* lb = getenv("CUDA_LAUNCH_BLOCKING");
    if (lb && atoi(lb) != 0) return true;
    const char* inj = getenv("CUDA_INJECTION64_PATH");
    if (inj && *inj) return true;
    if (getenv("OSR_FEED_SAFE") != nullptr) return true;
    // Nsight Compute injects its interceptor without that variable; a kernel it replays sees the `ready` word of
    // the first pass again and would wait out the feed timeout in every pass (observed: 40 passes x 20 s)
    if (FILE* f = fopen("/proc/self/maps", "r")) {
      char line[512];
      bool hit = false;
      while (!hit && fgets(line, sizeof line, f))
        hit = strstr(line, "InjectionTarget") || strstr(line, "nsight-compute") || strstr(line, "cuda-injection");
      fclose(f);
      if (hit) return true;
    }
    return false;
  }();
  return v;
}
static osr::DfsParams dfs_params(const osr_world* w, const Slot& s) {
  osr::DfsParams P;
  P.env_q = s.env_qnodes; P.rob_q = s.rob_qnodes;
  P.env_tris = s.env_tris; P.rob_tris = s.rob_tris;
  P.env_nodes = w->env.bvh.num_nodes; P.rob_nodes = w->rob.bvh.num_nodes;
  P.env_smem_nodes = s.dfs_env_smem; P.rob_smem_nodes = s.dfs_rob_smem;
  osr::state_to_rigid(w->perturb, P.env_tf);
  for (int k = 0; k < 4; k++) P.env_quat[k] = w->perturb[3 + k];
  P.r_rob = w->rob.bvh.max_radius;
  P.counters = s.counters;
  P.work_counter = s.work;
  P.error_flag = s.err;
  P.ready = nullptr;
  P.ready_floor = 0;
  P.feed_timeout_ns = feed_timeout_ns();
  return P;
}

// `lane` 0: the handle's main stream with events around the kernel; 1 / 2: the two pipeline lanes of a host
// call (main stream / stream2), each with its own work queue word so that consecutive chunks may overlap
static int launch_valid(osr_world* w, Slot& s, const double* d_q, int64_t n, uint8_t* d_out, int lane = 0,
                        const long long* ready = nullptr) {
  cudaStream_t st = lane == 2 ? s.stream2 : s.stream;
  unsigned long long* work = lane == 0 ? s.work : s.work + 1 + lane;
  CU(cudaMemsetAsync(work, 0, (lane == 0 ? 2 : 1) * sizeof(unsigned long long), st));
  if (s.dfs_variant >= 0 && !w->force_tile_kernel) {
    if (n > 0x7fffffffLL) return fail(OSR_EINVAL, "more than 2^31-1 states in one launch");
    osr::DfsParams D = dfs_params(w, s);
    D.work_counter = work;
    D.ready = ready;
    if (lane != 0) {
      if (ready) {
        CU(cudaEventRecord(s.ev0, st));
        launch_dfs_variant<true, double>(s, D, d_q, (long long)n, d_out, st);
      } else launch_dfs_variant<false, double>(s, D, d_q, (long long)n, d_out, st);
      CU(cudaGetLastError());
      if (ready) CU(cudaEventRecord(s.ev1, st));
      w->host_counters.kernel_launches++;
      return OSR_OK;
    }
    CU(cudaEventRecord(s.ev0, s.stream));
    static const bool probe_feed = getenv("OSR_PROBE_FEED_KERNEL") != nullptr;  // A/B: the fed variant on resident input
    if (probe_feed) { D.ready_floor = (long long)n; launch_dfs_variant<true, double>(s, D, d_q, (long long)n, d_out, s.stream); }
    else
    launch_dfs_variant<false, double>(s, D, d_q, (long long)n, d_out, s.stream);
    CU(cudaGetLastError());
    CU(cudaEventRecord(s.ev1, s.stream));
    w->host_counters.kernel_launches++;
    return OSR_OK;
  }
  osr::KParams P = w->params(s);
  P.work_counter = work;
  if (lane == 0) CU(cudaEventRecord(s.ev0, s.stream));
  osr::k_is_valid<false><<<s.grid_valid, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, st>>>(P, d_q, (long long)n, d_out, 32);
  CU(cudaGetLastError());
  if (lane == 0) CU(cudaEventRecord(s.ev1, s.stream));
  w->host_counters.kernel_launches++;
  return OSR_OK;
}
static int launch_transit(osr_world* w, Slot& s, const double* d_from, const double* d_to, int64_t n,
                          const double* d_vd, int64_t vd_count, double vd_scalar, int transition_mode,
                          osr::SegOut O) {
  CU(cudaMemsetAsync(s.work, 0, 2 * sizeof(unsigned long long), s.stream));
  // K2 default = the tile kernel: on every measured segment workload (profiles/r01/README.md) the warp-shared
  // frontier keeps all 32 lanes busy on the mostly FREE steps of a walk, where depth-first early exit buys nothing.
  // OSR_SEG_KERNEL=dfs selects the lane-owned depth-first variant (kernels_seg.cuh) for A/B runs.
  if (s.seg_warps > 0 && w->seg_dfs && !O.to_idx) {   // (indexed segments exist in the tile kernel only)
    osr::DfsParams D = dfs_params(w, s);
    D.env_smem_nodes = s.seg_env_smem; D.rob_smem_nodes = s.seg_rob_smem;
    osr::SegArgs A = {d_from, d_to, (long long)n, d_vd, (long long)vd_count, vd_scalar, transition_mode, O};
    CU(cudaEventRecord(s.ev0, s.stream));
    const int nb = s.num_sms, nt = s.seg_warps * 32;
    const size_t sm = s.seg_smem_bytes;
    if (s.dfs_variant == 0 && s.seg_warps == 8 && s.seg_all_smem) osr::k_transit_dfs<uint32_t, 32, true, 8><<<nb, nt, sm, s.stream>>>(D, A);
    else if (s.dfs_variant == 0 && s.seg_warps == 8) osr::k_transit_dfs<uint32_t, 32, false, 8><<<nb, nt, sm, s.stream>>>(D, A);
    else if (s.dfs_variant == 0 && s.seg_all_smem) osr::k_transit_dfs<uint32_t, 32, true, 12><<<nb, nt, sm, s.stream>>>(D, A);
    else if (s.dfs_variant == 0) osr::k_transit_dfs<uint32_t, 32, false, 12><<<nb, nt, sm, s.stream>>>(D, A);
    else if (s.dfs_variant == 1) osr::k_transit_dfs<uint32_t, 64, false, 8><<<nb, nt, sm, s.stream>>>(D, A);
    else osr::k_transit_dfs<unsigned long long, 48, false, 8><<<nb, nt, sm, s.stream>>>(D, A);
    CU(cudaGetLastError());
    CU(cudaEventRecord(s.ev1, s.stream));
    w->host_counters.kernel_launches++;
    return OSR_OK;
  }
  osr::KParams P = w->params(s);
  CU(cudaEventRecord(s.ev0, s.stream));
  if (s.transit_minb == 4)
    osr::k_transit<4><<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, d_from, d_to, (long long)n, d_vd, (long long)vd_count, vd_scalar, transition_mode, O);
  else if (s.transit_minb == 3)
    osr::k_transit<3><<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, d_from, d_to, (long long)n, d_vd, (long long)vd_count, vd_scalar, transition_mode, O);
  else
    osr::k_transit<2><<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, d_from, d_to, (long long)n, d_vd, (long long)vd_count, vd_scalar, transition_mode, O);
  CU(cudaGetLastError());
  CU(cudaEventRecord(s.ev1, s.stream));
  w->host_counters.kernel_launches++;
  return OSR_OK;
}

// ---- validity ---------------------------------------------------------------------------------
// Host buffers in, host buffers out.  Every device gets a contiguous index range (shard()); inside a range the
// states move in chunks through a two-lane pipeline (main stream / stream2), so that the PCIe upload of chunk
// i+1 and the download of chunk i-1 run under the kernel of chunk i -- the call costs max(copy, kernel), not
// their sum, when the caller's buffers are pinned.
static int64_t host_chunk_states() {
  static int64_t v = [] {
    const char* e = getenv("OSR_HOST_CHUNK");
    long long x = e ? atoll(e) : 0;
    return (int64_t)(x > 0 ? x : (1 << 16));
  }();
  return v;
}

}  // extern "C" (templates cannot have C linkage)

// stream-fed K1 launch on the main stream (work word [2]); T = double or float states
template <typename T>
static int launch_valid_feed(osr_world* w, Slot& s, const T* d_q, int64_t n, uint8_t* d_out, const long long* ready,
                             long long ready_floor) {
  if (n > 0x7fffffffLL) return fail(OSR_EINVAL, "more than 2^31-1 states in one launch");
  unsigned long long* work = s.work + 2;
  CU(cudaMemsetAsync(work, 0, sizeof(unsigned long long), s.stream));
  osr::DfsParams D = dfs_params(w, s);
  D.work_counter = work;
  D.ready = ready;
  D.ready_floor = ready_floor;
  CU(cudaEventRecord(s.ev0, s.stream));
  launch_dfs_variant<true, T>(s, D, d_q, (long long)n, d_out, s.stream);
  CU(cudaGetLastError());
  CU(cudaEventRecord(s.ev1, s.stream));
  w->host_counters.kernel_launches++;
  return OSR_OK;
}

static int staging_threads() {
  static int v = [] {
    const char* e = getenv("OSR_STAGING_THREADS");
    // default: half of the host's hardware threads, between 2 and 8 (the staging is bound by the memory system)
    int x = e ? atoi(e) : std::max(2, std::min(8, (int)std::thread::hardware_concurrency() / 2));
    return std::max(0, std::min(16, x));
  }();
  return v;
}

// Copy into a pinned staging buffer with non-temporal stores: the buffer is read next by the DMA engine, not by a
// core, and ordinary stores would first read every destination line into the cache (the host's memory system, not
// its cores, bounds the staging of a pageable upload).
static void stage_copy(void* dst, const void* src, size_t bytes) {
#if defined(__x86_64__)
  if ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0) {
    const __m128i* s = (const __m128i*)src;
    __m128i* d = (__m128i*)dst;
    size_t n = bytes / 64;
    for (size_t i = 0; i < n; i++) {
      __m128i a = _mm_load_si128(s + 4 * i), b = _mm_load_si128(s + 4 * i + 1), c = _mm_load_si128(s + 4 * i + 2),
              e = _mm_load_si128(s + 4 * i + 3);
      _mm_stream_si128(d + 4 * i, a); _mm_stream_si128(d + 4 * i + 1, b);
      _mm_stream_si128(d + 4 * i + 2, c); _mm_stream_si128(d + 4 * i + 3, e);
    }
    _mm_sfence();
    const size_t done = n * 64;
    if (done < bytes) memcpy((char*)dst + done, (const char*)src + done, bytes - done);
    return;
  }
#endif
  memcpy(dst, src, bytes);
}

// Persistent staging workers (a host call on pageable arrays used to create and join its threads: ~0.3 ms of a 3 ms
// call).  Detached: they sleep on the queue and end with the process.
class StagePool {
 public:
  // runs job(0) .. job(n - 1) on the workers; returns false when a worker could not be created (caller runs them)
  bool run(int n, const std::function<void(int)>& job, std::atomic<int>& remaining) {
    std::unique_lock<std::mutex> lk(m_);
    while ((int)workers_ < n) {
      try {
        std::thread([this] { loop(); }).detach();
        workers_++;
      } catch (...) {
        return false;
      }
    }
    remaining.store(n);
    for (int t = 0; t < n; t++) q_.push_back([&job, &remaining, t] { job(t); remaining.fetch_sub(1, std::memory_order_release); });
    lk.unlock();
    cv_.notify_all();
    return true;
  }

 private:
  void loop() {
    for (;;) {
      std::function<void()> f;
      {
        std::unique_lock<std::mutex> lk(m_);
        cv_.wait(lk, [this] { return !q_.empty(); });
        f = std::move(q_.front());
        q_.pop_front();
      }
      f();
    }
  }
  std::mutex m_;
  std::condition_variable cv_;
  std::deque<std::function<void()>> q_;
  size_t workers_ = 0;
};
static StagePool& stage_pool() {
  static StagePool* p = new StagePool();   // never destroyed: its workers may outlive static destructors
  return *p;
}

static bool is_pageable(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
  return a.type == cudaMemoryTypeUnregistered;
}

// Stream-fed launch from PAGEABLE host memory.  cudaMemcpyAsync from pageable memory is staged by the driver on
// the calling thread (~9 GB/s measured, and the launch behind it waits for the whole upload).  Here the kernel is
// launched first and polls `ready`; worker threads copy the chunks into a ring of pinned staging buffers in
// parallel and queue the DMA from there; this thread advances `ready` behind the chunks in index order.
template <typename T>
static int feed_pageable(osr_world* w, Slot& s, const T* q, int64_t m, int64_t ck, uint8_t* out) {
  const int nchunks = (int)((m + ck - 1) / ck);
  const size_t cbytes = (size_t)ck * 7 * sizeof(T);
  const int R = std::min(nchunks, 8);
  if (s.stage_cap < (size_t)R * cbytes) {
    if (s.stage) cudaFreeHost(s.stage);
    s.stage = nullptr; s.stage_cap = 0;
    CU(cudaHostAlloc((void**)&s.stage, (size_t)R * cbytes, cudaHostAllocDefault));
    s.stage_cap = (size_t)R * cbytes;
  }
  for (int c = 0; c < nchunks; c++)
    if (!s.feed_ev[c]) CU(cudaEventCreateWithFlags(&s.feed_ev[c], cudaEventDisableTiming));
  CU(cudaEventRecord(s.evp, s.stream));
  CU(cudaStreamWaitEvent(s.stream2, s.evp, 0));
  CU(cudaStreamWaitEvent(s.stream3, s.evp, 0));
  CU(cudaMemsetAsync(s.ready, 0, sizeof(long long), s.stream3));
  CU(cudaEventRecord(s.evq, s.stream3));
  CU(cudaStreamWaitEvent(s.stream, s.evq, 0));          // the kernel starts behind the reset of `ready`
  int rc = launch_valid_feed<T>(w, s, (const T*)s.in0.p, m, (uint8_t*)s.out[0].p, s.ready, 0);
  if (rc) return rc;
  std::atomic<int> enq[MAX_FEED_CHUNKS];
  for (int c = 0; c < nchunks; c++) enq[c].store(0, std::memory_order_relaxed);
  std::atomic<int> err(0);
  const int nthreads = std::max(1, std::min(nchunks, staging_threads()));
  auto work = [&](int t) {
    if (cudaSetDevice(s.dev) != cudaSuccess) err.store(1);
    for (int c = t; c < nchunks; c += nthreads) {
      if (c >= R) {   // the staging buffer is free once the DMA of its previous chunk has completed
        while (!enq[c - R].load(std::memory_order_acquire)) std::this_thread::yield();
        if (cudaEventSynchronize(s.feed_ev[c - R]) != cudaSuccess) err.store(1);
      }
      char* st = s.stage + (size_t)(c % R) * cbytes;
      const int64_t c0 = (int64_t)c * ck, len = std::min(ck, m - c0);
      stage_copy(st, q + 7 * c0, (size_t)len * 7 * sizeof(T));
      if (cudaMemcpyAsync((T*)s.in0.p + 7 * c0, st, (size_t)len * 7 * sizeof(T), cudaMemcpyHostToDevice, s.stream2) != cudaSuccess ||
          cudaEventRecord(s.feed_ev[c], s.stream2) != cudaSuccess)
        err.store(1);
      enq[c].store(1, std::memory_order_release);
    }
  };
  std::atomic<int> remaining(0);
  const std::function<void(int)> job = work;
  if (!stage_pool().run(nthreads, job, remaining)) {
    // no thread to be had: this thread stages everything itself (the DMA does not depend on `ready`)
    for (int t = 0; t < nthreads; t++) work(t);
  }
  StreamWriteValue64 wv = stream_write_value64();
  cudaError_t ce = cudaSuccess;
  for (int c = 0; c < nchunks; c++) {
    while (!enq[c].load(std::memory_order_acquire)) std::this_thread::yield();
    const int64_t upto = std::min((int64_t)(c + 1) * ck, m);
    if (ce == cudaSuccess) ce = cudaStreamWaitEvent(s.stream3, s.feed_ev[c], 0);
    s.h_ready[c] = upto;
    if (ce == cudaSuccess && (!wv || wv(s.stream3, (unsigned long long)(uintptr_t)s.ready, (unsigned long long)upto, 0) != 0))
      ce = cudaMemcpyAsync(s.ready, s.h_ready + c, sizeof(long long), cudaMemcpyHostToDevice, s.stream3);
  }
  while (remaining.load(std::memory_order_acquire) > 0) std::this_thread::yield();   // `work` reads this frame
  if (ce != cudaSuccess || err.load()) {
    // the kernel may still be waiting for input that will not come: release it before reporting
    long long all = m;
    cudaMemcpyAsync(s.ready, &all, sizeof(long long), cudaMemcpyHostToDevice, s.stream3);
    cudaStreamSynchronize(s.stream3);
    cudaStreamSynchronize(s.stream);
    return fail(OSR_ECUDA, std::string("staged upload failed: ") + cudaGetErrorString(ce != cudaSuccess ? ce : cudaGetLastError()));
  }
  CU(cudaMemcpyAsync(out, s.out[0].p, (size_t)m, cudaMemcpyDeviceToHost, s.stream));
  CU(cudaEventRecord(s.evq, s.stream3));
  CU(cudaStreamWaitEvent(s.stream, s.evq, 0));
  return OSR_OK;
}

// `safe`: every upload is queued in front of the launch and the kernel starts behind the last one (it never
// polls) -- the form that cannot starve under launch-serialising tools, a time-sliced GPU or a slow source.
template <typename T>
static int valid_states_host_impl(osr_world* w, const T* q, int64_t n, uint8_t* out, bool safe) {
  int rc;
  std::lock_guard<std::mutex> lk(w->mu);
  int parts = (int)w->slots.size();
  const int64_t chunk = host_chunk_states();
  static const bool no_feed = getenv("OSR_NO_STREAM_FEED") != nullptr;  // A/B: chunked two-lane pipeline instead
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(n, parts, i, lo, hi);
    if (hi == lo) continue;
    CU(cudaSetDevice(s.dev));
    CU(s.in0.need((size_t)(hi - lo) * 7 * sizeof(T)));
    CU(s.out[0].need((size_t)(hi - lo)));
    int64_t nchunks = (hi - lo + chunk - 1) / chunk;
    const bool dfs_ok = s.dfs_variant >= 0 && !w->force_tile_kernel;
    if (!dfs_ok && !std::is_same<T, double>::value)
      return fail(OSR_EINVAL, "float32 states need the depth-first validity kernel (trees too deep for it)");
    if (dfs_ok && ((nchunks > 1 && !no_feed) || !std::is_same<T, double>::value)) {
      // Stream-fed launch: ONE persistent kernel over the whole range.  The upload is queued chunk by chunk on
      // the copy stream, each chunk followed by an 8-byte copy that advances the device word `ready`; the
      // kernel's warps pull states in index order and wait on that word, so compute runs under the PCIe
      // transfer and pays the kernel's ramp-up and tail once.  All copies are enqueued BEFORE the launch, so a
      // tool that serialises launches (ncu) cannot starve the kernel of its input.
      int64_t m = hi - lo;
      int64_t ck = std::max(chunk, (m + MAX_FEED_CHUNKS - 1) / MAX_FEED_CHUNKS);
      ck = (ck + 1023) / 1024 * 1024;
      if (!safe && staging_threads() > 0 && m > ck && is_pageable(q + 7 * lo)) {
        rc = feed_pageable<T>(w, s, q + 7 * lo, m, ck, out + lo);
        if (rc) return rc;
        continue;
      }
      // The word is advanced from a third stream that waits on an event behind each chunk: a write queued on
      // the copy stream itself would drain the copy engine's pipeline between chunks (~20 us per chunk).
      CU(cudaEventRecord(s.evp, s.stream));
      CU(cudaStreamWaitEvent(s.stream2, s.evp, 0));
      CU(cudaStreamWaitEvent(s.stream3, s.evp, 0));
      CU(cudaMemsetAsync(s.ready, 0, sizeof(long long), s.stream3));
      StreamWriteValue64 wv = stream_write_value64();
      // Chunk schedule.  The kernel is faster than the upload, so the call costs (upload) + (kernel work on the LAST
      // chunk) + (a few us of event / flag traffic per chunk): uniform 64 Ki chunks paid 16 of those gaps and 30 us
      // of tail.  Default: a first chunk of 64 Ki states (the kernel starts behind it; with float32 states the upload
      // is no faster than the kernel and a late start would show), then every chunk is half of what is left, down
      // to 8 Ki states -- 9 chunks for 2^20 states.  OSR_HOST_CHUNK set = uniform chunks of that size (A/B).
      static const bool uniform_chunks = getenv("OSR_HOST_CHUNK") != nullptr;
      static const bool trace = getenv("OSR_FEED_TRACE") != nullptr;   // development: where a call's time goes
      cudaEvent_t* tr = s.trace_ev;   // per device slot
      const auto host_t0 = std::chrono::steady_clock::now();
      if (trace) {
        for (int k = 0; k < 3; k++)
          if (!tr[k]) CU(cudaEventCreate(&tr[k]));
        CU(cudaEventRecord(tr[0], s.stream2));
      }
      static const int64_t min_chunk = [] { const char* e = getenv("OSR_FEED_MIN_CHUNK"); return e ? std::max<int64_t>(1024, atoll(e)) : (int64_t)8192; }();
      int64_t first_len = 0;
      int c = 0;
      for (int64_t c0 = 0, len = 0; c0 < m; c0 += len, c++) {
        if (uniform_chunks || c == MAX_FEED_CHUNKS - 1) len = c == MAX_FEED_CHUNKS - 1 ? m - c0 : std::min(ck, m - c0);
        else {
          len = c == 0 ? std::min<int64_t>(65536, m) : std::max(min_chunk, ((m - c0) / 2 + 1023) / 1024 * 1024);
          if (m - c0 - len < min_chunk / 2) len = m - c0;
        }
        if (c == 0) first_len = len;
        if (!s.feed_ev[c]) CU(cudaEventCreateWithFlags(&s.feed_ev[c], cudaEventDisableTiming));
        CU(cudaMemcpyAsync((T*)s.in0.p + 7 * c0, q + 7 * (lo + c0), (size_t)len * 7 * sizeof(T), cudaMemcpyHostToDevice, s.stream2));
        CU(cudaEventRecord(s.feed_ev[c], s.stream2));
        CU(cudaStreamWaitEvent(s.stream3, s.feed_ev[c], 0));
        s.h_ready[c] = c0 + len;
        if (!wv || wv(s.stream3, (unsigned long long)(uintptr_t)s.ready, (unsigned long long)(c0 + len), 0) != 0)
          CU(cudaMemcpyAsync(s.ready, s.h_ready + c, sizeof(long long), cudaMemcpyHostToDevice, s.stream3));
      }
      static const bool wait_all_env = getenv("OSR_FEED_WAIT_ALL") != nullptr;  // probe: kernel starts after the last chunk
      const bool wait_all = wait_all_env || safe;
      CU(cudaStreamWaitEvent(s.stream, s.feed_ev[wait_all ? c - 1 : 0], 0));
      rc = launch_valid_feed<T>(w, s, (const T*)s.in0.p, m, (uint8_t*)s.out[0].p, s.ready,
                                (long long)(wait_all ? m : first_len));  // the launch waits for that event
      if (rc) return rc;
      CU(cudaMemcpyAsync(out + lo, s.out[0].p, (size_t)m, cudaMemcpyDeviceToHost, s.stream));
      CU(cudaEventRecord(s.evq, s.stream3));
      CU(cudaStreamWaitEvent(s.stream, s.evq, 0));
      if (trace) {
        CU(cudaEventRecord(tr[1], s.stream2));
        CU(cudaEventRecord(tr[2], s.stream));
        const auto host_t1 = std::chrono::steady_clock::now();
        CU(cudaStreamSynchronize(s.stream));
        CU(cudaStreamSynchronize(s.stream2));
        const auto host_t2 = std::chrono::steady_clock::now();
        float up = 0, k0 = 0, ks = 0, tail = 0, d2h = 0, all = 0;
        cudaEventElapsedTime(&up, tr[0], tr[1]);
        cudaEventElapsedTime(&k0, tr[0], s.ev0);
        cudaEventElapsedTime(&ks, s.ev0, s.ev1);
        cudaEventElapsedTime(&tail, tr[1], s.ev1);
        cudaEventElapsedTime(&d2h, s.ev1, tr[2]);
        cudaEventElapsedTime(&all, tr[0], tr[2]);
        fprintf(stderr, "[feed trace] m=%lld chunks=%d: host enqueue %.0f us, host total %.0f us | device: upload %.0f us, "
                "kernel starts +%.0f us, kernel span %.0f us, last chunk -> kernel end %.0f us, kernel end -> labels on "
                "host %.0f us, first copy -> labels %.0f us\n", (long long)m, c,
                std::chrono::duration<double, std::micro>(host_t1 - host_t0).count(),
                std::chrono::duration<double, std::micro>(host_t2 - host_t0).count(), up * 1e3, k0 * 1e3, ks * 1e3,
                tail * 1e3, d2h * 1e3, all * 1e3);
      }
      continue;
    }
    if (nchunks > 1) {  // stream2 joins after whatever the main stream still has queued
      CU(cudaEventRecord(s.evp, s.stream));
      CU(cudaStreamWaitEvent(s.stream2, s.evp, 0));
    }
    for (int64_t c = 0; c < nchunks; c++) {
      int64_t c0 = c * chunk, m = std::min(chunk, hi - lo - c0);
      int lane = nchunks == 1 ? 1 : 1 + (int)(c & 1);
      cudaStream_t st = lane == 2 ? s.stream2 : s.stream;
      double* d_in = (double*)s.in0.p + 7 * c0;  // (T == double here: float32 input always takes the fed path)
      uint8_t* d_out = (uint8_t*)s.out[0].p + c0;
      CU(cudaMemcpyAsync(d_in, (const void*)(q + 7 * (lo + c0)), (size_t)m * 56, cudaMemcpyHostToDevice, st));
      rc = launch_valid(w, s, d_in, m, d_out, lane);
      if (rc) return rc;
      CU(cudaMemcpyAsync(out + lo + c0, d_out, (size_t)m, cudaMemcpyDeviceToHost, st));
    }
    if (nchunks > 1) {
      CU(cudaEventRecord(s.evp, s.stream2));
      CU(cudaStreamWaitEvent(s.stream, s.evp, 0));
    }
  }
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    rc = check_device_error(w, s);
    if (rc) return rc;
  }
  return OSR_OK;
}

// Small host calls (a planner or a Python loop asking for a handful of states): what counts is the round trip, not
// throughput.  The states are copied into a mapped page-locked buffer, ONE launch reads them over PCIe and writes the
// labels into a second mapped buffer, one synchronisation: a memset, a launch and a wait instead of two copies, a
// memset, two events, a launch, a flag copy and a wait (72 -> ~35 us for one state).  As many blocks as the states
// need (a block preloads both trees: 129 KB), one state per lane.
static int valid_small(osr_world* w, const double* q, int64_t n, uint8_t* out) {
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  memcpy(s.small_in, q, (size_t)n * 56);
  static const int64_t tile_n = [] { const char* e = getenv("OSR_SMALL_TILE_N"); return e ? (int64_t)atoll(e) : SMALL_STATES; }();
  bool tile_kernel = n <= tile_n;
  if (tile_kernel) {
    // a handful of states: one WARP per state on the warp-shared frontier of the tile kernel (32 lanes on one pose's
    // node pairs) -- the lane-owned traversal would walk each pose's tree with a single lane (40 us for one state)
    osr::KParams P = w->params(s);
    P.work_counter = nullptr;                           // a warp per tile (the grid below covers every tile)
    int ppt = 1;                                        // poses per warp: one while the warps fit the GPU at once
    while (ppt < 32 && (n + ppt - 1) / ppt > (int64_t)s.grid_valid * osr::WARPS_PER_BLOCK) ppt *= 2;
    const int64_t tiles = (n + ppt - 1) / ppt;
    const int grid = (int)std::min<int64_t>(s.grid_valid, (tiles + osr::WARPS_PER_BLOCK - 1) / osr::WARPS_PER_BLOCK);
    if (tiles > (int64_t)grid * osr::WARPS_PER_BLOCK) {  // (cannot happen below SMALL_STATES; kept for safety)
      CU(cudaMemsetAsync(s.work, 0, 2 * sizeof(unsigned long long), s.stream));
      P.work_counter = s.work;
    }
    // few poses per warp: a round descends BOTH trees (half the rounds; OSR_SMALL_BOTH=0 restores one tree a round)
    static const int both_max = [] { const char* e = getenv("OSR_SMALL_BOTH"); return e ? atoi(e) : 4; }();
    if (ppt <= both_max)
      osr::k_is_valid<true><<<grid, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(P, s.small_in, (long long)n, s.small_out, ppt);
    else
      osr::k_is_valid<false><<<grid, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(P, s.small_in, (long long)n, s.small_out, ppt);
  } else {
    CU(cudaMemsetAsync(s.work, 0, 2 * sizeof(unsigned long long), s.stream));
    osr::DfsParams D = dfs_params(w, s);
    D.work_counter = s.work;
    const int grid = (int)std::min<int64_t>(s.num_sms, (n + osr::DFS_THREADS - 1) / osr::DFS_THREADS);
    launch_dfs_variant<false, double>(s, D, s.small_in, (long long)n, s.small_out, s.stream, grid);
  }
  CU(cudaGetLastError());
  w->host_counters.kernel_launches++;
  if (tile_kernel) {                     // the tile kernel can exhaust its spill area: read the flag with the labels
    int rc = check_device_error(w, s);
    if (rc) return rc;
  } else {
    CU(cudaStreamSynchronize(s.stream));   // (the resident lane-owned kernel raises no device error flag)
  }
  memcpy(out, s.small_out, (size_t)n);
  return OSR_OK;
}

template <typename T>
static int valid_states_host(osr_world* w, const T* q, int64_t n, uint8_t* out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!q || !out))) return fail(OSR_EINVAL, "bad arguments");
  if (n == 0) return OSR_OK;
  if (std::is_same<T, double>::value && n <= SMALL_STATES && w->slots[0].dfs_variant >= 0 && !w->force_tile_kernel) {
    static const bool no_small = getenv("OSR_NO_SMALL_PATH") != nullptr;   // A/B
    if (!no_small) return valid_small(w, (const double*)q, n, out);
  }
  g_feed_timed_out = false;
  rc = valid_states_host_impl<T>(w, q, n, out, launches_may_block());
  if (rc == OSR_ECUDA && g_feed_timed_out) {   // the fed kernel starved: same call, nothing to wait for
    g_feed_timed_out = false;
    for (auto& s : w->slots) {
      if (cudaSetDevice(s.dev) == cudaSuccess) cudaStreamSynchronize(s.stream);
    }
    rc = valid_states_host_impl<T>(w, q, n, out, true);
  }
  return rc;
}

extern "C" {

int osr_is_valid_states(osr_world* w, const double* q, int64_t n, uint8_t* out) {
  return valid_states_host<double>(w, q, n, out);
}
int osr_is_valid_states_f32(osr_world* w, const float* q, int64_t n, uint8_t* out) {
  return valid_states_host<float>(w, q, n, out);
}

int osr_is_valid_states_device(osr_world* w, const double* d_q, int64_t n, uint8_t* d_out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n <= 0) return n == 0 ? OSR_OK : fail(OSR_EINVAL, "bad n");
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  return launch_valid(w, s, d_q, n, d_out);
}

// ---- segments ---------------------------------------------------------------------------------
// Host buffers in, host buffers out.  Every device gets a contiguous index range; inside a range the segments move
// in chunks: the upload of chunk c runs on a copy stream while the kernel of chunk c - 1 works, and the download of
// chunk c - 1 is issued after the kernel of chunk c has been queued -- so that a device always has its next kernel
// waiting, also when the caller's arrays are pageable and every copy blocks this thread until it is done (then the
// copies used to be 15 % of a 2^20-segment touch call, and the devices of one handle took turns).
static int transit_host(osr_world* w, const double* from, const double* to, int64_t n, const double* vd,
                        int64_t vd_count, double vd_scalar, int transition_mode, double* last_free,
                        double* first_hit, uint8_t* is_free, double* free_tau, double* touch_tau, int64_t* steps) {
  const int parts = (int)w->slots.size();
  int rc;
  std::vector<int64_t> los(parts, 0), ms(parts, 0), cks(parts, 0);
  std::vector<int> ncs(parts, 0);
  int rounds = 0;
  const bool per_seg_vd = vd && vd_count > 1;
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(n, parts, i, lo, hi);
    const int64_t m = hi - lo;
    los[i] = lo; ms[i] = m;
    if (m == 0) continue;
    // Few chunks: a chunk's kernel ends with its slowest segment (walks that graze an obstacle take 100x the
    // average; measured 2-3 ms per extra chunk), so a range is split only where the copies it hides are worth it:
    // estimated copy time (pageable arrays move at ~9 GB/s) / 10 ms chunks, at most 16.  Two or three become
    // 1/8, 3/4, 1/8 of the range, which exposes only a small first upload and last download.  OSR_SEG_CHUNKS=1
    // forces the single launch (A/B).
    static const int max_chunks = [] { const char* e = getenv("OSR_SEG_CHUNKS"); return e ? std::max(1, std::min(16, atoi(e))) : 16; }();
    const double bytes_per_seg = 112.0 + (per_seg_vd ? 8 : 0) + (last_free ? 56 : 0) + (first_hit ? 56 : 0) + (is_free ? 1 : 0) +
                                 (free_tau ? 8 : 0) + (touch_tau ? 8 : 0) + (steps ? 8 : 0);
    const int want = (int)std::min<double>(16.0, (double)m * bytes_per_seg / 9.0e6 / 10.0);
    ncs[i] = std::min(max_chunks, want <= 1 ? 1 : (want <= 3 ? 3 : want));
    cks[i] = (m + ncs[i] - 1) / ncs[i];     // (uniform; the three-chunk split is in chunk_range)
    rounds = std::max(rounds, ncs[i]);
    CU(cudaSetDevice(s.dev));
    CU(s.in0.need((size_t)m * 56));
    CU(s.in1.need((size_t)m * 56));
    if (per_seg_vd) CU(s.vd.need((size_t)m * 8));
    if (last_free) CU(s.out[0].need((size_t)m * 56));
    if (first_hit) CU(s.out[1].need((size_t)m * 56));
    if (is_free) CU(s.out[2].need((size_t)m));
    if (free_tau) CU(s.out[3].need((size_t)m * 8));
    if (touch_tau) CU(s.out[4].need((size_t)m * 8));
    if (steps) CU(s.out[5].need((size_t)m * 8));
    for (int c = 0; c < 2 * ncs[i]; c++)
      if (!s.feed_ev[c]) CU(cudaEventCreateWithFlags(&s.feed_ev[c], cudaEventDisableTiming));
    CU(cudaEventRecord(s.evp, s.stream));              // the copy streams join behind what the main stream holds
    CU(cudaStreamWaitEvent(s.stream2, s.evp, 0));
    CU(cudaStreamWaitEvent(s.stream3, s.evp, 0));
  }
  auto chunk_range = [&](int i, int c, int64_t& c0, int64_t& len) {
    if (ncs[i] == 3) {
      const int64_t e = (ms[i] / 8 + 1023) / 1024 * 1024;
      c0 = c == 0 ? 0 : (c == 1 ? e : ms[i] - e);
      len = c == 1 ? ms[i] - 2 * e : e;
    } else {
      c0 = (int64_t)c * cks[i];
      len = std::min(cks[i], ms[i] - c0);
    }
  };
  auto download = [&](int i, int c) -> int {
    Slot& s = w->slots[i];
    int64_t c0, len;
    chunk_range(i, c, c0, len);
    const int64_t g0 = los[i] + c0;
    CU(cudaSetDevice(s.dev));
    CU(cudaStreamWaitEvent(s.stream3, s.feed_ev[2 * c + 1], 0));
    if (last_free) CU(cudaMemcpyAsync(last_free + 7 * g0, (double*)s.out[0].p + 7 * c0, (size_t)len * 56, cudaMemcpyDeviceToHost, s.stream3));
    if (first_hit) CU(cudaMemcpyAsync(first_hit + 7 * g0, (double*)s.out[1].p + 7 * c0, (size_t)len * 56, cudaMemcpyDeviceToHost, s.stream3));
    if (is_free) CU(cudaMemcpyAsync(is_free + g0, (uint8_t*)s.out[2].p + c0, (size_t)len, cudaMemcpyDeviceToHost, s.stream3));
    if (free_tau) CU(cudaMemcpyAsync(free_tau + g0, (double*)s.out[3].p + c0, (size_t)len * 8, cudaMemcpyDeviceToHost, s.stream3));
    if (touch_tau) CU(cudaMemcpyAsync(touch_tau + g0, (double*)s.out[4].p + c0, (size_t)len * 8, cudaMemcpyDeviceToHost, s.stream3));
    if (steps) CU(cudaMemcpyAsync(steps + g0, (long long*)s.out[5].p + c0, (size_t)len * 8, cudaMemcpyDeviceToHost, s.stream3));
    return OSR_OK;
  };
  for (int c = 0; c <= rounds; c++) {
    for (int i = 0; i < parts && c < rounds; i++) {      // upload + kernel of chunk c on every device
      if (c >= ncs[i]) continue;
      Slot& s = w->slots[i];
      int64_t c0, len;
      chunk_range(i, c, c0, len);
      const int64_t g0 = los[i] + c0;
      CU(cudaSetDevice(s.dev));
      CU(cudaMemcpyAsync((double*)s.in0.p + 7 * c0, from + 7 * g0, (size_t)len * 56, cudaMemcpyHostToDevice, s.stream2));
      CU(cudaMemcpyAsync((double*)s.in1.p + 7 * c0, to + 7 * g0, (size_t)len * 56, cudaMemcpyHostToDevice, s.stream2));
      const double* d_vd = nullptr;
      if (per_seg_vd) {
        CU(cudaMemcpyAsync((double*)s.vd.p + c0, vd + g0, (size_t)len * 8, cudaMemcpyHostToDevice, s.stream2));
        d_vd = (const double*)s.vd.p + c0;
      }
      CU(cudaEventRecord(s.feed_ev[2 * c], s.stream2));
      CU(cudaStreamWaitEvent(s.stream, s.feed_ev[2 * c], 0));
      osr::SegOut O = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
      if (last_free) O.last_free = (double*)s.out[0].p + 7 * c0;
      if (first_hit) O.first_hit = (double*)s.out[1].p + 7 * c0;
      if (is_free) O.is_free = (uint8_t*)s.out[2].p + c0;
      if (free_tau) O.free_tau = (double*)s.out[3].p + c0;
      if (touch_tau) O.touch_tau = (double*)s.out[4].p + c0;
      if (steps) O.steps = (long long*)s.out[5].p + c0;
      rc = launch_transit(w, s, (const double*)s.in0.p + 7 * c0, (const double*)s.in1.p + 7 * c0, len, d_vd, d_vd ? len : 1,
                          (vd && vd_count == 1) ? vd[0] : vd_scalar, transition_mode, O);
      if (rc) return rc;
      CU(cudaEventRecord(s.feed_ev[2 * c + 1], s.stream));
    }
    for (int i = 0; i < parts && c > 0; i++) {           // download of chunk c - 1
      if (c - 1 >= ncs[i]) continue;
      rc = download(i, c - 1);
      if (rc) return rc;
    }
  }
  for (int i = 0; i < parts; i++) {
    if (ms[i] == 0) continue;
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    CU(cudaEventRecord(s.evq, s.stream3));
    CU(cudaStreamWaitEvent(s.stream, s.evq, 0));
    rc = check_device_error(w, s);
    if (rc) return rc;
  }
  return OSR_OK;
}

// Small segment calls (a Python loop walking one ray at a time): inputs and outputs in one mapped page-locked arena,
// one launch with as many blocks as the segments need, one wait -- no staging copies, no events.
static int transit_small(osr_world* w, const double* from, const double* to, int64_t n, const double* vd,
                         int64_t vd_count, double vd_scalar, int transition_mode, double* last_free,
                         double* first_hit, uint8_t* is_free, double* free_tau, double* touch_tau, int64_t* steps) {
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  unsigned char* a = s.small_seg;     // layout (n segments): from | to | vd | last_free | first_hit | taus | steps | flags
  double* m_from = (double*)a;
  double* m_to = m_from + 7 * n;
  double* m_vd = m_to + 7 * n;
  double* m_lf = m_vd + n;
  double* m_fh = m_lf + 7 * n;
  double* m_ft = m_fh + 7 * n;
  double* m_tt = m_ft + n;
  long long* m_st = (long long*)(m_tt + n);
  uint8_t* m_if = (uint8_t*)(m_st + n);
  memcpy(m_from, from, (size_t)n * 56);
  memcpy(m_to, to, (size_t)n * 56);
  const bool per_seg = vd && vd_count > 1;
  if (per_seg) memcpy(m_vd, vd, (size_t)n * 8);
  osr::SegOut O = {last_free ? m_lf : nullptr, first_hit ? m_fh : nullptr, is_free ? m_if : nullptr,
                   free_tau ? m_ft : nullptr, touch_tau ? m_tt : nullptr, steps ? m_st : nullptr};
  osr::KParams P = w->params(s);
  CU(cudaMemsetAsync(s.work, 0, 2 * sizeof(unsigned long long), s.stream));
  const int grid = (int)std::min<int64_t>(s.grid_transit, (n + osr::WARPS_PER_BLOCK - 1) / osr::WARPS_PER_BLOCK);
  const double vds = (vd && vd_count == 1) ? vd[0] : vd_scalar;
  if (s.transit_minb == 4)
    osr::k_transit<4><<<grid, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, m_from, m_to, (long long)n, per_seg ? m_vd : nullptr, per_seg ? (long long)n : 1LL, vds, transition_mode, O);
  else if (s.transit_minb == 3)
    osr::k_transit<3><<<grid, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, m_from, m_to, (long long)n, per_seg ? m_vd : nullptr, per_seg ? (long long)n : 1LL, vds, transition_mode, O);
  else
    osr::k_transit<2><<<grid, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, m_from, m_to, (long long)n, per_seg ? m_vd : nullptr, per_seg ? (long long)n : 1LL, vds, transition_mode, O);
  CU(cudaGetLastError());
  w->host_counters.kernel_launches++;
  int rc = check_device_error(w, s);   // the flag and the wait
  if (rc) return rc;
  if (last_free) memcpy(last_free, m_lf, (size_t)n * 56);
  if (first_hit) memcpy(first_hit, m_fh, (size_t)n * 56);
  if (is_free) memcpy(is_free, m_if, (size_t)n);
  if (free_tau) memcpy(free_tau, m_ft, (size_t)n * 8);
  if (touch_tau) memcpy(touch_tau, m_tt, (size_t)n * 8);
  if (steps) memcpy(steps, m_st, (size_t)n * 8);
  return OSR_OK;
}
static bool small_segments_ok(const osr_world* w, int64_t n) {
  static const bool off = getenv("OSR_NO_SMALL_PATH") != nullptr;
  return !off && n <= SMALL_SEGMENTS && !w->seg_dfs;
}

int osr_transit_states_to_with_contact(osr_world* w, const double* from, const double* to, int64_t n,
                                       const double* vd, int64_t vd_count, double* last_free, double* first_hit,
                                       uint8_t* is_free, double* free_tau, double* touch_tau, int64_t* steps) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!from || !to || !vd)) || (vd_count != 1 && vd_count != n))
    return fail(OSR_EINVAL, "bad arguments");
  if (n == 0) return OSR_OK;
  // verify_delta < 0 never terminates in the reference (unit_world.cc:445-470); 0 is legal only on zero-length
  // segments (the kernel reports those that are not); NaN falls through the reference's comparisons and is passed on
  for (int64_t i = 0; i < vd_count; i++)
    if (vd[i] < 0.0) return fail(OSR_EINVAL, "verify_delta must not be negative");
  std::lock_guard<std::mutex> lk(w->mu);
  if (small_segments_ok(w, n))
    return transit_small(w, from, to, n, vd, vd_count, 0.0, 0, last_free, first_hit, is_free, free_tau, touch_tau, steps);
  return transit_host(w, from, to, n, vd, vd_count, 0.0, 0, last_free, first_hit, is_free, free_tau, touch_tau, steps);
}

int osr_is_valid_transitions(osr_world* w, const double* from, const double* to, int64_t n,
                             double initial_verify_delta, uint8_t* out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!from || !to || !out))) return fail(OSR_EINVAL, "bad arguments");
  if (initial_verify_delta < 0.0) return fail(OSR_EINVAL, "verify_delta must not be negative");
  if (n == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  if (small_segments_ok(w, n))
    return transit_small(w, from, to, n, nullptr, 1, initial_verify_delta, 1, nullptr, nullptr, out, nullptr, nullptr, nullptr);
  return transit_host(w, from, to, n, nullptr, 1, initial_verify_delta, 1, nullptr, nullptr, out, nullptr, nullptr,
                      nullptr);
}

int osr_transit_states_device(osr_world* w, const double* d_from, const double* d_to, int64_t n, double vd,
                              double* d_last_free, double* d_first_hit, uint8_t* d_is_free, double* d_free_tau,
                              double* d_touch_tau, int64_t* d_steps) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (vd < 0.0) return fail(OSR_EINVAL, "verify_delta must not be negative");
  if (n <= 0) return n == 0 ? OSR_OK : fail(OSR_EINVAL, "bad n");
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  osr::SegOut O = {d_last_free, d_first_hit, d_is_free, d_free_tau, d_touch_tau, (long long*)d_steps};
  return launch_transit(w, s, d_from, d_to, n, nullptr, 1, vd, 0, O);
}
int osr_is_valid_transitions_device(osr_world* w, const double* d_from, const double* d_to, int64_t n,
                                    double initial_verify_delta, uint8_t* d_out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (initial_verify_delta < 0.0) return fail(OSR_EINVAL, "verify_delta must not be negative");
  if (n <= 0) return n == 0 ? OSR_OK : fail(OSR_EINVAL, "bad n");
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  osr::SegOut O = {nullptr, nullptr, d_out, nullptr, nullptr, nullptr};
  return launch_transit(w, s, d_from, d_to, n, nullptr, 1, initial_verify_delta, 1, O);
}

// Allocation-free variant of osr_is_valid_transitions_device (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops).  Same kernel, same
// labels; K2's scratch -- the work-queue word and the per-warp spill areas of the traversal stack -- lives in the
// static array g_noalloc_scratch instead of the handle's cudaMalloc'ed buffers.  Calls no cudaMalloc / cudaFree and
// no allocating library routine: a memset of the 256-byte header, two event records, one launch.  Every call (of every
// handle in the process) on one GPU shares the array: calls that overlap on that GPU race.
int osr_is_valid_transitions_device_noalloc(osr_world* w, const double* d_from, const double* d_to, int64_t n,
                                            double initial_verify_delta, uint8_t* d_out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (initial_verify_delta < 0.0) return fail(OSR_EINVAL, "verify_delta must not be negative");
  if (n < 0 || (n > 0 && (!d_from || !d_to || !d_out))) return fail(OSR_EINVAL, "bad arguments");
  if (n > (int64_t)OSR_NOALLOC_MAX_SEGMENTS)
    return fail(OSR_EINVAL, "osr_is_valid_transitions_device_noalloc: more than OSR_NOALLOC_MAX_SEGMENTS (" +
                                std::to_string((long long)OSR_NOALLOC_MAX_SEGMENTS) + ") segments in one call");
  if (n == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  if (!s.noalloc) return fail(OSR_ECUDA, "static scratch not resolved on this device");
  if ((size_t)s.grid_transit * osr::WARPS_PER_BLOCK > (size_t)OSR_NOALLOC_MAX_WARPS)
    return fail(OSR_EOVERFLOW, "osr_is_valid_transitions_device_noalloc: the launch grid has more warps than the static "
                               "scratch holds spill areas for");
  osr::KParams P = w->params(s);   // error flag and counters stay the handle's (osr_sync, osr_get_counters)
  P.work_counter = reinterpret_cast<unsigned long long*>(s.noalloc);
  P.spill = reinterpret_cast<unsigned long long*>(s.noalloc + OSR_NOALLOC_HEADER_BYTES);
  P.spill_cap = OSR_NOALLOC_SPILL_CAP;
  // the array keeps the last call's contents: the queue word must start at zero, the spill areas need no clearing
  CU(cudaMemsetAsync(s.noalloc, 0, OSR_NOALLOC_HEADER_BYTES, s.stream));
  osr::SegOut O = {nullptr, nullptr, d_out, nullptr, nullptr, nullptr};
  CU(cudaEventRecord(s.ev0, s.stream));
  if (s.transit_minb == 4)
    osr::k_transit<4><<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, d_from, d_to, (long long)n, nullptr, 1LL, initial_verify_delta, 1, O);
  else if (s.transit_minb == 3)
    osr::k_transit<3><<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, d_from, d_to, (long long)n, nullptr, 1LL, initial_verify_delta, 1, O);
  else
    osr::k_transit<2><<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, d_from, d_to, (long long)n, nullptr, 1LL, initial_verify_delta, 1, O);
  CU(cudaGetLastError());
  CU(cudaEventRecord(s.ev1, s.stream));
  w->host_counters.kernel_launches++;
  return OSR_OK;
}

// Page-locked host memory for sample / result arrays: a host call whose buffers are pageable is bounded by the
// driver's staging copies (~10 GB/s), one on pinned buffers by PCIe.
// The block is anonymous memory first touched on the NUMA node of the CURRENT device (NumaScope) and then page-locked
// with cudaHostRegister; cudaHostAlloc is the fallback.
static std::mutex g_host_mu;
static std::map<void*, size_t> g_host_blocks;   // registered blocks -> mapped size
void* osr_host_alloc(size_t bytes) {
  if (bytes == 0) {
    g_err = "osr_host_alloc: zero bytes";
    return nullptr;
  }
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    g_err = "osr_host_alloc: no CUDA device";
    return nullptr;
  }
  {
    NumaScope numa(dev);
    if (numa.changed) {
      const size_t len = (bytes + 4095) & ~(size_t)4095;
      void* p = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
      if (p != MAP_FAILED) {
        memset(p, 0, len);   // first touch: on this thread's (the device's) node
        if (cudaHostRegister(p, len, cudaHostRegisterPortable) == cudaSuccess) {
          std::lock_guard<std::mutex> lk(g_host_mu);
          g_host_blocks[p] = len;
          return p;
        }
        cudaGetLastError();
        munmap(p, len);
      }
    }
    void* p = nullptr;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) == cudaSuccess) return p;
  }
  cudaGetLastError();
  g_err = "cudaHostAlloc failed";
  return nullptr;
}
void osr_host_free(void* p) {
  if (!p) return;
  size_t len = 0;
  {
    std::lock_guard<std::mutex> lk(g_host_mu);
    auto it = g_host_blocks.find(p);
    if (it != g_host_blocks.end()) {
      len = it->second;
      g_host_blocks.erase(it);
    }
  }
  if (len) {
    cudaHostUnregister(p);
    munmap(p, len);
  } else {
    cudaFreeHost(p);
  }
}

void* osr_stream(osr_world* w) { return (w && !w->slots.empty()) ? (void*)w->slots[0].stream : nullptr; }
int osr_sync(osr_world* w) {
  int rc = check_ready(w);
  if (rc) return rc;
  for (auto& s : w->slots) {   // also surfaces the device error flag of launch-only (_device) calls
    CU(cudaSetDevice(s.dev));
    rc = check_device_error(w, s);
    if (rc) return rc;
  }
  return OSR_OK;
}
int osr_last_kernel_ms(osr_world* w, float* ms) {
  int rc = check_ready(w);
  if (rc) return rc;
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  CU(cudaEventSynchronize(s.ev1));
  CU(cudaEventElapsedTime(ms, s.ev0, s.ev1));
  return OSR_OK;
}

// ---- disentangled -----------------------------------------------------------------------------
int osr_is_disentangled_states(osr_world* w, const double* q, int64_t n, uint8_t* out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!q || !out))) return fail(OSR_EINVAL, "bad arguments");
  if (n == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  osr::BoxPair bp;
  for (int k = 0; k < 3; k++) {  // fcl AABB::center() = (min+max)*0.5, width = max-min
    bp.rob_c[k] = (w->rob.aabb_min[k] + w->rob.aabb_max[k]) * 0.5;
    bp.rob_side[k] = w->rob.aabb_max[k] - w->rob.aabb_min[k];
    bp.env_c[k] = (w->env.aabb_min[k] + w->env.aabb_max[k]) * 0.5;
    bp.env_side[k] = w->env.aabb_max[k] - w->env.aabb_min[k];
  }
  osr::Rigid env_tf;
  osr::state_to_rigid(w->perturb, env_tf);
  static const bool no_small = getenv("OSR_NO_SMALL_PATH") != nullptr;
  if (n <= SMALL_STATES && !no_small) {   // small call: mapped buffers, one launch, one wait (as valid_small)
    Slot& s = w->slots[0];
    CU(cudaSetDevice(s.dev));
    memcpy(s.small_in, q, (size_t)n * 56);
    osr::k_disentangled<<<(unsigned)((n + 255) / 256), 256, 0, s.stream>>>(env_tf, bp, s.small_in, (long long)n, s.small_out);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    CU(cudaStreamSynchronize(s.stream));
    memcpy(out, s.small_out, (size_t)n);
    return OSR_OK;
  }
  int parts = (int)w->slots.size();
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(n, parts, i, lo, hi);
    int64_t m = hi - lo;
    if (m == 0) continue;
    CU(cudaSetDevice(s.dev));
    CU(s.in0.need((size_t)m * 56));
    CU(s.out[0].need((size_t)m));
    CU(cudaMemcpyAsync(s.in0.p, q + 7 * lo, (size_t)m * 56, cudaMemcpyHostToDevice, s.stream));
    osr::k_disentangled<<<(unsigned)((m + 255) / 256), 256, 0, s.stream>>>(env_tf, bp, (const double*)s.in0.p,
                                                                           (long long)m, (uint8_t*)s.out[0].p);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    CU(cudaMemcpyAsync(out + lo, s.out[0].p, (size_t)m, cudaMemcpyDeviceToHost, s.stream));
  }
  for (auto& s : w->slots) {
    CU(cudaSetDevice(s.dev));
    CU(cudaStreamSynchronize(s.stream));
  }
  return OSR_OK;
}

// ---- state-frame maps (host, fp64) --------------------------------------------------------------
static void apply_perturbation(const double* pert, const double* in, double* out) {  // unit_world.cc:713-724
  double r[4] = {in[3], in[4], in[5], in[6]}, p[4] = {pert[3], pert[4], pert[5], pert[6]};
  osr::quat_normalize(r[0], r[1], r[2], r[3]);
  osr::quat_normalize(p[0], p[1], p[2], p[3]);
  double qr[4], R[9];
  osr::quat_mul(p, r, qr);
  osr::quat_to_rot(p[0], p[1], p[2], p[3], R);
  double t[3];
  for (int k = 0; k < 3; k++) t[k] = pert[k] + (R[3 * k] * in[0] + R[3 * k + 1] * in[1] + R[3 * k + 2] * in[2]);
  osr::quat_normalize(qr[0], qr[1], qr[2], qr[3]);
  out[0] = t[0]; out[1] = t[1]; out[2] = t[2];
  out[3] = qr[0]; out[4] = qr[1]; out[5] = qr[2]; out[6] = qr[3];
}
static void unapply_perturbation(const double* pert, const double* in, double* out) {  // unit_world.cc:726-738
  double r[4] = {in[3], in[4], in[5], in[6]}, p[4] = {pert[3], pert[4], pert[5], pert[6]};
  osr::quat_normalize(r[0], r[1], r[2], r[3]);
  osr::quat_normalize(p[0], p[1], p[2], p[3]);
  double n2 = p[0] * p[0] + p[1] * p[1] + p[2] * p[2] + p[3] * p[3];
  double pi[4] = {p[0] / n2, -p[1] / n2, -p[2] / n2, -p[3] / n2};  // Eigen inverse(): conj / squaredNorm
  double R[9];
  osr::quat_to_rot(pi[0], pi[1], pi[2], pi[3], R);
  double d[3] = {-pert[0] + in[0], -pert[1] + in[1], -pert[2] + in[2]}, t[3], qr[4];
  for (int k = 0; k < 3; k++) t[k] = R[3 * k] * d[0] + R[3 * k + 1] * d[1] + R[3 * k + 2] * d[2];
  osr::quat_mul(pi, r, qr);
  osr::quat_normalize(qr[0], qr[1], qr[2], qr[3]);
  out[0] = t[0]; out[1] = t[1]; out[2] = t[2];
  out[3] = qr[0]; out[4] = qr[1]; out[5] = qr[2]; out[6] = qr[3];
}
static void to_unit_one(const osr_world* w, const double* q, double* out) {  // unit_world.cc:526-539
  double s = (double)w->scene_scale;
  double tmp[7] = {s * q[0], s * q[1], s * q[2], q[3], q[4], q[5], q[6]};
  apply_perturbation(w->perturb, tmp, out);
}
int osr_translate_to_unit(const osr_world* w, const double* q, int64_t n, double* out) {
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  for (int64_t i = 0; i < n; i++) to_unit_one(w, q + 7 * i, out + 7 * i);
  return OSR_OK;
}
int osr_translate_from_unit(const osr_world* w, const double* q, int64_t n, double* out) {  // unit_world.cc:541-555
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  double inv = 1.0 / (double)w->scene_scale;
  for (int64_t i = 0; i < n; i++) {
    double tmp[7];
    unapply_perturbation(w->perturb, q + 7 * i, tmp);
    double* o = out + 7 * i;
    o[0] = inv * tmp[0]; o[1] = inv * tmp[1]; o[2] = inv * tmp[2];
    o[3] = tmp[3]; o[4] = tmp[4]; o[5] = tmp[5]; o[6] = tmp[6];
  }
  return OSR_OK;
}
// Eigen quaternion * vector: v + w*(2 u x v) + u x (2 u x v)
using osr::quat_rotate;
int osr_translate_ompl_to_unit(const osr_world* w, const double* qin, int64_t n, double* out) {  // unit_world.cc:689-711
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  double dc[3];
  for (int k = 0; k < 3; k++) dc[k] = (double)(w->ompl_centre[k] - w->centre[k]) * (double)w->scene_scale;
  for (int64_t i = 0; i < n; i++) {
    const double* q = qin + 7 * i;
    double wf[7] = {q[0], q[1], q[2], q[6], q[3], q[4], q[5]}, u[7];
    to_unit_one(w, wf, u);
    double r[4] = {u[3], u[4], u[5], u[6]}, rot[3];
    osr::quat_normalize(r[0], r[1], r[2], r[3]);
    quat_rotate(r, dc, rot);
    double* o = out + 7 * i;
    for (int k = 0; k < 3; k++) o[k] = u[k] - rot[k];
    o[3] = r[0]; o[4] = r[1]; o[5] = r[2]; o[6] = r[3];
  }
  return OSR_OK;
}
int osr_translate_unit_to_ompl(const osr_world* w, const double* qin, int64_t n, double* out) {  // unit_world.cc:557-591
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  double dc[3];
  for (int k = 0; k < 3; k++) dc[k] = (double)(w->ompl_centre[k] - w->centre[k]) * (double)w->scene_scale;
  for (int64_t i = 0; i < n; i++) {
    const double* q = qin + 7 * i;
    double r[4] = {q[3], q[4], q[5], q[6]}, rot[3];
    osr::quat_normalize(r[0], r[1], r[2], r[3]);
    quat_rotate(r, dc, rot);
    double shifted[7] = {q[0] + rot[0], q[1] + rot[1], q[2] + rot[2], r[0], r[1], r[2], r[3]}, v[7];
    osr_translate_from_unit(w, shifted, 1, v);
    double* o = out + 7 * i;
    o[0] = v[0]; o[1] = v[1]; o[2] = v[2];
    o[3] = v[4]; o[4] = v[5]; o[5] = v[6]; o[6] = v[3];
  }
  return OSR_OK;
}

double osr_distance(const double a[7], const double b[7]) { return osr::distance(a, b); }
void osr_interpolate(const double a[7], const double b[7], double tau, double out[7]) { osr::interpolate(a, b, tau, out); }
void osr_apply(const double from[7], const double tr[3], const double aa[3], double out[7]) { osr::apply(from, tr, aa, out); }
int osr_walk_deltas(double* delta, double verify_delta, double dist, int nsteps, double* last_valid) {
  double last = *delta;
  int valid = osr::advance_chain(*delta, verify_delta, dist, nsteps, last);
  if (last_valid) *last_valid = last;
  return valid;
}

// Host build of the two triangle-pair predicates of the traversal kernels (sat17.h), for the parity tests:
// the vertices are taken relative to P1 and rounded to fp32 exactly as the kernels do.
int osr_triangle_pair_predicates(const double* tris, int64_t n, uint8_t* filter, uint8_t* exact) {
  if (n < 0 || (n > 0 && !tris)) return fail(OSR_EINVAL, "bad arguments");
  for (int64_t i = 0; i < n; i++) {
    const double* t = tris + 18 * i;
    osr::TriPair<double> td;
    for (int k = 0; k < 3; k++) {
      td.p2[k] = t[3 + k] - t[k]; td.p3[k] = t[6 + k] - t[k];
      td.q1[k] = t[9 + k] - t[k]; td.q2[k] = t[12 + k] - t[k]; td.q3[k] = t[15 + k] - t[k];
    }
    osr::TriPair<float> tf;
    for (int k = 0; k < 3; k++) {
      tf.p2[k] = (float)td.p2[k]; tf.p3[k] = (float)td.p3[k];
      tf.q1[k] = (float)td.q1[k]; tf.q2[k] = (float)td.q2[k]; tf.q3[k] = (float)td.q3[k];
    }
    if (filter) filter[i] = (uint8_t)osr::sat17_filter_c(tf);
    if (exact) exact[i] = osr::sat17_exact_c(td) ? 1 : 0;
  }
  return OSR_OK;
}

// ---- visibility (batched isValidTransition) -----------------------------------------------------
static int visibility_common(osr_world* w, const double* q0, int64_t m, int unit0, const double* q1, int64_t n,
                             int unit1, double vmag, int8_t* out, bool pairwise) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (m <= 0 || n <= 0 || !q0 || !q1 || !out) return fail(OSR_EINVAL, "bad arguments");
  std::vector<double> u0((size_t)m * 7), u1((size_t)n * 7);
  if (unit0) memcpy(u0.data(), q0, (size_t)m * 56); else osr_translate_to_unit(w, q0, m, u0.data());
  if (unit1) memcpy(u1.data(), q1, (size_t)n * 56); else osr_translate_to_unit(w, q1, n, u1.data());
  int64_t total = pairwise ? std::max(m, n) : m * n;
  std::vector<double> fr((size_t)total * 7), to((size_t)total * 7);
  if (pairwise) {  // unit_world.cc:822-827: rows i % M and i % N
    for (int64_t i = 0; i < total; i++) {
      memcpy(&fr[(size_t)7 * i], &u0[(size_t)7 * (i % m)], 56);
      memcpy(&to[(size_t)7 * i], &u1[(size_t)7 * (i % n)], 56);
    }
  } else {
    for (int64_t i = 0; i < m; i++)
      for (int64_t j = 0; j < n; j++) {
        memcpy(&fr[(size_t)7 * (i * n + j)], &u0[(size_t)7 * i], 56);
        memcpy(&to[(size_t)7 * (i * n + j)], &u1[(size_t)7 * j], 56);
      }
  }
  std::vector<uint8_t> res((size_t)total);
  rc = osr_is_valid_transitions(w, fr.data(), to.data(), total, vmag, res.data());
  if (rc) return rc;
  for (int64_t i = 0; i < total; i++) out[i] = (int8_t)res[i];
  return OSR_OK;
}
int osr_visibility_matrix2(osr_world* w, const double* q0, int64_t m, int is_unit0, const double* q1, int64_t n,
                           int is_unit1, double vmag, int8_t* out) {
  return visibility_common(w, q0, m, is_unit0, q1, n, is_unit1, vmag, out, false);
}
int osr_visibility_pair(osr_world* w, const double* q0, int64_t m, int is_unit0, const double* q1, int64_t n,
                        int is_unit1, double vmag, int8_t* out) {
  return visibility_common(w, q0, m, is_unit0, q1, n, is_unit1, vmag, out, true);
}

// ---- contacts -----------------------------------------------------------------------------------
int osr_intersecting_segments(osr_world* w, const double* q, int64_t n, int64_t cap, int64_t* offsets, double* seg,
                              int32_t* faces, uint8_t* coplanar, double* bary, uint8_t* overflow, int64_t* total) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || cap < 0 || !offsets || (n > 0 && !q) || (cap > 0 && (!seg || !faces || !coplanar)))
    return fail(OSR_EINVAL, "bad arguments");
  std::lock_guard<std::mutex> lk(w->mu);
  offsets[0] = 0;
  if (overflow) memset(overflow, 0, (size_t)n);
  if (total) *total = 0;
  if (n == 0) return OSR_OK;
  // Pass 1 (every device, index-range shards): all intersecting (state, env tri, rob tri) records, then their
  // number and their FCL row count per state (K3c; the records stay on the device).
  int parts = (int)w->slots.size();
  std::vector<int64_t> los(parts), nrec(parts, 0);
  std::vector<std::vector<int>> h_cnt(parts), h_msum(parts);
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(n, parts, i, lo, hi);
    los[i] = lo;
    int64_t m = hi - lo;
    if (m == 0) continue;
    CU(cudaSetDevice(s.dev));
    CU(s.in0.need((size_t)m * 56));
    CU(cudaMemcpyAsync(s.in0.p, q + 7 * lo, (size_t)m * 56, cudaMemcpyHostToDevice, s.stream));
    unsigned long long rec_cap = (unsigned long long)std::max<int64_t>(2 * cap + 1024, 1 << 16);
    for (int attempt = 0; attempt < 2; attempt++) {
      CU(s.recs.need((size_t)rec_cap * sizeof(osr::ContactRec)));
      CU(cudaMemsetAsync(s.work, 0, 2 * sizeof(unsigned long long), s.stream));
      osr::KParams P = w->params(s);
      CU(cudaEventRecord(s.ev0, s.stream));
      osr::k_contacts<<<s.grid_contacts, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
          P, (const double*)s.in0.p, (long long)m, (osr::ContactRec*)s.recs.p, s.work + 1, rec_cap);
      CU(cudaGetLastError());
      CU(cudaEventRecord(s.ev1, s.stream));
      w->host_counters.kernel_launches++;
      unsigned long long found = 0;
      CU(cudaMemcpyAsync(&found, s.work + 1, 8, cudaMemcpyDeviceToHost, s.stream));
      rc = check_device_error(w, s);
      if (rc) return rc;
      if (found <= rec_cap) {
        nrec[i] = (int64_t)found;
        break;
      }
      if (attempt == 1) return fail(OSR_EOVERFLOW, "contact record buffer overflow");
      rec_cap = found + 1024;  // exact size known now: run again
    }
    h_cnt[i].assign((size_t)m, 0);
    h_msum[i].assign((size_t)m, 0);
    if (nrec[i] == 0) continue;
    CU(s.c_cnt.need((size_t)m * 4));
    CU(s.c_msum.need((size_t)m * 4));
    CU(cudaMemsetAsync(s.c_cnt.p, 0, (size_t)m * 4, s.stream));
    CU(cudaMemsetAsync(s.c_msum.p, 0, (size_t)m * 4, s.stream));
    osr::k_rec_count<<<(unsigned)((nrec[i] + 255) / 256), 256, 0, s.stream>>>(
        (const osr::ContactRec*)s.recs.p, (long long)nrec[i], (int*)s.c_cnt.p, (int*)s.c_msum.p);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    CU(cudaMemcpyAsync(h_cnt[i].data(), s.c_cnt.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaMemcpyAsync(h_msum[i].data(), s.c_msum.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s.stream));
  }
  for (auto& s : w->slots) {
    CU(cudaSetDevice(s.dev));
    CU(cudaStreamSynchronize(s.stream));
  }
  // Host: CSR offsets of the rows with the output capacity applied (n integers; the records are not touched).
  int64_t rows_found = 0, rows_written = 0;
  for (int i = 0; i < parts; i++)
    for (size_t j = 0; j < h_msum[i].size(); j++) {
      const int64_t st = los[i] + (int64_t)j, have = h_msum[i][j];
      rows_found += have;
      const int64_t take = std::min(have, std::max<int64_t>(cap - rows_written, 0));
      if (take < have && overflow) overflow[st] = 1;
      rows_written += take;
      offsets[st + 1] = rows_written;
    }
  if (total) *total = rows_found;
  // Pass 2 (per device): records into per-state buckets, ordered inside each bucket, row list laid out (K3c); then
  // one thread per output row (Moeller segment + optional barycentrics).
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    const int64_t m = (int64_t)h_cnt[i].size();
    if (m == 0 || nrec[i] == 0) continue;
    const int64_t first_row = offsets[los[i]], nrows = offsets[los[i] + m] - first_row;
    if (nrows == 0) continue;
    std::vector<long long> rec_off((size_t)m + 1), row_off((size_t)m + 1);
    rec_off[0] = 0;
    for (int64_t j = 0; j < m; j++) {
      rec_off[(size_t)j + 1] = rec_off[(size_t)j] + h_cnt[i][(size_t)j];
      row_off[(size_t)j] = offsets[los[i] + j] - first_row;
    }
    row_off[(size_t)m] = nrows;
    CU(cudaSetDevice(s.dev));
    CU(s.c_recoff.need((size_t)(m + 1) * 8));
    CU(s.c_rowoff.need((size_t)(m + 1) * 8));
    CU(s.c_cur.need((size_t)m * 4));
    CU(s.c_recs2.need((size_t)nrec[i] * sizeof(osr::ContactRec)));
    CU(s.rowrec.need((size_t)nrows * 4));
    CU(cudaMemcpyAsync(s.c_recoff.p, rec_off.data(), (size_t)(m + 1) * 8, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.c_rowoff.p, row_off.data(), (size_t)(m + 1) * 8, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemsetAsync(s.c_cur.p, 0, (size_t)m * 4, s.stream));
    osr::k_rec_scatter<<<(unsigned)((nrec[i] + 255) / 256), 256, 0, s.stream>>>(
        (const osr::ContactRec*)s.recs.p, (long long)nrec[i], (const long long*)s.c_recoff.p, (int*)s.c_cur.p,
        (osr::ContactRec*)s.c_recs2.p);
    CU(cudaGetLastError());
    const unsigned order_blocks = (unsigned)std::min<int64_t>((m + 3) / 4, (int64_t)s.num_sms * 16);
    osr::k_rec_order<<<order_blocks, 128, 0, s.stream>>>(
        (const osr::ContactRec*)s.c_recs2.p, (osr::ContactRec*)s.recs.p, (const long long*)s.c_recoff.p, (long long)m,
        (const long long*)s.c_rowoff.p, (int*)s.rowrec.p, w->rec_warp_max);
    CU(cudaGetLastError());
    std::vector<int> big;
    for (int64_t j = 0; j < m; j++)
      if (h_cnt[i][(size_t)j] > w->rec_warp_max) big.push_back((int)j);
    if (!big.empty()) {
      CU(s.c_cnt.need(big.size() * 4));   // (the per-state counts on the device are no longer needed)
      CU(cudaMemcpyAsync(s.c_cnt.p, big.data(), big.size() * 4, cudaMemcpyHostToDevice, s.stream));
      osr::k_rec_order_big<<<(unsigned)big.size(), 1024, 0, s.stream>>>(
          (const osr::ContactRec*)s.c_recs2.p, (osr::ContactRec*)s.recs.p, (const long long*)s.c_recoff.p,
          (const int*)s.c_cnt.p, (const long long*)s.c_rowoff.p, (int*)s.rowrec.p);
      CU(cudaGetLastError());
      w->host_counters.kernel_launches++;
    }
    CU(cudaStreamSynchronize(s.stream));   // rec_off / row_off / big (host vectors of this iteration) have been consumed
    CU(s.seg.need((size_t)nrows * 48));
    CU(s.faces.need((size_t)nrows * 8));
    CU(s.cop.need((size_t)nrows));
    if (bary) CU(s.bary.need((size_t)nrows * 48));
    osr::KParams P = w->params(s);
    osr::k_contact_rows<<<(unsigned)((nrows + 127) / 128), 128, 0, s.stream>>>(
        P, (const double*)s.in0.p, (const osr::ContactRec*)s.recs.p, (const int*)s.rowrec.p, (long long)nrows,
        (double*)s.seg.p, (int*)s.faces.p, (uint8_t*)s.cop.p, bary ? (double*)s.bary.p : nullptr);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches += 3;
    CU(cudaMemcpyAsync(seg + 6 * first_row, s.seg.p, (size_t)nrows * 48, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaMemcpyAsync(faces + 2 * first_row, s.faces.p, (size_t)nrows * 8, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaMemcpyAsync(coplanar + first_row, s.cop.p, (size_t)nrows, cudaMemcpyDeviceToHost, s.stream));
    if (bary) CU(cudaMemcpyAsync(bary + 6 * first_row, s.bary.p, (size_t)nrows * 48, cudaMemcpyDeviceToHost, s.stream));
  }
  for (auto& s : w->slots) {
    CU(cudaSetDevice(s.dev));
    CU(cudaStreamSynchronize(s.stream));
  }
  return OSR_OK;
}

// ---- surface mapping of an intersection mesh (unit_world.cc:928-957, 1092-1176) -------------------
int osr_intersecting_to_surface(osr_world* w, int which, const double* q, const double* V, int64_t nV,
                                const int32_t* F, int64_t nF, int64_t cap, int32_t* out_faces, double* out_bary,
                                int64_t* total) {
  int rc = check_ready(w);
  if (rc) return rc;
  if ((which != 0 && which != 1) || nV < 0 || nF < 0 || cap < 0 || !total || (which == 1 && !q) ||
      (nF > 0 && (!V || !F)) || (cap > 0 && (!out_faces || !out_bary)))
    return fail(OSR_EINVAL, "bad arguments");
  for (int64_t i = 0; i < 3 * nF; i++)
    if (F[i] < 0 || F[i] >= nV) return fail(OSR_EINVAL, "face index out of range");
  *total = 0;
  if (nF == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  const osr::UnitModel& tm = which ? w->rob : w->env;
  const double* d_tris = which ? s.rob_tris : s.env_tris;
  osr::Rigid tf;
  if (which) osr::state_to_rigid(q, tf);          // getRobotGeometry: translate_state_to_transform(q) * V
  else osr::state_to_rigid(w->perturb, tf);       // getSceneGeometry: the environment transform of getCDTransforms
  const long long nT = tm.nf(), pairs = nT * (long long)nF;
  const long long nblocks = (pairs + osr::SURF_BLOCK - 1) / osr::SURF_BLOCK;
  if (nblocks > 0x7fffffffLL) return fail(OSR_EINVAL, "too many triangle pairs for one call");
  CU(s.surf_v.need((size_t)nV * 24));
  CU(s.surf_f.need((size_t)nF * 12));
  CU(s.surf_flags.need((size_t)pairs));
  CU(s.surf_cnt.need((size_t)nblocks * 4));
  CU(s.surf_off.need((size_t)nblocks * 8));
  CU(cudaMemcpyAsync(s.surf_v.p, V, (size_t)nV * 24, cudaMemcpyHostToDevice, s.stream));
  CU(cudaMemcpyAsync(s.surf_f.p, F, (size_t)nF * 12, cudaMemcpyHostToDevice, s.stream));
  osr::k_surface_flags<<<(unsigned)nblocks, osr::SURF_BLOCK, 0, s.stream>>>(
      d_tris, tf, nT, (const double*)s.surf_v.p, (const int*)s.surf_f.p, (long long)nF, (uint8_t*)s.surf_flags.p,
      (int*)s.surf_cnt.p);
  CU(cudaGetLastError());
  w->host_counters.kernel_launches++;
  std::vector<int> cnt((size_t)nblocks);
  CU(cudaMemcpyAsync(cnt.data(), s.surf_cnt.p, (size_t)nblocks * 4, cudaMemcpyDeviceToHost, s.stream));
  CU(cudaStreamSynchronize(s.stream));
  std::vector<long long> off((size_t)nblocks);
  long long found = 0;
  for (long long b = 0; b < nblocks; b++) { off[(size_t)b] = found; found += cnt[(size_t)b]; }
  *total = found;
  long long rows = std::min<long long>(found, cap);
  if (rows == 0) return OSR_OK;
  CU(s.surf_face.need((size_t)rows * 4));
  CU(s.surf_l.need((size_t)rows * 72));
  CU(cudaMemcpyAsync(s.surf_off.p, off.data(), (size_t)nblocks * 8, cudaMemcpyHostToDevice, s.stream));
  osr::k_surface_emit<<<(unsigned)nblocks, osr::SURF_BLOCK, 0, s.stream>>>(
      d_tris, tf, nT, (const double*)s.surf_v.p, (const int*)s.surf_f.p, (long long)nF, (const uint8_t*)s.surf_flags.p,
      (const long long*)s.surf_off.p, rows, (int*)s.surf_face.p, (double*)s.surf_l.p);
  CU(cudaGetLastError());
  w->host_counters.kernel_launches++;
  std::vector<int> face((size_t)rows);
  CU(cudaMemcpyAsync(face.data(), s.surf_face.p, (size_t)rows * 4, cudaMemcpyDeviceToHost, s.stream));
  CU(cudaMemcpyAsync(out_bary, s.surf_l.p, (size_t)rows * 72, cudaMemcpyDeviceToHost, s.stream));
  CU(cudaStreamSynchronize(s.stream));
  for (long long r = 0; r < rows; r++)  // retF.row = targetF.row(target_fi)
    for (int k = 0; k < 3; k++) out_faces[3 * r + k] = tm.F[(size_t)3 * face[(size_t)r] + k];
  return OSR_OK;
}

// ---- windowed k-NN under the state metric (config 3 edge generation) ------------------------------
int osr_knn_windowed(osr_world* w, const double* q, int64_t n, int k, int64_t window, int32_t* out_idx,
                     double* out_dist) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!q || !out_idx)) || k <= 0 || k > osr::KNN_KMAX || window <= 0 ||
      window % osr::KNN_TILE != 0 || n > 0x7fffffffLL)
    return fail(OSR_EINVAL, "bad arguments (1 <= k <= 16, window a positive multiple of 256)");
  if (n == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  CU(s.in0.need((size_t)n * 56));
  CU(s.faces.need((size_t)n * k * 4));
  if (out_dist) CU(s.seg.need((size_t)n * k * 8));
  CU(cudaMemcpyAsync(s.in0.p, q, (size_t)n * 56, cudaMemcpyHostToDevice, s.stream));
  CU(cudaEventRecord(s.ev0, s.stream));
  osr::k_knn_window<<<(unsigned)((n + osr::KNN_TILE - 1) / osr::KNN_TILE), osr::KNN_TILE, 0, s.stream>>>(
      (const double*)s.in0.p, (long long)n, k, (long long)window, (int*)s.faces.p, out_dist ? (double*)s.seg.p : nullptr);
  CU(cudaGetLastError());
  CU(cudaEventRecord(s.ev1, s.stream));
  w->host_counters.kernel_launches++;
  CU(cudaMemcpyAsync(out_idx, s.faces.p, (size_t)n * k * 4, cudaMemcpyDeviceToHost, s.stream));
  if (out_dist) CU(cudaMemcpyAsync(out_dist, s.seg.p, (size_t)n * k * 8, cudaMemcpyDeviceToHost, s.stream));
  CU(cudaStreamSynchronize(s.stream));
  return OSR_OK;
}

// ---- k-NN edges validated where they are --------------------------------------
// The planners' pattern (GNAT k-NN, then checkMotion on every edge found: lib/pyse3ompl/ompldriver.cc:176-260) and
// config 3 of BASELINE.json in one call: K6 leaves the neighbour indices on the device and K2 walks edge
// (i, idx[i][s]) straight from the milestone array.  56 bytes per milestone go up, k (4 + 1) bytes come back; the
// (from, to) arrays of the edges -- 112 bytes per edge, 3.7 GB for 4M milestones x 8 -- are never built or moved.
int osr_knn_validate_edges(osr_world* w, const double* q, int64_t n, int k, int64_t window, double initial_verify_delta,
                           int32_t* out_idx, uint8_t* out_valid) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!q || !out_valid)) || k <= 0 || k > osr::KNN_KMAX || window <= 0 ||
      window % osr::KNN_TILE != 0 || n > 0x7fffffffLL / osr::KNN_KMAX)
    return fail(OSR_EINVAL, "bad arguments (1 <= k <= 16, window a positive multiple of 256, n k < 2^31)");
  if (initial_verify_delta < 0.0) return fail(OSR_EINVAL, "verify_delta must not be negative");
  if (n == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  // k-NN is local to a window, so whole windows are sharded over the devices of the handle: each device gets the
  // milestones of its windows only and works on local indices; the window's first index is added on the way out.
  const int parts = (int)w->slots.size();
  const int64_t nwin = (n + window - 1) / window;
  std::vector<int64_t> los(parts, 0), his(parts, 0);
  for (int i = 0; i < parts; i++) {
    int64_t wlo, whi;
    shard(nwin, parts, i, wlo, whi);
    los[i] = std::min(n, wlo * window);
    his[i] = std::min(n, whi * window);
    const int64_t nl = his[i] - los[i];
    if (nl == 0) continue;
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    const int64_t m = nl * k;
    CU(s.in0.need((size_t)nl * 56));
    CU(s.faces.need((size_t)m * 4));
    CU(s.out[2].need((size_t)m));
    CU(cudaMemcpyAsync(s.in0.p, q + 7 * los[i], (size_t)nl * 56, cudaMemcpyHostToDevice, s.stream));
    osr::k_knn_window<<<(unsigned)((nl + osr::KNN_TILE - 1) / osr::KNN_TILE), osr::KNN_TILE, 0, s.stream>>>(
        (const double*)s.in0.p, (long long)nl, k, (long long)window, (int*)s.faces.p, nullptr);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    if (out_idx) CU(cudaMemcpyAsync(out_idx + los[i] * k, s.faces.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s.stream));
    osr::SegOut O = {nullptr, nullptr, (uint8_t*)s.out[2].p, nullptr, nullptr, nullptr, (const int*)s.faces.p, k};
    rc = launch_transit(w, s, (const double*)s.in0.p, (const double*)s.in0.p, m, nullptr, 1, initial_verify_delta, 1, O);
    if (rc) return rc;
    CU(cudaMemcpyAsync(out_valid + los[i] * k, s.out[2].p, (size_t)m, cudaMemcpyDeviceToHost, s.stream));
  }
  for (int i = 0; i < parts; i++) {
    if (his[i] == los[i]) continue;
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    rc = check_device_error(w, s);
    if (rc) return rc;
    if (out_idx && los[i] > 0) {
      const int32_t base = (int32_t)los[i];
      int32_t* p = out_idx + los[i] * k;
      const int64_t cnt = (his[i] - los[i]) * k;
      for (int64_t j = 0; j < cnt; j++)
        if (p[j] >= 0) p[j] += base;
    }
  }
  return OSR_OK;
}

// ---- clearance statistics of key-configuration candidates ----------------------
// calc_touch + multi_distance + [median, max, min, mean, std] per candidate (preprocess_key.py:174-184,275-281):
// candidates are sharded over the devices, every device walks its candidates' touch segments (K2) and reduces
// the last free states in place (K8); 40 bytes per candidate come back instead of npoint x 15 doubles.
int osr_clearance_stats(osr_world* w, const double* q0, int64_t m, const double* to, int npoint, double verify_delta,
                        double* stats, double* distances) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (m < 0 || npoint <= 0 || npoint > osr::CLR_MAX_POINTS || (m > 0 && (!q0 || !to || !stats)) || !(verify_delta > 0.0))
    return fail(OSR_EINVAL, "bad arguments (1 <= npoint <= 8192, verify_delta > 0)");
  if (m == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  int padded = 1;
  while (padded < npoint) padded <<= 1;
  const size_t smem = (size_t)padded * 8;
  int parts = (int)w->slots.size();
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(m, parts, i, lo, hi);
    const int64_t mc = hi - lo, n = mc * npoint;
    if (mc == 0) continue;
    CU(cudaSetDevice(s.dev));
    CU(s.clr_q0.need((size_t)mc * 56));
    CU(s.in0.need((size_t)n * 56));
    CU(s.in1.need((size_t)n * 56));
    CU(s.out[0].need((size_t)n * 56));
    CU(s.clr_stats.need((size_t)mc * 40));
    if (distances) CU(s.clr_dist.need((size_t)n * 8));
    CU(cudaMemcpyAsync(s.clr_q0.p, q0 + 7 * lo, (size_t)mc * 56, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.in1.p, to + 7 * lo * npoint, (size_t)n * 56, cudaMemcpyHostToDevice, s.stream));
    const long long total = (long long)n * 7;
    osr::k_repeat_states<<<(unsigned)((total + 255) / 256), 256, 0, s.stream>>>((const double*)s.clr_q0.p, (long long)mc,
                                                                               npoint, (double*)s.in0.p);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    osr::SegOut O = {(double*)s.out[0].p, nullptr, nullptr, nullptr, nullptr, nullptr};
    rc = launch_transit(w, s, (const double*)s.in0.p, (const double*)s.in1.p, n, nullptr, 1, verify_delta, 0, O);
    if (rc) return rc;
    if (smem > 48 * 1024)
      CU(cudaFuncSetAttribute(osr::k_clearance_stats, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    osr::k_clearance_stats<<<(unsigned)mc, osr::CLR_THREADS, smem, s.stream>>>(
        (const double*)s.clr_q0.p, (const double*)s.out[0].p, npoint, padded, (double*)s.clr_stats.p,
        distances ? (double*)s.clr_dist.p : nullptr);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    CU(cudaMemcpyAsync(stats + 5 * lo, s.clr_stats.p, (size_t)mc * 40, cudaMemcpyDeviceToHost, s.stream));
    if (distances)
      CU(cudaMemcpyAsync(distances + lo * npoint, s.clr_dist.p, (size_t)n * 8, cudaMemcpyDeviceToHost, s.stream));
  }
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    rc = check_device_error(w, s);
    if (rc) return rc;
  }
  return OSR_OK;
}

// ---- barycentric UV-chart splat (Renderer::renderBarycentric, osr_render.cc:590-759) ---------------
int osr_rasterize_barycentric(osr_world* w, const float* uv, const float* bary, const float* weight, int64_t nprim,
                              int res_x, int res_y, float* out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (nprim < 0 || res_x <= 0 || res_y <= 0 || !out || (nprim > 0 && (!uv || !bary || !weight)))
    return fail(OSR_EINVAL, "bad arguments");
  std::lock_guard<std::mutex> lk(w->mu);
  Slot& s = w->slots[0];
  CU(cudaSetDevice(s.dev));
  const size_t img = (size_t)res_x * res_y * 4;
  CU(s.surf_l.need(img));
  CU(cudaMemsetAsync(s.surf_l.p, 0, img, s.stream));   // bary_tex_->clear()
  if (nprim > 0) {
    CU(s.surf_v.need((size_t)nprim * 24));
    CU(s.surf_f.need((size_t)nprim * 36));
    CU(s.surf_off.need((size_t)nprim * 12));
    CU(cudaMemcpyAsync(s.surf_v.p, uv, (size_t)nprim * 24, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.surf_f.p, bary, (size_t)nprim * 36, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.surf_off.p, weight, (size_t)nprim * 12, cudaMemcpyHostToDevice, s.stream));
    const long long blocks = (nprim * 32 + 255) / 256;
    if (blocks > 0x7fffffffLL) return fail(OSR_EINVAL, "too many primitives for one call");
    CU(cudaEventRecord(s.ev0, s.stream));
    osr::k_bary_raster<<<(unsigned)blocks, 256, 0, s.stream>>>((const float*)s.surf_v.p, (const float*)s.surf_f.p,
                                                              (const float*)s.surf_off.p, (long long)nprim, res_x,
                                                              res_y, (float*)s.surf_l.p);
    CU(cudaGetLastError());
    CU(cudaEventRecord(s.ev1, s.stream));
    w->host_counters.kernel_launches++;
  }
  CU(cudaMemcpyAsync(out, s.surf_l.p, img, cudaMemcpyDeviceToHost, s.stream));
  CU(cudaStreamSynchronize(s.stream));
  return OSR_OK;
}

// ---- OMPL-frame batched validators (OmplDriver::validateStates / validateMotionPairs, ompldriver.cc:566-621) ----
static int ompl_bounds_ok(const double* bounds) {
  if (!bounds) return 0;
  for (int k = 0; k < 3; k++)
    if (!(bounds[3 + k] >= bounds[k])) return 0;
  return 1;
}
int osr_ompl_validate_states(osr_world* w, const double* q_ompl, int64_t n, const double* bounds, uint8_t* out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!q_ompl || !out)) || (bounds && !ompl_bounds_ok(bounds)))
    return fail(OSR_EINVAL, "bad arguments");
  if (n == 0) return OSR_OK;
  std::vector<double> u((size_t)n * 7);
  rc = osr_translate_ompl_to_unit(w, q_ompl, n, u.data());
  if (rc) return rc;
  rc = osr_is_valid_states(w, u.data(), n, out);
  if (rc || !bounds) return rc;
  // FCLStateValidityChecker::isValid = si_->satisfiesBounds(state) && collision free
  for (int64_t i = 0; i < n; i++)
    if (out[i] && !osr::ompl_satisfies_bounds(bounds, bounds + 3, q_ompl + 7 * i)) out[i] = 0;
  return OSR_OK;
}

int osr_ompl_check_motions(osr_world* w, const double* q0, const double* q1, int64_t n, const double* bounds,
                           double resolution, uint8_t* out, int64_t* nd_out) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!q0 || !q1 || !out)) || !ompl_bounds_ok(bounds) || !(resolution > 0.0))
    return fail(OSR_EINVAL, "bad arguments");
  // RealVectorStateSpace::getMaximumExtent (spaces/src/RealVectorStateSpace.cpp:167-176)
  double trans_extent = 0.0;
  for (int k = 0; k < 3; k++) {
    const double d = bounds[3 + k] - bounds[k];
    trans_extent += d * d;
  }
  trans_extent = std::sqrt(trans_extent);
  if (!(trans_extent > 0.0)) return fail(OSR_EINVAL, "translation bounds have no extent");
  const double ident[7] = {0, 0, 0, 1, 0, 0, 0};
  if (memcmp(w->perturb, ident, sizeof(ident)) != 0)
    return fail(OSR_EINVAL, "OMPL-frame motion checks need the identity perturbation");
  if (n == 0) return OSR_OK;
  std::lock_guard<std::mutex> lk(w->mu);
  osr::OmplFrame F;
  F.scale = (double)w->scene_scale;
  for (int k = 0; k < 3; k++) F.dc[k] = (double)(w->ompl_centre[k] - w->centre[k]) * (double)w->scene_scale;
  for (int k = 0; k < 3; k++) { F.lo[k] = bounds[k]; F.hi[k] = bounds[3 + k]; }
  F.seg_trans = resolution * trans_extent;
  F.seg_rot = resolution * (0.5 * 3.14159265358979323846);
  int parts = (int)w->slots.size();
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(n, parts, i, lo, hi);
    int64_t m = hi - lo;
    if (m == 0) continue;
    CU(cudaSetDevice(s.dev));
    CU(s.in0.need((size_t)m * 56));
    CU(s.in1.need((size_t)m * 56));
    CU(s.out[0].need((size_t)m));
    CU(s.out[5].need((size_t)m * 8));
    CU(cudaMemcpyAsync(s.in0.p, q0 + 7 * lo, (size_t)m * 56, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(s.in1.p, q1 + 7 * lo, (size_t)m * 56, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemsetAsync(s.work, 0, 2 * sizeof(unsigned long long), s.stream));
    osr::KParams P = w->params(s);
    CU(cudaEventRecord(s.ev0, s.stream));
    osr::k_ompl_motion<<<s.grid_transit, osr::WARPS_PER_BLOCK * 32, SMEM_PER_BLOCK, s.stream>>>(
        P, F, (const double*)s.in0.p, (const double*)s.in1.p, (long long)m, (uint8_t*)s.out[0].p,
        nd_out ? (long long*)s.out[5].p : nullptr);
    CU(cudaGetLastError());
    CU(cudaEventRecord(s.ev1, s.stream));
    w->host_counters.kernel_launches++;
    CU(cudaMemcpyAsync(out + lo, s.out[0].p, (size_t)m, cudaMemcpyDeviceToHost, s.stream));
    if (nd_out) CU(cudaMemcpyAsync(nd_out + lo, s.out[5].p, (size_t)m * 8, cudaMemcpyDeviceToHost, s.stream));
  }
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    rc = check_device_error(w, s);
    if (rc) return rc;
  }
  return OSR_OK;
}

// ---- key-configuration enumeration (kernels_keyconf.cuh) -------------------------
static void keyconf_half_angles(int den, std::vector<double>& cs) {   // unit_world.cc:1358-1362: i * (2 pi / den)
  const double PI2 = 2.0 * 3.14159265358979323846;
  const double delta = PI2 / (double)den;
  cs.resize((size_t)2 * den);
  for (int i = 0; i < den; i++) {
    double ang = (double)i * delta;
    cs[2 * i] = std::cos(0.5 * ang);
    cs[2 * i + 1] = std::sin(0.5 * ang);
  }
}

// host construction of the candidates (no device work): the states osr_enum_free_configurations labels
int osr_enum_configuration_states(const double* pts, int64_t b, double margin, int den, double* out_q) {
  if (b < 0 || den <= 0 || (b > 0 && (!pts || !out_q))) return fail(OSR_EINVAL, "bad arguments");
  std::vector<double> cs;
  keyconf_half_angles(den, cs);
  for (int64_t r = 0; r < b; r++)
    for (int i = 0; i < den; i++) osr::keyconf_state(pts + 12 * r, margin, cs[2 * i], cs[2 * i + 1], out_q + 7 * (r * den + i));
  return OSR_OK;
}

int osr_enum_free_configurations(osr_world* w, const double* pts, int64_t b, double margin, int den, int only_median,
                                 double* out_q, uint8_t* out_valid, int32_t* counts) {
  int rc = check_ready(w);
  if (rc) return rc;
  if (b <= 0 || den <= 0 || !pts || !out_q || !out_valid) return fail(OSR_EINVAL, "bad arguments");
  std::lock_guard<std::mutex> lk(w->mu);
  std::vector<double> cs;
  keyconf_half_angles(den, cs);
  int parts = (int)w->slots.size();
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    int64_t lo, hi;
    shard(b, parts, i, lo, hi);
    const int64_t mb = hi - lo, n = mb * den;
    if (mb == 0) continue;
    CU(cudaSetDevice(s.dev));
    CU(s.in1.need((size_t)mb * 96 + (size_t)den * 16));
    CU(s.in0.need((size_t)n * 56));
    CU(s.out[2].need((size_t)n));
    CU(s.out[5].need((size_t)mb * 4));
    double* d_pts = (double*)s.in1.p;
    double* d_cs = d_pts + 12 * mb;
    CU(cudaMemcpyAsync(d_pts, pts + 12 * lo, (size_t)mb * 96, cudaMemcpyHostToDevice, s.stream));
    CU(cudaMemcpyAsync(d_cs, cs.data(), (size_t)den * 16, cudaMemcpyHostToDevice, s.stream));
    osr::k_keyconf_states<<<(unsigned)((n + 255) / 256), 256, 0, s.stream>>>(d_pts, (long long)mb, den, margin, d_cs,
                                                                           (double*)s.in0.p);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    rc = launch_valid(w, s, (const double*)s.in0.p, n, (uint8_t*)s.out[2].p);
    if (rc) return rc;
    osr::k_keyconf_filter<<<(unsigned)((mb + 127) / 128), 128, 0, s.stream>>>((uint8_t*)s.out[2].p, (long long)mb, den,
                                                                            only_median, (int*)s.out[5].p);
    CU(cudaGetLastError());
    w->host_counters.kernel_launches++;
    CU(cudaMemcpyAsync(out_q + 7 * lo * den, s.in0.p, (size_t)n * 56, cudaMemcpyDeviceToHost, s.stream));
    CU(cudaMemcpyAsync(out_valid + lo * den, s.out[2].p, (size_t)n, cudaMemcpyDeviceToHost, s.stream));
    if (counts) CU(cudaMemcpyAsync(counts + lo, s.out[5].p, (size_t)mb * 4, cudaMemcpyDeviceToHost, s.stream));
  }
  for (int i = 0; i < parts; i++) {
    Slot& s = w->slots[i];
    CU(cudaSetDevice(s.dev));
    rc = check_device_error(w, s);
    if (rc) return rc;
  }
  return OSR_OK;
}

int osr_get_bvh_nodes(const osr_world* w, int which, float* out) {
  if (!w || !w->host_ready) return fail(OSR_EINVAL, "geometry not built: call osr_finalize()");
  const osr::UnitModel& m = which ? w->rob : w->env;
  memcpy(out, m.bvh.nodes.data(), m.bvh.nodes.size() * 4);
  return OSR_OK;
}

}  // extern "C"

/* osrgpu.h -- C ABI of libosrgpu.so, the B200 (sm_100a) batched rigid-body collision engine
 * behind the pyosr `UnitWorld` surface.
 *
 * The reference has no C ABI for this path: its boundary is the pybind11 class
 * `pyosr.UnitWorld` (/root/reference/lib/pyosr/module.cc:127-250) over C++ member functions of
 * osr::UnitWorld (/root/reference/lib/osr/unit_world.h).  Each entry point below names the
 * member function(s) it replaces; INTEGRATION.md shows the pybind11 stub a maintainer would
 * put in lib/pyosr/module.cc to call it.
 *
 * Conventions
 *   - every function returns 0 on success or a negative OSR_E* code; osr_last_error() gives
 *     the message of the calling thread's last failure (the reference throws
 *     std::runtime_error, module.cc relies on pybind11's translation).
 *   - states are `double[N][7]`, row-major, [tx ty tz qw qx qy qz] (lib/osr/osr_state.h:20-33),
 *     in UNIT-world coordinates unless stated.
 *   - all arrays are caller-owned HOST memory (pageable or pinned); results are complete when
 *     the call returns.  The *_device variants take device pointers on device 0 of the handle
 *     and are ordered on the handle's stream; they exist for callers that keep samples in HBM.
 *   - there is NO CPU fallback: every query runs on the GPU(s) selected at osr_finalize().
 */
#ifndef OSRGPU_H
#define OSRGPU_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct osr_world osr_world;

enum {
  OSR_OK = 0,
  OSR_EINVAL = -1,   /* bad argument / call order            */
  OSR_EIO = -2,      /* mesh file unreadable                  */
  OSR_ECUDA = -3,    /* CUDA runtime failure (message kept)   */
  OSR_ENODEV = -4,   /* no usable CUDA device                 */
  OSR_EOVERFLOW = -5 /* device work stack exhausted           */
};

const char* osr_last_error(void);

/* UnitWorld::UnitWorld / ~UnitWorld (lib/osr/unit_world.cc:33-63) */
int osr_create(osr_world** out);
void osr_destroy(osr_world* w);

/* UnitWorld::loadModelFromFile / loadRobotFromFile (unit_world.cc:94-110) */
int osr_load_env_obj(osr_world* w, const char* path);
int osr_load_rob_obj(osr_world* w, const char* path);
/* same, from memory (float32 vertices, int32 triangles) -- for synthetic meshes */
int osr_set_env_mesh(osr_world* w, const float* V, int nV, const int32_t* F, int nF);
int osr_set_rob_mesh(osr_world* w, const float* V, int nV, const int32_t* F, int nF);
/* UnitWorld::enforceRobotCenter (unit_world.cc:113-117) */
int osr_enforce_robot_center(osr_world* w, const double c[3]);
/* UnitWorld::scaleToUnit + angleModel(0,0) (unit_world.cc:123-161): computes scene_scale, the
 * unit-space meshes and the BVHs, uploads them to `ndev` devices listed in `devices`
 * (ndev<=0: current device only).  Without a CUDA device it returns OSR_ENODEV after building the
 * host-side geometry, so the getters below still work; every query then fails (no CPU fallback). */
int osr_finalize(osr_world* w, const int* devices, int ndev);
/* UnitWorld::setPerturbation (unit_world.cc:165-176) */
int osr_set_perturbation(osr_world* w, const double q[7]);
/* properties scene_scale / recommended_cres (unit_world.h:322-346) and the robot centre used */
double osr_scene_scale(const osr_world* w);
double osr_get_recommended_cres(const osr_world* w);
int osr_set_recommended_cres(osr_world* w, double v);
int osr_get_robot_center(const osr_world* w, double c[3]);
/* UnitWorld::getOMPLCenter(geo) (unit_world.cc:1562-1565): mean of the mesh's vertices, which=0 env, 1 robot */
int osr_get_ompl_center(const osr_world* w, int which, double c[3]);
int osr_num_devices(const osr_world* w);
/* geometry read-back (CDModel::vertices()/faces(), cdmodel.cc:232-246): which 0=env 1=robot */
int osr_mesh_sizes(const osr_world* w, int which, int* nV, int* nF, int* nNodes, int* depth);
int osr_get_unit_mesh(const osr_world* w, int which, double* V, int32_t* F);
/* the engine's own OBB tree (16 floats per node, layout in csrc/geometry.h); for conservativeness tests */
int osr_get_bvh_nodes(const osr_world* w, int which, float* out);

/* Work counters accumulated over the calls made since the last reset (FCL's num_bv_tests /
 * num_leaf_tests, mesh_collision_traversal_node-inl.h:497,547, plus the engine's own). */
typedef struct osr_counters {
  uint64_t bv_tests;      /* OBB-pair tests executed                         */
  uint64_t leaf_tests;    /* triangle-pair tests executed (fp32 filter)      */
  uint64_t fp64_rechecks; /* triangle pairs re-tested with the exact SAT-17  */
  uint64_t poses;         /* validity checks (configs or segment steps)      */
  uint64_t spills;        /* stack spill events to global memory             */
  uint64_t kernel_launches;
  /* lane occupancy of the depth-first traversal kernels: per warp iteration 32 lanes can each run one box
   * test; lanes_active did, lanes_waiting held a pose whose queued triangle tests were pending,
   * lanes_idle had no pose (input exhausted).  warp efficiency = lanes_active / (32 * warp_iters). */
  uint64_t warp_iters, lanes_active, lanes_waiting, lanes_idle;
} osr_counters;
int osr_get_counters(osr_world* w, osr_counters* out, int reset);

/* UnitWorld::isValid per row (unit_world.cc:215-247): out[i]=1 iff state i is collision free */
int osr_is_valid_states(osr_world* w, const double* q, int64_t n, uint8_t* out);
/* Same labels for float32 states: the reference's callers hold float32 samples (src/GP/pipeline/uw_random.py:15-17)
 * that pybind11 widens to float64 on the host (module.cc:138); here the widening -- exact -- happens on the device
 * and the upload is half the size. */
int osr_is_valid_states_f32(osr_world* w, const float* q, int64_t n, uint8_t* out);
/* UnitWorld::isDisentangled per row (unit_world.cc:249-264) */
int osr_is_disentangled_states(osr_world* w, const double* q, int64_t n, uint8_t* out);

/* UnitWorld::transitStateToWithContact per row (unit_world.cc:405-474).  `vd` holds one value
 * (vd_count==1) or n values.  Outputs (any may be NULL): last_free[n][7], first_hit[n][7],
 * is_free[n], free_tau[n], touch_tau[n] (taus as double; the reference narrows them to float
 * in its return tuple), steps[n] = number of validity checks the reference would have made. */
int osr_transit_states_to_with_contact(osr_world* w, const double* from, const double* to, int64_t n,
                                       const double* vd, int64_t vd_count, double* last_free,
                                       double* first_hit, uint8_t* is_free, double* free_tau,
                                       double* touch_tau, int64_t* steps);
/* UnitWorld::isValidTransition per row (unit_world.cc:477-486): vdelta=min(d/2, initial_vd) */
int osr_is_valid_transitions(osr_world* w, const double* from, const double* to, int64_t n,
                             double initial_verify_delta, uint8_t* out);
/* UnitWorld::calculateVisibilityMatrix2 (unit_world.cc:764-797): out[m][n] int8, row-major.
 * is_unit0/1 == 0 -> rows are first mapped with translateToUnitState. */
int osr_visibility_matrix2(osr_world* w, const double* q0, int64_t m, int is_unit0, const double* q1,
                           int64_t n, int is_unit1, double verify_magnitude, int8_t* out);
/* UnitWorld::calculateVisibilityPair (unit_world.cc:799-840): out[max(m,n)] */
int osr_visibility_pair(osr_world* w, const double* q0, int64_t m, int is_unit0, const double* q1,
                        int64_t n, int is_unit1, double verify_magnitude, int8_t* out);

/* UnitWorld::intersectingSegments per row (unit_world.cc:959-1012; CDModel::collideForDetails,
 * cdmodel.cc:200-229; tritri::TriTri, lib/tritri/tritri_igl.hh:15-67).  CSR output:
 * rows offsets[i]..offsets[i+1] belong to state i; faces[r] = (env triangle, robot triangle);
 * seg[r] = 6 doubles (two end points, zeros when Moeller reports none); coplanar[r].
 * Optional bary[r] = 6 doubles: barycentric coordinates of the segment mid point on the env
 * triangle then on the robot triangle.  At most `cap` rows are written; *total gets the number
 * found; overflow[i]=1 for states whose rows were truncated. */
int osr_intersecting_segments(osr_world* w, const double* q, int64_t n, int64_t cap, int64_t* offsets,
                              double* seg, int32_t* faces, uint8_t* coplanar, double* bary,
                              uint8_t* overflow, int64_t* total);

/* UnitWorld::enumFreeConfiguration per row (unit_world.cc:1340-1389): pts[b] = {rob_o, rob_n,
 * env_o, env_n} (4x3 doubles).  out_q[b][den][7], out_valid[b][den] (after the optional
 * median-of-run filter), counts[b] = number of valid entries.  The candidates are built, labelled and
 * filtered on the device (pairs sharded over the devices of the handle); 96 bytes per pair go up. */
int osr_enum_free_configurations(osr_world* w, const double* pts, int64_t b, double margin, int den,
                                 int only_median, double* out_q, uint8_t* out_valid, int32_t* counts);

/* State-frame maps (unit_world.cc:526-555, 557-591, 689-738), n rows each */
int osr_translate_to_unit(const osr_world* w, const double* q, int64_t n, double* out);
int osr_translate_from_unit(const osr_world* w, const double* q, int64_t n, double* out);
int osr_translate_ompl_to_unit(const osr_world* w, const double* q_wlast, int64_t n, double* out);
int osr_translate_unit_to_ompl(const osr_world* w, const double* q, int64_t n, double* out_wlast);

/* Page-locked host memory for the arrays handed to the batched calls (optional: every call accepts ordinary
 * pageable memory, but then the driver's staging copies, not PCIe, bound its throughput).  NULL on failure. */
void* osr_host_alloc(size_t bytes);
void osr_host_free(void* p);

/* module-level state functions (lib/pyosr/module.cc:77-99; lib/osr/osr_state.cc) */
double osr_distance(const double a[7], const double b[7]);
/* The step deltas of the segment walk (unit_world.cc:445-470: delta <- delta < dist ? min(dist, delta + vd)
 * : delta + vd, a running fp64 sum) advanced by up to `nsteps` steps from *delta, stopping at the first value beyond
 * dist; every rounding of the one-add-per-step loop is reproduced (host build of the routine the segment kernel
 * uses to advance over swept windows).  Returns the number of steps whose delta is <= dist; *delta = value after the
 * last executed step, *last_valid = delta of the last valid step (the input delta when there is none). */
int osr_walk_deltas(double* delta, double verify_delta, double dist, int nsteps, double* last_valid);

/* Host build of the kernels' triangle-pair predicates (csrc/sat17.h), n pairs of 18 doubles P1 P2 P3 Q1 Q2 Q3 with Q
 * already in P's frame (what Intersect::intersect_Triangle sees after Q' = R Q + T, intersect-inl.h:598-617):
 * filter[i] = fp32 SAT-17 filter (0 separated, 1 intersecting, 2 undecided), exact[i] = fp64 SAT-17 in the reference's
 * operation order (the label).  Either output may be NULL.  Needs no device. */
int osr_triangle_pair_predicates(const double* tris, int64_t n, uint8_t* filter, uint8_t* exact);

void osr_interpolate(const double a[7], const double b[7], double tau, double out[7]);
void osr_apply(const double from[7], const double tr[3], const double aa[3], double out[7]);

/* UnitWorld::intersectingToModelSurface (which=0) / intersectingToRobotSurface (which=1)
 * (unit_world.cc:928-957 -> intersectingToSurface :1092-1176): every (surface face i, mesh face j) pair that
 * Moeller97 reports as intersecting AND coplanar, in the reference's order (j outer, i inner).  Row r gets
 * out_faces[r][3] = vertex ids of surface face i and out_bary[3r+v][3] = barycentric coordinates (libigl
 * formula) of vertex v of mesh face j w.r.t. the world-space surface triangle.  `q` is the UNIT state of the
 * robot (ignored for which=0, where the environment transform is the perturbation).  *total = rows found;
 * at most `cap` rows are written. */
int osr_intersecting_to_surface(osr_world* w, int which, const double* q, const double* V, int64_t nV,
                                const int32_t* F, int64_t nF, int64_t cap, int32_t* out_faces,
                                double* out_bary, int64_t* total);

/* The candidates osr_enum_free_configurations labels, built on the host (no device, no world): out_q[b][den][7].
 * Same arithmetic as the device construction, bit for bit (csrc/se3.h keyconf_state). */
int osr_enum_configuration_states(const double* pts, int64_t b, double margin, int den, double* out_q);

/* Clearance statistics of m key-configuration candidates (replaces the host loop touchq_util.calc_touch +
 * pyosr.multi_distance + numpy statistics: src/GP/pipeline/touchq_util.py:8-27, preprocess_key.py:174-184 and
 * :275-281, keyconf.py:312-315; lib/osr/osr_state.cc:157-167).  Candidate c (unit state q0[c]) owns the touch
 * segments q0[c] -> to[c * npoint + i], i < npoint, walked as by osr_transit_states_to_with_contact with
 * `verify_delta`; d_i = osr::distance(q0[c], last free state of segment i).
 * stats[c][5] = {median, max, min, mean, std} of d (numpy conventions: median of an even count is the mean of
 * the two middle values, std is the population deviation); distances[m * npoint] optional (may be NULL).
 * 1 <= npoint <= 8192.  Candidates are sharded over the devices of the handle. */
int osr_clearance_stats(osr_world* w, const double* q0, int64_t m, const double* to, int npoint,
                        double verify_delta, double* stats, double* distances);

/* k nearest states under osr::distance (lib/osr/osr_state.cc:130-154) inside windows of `window` consecutive
 * indices (a positive multiple of 256; the last window may be short), the query itself excluded: the k-NN edge
 * generator of the RDT-forest edge-validation workload.  out_idx[n][k] = global indices ordered by
 * (distance, index), -1 where a window holds fewer than k other states; out_dist[n][k] optional.  1 <= k <= 16. */
int osr_knn_windowed(osr_world* w, const double* q, int64_t n, int k, int64_t window, int32_t* out_idx,
                     double* out_dist);

/* osr_knn_windowed + osr_is_valid_transitions on the edges it finds, without building them: out_valid[n][k] = 1 iff
 * is_valid_transition(q[i], q[out_idx[i][s]], initial_verify_delta) (unit_world.cc:477-486), 0 where out_idx is -1.
 * The planners' pattern -- nearest neighbours, then checkMotion on every edge (lib/pyse3ompl/ompldriver.cc:176-260,
 * RDT-forest edge validation of BASELINE config 3).  Only the milestones are uploaded.  out_idx optional.  Labels equal
 * those of the two separate calls.  Whole windows are sharded over the devices of the handle; n k < 2^31. */
int osr_knn_validate_edges(osr_world* w, const double* q, int64_t n, int k, int64_t window, double initial_verify_delta,
                           int32_t* out_idx, uint8_t* out_valid);

/* Renderer::renderBarycentric (lib/osr/osr_render.cc:590-759, shaders lib/osr/shader/bary.{vert,geom,frag}) as a
 * headless rasteriser: primitive p has three vertices with uv[3p+v][2] (UV corners of its surface face),
 * bary[3p+v][3] and weight[3p+v]; vertex v is drawn at sum_k uv[3p+k]*bary[3p+v][k] in a res_x x res_y chart,
 * fragments whose interpolated barycentrics leave [0,1] are discarded, the rest ADD their weight.  out[res_y][res_x]
 * (row y bottom-up as glReadPixels returns it) is cleared by the call. */
int osr_rasterize_barycentric(osr_world* w, const float* uv, const float* bary, const float* weight, int64_t nprim,
                              int res_x, int res_y, float* out);

/* OMPL-frame batched validators (rows x y z qx qy qz qw, StateSpace::copyFromEigen3):
 * OmplDriver::validateStates (lib/pyse3ompl/ompldriver.cc:566-589) = isValid after translateOMPLStateToUnitState;
 * OmplDriver::validateMotionPairs (:592-621) = DiscreteMotionValidator::checkMotion
 * (third-party/ompl.app/ompl/src/ompl/base/src/DiscreteMotionValidator.cpp:96-150): out[i]=1 iff s2 and every
 * interior point m/nd of OMPL's SE(3) interpolation are valid, nd = max(ceil(|dt| / (resolution*extent)),
 * ceil(acos|q1.q2| / (resolution*pi/2))).  "Valid" is FCLStateValidityChecker::isValid
 * (third-party/ompl.app/src/omplapp/geometry/detail/FCLStateValidityChecker.h:103): satisfiesBounds AND collision
 * free.  bounds = {lo x y z, hi x y z} of the planner's RealVectorBounds (OmplDriver::setBounds, OMPL frame);
 * extent = its diagonal (RealVectorStateSpace::getMaximumExtent); the SO(3) part of satisfiesBounds is
 * |norm(q) - 1| < 1e-9.  osr_ompl_validate_states accepts bounds = NULL (collision only); check_motions needs them.
 * resolution = the state-validity-checking resolution (OmplDriver::setCDRes).  nd_out optional. */
int osr_ompl_validate_states(osr_world* w, const double* q_ompl, int64_t n, const double* bounds, uint8_t* out);
int osr_ompl_check_motions(osr_world* w, const double* q0, const double* q1, int64_t n, const double* bounds,
                           double resolution, uint8_t* out, int64_t* nd_out);

/* Device-resident variants (pointers on the handle's first device).  Launch only; the caller
 * synchronises with osr_sync() (or its own stream wait on osr_stream()). */
int osr_is_valid_states_device(osr_world* w, const double* d_q, int64_t n, uint8_t* d_out);
int osr_transit_states_device(osr_world* w, const double* d_from, const double* d_to, int64_t n,
                              double vd, double* d_last_free, double* d_first_hit, uint8_t* d_is_free,
                              double* d_free_tau, double* d_touch_tau, int64_t* d_steps);
int osr_is_valid_transitions_device(osr_world* w, const double* d_from, const double* d_to, int64_t n,
                                    double initial_verify_delta, uint8_t* d_out);
/* Allocation-free variant of osr_is_valid_transitions_device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  Same kernel and labels; the kernel's scratch (work-queue word, per-warp spill
 * areas of the traversal stack: 1 024 entries per warp instead of 8 192) lives in a static device array of
 * 16 MiB + 256 B that every GPU loading the library holds until exit.  The call allocates nothing (no cudaMalloc /
 * cudaFree, no allocating library routine); at most OSR_NOALLOC_MAX_SEGMENTS segments per call, more is OSR_EINVAL.
 * All calls on one GPU share the array: calls that overlap on the GPU (two handles, two processes' worth of streams)
 * race -- serialise them.  Launch only, like the other *_device calls: osr_sync() surfaces a spill overflow. */
#define OSR_NOALLOC_MAX_SEGMENTS (1 << 22)
int osr_is_valid_transitions_device_noalloc(osr_world* w, const double* d_from, const double* d_to, int64_t n,
                                            double initial_verify_delta, uint8_t* d_out);
void* osr_stream(osr_world* w);   /* cudaStream_t of device slot 0 */
int osr_sync(osr_world* w);
/* elapsed GPU milliseconds of the most recent *_device launch on slot 0 (CUDA events recorded
 * around the kernel on the launching stream); valid after osr_sync(). */
int osr_last_kernel_ms(osr_world* w, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* OSRGPU_H */

// osrgpu.hpp -- header-only C++ host side over the C ABI of include/osrgpu.h, for C++ callers of the reference.
//
// Mirrors osr::UnitWorld (/root/reference/lib/osr/unit_world.h:25-380) for the hot path: the same member names,
// argument meaning, return tuples and error behaviour (std::runtime_error with the library's message), plus the
// batched members (one device call per array) that the hot path is about.  Needs a C++14 compiler and libosrgpu.so at
// link time; no CUDA toolkit, no Eigen.  The reference's Eigen types map as
//     StateVector   (Eigen::Matrix<double,7,1>)   -> std::array<double,7>   [tx ty tz qw qx qy qz], osr_state.h:20-33
//     ArrayOfStates (Eigen::Matrix<double,-1,7>)  -> std::vector<StateVector>   (one state per row, contiguous)
//     StateTrans    (Eigen::Vector3d)             -> std::array<double,3>
// A maintainer with Eigen passes `Eigen::Matrix<double,-1,7,Eigen::RowMajor>::data()` to the pointer overloads.
// There is no CPU path: without a CUDA device angleModel() throws and so does every query.
// tests/test_cxx_host.py builds tests/cxx/host_check.cc against this header with g++ and runs it on the GPU.
#ifndef OSRGPU_HPP
#define OSRGPU_HPP
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "osrgpu.h"

namespace osrgpu {

using StateVector = std::array<double, 7>;
using ArrayOfStates = std::vector<StateVector>;
using StateTrans = std::array<double, 3>;
using AngleAxisVector = std::array<double, 3>;
static_assert(sizeof(StateVector) == 7 * sizeof(double), "states must be 7 packed doubles");

constexpr uint32_t GEO_ENV = 0;   // unit_world.h:27 (UnitWorld::GEO_ENV / GEO_ROB)
constexpr uint32_t GEO_ROB = 1;

inline void check(int rc) {
  if (rc != OSR_OK) throw std::runtime_error(std::string("osrgpu: ") + osr_last_error());
}

inline double distance(const StateVector& a, const StateVector& b) { return osr_distance(a.data(), b.data()); }
inline StateVector interpolate(const StateVector& a, const StateVector& b, double tau) {
  StateVector o;
  osr_interpolate(a.data(), b.data(), tau, o.data());
  return o;
}

// What transitStatesToWithContact returns for n segments (the reference's 5-tuple, one array per member).
struct ContactWalks {
  ArrayOfStates last_free, first_hit;
  std::vector<uint8_t> is_free;
  std::vector<double> free_tau, touch_tau;
};

// CSR rows of intersectingSegments over a batch: rows offsets[i]..offsets[i+1] belong to state i.
struct ContactRows {
  std::vector<int64_t> offsets;
  std::vector<std::array<double, 6>> segments;   // two end points
  std::vector<std::array<int32_t, 2>> faces;     // (env triangle, robot triangle)
  std::vector<uint8_t> coplanar;
};

class UnitWorld {
 public:
  // devices: the GPUs to replicate the geometry on (empty = the current device)
  explicit UnitWorld(std::vector<int> devices = {}) : devices_(std::move(devices)) { check(osr_create(&w_)); }
  ~UnitWorld() { if (w_) osr_destroy(w_); }
  UnitWorld(const UnitWorld&) = delete;
  UnitWorld& operator=(const UnitWorld&) = delete;
  osr_world* handle() const { return w_; }

  // ---- set-up (unit_world.cc:94-176) ----
  void loadModelFromFile(const std::string& fn) { check(osr_load_env_obj(w_, fn.c_str())); }
  void loadRobotFromFile(const std::string& fn) { check(osr_load_rob_obj(w_, fn.c_str())); }
  void enforceRobotCenter(const StateTrans& c) { check(osr_enforce_robot_center(w_, c.data())); }
  void scaleToUnit() { scaled_ = true; }   // the scale is computed with the unit meshes in angleModel()
  void angleModel(float latitude, float longitude) {
    if (latitude != 0.0f || longitude != 0.0f)   // the reference's own check (unit_world.cc:143-147)
      throw std::runtime_error("The arguments of UnitWorld::angleModel are resevered for compatibility reason. "
                               "Both values must be 0.");
    check(osr_finalize(w_, devices_.empty() ? nullptr : devices_.data(), (int)devices_.size()));
  }
  void setPerturbation(const StateVector& pert) { check(osr_set_perturbation(w_, pert.data())); }
  double getSceneScale() const { return osr_scene_scale(w_); }
  double getRecommendedCres() const { return osr_get_recommended_cres(w_); }
  void setRecommendedCres(double v) { check(osr_set_recommended_cres(w_, v)); }
  StateTrans getOMPLCenter(uint32_t geo = GEO_ROB) const {
    StateTrans c;
    check(osr_get_ompl_center(w_, (int)geo, c.data()));
    return c;
  }

  // ---- one state / one segment (unit_world.cc:215-264, 405-506) ----
  bool isValid(const StateVector& state) const {
    uint8_t o = 0;
    check(osr_is_valid_states(w_, state.data(), 1, &o));
    return o != 0;
  }
  bool isDisentangled(const StateVector& state) const {
    uint8_t o = 0;
    check(osr_is_disentangled_states(w_, state.data(), 1, &o));
    return o != 0;
  }
  std::tuple<StateVector, StateVector, bool, float, float> transitStateToWithContact(const StateVector& from,
                                                                                     const StateVector& to,
                                                                                     double verify_delta) const {
    StateVector lf, fh;
    uint8_t isf = 0;
    double ft = 0, tt = 0;
    check(osr_transit_states_to_with_contact(w_, from.data(), to.data(), 1, &verify_delta, 1, lf.data(), fh.data(),
                                             &isf, &ft, &tt, nullptr));
    return std::make_tuple(lf, fh, isf != 0, (float)ft, (float)tt);   // the reference's taus are C floats
  }
  std::tuple<StateVector, bool, float> transitStateTo(const StateVector& from, const StateVector& to,
                                                      double verify_delta) const {
    auto t = transitStateToWithContact(from, to, verify_delta);
    return std::make_tuple(std::get<0>(t), std::get<2>(t), std::get<3>(t));
  }
  bool isValidTransition(const StateVector& from, const StateVector& to, double initial_verify_delta) const {
    uint8_t o = 0;
    check(osr_is_valid_transitions(w_, from.data(), to.data(), 1, initial_verify_delta, &o));
    return o != 0;
  }
  // unit_world.cc:489-506: target = (from.t + tr, AngleAxis(|aa|, aa/|aa|) * from.q), NOT renormalised
  std::tuple<StateVector, bool, float> transitStateBy(const StateVector& from, const StateTrans& tr,
                                                      const AngleAxisVector& aa, double verify_delta) const {
    const double ang = std::sqrt(aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2]);
    const double den = ang > 0 ? ang : 1.0;   // Eigen's normalized() leaves a zero vector unchanged
    const double s = std::sin(0.5 * ang);
    const double w0 = std::cos(0.5 * ang), x0 = s * (aa[0] / den), y0 = s * (aa[1] / den), z0 = s * (aa[2] / den);
    const double w1 = from[3], x1 = from[4], y1 = from[5], z1 = from[6];
    StateVector to = {from[0] + tr[0], from[1] + tr[1], from[2] + tr[2],
                      w0 * w1 - x0 * x1 - y0 * y1 - z0 * z1, w0 * x1 + x0 * w1 + y0 * z1 - z0 * y1,
                      w0 * y1 + y0 * w1 + z0 * x1 - x0 * z1, w0 * z1 + z0 * w1 + x0 * y1 - y0 * x1};
    return transitStateTo(from, to, verify_delta);
  }

  // ---- batched members: one device call per array (pyosr's batched surface, lib/pyosr/module.cc:138-250) ----
  std::vector<uint8_t> isValidStates(const double* q, int64_t n) const {
    std::vector<uint8_t> o((size_t)n);
    check(osr_is_valid_states(w_, q, n, o.data()));
    return o;
  }
  std::vector<uint8_t> isValidStates(const ArrayOfStates& qs) const {
    return isValidStates(qs.empty() ? nullptr : qs[0].data(), (int64_t)qs.size());
  }
  std::vector<uint8_t> isDisentangledStates(const ArrayOfStates& qs) const {
    std::vector<uint8_t> o(qs.size());
    check(osr_is_disentangled_states(w_, qs.empty() ? nullptr : qs[0].data(), (int64_t)qs.size(), o.data()));
    return o;
  }
  std::vector<uint8_t> isValidTransitions(const ArrayOfStates& from, const ArrayOfStates& to,
                                          double initial_verify_delta) const {
    same_rows(from, to);
    std::vector<uint8_t> o(from.size());
    check(osr_is_valid_transitions(w_, ptr(from), ptr(to), (int64_t)from.size(), initial_verify_delta, o.data()));
    return o;
  }
  ContactWalks transitStatesToWithContact(const ArrayOfStates& from, const ArrayOfStates& to,
                                          double verify_delta) const {
    same_rows(from, to);
    const size_t n = from.size();
    ContactWalks r;
    r.last_free.resize(n); r.first_hit.resize(n); r.is_free.resize(n); r.free_tau.resize(n); r.touch_tau.resize(n);
    check(osr_transit_states_to_with_contact(w_, ptr(from), ptr(to), (int64_t)n, &verify_delta, 1,
                                             n ? r.last_free[0].data() : nullptr, n ? r.first_hit[0].data() : nullptr,
                                             r.is_free.data(), r.free_tau.data(), r.touch_tau.data(), nullptr));
    return r;
  }
  // calculateVisibilityMatrix2 (unit_world.cc:764-797): out[m][n], row-major
  std::vector<int8_t> calculateVisibilityMatrix2(const ArrayOfStates& qs0, bool qs0_is_unit, const ArrayOfStates& qs1,
                                                 bool qs1_is_unit, double verify_magnitude) const {
    std::vector<int8_t> o(qs0.size() * qs1.size());
    check(osr_visibility_matrix2(w_, ptr(qs0), (int64_t)qs0.size(), qs0_is_unit, ptr(qs1), (int64_t)qs1.size(),
                                 qs1_is_unit, verify_magnitude, o.data()));
    return o;
  }
  // calculateVisibilityPair (unit_world.cc:799-840)
  std::vector<int8_t> calculateVisibilityPair(const ArrayOfStates& qs0, bool qs0_is_unit, const ArrayOfStates& qs1,
                                              bool qs1_is_unit, double verify_magnitude) const {
    std::vector<int8_t> o(std::max(qs0.size(), qs1.size()));
    check(osr_visibility_pair(w_, ptr(qs0), (int64_t)qs0.size(), qs0_is_unit, ptr(qs1), (int64_t)qs1.size(),
                              qs1_is_unit, verify_magnitude, o.data()));
    return o;
  }
  // intersectingSegments (unit_world.cc:959-1012) for a batch of unit states; `cap` bounds the rows kept
  ContactRows intersectingSegments(const ArrayOfStates& qs, int64_t cap = 1 << 20) const {
    ContactRows r;
    const size_t n = qs.size();
    r.offsets.resize(n + 1);
    r.segments.resize((size_t)cap); r.faces.resize((size_t)cap); r.coplanar.resize((size_t)cap);
    std::vector<uint8_t> over(n);
    int64_t total = 0;
    check(osr_intersecting_segments(w_, ptr(qs), (int64_t)n, cap, r.offsets.data(), r.segments[0].data(),
                                    r.faces[0].data(), r.coplanar.data(), nullptr, over.data(), &total));
    for (uint8_t o : over)
      if (o) throw std::runtime_error("osrgpu: intersectingSegments: more rows than `cap`");
    const size_t rows = (size_t)r.offsets[n];
    r.segments.resize(rows); r.faces.resize(rows); r.coplanar.resize(rows);
    return r;
  }
  // the reference's single-state form (unit_world.h:221-227): segment beginnings, ends, magnitudes, (env, rob) faces
  std::tuple<std::vector<std::array<double, 3>>, std::vector<std::array<double, 3>>, std::vector<double>,
             std::vector<std::array<int32_t, 2>>>
  intersectingSegments(const StateVector& unitq) const {
    ContactRows r = intersectingSegments(ArrayOfStates{unitq}, 1 << 16);
    const size_t m = r.faces.size();
    std::vector<std::array<double, 3>> b(m), e(m);
    std::vector<double> mag(m);
    for (size_t i = 0; i < m; i++) {
      const auto& sg = r.segments[i];
      b[i] = {sg[0], sg[1], sg[2]};
      e[i] = {sg[3], sg[4], sg[5]};
      const double dx = sg[3] - sg[0], dy = sg[4] - sg[1], dz = sg[5] - sg[2];
      mag[i] = std::sqrt(dx * dx + dy * dy + dz * dz);
    }
    return std::make_tuple(std::move(b), std::move(e), std::move(mag), std::move(r.faces));
  }

  // ---- state-frame maps (unit_world.cc:526-591, 689-738) ----
  StateVector translateToUnitState(const StateVector& q) const { return map1(osr_translate_to_unit, q); }
  StateVector translateFromUnitState(const StateVector& q) const { return map1(osr_translate_from_unit, q); }
  ArrayOfStates translateOMPLStateToUnitState(const ArrayOfStates& qs) const { return mapn(osr_translate_ompl_to_unit, qs); }
  ArrayOfStates translateUnitStateToOMPLState(const ArrayOfStates& qs) const { return mapn(osr_translate_unit_to_ompl, qs); }

 private:
  static const double* ptr(const ArrayOfStates& a) { return a.empty() ? nullptr : a[0].data(); }
  static void same_rows(const ArrayOfStates& a, const ArrayOfStates& b) {
    if (a.size() != b.size()) throw std::runtime_error("osrgpu: from/to shape mismatch");
  }
  template <class F>
  StateVector map1(F f, const StateVector& q) const {
    StateVector o;
    check(f(w_, q.data(), 1, o.data()));
    return o;
  }
  template <class F>
  ArrayOfStates mapn(F f, const ArrayOfStates& qs) const {
    ArrayOfStates o(qs.size());
    if (!qs.empty()) check(f(w_, qs[0].data(), (int64_t)qs.size(), o[0].data()));
    return o;
  }
  osr_world* w_ = nullptr;
  std::vector<int> devices_;
  bool scaled_ = false;
};

// The batched validators of OmplDriver (/root/reference/lib/pyse3ompl/ompldriver.cc:566-621, setters
// ompldriver.h:107-116) on a UnitWorld: rows are OMPL-frame states x y z qx qy qz qw, the results int 0/1 like the
// reference's Eigen::VectorXi.  validateStates = FCLStateValidityChecker::isValid (bounds AND collision free);
// validateMotionPairs = DiscreteMotionValidator::checkMotion at the resolution given to setCDRes.
class OmplValidators {
 public:
  explicit OmplValidators(const UnitWorld& w) : w_(w) {}
  void setBB(const StateTrans& mins, const StateTrans& maxs) {
    for (int k = 0; k < 3; k++) { bounds_[k] = mins[k]; bounds_[3 + k] = maxs[k]; }
    have_bounds_ = true;
  }
  void setCDRes(double cdres) { cdres_ = cdres; }
  std::vector<int> validateStates(const ArrayOfStates& qs0) const {
    std::vector<uint8_t> o(qs0.size());
    check(osr_ompl_validate_states(w_.handle(), qs0.empty() ? nullptr : qs0[0].data(), (int64_t)qs0.size(),
                                   have_bounds_ ? bounds_ : nullptr, o.data()));
    return std::vector<int>(o.begin(), o.end());
  }
  std::vector<int> validateMotionPairs(const ArrayOfStates& qs0, const ArrayOfStates& qs1) const {
    if (!have_bounds_) throw std::runtime_error("osrgpu: validateMotionPairs needs setBB (the space's extent)");
    const size_t n = std::min(qs0.size(), qs1.size());   // the reference's loop bound
    std::vector<uint8_t> o(n);
    check(osr_ompl_check_motions(w_.handle(), n ? qs0[0].data() : nullptr, n ? qs1[0].data() : nullptr, (int64_t)n,
                                 bounds_, cdres_, o.data(), nullptr));
    return std::vector<int>(o.begin(), o.end());
  }

 private:
  const UnitWorld& w_;
  double bounds_[6] = {0, 0, 0, 0, 0, 0};
  bool have_bounds_ = false;
  double cdres_ = 0.01;   // OMPL's default state-validity-checking resolution (1 % of the extent)
};

}  // namespace osrgpu
#endif  // OSRGPU_HPP

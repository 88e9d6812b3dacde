// host_check.cc -- a C++ caller of the reference's UnitWorld interface, built against include/osrgpu.hpp only
// (g++, no CUDA toolkit, no Eigen) and linked to libosrgpu.so.  tests/test_cxx_host.py compiles it, runs it on the
// GPU and compares what it writes with the Python layer's answers on the same states.
//   usage: host_check <env.obj> <rob.obj> <states.f64> <n> <out.bin> [centre_x centre_y centre_z]
//   usage: host_check --no-device <env.obj> <rob.obj>     (set-up and error behaviour without a CUDA device)
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "osrgpu.hpp"

using namespace osrgpu;

#define REQUIRE(c)                                                        \
  do {                                                                    \
    if (!(c)) {                                                           \
      std::fprintf(stderr, "host_check: %s:%d: %s\n", __FILE__, __LINE__, #c); \
      return 1;                                                           \
    }                                                                     \
  } while (0)

template <class T>
static void put(FILE* f, const T* p, size_t n) {
  if (n && std::fwrite(p, sizeof(T), n, f) != n) std::abort();
}

template <class F>
static bool throws(F f) {
  try {
    f();
  } catch (const std::runtime_error&) {
    return true;
  }
  return false;
}

int main(int argc, char** argv) {
  if (argc >= 4 && !std::strcmp(argv[1], "--no-device")) {
    UnitWorld w;
    REQUIRE(throws([&] { w.loadModelFromFile("/nonexistent/env.obj"); }));
    w.loadModelFromFile(argv[2]);
    w.loadRobotFromFile(argv[3]);
    w.scaleToUnit();
    REQUIRE(throws([&] { w.angleModel(1.0f, 0.0f); }));              // the reference's own argument check
    const bool no_dev = throws([&] { w.angleModel(0.0f, 0.0f); });   // OSR_ENODEV on a box without a GPU
    if (no_dev) {
      REQUIRE(w.getSceneScale() > 0);                                 // host-side geometry is still there
      REQUIRE(throws([&] { w.isValid(StateVector{0, 0, 0, 1, 0, 0, 0}); }));   // and there is no CPU path
    }
    std::printf("no-device checks passed (device %s), scene_scale %.9g\n", no_dev ? "absent" : "present",
                w.getSceneScale());
    return 0;
  }
  if (argc < 6) {
    std::fprintf(stderr, "usage: host_check <env.obj> <rob.obj> <states.f64> <n> <out.bin> [cx cy cz]\n");
    return 2;
  }
  const size_t n = (size_t)std::atoll(argv[4]);
  ArrayOfStates qs(n);
  {
    FILE* f = std::fopen(argv[3], "rb");
    REQUIRE(f && std::fread(qs[0].data(), sizeof(double), 7 * n, f) == 7 * n);
    std::fclose(f);
  }
  UnitWorld w;
  REQUIRE(throws([&] { w.isValid(qs[0]); }));   // queries before set-up fail, they do not guess
  w.loadModelFromFile(argv[1]);
  w.loadRobotFromFile(argv[2]);
  if (argc >= 9) w.enforceRobotCenter(StateTrans{std::atof(argv[6]), std::atof(argv[7]), std::atof(argv[8])});
  w.scaleToUnit();
  w.angleModel(0.0f, 0.0f);

  FILE* out = std::fopen(argv[5], "wb");
  REQUIRE(out);
  const double scale = w.getSceneScale(), cres = w.getRecommendedCres();
  put(out, &scale, 1);
  put(out, &cres, 1);

  // batched validity, and the scalar member on the first 16 states
  std::vector<uint8_t> valid = w.isValidStates(qs);
  for (size_t i = 0; i < 16 && i < n; i++) REQUIRE(w.isValid(qs[i]) == (valid[i] != 0));
  put(out, valid.data(), n);
  std::vector<uint8_t> dis = w.isDisentangledStates(qs);
  REQUIRE(w.isDisentangled(qs[0]) == (dis[0] != 0));
  put(out, dis.data(), n);

  // segments between the free states, in order: walks with first contact, and edge labels
  ArrayOfStates fr, to;
  for (size_t i = 0, last = n; i < n && fr.size() < 256; i++) {
    if (!valid[i]) continue;
    if (last != n) { fr.push_back(qs[last]); to.push_back(qs[i]); }
    last = i;
  }
  const uint64_t m = fr.size();
  put(out, &m, 1);
  const double vd = 0.0125 / 8;
  ContactWalks cw = w.transitStatesToWithContact(fr, to, vd);
  std::vector<uint8_t> edge = w.isValidTransitions(fr, to, vd);
  put(out, m ? fr[0].data() : nullptr, 7 * m);
  put(out, m ? to[0].data() : nullptr, 7 * m);
  put(out, m ? cw.last_free[0].data() : nullptr, 7 * m);
  put(out, m ? cw.first_hit[0].data() : nullptr, 7 * m);
  put(out, cw.is_free.data(), m);
  put(out, cw.free_tau.data(), m);
  put(out, cw.touch_tau.data(), m);
  put(out, edge.data(), m);
  // the scalar members agree with the batch, tuple member by tuple member
  for (size_t i = 0; i < 4 && i < m; i++) {
    auto t = w.transitStateToWithContact(fr[i], to[i], vd);
    REQUIRE(std::get<0>(t) == cw.last_free[i] && std::get<1>(t) == cw.first_hit[i]);
    REQUIRE(std::get<2>(t) == (cw.is_free[i] != 0));
    REQUIRE(std::get<3>(t) == (float)cw.free_tau[i] && std::get<4>(t) == (float)cw.touch_tau[i]);
    auto t3 = w.transitStateTo(fr[i], to[i], vd);
    REQUIRE(std::get<0>(t3) == cw.last_free[i] && std::get<1>(t3) == std::get<2>(t));
    REQUIRE(w.isValidTransition(fr[i], to[i], vd) == (edge[i] != 0));
  }
  // transitStateBy: a pure translation is transitStateTo to the translated state
  if (m) {
    const StateTrans tr = {0.01, -0.02, 0.015};
    StateVector tgt = fr[0];
    for (int k = 0; k < 3; k++) tgt[k] += tr[k];
    auto a = w.transitStateBy(fr[0], tr, AngleAxisVector{0, 0, 0}, vd);
    auto b = w.transitStateTo(fr[0], tgt, vd);
    REQUIRE(a == b);
  }

  // contact rows of the colliding states among the first 64
  ArrayOfStates hit;
  for (size_t i = 0; i < n && i < 64; i++)
    if (!valid[i]) hit.push_back(qs[i]);
  ContactRows rows = w.intersectingSegments(hit);
  const uint64_t nh = hit.size(), nr = rows.faces.size();
  put(out, &nh, 1);
  put(out, &nr, 1);
  put(out, rows.offsets.data(), nh + 1);
  put(out, nr ? rows.faces[0].data() : nullptr, 2 * nr);
  put(out, nr ? rows.segments[0].data() : nullptr, 6 * nr);
  if (nh) {
    auto one = w.intersectingSegments(hit[0]);
    REQUIRE((int64_t)std::get<3>(one).size() == rows.offsets[1] && std::get<2>(one).size() == std::get<3>(one).size());
    for (size_t i = 0; i < std::get<3>(one).size(); i++) REQUIRE(std::get<3>(one)[i] == rows.faces[i]);
  }

  // frame maps round trip through the library
  StateVector u = w.translateToUnitState(qs[0]);
  StateVector back = w.translateFromUnitState(u);
  put(out, u.data(), 7);
  put(out, back.data(), 7);
  ArrayOfStates om = w.translateUnitStateToOMPLState(ArrayOfStates(qs.begin(), qs.begin() + 4));
  ArrayOfStates un = w.translateOMPLStateToUnitState(om);
  put(out, om[0].data(), 28);
  put(out, un[0].data(), 28);

  // OmplDriver's batched validators on the OMPL-frame images of the first 512 states and of the segments
  {
    OmplValidators ov(w);
    const size_t k = std::min<size_t>(n, 512);
    ArrayOfStates oq = w.translateUnitStateToOMPLState(ArrayOfStates(qs.begin(), qs.begin() + k));
    REQUIRE(throws([&] { ov.validateMotionPairs(oq, oq); }));          // no extent yet
    std::vector<int> v0 = ov.validateStates(oq);                       // collision only
    ov.setBB(StateTrans{-100, -100, -100}, StateTrans{100, 100, 100});
    ov.setCDRes(0.002);
    std::vector<int> v1 = ov.validateStates(oq);                       // bounds AND collision
    ArrayOfStates of = w.translateUnitStateToOMPLState(fr), ot = w.translateUnitStateToOMPLState(to);
    std::vector<int> mv = ov.validateMotionPairs(of, ot);
    const uint64_t kk = k;
    put(out, &kk, 1);
    put(out, v0.data(), k);
    put(out, v1.data(), k);
    put(out, mv.data(), (size_t)m);
  }

  REQUIRE(throws([&] { w.isValidTransitions(fr, ArrayOfStates(), vd); }) || m == 0);   // shape mismatch
  std::fclose(out);
  std::printf("host_check: %zu states, %llu segments, %llu contact rows\n", n, (unsigned long long)m,
              (unsigned long long)nr);
  return 0;
}

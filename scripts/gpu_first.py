"""Ad-hoc first GPU check: parity of validity / segments / contacts against the oracle + timing."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O
from puzzletunneldiscovery_b200 import workloads as W, pyosr

A = W.assets_dir()
which = sys.argv[1] if len(sys.argv) > 1 else "org"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
envf, robf = W.REF_ALPHA_FILES[which]

def mk(cls):
    w = cls()
    w.loadModelFromFile(A + "/" + envf); w.loadRobotFromFile(A + "/" + robf)
    if which == "wt2": w.enforceRobotCenter(W.ROB_OMPL_CENTER)
    w.scaleToUnit(); w.angleModel(0.0, 0.0)
    return w
t = time.time(); uw = mk(pyosr.UnitWorld); print("gpu world", time.time() - t, uw.mesh_sizes(0), uw.mesh_sizes(1))
ow = mk(O.OracleWorld)
q = W.random_states(N, 0)
uw.is_valid_states(q[:1024])
uw.counters(reset=True)
t = time.time(); v = uw.is_valid_states(q); dt = time.time() - t
c = uw.counters(reset=True)
print("GPU validity: %d states in %.4fs = %.3e cfg/s ; valid frac %.4f" % (N, dt, N / dt, v.mean()), c)
print("  per pose: bv %.1f leaf %.2f fp64 %.4f" % (c["bv_tests"] / c["poses"], c["leaf_tests"] / c["poses"], c["fp64_rechecks"] / c["poses"]))
M = min(N, 1 << 18)
t = time.time(); vo, cnt = ow.is_valid_states(q[:M], counters=True); dt = time.time() - t
print("oracle: %d in %.3fs = %.3e cfg/s (%d thr); mean bv %.1f leaf %.2f" % (M, dt, M / dt, O.lib().orc_max_threads(), cnt[:, 0].mean(), cnt[:, 1].mean()))
bad = np.nonzero(vo.astype(bool) != v[:M])[0]
print("validity mismatches:", len(bad), bad[:10])
# disentangled
d = uw.is_disentangled_states(q[:M]); print("disentangled frac", d.mean())
# segments
free = q[:M][vo.astype(bool) & ~d][:1024] if which == "org" else None
if which == "wt2":
    tv = np.load(A + "/alpha-1.2.org.tunnel.npz")["TUNNEL_V"]
    free = uw.translate_to_unit_states(tv) if False else np.concatenate([tv[:, :3] * uw.scene_scale, tv[:, 3:]], axis=1)
    print("tunnel valid (gpu):", uw.is_valid_states(free), "oracle:", ow.is_valid_states(free))
S = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 14
fr, to = W.touch_rays(free, S, 1)
vd = 0.0125 / 8
t = time.time(); lf, fh, isf, ft, tt, st = uw.transit_states_to_with_contact(fr, to, vd, with_steps=True); dt = time.time() - t
c = uw.counters(reset=True)
print("GPU segments: %d in %.4fs = %.3e seg/s; steps total %d (%.3e checks/s), free frac %.4f" % (S, dt, S / dt, st.sum(), st.sum() / dt, isf.mean()), c)
So = min(S, 2048)
t = time.time(); olf, ofh, oisf, oft, ott, ocnt = ow.transit_states_to_with_contact(fr[:So], to[:So], vd, counters=True); dt = time.time() - t
print("oracle segments: %d in %.3fs = %.3e seg/s, steps %d" % (So, dt, So / dt, ocnt[:, 0].sum()))
print("  step mismatches:", (ocnt[:, 0] != st[:So]).sum(), "is_free mism", (oisf.astype(bool) != isf[:So]).sum(),
      "max |tau diff|", np.abs(ott - tt[:So]).max(), np.abs(oft - ft[:So]).max(),
      "max state diff", np.abs(olf - lf[:So]).max(), np.abs(ofh - fh[:So]).max())
# transitions
t = time.time(); vt = uw.is_valid_transitions(fr[:So], to[:So], 0.01); dt = time.time() - t
ovt = ow.is_valid_transitions(fr[:So], to[:So], 0.01)
print("transitions mismatches", (ovt.astype(bool) != vt).sum(), "free", vt.mean())
# contacts
hit = fh[~isf][:256]
r = uw.intersecting_segments_batch(hit, cap=1 << 18, with_bary=True)
ooff, osegs, ofaces, ocop, ototal = ow.intersecting_segments_batch(hit, cap=1 << 18)
print("contacts: gpu rows", r["total"], "oracle rows", ototal)
ok = True
for i in range(len(hit)):
    g = r["faces"][r["offsets"][i]:r["offsets"][i + 1]]; o = ofaces[ooff[i]:ooff[i + 1]]
    gs = sorted(map(tuple, g)); os_ = sorted(map(tuple, o))
    if gs != os_:
        ok = False; print(" state", i, "pair sets differ", len(gs), len(os_)); break
print("contact pair multisets equal:", ok)
if ok and ototal:
    # compare segments after canonical ordering
    def canon(off, faces, segs, cop):
        rows = []
        for i in range(len(off) - 1):
            for k in range(off[i], off[i + 1]):
                rows.append((i, faces[k][0], faces[k][1]) + tuple(segs[k]) + (int(cop[k]),))
        return np.array(sorted(rows))
    a = canon(r["offsets"], r["faces"], r["seg"], r["coplanar"]); b = canon(ooff, ofaces, osegs, ocop)
    print("  seg max diff", np.abs(a - b).max())

"""Ad-hoc GPU sanity run (development aid): product vs oracle on small scenes."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import oracle_lib as O
from slam_sam_b200 import synth, ndtb

def cmp_grid(a, b, name):
    ok = np.array_equal(a['idx'], b['idx']) and np.array_equal(a['n'], b['n'])
    print(name, "V", len(a['idx']), len(b['idx']), "idx/n equal:", ok, "meta", np.array_equal(a['min_b'], b['min_b']), np.array_equal(a['div_b'], b['div_b']))
    if ok:
        v = a['n'] >= 1
        print("  mean bit-equal:", np.array_equal(a['mean'], b['mean']), " icov bit-equal(valid):", np.array_equal(a['icov'], b['icov']),
              " cov:", np.array_equal(a['cov'], b['cov']), " evals:", np.array_equal(a['evals'], b['evals']))
        if not np.array_equal(a['icov'], b['icov']):
            d = np.abs(a['icov'] - b['icov']); print("  max icov diff", d.max(), "rel", (d / (np.abs(a['icov']) + 1e-300)).max())

def rel(a, b):
    a = np.asarray(a); b = np.asarray(b)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)

sc = synth.two_plane_scene()
h = ndtb.Handle(0)
h.set_params(resolution=1.0, max_iterations=50, trans_epsilon=1e-4, step_size=0.1)
t = time.time(); rc = h.set_target(sc['target']); print("set_target rc", rc, time.time() - t, h.timings())
h.set_source(sc['source'])
o = O.OracleNdt(resolution=1.0, max_iterations=50, trans_epsilon=1e-4, step_size=0.1, num_threads=8)
o.set_target(sc['target']); o.set_source(sc['source'])
cmp_grid(h.voxels(), o.grid(), "pclomp two-plane")
ro = o.align(sc['guess'].astype(np.float32))
for k in (0, 1, len(ro['trace']) - 1):
    p = ro['trace'][k, :6]
    T = ro['transformation_array'][min(k, len(ro['transformation_array'])-1)] if k == 0 else None
    s, g, H, nv = h.compute_derivatives(p, T=T)
    so, go, Ho, nvo = ro['trace'][k, 6], ro['trace'][k, 7:13], ro['trace'][k, 13:49].reshape(6, 6), ro['trace'][k, 49]
    print("eval", k, "score rel", abs(s - so) / abs(so), "g rel", rel(g, go), "H rel", rel(H, Ho), "nvtl", abs(nv - nvo), "hits", h.last_hits())
t = time.time(); rg = h.align(sc['guess'].astype(np.float32)); print("gpu align", time.time() - t, rg['converged'], rg['iterations'], "oracle", ro['converged'], ro['iterations'], h.timings())
print("final T diff", np.abs(rg['T'] - ro['T']).max())
tg = h.trace(); n = min(len(tg), len(ro['trace']))
print("trace len", len(tg), len(ro['trace']), "max rel score diff over trace", max(abs(tg[i, 6] - ro['trace'][i, 6]) / abs(ro['trace'][i, 6]) for i in range(n)),
      "p diff", np.abs(tg[:n, :6] - ro['trace'][:n, :6]).max())
# svn flavour grid
h2 = ndtb.Handle(0); h2.set_params(grid_flavour=1, min_points_per_voxel=3, resolution=1.0)
print("svn set_target", h2.set_target(sc['target'])); h2.set_source(sc['source'])
os_ = O.OracleSvn(resolution=1.0, min_pts=3, K=4); os_.set_target(sc['target']); os_.set_source(sc['source'])
cmp_grid(h2.voxels(), os_.grid(), "svn two-plane")
P = O.OracleSvn.sample_particles(sc['guess'], 4, 7)
for gn in (1, 0):
    h2.set_svn_params(particle_count=4, use_gauss_newton=gn)
    os2 = O.OracleSvn(resolution=1.0, min_pts=3, K=4, use_gn=bool(gn)); os2.set_target(sc['target']); os2.set_source(sc['source'])
    s, g, H = h2.svn_particle_derivatives(P)
    for k in range(4):
        so, go, Ho = os2.particle_derivatives(P[k])
        print("svn gn", gn, "k", k, "score", abs(s[k] - so) / abs(so), "g", rel(g[k], go), "H", rel(H[k] - 1e-6 * np.eye(6), Ho - 1e-6*np.eye(6)))
    Pn, s2, g2, H2, u2 = h2.svn_step(P)
    Po, so, go, Ho, uo = os2.step(P)
    print("svn step gn", gn, "upd rel", rel(u2, uo), "pose diff", np.abs(Pn - Po).max())
# synthetic ouster
scene = synth.Scene()
t = time.time(); A = synth.trajectory_pose(40.0); B = A @ synth.pose_matrix((0.5, 0.1, 0.02), np.deg2rad((0.5, 0.3, 2.0)))
tgt = synth.ouster_scan(scene, A, seed=1); src = synth.ouster_scan(scene, B, seed=2); print("scan gen", time.time() - t, tgt.shape)
h.set_params(resolution=1.0, max_iterations=35, trans_epsilon=0.01)
print("set_target", h.set_target(tgt), h.timings()); h.set_source(src)
o = O.OracleNdt(resolution=1.0, max_iterations=35, trans_epsilon=0.01, num_threads=8); t = time.time(); o.set_target(tgt); print("oracle build", time.time() - t); o.set_source(src)
cmp_grid(h.voxels(), o.grid(), "pclomp ouster")
t = time.time(); ro = o.align(); print("oracle align", time.time() - t, ro['iterations'], ro['converged'])
t = time.time(); rg = h.align(); print("gpu align", time.time() - t, rg['iterations'], rg['converged'], h.timings(), "hits", h.last_hits())
print("final T diff", np.abs(rg['T'] - ro['T']).max()); print(rg['T']); print(np.linalg.inv(A) @ B)
tg = h.trace(); n = min(len(tg), len(ro['trace']))
for i in range(n):
    print(i, "score rel", abs(tg[i, 6] - ro['trace'][i, 6]) / abs(ro['trace'][i, 6]), "g", rel(tg[i, 7:13], ro['trace'][i, 7:13]), "H", rel(tg[i, 13:49], ro['trace'][i, 13:49]), "p", np.abs(tg[i, :6] - ro['trace'][i, :6]).max())

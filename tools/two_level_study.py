"""CPU study (numpy/scipy, no GPU): PCG iteration counts on the reduced camera system with
block-Jacobi versus block-Jacobi + a rigid-motion coarse space over camera clusters.

usage: python tools/two_level_study.py c3/4 [cams_per_cluster ...]
"""
import os, sys, time
import numpy as np
import scipy.sparse as sp
import scipy.linalg as sl

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle
from visual_marker_mapping_b200 import synth


def blocks(scene, use_gt=False):
    s = scene.at_ground_truth() if use_gt else scene
    ev = oracle.evaluate(s, oracle.LOSS_HUBER, 1.0, gradient=False)
    J = ev.jac_local.reshape(-1, 8, 12)          # per tag obs: 8 rows
    r = ev.residuals.reshape(-1, 8)
    Jc, Jm = J[:, :, :6], J[:, :, 6:]
    Nc, Nm = s.n_cameras, s.n_markers
    Bk = np.einsum("kra,krb->kab", Jc, Jc)
    Ck = np.einsum("kra,krb->kab", Jm, Jm)
    E = np.einsum("kra,krb->kab", Jc, Jm)
    B = np.zeros((Nc, 6, 6)); C = np.zeros((Nm, 6, 6))
    np.add.at(B, s.to_cam, Bk); np.add.at(C, s.to_tag, Ck)
    gc = np.zeros((Nc, 6)); gm = np.zeros((Nm, 6))
    np.add.at(gc, s.to_cam, np.einsum("kra,kr->ka", Jc, r))
    np.add.at(gm, s.to_tag, np.einsum("kra,kr->ka", Jm, r))
    return B, C, E, gc, gm


def kd_clusters(pts, target):
    """balanced kd split of points into clusters of ~target points; returns cluster id per point"""
    ids = np.zeros(len(pts), dtype=np.int64)
    stack = [np.arange(len(pts))]
    out = []
    while stack:
        idx = stack.pop()
        if len(idx) <= target * 1.5:
            out.append(idx); continue
        p = pts[idx]
        ax = int(np.argmax(p.max(0) - p.min(0)))
        order = np.argsort(p[:, ax], kind="stable")
        h = len(idx) // 2
        stack.append(idx[order[:h]]); stack.append(idx[order[h:]])
    for c, idx in enumerate(out):
        ids[idx] = c
    return ids, len(out)


def skew(v):
    z = np.zeros(v.shape[:-1])
    return np.stack([np.stack([z, -v[..., 2], v[..., 1]], -1), np.stack([v[..., 2], z, -v[..., 0]], -1),
                     np.stack([-v[..., 1], v[..., 0], z], -1)], -2)


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c3/4"
    targets = [int(a) for a in sys.argv[2:]] or [100]
    s = synth.make_config(name)
    print(name, s.n_markers, s.n_cameras, s.n_tag_obs, "side", s.meta["site_side_m"], flush=True)
    Nc, Nm = s.n_cameras, s.n_markers
    t0 = time.time()
    state = os.environ.get("STATE", "gt")
    if state.startswith("it"):
        r = oracle.solve(s, max_num_iterations=int(state[2:]), loss_kind=oracle.LOSS_HUBER, loss_scale=1.0)
        print("oracle", r.summary.num_iterations, r.summary.final_cost, time.time() - t0, flush=True)
        s.cam_q, s.cam_t, s.tag_q, s.tag_t = r.cam_q, r.cam_t, r.tag_q, r.tag_t
    B, C, E, gc, gm = blocks(s, use_gt=(state == "gt"))
    print("blocks", time.time() - t0, flush=True)
    active_m = np.ones(Nm, bool); active_m[s.fixed_tag] = False
    # Jacobi scaling + LM damping as in the engine: lam = clamp(s^2 H) / (radius s^2)
    hc = np.einsum("kaa->ka", B); hm = np.einsum("kaa->ka", C)
    sc = 1 / (1 + np.sqrt(hc)); sm = 1 / (1 + np.sqrt(hm))
    R = synth.quat_to_rot(s.cam_q)
    centres = -np.einsum("nji,nj->ni", R, s.cam_t)
    for radius in (1e4, 1e6, 1e8):
        lam_c = np.clip(hc * sc * sc, 1e-6, 1e32) / (radius * sc * sc)
        lam_m = np.clip(hm * sm * sm, 1e-6, 1e32) / (radius * sm * sm)
        Bl = B + np.einsum("ka,ab->kab", lam_c, np.eye(6))
        Cl = C + np.einsum("ka,ab->kab", lam_m, np.eye(6))
        Cinv = np.zeros_like(Cl)
        Cinv[active_m] = np.linalg.inv(Cl[active_m])
        Eb = sp.bsr_matrix((E, s.to_tag.astype(np.int64), np.concatenate([[0], np.cumsum(np.bincount(s.to_cam, minlength=Nc))])),
                           shape=(6 * Nc, 6 * Nm)) if np.all(np.diff(s.to_cam) >= 0) else None
        assert Eb is not None
        Ecsr = Eb.tocsr(); Et = Ecsr.T.tocsr()
        Bd = sp.bsr_matrix((Bl, np.arange(Nc), np.arange(Nc + 1)), shape=(6 * Nc, 6 * Nc)).tocsr()
        Cd = sp.bsr_matrix((Cinv, np.arange(Nm), np.arange(Nm + 1)), shape=(6 * Nm, 6 * Nm)).tocsr()
        Minv = np.linalg.inv(Bl)

        def S(x):
            return Bd @ x - Ecsr @ (Cd @ (Et @ x))

        rhs = -(gc.reshape(-1)) + Ecsr @ (Cd @ gm.reshape(-1))
        # random smooth-ish rhs as well: at ground truth the gradient is noise-only; use it
        def pcg(apply_prec, tol, maxit=5000):
            x = np.zeros_like(rhs); r = rhs.copy(); z = apply_prec(r); p = z.copy(); rz = r @ z; rz0 = rz
            for it in range(1, maxit + 1):
                q = S(p); a = rz / (p @ q); x += a * p; r -= a * q
                z = apply_prec(r); rzn = r @ z
                if rzn <= tol * tol * rz0:
                    return it
                p = z + (rzn / rz) * p; rz = rzn
            return maxit

        def jac(r):
            return np.einsum("kab,kb->ka", Minv, r.reshape(-1, 6)).reshape(-1)

        t0 = time.time()
        res = {"jacobi": [pcg(jac, 1e-2), pcg(jac, 1e-6)]}
        print(f"radius {radius:g} jacobi iters (1e-2, 1e-6): {res['jacobi']}  {time.time()-t0:.1f}s", flush=True)
        for target in targets:
            cid, ncl = kd_clusters(centres[:, :2], target)
            xbar = np.zeros((ncl, 3)); np.add.at(xbar, cid, centres); xbar /= np.bincount(cid, minlength=ncl)[:, None]
            # P_i = [[-R, -R [xbar]x], [0, -R/2]]
            Pi = np.zeros((Nc, 6, 6))
            Pi[:, :3, :3] = -R
            Pi[:, :3, 3:] = -np.einsum("nij,njk->nik", R, skew(xbar[cid] - 0 * centres))
            Pi[:, 3:, 3:] = -0.5 * R
            P = sp.bsr_matrix((Pi, cid, np.arange(Nc + 1)), shape=(6 * Nc, 6 * ncl)).tocsr()
            W = (Et @ P).tocsr()                     # 6Nm x 6ncl
            Ac = (P.T @ (Bd @ P) - W.T @ (Cd @ W)).toarray()
            nnzb = (np.abs(Ac.reshape(ncl, 6, ncl, 6)).sum((1, 3)) > 0).sum()
            cf = sl.cho_factor(Ac)

            def two_add(r):
                return jac(r) + P @ sl.cho_solve(cf, P.T @ r)

            t0 = time.time()
            it2 = [pcg(two_add, 1e-2), pcg(two_add, 1e-6)]
            print(f"   additive two-level, {ncl} clusters (~{target} cams, coarse n={6*ncl}, {nnzb/ncl:.1f} nz blocks/row): {it2}  {time.time()-t0:.1f}s", flush=True)


if __name__ == "__main__":
    main()

"""Quadric index of the accelerated mode (csrc/rt_qindex.cuh): the host classification
and the traversal's node test, driven through the C ABI diagnostics against the oracle's restatement of the
reference's quadric intersector (C/raycast.c:614-656).  The index may only drop a quadric for a ray when the
reference would not report it; these tests hunt for a counterexample with millions of adversarial rays."""
import ctypes as C

import numpy as np
import pytest

import rtlib

FP = C.POINTER(C.c_float)
IP = C.POINTER(C.c_int)


class QInfo(C.Structure):
    _fields_ = [("centre", C.c_double * 3), ("K", C.c_double), ("lmin", C.c_double), ("lmax", C.c_double),
                ("delta", C.c_double * 3), ("ball", C.c_float * 4), ("consts", C.c_float * 5)]


def _lib():
    lib = rtlib.pkg.load_library()
    lib.rt_dbg_qindex_classify.argtypes = [FP, C.POINTER(QInfo)]
    lib.rt_dbg_qindex_prune.argtypes = [FP, FP, FP, C.c_int, C.c_int, FP, FP, IP]
    o = rtlib.oracle()
    o.orc_isect_quadric_batch.argtypes = [FP, FP, FP, C.c_int, IP, FP]
    o.orc_isect_quadric_batch.restype = None
    return lib, o


def _f(a):
    return a.ctypes.data_as(FP)


def classify(lib, q):
    info = QInfo()
    ok = lib.rt_dbg_qindex_classify(_f(q), C.byref(info))
    return info if ok else None


def delta_of(info, s):
    return info.delta[0] + info.delta[1] * s + info.delta[2] * s * s


def prune(lib, info, P, d, nearest=False, best_t=None):
    n = len(P)
    node = np.array(list(info.ball), np.float32)
    k = np.array(list(info.consts), np.float32)
    codes = np.zeros(n, np.int32)
    bt = np.ascontiguousarray(best_t, np.float32) if best_t is not None else np.zeros(n, np.float32)
    lib.rt_dbg_qindex_prune(_f(node), _f(P), _f(d), n, 1 if nearest else 0, _f(bt), _f(k), codes.ctypes.data_as(IP))
    return codes


def isect(o, q, P, d):
    n = len(P)
    hit = np.zeros(n, np.int32)
    t = np.zeros(n, np.float32)
    o.orc_isect_quadric_batch(_f(P), _f(d), _f(q), n, hit.ctypes.data_as(IP), _f(t))
    return hit.astype(bool), t


def make_ellipsoid(rng, centre_scale, axes, rotate):
    """float32 coefficients a..j of the ellipsoid with the given semi-axes, random centre and rotation."""
    c = rng.uniform(-centre_scale, centre_scale, 3)
    R = np.eye(3)
    if rotate:
        R, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    M = R @ np.diag(1.0 / np.asarray(axes, float) ** 2) @ R.T
    s = rng.uniform(0.5, 8.0)                     # overall scale of the equation (the scene generator's K != 1)
    M = M * s
    g = -2.0 * M @ c
    j = c @ M @ c - s
    q = np.array([M[0, 0], M[1, 1], M[2, 2], 2 * M[0, 1], 2 * M[0, 2], 2 * M[1, 2], g[0], g[1], g[2], j], np.float32)
    return q


def exact_parts(q, P, d):
    """A, B, C of q(P + t d) in float64 from the float32 inputs."""
    q = q.astype(np.float64); P = P.astype(np.float64); d = d.astype(np.float64)
    M = np.array([[q[0], q[3] / 2, q[4] / 2], [q[3] / 2, q[1], q[5] / 2], [q[4] / 2, q[5] / 2, q[2]]])
    g = q[6:9]
    A = np.einsum("ni,ij,nj->n", d, M, d)
    B = 2 * np.einsum("ni,ij,nj->n", P, M, d) + d @ g
    Cc = np.einsum("ni,ij,nj->n", P, M, P) + P @ g + q[9]
    return A, B, Cc


def unit(v):
    v = v / np.linalg.norm(v, axis=1, keepdims=True)
    v = v.astype(np.float32)
    # renormalise in float the way the tracer does (lengths within a few ulp of 1)
    l = np.sqrt((v.astype(np.float64) ** 2).sum(1)).astype(np.float32)
    return (v / l[:, None]).astype(np.float32)


def rays_at_offset(rng, P, centre, dist):
    """Directions from P whose line passes at distance `dist` (array) from `centre`."""
    w = centre[None, :] - P.astype(np.float64)
    L = np.linalg.norm(w, axis=1)
    u = w / L[:, None]
    r = rng.normal(size=w.shape)
    r -= (r * u).sum(1)[:, None] * u
    r /= np.linalg.norm(r, axis=1)[:, None]
    s = np.clip(dist / L, 0, 0.999999)
    d = u * np.sqrt(1 - s * s)[:, None] + r * s[:, None]
    return d


CASES = [  # (centre scale, semi-axes, rotated, box)
    (55.0, (0.33, 0.75, 0.89), False, 240.0),     # the C5 generator's ellipsoids
    (55.0, (0.5, 1.0, 2.5), True, 240.0),
    (10.0, (1.0, 1.0, 1.0), False, 40.0),
    (3.0, (0.2, 0.25, 0.3), True, 12.0),
    (400.0, (5.0, 9.0, 20.0), True, 2000.0),
    (0.0, (2.0, 3.0, 4.0), True, 16.0),
]


def test_classify_geometry():
    lib, _ = _lib()
    rng = np.random.default_rng(11)
    for cs, axes, rot, box in CASES:
        for _ in range(20):
            q = make_ellipsoid(rng, cs, axes, rot)
            info = classify(lib, q)
            assert info is not None, (cs, axes)
            q64 = q.astype(np.float64)
            M = np.array([[q64[0], q64[3] / 2, q64[4] / 2], [q64[3] / 2, q64[1], q64[5] / 2], [q64[4] / 2, q64[5] / 2, q64[2]]])
            c = np.linalg.solve(2 * M, -q64[6:9])
            ev = np.linalg.eigvalsh(M)
            assert np.allclose(list(info.centre), c, rtol=1e-9, atol=1e-9)
            assert np.isclose(info.lmin, ev[0], rtol=1e-6) and np.isclose(info.lmax, ev[2], rtol=1e-6)
            K = -(c @ M @ c + q64[6:9] @ c + q64[9])
            assert np.isclose(info.K, K, rtol=1e-6, atol=1e-9 * abs(q64[9]) + 1e-12)
            # the stored ball plus the reach term of the node test holds {q <= delta(s)} for every origin distance s
            off = np.linalg.norm(np.array(list(info.ball)[:3]) - c)
            for sdist in (0.0, 1.0, 10.0, 1e3, 1e6):
                need = np.sqrt((K + delta_of(info, sdist)) / ev[0]) + off
                assert info.ball[3] * 1.0001 + info.consts[0] * (max(sdist - off, 0.0) + info.ball[3]) >= need * (1 - 1e-9), sdist
            assert all(v >= 0 for v in info.consts) and 0 < info.consts[0] < 0.07 and info.consts[3] < 0.125


def test_non_ellipsoids_are_not_indexed():
    lib, _ = _lib()
    bad = [
        [1, 1, 0, 0, 0, 0, 0, 0, 0, -1],          # cylinder
        [1, 1, -1, 0, 0, 0, 0, 0, 0, 0],          # cone
        [1, 1, -1, 0, 0, 0, 0, 0, 0, -1],         # hyperboloid
        [1, 1, 0, 0, 0, 0, 0, 0, -1, 0],          # paraboloid
        [0, 0, 0, 0, 0, 0, 0, 1, 0, 3],           # plane
        [1, 1, 1, 0, 0, 0, 0, 0, 0, 1],           # no real points
        [1, 1, 1, 0, 0, 0, 0, 0, 0, 0],           # a single point
        [-1, -1, -1, 0, 0, 0, 0, 0, 0, 1],        # sign-flipped sphere (negative definite)
        [1, 1, 1e-4, 0, 0, 0, 0, 0, 0, -1],       # needle: condition 1e4
        [1, 1, 1, 2.5, 0, 0, 0, 0, 0, -1],        # indefinite through the cross term
        [float("nan"), 1, 1, 0, 0, 0, 0, 0, 0, -1], [1, 1, 1, 0, 0, 0, float("inf"), 0, 0, -1],
        [1e-9, 1e-9, 1e-9, 0, 0, 0, 0, 0, 0, -1e-9],
    ]
    for q in bad:
        assert classify(lib, np.array(q, np.float32)) is None, q
    assert classify(lib, np.array([1, 1, 1, 0, 0, 0, 0, 0, 0, -1], np.float32)) is not None


def _origins(rng, n, box, centre, r):
    """Origins: uniform in a box, near the ellipsoid, at 6..200 sizes from it, on the box faces, and far outside."""
    m = n // 5
    parts = [rng.uniform(-box, box, (m, 3)),
             centre + rng.normal(size=(m, 3)) * r * rng.uniform(1.0, 6.0, (m, 1)),
             centre + rng.normal(size=(m, 3)) * r * rng.uniform(6.0, 200.0, (m, 1)),
             rng.uniform(-box, box, (m, 3)) * 10.0 ** rng.uniform(0, 3, (m, 1))]
    f = rng.uniform(-box, box, (n - 4 * m, 3))
    ax = rng.integers(0, 3, len(f))
    f[np.arange(len(f)), ax] = box * rng.choice([-1.0, 1.0], len(f))
    parts.append(f)
    return np.ascontiguousarray(np.concatenate(parts).astype(np.float32))


@pytest.mark.timeout(600)
def test_lines_outside_inflated_ellipsoid_are_misses():
    """Claim (1): a line whose minimum of q is above delta is a miss in the reference's arithmetic -- probed right
    at the boundary (qmin between delta and 4 delta), where the bounding sphere never prunes."""
    lib, o = _lib()
    rng = np.random.default_rng(5)
    total = 0
    for cs, axes, rot, box in CASES:
        for _ in range(6):
            q = make_ellipsoid(rng, cs, axes, rot)
            info = classify(lib, q)
            assert info is not None
            c = np.array(list(info.centre))
            P = _origins(rng, 40000, box, c, max(axes))
            # aim at the surface region: offsets spread over the ellipsoid's cross-section sizes, then keep the
            # rays whose exact minimum of q lands in the band
            dist = rng.uniform(0.2 * min(axes), 1.3 * max(axes), len(P))
            d = unit(rays_at_offset(rng, P, c, dist))
            d = np.ascontiguousarray(d)
            A, B, Cc = exact_parts(q, P, d)
            qmin = Cc - B * B / (4 * A)
            dl = delta_of(info, np.linalg.norm(P.astype(np.float64) - c[None, :], axis=1))
            band = (qmin > dl) & (qmin < 4 * dl) & (Cc > dl)
            if not band.any():
                continue
            Pb, db = np.ascontiguousarray(P[band]), np.ascontiguousarray(d[band])
            hit, t = isect(o, q, Pb, db)
            assert not hit.any(), (cs, axes, int(hit.sum()), float(qmin[band][hit].min()))
            total += int(band.sum())
    assert total > 20000


@pytest.mark.timeout(900)
def test_node_test_never_drops_a_reported_quadric():
    """The traversal's test on the item's own ball: 'line misses' and 'behind' only where the reference reports no
    hit with t > 0, 'beyond best_t' never for best_t = the t the reference reports (rays aimed at the rims)."""
    lib, o = _lib()
    rng = np.random.default_rng(9)
    n_pruned = np.zeros(4, np.int64)
    for cs, axes, rot, box in CASES:
        for _ in range(6):
            q = make_ellipsoid(rng, cs, axes, rot)
            info = classify(lib, q)
            assert info is not None
            ball = np.array(list(info.ball), np.float64)
            c, r = ball[:3], ball[3]
            P = _origins(rng, 60000, box, c, max(axes))
            n = len(P)
            L = np.linalg.norm(c[None, :] - P.astype(np.float64), axis=1)
            reach = L + r
            # offsets: inside, on the rim of the ball and of the margin the test adds, and well outside
            rim = r * 1.0001 + info.consts[0] * reach
            kind = rng.integers(0, 5, n)
            dist = np.where(kind == 0, rng.uniform(0, r, n),
                   np.where(kind == 1, rim * rng.uniform(0.999, 1.001, n),
                   np.where(kind == 2, rng.uniform(0.5 * min(axes), 1.2 * max(axes), n),
                   np.where(kind == 3, rim * rng.uniform(1.0, 1.2, n), rng.uniform(0, 4 * r, n)))))
            d = rays_at_offset(rng, P, c, dist)
            flip = rng.random(n) < 0.35                              # pointing away: the "behind" rule
            d[flip] *= -1.0
            side = rng.random(n) < 0.1                               # nearly perpendicular to the line of sight
            if side.any():
                w = (c[None, :] - P[side].astype(np.float64)) / L[side, None]
                p = rng.normal(size=w.shape); p -= (p * w).sum(1)[:, None] * w; p /= np.linalg.norm(p, axis=1)[:, None]
                d[side] = p + w * rng.normal(size=(int(side.sum()), 1)) * 0.02
            d = np.ascontiguousarray(unit(d))
            hit, t = isect(o, q, P, d)
            reported = hit & (t > 0) & np.isfinite(t)
            codes = prune(lib, info, P, d)
            bad = reported & (codes != 0)
            assert not bad.any(), (cs, axes, codes[bad][:5], t[bad][:5], P[bad][:2], d[bad][:2])
            for k in (1, 2):
                n_pruned[k] += int((codes == k).sum())
            # nearest-hit search: with best_t = the reported t itself the node must still be entered (an equal t
            # with a lower index wins the tie, C/raycast.c:456), and for any larger best_t as well
            idx = np.nonzero(reported)[0]
            if len(idx):
                Pi, di = np.ascontiguousarray(P[idx]), np.ascontiguousarray(d[idx])
                for scale in (1.0, 1.0 + 1e-6, 1.5):
                    cn = prune(lib, info, Pi, di, nearest=True, best_t=(t[idx] * np.float32(scale)).astype(np.float32))
                    assert not (cn != 0).any(), (cs, axes, scale)
                # and it does prune when the best hit is far in front of the ball
                cn = prune(lib, info, Pi, di, nearest=True, best_t=np.full(len(idx), 1e-3, np.float32))
                n_pruned[3] += int((cn == 3).sum())
    assert n_pruned[1] > 10000 and n_pruned[2] > 10000 and n_pruned[3] > 1000, n_pruned


def _scene_quadrics(path):
    sc = rtlib.pkg.load_scene(path)
    qs = [list(sc.shapes[i].params)[:10] for i in range(sc.num_shapes) if sc.shapes[i].type == 2]
    sph = [list(sc.shapes[i].position) + [sc.shapes[i].params[0]] for i in range(sc.num_shapes) if sc.shapes[i].type == 0]
    lights = [list(sc.lights[i].position) for i in range(sc.num_lights)]
    return np.array(qs, np.float32), np.array(sph, np.float64), np.array(lights, np.float64)


@pytest.mark.timeout(900)
@pytest.mark.parametrize("scene,n_rays", [("scenes/c5_syn4096.scene", 1500), ("synthetic", 6000)])
def test_hierarchy_visits_every_reported_quadric(scene, n_rays):
    """The whole index as rt_scene_upload builds it (classification, hierarchy, per-node margins, leaf order), walked
    on the host for rays like the tracer's: every quadric the reference reports with t > 0 is visited (shadow
    queries), and the nearest one is visited even when the search already holds its t (nearest-hit queries)."""
    import os
    lib, o = _lib()
    lib.rt_dbg_qindex_visit.argtypes = [FP, C.c_int, FP, FP, C.c_int, C.c_int, FP, C.POINTER(C.c_ubyte)]
    if scene == "synthetic":     # rotated ellipsoids of mixed sizes plus quadrics the index cannot hold
        r0 = np.random.default_rng(77)
        qs = [make_ellipsoid(r0, 30.0, r0.uniform(0.3, 3.0, 3), True) for _ in range(150)]
        qs += [np.array(v, np.float32) for v in ([1, 1, 0, 0, 0, 0, -8, 4, 0, 19], [1, -1, 1, 0, 0, 0, 0, 6, 0, -9],
                                                 [0.5, 0, 0.5, 0, 0, 0, 0, -1, 0, 2])]
        qs = np.array(qs, np.float32)
        sph = np.zeros((0, 4))
        lights = r0.uniform(-40, 40, (6, 3))
    else:
        qs, sph, lights = _scene_quadrics(os.path.join(rtlib.ROOT, scene))
    nq = len(qs)
    if nq < 8:
        pytest.skip("scene has too few quadrics for an index")
    rng = np.random.default_rng(21)
    infos = [classify(lib, q) for q in qs]
    cen = np.array([list(i.centre) if i else [0, 0, 0] for i in infos])
    rad = np.array([i.ball[3] if i else 1.0 for i in infos])
    # origins: the camera, points on the quadrics' bounding spheres and on the spheres, points far out on a floor
    m = n_rays // 4
    k = rng.integers(0, nq, m)
    u = rng.normal(size=(m, 3)); u /= np.linalg.norm(u, axis=1)[:, None]
    parts = [np.zeros((m, 3)), cen[k] + u * rad[k][:, None] * rng.uniform(0.7, 1.5, (m, 1))]
    if len(sph):
        k2 = rng.integers(0, len(sph), m)
        u2 = rng.normal(size=(m, 3)); u2 /= np.linalg.norm(u2, axis=1)[:, None]
        parts.append(sph[k2, :3] + u2 * sph[k2, 3:4])
    far = rng.uniform(-1, 1, (n_rays - sum(len(p) for p in parts), 3)) * np.array([600.0, 5.0, 600.0]) + np.array([0, -40.0, 0])
    parts.append(far)
    P = np.ascontiguousarray(np.concatenate(parts).astype(np.float32))
    n = len(P)
    # directions: at a light, at a quadric (centre or rim), or anywhere
    tgt = np.where(rng.random((n, 1)) < 0.4, lights[rng.integers(0, len(lights), n)] if len(lights) else 0.0,
                   cen[rng.integers(0, nq, n)] + rng.normal(size=(n, 3)) * rad[rng.integers(0, nq, n)][:, None] * 0.7)
    d = tgt - P
    rnd = rng.random(n) < 0.2
    d[rnd] = rng.normal(size=(int(rnd.sum()), 3))
    d = np.ascontiguousarray(unit(d))
    hits = np.zeros((n, nq), bool)
    ts = np.full((n, nq), np.inf, np.float32)
    for kq in range(nq):
        h, t = isect(o, np.ascontiguousarray(qs[kq]), P, d)
        rep = h & (t > 0) & np.isfinite(t)
        hits[:, kq] = rep
        ts[rep, kq] = t[rep]
    visited = np.zeros((n, nq), np.uint8)
    n_idx = lib.rt_dbg_qindex_visit(_f(np.ascontiguousarray(qs)), nq, _f(P), _f(d), n, 0, None, visited.ctypes.data_as(C.POINTER(C.c_ubyte)))
    assert n_idx == sum(i is not None for i in infos) or n_idx > 0
    missed = hits & (visited == 0)
    assert not missed.any(), (int(missed.sum()), np.argwhere(missed)[:5])
    assert hits.sum() > n // 2                                # the rays do hit things
    assert visited.sum() < 0.25 * n * nq or n_idx < nq // 2   # and the index does prune
    best = ts.min(axis=1)
    has = np.isfinite(best)
    Pn, dn, bn = np.ascontiguousarray(P[has]), np.ascontiguousarray(d[has]), np.ascontiguousarray(best[has])
    vis2 = np.zeros((len(Pn), nq), np.uint8)
    lib.rt_dbg_qindex_visit(_f(np.ascontiguousarray(qs)), nq, _f(Pn), _f(dn), len(Pn), 1, _f(bn), vis2.ctypes.data_as(C.POINTER(C.c_ubyte)))
    winners = ts[has] == bn[:, None]
    assert not (winners & (vis2 == 0)).any()
    assert vis2.sum() <= visited[has].sum()


@pytest.mark.timeout(600)
def test_lopsided_hierarchy_stays_within_depth_and_exact():
    """Ellipsoids in a geometric progression (each as large as all the smaller ones together): the sweep split peels
    them off one at a time, the builder has to switch to median splits before the traversal stack (32 entries) could
    overflow, and the index still visits everything the reference reports."""
    lib, o = _lib()
    lib.rt_dbg_qindex_visit.argtypes = [FP, C.c_int, FP, FP, C.c_int, C.c_int, FP, C.POINTER(C.c_ubyte)]
    qs = []
    for k in range(70):
        x, r = 1.12 ** k, 0.2 * 1.12 ** k
        s = 1.0 / (r * r)
        qs.append([s, s, s * 1.3, 0, 0, 0, -2 * s * x, 0, 0, s * x * x - 1.0])
    qs = np.array(qs, np.float32)
    nq = len(qs)
    rng = np.random.default_rng(2)
    n = 1500
    P = np.zeros((n, 3), np.float32)
    P[n // 2:] = (rng.normal(size=(n - n // 2, 3)) * 10.0 ** rng.uniform(-1, 5, (n - n // 2, 1))).astype(np.float32)
    d = np.array([1.0, 0, 0]) + rng.normal(size=(n, 3)) * 10.0 ** rng.uniform(-4, -0.5, (n, 1))
    d = np.ascontiguousarray(unit(d))
    hits = np.zeros((n, nq), bool)
    for kq in range(nq):
        h, t = isect(o, np.ascontiguousarray(qs[kq]), P, d)
        hits[:, kq] = h & (t > 0) & np.isfinite(t)
    visited = np.zeros((n, nq), np.uint8)
    n_idx = lib.rt_dbg_qindex_visit(_f(np.ascontiguousarray(qs)), nq, _f(P), _f(d), n, 0, None, visited.ctypes.data_as(C.POINTER(C.c_ubyte)))
    assert n_idx == nq                                       # all indexed, hierarchy built (depth within the limit)
    assert not (hits & (visited == 0)).any()
    assert hits.sum() > n

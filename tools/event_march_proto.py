"""Design prototype (numpy, CPU): an EVENT-DRIVEN form of tracing algorithm 3 (compute.glsl:34-68) for large scenes,
checked bit for bit against the eager form on sampled rays of the 10,000-sphere scene, with its work counted.

The eager algorithm touches every object on every march trip (d[i] -= last; compare; min).  The event-driven form
keeps, per object, the exact cached value d0[i] as of trip j0[i] and defers the subtractions.  What makes that exact:

  * the chain d <- fl(d - last_k) rounds at most 2^-24 |d0| per step (its values stay in [0, d0] up to rounding), so after
    J deferred steps   |chain - (d0 - sum of the J last values)| <= J 2^-24 |d0| (+ the rounding of the sum, kept in
    float64);
  * a refresh whose outcome is SURE (upper bound <= radius) replaces the value by sphereSDF(position): the deferred
    chain never needs evaluating;  sure-not-due objects (lower bound > radius) that cannot be the minimum (lower bound
    > an upper bound of the minimum) are not touched at all;
  * only the ambiguous ones and the candidates of the minimum replay their deferred subtractions exactly.

Usage: python tools/event_march_proto.py [rays]     (no GPU, no oracle: the eager form is restated here in numpy)
"""
import sys
import os
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracing_engine_b200 import camera, scene
from raytracing_engine_b200.engine import host_level_dims

f32 = np.float32
RD = f32(1000.0)
sph = scene.synthetic_spheres(10000, seed=0xB200)
C = sph[:, :3].astype(f32)
R = sph[:, 3].astype(f32)
N = len(R)


def sdf_some(p, idx):
    d = p[None, :] - C[idx]
    return np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]).astype(f32) - R[idx]


ALL = np.arange(N)


def eager(origin, dirv, thr):
    d = sdf_some(origin, ALL)
    ln, last, trips, evals = f32(0), f32(0), 0, N
    while ln < RD:
        p = origin + dirv * ln
        radius = (ln + f32(1)) * thr
        d = d - last
        due = np.nonzero(d <= radius)[0]
        if len(due):
            d[due] = sdf_some(p, due)
            evals += len(due)
        dist = min(RD, d.min())
        trips += 1
        last = max(dist, f32(0))
        ln = f32(ln + last)
        if dist <= radius:
            ln = f32(ln - radius)
            break
    return ln, trips, evals


def event_driven(origin, dirv, thr):
    d0 = sdf_some(origin, ALL)            # exact cached value as of trip j0
    j0 = np.zeros(N, np.int64)
    hist = [f32(0)]                       # hist[k] = `last` subtracted at trip k (k = 1, 2, ...); trip numbering from 1
    Ld = [0.0]                            # Ld[k] = sum of hist[1..k] in float64
    ln, last, k = f32(0), f32(0), 0
    stats = dict(trips=0, refresh_sure=0, replayed_objects=0, replayed_subtractions=0, evals=N)
    EPS = 2.0 ** -24

    def replay(idx):
        """Exact chain for objects idx from their j0 to trip k."""
        for i in idx:
            v = d0[i]
            for t in range(j0[i] + 1, k + 1):
                v = f32(v - hist[t])
            stats["replayed_subtractions"] += k - j0[i]
            d0[i] = v
            j0[i] = k
        stats["replayed_objects"] += len(idx)

    while ln < RD:
        k += 1
        hist.append(last)
        Ld.append(Ld[-1] + float(last))
        p = origin + dirv * ln
        radius = (ln + f32(1)) * thr
        J = k - j0
        est = d0.astype(np.float64) - (Ld[k] - np.asarray(Ld)[j0])
        err = (J + 2) * EPS * np.abs(d0.astype(np.float64)) * 1.0001 + 1e-30 + 1e-12 * Ld[k]
        lo, hi = est - err, est + err
        rad = float(radius)
        sure = np.nonzero(hi <= rad)[0]
        amb = np.nonzero((lo <= rad) & (hi > rad))[0]
        if len(amb):
            replay(amb)
            due_amb = amb[d0[amb] <= radius]
        else:
            due_amb = amb
        due = np.concatenate([sure, due_amb])
        if len(due):
            d0[due] = sdf_some(p, due)
            j0[due] = k
            stats["evals"] += len(due)
            stats["refresh_sure"] += len(sure)
        # exact minimum: objects whose value is exact as of this trip, plus every object whose lower bound does not
        # exceed the smallest upper bound
        exact_now = j0 == k
        upper = np.where(exact_now, d0.astype(np.float64), hi)
        m_up = min(float(RD), upper.min())
        cand = np.nonzero(~exact_now & (lo <= m_up))[0]
        if len(cand):
            replay(cand)
        exact_now = j0 == k
        dist = min(RD, d0[exact_now].min()) if exact_now.any() else RD
        stats["trips"] += 1
        last = max(dist, f32(0))
        ln = f32(ln + last)
        if dist <= radius:
            ln = f32(ln - radius)
            break
    return ln, stats


def run(rays=40, seed=3):
    """Returns (mismatches, per-level totals)."""
    W, H = 1920, 1080
    n = 8
    dims = host_level_dims(n, W, H)
    rots, poss = camera.static_path(1)
    ratio = scene.view_ratio(W, H)
    cam = np.array(poss[0], f32)
    rs = np.random.RandomState(seed)
    tot = {}
    bad = 0
    for s in range(rays):
        X, Y = int(rs.randint(0, W)), int(rs.randint(0, H))
        start = f32(1.0)
        for l in range(n):
            x, y = X >> (n - 1 - l), Y >> (n - 1 - l)
            isx = f32(2.0 ** (n - 1 - l)) / f32(W)
            isy = f32(2.0 ** (n - 1 - l)) / f32(H)
            ncx = (f32(2 * x + 1) * isx - f32(1)) * f32(ratio[0])
            ncy = (f32(2 * y + 1) * isy - f32(1)) * f32(ratio[1])
            v = np.array([ncx, 1, ncy], f32)
            v = v / np.sqrt((v * v).sum()).astype(f32)
            thr = f32(1.4142135) * f32(8) * isx
            origin = cam + v * start
            a, trips, evals = eager(origin, v, thr)
            b, st = event_driven(origin, v, thr)
            if np.float32(a).view(np.uint32) != np.float32(b).view(np.uint32) or trips != st["trips"] or evals != st["evals"]:
                bad += 1
            t = tot.setdefault(l, dict(rays=0, trips=0, evals=0, refresh_sure=0, replayed_objects=0, replayed_subtractions=0))
            t["rays"] += 1
            for key in ("trips", "evals", "refresh_sure", "replayed_objects", "replayed_subtractions"):
                t[key] += st[key]
            start = max(f32(start + a), f32(0))
    return bad, tot, n


def main():
    rays = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    bad, tot, n = run(rays)
    print(f"{rays} ray chains x {n} levels: mismatches against the eager form (len bits, trips, SDF evaluations): {bad}")
    for l in sorted(tot):
        t = tot[l]
        tr = max(t["trips"], 1)
        print(f" level {l}: trips/ray {t['trips'] / t['rays']:.1f}; per trip: refreshes {(t['evals'] - N * t['rays']) / tr:.1f} "
              f"(sure, no replay: {t['refresh_sure'] / tr:.1f}), objects replayed {t['replayed_objects'] / tr:.2f}, "
              f"subtractions replayed {t['replayed_subtractions'] / tr:.1f}  (eager: {N} subtractions per trip)")


if __name__ == "__main__":
    main()

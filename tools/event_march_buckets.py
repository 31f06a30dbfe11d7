"""Design prototype, second step (numpy + Python dicts, CPU): the event-driven march of tools/event_march_proto.py with
its objects held in a CALENDAR QUEUE over the absolute ray length, so that a trip only examines the objects whose
conservative key has come within reach — the data structure the CUDA kernel of the next round would use.

  key_lo[i] = d0[i] + Ld[j0[i]] - e[i]      absolute length before which object i can neither be due nor be the minimum
  e[i]      = (B + 2) 2^-24 |d0[i]| (+ slack)  bound of the deferred chain's rounding for at most B deferred trips
  bucket    = floor(key_lo / DELTA); a trip examines the buckets up to the horizon
              max(Ld_k + radius_k, smallest key_hi in the queue): exact replay -> due? -> refresh -> re-insert.
  every B trips all chains are brought up to date (eager re-base) so that e[] stays valid.

Checked bit for bit against the eager march (final length, trips, SDF evaluations); prints what a trip costs.
Usage: python tools/event_march_buckets.py [rays] [DELTA] [B]
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import event_march_proto as P   # scene, eager march, helpers

f32 = np.float32
EPS = 2.0 ** -24


def bucketed(origin, dirv, thr, delta, B):
    N = P.N
    d0 = P.sdf_some(origin, P.ALL)
    j0 = np.zeros(N, np.int64)
    hist, Ld = [f32(0)], [0.0]
    k = 0
    ln, last = f32(0), f32(0)
    st = dict(trips=0, evals=N, examined=0, replayed_subtractions=0, rebase_subtractions=0, buckets_visited=0)

    def err_of(idx):
        return (B + 2) * EPS * np.abs(d0[idx].astype(np.float64)) * 1.0001 + 1e-30 + 1e-12 * (Ld[-1] + 1000.0)

    key_lo = np.empty(N, np.float64)
    key_hi = np.empty(N, np.float64)
    buckets = {}

    def insert(idx):
        e = err_of(idx)
        base = d0[idx].astype(np.float64) + np.asarray(Ld)[j0[idx]]
        key_lo[idx] = base - e
        key_hi[idx] = base + e
        for i, b in zip(idx, np.floor(key_lo[idx] / delta).astype(np.int64)):
            buckets.setdefault(int(b), []).append(int(i))

    insert(P.ALL)

    def replay(idx):
        for i in idx:
            v = d0[i]
            for t in range(j0[i] + 1, k + 1):
                v = f32(v - hist[t])
            st["replayed_subtractions"] += k - j0[i]
            d0[i] = v
            j0[i] = k

    since_rebase = 0
    while ln < P.RD:
        k += 1
        hist.append(last)
        Ld.append(Ld[-1] + float(last))
        since_rebase += 1
        if since_rebase > B:   # eager re-base: every chain brought up to date, every key rebuilt
            stale = np.nonzero(j0 < k - 1)[0]
            # (replay to k - 1 so that this trip's subtraction is still pending, as for everyone else)
            for i in stale:
                v = d0[i]
                for t in range(j0[i] + 1, k):
                    v = f32(v - hist[t])
                d0[i] = v
            st["rebase_subtractions"] += int((k - 1 - j0[stale]).sum())
            j0[stale] = k - 1
            buckets.clear()
            insert(P.ALL)
            since_rebase = 1
        p = origin + dirv * ln
        radius = (ln + f32(1)) * thr
        rad_abs = Ld[k] + float(radius)
        exact_vals = []          # exact post-refresh values seen this trip
        taken = []
        pending_hi = []          # upper bounds (absolute) of examined objects left untouched this trip
        while True:
            # horizon: due objects, and every object that could still undercut the best upper bound of the minimum
            m_up_abs = min([Ld[k] + float(v) for v in exact_vals] + [Ld[k] + float(P.RD)] + pending_hi)
            if buckets:
                b0 = min(buckets)
                m_up_abs = min(m_up_abs, min(key_hi[i] for i in buckets[b0]))
            horizon = max(rad_abs, m_up_abs)
            todo = [b for b in buckets if b * delta <= horizon]
            if not todo:
                break
            idx = []
            for b in todo:
                idx += buckets.pop(b)
            st["buckets_visited"] += len(todo)
            idx = np.asarray(idx, np.int64)
            st["examined"] += len(idx)
            near = idx[key_lo[idx] <= horizon]
            far = idx[key_lo[idx] > horizon]
            if len(far):             # same bucket, key still beyond the horizon: back into the queue untouched
                for i, b in zip(far, np.floor(key_lo[far] / delta).astype(np.int64)):
                    buckets.setdefault(int(b), []).append(int(i))
                if len(near) == 0:
                    # nothing within reach in these buckets: they stay; stop looking
                    break
            # tight bounds with the ACTUAL number of deferred trips (the bucket keys use the cap B)
            base = d0[near].astype(np.float64) + np.asarray(Ld)[j0[near]]
            e_act = (k - j0[near] + 2) * EPS * np.abs(d0[near].astype(np.float64)) * 1.0001 + 1e-30 + 1e-12 * (Ld[k] + 1000.0)
            lo_t, hi_t = base - e_act, base + e_act
            sure = near[hi_t <= rad_abs]
            amb = near[(lo_t <= rad_abs) & (hi_t > rad_abs)]
            other = (lo_t > rad_abs)
            replay(amb)
            due = np.concatenate([sure, amb[d0[amb] <= radius]])
            if len(due):
                d0[due] = P.sdf_some(p, due)
                j0[due] = k
                st["evals"] += len(due)
            exact_vals += [d0[i] for i in np.concatenate([sure, amb])]
            # candidates of the minimum among the surely-not-due ones
            oth, oth_lo, oth_hi = near[other], lo_t[other], hi_t[other]
            if len(oth):
                m_now = min([Ld[k] + float(v) for v in exact_vals] + [Ld[k] + float(P.RD), float(oth_hi.min())])
                if buckets:
                    m_now = min(m_now, min(key_hi[i] for i in buckets[min(buckets)]))
                cand = oth[oth_lo <= m_now]
                replay(cand)
                exact_vals += [d0[i] for i in cand]
                pending_hi.extend(float(h) for h, c in zip(oth_hi, oth_lo <= m_now) if not c)
            taken += list(near)
        if taken:
            insert(np.asarray(taken, np.int64))
        dist = min([P.RD] + exact_vals)
        st["trips"] += 1
        last = max(dist, f32(0))
        ln = f32(ln + last)
        if dist <= radius:
            ln = f32(ln - radius)
            break
    return ln, st


def run(rays=6, delta=0.5, B=32, seed=5):
    from raytracing_engine_b200 import camera, scene
    from raytracing_engine_b200.engine import host_level_dims
    W, H, n = 1920, 1080, 8
    host_level_dims(n, W, H)
    rots, poss = camera.static_path(1)
    ratio = scene.view_ratio(W, H)
    cam = np.array(poss[0], f32)
    rs = np.random.RandomState(seed)
    tot, bad = {}, 0
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
            a, trips, evals = P.eager(origin, v, thr)
            b, st = bucketed(origin, v, thr, delta, B)
            if np.float32(a).view(np.uint32) != np.float32(b).view(np.uint32) or trips != st["trips"] or evals != st["evals"]:
                bad += 1
            t = tot.setdefault(l, dict(rays=0))
            t["rays"] += 1
            for key, val in st.items():
                t[key] = t.get(key, 0) + val
            start = max(f32(start + a), f32(0))
    return bad, tot, n


def main():
    rays = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    delta = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
    B = int(sys.argv[3]) if len(sys.argv) > 3 else 32
    bad, tot, n = run(rays, delta, B)
    print(f"{rays} ray chains x {n} levels, DELTA = {delta}, re-base every {B} trips: mismatches against the eager form: {bad}")
    for l in sorted(tot):
        t = tot[l]
        tr = max(t["trips"], 1)
        print(f" level {l}: trips/ray {t['trips'] / t['rays']:.1f}; per trip: refreshes {(t['evals'] - P.N * t['rays']) / tr:.1f}, "
              f"objects examined {t['examined'] / tr:.1f} in {t['buckets_visited'] / tr:.1f} buckets, subtractions replayed "
              f"{t['replayed_subtractions'] / tr:.1f} + re-base {t['rebase_subtractions'] / tr:.0f}  (eager: {P.N} per trip)")


if __name__ == "__main__":
    main()
